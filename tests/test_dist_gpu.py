"""Node sharding on real kernels: several shards as separate contexts on one GPU (extrema
combined with a device max, standing in for the all-reduce), and — when the box has two GPUs —
two ranks over NCCL."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 3])
def test_shards_as_contexts_on_one_gpu(pkg, oracle_mod, synth, world):
    import torch
    from k8s_gpu_scheduler_b200 import dist as gsd

    dev = torch.device("cuda", 0)
    preds, cluster, pods = synth.make_workload(500, 211, gpus_per_node=8, n_apps=64)
    P = pods.n_pods
    d_app = torch.from_numpy(pods.app.view(np.int32)).to(dev)
    d_slo = torch.from_numpy(pods.slo).to(dev)
    stream = torch.cuda.Stream(device=dev)
    shards = [gsd.ShardedScheduler(pkg, preds, cluster, 0, rank=r, world=world) for r in range(world)]
    try:
        with torch.cuda.stream(stream):
            outs = [s.alloc_outputs(P) for s in shards]
            for s, o in zip(shards, outs):   # phase 1 on every shard
                s.sched.score_raw_dev(d_app, d_slo, P, o["score"], o["sel"], o["ext"][0], o["ext"][1],
                                      stream=stream.cuda_stream)
            ext = outs[0]["ext"].clone()
            for o in outs[1:]:
                ext = torch.maximum(ext, o["ext"])   # what all_reduce(MAX) computes
            for s, o in zip(shards, outs):   # phase 2 with the global extrema
                s.sched.normalize_dev(P, o["score"], ext[0], ext[1], o["best"], stream=stream.cuda_stream)
            stream.synchronize()
        score = np.concatenate([o["score"].cpu().numpy() for o in outs], axis=1)
        sel = np.concatenate([o["sel"].cpu().numpy() for o in outs], axis=1)
        want = oracle_mod.Oracle(preds, cluster).score_batch(pods, normalize=True)
        assert np.array_equal(score, want[0]) and np.array_equal(sel, want[1])
        for o in outs:
            assert np.array_equal(o["best"].cpu().numpy(), want[2])
    finally:
        for s in shards:
            s.close()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _nccl_worker(rank, world, port, errs):
    try:
        sys.path.insert(0, ROOT)
        import torch
        import torch.distributed as dist

        import __graft_entry__ as entry
        from oracle.score_oracle import Oracle

        pkg = entry.load_package()
        from k8s_gpu_scheduler_b200 import dist as gsd

        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world, device_id=dev)
        preds, cluster, pods = pkg.synth.make_workload(800, 301, gpus_per_node=8, n_apps=64)
        sh = gsd.ShardedScheduler(pkg, preds, cluster, rank, rank=rank, world=world)
        P = pods.n_pods
        d_app = torch.from_numpy(pods.app.view(np.int32)).to(dev)
        d_slo = torch.from_numpy(pods.slo).to(dev)
        stream = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(stream):
            out = sh.alloc_outputs(P)
            sh.score(d_app, d_slo, P, out)
            stream.synchronize()
        want = Oracle(preds, cluster).score_batch(pods, normalize=True)
        assert np.array_equal(out["score"].cpu().numpy(), want[0][:, sh.begin:sh.end])
        assert np.array_equal(out["sel"].cpu().numpy(), want[1][:, sh.begin:sh.end])
        assert np.array_equal(out["best"].cpu().numpy(), want[2])
        sh.close()
        dist.destroy_process_group()
    except Exception as ex:  # noqa: BLE001
        errs.put(f"rank {rank}: {type(ex).__name__}: {ex}")
        raise


def test_two_ranks_nccl():
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    errs = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, errs)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=600)
    msgs = []
    while not errs.empty():
        msgs.append(errs.get())
    assert not msgs, msgs
    assert all(p.exitcode == 0 for p in procs)
