"""World-size-2 gloo test (CPU) of the node-sharding protocol in k8s-gpu-scheduler_b200/dist.py:
shard bounds, the packed per-pod extrema, the single all-reduce(MAX), NormalizeScore on column
shards with global extrema, global best node.  Per-shard raw scores come from the CPU oracle."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_pods, n_nodes, errs):
    try:
        sys.path.insert(0, ROOT)
        import torch
        import torch.distributed as dist

        import __graft_entry__ as entry
        from oracle.score_oracle import Oracle

        pkg = entry.load_package()
        from k8s_gpu_scheduler_b200 import dist as gsd

        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
        preds, cluster, pods = pkg.synth.make_workload(n_pods, n_nodes, gpus_per_node=8, n_apps=40)
        b, e = gsd.shard_bounds(n_nodes, world, rank)
        local = cluster.shard(b, e)
        raw, sel, _ = Oracle(preds, local).score_batch(pods, normalize=False)
        nodes = np.arange(b, e, dtype=np.int64)
        packed = (raw.astype(np.int64) << 32) | (0xFFFFFFFF - nodes)[None, :]
        ext = np.zeros((2, n_pods), dtype=np.int64)
        if e > b:
            ext[0] = packed.max(axis=1)
            ext[1] = (100 - raw.astype(np.int64)).max(axis=1)
        t = torch.from_numpy(ext)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        hi, lo, best = gsd.decode_extrema(t.numpy())
        got = gsd.normalize_columns(raw, hi, lo)
        want_score, want_sel, want_best = Oracle(preds, cluster).score_batch(pods, normalize=True)
        assert np.array_equal(got, want_score[:, b:e]), "normalised shard differs"
        assert np.array_equal(sel, want_sel[:, b:e]), "selection shard differs"
        assert np.array_equal(best, want_best), "best node differs"
        dist.destroy_process_group()
    except Exception as ex:  # noqa: BLE001
        errs.put(f"rank {rank}: {type(ex).__name__}: {ex}")
        raise


@pytest.mark.parametrize("n_nodes", [37, 2, 1])
def test_sharded_protocol_world2_gloo(n_nodes):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    errs = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 60, n_nodes, errs)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    msgs = []
    while not errs.empty():
        msgs.append(errs.get())
    assert not msgs, msgs
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]


def test_shard_bounds_cover_and_balance(pkg):
    from k8s_gpu_scheduler_b200 import dist as gsd

    for n in (0, 1, 7, 8, 9, 5000, 20000):
        for w in (1, 2, 4, 8):
            b = [gsd.shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [e - s for s, e in b]
            assert max(sizes) - min(sizes) <= 1


def test_cluster_shard_rebases_offsets(pkg):
    cl = pkg.synth.make_cluster(50, 16, 8)
    s = cl.shard(13, 29)
    assert s.n_nodes == 16 and s.gpu_off[0] == 0 and s.gpu_off[-1] == s.res_app.size
    lo = int(cl.gpu_off[13 * 8])
    assert np.array_equal(s.res_app, cl.res_app[lo:lo + s.res_app.size])
