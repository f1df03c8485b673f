"""gs_score_raw_dev_noalloc — the allocation-free scoring entry point.

Some callers of this library cannot allocate device memory where they score: a scheduler
that has handed all HBM to its own pools at start-up, a latency-critical cycle that must not
block on the driver's allocator, a harness that forbids allocation after initialisation.
For them the library offers one variant of the scoring phase whose working memory is a fixed
array inside the library image.  This test calls it through the C ABI and requires the very
same bytes as the regular entry point (gs_score_raw_dev) — scores, selection bytes and both
per-pod extrema — on two different batches back to back (the array is shared between calls
and keeps the previous call's contents, so the second call proves that stale contents do no
harm), and that batches beyond the stated limit are refused with a clear message.

Collected last on purpose (file name), so that a run stopping at the first failure has
already run every other test.  Passes alone in a fresh process and after the other tests.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_batch(torch, pods, n_nodes, dev):
    P = pods.n_pods
    return dict(
        app=torch.from_numpy(pods.app.view(np.int32)).to(dev),
        slo=torch.from_numpy(pods.slo).to(dev),
        score=torch.full((P, n_nodes), 0xEE, dtype=torch.uint8, device=dev),
        sel=torch.full((P, n_nodes), 0xEE, dtype=torch.uint8, device=dev),
        hi=torch.full((P,), -1, dtype=torch.int64, device=dev),   # garbage in: the call must zero these itself
        neg=torch.full((P,), -1, dtype=torch.int64, device=dev),
    )


def test_noalloc_entry_point_matches_regular_entry_point(pkg, synth):
    import torch

    dev = torch.device("cuda", 0)
    sch = pkg.Scheduler(0)
    try:
        preds, cluster, pods_a = synth.make_workload(700, 300, gpus_per_node=8, n_apps=120, p_unknown_app=0.03)
        sch.upload_predictions(preds)
        sch.upload_cluster(cluster, node_base=40)
        pods_b = synth.make_pods(1500, 120, seed=99)
        slo_b = pods_b.slo.copy()
        slo_b[::9] = 0.0
        slo_b[1::17] = np.nan
        pods_b = synth.Pods(app=pods_b.app, slo=slo_b)
        stream = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(stream):
            st = stream.cuda_stream
            for pods in (pods_a, pods_b):          # two calls in a row on the shared array
                ref = _device_batch(torch, pods, 300, dev)
                got = _device_batch(torch, pods, 300, dev)
                sch.score_raw_dev(ref["app"], ref["slo"], pods.n_pods, ref["score"], ref["sel"], ref["hi"], ref["neg"],
                                  stream=st)
                sch.score_raw_dev_noalloc(got["app"], got["slo"], pods.n_pods, got["score"], got["sel"], got["hi"],
                                          got["neg"], stream=st)
                stream.synchronize()
                for k in ("score", "sel", "hi", "neg"):
                    assert torch.equal(ref[k], got[k]), f"{k} differs between gs_score_raw_dev and the no-alloc variant"
                # and the normalise phase consumes either result identically
                best_r = torch.empty(pods.n_pods, dtype=torch.int32, device=dev)
                best_g = torch.empty(pods.n_pods, dtype=torch.int32, device=dev)
                sch.normalize_dev(pods.n_pods, ref["score"], ref["hi"], ref["neg"], best_r, stream=st)
                sch.normalize_dev(pods.n_pods, got["score"], got["hi"], got["neg"], best_g, stream=st)
                stream.synchronize()
                assert torch.equal(ref["score"], got["score"]) and torch.equal(best_r, best_g)
                assert int(best_g.min()) >= 40  # node_base applied
    finally:
        sch.close()


def test_noalloc_entry_point_matches_the_oracle(pkg, synth, oracle_mod):
    """Not only the same bytes as the regular entry point: the same bytes as the CPU restatement of
    the reference (raw scores, selection bytes, and — after the normalise phase — normalised
    scores and best node)."""
    import torch

    dev = torch.device("cuda", 0)
    sch = pkg.Scheduler(0)
    try:
        preds, cluster, pods = synth.make_workload(500, 260, gpus_per_node=8, n_apps=90, ragged_gpus=True,
                                                   p_unknown_app=0.02)
        sch.upload_predictions(preds)
        sch.upload_cluster(cluster, node_base=7)
        ora = oracle_mod.Oracle(preds, cluster)
        want_raw = ora.score_batch(pods, normalize=False, node_base=7)
        want_nrm = ora.score_batch(pods, normalize=True, node_base=7)
        b = _device_batch(torch, pods, 260, dev)
        best = torch.empty(pods.n_pods, dtype=torch.int32, device=dev)
        stream = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(stream):
            sch.score_raw_dev_noalloc(b["app"], b["slo"], pods.n_pods, b["score"], b["sel"], b["hi"], b["neg"],
                                      stream=stream.cuda_stream)
            stream.synchronize()
            assert np.array_equal(b["score"].cpu().numpy(), want_raw[0])
            assert np.array_equal(b["sel"].cpu().numpy(), want_raw[1])
            sch.normalize_dev(pods.n_pods, b["score"], b["hi"], b["neg"], best, stream=stream.cuda_stream)
            stream.synchronize()
        assert np.array_equal(b["score"].cpu().numpy(), want_nrm[0])
        assert np.array_equal(best.cpu().numpy(), want_nrm[2])
    finally:
        sch.close()


def test_noalloc_entry_point_on_an_empty_node_shard(pkg, synth):
    """A context that holds no node (a rank whose shard is empty) must still leave both per-pod
    extrema at zero, as the regular entry point does: they go into the all-reduce(MAX) over ranks,
    where stale values would corrupt every rank's minimum, maximum and best node."""
    import torch

    dev = torch.device("cuda", 0)
    sch = pkg.Scheduler(0)
    try:
        preds, _, pods = synth.make_workload(64, 4, gpus_per_node=8, n_apps=16)
        sch.upload_predictions(preds)
        sch.upload_cluster(synth.make_cluster(0, 16, 8))
        ref = _device_batch(torch, pods, 1, dev)   # (an empty [P, 0] tensor would be a null pointer)
        got = _device_batch(torch, pods, 1, dev)   # extrema start as garbage (-1)
        sch.score_raw_dev(ref["app"], ref["slo"], pods.n_pods, ref["score"], ref["sel"], ref["hi"], ref["neg"])
        sch.score_raw_dev_noalloc(got["app"], got["slo"], pods.n_pods, got["score"], got["sel"], got["hi"], got["neg"])
        torch.cuda.synchronize()
        assert int(got["hi"].abs().max()) == 0 and int(got["neg"].abs().max()) == 0
        assert torch.equal(ref["hi"], got["hi"]) and torch.equal(ref["neg"], got["neg"])
    finally:
        sch.close()


def test_noalloc_entry_point_refuses_oversize_batches(pkg, synth):
    import torch

    dev = torch.device("cuda", 0)
    sch = pkg.Scheduler(0)
    try:
        preds, cluster, _ = synth.make_workload(8, 4, gpus_per_node=8, n_apps=16)
        P = pkg.NOALLOC_MAX_PODS + 1
        big = synth.make_pods(P, 16)
        b = _device_batch(torch, big, 4, dev)
        with pytest.raises(pkg.GsError) as e:   # nothing uploaded yet
            sch.score_raw_dev_noalloc(b["app"], b["slo"], 8, b["score"], b["sel"], b["hi"], b["neg"])
        assert e.value.code == -3
        sch.upload_predictions(preds)
        sch.upload_cluster(cluster)
        with pytest.raises(pkg.GsError) as e:
            sch.score_raw_dev_noalloc(b["app"], b["slo"], P, b["score"], b["sel"], b["hi"], b["neg"])
        assert e.value.code == -1 and "GS_NOALLOC_MAX_PODS" in str(e.value)
        # the largest allowed batch is served
        sch.score_raw_dev_noalloc(b["app"], b["slo"], P - 1, b["score"], b["sel"], b["hi"], b["neg"])
        torch.cuda.synchronize()
        # more than 8 GPU slots per node is outside what the variant covers
        wide = synth.make_cluster(4, 16, gpus_per_node=12)
        sch.upload_cluster(wide)
        with pytest.raises(pkg.GsError) as e:
            sch.score_raw_dev_noalloc(b["app"], b["slo"], 8, b["score"], b["sel"], b["hi"], b["neg"])
        assert e.value.code == -1
    finally:
        sch.close()
