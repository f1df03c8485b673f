"""Several GPUs through the C ABI alone (gs_create_multi / gs_mscore_batch / gs_mscore_decisions): no
torch, no NCCL in the data plane.  Results must equal the CPU oracle over the whole cluster bit for bit,
whatever the number of devices (one device runs everywhere; two or more when the box has them)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _n_devices(pkg):
    import torch

    return torch.cuda.device_count()


@pytest.mark.parametrize("want", [1, 2, 4, 8])
def test_multi_device_scoring_matches_the_oracle(pkg, synth, oracle_mod, want):
    if _n_devices(pkg) < want:
        pytest.skip(f"needs {want} GPUs")
    preds, cluster, pods = synth.make_workload(1500, 1003, gpus_per_node=8, n_apps=120, p_unknown_app=0.02,
                                               ragged_gpus=True, seed=4242 + want)
    slo = pods.slo.copy()
    slo[::41] = np.nan
    pods = synth.Pods(app=pods.app, slo=slo)
    ora = oracle_mod.Oracle(preds, cluster)
    ms = pkg.MultiScheduler(list(range(want)))
    try:
        ms.upload_predictions(preds)
        ms.upload_cluster(cluster)
        for normalize in (True, False):
            w = ora.score_batch(pods, normalize=normalize)
            g = ms.score_batch(pods, normalize=normalize)
            for a, b, name in zip(w, g, ("score", "sel", "best")):
                assert np.array_equal(a, b), f"{name} differs on {want} device(s), normalize={normalize}"
        raw, sel, best = ora.score_batch(pods, normalize=False)
        dec = ms.score_decisions(pods)
        assert np.array_equal(dec["node"], best)
        assert np.array_equal(dec["sel"], sel[np.arange(pods.n_pods), best])
        assert np.array_equal(dec["raw_hi"], raw.max(axis=1)) and np.array_equal(dec["raw_lo"], raw.min(axis=1))
        # fewer nodes than devices: some shards are empty
        tiny = synth.make_cluster(max(1, want - 1), 120, 8, seed=5)
        ms.upload_cluster(tiny)
        w = oracle_mod.Oracle(preds, tiny).score_batch(pods, normalize=True)
        g = ms.score_batch(pods, normalize=True)
        assert all(np.array_equal(a, b) for a, b in zip(w, g))
    finally:
        ms.close()


def test_multi_errors(pkg, synth):
    with pytest.raises(pkg.GsError):
        pkg.MultiScheduler([0, 0])                       # a device twice
    ms = pkg.MultiScheduler([0])
    try:
        _, _, pods = synth.make_workload(8, 4, n_apps=8)
        with pytest.raises(pkg.GsError) as e:
            ms.score_batch(pods)
        assert e.value.code == -3                        # nothing uploaded
    finally:
        ms.close()
