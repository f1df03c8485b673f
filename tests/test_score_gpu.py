"""GPU parity: the CUDA scoring path, through the C ABI, against the CPU oracle.

Bit-exact on raw scores, normalised scores, packed selection bytes (selected GPU + F-bits)
and best node.  Both kernels (restructured and direct) are held to the same bar.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


# every way the library can produce the bytes: the lanes-are-pods kernel (default where the batch is
# large enough for the residents-sum table), the one-thread-per-node kernel, without class dedupe,
# and the op-for-op kernel
VARIANTS = (dict(), dict(flags=8), dict(flags=4), dict(flags=4 | 8), dict(direct=True))


def _check(sched, oracle_mod, preds, cluster, pods, node_base=0, variants=VARIANTS):
    sched.upload_predictions(preds)
    sched.upload_cluster(cluster, node_base=node_base)
    ora = oracle_mod.Oracle(preds, cluster)
    for normalize in (False, True):
        want = ora.score_batch(pods, normalize=normalize, node_base=node_base)
        for kw in variants:
            got = sched.score_batch(pods, normalize=normalize, **kw)
            for w, g, name in zip(want, got, ("score", "sel", "best")):
                bad = np.argwhere(w != g)
                assert bad.size == 0, (f"{name} differs at {bad[:5].tolist()} (normalize={normalize}, {kw}): "
                                       f"want {w[tuple(bad[0])]}, got {g[tuple(bad[0])]}")


def test_config1_1k_pods_100_nodes_8_gpus(sched, oracle_mod, synth):
    """BASELINE.json configs[0]: the reference's own CPU-runnable case."""
    preds, cluster, pods = synth.make_workload(1000, 100, gpus_per_node=8)
    _check(sched, oracle_mod, preds, cluster, pods)


def test_no_8p_column_variant(sched, oracle_mod, synth):
    """The reference tables stop at 4P: every resident is skipped (:585-587), A30 candidates get conf 0 (:637)."""
    preds, cluster, pods = synth.make_workload(400, 150, gpus_per_node=8, with_8p=False)
    _check(sched, oracle_mod, preds, cluster, pods)


def test_ragged_gpus_and_unknown_apps(sched, oracle_mod, synth):
    preds, cluster, pods = synth.make_workload(700, 333, gpus_per_node=8, n_apps=97, ragged_gpus=True,
                                               p_unknown_app=0.07)
    _check(sched, oracle_mod, preds, cluster, pods)


@pytest.mark.parametrize("n_pods,n_nodes,gpn", [(1, 1, 1), (3, 17, 2), (65, 129, 4), (257, 5, 8), (130, 1000, 8)])
def test_odd_shapes(sched, oracle_mod, synth, n_pods, n_nodes, gpn):
    preds, cluster, pods = synth.make_workload(n_pods, n_nodes, gpus_per_node=gpn, n_apps=33, seed=1234 + n_pods)
    _check(sched, oracle_mod, preds, cluster, pods, node_base=1000)


@pytest.mark.parametrize("n_pods,n_nodes,gpn,n_apps", [(300, 33, 8, 7), (777, 250, 8, 40), (1024, 97, 4, 20),
                                                       (513, 64, 2, 5), (2000, 1001, 8, 300), (260, 31, 8, 1)])
def test_lanes_are_pods_kernel_shapes(sched, oracle_mod, synth, n_pods, n_nodes, gpn, n_apps):
    """Batches large enough for the residents-sum table (>= 256 pods) take the lanes-are-pods kernel:
    node counts that are no multiple of 8 or 32 (byte-wise row stores, short last slice), 2/4/8 GPU
    slots, ragged GPU counts, unknown apps, tiny and huge app buckets."""
    preds, cluster, pods = synth.make_workload(n_pods, n_nodes, gpus_per_node=gpn, n_apps=n_apps, seed=77 + n_pods,
                                               ragged_gpus=True, p_unknown_app=0.03)
    if gpn < 8:
        preds.col_of[:, gpn] = preds.col_of[:, 8]       # give the tables a column for this GPU count
        preds.col_of[:, 1] = preds.col_of[:, 8]
    _check(sched, oracle_mod, preds, cluster, pods, node_base=5)


def test_class_dedupe(sched, oracle_mod, synth):
    """Pods with equal (app, SLO bits) are scored once and copied: many duplicates, duplicates of the
    special values (0, -0, NaN, Inf, denormal, negative), the unknown app, and the one key that
    collides with the hash table's empty marker (unknown app + all-ones NaN)."""
    preds, cluster, _ = synth.make_workload(10, 180, gpus_per_node=8, n_apps=12)
    rng = np.random.default_rng(9)
    P = 3000
    slo_values = np.array([10.0, 35.5, 120.25, 300.0, 599.0, 0.0, -0.0, np.nan, np.inf, 1e-40, -77.0, 0.01],
                          dtype=np.float32)
    app = rng.integers(0, 14, size=P).astype(np.uint32)
    app[app >= 12] = synth.APP_UNKNOWN
    slo = slo_values[rng.integers(0, slo_values.size, size=P)].copy()
    bits = slo.view(np.uint32)
    bits[::97] = 0xFFFFFFFF                               # a NaN whose bits are all ones
    pods = synth.Pods(app=app, slo=bits.view(np.float32))
    assert synth.distinct_classes(pods) < 200
    _check(sched, oracle_mod, preds, cluster, pods)
    uniq = synth.Pods(app=app, slo=rng.uniform(10, 600, size=P).astype(np.float32))   # no duplicates at all
    _check(sched, oracle_mod, preds, cluster, uniq, variants=(dict(), dict(flags=4)))


def test_more_than_8_gpus_per_node_uses_direct_kernel(sched, oracle_mod, synth):
    preds, cluster, pods = synth.make_workload(200, 60, gpus_per_node=16, n_apps=50, with_8p=False, ragged_gpus=True)
    # give the table columns that exist for some of the ragged GPU counts
    _check(sched, oracle_mod, preds, cluster, pods)


def test_single_app_many_pods_and_all_unique_apps(sched, oracle_mod, synth):
    preds, cluster, pods = synth.make_workload(900, 140, gpus_per_node=8, n_apps=300)
    one = synth.Pods(app=np.full(900, 5, dtype=np.uint32), slo=pods.slo)
    _check(sched, oracle_mod, preds, cluster, one)
    uniq = synth.Pods(app=(np.arange(900) % 300).astype(np.uint32), slo=pods.slo)
    _check(sched, oracle_mod, preds, cluster, uniq)


def test_special_slo_values(sched, oracle_mod, synth):
    """SLO 0 (+Inf reciprocal), NaN, Inf, negative and denormal SLOs take the same IEEE path as Go."""
    preds, cluster, pods = synth.make_workload(64, 120, gpus_per_node=8, n_apps=40)
    slo = pods.slo.copy()
    slo[:10] = [0.0, -0.0, np.nan, np.inf, -np.inf, -50.0, 1e-40, 3.4e38, 1e-30, 599.99]
    _check(sched, oracle_mod, preds, cluster, synth.Pods(app=pods.app, slo=slo))
    res = cluster.res_slo.copy()
    res[::7] = np.nan
    res[1::11] = -25.0
    res[2::13] = np.inf
    cluster.res_slo = res
    _check(sched, oracle_mod, preds, cluster, pods)


def test_redis_unreachable_scores_zero(sched, oracle_mod, synth):
    preds, cluster, pods = synth.make_workload(50, 40, gpus_per_node=8, n_apps=20)
    cluster.redis_reachable = 0
    _check(sched, oracle_mod, preds, cluster, pods)
    score, sel, best = sched.score_batch(pods, normalize=True)
    assert not score.any() and (sel == 31).all() and (best == 0).all()


def test_self_exclusion(sched, oracle_mod, synth):
    """Pods that are themselves residents (gpu_plugins.go:596,:704)."""
    preds, cluster, pods = synth.make_workload(300, 90, gpus_per_node=8, n_apps=30)
    rng = np.random.default_rng(5)
    R = cluster.res_app.size
    self_slot = np.where(rng.random(300) < 0.5, rng.integers(0, R, size=300), -1).astype(np.int32)
    app = pods.app.copy()
    slo = pods.slo.copy()
    m = self_slot >= 0
    app[m] = cluster.res_app[self_slot[m]]
    slo[m] = cluster.res_slo[self_slot[m]]
    _check(sched, oracle_mod, preds, cluster, synth.Pods(app=app, slo=slo, self_slot=self_slot))
    # an all -1 table is the same as none
    _check(sched, oracle_mod, preds, cluster, synth.Pods(app=pods.app, slo=pods.slo, self_slot=np.full(300, -1, np.int32)))


def test_empty_inputs(sched, synth):
    preds, cluster, pods = synth.make_workload(10, 20, gpus_per_node=8, n_apps=8)
    sched.upload_predictions(preds)
    sched.upload_cluster(cluster)
    score, sel, best = sched.score_batch(synth.Pods(app=np.zeros(0, np.uint32), slo=np.zeros(0, np.float32)))
    assert score.shape == (0, 20) and best.shape == (0,)
    empty = synth.make_cluster(0, 8, 8)
    sched.upload_cluster(empty)
    score, sel, best = sched.score_batch(pods)
    assert score.shape == (10, 0) and (best == -1).all()


def test_errors(pkg, synth):
    s = pkg.Scheduler(0)
    preds, cluster, pods = synth.make_workload(10, 20, gpus_per_node=8, n_apps=8)
    with pytest.raises(pkg.GsError) as e:
        s.score_batch(pods)
    assert e.value.code == -3  # GS_ERR_STATE: nothing uploaded
    bad = synth.make_cluster(5, 8, 8)
    bad.node_model = bad.node_model.copy()
    bad.node_model[0] = 9
    with pytest.raises(pkg.GsError) as e:
        s.upload_cluster(bad)
    assert e.value.code == -1
    s.close()


def test_normalize_scores_list(sched, oracle_mod):
    rng = np.random.default_rng(0)
    for lst in ([10, 20, 30], [7, 7, 7], [0, 1, 3], [5], [], rng.integers(-1000, 1000, size=5000).tolist(),
                [-(2 ** 40), 2 ** 40, 17]):
        assert sched.normalize_scores(lst).tolist() == oracle_mod.normalize(lst).tolist()


def test_select_mig(sched, oracle_mod, synth):
    preds, cluster, pods = synth.make_workload(5000, 10, n_apps=200, p_unknown_app=0.05)
    sched.upload_predictions(preds)
    got = sched.select_mig(pods)
    want = oracle_mod.Oracle(preds, cluster).select_mig(pods)
    assert np.array_equal(got, want)
    assert set(np.unique(got)) <= {0, 1, 2}
    # tables without any A30 column: always configuration 0
    preds2 = synth.make_predictions(200)
    preds2.col_of[0, :] = -1
    sched.upload_predictions(preds2)
    assert not sched.select_mig(pods).any()


def test_full_size_properties_fast_vs_direct(sched, oracle_mod, synth):
    """At a size the oracle cannot finish quickly: the two independent CUDA kernels agree byte for
    byte, NormalizeScore's invariants hold, and a sampled set of rows matches the oracle."""
    preds, cluster, pods = synth.make_workload(6000, 2500, gpus_per_node=8)
    sched.upload_predictions(preds)
    sched.upload_cluster(cluster)
    raw_f = sched.score_batch(pods, normalize=False)
    raw_d = sched.score_batch(pods, normalize=False, direct=True)
    for a, b in zip(raw_f, raw_d):
        assert np.array_equal(a, b)
    nrm, sel, best = sched.score_batch(pods, normalize=True)
    raw = raw_f[0].astype(np.int64)
    lo, hi = raw.min(axis=1, keepdims=True), raw.max(axis=1, keepdims=True)
    rng_ = hi - lo
    want = np.where(rng_ == 0, 0, (raw - lo) * 100 // np.where(rng_ == 0, 1, rng_))
    assert np.array_equal(nrm, want.astype(np.uint8))
    assert np.array_equal(best, raw.argmax(axis=1).astype(np.int32))          # lowest index among maxima
    assert np.array_equal((sel & 31) == 31, raw_f[0] == 0)                     # a GPU is selected iff score >= 1
    rows = np.random.default_rng(1).choice(6000, size=40, replace=False)
    sub = synth.Pods(app=pods.app[rows], slo=pods.slo[rows])
    w = oracle_mod.Oracle(preds, cluster).score_batch(sub, normalize=True)
    assert np.array_equal(w[0], nrm[rows]) and np.array_equal(w[1], sel[rows]) and np.array_equal(w[2], best[rows])


def test_streamlined_arithmetic_matches_ieee_intrinsics(sched):
    """rcp_ge1 / div_small_int / floor_nonneg (score_math.cuh) against __drcp_rn / __ddiv_rn /
    __double2int_rz on 3 x 2*10^8 pseudo-random operands: not one differing bit."""
    assert sched.selftest(seed=0xB2005EED, n_million=200) == 0
    assert sched.selftest(seed=12345, n_million=50) == 0


def test_float32_filter_against_float64_on_the_device(sched):
    """The filter's error bound, measured with the hardware's own rcp.approx: over more than 10^9
    drawn samples — half of them aimed within 1e-4 of an integer score — the float32 estimate never
    strays from the float64 tmpScore by as much as kFilterDelta, and no estimate outside the band
    floors differently.  The measured worst case must leave real margin (below 70 % of delta)."""
    tot = {"evaluated": 0, "in_band": 0, "wrong_floors": 0, "max_dev": 0.0}
    for seed in (0xB2005EED, 7):
        r = sched.selftest_filter(seed=seed, n_million=700)
        tot["evaluated"] += r["evaluated"]
        tot["in_band"] += r["in_band"]
        tot["wrong_floors"] += r["wrong_floors"]
        tot["max_dev"] = max(tot["max_dev"], r["max_dev"])
        delta = r["delta"]
    print(f"filter selftest: {tot}, delta {delta:.3e}, max_dev/delta {tot['max_dev'] / delta:.3f}")
    assert tot["evaluated"] >= 1_000_000_000
    assert tot["wrong_floors"] == 0
    assert tot["in_band"] > 1000               # the near-integer stratum really lands inside the band
    assert tot["max_dev"] < 0.7 * delta


def test_overflowing_terms_take_the_general_path(sched, oracle_mod, synth):
    """Tiny SLOs make the float32 term overflow to Inf inside otherwise ordinary pods/residents."""
    preds, cluster, pods = synth.make_workload(200, 150, gpus_per_node=8, n_apps=20)
    slo = pods.slo.copy()
    slo[::3] = 1e-37
    slo[1::5] = 3e-39   # denormal
    slo[2::7] = -1e-36
    res = cluster.res_slo.copy()
    res[::5] = 1e-37
    cluster.res_slo = res
    preds.conf[3, :] = np.inf
    preds.conf[4, :] = np.nan
    preds.interf[0, 5, :] = 3e38
    preds.interf[1, :, 6] = -3e38
    _check(sched, oracle_mod, preds, cluster, synth.Pods(app=pods.app, slo=slo))


def test_many_residents_per_gpu(sched, oracle_mod, synth):
    """Counts beyond the small-integer table (>= 63 residents on a GPU) and mid-size counts."""
    preds, cluster, pods = synth.make_workload(150, 40, gpus_per_node=2, n_apps=25)
    rng = np.random.default_rng(3)
    counts = rng.integers(0, 12, size=40 * 2)
    counts[5] = 70
    counts[17] = 64
    counts[18] = 62
    off = np.zeros(81, dtype=np.uint32)
    off[1:] = np.cumsum(counts)
    R = int(off[-1])
    cluster.gpu_off = off
    cluster.node_ngpu = np.full(40, 2, dtype=np.uint8)
    cluster.res_app = rng.integers(0, 25, size=R).astype(np.uint32)
    cluster.res_slo = np.round(rng.uniform(10, 600, size=R), 2).astype(np.float32)
    preds.col_of[:, 2] = preds.col_of[:, 8]  # give the tables a 2P column
    _check(sched, oracle_mod, preds, cluster, pods)


def test_without_residents_sum_table(pkg, oracle_mod, synth, monkeypatch):
    """GS_STAGE_TABLE_MB=0 disables the per-batch residents-sum table: the scoring kernel then walks
    the resident lists itself.  Same bytes."""
    monkeypatch.setenv("GS_STAGE_TABLE_MB", "0")
    s = pkg.Scheduler(0)
    try:
        for kw in (dict(n_pods=1000, n_nodes=100, gpus_per_node=8),
                   dict(n_pods=300, n_nodes=333, gpus_per_node=8, n_apps=97, ragged_gpus=True, p_unknown_app=0.07),
                   dict(n_pods=200, n_nodes=150, gpus_per_node=4, n_apps=31, with_8p=False)):
            preds, cluster, pods = synth.make_workload(**kw)
            _check(s, oracle_mod, preds, cluster, pods)
    finally:
        s.close()


def test_sliced_host_path_large_batch(sched, oracle_mod, synth):
    """gs_score_batch slices large batches (>= 8192 pods, >= 64 MB of results) and overlaps the
    device->host copies with the scoring of the next slice: same bytes as the one-shot op-for-op
    kernel, and sampled rows match the oracle."""
    preds, cluster, pods = synth.make_workload(9000, 4000, gpus_per_node=8, n_apps=300, p_unknown_app=0.01)
    sched.upload_predictions(preds)
    sched.upload_cluster(cluster)
    for normalize in (False, True):
        fast = sched.score_batch(pods, normalize=normalize)
        slow = sched.score_batch(pods, normalize=normalize, direct=True)
        for a, b, name in zip(fast, slow, ("score", "sel", "best")):
            assert np.array_equal(a, b), name
    rows = np.random.default_rng(3).choice(9000, size=24, replace=False)
    w = oracle_mod.Oracle(preds, cluster).score_batch(synth.Pods(app=pods.app[rows], slo=pods.slo[rows]), normalize=True)
    assert np.array_equal(w[0], fast[0][rows]) and np.array_equal(w[1], fast[1][rows]) and np.array_equal(w[2], fast[2][rows])


def test_baseline_config2_full_size_against_the_oracle(sched, oracle_mod, synth):
    """BASELINE.json configs[1] at full size (50 000 pods x 5 000 nodes x 8 GPUs): EVERY one of the
    2.5*10^8 (pod, node) pairs against the CPU oracle — raw score, normalised score, selection byte
    (selected GPU + F-bits) and every pod's best node — in pod chunks to bound memory (the oracle
    does about 4*10^7 evaluations per second on 16 cores).  On top of that the op-for-op kernel
    must produce the same bytes, and the F-bits must be consistent with the node tables."""
    preds, cluster, pods = synth.make_workload(50_000, 5_000, gpus_per_node=8)
    sched.upload_predictions(preds)
    sched.upload_cluster(cluster)
    raw, sel, best = sched.score_batch(pods, normalize=False)
    nrm, sel_n, best_n = sched.score_batch(pods, normalize=True)
    assert np.array_equal(sel, sel_n) and np.array_equal(best, best_n)
    ora = oracle_mod.Oracle(preds, cluster)
    for rows in np.array_split(np.arange(50_000), 20):
        sub = synth.Pods(app=pods.app[rows], slo=pods.slo[rows])
        w_raw, w_sel, w_best = ora.score_batch(sub, normalize=False)
        w_nrm, w_sel_n, w_best_n = ora.score_batch(sub, normalize=True)
        for got, want, name in ((raw[rows], w_raw, "raw score"), (sel[rows], w_sel, "selection byte"),
                                (best[rows], w_best, "best node"), (nrm[rows], w_nrm, "normalised score"),
                                (sel[rows], w_sel_n, "selection byte (normalised call)"),
                                (best[rows], w_best_n, "best node (normalised call)")):
            bad = np.argwhere(got != want)
            assert bad.size == 0, f"{name}: {bad.shape[0]} mismatches in pods {rows[0]}..{rows[-1]}, first at {bad[0].tolist()}"
    raw_d, sel_d, best_d = sched.score_batch(pods, normalize=False, direct=True)
    assert np.array_equal(raw, raw_d) and np.array_equal(sel, sel_d) and np.array_equal(best, best_d)
    del raw_d, sel_d
    eligible = (cluster.node_model != 0) & (cluster.node_ngpu > 0)
    assert np.array_equal((sel >> 7).astype(bool), np.broadcast_to(eligible, sel.shape))
    assert not raw[:, ~eligible].any()
    assert raw.max() <= 100
    for rows in np.array_split(np.arange(50_000), 10):
        assert np.array_equal((sel[rows] & 31) == 31, raw[rows] == 0)      # a GPU is selected iff the score is >= 1


def test_exact_integer_boundaries(sched, oracle_mod, synth):
    """Scores that land EXACTLY on integers: SLO == conf - interf gives t = 0, so the term is 1 (pos:
    tmp = 100, 50, 33.33.., 25 ...) — and near misses one float32 ulp away.  Truncation must agree bit
    for bit wherever the float64 value sits relative to the integer."""
    rng = np.random.default_rng(17)
    A = 48
    preds = synth.make_predictions(A)
    # round the tables to few bits so that many exact cancellations occur
    preds.conf = np.round(preds.conf / 8).astype(np.float32) * 8
    preds.interf = np.round(preds.interf / 4).astype(np.float32) * 4
    cluster = synth.make_cluster(200, A, 8, seed=23)
    cluster.res_slo = (np.round(cluster.res_slo / 8) * 8).astype(np.float32)
    app = rng.integers(0, A, size=1500).astype(np.uint32)
    # candidate SLOs drawn from the set {conf - k*4}: hits t == 0 on empty GPUs and small multiples elsewhere
    col = rng.integers(0, preds.conf.shape[1], size=1500)
    slo = preds.conf[app, col] - 4.0 * rng.integers(-3, 4, size=1500)
    slo = np.where(rng.random(1500) < 0.3, np.nextafter(slo.astype(np.float32), np.float32(np.inf)), slo)
    slo = np.where(rng.random(1500) < 0.3, np.nextafter(slo.astype(np.float32), np.float32(-np.inf)), slo)
    pods = synth.Pods(app=app, slo=np.maximum(slo, 1.0).astype(np.float32))
    _check(sched, oracle_mod, preds, cluster, pods)
    raw, _, _ = sched.score_batch(pods, normalize=False)
    assert (raw == 100).any() and (raw == 50).any()      # the exact cases are really present


def test_best_only_and_thread_safety(pkg, sched, oracle_mod, synth):
    """out_score/out_sel may be NULL (decisions only); entry points are safe to call from several
    threads on one context (the framework calls Score from 16 goroutines)."""
    import threading

    preds, cluster, pods = synth.make_workload(400, 120, gpus_per_node=8, n_apps=40)
    sched.upload_predictions(preds)
    sched.upload_cluster(cluster)
    want = oracle_mod.Oracle(preds, cluster).score_batch(pods, normalize=True)
    score, sel, best = sched.score_batch(pods, normalize=True, want_matrices=False)
    assert score is None and sel is None and np.array_equal(best, want[2])
    errors = []

    def worker(seed):
        try:
            for _ in range(5):
                got = sched.score_batch(pods, normalize=True, direct=bool(seed & 1))
                assert all(np.array_equal(a, b) for a, b in zip(want, got))
                assert sched.normalize_scores([3, 9, 6]).tolist() == [0, 100, 50]
        except Exception as ex:  # noqa: BLE001
            errors.append(ex)

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(8)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=120)
    assert not errors, errors


def test_update_node_equals_a_rebuilt_snapshot(sched, oracle_mod, synth):
    """gs_update_node splices one node's GPUs/residents into the uploaded snapshot: after a series of
    updates (grow, shrink, empty, fewer GPUs, first and last node) the scores equal those of the oracle
    on the same cluster rebuilt from scratch."""
    preds, cluster, pods = synth.make_workload(600, 90, gpus_per_node=8, n_apps=40)
    sched.upload_predictions(preds)
    sched.upload_cluster(cluster)
    rng = np.random.default_rng(21)
    G = 8
    lists = [[(cluster.res_app[cluster.gpu_off[n * G + g]:cluster.gpu_off[n * G + g + 1]].copy(),
               cluster.res_slo[cluster.gpu_off[n * G + g]:cluster.gpu_off[n * G + g + 1]].copy()) for g in range(G)]
             for n in range(90)]
    ngpu = cluster.node_ngpu.copy()
    for node, k, fill in [(0, 8, 5), (89, 8, 0), (17, 3, 2), (17, 8, 9), (44, 0, 0), (45, 8, 1), (0, 1, 3), (60, 8, 40)]:
        counts = rng.integers(0, fill + 1, size=k) if k else np.zeros(0, np.int64)
        new = [(rng.integers(0, 40, size=int(c)).astype(np.uint32),
                np.round(rng.uniform(10, 600, size=int(c)), 2).astype(np.float32)) for c in counts]
        lists[node] = new + [(np.zeros(0, np.uint32), np.zeros(0, np.float32))] * (G - k)
        ngpu[node] = k
        sched.update_node(node, counts, np.concatenate([a for a, _ in new] + [np.zeros(0, np.uint32)]),
                          np.concatenate([s for _, s in new] + [np.zeros(0, np.float32)]))
        off = np.zeros(90 * G + 1, dtype=np.uint32)
        off[1:] = np.cumsum([len(a) for n in range(90) for a, _ in lists[n]])
        rebuilt = synth.Cluster(node_model=cluster.node_model, node_ngpu=ngpu.copy(), gpu_off=off,
                                res_app=np.concatenate([a for n in range(90) for a, _ in lists[n]]).astype(np.uint32),
                                res_slo=np.concatenate([s for n in range(90) for _, s in lists[n]]).astype(np.float32),
                                max_gpus=G)
        want = oracle_mod.Oracle(preds, rebuilt).score_batch(pods, normalize=True)
        for kw in (dict(), dict(direct=True)):
            got = sched.score_batch(pods, normalize=True, **kw)
            assert all(np.array_equal(a, b) for a, b in zip(want, got)), f"after updating node {node} ({kw})"
    with pytest.raises(Exception):
        sched.update_node(90, [1], [0], [10.0])
