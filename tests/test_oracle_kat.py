"""Known-answer tests that pin the CPU oracle (oracle/score_oracle.c).

The reference has no test for its scoring path, so these answers are derived by hand from
the formulas in pkg/plugins/gpu_plugin/gpu_plugins.go (line numbers in each test) with
explicit float32/float64 scalar arithmetic written out here, independently of the oracle's
code.  They run on the CPU.
"""
import numpy as np
import pytest

f32, f64 = np.float32, np.float64
A30, V100, NONE = 1, 2, 0
UNKNOWN = 0xFFFFFFFF


def tiny(pkg, conf_rows, columns, interf=None, nodes=(), redis=1, max_gpus=4):
    """nodes: list of (model, [ [ (app, slo), ... ] per GPU ])"""
    synth = pkg.synth
    conf = np.array(conf_rows, dtype=np.float32).reshape(len(conf_rows), len(columns))
    A = conf.shape[0]
    if interf is None:
        interf = np.zeros((2, A, A), dtype=np.float32)
    col_of = np.full((2, synth.MAX_GPUS_PER_NODE + 1), -1, dtype=np.int32)
    for c, name in enumerate(columns):
        p, m = name.split("P_")
        col_of[(A30 if m == "A30" else V100) - 1, int(p)] = c
    preds = synth.Predictions(conf=conf, interf=np.asarray(interf, dtype=np.float32), col_of=col_of, columns=list(columns))
    node_model, node_ngpu, off, apps, slos = [], [], [0], [], []
    for model, gpus in nodes:
        node_model.append(model)
        node_ngpu.append(len(gpus))
        for g in range(max_gpus):
            for app, slo in (gpus[g] if g < len(gpus) else []):
                apps.append(app)
                slos.append(slo)
            off.append(len(apps))
    cl = synth.Cluster(node_model=np.array(node_model, dtype=np.uint8), node_ngpu=np.array(node_ngpu, dtype=np.uint8),
                       gpu_off=np.array(off, dtype=np.uint32), res_app=np.array(apps, dtype=np.uint32),
                       res_slo=np.array(slos, dtype=np.float32), max_gpus=max_gpus, redis_reachable=redis)
    return preds, cl


def term(slo, conf, interf):
    """gpu_plugins.go:616-622 / :717-723 with Go's float32/float64 typing; returns (is_negative, value)."""
    slo, conf, interf = f32(slo), f32(conf), f32(interf)
    with np.errstate(all="ignore"):
        d = f32(conf - interf)
        t = f32(f32(f32(1) / slo) * f32(slo - d))
        x = abs(f64(t))
        if slo > d:
            return True, f64(1) / (f64(1) + (x + f64(1)) * (x + f64(1)))
        return False, f64(1) / (f64(1) + x)


def go_int(v):
    return 0 if np.isnan(v) else int(v)  # a NaN never beats the running score (>= 0)


COLS6 = ["1P_A30", "2P_A30", "4P_A30", "1P_V100", "2P_V100", "4P_V100"]


def test_empty_a30_positive_branch(pkg, oracle_mod):
    # one empty GPU, conf 120 >= SLO 100 -> positive branch: 100 * 1/(1+|0.01f*(100-120)|) = 83.33 -> 83
    preds, cl = tiny(pkg, [[120, 0, 0, 0, 0, 0]], COLS6, nodes=[(A30, [[]])])
    s, sel = oracle_mod.Oracle(preds, cl).logic(0, 100.0, 0)
    neg, v = term(100, 120, 0)
    assert not neg and s == go_int(100 * v) == 83
    assert sel & 31 == 0 and sel >> 5 == 0b111  # GPU 0 selected, F6, F7, eligible


def test_empty_a30_negative_branch(pkg, oracle_mod):
    # conf 80 < SLO 100 -> negative branch: 100 * 1/(1+(0.2+1)^2) = 40.98 -> 40
    preds, cl = tiny(pkg, [[80, 0, 0, 0, 0, 0]], COLS6, nodes=[(A30, [[]])])
    s, sel = oracle_mod.Oracle(preds, cl).logic(0, 100.0, 0)
    neg, v = term(100, 80, 0)
    assert neg and s == go_int(100 * v) == 40 and sel & 31 == 0


def test_slo_zero_gives_inf_then_zero(pkg, oracle_mod):
    # currSLO == 0: 1/SLO = +Inf, t = -Inf, 1/(1+Inf) = 0 -> score 0, nothing selected (:719-722)
    preds, cl = tiny(pkg, [[80, 0, 0, 0, 0, 0]], COLS6, nodes=[(A30, [[]])])
    s, sel = oracle_mod.Oracle(preds, cl).logic(0, 0.0, 0)
    assert s == 0 and sel & 31 == 31 and sel >> 5 == 0b111


def test_slo_zero_missing_column_is_nan(pkg, oracle_mod):
    # A30 node listing 3 UUIDs: no "3P_A30" column -> conf = 0 (Go zero value, :637), SLO 0 ->
    # t = Inf*(0-0) = NaN -> tmpScore NaN -> int64(NaN) never > score
    preds, cl = tiny(pkg, [[80, 70, 60, 0, 0, 0]], COLS6, nodes=[(A30, [[], [], []])])
    s, sel = oracle_mod.Oracle(preds, cl).logic(0, 0.0, 0)
    assert s == 0 and sel & 31 == 31
    assert (sel >> 5) & 1 == 0 and (sel >> 6) & 1 == 1  # F6 clear (no column), F7 set (conf=0 != -1)


def test_a30_missing_column_conf_zero(pkg, oracle_mod):
    # same node, SLO 50: conf = 0 -> SLO > 0-0 -> negative: t = (1/50)*(50-0) -> x=1 -> 1/(1+4) = 0.2 -> 20
    preds, cl = tiny(pkg, [[80, 70, 60, 0, 0, 0]], COLS6, nodes=[(A30, [[], [], []])])
    s, sel = oracle_mod.Oracle(preds, cl).logic(0, 50.0, 0)
    neg, v = term(50, 0, 0)
    assert neg and s == go_int(100 * v) == 20 and sel & 31 == 0  # all 3 GPUs tie: first wins (strict >)


def test_v100_picks_smallest_config_above_slo(pkg, oracle_mod):
    # :645-651 — predictions 300/200/120, SLO 150: candidates above SLO are 300 and 200 -> 200
    preds, cl = tiny(pkg, [[0, 0, 0, 300, 200, 120]], COLS6, nodes=[(V100, [[]])])
    s, sel = oracle_mod.Oracle(preds, cl).logic(0, 150.0, 0)
    neg, v = term(150, 200, 0)
    assert not neg and s == go_int(100 * v) == 74 and sel & 31 == 0  # float32(1/150)*50 < 1/3 -> 74.99.. not 75


def test_v100_no_config_above_slo_residents_only(pkg, oracle_mod):
    # :667-696 — nothing above SLO 500 -> conf = -1 -> score from the residents alone.
    # resident app 1 (SLO 100, conf 1P_V100 = 90); interference row(app1,V100) col(app0) = 5:
    # d = 90-5 = 85 < 100 -> negative
    interf = np.zeros((2, 2, 2), dtype=np.float32)
    interf[V100 - 1, 1, 0] = 5
    preds, cl = tiny(pkg, [[0, 0, 0, 300, 200, 120], [0, 0, 0, 90, 80, 70]], COLS6, interf=interf,
                     nodes=[(V100, [[(1, 100.0)]])])
    s, sel = oracle_mod.Oracle(preds, cl).logic(0, 500.0, 0)
    neg, v = term(100, 90, 5)
    assert neg and s == go_int(100 * v) and sel & 31 == 0
    assert (sel >> 6) & 1 == 0  # F7 clear: no candidate term


def test_both_branches_weighted_combine(pkg, oracle_mod):
    # one resident (negative) + candidate (positive): k = 1/2;
    # tmp = 100*((1-k)*pos/1) + 100*(k*neg/1)   (:733-736)
    interf = np.zeros((2, 2, 2), dtype=np.float32)
    interf[A30 - 1, 1, 0] = 7    # resident row, candidate column
    interf[A30 - 1, 0, 1] = 3    # candidate row, resident column
    preds, cl = tiny(pkg, [[150, 0, 0, 0, 0, 0], [60, 0, 0, 0, 0, 0]], COLS6, interf=interf,
                     nodes=[(A30, [[(1, 100.0)]])])
    s, sel = oracle_mod.Oracle(preds, cl).logic(0, 120.0, 0)
    n1, v1 = term(100, 60, 7)
    n2, v2 = term(120, 150, 3)
    assert n1 and not n2
    k = f64(1) / f64(2)
    tmp = f64(100) * ((f64(1) - k) * v2 / f64(1)) + f64(100) * (k * v1 / f64(1))
    assert s == go_int(tmp) and sel & 31 == 0


def test_coresident_interference_order_and_zero_slo_neighbours(pkg, oracle_mod):
    # GPU with residents r0 (app1, SLO 0), r1 (app2, SLO 200), r2 (app3, SLO 90).
    # r0 is skipped (:572) but still interferes with r1, r2 and the candidate (:595, :703).
    rng = np.random.default_rng(7)
    interf = rng.normal(5, 15, size=(2, 4, 4)).astype(np.float32)
    conf = [[300, 0, 0, 0, 0, 0], [100, 0, 0, 0, 0, 0], [250, 0, 0, 0, 0, 0], [80, 0, 0, 0, 0, 0]]
    preds, cl = tiny(pkg, conf, COLS6, interf=interf, nodes=[(A30, [[(1, 0.0), (2, 200.0), (3, 90.0)]])])
    s, sel = oracle_mod.Oracle(preds, cl).logic(0, 210.0, 0)
    I = interf[A30 - 1]
    i1 = f32(f32(f32(0) + I[2, 1]) + I[2, 3]); i1 = f32(i1 + I[2, 0])   # co-residents in order, candidate last
    i2 = f32(f32(f32(0) + I[3, 1]) + I[3, 2]); i2 = f32(i2 + I[3, 0])
    ic = f32(f32(f32(f32(0) + I[0, 1]) + I[0, 2]) + I[0, 3])
    neg_sum = pos_sum = f64(0)
    cn = cp = 0
    for slo, c, i in ((200, 250, i1), (90, 80, i2), (210, 300, ic)):
        n, v = term(slo, c, i)
        if n:
            neg_sum, cn = neg_sum + v, cn + 1
        else:
            pos_sum, cp = pos_sum + v, cp + 1
    if cn and cp:
        k = f64(cn) / f64(cn + cp)
        tmp = f64(100) * ((f64(1) - k) * pos_sum / f64(cp)) + f64(100) * (k * neg_sum / f64(cn))
    elif cn:
        tmp = f64(100) * (neg_sum / f64(cn))
    else:
        tmp = f64(100) * (pos_sum / f64(cp))
    assert s == go_int(tmp)


def test_self_exclusion(pkg, oracle_mod):
    # the pod being scheduled is itself resident r0: excluded as a co-resident (:596) and from
    # its own candidate interference (:704), but its resident term is still evaluated
    interf = np.full((2, 2, 2), 9, dtype=np.float32)
    preds, cl = tiny(pkg, [[150, 0, 0, 0, 0, 0], [60, 0, 0, 0, 0, 0]], COLS6, interf=interf,
                     nodes=[(A30, [[(0, 120.0), (1, 100.0)]])])
    o = oracle_mod.Oracle(preds, cl)
    s_self, _ = o.logic(0, 120.0, 0, pod_self=0)
    # with self-exclusion: r0: interf = I[0,1] + I[0,0](cand) = 18 ; r1: interf = I[1,0](cand) = 9 ; cand: I[0,1] = 9
    terms = [term(120, 150, 18), term(100, 60, 9), term(120, 150, 9)]
    neg = [v for n, v in terms if n]
    pos = [v for n, v in terms if not n]
    cn, cp = len(neg), len(pos)
    k = f64(cn) / f64(cn + cp)
    tmp = f64(100) * ((1 - k) * (pos[0] + pos[1]) / f64(cp)) + f64(100) * (k * neg[0] / f64(cn))
    assert s_self == go_int(tmp)


def test_gates(pkg, oracle_mod):
    conf = [[120, 110, 100, 90, 80, 70]]
    preds, cl = tiny(pkg, conf, COLS6, nodes=[(NONE, [[]]), (A30, []), (A30, [[]])])
    o = oracle_mod.Oracle(preds, cl)
    assert o.logic(0, 100.0, 0) == (0, 31)          # F1: no model -> no flags at all
    assert o.logic(0, 100.0, 1) == (0, 31)          # F3: no UUIDs in Redis
    assert o.logic(0, 100.0, 2)[0] > 0
    assert o.logic(UNKNOWN, 100.0, 2)[0] == 20      # unknown app on A30: empty map -> conf 0 -> like missing column
    preds, cl = tiny(pkg, conf, COLS6, nodes=[(A30, [[]])], redis=0)
    assert oracle_mod.Oracle(preds, cl).logic(0, 100.0, 0) == (0, 31)  # F2: :508-527 returns 0


def test_first_gpu_wins_ties_and_strict_greater(pkg, oracle_mod):
    # two identical empty GPUs: same integer score -> `>` keeps the first (:752)
    preds, cl = tiny(pkg, [[120, 120, 0, 0, 0, 0]], COLS6, nodes=[(A30, [[], []])])
    s, sel = oracle_mod.Oracle(preds, cl).logic(0, 100.0, 0)
    assert sel & 31 == 0 and s == 83


def test_normalize(oracle_mod):
    # gpu_plugins.go:816-841
    assert oracle_mod.normalize([10, 20, 30]).tolist() == [0, 50, 100]
    assert oracle_mod.normalize([7, 7, 7]).tolist() == [0, 0, 0]          # oldRange == 0
    assert oracle_mod.normalize([0, 1, 3]).tolist() == [0, 33, 100]       # truncating division
    assert oracle_mod.normalize([5]).tolist() == [0]
    assert oracle_mod.normalize([]).tolist() == []
    assert oracle_mod.normalize([0, 99, 100, 37]).tolist() == [0, 99, 100, 37]


@pytest.mark.parametrize("slo,conf3", [
    (100.0, (150, 140, 130)),   # all above SLO: 100/(1+(SLO-c)) is negative for c > SLO+1, never >= 0
    (100.0, (90, 80, 70)),
    (100.0, (100.5, 99, 250)),
    (0.0, (10, 20, 30)),
])
def test_select_mig_examples(pkg, oracle_mod, slo, conf3):
    preds, cl = tiny(pkg, [list(conf3) + [0, 0, 0]], COLS6, nodes=[(A30, [[]])])
    got = int(oracle_mod.Oracle(preds, cl).select_mig(pkg.synth.Pods(app=np.array([0], dtype=np.uint32),
                                                                   slo=np.array([slo], dtype=np.float32)))[0])
    # gpu_plugins.go:366-399 restated
    opt, mx = 0, f64(0)
    with np.errstate(all="ignore"):
        for part, c in zip((1, 2, 4), conf3):
            s_, c_ = f32(slo), f32(c)
            if s_ > c_:
                t = f64(f32(f32(s_ - c_) + f32(1)))
                sc = f64(100) / (f64(1) + t * t)
            else:
                sc = f64(100) / (f64(1) + f64(f32(s_ - c_)))
            if sc >= mx:
                mx, opt = sc, part - 1
    if opt == 3:
        opt = 2
    assert got == opt


def test_select_mig_edge_cases(pkg, oracle_mod):
    o_pods = lambda a, s: pkg.synth.Pods(app=np.array([a], dtype=np.uint32), slo=np.array([s], dtype=np.float32))
    # SLO - conf == -1 -> 100/(1-1) = +Inf >= max -> chosen; with 4P -> index 3 -> folded to 2 (:397-399)
    preds, cl = tiny(pkg, [[50, 50, 101, 0, 0, 0]], COLS6, nodes=[(A30, [[]])])
    assert int(oracle_mod.Oracle(preds, cl).select_mig(o_pods(0, 100.0))[0]) == 2
    # unknown app: no predictions -> optimalConfiguration stays 0
    assert int(oracle_mod.Oracle(preds, cl).select_mig(o_pods(UNKNOWN, 100.0))[0]) == 0
    # SLO above everything: 100/(1+(SLO-c+1)^2) largest for the largest c
    preds, cl = tiny(pkg, [[90, 95, 80, 0, 0, 0]], COLS6, nodes=[(A30, [[]])])
    assert int(oracle_mod.Oracle(preds, cl).select_mig(o_pods(0, 100.0))[0]) == 1
