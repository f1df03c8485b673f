"""A second, independent restatement of Logic() in pure Python (numpy float32/float64 scalars, every
operation typed and rounded like Go on amd64), run against the C oracle on random small snapshots.
Catches mistakes a single restatement cannot see in itself: C type promotion, FMA contraction,
loop-order slips.  Follows pkg/plugins/gpu_plugin/gpu_plugins.go:455-777 top to bottom."""
import numpy as np
import pytest

f32, f64 = np.float32, np.float64
UNKNOWN = 0xFFFFFFFF


def logic_py(preds, cl, pod_app, cur_slo, node, pod_self=-1):
    A = preds.conf.shape[0]
    score, sel, flags = 0, 31, 0
    if not cl.redis_reachable:                                     # :508-527
        return 0, 31
    model, n_uuids, G = int(cl.node_model[node]), int(cl.node_ngpu[node]), cl.max_gpus
    if model != 0 and n_uuids > 0:
        flags |= 0x80

    def conf_of(app, k):                                           # map lookup "%dP_%s"
        if app == UNKNOWN or app >= A or not (0 <= k <= 31):
            return None
        c = int(preds.col_of[model - 1, k])
        return None if c < 0 else f32(preds.conf[app, c])

    def interf_of(row, col):
        if row == UNKNOWN or row >= A or col == UNKNOWN or col >= A:
            return None
        return f32(preds.interf[model - 1, row, col])

    def term(slo, conf, interf, acc):                              # :616-622 / :717-723
        with np.errstate(all="ignore"):
            d = f32(conf - interf)
            t = f32(f32(f32(1) / slo) * f32(slo - d))
            x = abs(f64(t))
            if slo > d:
                acc["cn"] += 1
                y = f64(x + f64(1))
                acc["neg"] = f64(acc["neg"] + f64(1) / f64(f64(1) + f64(y * y)))
            else:
                acc["cp"] += 1
                acc["pos"] = f64(acc["pos"] + f64(1) / f64(f64(1) + x))

    def combine(acc):                                              # :675-688 / :731-743
        cn, cp = acc["cn"], acc["cp"]
        with np.errstate(all="ignore"):
            if cn > 0 and cp > 0:
                k = f64(cn) / f64(cn + cp)
                return f64(f64(100) * f64(f64(f64(f64(1) - k) * acc["pos"]) / f64(cp))) + \
                    f64(f64(100) * f64(f64(k * acc["neg"]) / f64(cn)))
            if cn > 0:
                return f64(100) * f64(acc["neg"] / f64(cn))
            if cp > 0:
                return f64(100) * f64(acc["pos"] / f64(cp))
        return f64(0)

    if model != 0:                                                 # :558
        for g in range(n_uuids):                                   # :562
            lo, hi = int(cl.gpu_off[node * G + g]), int(cl.gpu_off[node * G + g + 1])
            acc = {"neg": f64(0), "pos": f64(0), "cn": 0, "cp": 0}
            for r in range(lo, hi):                                # :571
                slo = f32(cl.res_slo[r])
                if slo == 0:
                    continue
                conf = conf_of(int(cl.res_app[r]), n_uuids)
                if conf is None:
                    continue
                interf = f32(0)
                for c in range(lo, hi):                            # :595
                    if c == r or c == pod_self:
                        continue
                    v = interf_of(int(cl.res_app[r]), int(cl.res_app[c]))
                    if v is not None:
                        interf = f32(interf + v)
                v = interf_of(int(cl.res_app[r]), pod_app)          # :607-612
                if v is not None:
                    interf = f32(interf + v)
                term(slo, conf, interf, acc)
            conf = f32(-1)                                         # :634
            if model == 1:
                v = conf_of(pod_app, n_uuids)
                conf = f32(0) if v is None else v
                if v is not None:
                    flags |= 0x20
            else:
                for part in (1, 2, 4):                             # :645-651
                    v = conf_of(pod_app, part)
                    if v is not None:
                        flags |= 0x20
                        if v > cur_slo and (conf == f32(-1) or v < conf):
                            conf = v
            if conf != f32(-1):
                flags |= 0x40
                interf = f32(0)
                for c in range(lo, hi):                            # :703
                    if c == pod_self:
                        continue
                    v = interf_of(pod_app, int(cl.res_app[c]))
                    if v is not None:
                        interf = f32(interf + v)
                term(f32(cur_slo), conf, interf, acc)
            tmp = combine(acc)
            si = 0 if np.isnan(tmp) else int(tmp)                  # int64(NaN) never beats score
            if si > score:
                score, sel = si, g
    return score, sel | flags


@pytest.mark.parametrize("seed,kw", [(1, {}), (2, dict(with_8p=False)), (3, dict(p_unknown_app=0.15, ragged_gpus=True)),
                                     (4, dict(gpus_per_node=3, n_apps=7))])
def test_c_oracle_equals_python_restatement(pkg, oracle_mod, seed, kw):
    args = dict(n_pods=40, n_nodes=30, gpus_per_node=kw.pop("gpus_per_node", 8), n_apps=kw.pop("n_apps", 24), seed=seed)
    preds, cl, pods = pkg.synth.make_workload(**args, **kw)
    slo = pods.slo.copy()
    slo[:6] = [0.0, np.nan, np.inf, 1e-39, -20.0, 3e38]
    ora = oracle_mod.Oracle(preds, cl)
    rng = np.random.default_rng(seed)
    for p in range(pods.n_pods):
        for n in rng.choice(cl.n_nodes, size=8, replace=False):
            self_slot = -1
            if p % 5 == 0 and cl.res_app.size:
                self_slot = int(rng.integers(0, cl.res_app.size))
            want = logic_py(preds, cl, int(pods.app[p]), f32(slo[p]), int(n), self_slot)
            got = ora.logic(int(pods.app[p]), float(slo[p]) if not np.isnan(slo[p]) else float("nan"), int(n), self_slot)
            assert got == want, (p, int(n), got, want)
