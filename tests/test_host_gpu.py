"""The plugin surface end to end (PreScore -> Score per node -> NormalizeScore) over a cluster
described the way the reference sees it — node names, Redis JSON values, pods with ConfigMap /
env, the recommender's named tables — against the CPU oracle evaluated on the tables the host
layer built.  Written the way a _test.go next to gpu_plugins.go would read."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def json_list(items):
    return "[" + ",".join('"%s"' % x for x in items) + "]"


@pytest.fixture()
def world(pkg):
    from k8s_gpu_scheduler_b200 import host

    names = np.load(os.path.join(GOLD, "recommender_table_names.npz"))
    conf = np.load(os.path.join(GOLD, "impute_configurations.npz"))["completed"]
    interf = np.load(os.path.join(GOLD, "impute_interference.npz"))["completed"]
    b = host.SnapshotBuilder()
    b.set_configurations(names["configurations_index"], names["configurations_columns"], conf)
    b.set_interference(names["interference_index"], names["interference_columns"], interf)
    apps = [a.replace("_", "-") for a in names["configurations_index"]]
    rng = np.random.default_rng(11)
    node_names, uuids = [], {}
    for i in range(24):
        kind = ["k8s-a30-%d", "gpu-node-%d", "cpu-node-%d"][i % 3 if i % 7 else 2]
        name = kind % i
        k = [1, 2, 4][i % 3]
        u = [f"{'MIG' if 'a30' in name else 'GPU'}-{i:02d}-{g}" for g in range(k)]
        uuids[name] = u
        b.add_node(name, json_list(u) if i != 5 else None)
        node_names.append(name)
    for p in range(60):
        node = node_names[int(rng.integers(0, 24))]
        if not uuids[node]:
            continue
        b.add_pod(f"{apps[int(rng.integers(0, 18))]}-{p:03d}", "Running" if p % 9 else "Pending",
                  uuids[node][int(rng.integers(0, len(uuids[node])))], ["150", "40.25", "0", "300", "75.5"][p % 5])
    return host, b, node_names, uuids, apps


def test_score_per_node_matches_logic(pkg, sched, oracle_mod, world):
    host, b, node_names, uuids, apps = world
    plugin = host.Plugin(sched, b)
    preds, cluster = b.build()
    ora = oracle_mod.Oracle(preds, cluster)
    for pod_name, slo in [(f"{apps[3]}-new", "100"), (f"{apps[10]}-new", "12.5"), (f"{apps[0]}-new", None),
                          ("busybox", "10"), (f"{apps[7]}-new", "garbage"), (f"{apps[16]}-new", "650")]:
        rc, msg = plugin.pre_score(pod_name, slo)
        assert rc == 0, msg
        app = b.app_of(pod_name)
        cur, _ = host.parse_slo(slo)
        scores = []
        for n, node in enumerate(node_names):
            got, code, status, uuid = plugin.score(node)
            want, sel = ora.logic(app, cur, n)
            assert code == 0 and status == ""
            assert got == want, (pod_name, node)
            g = sel & 31
            assert uuid == ("" if g == 31 else uuids[node][g])      # selectedUUID, gpu_plugins.go:458,752-755
            scores.append(got)
        assert plugin.normalize_score(scores).tolist() == oracle_mod.normalize(scores).tolist()
    plugin.close()
    b.close()


def test_score_error_statuses(pkg, sched, world):
    host, b, node_names, uuids, apps = world
    # a resident whose SLO does not parse: GetSLOs returns the error, Score fails for that node (:141-144,:544-548)
    b.add_pod(f"{apps[2]}-broken", "Running", uuids[node_names[1]][0], "fast")
    plugin = host.Plugin(sched, b)
    score, code, status, _ = plugin.score(node_names[0])
    assert code == 1 and "PreScore" in status                         # no cycle state yet
    assert plugin.pre_score(f"{apps[1]}-x", "90")[0] == 0
    score, code, status, uuid = plugin.score("no-such-node")
    assert code == 1 and score == 0 and 'getting node "no-such-node" from Snapshot' in status   # :798-801
    score, code, status, uuid = plugin.score(node_names[1])
    assert code == 1 and score == 0 and status.startswith("Error in Logic() in Score():")       # :805
    score, code, status, uuid = plugin.score(node_names[0])
    assert code == 0 and status == ""
    plugin.close()
    b.close()


def test_redis_unreachable_returns_zero_everywhere(pkg, sched, world):
    host, b, node_names, uuids, apps = world
    b.set_redis_reachable(False)                                      # FindNodesIPFromPod found no redis pod (:508-527)
    plugin = host.Plugin(sched, b)
    assert plugin.pre_score(f"{apps[4]}-x", "90")[0] == 0
    for node in node_names:
        score, code, status, uuid = plugin.score(node)
        assert (score, code, uuid) == (0, 0, "")
    plugin.close()
    b.close()


def test_postbind_record_and_decisions(pkg, sched, oracle_mod, world):
    """PostBind's data path (gpu_plugins.go:843-926): for the bound node, CUDA_VISIBLE_DEVICES is the
    UUID Logic selected there, the MPS limits follow the V100 configuration Logic recorded; and the
    batched decisions call returns the same facts in 8 bytes per pod."""
    host, b, node_names, uuids, apps = world
    plugin = host.Plugin(sched, b)
    preds, cluster = b.build()
    ora = oracle_mod.Oracle(preds, cluster)
    v100_cols = [int(preds.col_of[1, k]) for k in (1, 2, 4)]
    names, slos = [f"{apps[3]}-new", f"{apps[10]}-new", f"{apps[0]}-new", "busybox", f"{apps[16]}-new"], \
        ["100", "12.5", None, "10", "650"]
    pod_app = np.array([b.app_of(n) for n in names], dtype=np.uint32)
    pod_slo = np.array([host.parse_slo(s)[0] for s in slos], dtype=np.float32)
    dec = sched.score_decisions(pkg.Pods(app=pod_app, slo=pod_slo))
    raw, sel, best = ora.score_batch(pkg.Pods(app=pod_app, slo=pod_slo), normalize=False)
    for i, (pod_name, slo) in enumerate(zip(names, slos)):
        rc, msg = plugin.pre_score(pod_name, slo)
        assert rc == 0, msg
        # expected MPS configuration: "1P_V100" unless a smaller prediction above the SLO exists (:640-651)
        which, cur = 1, -1.0
        if pod_app[i] != pkg.APP_UNKNOWN:
            for k, col in zip((1, 2, 4), v100_cols):
                v = preds.conf[pod_app[i], col] if col >= 0 else None
                if v is not None and v > pod_slo[i] and (cur == -1.0 or v < cur):
                    which, cur = k, v
        for n, node in enumerate(node_names):
            would, uuid, mem, thr = plugin.postbind(node)
            g = int(sel[i, n]) & 31
            want_uuid = uuids[node][g] if g != 31 and node != node_names[5] else ""
            assert uuid == want_uuid and would == (want_uuid != "")
            is_v100 = cluster.node_model[n] == pkg.MODEL_V100 and cluster.node_ngpu[n] > 0
            want_mps = {2: ("0=16350MB", "50"), 4: ("0=8175MB", "25")}.get(which if is_v100 else 0, ("", ""))
            assert (mem, thr) == want_mps
        # the batched decision of the same pod
        assert dec["node"][i] == best[i]
        assert dec["sel"][i] == sel[i, best[i]]
        assert dec["raw_hi"][i] == raw[i].max() and dec["raw_lo"][i] == raw[i].min()
        bn = int(best[i])
        is_v100 = cluster.node_model[bn] == pkg.MODEL_V100 and cluster.node_ngpu[bn] > 0
        assert dec["mps"][i] == (which if is_v100 else 0)
    with pytest.raises(RuntimeError):
        plugin.postbind("no-such-node")
