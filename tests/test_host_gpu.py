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
