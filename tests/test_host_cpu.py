"""Host-side mirror of the reference plugin (include/gpusched_host.h): the string world of
gpu_plugins.go -> structure-of-arrays tables.  No GPU needed: nothing here computes scores."""
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
# pkg/profiler/test.sh:3 — the profiler's example nvidia-smi list; on a MIG node only MIG UUIDs are kept
MIG_UUIDS = ["MIG-7a700938-2114-5d88-a93c-167c2a498910", "MIG-c8956631-ac65-59d3-9065-c8764799febf",
             "MIG-266f4973-a8c6-5697-a549-dfc880e350c1", "MIG-41e35365-768c-587c-a36e-65197e7a7f93"]


@pytest.fixture()
def host(pkg):
    from k8s_gpu_scheduler_b200 import host as h

    return h


@pytest.fixture()
def tables():
    names = np.load(os.path.join(GOLD, "recommender_table_names.npz"))
    conf = np.load(os.path.join(GOLD, "impute_configurations.npz"))["completed"]
    interf = np.load(os.path.join(GOLD, "impute_interference.npz"))["completed"]
    return names, conf, interf


def json_list(items):
    return "[" + ",".join('"%s"' % x for x in items) + "]"


def test_header_and_binding_agree(host):
    import ctypes
    import re

    text = open(os.path.join(ROOT, "include", "gpusched_host.h")).read()
    declared = sorted(set(re.findall(r"GS_API\s+[\w\s\*]+?\b(gsp_\w+)\s*\(", text)))
    assert declared == sorted(host.HOST_EXPORTS)
    lib = ctypes.CDLL(os.path.join(ROOT, "k8s-gpu-scheduler_b200", "libgpusched.so"))
    for name in declared:
        assert hasattr(lib, name), name
    assert host.Plugin.name() == "GPU"   # gpu_plugins.go:50


def test_parse_slo_follows_logic(host):
    # gpu_plugins.go:459-469: "" -> 0; ParseFloat error -> 0; else float32(value)
    assert host.parse_slo(None) == (0.0, False)
    assert host.parse_slo("") == (0.0, False)
    assert host.parse_slo("10") == (10.0, True)            # deploy/busybox/busybox.yaml:26-28
    assert host.parse_slo("25.37") == (float(np.float32(25.37)), True)
    assert host.parse_slo("abc") == (0.0, False)
    assert host.parse_slo("12abc") == (0.0, False)
    assert host.parse_slo(" 12") == (0.0, False)            # ParseFloat rejects white space
    assert host.parse_slo("1e40") == (0.0, False)           # out of float32 range: ErrRange -> 0
    assert host.parse_slo("0") == (0.0, True)
    v, ok = host.parse_slo("1e-3")
    assert ok and v == float(np.float32(1e-3))


def test_real_recommender_tables_become_soa(host, tables):
    names, conf, interf = tables
    b = host.SnapshotBuilder()
    b.set_configurations(names["configurations_index"], names["configurations_columns"], conf)
    b.set_interference(names["interference_index"], names["interference_columns"], interf)
    # app ids: first training index contained in the pod name with '-' -> '_' (recom_server.py:67-71,155-156)
    assert b.app_of("onnx-mobilenet-1024-7d9f8") == 0
    assert b.app_of("tensorflow-ssd-mobilenet-4096") == 17
    assert b.app_of("prefix-onnx-resnet50-2048-x") == list(names["configurations_index"]).index("onnx_resnet50_2048")
    assert b.app_of("busybox") == 0xFFFFFFFF
    preds, cl = b.build()
    assert preds.conf.shape == (18, 6) and np.array_equal(preds.conf, conf.astype(np.float32))
    cols = list(names["configurations_columns"])
    assert preds.col_of[0, 1] == cols.index("1P_A30") and preds.col_of[0, 4] == cols.index("4P_A30")
    assert preds.col_of[1, 2] == cols.index("2P_V100") and preds.col_of[0, 8] == -1 and preds.col_of[1, 3] == -1
    idx = list(names["interference_index"])
    icol = list(names["interference_columns"])
    r = idx.index("onnx_resnet50_1024_A30")
    c = icol.index("tensorflow_mobilenet_2048")
    a, bb = list(names["configurations_index"]).index("onnx_resnet50_1024"), list(names["configurations_index"]).index("tensorflow_mobilenet_2048")
    assert preds.interf[0, a, bb] == np.float32(interf[r, c])        # model A30 = plane 0
    r = idx.index("onnx_resnet50_1024_V100")
    assert preds.interf[1, a, bb] == np.float32(interf[r, c])
    b.close()


def test_nodes_redis_values_and_residents(host, tables):
    names, conf, interf = tables
    b = host.SnapshotBuilder()
    b.set_configurations(names["configurations_index"], names["configurations_columns"], conf)
    b.set_interference(names["interference_index"], names["interference_columns"], interf)
    n0 = b.add_node("k8s-a30-worker", json_list(MIG_UUIDS))          # "a30" tested first
    n1 = b.add_node("gpu-a30-node", json_list(["GPU-1"]))            # contains both: A30 wins (:478)
    n2 = b.add_node("gpu-node-2", json_list(["GPU-2a", "GPU-2b"]))   # "gpu" -> V100
    n3 = b.add_node("cpu-only", json_list(["GPU-x"]))                # no model
    n4 = b.add_node("gpu-node-nokey", None)                          # redis.Nil: empty list
    n5 = b.add_node("gpu-node-badjson", "not json")                  # Unmarshal error ignored: empty list
    assert (n0, n1, n2, n3, n4, n5) == (0, 1, 2, 3, 4, 5)
    # GetSLOs (gpu_plugins.go:104-157)
    b.add_pod("onnx-mobilenet-1024-aaa", "Running", MIG_UUIDS[1], "120.5")
    b.add_pod("tensorflow-resnet50-2048-bbb", "Running", MIG_UUIDS[1], "0")       # SLO "0": resident with SLO 0
    b.add_pod("onnx-ssd-mobilenet-1024-ccc", "Pending", MIG_UUIDS[1], "50")       # not Running (:105)
    b.add_pod("onnx-ssd-mobilenet-1024-ddd", "Running", None, "50")               # no CUDA_VISIBLE_DEVICES (:131)
    b.add_pod("onnx-ssd-mobilenet-1024-eee", "Running", "GPU-unknown", "50")      # UUID of no node (:135)
    b.add_pod("onnx-ssd-mobilenet-1024-fff", "Running", "GPU-2b", None)           # no SLO env (:137-140)
    b.add_pod("unknown-workload", "Running", "GPU-2b", "33")                      # unknown app still resides
    b.add_pod("onnx-resnet50-4096-ggg", "Running", "GPU-2a", "77", namespace="kube-system")
    preds, cl = b.build()
    assert cl.node_model.tolist() == [1, 1, 2, 0, 2, 2]
    assert cl.node_ngpu.tolist() == [4, 1, 2, 1, 0, 0]
    assert cl.max_gpus == 4 and cl.gpu_off.size == 6 * 4 + 1
    off = cl.gpu_off.reshape(-1)
    # node 0, GPU 1 holds two residents in arrival order
    assert off[1] == 0 and off[2] == 2
    assert cl.res_app[:2].tolist() == [0, list(names["configurations_index"]).index("tensorflow_resnet50_2048")]
    assert cl.res_slo[:2].tolist() == [120.5, 0.0]
    # node 2, GPU 1 (GPU-2b) holds the unknown-app resident only
    s = 2 * 4 + 1
    assert off[s + 1] - off[s] == 1 and cl.res_app[off[s]] == 0xFFFFFFFF and cl.res_slo[off[s]] == 33.0
    assert cl.res_app.size == 3
    b.close()


def test_more_than_31_uuids_is_refused(host):
    b = host.SnapshotBuilder()
    with pytest.raises(ValueError):
        b.add_node("gpu-big", json_list([f"GPU-{i}" for i in range(32)]))
    b.close()


def test_redis_value_parsing_edge_cases(host):
    """json.Unmarshal of the Redis value (gpu_plugins.go:541-542): arrays of strings only; anything
    else leaves the node without UUIDs (the reference ignores the Unmarshal error)."""
    b = host.SnapshotBuilder()
    b.set_configurations(["app_a"], ["1P_V100"], np.ones((1, 1), np.float32))
    cases = [
        ('["GPU-1","GPU-2"]', 2), (' [ "GPU-1" ,\n "GPU-2" ] ', 2), ("[]", 0), ("", 0), ("null", 0), ('{"a":1}', 0),
        ('["GPU-1", 7]', 2),            # UnmarshalTypeError: the element keeps its zero value "", decoding goes on
        ('["GPU-\\u0031","MIG-\\"x\\""]', 2), ('["unterminated', 0),
        ('["a"] trailing', 0),          # invalid JSON: checkValid fails before anything is decoded
        # malformed escapes make the document invalid (and must never be read past the terminator)
        ('["\\u1', 0), ('["\\u12"]', 0), ('["\\uZZZZ"]', 0), ('["\\', 0), ('["\\u', 0), ('["\\x41"]', 0),
        ('["a\tb"]', 0),                # raw control character inside a string
        ('["\\u00e9\\ud83d\\ude00\\/"]', 1),   # UTF-8, a surrogate pair, the escaped solidus
    ]
    for i, (value, _) in enumerate(cases):
        b.add_node(f"gpu-node-{i}", value)
    _, cl = b.build()
    assert cl.node_ngpu.tolist() == [n for _, n in cases]
    b.close()


def test_recommender_table_reader(tmp_path, pkg):
    """The training tables are TSV with an `index` column and empty cells for unprofiled entries
    (recom_server.py:98-102, 218-220)."""
    from k8s_gpu_scheduler_b200 import recom_service as rec

    p = tmp_path / "t.ods"
    p.write_text("index\t1P_A30\t2P_A30\nonnx_a\t1.5\t\nonnx_b\t\t2.25\n")
    index, columns, X = rec.read_table(str(p))
    assert index == ["onnx_a", "onnx_b"] and columns == ["1P_A30", "2P_A30"]
    assert X[0, 0] == 1.5 and np.isnan(X[0, 1]) and np.isnan(X[1, 0]) and X[1, 1] == 2.25


def _prom_body(samples):
    """A Prometheus instant-query response of the shape of pkg/prom/test_data/prom_response_mock.txt."""
    import json

    res = []
    for name, pod, value, uuid, gid in samples:
        m = {"__name__": name, "pod": pod, "container": "exporter", "gpu": "0", "Hostname": pod}
        if uuid is not None:
            m["UUID"] = uuid
        if gid is not None:
            m["GPU_I_ID"] = gid
        res.append({"metric": m, "value": [1674051092.003, value]})
    return json.dumps({"status": "success", "data": {"resultType": "vector", "result": res}})


def test_dcgm_columns_follow_the_prometheus_contract(host):
    """pkg/prom ParseResponse (prom_metrics.go:22-61) + GetDcgmMetricsForUUIDS' keying
    (gpu_plugins.go:193-228): samples by UUID on plain GPUs, by the k-th distinct GPU_I_ID on MIG nodes,
    float32 values, defaults where nothing was reported."""
    b = host.SnapshotBuilder()
    b.set_configurations(["app_a"], ["1P_V100"], np.ones((1, 1), np.float32))
    b.add_node("gpu-node-0", json_list(["GPU-aa", "GPU-bb", "GPU-cc"]))
    b.add_node("k8s-a30-1", json_list(["MIG-x", "MIG-y"]))
    b.add_node("gpu-node-2", json_list(["GPU-dd"]))
    assert b.add_dcgm_response("gpu-node-0", _prom_body([
        ("DCGM_FI_DEV_FB_FREE", "dcgm-0", "4005", "GPU-aa", None), ("DCGM_FI_DEV_FB_FREE", "dcgm-0", "77.5", "GPU-cc", None),
        ("DCGM_FI_DEV_FB_FREE", "dcgm-0", "12", "GPU-zz", None)])) == 3          # an unknown UUID is parsed, fills nothing
    assert b.add_dcgm_response("gpu-node-0", _prom_body([
        ("DCGM_FI_PROF_GR_ENGINE_ACTIVE", "dcgm-0", "0.25", "GPU-bb", None),
        ("DCGM_FI_PROF_GR_ENGINE_ACTIVE", "dcgm-0", "0.5", "GPU-bb", None)])) == 2  # the later sample wins
    assert b.add_dcgm_response("k8s-a30-1", _prom_body([
        ("DCGM_FI_DEV_FB_FREE", "dcgm-1", "100", "GPU-parent", "9"), ("DCGM_FI_DEV_FB_FREE", "dcgm-1", "200", "GPU-parent", "13")])) == 2
    assert b.add_dcgm_response("gpu-node-2", "") == 0                             # empty body: no samples
    assert b.add_dcgm_response("gpu-node-2", '{"status":"success","data":{"resultType":"vector","result":[]}}') == 0
    with pytest.raises(ValueError):
        b.add_dcgm_response("gpu-node-2", "not json")
    with pytest.raises(ValueError):
        b.add_dcgm_response("gpu-node-2", '{"data":{"result":[{"metric":{"pod":"p"},"value":[1,"2"]}]}}')   # no __name__
    with pytest.raises(ValueError):
        b.add_dcgm_response("no-such-node", "")
    _, cl = b.build()
    G = cl.max_gpus
    free = b.dcgm_column("DCGM_FI_DEV_FB_FREE", 1.0, 3, G)
    assert free[0, :3].tolist() == [4005.0, 1.0, 77.5]
    assert free[1, :2].tolist() == [200.0, 100.0]      # GPU_I_ID sorted as STRINGS: "13" < "9" -> MIG-x = 13, MIG-y = 9
    assert free[2, 0] == 1.0 and (free[:, 3:] == 1.0).all() if G > 3 else True
    util = b.dcgm_column("DCGM_FI_PROF_GR_ENGINE_ACTIVE", 0.0, 3, G)
    assert util[0, :3].tolist() == [0.0, 0.5, 0.0] and not util[1:].any()
    b.close()
