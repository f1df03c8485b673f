"""The recommender's serving logic over the GPU completion: replies for the reference's own two
training tables equal what the reference's Imputer serves (goldens), unknown apps get [0] without
columns, a changed file is re-fitted, and the completed tables feed the scorer end to end."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _write_tsv(path, index, columns, X):
    with open(path, "w") as f:
        f.write("index\t" + "\t".join(columns) + "\n")
        for name, row in zip(index, X):
            f.write(name + "\t" + "\t".join("" if np.isnan(v) else repr(float(v)) for v in row) + "\n")


def test_find_index_for_request(pkg):
    from k8s_gpu_scheduler_b200 import recom_service as rec

    idx = ["onnx_mobilenet_1024", "onnx_mobilenet_2048", "tensorflow_resnet50_1024"]
    assert rec.find_index_for_request("onnx_mobilenet_2048_abcde", idx) == "onnx_mobilenet_2048"
    assert rec.find_index_for_request("busybox", idx) == ""
    assert rec.find_index_for_request("x_onnx_mobilenet_1024_onnx_mobilenet_2048", idx) == "onnx_mobilenet_1024"  # first in index order


@pytest.mark.gpu
def test_replies_match_reference_imputer_and_reload(pkg, sched, oracle_mod, tmp_path):
    from k8s_gpu_scheduler_b200 import host
    from k8s_gpu_scheduler_b200 import recom_service as rec

    names = np.load(os.path.join(GOLD, "recommender_table_names.npz"))
    gc = np.load(os.path.join(GOLD, "impute_configurations.npz"))
    gi = np.load(os.path.join(GOLD, "impute_interference.npz"))
    pc, pi = str(tmp_path / "configurations_train.ods"), str(tmp_path / "interference_train.ods")
    _write_tsv(pc, names["configurations_index"], names["configurations_columns"], gc["X"])
    _write_tsv(pi, names["interference_index"], names["interference_columns"], gi["X"])
    r = rec.Recommender(pkg, sched)
    r.fit_file("configurations", pc)
    r.fit_file("interference", pi)
    # ImputeConfigurations(pod name) / ImputeInterference("<pod>_<model>")
    res, cols = r.impute_configurations("onnx-resnet50-2048-7c4d")
    k = list(names["configurations_index"]).index("onnx_resnet50_2048")
    assert cols == list(names["configurations_columns"]) and res.dtype == np.float32
    assert np.allclose(res, gc["completed"][k].astype(np.float32), rtol=1e-5)
    res, cols = r.impute_interference("tensorflow-mobilenet-4096_A30")
    k = list(names["interference_index"]).index("tensorflow_mobilenet_4096_A30")
    assert cols == list(names["interference_columns"])
    assert np.allclose(res, gi["completed"][k].astype(np.float32), rtol=1e-5)     # incl. the imputed A30 diagonal entry
    res, cols = r.impute_configurations("busybox")                                  # recom_server.py:157-160
    assert res.tolist() == [0.0] and cols == []
    # the same replies over the wire: recom.proto served from the GPU-completed tables
    from k8s_gpu_scheduler_b200 import recom_grpc

    server, port = recom_grpc.make_server(r, port=0)
    server.start()
    try:
        cl = recom_grpc.Client(f"127.0.0.1:{port}")
        res_w, cols_w = cl.impute_interference("tensorflow-mobilenet-4096_A30")
        assert cols_w == list(names["interference_columns"])
        assert np.array_equal(np.float32(res_w), r.impute_interference("tensorflow-mobilenet-4096_A30")[0])
        assert cl.impute_configurations("busybox") == ([0.0], [])
        cl.close()
    finally:
        server.stop(0)
    # hot reload on md5 change (recom_server.py:84-97)
    assert r.maybe_reload() == []
    X2 = gc["X"].copy()
    X2[0, 0] += 10.0
    _write_tsv(pc, names["configurations_index"], names["configurations_columns"], X2)
    assert r.maybe_reload() == ["configurations"]
    res, _ = r.impute_configurations("onnx-mobilenet-1024")
    assert abs(float(res[0]) - float(np.float32(X2[0, 0]))) < 1e-3
    # the completed tables feed the scorer: strings -> tables -> GPU scores == oracle on the same tables
    b = host.SnapshotBuilder()
    r.fill_builder(b)
    b.add_node("k8s-a30-0", '["MIG-a","MIG-b"]')
    b.add_node("gpu-node-1", '["GPU-c"]')
    b.add_pod("tensorflow-resnet50-1024-x1", "Running", "MIG-b", "120")
    b.add_pod("onnx-ssd-mobilenet-2048-x2", "Running", "GPU-c", "45.5")
    plugin = host.Plugin(sched, b)
    preds, cluster = b.build()
    ora = oracle_mod.Oracle(preds, cluster)
    assert plugin.pre_score("onnx-mobilenet-4096-new", "200")[0] == 0
    for n, node in enumerate(["k8s-a30-0", "gpu-node-1"]):
        got, code, _, _ = plugin.score(node)
        assert code == 0 and got == ora.logic(b.app_of("onnx-mobilenet-4096-new"), 200.0, n)[0]
    plugin.close()
    b.close()
