"""recom.proto on the wire (pkg/recommender/protos/recom.proto:5-17): the hand-written message codec
against the reference's own protoc-generated classes (where /root/reference exists), and the two RPCs
end to end over a real grpc server on 127.0.0.1."""
import os
import sys

import numpy as np
import pytest

REF = "/root/reference/pkg/recommender"


@pytest.fixture(scope="module")
def rg(pkg):
    from k8s_gpu_scheduler_b200 import recom_grpc

    return recom_grpc


def test_codec_round_trip(rg):
    for idx in ("", "onnx-mobilenet-1024-7d9f", "x" * 300, "ünïcode"):
        assert rg.decode_request(rg.encode_request(idx)) == idx
    vals = [0.0, 1.5, -3.25e10, float(np.float32(86.57793519459662))]
    cols = ["1P_A30", "2P_A30", "", "4P_V100"]
    r, c = rg.decode_reply(rg.encode_reply(vals, cols))
    assert np.array_equal(np.float32(r), np.float32(vals)) and c == cols
    assert rg.decode_reply(rg.encode_reply([0.0], [])) == ([0.0], [])       # the unknown-app reply
    assert rg.decode_reply(b"") == ([], [])
    with pytest.raises(ValueError):
        rg.decode_reply(b"\x0a\x10abc")                                      # truncated field


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "recom_pb2.py")), reason="reference tree not present")
def test_codec_matches_the_reference_generated_classes(rg):
    sys.dont_write_bytecode = True
    sys.path.insert(0, REF)
    try:
        import recom_pb2
    finally:
        sys.path.remove(REF)
    for idx in ("", "tensorflow-resnet50-2048-abc", "ünï"):
        assert recom_pb2.Request(index=idx).SerializeToString() == rg.encode_request(idx)
        assert rg.decode_request(recom_pb2.Request(index=idx).SerializeToString()) == idx
    rng = np.random.default_rng(0)
    for n, cols in ((1, []), (6, [f"{p}P_{m}" for m in ("A30", "V100") for p in (1, 2, 4)]), (18, ["a"] * 18)):
        vals = rng.normal(size=n).astype(np.float32).tolist()
        theirs = recom_pb2.Reply(result=vals, columns=cols).SerializeToString()
        assert theirs == rg.encode_reply(vals, cols)
        msg = recom_pb2.Reply()
        msg.ParseFromString(rg.encode_reply(vals, cols))
        assert list(msg.result) == vals and list(msg.columns) == cols
        assert rg.decode_reply(theirs) == (vals, cols)


class _FakeRecommender:
    """Stands in for recom_service.Recommender (which needs a GPU to fit): same serving surface."""

    def __init__(self):
        self.reloads = 0
        self.conf = {"onnx_mobilenet": (np.array([1.5, 2.5, 3.5], np.float32), ["1P_A30", "2P_A30", "4P_A30"])}

    def maybe_reload(self):
        self.reloads += 1
        return []

    def _get(self, index):
        key = index.replace("-", "_")
        for k, v in self.conf.items():
            if k in key:
                return v
        return np.array([0.0], np.float32), []

    def impute_configurations(self, index):
        return self._get(index)

    def impute_interference(self, index):
        return np.array([7.0], np.float32), ["onnx_mobilenet"]


def test_both_rpcs_over_a_real_server(rg):
    rec = _FakeRecommender()
    server, port = rg.make_server(rec, port=0)
    server.start()
    try:
        cl = rg.Client(f"127.0.0.1:{port}")
        assert cl.impute_configurations("onnx-mobilenet-1024-x7") == ([1.5, 2.5, 3.5], ["1P_A30", "2P_A30", "4P_A30"])
        assert cl.impute_configurations("busybox") == ([0.0], [])        # recom_server.py:157-160
        assert cl.impute_interference("onnx-mobilenet_A30") == ([7.0], ["onnx_mobilenet"])
        assert rec.reloads == 3
        cl.close()
    finally:
        server.stop(0)
