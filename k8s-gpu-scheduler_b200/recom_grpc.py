"""The recommender's gRPC wire contract, served from the GPU-completed tables.

Reference: pkg/recommender/protos/recom.proto:5-17 (service `recommender.recommender`, methods
ImputeConfigurations / ImputeInterference, `Request{string index = 1}`,
`Reply{repeated float result = 1; repeated string columns = 2}`) and the servicer
pkg/recommender/recom_server.py:137-212.  The reference generates its message classes with protoc;
there is no protoc (grpcio-tools) in this image and generated code may not be copied, so the two
messages are encoded and decoded here by hand — they are two fields each — and registered with
grpcio's generic handlers.  tests/test_recom_grpc.py checks the bytes against the reference's own
generated classes where the reference tree is present.

What is served is `Recommender` (recom_service.py): rows of the training tables completed by
gs_complete_fit on the GPU, `[0]` without columns for an unknown app (recom_server.py:157-160), hot
reload on file change under a lock.
"""
from __future__ import annotations

import struct
from concurrent import futures

SERVICE = "recommender.recommender"
METHODS = ("ImputeConfigurations", "ImputeInterference")


# ---- protobuf wire format, just enough for the two messages ---------------------------------

def _varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf: bytes, pos: int):
    shift = result = 0
    while True:
        if pos >= len(buf):
            raise ValueError("truncated varint")
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7
        if shift > 63:
            raise ValueError("varint too long")


def _fields(buf: bytes):
    """Yields (field number, wire type, value) — value is bytes for length-delimited fields."""
    pos = 0
    while pos < len(buf):
        key, pos = _read_varint(buf, pos)
        field, wt = key >> 3, key & 7
        if wt == 0:
            val, pos = _read_varint(buf, pos)
        elif wt == 1:
            val, pos = buf[pos:pos + 8], pos + 8
        elif wt == 2:
            n, pos = _read_varint(buf, pos)
            if pos + n > len(buf):
                raise ValueError("truncated field")
            val, pos = buf[pos:pos + n], pos + n
        elif wt == 5:
            val, pos = buf[pos:pos + 4], pos + 4
        else:
            raise ValueError(f"unsupported wire type {wt}")
        yield field, wt, val


def encode_request(index: str) -> bytes:
    raw = index.encode("utf-8")
    return (b"\x0a" + _varint(len(raw)) + raw) if raw else b""     # proto3 omits default values


def decode_request(buf: bytes) -> str:
    index = ""
    for field, wt, val in _fields(buf):
        if field == 1 and wt == 2:
            index = val.decode("utf-8")
    return index


def encode_reply(result, columns) -> bytes:
    out = bytearray()
    vals = [float(x) for x in result]
    if vals:                                                          # packed repeated float, field 1
        raw = struct.pack(f"<{len(vals)}f", *vals)
        out += b"\x0a" + _varint(len(raw)) + raw
    for c in columns:                                                 # repeated string, field 2
        raw = str(c).encode("utf-8")
        out += b"\x12" + _varint(len(raw)) + raw
    return bytes(out)


def decode_reply(buf: bytes):
    result, columns = [], []
    for field, wt, val in _fields(buf):
        if field == 1 and wt == 2:
            result.extend(struct.unpack(f"<{len(val) // 4}f", val))
        elif field == 1 and wt == 5:                                  # unpacked form, accepted by every parser
            result.append(struct.unpack("<f", val)[0])
        elif field == 2 and wt == 2:
            columns.append(val.decode("utf-8"))
    return result, columns


# ---- server and client ------------------------------------------------------------------------

def make_server(recommender, port: int = 0, host: str = "127.0.0.1", max_workers: int = 10, reload_on_request=True):
    """grpc.server serving `recommender` (recom_service.Recommender).  max_workers = 10 as the reference
    (recom_server.py:208).  Returns (server, bound port); call server.start()."""
    import grpc

    def handler(which):
        def call(request_index, context):
            if reload_on_request:
                recommender.maybe_reload()        # the reference swaps in retrained models here (recom_server.py:141-148)
            result, columns = (recommender.impute_configurations if which == "ImputeConfigurations"
                               else recommender.impute_interference)(request_index)
            return result, columns
        return grpc.unary_unary_rpc_method_handler(call, request_deserializer=decode_request,
                                                   response_serializer=lambda rc: encode_reply(rc[0], rc[1]))

    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers))
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(SERVICE, {m: handler(m) for m in METHODS}),))
    bound = server.add_insecure_port(f"{host}:{port}")
    return server, bound


class Client:
    """What pkg/recommender/go_client/pkg/client_call.go:11-37 does, for tests: dial, one call, close."""

    def __init__(self, target: str):
        import grpc

        self._ch = grpc.insecure_channel(target)
        self._calls = {m: self._ch.unary_unary(f"/{SERVICE}/{m}", request_serializer=encode_request,
                                               response_deserializer=decode_reply) for m in METHODS}

    def impute_configurations(self, index: str, timeout: float = 10.0):
        return self._calls["ImputeConfigurations"](index, timeout=timeout)

    def impute_interference(self, index: str, timeout: float = 10.0):
        return self._calls["ImputeInterference"](index, timeout=timeout)

    def close(self):
        self._ch.close()
