"""CPU tests of the C1 baseline (BASELINE.json configs[0]): GPT-2 fp32 in two layer-range stages behind
localhost gRPC (`oracle/grpc_stages.py`, conventions of the reference's Code/gRPC/server.py:13-19 and
client.py:7-11).  Checks: the hand-written proto3 wire format against the protobuf runtime, and that the
staged pipeline decodes exactly the tokens HF's own `generate` produces on the unsplit model."""
import numpy as np
import pytest
import torch

from oracle import grpc_stages as gs


def _runtime_message_class():
    """TensorMsg built at run time from a FileDescriptorProto (no protoc in the image)."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    fd = descriptor_pb2.FileDescriptorProto(name="b2_test_stage.proto", syntax="proto3")
    m = fd.message_type.add(name="TensorMsg")
    F = descriptor_pb2.FieldDescriptorProto
    m.field.add(name="data", number=1, type=F.TYPE_BYTES, label=F.LABEL_OPTIONAL)
    m.field.add(name="shape", number=2, type=F.TYPE_INT64, label=F.LABEL_REPEATED)
    m.field.add(name="dtype", number=3, type=F.TYPE_STRING, label=F.LABEL_OPTIONAL)
    m.field.add(name="request_id", number=4, type=F.TYPE_INT64, label=F.LABEL_OPTIONAL)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    return message_factory.GetMessageClass(pool.FindMessageTypeByName("TensorMsg"))


@pytest.mark.parametrize("arr,rid", [
    (np.arange(6, dtype=np.float32).reshape(2, 3), 7),
    (np.array([5, 1 << 40, 3], dtype=np.int64), 1 << 50),
    (np.zeros((0,), np.int64), 0),
    (np.array([[1.5, -2.25]], dtype=np.float16), -3),
    (np.float32(3.0).reshape(()), 300),
])
def test_wire_format_matches_protobuf_runtime(arr, rid):
    Msg = _runtime_message_class()
    raw = gs.encode_msg(arr, rid)
    pb = Msg.FromString(raw)                                     # protobuf parses what we wrote
    assert pb.data == arr.tobytes() and list(pb.shape) == list(arr.shape)
    assert pb.dtype == arr.dtype.name and pb.request_id == rid
    assert raw == Msg(data=arr.tobytes(), shape=list(arr.shape), dtype=arr.dtype.name, request_id=rid).SerializeToString()
    back, rid2 = gs.decode_msg(pb.SerializeToString())           # and we parse what protobuf writes
    assert rid2 == rid and back.dtype == arr.dtype and back.shape == arr.shape and np.array_equal(back, arr)


def test_decode_accepts_unpacked_shape_and_unknown_fields():
    # shape as two unpacked varints (field 2, wire type 0), an unknown varint field 9, an unknown fixed32 field 10
    data = np.arange(4, dtype=np.float32).tobytes()
    raw = b"\x10\x02\x10\x02" + b"\x48\x05" + b"\x55\x00\x00\x00\x00" + b"\x0a\x10" + data + b"\x1a\x07float32"
    arr, rid = gs.decode_msg(raw)
    assert rid == 0 and arr.shape == (2, 2) and np.array_equal(arr.ravel(), np.arange(4, dtype=np.float32))
    with pytest.raises(ValueError):
        gs.decode_msg(b"\x1a\x04int3")


def test_split_layers():
    assert gs.split_layers(12, 2) == [(0, 6), (6, 12)]
    assert gs.split_layers(4, 3) == [(0, 2), (2, 3), (3, 4)]


@pytest.mark.parametrize("n_stages", [2, 3])
def test_staged_grpc_pipeline_decodes_hf_generate_tokens(n_stages):
    torch.set_num_threads(2)
    name, new = "gpt2-tiny", 12
    cfg = gs.gpt2_config(name)
    servers, ports = [], []
    try:
        for lo, hi in gs.split_layers(cfg.n_layer, n_stages):
            srv, port = gs.make_server(gs.StageModel(name, lo, hi, seed=3), 0)      # ephemeral port
            servers.append(srv)
            ports.append(port)
        client = gs.StageClient(ports, ready_timeout=30)
        ids = np.random.default_rng(5).integers(0, cfg.vocab_size, size=(17,), dtype=np.int64)
        toks, _, _ = client.generate(ids, new, request_id=11)
        toks2, _, _ = client.generate(ids, new, request_id=12)              # a second request id: fresh server-side cache
        # two requests interleaved on the same servers keep their caches apart
        a = int(np.argmax(client.forward(ids, 21)[0]))
        b = int(np.argmax(client.forward(ids[:9], 22)[0]))
        a2 = int(np.argmax(client.forward(np.array([a]), 21)[0]))
        client.release(21), client.release(22)
        client.close()
    finally:
        for s in servers:
            s.stop(0)
    full = gs.build_model(name, seed=3)
    t = torch.from_numpy(ids)[None]
    with torch.no_grad():
        ref = full.generate(input_ids=t, attention_mask=torch.ones_like(t), max_new_tokens=new, min_new_tokens=new,
                            do_sample=False, pad_token_id=0)[0, ids.size:].numpy()
        ref_short = int(full(torch.from_numpy(ids[:9])[None]).logits[0, -1].argmax())
    assert np.array_equal(toks, ref) and np.array_equal(toks2, ref)
    assert a == ref[0] and a2 == ref[1] and b == ref_short
