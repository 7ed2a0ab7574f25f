"""The gRPC front door (`b200stage.rpc`; conventions of the reference's Code/gRPC/server.py:13-19 and
client.py:7-11).  CPU: wire format against the protobuf runtime and against the baseline's independent codec,
servicer logic over a real localhost channel with a stand-in stage.  GPU: a real Stage / Pipeline behind the
service returns exactly what the direct calls return."""
import types

import numpy as np
import pytest
import torch

from b200stage import rpc


def _pb_classes():
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    F = descriptor_pb2.FieldDescriptorProto
    fd = descriptor_pb2.FileDescriptorProto(name="b2_test_rpc.proto", syntax="proto3")

    def msg(name, fields):
        m = fd.message_type.add(name=name)
        for fname, num, typ, rep in fields:
            m.field.add(name=fname, number=num, type=typ, label=F.LABEL_REPEATED if rep else F.LABEL_OPTIONAL)
    msg("TensorMsg", [("data", 1, F.TYPE_BYTES, 0), ("shape", 2, F.TYPE_INT64, 1), ("dtype", 3, F.TYPE_STRING, 0),
                      ("request_id", 4, F.TYPE_INT64, 0)])
    msg("GenerateRequest", [("input_ids", 1, F.TYPE_INT64, 1), ("max_new_tokens", 2, F.TYPE_INT64, 0),
                            ("do_sample", 3, F.TYPE_BOOL, 0), ("temperature", 4, F.TYPE_DOUBLE, 0),
                            ("top_k", 5, F.TYPE_INT64, 0), ("top_p", 6, F.TYPE_DOUBLE, 0),
                            ("repetition_penalty", 7, F.TYPE_DOUBLE, 0), ("seed", 8, F.TYPE_INT64, 0),
                            ("prompt", 9, F.TYPE_STRING, 0)])
    msg("GenerateReply", [("output_ids", 1, F.TYPE_INT64, 1), ("seconds", 2, F.TYPE_DOUBLE, 0),
                          ("tokens_per_second", 3, F.TYPE_DOUBLE, 0), ("text", 4, F.TYPE_STRING, 0)])
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    return {n: message_factory.GetMessageClass(pool.FindMessageTypeByName(n)) for n in ("TensorMsg", "GenerateRequest", "GenerateReply")}


def test_tensor_wire_format_against_protobuf_and_the_baseline_codec():
    from oracle import grpc_stages as gs
    T = _pb_classes()["TensorMsg"]
    for arr, rid in [(np.arange(12, dtype=np.float16).reshape(3, 4), 9), (np.array([3, 1 << 45], dtype=np.int64), -2),
                     (np.zeros((0,), np.int64), 0), (np.ones((1, 5), np.float32), 1 << 40)]:
        raw = rpc.encode_tensor(arr, rid)
        assert raw == T(data=arr.tobytes(), shape=list(arr.shape), dtype=arr.dtype.name, request_id=rid).SerializeToString()
        assert raw == gs.encode_msg(arr, rid)                       # two independent writers, same bytes
        for dec in (rpc.decode_tensor, gs.decode_msg):
            back, r2 = dec(raw)
            assert r2 == rid and back.dtype == arr.dtype and np.array_equal(back, arr)
    with pytest.raises(ValueError):
        rpc.decode_tensor(T(data=b"\0" * 6, shape=[2, 2], dtype="float16").SerializeToString())   # 6 bytes for 4 halves
    with pytest.raises(ValueError):
        rpc.decode_tensor(rpc.encode_tensor(np.ones(4, np.float32), 1)[:-3])                        # truncated


def test_generate_messages_against_protobuf():
    C = _pb_classes()
    kw = dict(input_ids=[5, 0, 70000, 1], max_new_tokens=7, do_sample=True, temperature=0.7, top_k=50, top_p=0.9,
              repetition_penalty=1.1, seed=3, prompt="héllo")
    raw = rpc.encode_generate_request(**kw)
    assert raw == C["GenerateRequest"](**kw).SerializeToString()
    assert rpc.decode_generate_request(raw) == kw
    d = rpc.decode_generate_request(C["GenerateRequest"](input_ids=[1, 2]).SerializeToString())
    assert d["input_ids"] == [1, 2] and d["max_new_tokens"] == 0 and d["do_sample"] is False and d["temperature"] == 0.0
    rep = dict(output_ids=[1, 2, 3], seconds=0.25, tokens_per_second=12.0, text="ok")
    raw = rpc.encode_generate_reply(**rep)
    assert raw == C["GenerateReply"](**rep).SerializeToString() and rpc.decode_generate_reply(raw) == rep


class _FakeCache:
    def __init__(self):
        self.lens = [0]

    def reset(self):
        self.lens = [0]


class _FakeStage:
    """Stand-in with the Stage interface the servicer uses; output encodes (input, positions) so the test can
    see that positions are counted per request on the server."""

    def __init__(self, first, last, d=8, vocab=16):
        self.first, self.last, self.device = first, last, "cpu"
        self.cfg = types.SimpleNamespace(d_model=d, vocab=vocab)
        self.made = 0

    def new_kv_cache(self, rows, max_seq):
        self.made += 1
        return _FakeCache()

    def forward(self, x, cache, positions, logits="last"):
        pos = torch.as_tensor(np.asarray(positions), dtype=torch.float32)
        h = (x.float().view(-1, 1) if self.first else x.float()[:, :1]) + 100.0 * pos.view(-1, 1)
        cache.lens[0] = int(positions[-1]) + 1
        if self.last:
            return (h[-1:] + torch.arange(self.cfg.vocab)).float()
        return h.expand(-1, self.cfg.d_model).half().contiguous()


def test_stage_service_over_localhost_with_a_stand_in_stage():
    import grpc
    first = rpc.StageServicer(_FakeStage(True, False), max_requests=2)
    last = rpc.StageServicer(_FakeStage(False, True))
    s0, p0 = rpc.serve(first, port=0)
    s1, p1 = rpc.serve(last, port=0)
    try:
        c0, c1 = rpc.Client(f"127.0.0.1:{p0}"), rpc.Client(f"127.0.0.1:{p1}")
        h = c0.forward(np.array([1, 2, 3], dtype=np.int64), 5)
        assert h.dtype == np.float16 and h.shape == (3, 8) and h[:, 0].tolist() == [1.0, 102.0, 203.0]
        h2 = c0.forward(np.array([4], dtype=np.int64), 5)                 # continues request 5 at position 3
        assert h2[0, 0] == 304.0
        assert c0.forward(np.array([7], dtype=np.int64), 6)[0, 0] == 7.0  # another request starts at 0
        lg = c1.forward(h, 5)
        assert lg.dtype == np.float32 and lg.shape == (1, 16) and lg[0, 0] == 203.0 + 200.0
        with pytest.raises(grpc.RpcError) as e:                           # a third open request: refused
            c0.forward(np.array([1], dtype=np.int64), 7)
        assert e.value.code() == grpc.StatusCode.RESOURCE_EXHAUSTED
        with pytest.raises(grpc.RpcError) as e:                           # hidden states sent to the first stage
            c0.forward(np.zeros((1, 8), np.float16), 5)
        assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
        c0.release(6)
        assert c0.forward(np.array([9], dtype=np.int64), 7)[0, 0] == 9.0  # the released cache is reused, from 0
        assert first.stage.made == 2
        c0.close(), c1.close()
    finally:
        s0.stop(0), s1.stop(0)


class _FakePipe:
    """generate() appends max_new_tokens copies of (last id + 1); enough to see every field arrive."""
    cfg = types.SimpleNamespace(name="fake-model")
    stages = [types.SimpleNamespace(device="cpu")]

    def generate(self, ids, max_new_tokens=32, **kw):
        self.kw = kw
        return torch.cat([ids, torch.full((max_new_tokens,), int(ids[-1]) + 1)])


def test_generate_service_and_rest_twin_with_a_stand_in_pipeline(monkeypatch):
    import json
    import socket
    import threading
    import urllib.error
    import urllib.request
    import uvicorn
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    pipe = _FakePipe()
    sv = rpc.GenerateServicer(pipe)
    server, port = rpc.serve(generate_servicer=sv, port=0)
    try:
        c = rpc.Client(f"127.0.0.1:{port}")
        r = c.generate([4, 5, 6], max_new_tokens=3)
        assert r["output_ids"] == [4, 5, 6, 7, 7, 7] and pipe.kw == {} and r["seconds"] > 0
        assert abs(r["tokens_per_second"] - 6 / r["seconds"]) < 1e-6          # prompt + new, as models_quant_updated.py:95-96
        c.generate([1], max_new_tokens=1, do_sample=True, temperature=0.7, top_p=0.9, seed=2)
        assert pipe.kw == dict(do_sample=True, temperature=0.7, top_k=50, top_p=0.9, repetition_penalty=1.0, seed=2)
        import grpc
        with pytest.raises(grpc.RpcError) as e:
            c.generate([], prompt="text needs a tokenizer")
        assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
        c.close()
    finally:
        server.stop(0)
    # REST twin on an ephemeral port
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        hport = s.getsockname()[1]
    us = uvicorn.Server(uvicorn.Config(rpc.rest_app(sv), host="127.0.0.1", port=hport, log_level="error"))
    th = threading.Thread(target=us.run, daemon=True)
    th.start()
    try:
        import time
        for _ in range(200):
            if us.started:
                break
            time.sleep(0.05)
        base = f"http://127.0.0.1:{hport}"
        st = json.load(urllib.request.urlopen(base + "/"))
        assert st["model"] == "fake-model" and len(st["current_time"]) == 19

        def post(body):
            rq = urllib.request.Request(base + "/generate", data=json.dumps(body).encode(), headers={"Content-Type": "application/json"})
            return json.load(urllib.request.urlopen(rq))
        r = post({"input_ids": [9, 10], "max_new_tokens": 2})
        assert r["output_ids"] == [9, 10, 11, 11] and r["tokens_per_second"] > 0
        for bad, code in (({"input_ids": [1], "num_beams": 4}, 422), ({"input_ids": []}, 400)):
            with pytest.raises(urllib.error.HTTPError) as e:
                post(bad)
            assert e.value.code == code
    finally:
        us.should_exit = True
        th.join(timeout=10)


# ---- GPU ---------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_gpu_stages_and_generate_behind_grpc_match_direct_calls():
    import b200stage
    cfg = b200stage.get_config("tiny-gqa-test")
    pipe = b200stage.Pipeline.load(cfg, n_stages=2, seed=7)
    ids = torch.randint(0, cfg.vocab, (21,), generator=torch.Generator().manual_seed(2), dtype=torch.int64)
    ref = pipe.generate(ids, max_new_tokens=9).cpu().tolist()

    servers = []
    try:
        clients = []
        for st in pipe.stages:
            srv, port = rpc.serve(rpc.StageServicer(st, max_seq=64), port=0)
            servers.append(srv)
            clients.append(rpc.Client(f"127.0.0.1:{port}"))
        gsrv, gport = rpc.serve(generate_servicer=rpc.GenerateServicer(pipe), port=0)
        servers.append(gsrv)

        def hop(x, rid):
            h = clients[0].forward(x, rid)
            assert h.dtype == np.float16 and h.shape == (len(x), cfg.d_model)
            return clients[1].forward(h, rid)
        # greedy decode driven from the client, hidden tiles relayed as bytes: same tokens as pipe.generate
        toks, x = [], ids.numpy()
        for _ in range(9):
            lg = hop(x, 42)
            assert lg.dtype == np.float32 and lg.shape == (1, cfg.vocab)
            toks.append(int(lg[0].argmax()))
            x = np.array([toks[-1]], dtype=np.int64)
        assert ids.tolist() + toks == ref
        for c in clients:
            c.release(42)
        # the whole pipeline behind Generate
        gc = rpc.Client(f"127.0.0.1:{gport}")
        r = gc.generate(ids.tolist(), max_new_tokens=9)
        assert r["output_ids"] == ref and r["seconds"] > 0 and abs(r["tokens_per_second"] - len(ref) / r["seconds"]) < 1e-6
        # sampling arguments travel too (the synthetic model is very peaked: a high temperature makes the draw matter)
        s1 = gc.generate(ids.tolist(), max_new_tokens=9, do_sample=True, temperature=25.0, top_k=20, top_p=0.9, seed=5)
        direct = pipe.generate(ids, max_new_tokens=9, do_sample=True, temperature=25.0, top_k=20, top_p=0.9, seed=5)
        assert s1["output_ids"] == direct.cpu().tolist() and s1["output_ids"] != ref
        gc.close()
        for c in clients:
            c.close()
    finally:
        for s in servers:
            s.stop(0)
