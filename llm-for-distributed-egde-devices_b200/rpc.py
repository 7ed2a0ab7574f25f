"""gRPC front door for a GPU stage / pipeline (SURVEY.md §8(b) "RPC contract", §8(f) rank 4).

The reference's only RPC code is a unary proto3 service behind
``grpc.server(futures.ThreadPoolExecutor(max_workers=10))`` on the insecure port ``[::]:50051``
(``Code/gRPC/server.py:13-19``), called through ``grpc.insecure_channel(host:port)`` and a stub
(``Code/gRPC/client.py:7-11``); its REST twin returns one JSON object (``rest_api.py:9-15``).  This module
keeps those conventions and puts them in front of the GPU path, so the Jetson-side client of the reference
can talk to a B200 box:

    syntax = "proto3";
    service Stage {
      rpc Forward  (TensorMsg)       returns (TensorMsg);       // one stage: hidden tile in, hidden tile / logits out
      rpc Generate (GenerateRequest) returns (GenerateReply);   // whole pipeline: the reference's get_response()
    }
    message TensorMsg       { bytes data = 1; repeated int64 shape = 2; string dtype = 3; int64 request_id = 4; }
    message GenerateRequest { repeated int64 input_ids = 1; int64 max_new_tokens = 2; bool do_sample = 3;
                              double temperature = 4; int64 top_k = 5; double top_p = 6;
                              double repetition_penalty = 7; int64 seed = 8; string prompt = 9; }
    message GenerateReply   { repeated int64 output_ids = 1; double seconds = 2; double tokens_per_second = 3;
                              string text = 4; }

``protoc`` / ``grpc_tools`` are not part of this image: messages are (de)serialised here in the proto3 wire
format (any protobuf client built from the schema above interoperates; the tests check that against the
protobuf runtime) and the methods are registered with ``grpc.method_handlers_generic_handler``.

Between GPUs of one box hidden tiles move over NCCL (``pipeline.DistPipeline``), never through this service;
the RPC is the outward-facing door only.  The GPU work of concurrent requests is serialised by a lock (one
stage = one GPU = one stream); the 10-thread pool only overlaps (de)serialisation.

``Generate`` mirrors ``get_response`` (``models_quant_updated.py:75-97``): generate, then
``tokens_per_second = len(output[0]) / elapsed`` (``:72-73, :95-96``: prompt + new tokens).  ``prompt`` /
``text`` are served only when the servicer was given a tokenizer (none is reachable offline).
"""
from __future__ import annotations

import struct
import threading
import time
from concurrent import futures

import numpy as np

SERVICE = "Stage"
DEFAULT_PORT = 50051                       # Code/gRPC/server.py:16
MAX_MSG = 256 << 20
_DTYPES = ("float16", "float32", "int64", "int32", "int8")


# ------------------------------------------------------------------------------------------------
# proto3 wire format
# ------------------------------------------------------------------------------------------------
class _Writer:
    def __init__(self):
        self.b = bytearray()

    def varint(self, v: int):
        v &= 0xFFFFFFFFFFFFFFFF
        while v > 0x7F:
            self.b.append((v & 0x7F) | 0x80)
            v >>= 7
        self.b.append(v)

    def int_field(self, num: int, v: int):
        if v:
            self.varint(num << 3)
            self.varint(int(v))

    def bytes_field(self, num: int, v: bytes):
        if v:
            self.varint((num << 3) | 2)
            self.varint(len(v))
            self.b += v

    def double_field(self, num: int, v: float):
        if v != 0.0:
            self.varint((num << 3) | 1)
            self.b += struct.pack("<d", v)

    def packed_int64(self, num: int, vals):
        if len(vals):
            inner = _Writer()
            for v in vals:
                inner.varint(int(v))
            self.bytes_field(num, bytes(inner.b))


def _fields(buf: bytes):
    """Yield (field number, wire type, value) over a serialised message; value is an int (varint, as
    signed 64-bit), bytes (length-delimited) or the raw 8 / 4 bytes of a fixed field."""
    i, n = 0, len(buf)

    def varint():
        nonlocal i
        v = s = 0
        while True:
            if i >= n:
                raise ValueError("truncated message")
            c = buf[i]
            i += 1
            v |= (c & 0x7F) << s
            if c < 0x80:
                return v
            s += 7
    while i < n:
        key = varint()
        num, wt = key >> 3, key & 7
        if wt == 0:
            v = varint()
            yield num, wt, v - (1 << 64) if v >> 63 else v
        elif wt == 2:
            ln = varint()
            if i + ln > n:
                raise ValueError("truncated message")
            yield num, wt, bytes(buf[i:i + ln])
            i += ln
        elif wt in (1, 5):
            ln = 8 if wt == 1 else 4
            yield num, wt, bytes(buf[i:i + ln])
            i += ln
        else:
            raise ValueError(f"unsupported wire type {wt}")


def _repeated_int64(acc: list, wt: int, v):
    """One occurrence of a `repeated int64` field: a lone varint, or a packed run of them."""
    if wt == 0:
        acc.append(v)
        return
    j = 0
    while j < len(v):
        x = s = 0
        while True:
            c = v[j]
            j += 1
            x |= (c & 0x7F) << s
            if c < 0x80:
                break
            s += 7
        acc.append(x - (1 << 64) if x >> 63 else x)


def encode_tensor(arr: np.ndarray, request_id: int = 0) -> bytes:
    arr = np.require(arr, requirements="C")
    if arr.dtype.name not in _DTYPES:
        raise ValueError(f"unsupported dtype {arr.dtype.name}")
    w = _Writer()
    w.bytes_field(1, arr.tobytes())
    w.packed_int64(2, arr.shape)
    w.bytes_field(3, arr.dtype.name.encode())
    w.int_field(4, request_id)
    return bytes(w.b)


def decode_tensor(buf: bytes):
    """-> (ndarray, request_id)"""
    data, shape, dtype, rid = b"", [], "float32", 0
    for num, wt, v in _fields(buf):
        if num == 1 and wt == 2:
            data = v
        elif num == 2:
            _repeated_int64(shape, wt, v)
        elif num == 3 and wt == 2:
            dtype = v.decode()
        elif num == 4 and wt == 0:
            rid = v
    if dtype not in _DTYPES:
        raise ValueError(f"unsupported dtype {dtype!r}")
    n = int(np.prod(shape, dtype=np.int64)) if shape else 1
    if n * np.dtype(dtype).itemsize != len(data):
        raise ValueError(f"shape {shape} of {dtype} does not match {len(data)} data bytes")
    return np.frombuffer(data, dtype=dtype).reshape(shape), rid


_GEN_DEFAULTS = dict(input_ids=(), max_new_tokens=0, do_sample=False, temperature=0.0, top_k=0, top_p=0.0,
                     repetition_penalty=0.0, seed=0, prompt="")


def encode_generate_request(input_ids=(), max_new_tokens: int = 32, do_sample: bool = False, temperature: float = 0.0,
                            top_k: int = 0, top_p: float = 0.0, repetition_penalty: float = 0.0, seed: int = 0,
                            prompt: str = "") -> bytes:
    """Zero / empty means "not set" (proto3 has no presence for scalars): the server then applies HF's
    defaults (temperature 1, top_k 50, top_p 1, repetition_penalty 1)."""
    w = _Writer()
    w.packed_int64(1, list(input_ids))
    w.int_field(2, max_new_tokens)
    w.int_field(3, 1 if do_sample else 0)
    w.double_field(4, float(temperature))
    w.int_field(5, top_k)
    w.double_field(6, float(top_p))
    w.double_field(7, float(repetition_penalty))
    w.int_field(8, seed)
    w.bytes_field(9, prompt.encode())
    return bytes(w.b)


def decode_generate_request(buf: bytes) -> dict:
    r = dict(_GEN_DEFAULTS)
    ids = []
    names = {2: "max_new_tokens", 5: "top_k", 8: "seed"}
    dbl = {4: "temperature", 6: "top_p", 7: "repetition_penalty"}
    for num, wt, v in _fields(buf):
        if num == 1:
            _repeated_int64(ids, wt, v)
        elif num in names and wt == 0:
            r[names[num]] = v
        elif num == 3 and wt == 0:
            r["do_sample"] = bool(v)
        elif num in dbl and wt == 1:
            r[dbl[num]] = struct.unpack("<d", v)[0]
        elif num == 9 and wt == 2:
            r["prompt"] = v.decode()
    r["input_ids"] = ids
    return r


def encode_generate_reply(output_ids, seconds: float, tokens_per_second: float, text: str = "") -> bytes:
    w = _Writer()
    w.packed_int64(1, list(output_ids))
    w.double_field(2, float(seconds))
    w.double_field(3, float(tokens_per_second))
    w.bytes_field(4, text.encode())
    return bytes(w.b)


def decode_generate_reply(buf: bytes) -> dict:
    r = dict(output_ids=[], seconds=0.0, tokens_per_second=0.0, text="")
    for num, wt, v in _fields(buf):
        if num == 1:
            _repeated_int64(r["output_ids"], wt, v)
        elif num == 2 and wt == 1:
            r["seconds"] = struct.unpack("<d", v)[0]
        elif num == 3 and wt == 1:
            r["tokens_per_second"] = struct.unpack("<d", v)[0]
        elif num == 4 and wt == 2:
            r["text"] = v.decode()
    return r


# ------------------------------------------------------------------------------------------------
# servicers
# ------------------------------------------------------------------------------------------------
class StageServicer:
    """`Forward` in front of one `Stage` (a contiguous layer range on one GPU).

    Request: first stage -> int64 token ids [T]; otherwise fp16 hidden [T, d].  The tokens continue the
    sequence `request_id` (its KV cache lives here, positions are counted here).  Reply: fp16 hidden [T, d],
    or on the last stage fp32 logits [1, V] of the final position.  An empty tensor releases the request."""

    def __init__(self, stage, max_seq: int = 2048, max_requests: int = 8):
        self.stage, self.max_seq, self.max_requests = stage, max_seq, max_requests
        self.caches, self.free = {}, []
        self.lock = threading.Lock()

    def forward(self, request: bytes, context=None) -> bytes:
        import torch
        arr, rid = decode_tensor(request)
        st = self.stage
        with self.lock:
            if arr.size == 0:
                c = self.caches.pop(rid, None)
                if c is not None:
                    c.reset()
                    self.free.append(c)
                return encode_tensor(np.zeros((0,), np.float16), rid)
            want = ("int64", 1) if st.first else ("float16", 2)
            if (arr.dtype.name, arr.ndim) != want or (arr.ndim == 2 and arr.shape[1] != st.cfg.d_model):
                raise ValueError(f"stage expects {want[0]} with {want[1]} dim(s), got {arr.dtype.name} {arr.shape}")
            c = self.caches.get(rid)
            if c is None:
                if len(self.caches) >= self.max_requests:
                    raise RuntimeError("too many open requests; release one with an empty tensor")
                c = self.caches[rid] = self.free.pop() if self.free else st.new_kv_cache(1, self.max_seq)
            T, p0 = int(arr.shape[0]), int(c.lens[0])
            import contextlib
            on_gpu = torch.device(st.device).type == "cuda"      # (tests drive the servicer with a CPU stand-in stage)
            # gRPC pool threads start with device 0 current
            with (torch.cuda.device(st.device) if on_gpu else contextlib.nullcontext()):
                x = torch.from_numpy(arr.copy()).to(st.device)
                y = st.forward(x, c, np.arange(p0, p0 + T), logits="last" if st.last else "none")
                out = y.view(1, -1).float().cpu().numpy() if st.last else y.cpu().numpy()
        return encode_tensor(out, rid)


class GenerateServicer:
    """`Generate` in front of a `Pipeline`: the reference's get_response() (models_quant_updated.py:75-97)."""

    def __init__(self, pipe, tokenizer=None):
        self.pipe, self.tokenizer = pipe, tokenizer
        self.lock = threading.Lock()

    def generate(self, request: bytes, context=None) -> bytes:
        import torch
        r = decode_generate_request(request)
        ids = r["input_ids"]
        if not ids and r["prompt"]:
            if self.tokenizer is None:
                raise ValueError("this server has no tokenizer: send input_ids")
            ids = self.tokenizer(r["prompt"])["input_ids"]
        if not ids:
            raise ValueError("empty request")
        kw = {}
        if r["do_sample"]:
            kw = dict(do_sample=True, temperature=r["temperature"] or 1.0, top_k=r["top_k"] or 50, top_p=r["top_p"] or 1.0,
                      repetition_penalty=r["repetition_penalty"] or 1.0, seed=r["seed"])
        with self.lock:
            torch.cuda.synchronize(self.pipe.stages[0].device)
            t0 = time.time()                                              # start_time = time.time()  (:83)
            out = self.pipe.generate(torch.tensor(ids, dtype=torch.int64), max_new_tokens=r["max_new_tokens"] or 32, **kw)
            out = out.cpu().tolist()
            dt = time.time() - t0
        text = self.tokenizer.decode(out, skip_special_tokens=True) if self.tokenizer is not None else ""
        return encode_generate_reply(out, dt, len(out) / dt if dt > 0 else 0.0, text)   # tokens_per_second (:72-73)


def _wrap(fn):
    """Turn Python exceptions into gRPC status codes instead of UNKNOWN with a stack trace."""
    import grpc

    def call(request, context):
        try:
            return fn(request, context)
        except ValueError as e:
            context.abort(grpc.StatusCode.INVALID_ARGUMENT, str(e))
        except RuntimeError as e:
            context.abort(grpc.StatusCode.RESOURCE_EXHAUSTED if "too many" in str(e) else grpc.StatusCode.INTERNAL, str(e))
    return call


def serve(stage_servicer: StageServicer | None = None, generate_servicer: GenerateServicer | None = None,
          port: int = DEFAULT_PORT, max_workers: int = 10):
    """Start the server the way Code/gRPC/server.py:13-19 does.  Returns (server, bound_port); the caller
    ends with server.wait_for_termination() or server.stop(grace)."""
    import grpc
    methods = {}
    if stage_servicer is not None:
        methods["Forward"] = grpc.unary_unary_rpc_method_handler(_wrap(stage_servicer.forward))
    if generate_servicer is not None:
        methods["Generate"] = grpc.unary_unary_rpc_method_handler(_wrap(generate_servicer.generate))
    if not methods:
        raise ValueError("nothing to serve")
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers),
                         options=[("grpc.max_receive_message_length", MAX_MSG), ("grpc.max_send_message_length", MAX_MSG)])
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(SERVICE, methods),))
    bound = server.add_insecure_port(f"[::]:{port}")
    if bound == 0:
        raise RuntimeError(f"could not bind port {port}")
    server.start()
    return server, bound


class Client:
    """Code/gRPC/client.py:7-11 for this service."""

    def __init__(self, target: str = f"127.0.0.1:{DEFAULT_PORT}", ready_timeout: float = 60.0):
        import grpc
        self.channel = grpc.insecure_channel(target, options=[("grpc.max_receive_message_length", MAX_MSG),
                                                              ("grpc.max_send_message_length", MAX_MSG)])
        grpc.channel_ready_future(self.channel).result(timeout=ready_timeout)
        self._forward = self.channel.unary_unary(f"/{SERVICE}/Forward")
        self._generate = self.channel.unary_unary(f"/{SERVICE}/Generate")

    def forward(self, arr: np.ndarray, request_id: int) -> np.ndarray:
        return decode_tensor(self._forward(encode_tensor(arr, request_id)))[0]

    def release(self, request_id: int):
        self._forward(encode_tensor(np.zeros((0,), np.int64), request_id))

    def generate(self, input_ids=(), **kw) -> dict:
        return decode_generate_reply(self._generate(encode_generate_request(input_ids, **kw)))

    def close(self):
        self.channel.close()


# ------------------------------------------------------------------------------------------------
# REST twin (Code/gRPC/rest_api.py:6-15: a FastAPI app, one JSON object per call, uvicorn on 0.0.0.0:8000)
# ------------------------------------------------------------------------------------------------
def rest_app(generate_servicer: GenerateServicer):
    """FastAPI app over the same GenerateServicer: `GET /` -> {"model", "current_time"} (the reference's
    endpoint, with the model name beside the time stamp), `POST /generate` with the GenerateRequest fields as
    JSON -> the GenerateReply fields as JSON.  Run it like the reference does: ``uvicorn.run(app,
    host="0.0.0.0", port=8000)``."""
    from fastapi import Body, FastAPI, HTTPException
    app = FastAPI()

    @app.get("/")
    def status():
        return {"model": generate_servicer.pipe.cfg.name, "current_time": time.strftime("%Y-%m-%d %H:%M:%S")}

    @app.post("/generate")
    def generate(req: dict = Body(...)):
        unknown = set(req) - set(_GEN_DEFAULTS)
        if unknown:
            raise HTTPException(status_code=422, detail=f"unknown fields: {sorted(unknown)}")
        try:
            return decode_generate_reply(generate_servicer.generate(encode_generate_request(**req)))
        except (ValueError, TypeError) as e:
            raise HTTPException(status_code=400, detail=str(e))
    return app
