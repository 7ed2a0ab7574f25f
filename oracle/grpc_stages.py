"""CPU baseline C1 (BASELINE.json configs[0]): GPT-2 fp32 split into layer-range stages, each a
gRPC server on localhost, hidden states handed from stage to stage as raw bytes in a unary RPC.

TEST / BASELINE INFRASTRUCTURE ONLY.  Nothing here is on the product path: only ``tests/`` and
``bench.py``'s CPU-baseline legs may import or execute it.  It is the CPU arm that
``BASELINE.json:north_star`` asks to be timed beside the GPU pipeline ("the reference's CPU gRPC
pipeline ... on the box's own host cores, with the core count stated").

What the reference really has (and what is kept): the only RPC code in the reference is
``Code/gRPC/server.py:13-19`` -- ``grpc.server(futures.ThreadPoolExecutor(max_workers=10))``, one
servicer, ``add_insecure_port("[::]:50051")``, ``start()`` -- and ``Code/gRPC/client.py:7-11`` --
``grpc.insecure_channel(host:port)``, a stub, one unary call; the service is declared in proto3
(``time_service.proto:5-7``).  Its payload is a timestamp; here the same conventions carry a tensor:

    syntax = "proto3";
    service Stage { rpc Forward (TensorMsg) returns (TensorMsg); }
    message TensorMsg { bytes data = 1; repeated int64 shape = 2; string dtype = 3; int64 request_id = 4; }

``protoc`` / ``grpc_tools`` are not in this image, so the message is (de)serialised by hand in the
proto3 wire format (``encode_msg`` / ``decode_msg``; checked in the tests against a descriptor
built at run time with the protobuf package) and the method is registered with
``grpc.method_handlers_generic_handler``.

The model arithmetic is HF's own: every stage instantiates ``GPT2LMHeadModel`` from the same
config under the same seed (no checkpoints offline), keeps the ``GPT2Block`` objects of its layer
range (transformers 5.5 ``models/gpt2/modeling_gpt2.py`` ``GPT2Block.forward``), stage 0 keeps
``wte``/``wpe``, the last stage ``ln_f`` + ``lm_head``; the KV cache (HF ``DynamicCache``) is held
server-side per ``request_id``.  The client sends token ids to stage 0, relays the hidden tile to
the next stage, takes the arg-max of the last stage's logits (greedy) and feeds it back.
"""
from __future__ import annotations

import argparse
import os
import socket
import subprocess
import sys
import time
from concurrent import futures

import numpy as np

SERVICE = "Stage"
METHOD = "Forward"
DEFAULT_PORTS = (50051, 50052)          # server.py:16 uses 50051; the second stage takes the next port
MAX_MSG = 256 << 20
_DTYPES = {"float32": np.float32, "int64": np.int64, "float16": np.float16}


# ------------------------------------------------------------------------------------------------
# proto3 wire format of TensorMsg, by hand
# ------------------------------------------------------------------------------------------------
def _varint(v: int) -> bytes:
    v &= (1 << 64) - 1                  # int64: negative values as 10-byte two's complement
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf: bytes, i: int):
    shift = v = 0
    while True:
        b = buf[i]
        i += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            break
        shift += 7
    if v >= 1 << 63:
        v -= 1 << 64
    return v, i


def encode_msg(arr: np.ndarray, request_id: int) -> bytes:
    """TensorMsg{data=1: bytes, shape=2: packed int64, dtype=3: string, request_id=4: int64}.
    proto3 omits fields holding their default (empty / zero)."""
    arr = np.require(arr, requirements="C")         # (ascontiguousarray would turn a 0-d array into 1-d)
    name = arr.dtype.name
    if name not in _DTYPES:
        raise ValueError(f"unsupported dtype {name}")
    data = arr.tobytes()
    out = bytearray()
    if data:
        out += b"\x0a" + _varint(len(data)) + data
    if arr.ndim:
        packed = b"".join(_varint(int(s)) for s in arr.shape)
        out += b"\x12" + _varint(len(packed)) + packed
    nb = name.encode()
    out += b"\x1a" + _varint(len(nb)) + nb
    if request_id:
        out += b"\x20" + _varint(int(request_id))
    return bytes(out)


def decode_msg(buf: bytes):
    """-> (ndarray, request_id).  Accepts packed and unpacked `shape`, any field order, and skips
    unknown fields, as a proto3 parser does."""
    data, shape, dtype, rid, i = b"", [], "float32", 0, 0
    while i < len(buf):
        key, i = _read_varint(buf, i)
        field, wt = key >> 3, key & 7
        if wt == 0:
            v, i = _read_varint(buf, i)
            if field == 2:
                shape.append(v)
            elif field == 4:
                rid = v
        elif wt == 2:
            n, i = _read_varint(buf, i)
            chunk = bytes(buf[i:i + n])
            i += n
            if field == 1:
                data = chunk
            elif field == 2:
                j = 0
                while j < len(chunk):
                    v, j = _read_varint(chunk, j)
                    shape.append(v)
            elif field == 3:
                dtype = chunk.decode()
        elif wt == 1:
            i += 8
        elif wt == 5:
            i += 4
        else:
            raise ValueError(f"bad wire type {wt}")
    if dtype not in _DTYPES:
        raise ValueError(f"unsupported dtype {dtype!r}")
    arr = np.frombuffer(data, dtype=_DTYPES[dtype]).reshape(shape)
    return arr, rid


# ------------------------------------------------------------------------------------------------
# a stage: HF GPT2Blocks [lo, hi) on CPU
# ------------------------------------------------------------------------------------------------
def gpt2_config(name: str = "gpt2-small"):
    """HF config from the public dims; `gpt2-tiny` is the test model."""
    from transformers import GPT2Config
    if name == "gpt2-small":
        return GPT2Config(vocab_size=50257, n_positions=1024, n_embd=768, n_layer=12, n_head=12)
    if name == "gpt2-tiny":
        return GPT2Config(vocab_size=512, n_positions=128, n_embd=64, n_layer=4, n_head=4)
    raise ValueError(name)


def build_model(name: str, seed: int = 0):
    """Random-init GPT2LMHeadModel (no checkpoints offline); dropout off, fp32, eager attention."""
    import torch
    from transformers import GPT2LMHeadModel
    cfg = gpt2_config(name)
    cfg.attn_pdrop = cfg.embd_pdrop = cfg.resid_pdrop = 0.0
    cfg._attn_implementation = "eager"
    torch.manual_seed(seed)
    return GPT2LMHeadModel(cfg).eval()


class StageModel:
    """Layers [lo, hi) of the model, with the embedding on the first and ln_f + lm_head on the last
    stage; one DynamicCache per request id."""

    def __init__(self, name: str, lo: int, hi: int, seed: int = 0, threads: int | None = None):
        import copy
        import torch
        if threads:
            torch.set_num_threads(threads)
        full = build_model(name, seed)
        n = full.config.n_layer
        assert 0 <= lo < hi <= n
        self.first, self.last = lo == 0, hi == n
        self.cfg = copy.deepcopy(full.config)
        self.cfg.n_layer = self.cfg.num_hidden_layers = hi - lo
        self.blocks = [full.transformer.h[i] for i in range(lo, hi)]
        for j, b in enumerate(self.blocks):            # the stage-local cache is indexed from 0
            b.attn.layer_idx = j
        self.wte, self.wpe = (full.transformer.wte, full.transformer.wpe) if self.first else (None, None)
        self.ln_f, self.lm_head = (full.transformer.ln_f, full.lm_head) if self.last else (None, None)
        self.caches = {}

    def forward(self, x: np.ndarray, request_id: int) -> np.ndarray:
        """first stage: int64 ids [T]; others: fp32 hidden [T, d].  Returns fp32 hidden [T, d], or
        the last position's logits [1, V] on the last stage.  An empty input drops the request's cache."""
        import torch
        from transformers import DynamicCache
        from transformers.masking_utils import create_causal_mask
        if x.size == 0:
            self.caches.pop(request_id, None)
            return np.zeros((0,), np.float32)
        cache = self.caches.get(request_id)
        if cache is None:
            cache = self.caches[request_id] = DynamicCache(config=self.cfg)
        with torch.no_grad():
            T = x.shape[0]
            pos = (torch.arange(T) + cache.get_seq_length()).unsqueeze(0)
            if self.first:
                ids = torch.from_numpy(x.astype(np.int64)).unsqueeze(0)
                h = self.wte(ids) + self.wpe(pos)
            else:
                h = torch.from_numpy(x.astype(np.float32)).unsqueeze(0)
            mask = create_causal_mask(config=self.cfg, inputs_embeds=h, attention_mask=None,
                                      past_key_values=cache, position_ids=pos)
            for blk in self.blocks:
                h = blk(h, cache, mask, use_cache=True, position_ids=pos)
                if isinstance(h, tuple):
                    h = h[0]
            if self.last:
                h = self.lm_head(self.ln_f(h[:, -1:, :]))
            return h[0].float().numpy()


def make_server(stage: StageModel, port: int):
    """server.py:13-19's recipe with a generic handler in place of the generated *_pb2_grpc glue.
    Returns (server, bound_port)."""
    import grpc

    def forward(request: bytes, context):
        arr, rid = decode_msg(request)
        return encode_msg(stage.forward(arr, rid), rid)

    handler = grpc.method_handlers_generic_handler(SERVICE, {
        METHOD: grpc.unary_unary_rpc_method_handler(forward, request_deserializer=None, response_serializer=None)})
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=10),
                         options=[("grpc.max_receive_message_length", MAX_MSG), ("grpc.max_send_message_length", MAX_MSG)])
    server.add_generic_rpc_handlers((handler,))
    bound = server.add_insecure_port(f"[::]:{port}")
    if bound == 0:
        raise RuntimeError(f"could not bind port {port}")
    server.start()
    return server, bound


class StageClient:
    """client.py:7-11's recipe: insecure channel + one unary call per hop."""

    def __init__(self, ports, host: str = "127.0.0.1", ready_timeout: float = 300.0):
        import grpc
        opts = [("grpc.max_receive_message_length", MAX_MSG), ("grpc.max_send_message_length", MAX_MSG)]
        self.channels = [grpc.insecure_channel(f"{host}:{p}", options=opts) for p in ports]
        for ch in self.channels:
            grpc.channel_ready_future(ch).result(timeout=ready_timeout)
        self.calls = [ch.unary_unary(f"/{SERVICE}/{METHOD}", request_serializer=None, response_deserializer=None)
                      for ch in self.channels]
        self.bytes_sent = 0

    def forward(self, ids: np.ndarray, request_id: int) -> np.ndarray:
        x = np.asarray(ids, dtype=np.int64)
        for call in self.calls:
            msg = encode_msg(x, request_id)
            self.bytes_sent += len(msg)
            x, _ = decode_msg(call(msg))
        return x                                              # logits [1, V]

    def release(self, request_id: int):
        for call in self.calls:
            call(encode_msg(np.zeros((0,), np.int64), request_id))

    def generate(self, prompt: np.ndarray, max_new_tokens: int, request_id: int = 1):
        """Greedy.  Returns (tokens [max_new_tokens], prefill seconds, decode seconds)."""
        t0 = time.perf_counter()
        tok = int(np.argmax(self.forward(prompt, request_id)[0]))
        t1 = time.perf_counter()
        out = [tok]
        for _ in range(max_new_tokens - 1):
            tok = int(np.argmax(self.forward(np.array([tok]), request_id)[0]))
            out.append(tok)
        t2 = time.perf_counter()
        self.release(request_id)
        return np.array(out, dtype=np.int64), t1 - t0, t2 - t1

    def close(self):
        for ch in self.channels:
            ch.close()


# ------------------------------------------------------------------------------------------------
# launcher: one server process per stage, used by bench.py
# ------------------------------------------------------------------------------------------------
def _free_port(preferred: int) -> int:
    for p in (preferred, 0):
        with socket.socket(socket.AF_INET6 if socket.has_ipv6 else socket.AF_INET, socket.SOCK_STREAM) as s:
            try:
                s.bind(("", p))
                return s.getsockname()[1]
            except OSError:
                continue
    raise RuntimeError("no free port")


def split_layers(n_layers: int, n_stages: int):
    base, rem = divmod(n_layers, n_stages)
    out, lo = [], 0
    for i in range(n_stages):
        hi = lo + base + (1 if i < rem else 0)
        out.append((lo, hi))
        lo = hi
    return out


def timed_run(model: str = "gpt2-small", n_stages: int = 2, prompt: int = 128, new: int = 32,
              steps: int = 5, warmup: int = 3, seed: int = 0) -> dict:
    """Start `n_stages` stage-server processes on localhost, run `warmup` + `steps` greedy
    generations of a seeded synthetic prompt through them, return the median timings."""
    import statistics
    cores = os.cpu_count() or 1
    cfg = gpt2_config(model)
    ranges = split_layers(cfg.n_layer, n_stages)
    ports = [_free_port(DEFAULT_PORTS[0] + i) for i in range(n_stages)]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    procs = []
    try:
        for (lo, hi), port in zip(ranges, ports):
            # the stages of one request never run at the same time, so each may use every core
            procs.append(subprocess.Popen(
                [sys.executable, "-m", "oracle.grpc_stages", "--model", model, "--layers", f"{lo}:{hi}", "--port", str(port),
                 "--seed", str(seed), "--threads", str(cores)], cwd=root, stdout=subprocess.DEVNULL))
        client = StageClient(ports)
        rng = np.random.default_rng(1234)
        ids = rng.integers(0, cfg.vocab_size, size=(prompt,), dtype=np.int64)
        pre, dec, toks = [], [], None
        for i in range(warmup + steps):
            toks, tp, td = client.generate(ids, new, request_id=i + 1)
            if i >= warmup:
                pre.append(tp)
                dec.append(td)
        sent = client.bytes_sent // (warmup + steps)
        client.close()
    finally:
        for p in procs:
            p.terminate()
        for p in procs:
            try:
                p.wait(timeout=20)
            except subprocess.TimeoutExpired:
                p.kill()
    tp, td = statistics.median(pre), statistics.median(dec)
    return dict(value=(prompt + new) / (tp + td), unit="tokens/s", cores=cores, kind="reference",
                prefill_tokens_per_s=prompt / tp, decode_tokens_per_s=(new - 1) / td, ms_per_step=(tp + td) * 1e3,
                request_bytes_per_generation=sent, tokens=[int(t) for t in toks],
                sample=(f"{model} fp32 (HF GPT2Block, random-init seed {seed}) in {n_stages} layer-range stage servers "
                        f"{ranges} on localhost gRPC (ports {ports}, unary Stage/Forward, 10-thread pools, server-side "
                        f"KV cache), {prompt}-token synthetic prompt + {new} greedy tokens; median of {steps} after "
                        f"{warmup} warm-ups; {cores} host cores, torch threads = {cores} per stage process"))


def main():
    ap = argparse.ArgumentParser(description="one CPU stage server (see module docstring)")
    ap.add_argument("--model", default="gpt2-small")
    ap.add_argument("--layers", required=True, help="lo:hi")
    ap.add_argument("--port", type=int, default=DEFAULT_PORTS[0])
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--threads", type=int, default=0)
    a = ap.parse_args()
    lo, hi = (int(v) for v in a.layers.split(":"))
    import warnings
    warnings.filterwarnings("ignore")
    stage = StageModel(a.model, lo, hi, a.seed, a.threads or None)
    server, port = make_server(stage, a.port)
    print(f"gRPC stage server: layers [{lo},{hi}) of {a.model} on port {port}", file=sys.stderr, flush=True)
    server.wait_for_termination()


if __name__ == "__main__":
    main()
