"""CPU tests: the C-ABI library loads and exports what include/b200stage.h declares (argument
checks only -- no compute without a GPU), model configs, layer partitioning, and the multi-stage
schedule on two gloo ranks with a stand-in stage."""
import ctypes as C
import os
import socket

import numpy as np
import pytest
import torch

import b200stage
from b200stage import _abi
from b200stage.pipeline import DistPipeline, split_layers


# ---- C ABI -----------------------------------------------------------------------------------
def test_library_exports_every_declared_symbol():
    L = _abi.lib()
    declared = _abi.header_functions()
    assert len(declared) >= 17 and "b2_gemm_w8a8" in declared and "b2_init" in declared
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/b200stage.h but not exported"
        assert name in _abi.SIGNATURES, f"{name} has no ctypes signature"
    assert set(_abi.SIGNATURES) <= set(declared)
    assert L.b2_version() == 100


def test_error_strings_and_argument_checks_need_no_gpu():
    L = _abi.lib()
    assert L.b2_strerror(0) == b"ok"
    for code in range(-8, 0):
        assert L.b2_strerror(code) not in (b"ok", b"unknown error")
    assert L.b2_strerror(-99) == b"unknown error"
    # empty problems succeed, null pointers / bad shapes are refused, before any CUDA call
    assert L.b2_gemm_w8a8(None, None, None, None, None, None, None, None, 0, 16, 16, None) == 0
    assert L.b2_gemm_w8a8(None, None, None, None, None, None, None, None, 4, 16, 16, None) == -3
    buf = (C.c_char * 4096)()
    p = C.cast(C.addressof(buf) + (-C.addressof(buf)) % 16, C.c_void_p)
    assert L.b2_gemm_w8a8(p, p, p, p, None, None, p, None, 4, 16, 20, None) == -1      # K % 16
    assert L.b2_gemm_w8a8(C.c_void_p(p.value + 1), p, p, p, None, None, p, None, 4, 16, 16, None) == -2
    assert L.b2_norm_quant(None, None, None, None, None, None, 0, 64, 1e-5, None) == 0
    assert L.b2_norm_quant(None, None, None, None, None, None, 3, 64, 1e-5, None) == -3
    assert L.b2_decode_attn(p, 64, p, p, p, None, p, 64, None, 0, 1, 6, 4, 64, 16, 8, 0.1, 0, None, 0, None, 0, None) == -1  # H % KV
    assert L.b2_decode_attn(p, 64, p, p, p, None, p, 64, None, 0, 1, 4, 4, 80, 16, 8, 0.1, 0, None, 0, None, 0, None) == -6  # head_dim
    assert L.b2_lm_head(None, None, None, 0, 8, 8, None) == 0


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_abi, "_lib", None)
    monkeypatch.setattr(_abi, "LIB_PATH", "/nonexistent/libb200stage.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _abi.lib()


def test_stage_has_no_cpu_path():
    with pytest.raises(RuntimeError, match="CUDA"):
        b200stage.Stage.load("tiny-llama-test", 0, 1, first=True, last=False, device="cpu")


# ---- configs and partitioning ------------------------------------------------------------------
def test_presets_match_public_dims():
    lin = {"tinyllama-1.1b": 44.04e6, "llama-2-7b": 202.4e6, "llama-2-13b": 317.2e6, "mistral-7b": 218.1e6,
           "gpt2": 7.08e6}
    for name, per_layer in lin.items():
        cfg = b200stage.get_config(name)
        assert abs(cfg.linear_params_per_layer() - per_layer) / per_layer < 2e-3, name
    c = b200stage.get_config({"hidden_size": 4096, "num_attention_heads": 32, "num_key_value_heads": 8,
                              "num_hidden_layers": 32, "intermediate_size": 14336, "vocab_size": 32000,
                              "sliding_window": 4096, "rms_norm_eps": 1e-5, "model_type": "mistral"})
    assert (c.n_kv_heads, c.head_dim, c.window, c.qkv_out) == (8, 128, 4096, 6144)
    g = b200stage.get_config({"n_embd": 768, "n_head": 12, "n_layer": 12, "vocab_size": 50257, "model_type": "gpt2"})
    assert (g.norm, g.rope, g.gated, g.ffn) == ("layernorm", False, False, 3072)


@pytest.mark.parametrize("L,P", [(32, 1), (32, 8), (40, 8), (22, 4), (12, 5)])
def test_split_layers_contiguous_and_balanced(L, P):
    r = split_layers(L, P)
    assert r[0][0] == 0 and r[-1][1] == L and all(a[1] == b[0] for a, b in zip(r, r[1:]))
    sizes = [e - b for b, e in r]
    assert max(sizes) - min(sizes) <= 1 and min(sizes) >= 1


def test_weights_are_seeded_and_stage_local():
    from b200stage import weights
    cfg = b200stage.get_config("tiny-llama-test")
    a = weights.build_stage_weights(cfg, 1, 3, False, False, 7)
    b = weights.build_stage_weights(cfg, 0, 4, True, True, 7)
    for k in a["layers"][0]:
        assert torch.equal(a["layers"][0][k], b["layers"][1][k]), k        # layer 1 is layer 1 in any shard
    assert "embed" not in a and "lm_head" not in a and b["lm_head"] is b["embed"]
    c = weights.build_stage_weights(cfg, 1, 2, False, False, 8)
    assert not torch.equal(a["layers"][0]["wqkv"], c["layers"][0]["wqkv"])
    lw = a["layers"][0]
    assert lw["wqkv"].dtype == torch.int8 and lw["wqkv"].abs().max() == 127
    assert lw["wgu"].shape == (2 * cfg.ffn, cfg.d_model) and lw["sm_d"].shape == (cfg.ffn,)


# ---- two gloo ranks, stand-in stage ------------------------------------------------------------
D, V = 16, 32


def _fake_tables():
    g = torch.Generator().manual_seed(0)
    return torch.randn(V, D, generator=g).half(), torch.randn(V, D, generator=g)


def _fake_stage(rank, world):
    emb, head = _fake_tables()

    def fwd(x, positions, seq_ids, logits):
        pos = torch.as_tensor(np.asarray(positions), dtype=torch.float32)[:, None]
        h = emb[x.view(-1)].float() if rank == 0 else x.float()
        h = (h * (1.0 + 0.25 * rank) + 0.01 * pos + 0.001 * float(np.asarray(seq_ids)[0])).half()
        if rank == world - 1 and logits != "none":
            hs = h if (len(positions) == 1 or logits == "all") else h[-1:]
            return hs.float() @ head.T
        return h
    return fwd


def _single_process_reference(prompts, chunk, new):
    """The same fake stages chained in one process: what the 2-rank run must reproduce."""
    f0, f1 = _fake_stage(0, 2), _fake_stage(1, 2)
    toks = []
    for p, ids in enumerate(prompts):
        T = len(ids)
        for c0 in range(0, T, chunk):
            c1 = min(T, c0 + chunk)
            pos = np.arange(c0, c1)
            y = f1(f0(ids[c0:c1], pos, np.full(c1 - c0, p), "none"), pos, np.full(c1 - c0, p), "last" if c1 == T else "none")
        toks.append([int(y.argmax(-1)[0])])
    for i in range(1, new):
        for p, ids in enumerate(prompts):
            pos = np.array([len(ids) + i - 1])
            y = f1(f0(torch.tensor([toks[p][-1]]), pos, [p], "none"), pos, [p], "last")
            toks[p].append(int(y.argmax(-1)[0]))
    return toks


def _worker(rank, world, port, prompts, chunk, new, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pipe = DistPipeline(rank, world, D, torch.device("cpu"), _fake_stage(rank, world), V)
        lens = [len(p) for p in prompts]
        it = iter(lens)
        pipe._bcast_len = lambda T: next(it)
        lgs = pipe.prefill_stream(prompts if pipe.first else [None] * len(prompts), chunk)
        toks = []
        for p in range(len(prompts)):
            t = lgs[p].argmax(-1).view(1) if pipe.last else (torch.empty(1, dtype=torch.int64) if pipe.first else None)
            toks.append(pipe.return_tokens(t))
        gen = [[int(t[0])] for t in toks] if (pipe.first or pipe.last) else None
        pipe.setup_groups()
        rows = [[p] for p in range(len(prompts))]
        # two steps through decode_round (one step per call), the rest as one continuous stream
        for i in range(1, 3):
            pos = [[lens[p] + i - 1] for p in range(len(prompts))]
            toks = pipe.decode_round(toks, pos, rows)
            if gen is not None:
                for p in range(len(prompts)):
                    gen[p].append(int(toks[p][0]))
        out = pipe.decode_stream(toks, [[lens[p] + 2] for p in range(len(prompts))], rows, new - 3)
        if gen is not None:
            for p in range(len(prompts)):
                gen[p] += [int(t[0]) for t in out[p]]
        pipe.drain()
        q.put((rank, gen))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_rank_pipeline_schedule_gloo():
    import torch.multiprocessing as mp
    g = torch.Generator().manual_seed(3)
    prompts = [torch.randint(0, V, (n,), generator=g) for n in (11, 7, 16)]     # ragged: chunks 4,4,3 / 4,3 / 4x4
    chunk, new = 4, 7
    want = _single_process_reference(prompts, chunk, new)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, prompts, chunk, new, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0] == want, "tokens returned to rank 0 differ from the single-process chain"
    assert got[1] == want, "tokens on the last rank differ from the single-process chain"


# ----------------------------------------------------------------------------------------------
# lm_head shared by all pipeline stages: the per-stage launch lists, run against each other
# ----------------------------------------------------------------------------------------------
def _run_ring_schedules(W, G, steps):
    """Execute the W stages' launch lists as W in-order streams whose waits block: returns the number of launches
    retired, or raises on a deadlock / a message overwritten before every reader has consumed it."""
    from b200stage.pipeline import ring_head_schedule, ring_head_ranks
    hr = ring_head_ranks(W)
    assert (hr == list(range(1, W - 1))) if W >= 4 else (hr == list(range(W)))
    lists = [list(ring_head_schedule(r, W, G, steps)) for r in range(W)]
    pc = [0] * W
    done = set()                                  # (rank, op, g, i)
    readers_left = {}                             # message -> readers that still have to consume it

    def ready(r, op, g, i):
        if op == "wait_h":
            return (r - 1, "send_h", g, i) in done
        if op == "slice":
            return (W - 1, "send_hn", g, i) in done
        if op == "merge":
            return all((q, "slice", g, i) in done for q in hr)
        return True

    def retire(r, op, g, i):
        # one slot per group and edge: writing step i needs step i-1 of the same group fully consumed
        if op in ("send_h", "send_hn") and i > 0:
            prev = ("h", r, g, i - 1) if op == "send_h" else ("hn", g, i - 1)
            assert readers_left.get(prev, 0) == 0, (f"stage {r} overwrites {prev} before it was read", W, G)
        if op == "slice" and i > 0:
            assert readers_left.get(("win", g, i - 1), 0) == 0, (f"stage {r} overwrites the winners of {(g, i - 1)}", W, G)
        if op == "send_h":
            readers_left[("h", r, g, i)] = 1
        elif op == "send_hn":
            readers_left[("hn", g, i)] = len(hr)
        elif op == "wait_h":
            readers_left[("h", r - 1, g, i)] -= 1
        elif op == "slice":
            readers_left[("hn", g, i)] -= 1
            readers_left.setdefault(("win", g, i), 1)
        elif op == "merge":
            readers_left[("win", g, i)] = 0
        done.add((r, op, g, i))

    total = sum(len(x) for x in lists)
    n = 0
    while n < total:
        moved = False
        for r in range(W):
            while pc[r] < len(lists[r]) and ready(r, *lists[r][pc[r]]):
                retire(r, *lists[r][pc[r]])
                pc[r] += 1
                n += 1
                moved = True
        assert moved, ("deadlock", W, G, steps, [lists[r][pc[r]] if pc[r] < len(lists[r]) else None for r in range(W)])
    return n, lists


@pytest.mark.parametrize("W,G,steps", [(2, 2, 3), (2, 3, 4), (3, 3, 1), (4, 4, 5), (4, 7, 3), (8, 8, 4), (8, 9, 2), (8, 32, 2)])
def test_ring_head_schedule_runs_to_completion(W, G, steps):
    n, lists = _run_ring_schedules(W, G, steps)
    for r in range(W):
        ops_r = lists[r]
        assert sum(1 for o in ops_r if o[0] == "layers") == G * steps
        from b200stage.pipeline import ring_head_ranks
        want_slices = sorted((g, i) for i in range(steps) for g in range(G)) if r in ring_head_ranks(W) else []
        assert sorted(o[1:] for o in ops_r if o[0] == "slice") == want_slices
        if r == 0:
            assert sorted(o[1:] for o in ops_r if o[0] == "merge") == sorted((g, i) for i in range(steps) for g in range(G))
            # a token is merged before the layers that consume it
            for g in range(G):
                for i in range(1, steps):
                    assert ops_r.index(("merge", g, i - 1)) < ops_r.index(("layers", g, i))


def test_ring_head_schedule_needs_as_many_groups_as_stages():
    """Fewer groups than stages: a stage would wait for a hidden state that depends on a token it has not helped to
    pick yet.  DistPipeline falls back to the two-way split there; the schedule itself must show the deadlock."""
    with pytest.raises(AssertionError, match="deadlock"):
        _run_ring_schedules(4, 2, 3)


@pytest.mark.parametrize("W", [2, 3, 4, 5, 8])
@pytest.mark.parametrize("V", [32000, 50257, 512, 50304])
def test_shared_head_slices_cover_the_vocabulary(W, V):
    """head_bounds: the slices of the stages in ring_head_ranks tile [0, V) in order, every cut on a multiple of 8
    rows (16-byte aligned fp16 rows for any d % 8 == 0), none empty."""
    from b200stage.pipeline import DistPipeline, ring_head_ranks
    hr = ring_head_ranks(W)
    pipes = [DistPipeline(r, W, 64, torch.device("cpu"), None, V) for r in hr]
    cuts = [p.head_bounds(p.rank) for p in pipes]
    assert cuts[0][0] == 0 and cuts[-1][1] == V
    for (lo, hi), (lo2, _) in zip(cuts, cuts[1:] + [(V, V)]):
        assert lo < hi and hi == lo2 and lo % 8 == 0
