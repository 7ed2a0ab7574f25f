"""b2_fused_linear (decode-time prologue + quantise + GEMV + epilogue in one launch) must be
bit-identical to the oracle composition and to the unfused entry points it replaces."""
import numpy as np
import pytest
import torch

import b200stage
from b200stage import ops
from b200stage.pipeline import Pipeline
from oracle import w8a8 as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def n(x):
    return x.detach().cpu().numpy()


def _w(r, N, K):
    w = r.integers(-127, 128, (N, K), dtype=np.int8)
    s = r.uniform(1e-4, 2e-3, N).astype(np.float32)
    return w, s


@pytest.mark.parametrize("M", [1, 2, 3, 4, 7, 8])
@pytest.mark.parametrize("N,K", [(4096, 4096), (2560, 2048), (1000, 768), (24, 144)])
@pytest.mark.parametrize("ln", [False, True])
def test_fused_norm_linear_exact(M, N, K, ln):
    r = np.random.default_rng(M * 100 + N + K)
    x = (r.standard_normal((M, K)) * r.uniform(0.2, 3, (M, 1))).astype(np.float16)
    g = (1 + 0.1 * r.standard_normal(K)).astype(np.float16)
    b = (0.05 * r.standard_normal(K)).astype(np.float16) if ln else None
    sm = r.uniform(0.3, 3, K).astype(np.float32)
    w, ws = _w(r, N, K)
    bias = (0.1 * r.standard_normal(N)).astype(np.float16)
    res = r.standard_normal((M, N)).astype(np.float16)
    q, s = O.norm_quant(x, g, b, sm, 1e-5)
    ref, _ = O.gemm_w8a8(q, w, s, ws, bias, res)
    out = torch.empty((M, N), dtype=torch.float16, device=DEV)
    ops.fused_linear(t(x), t(w), t(ws), x_kind=ops.X_LN if ln else ops.X_RMS, gamma=t(g), beta=None if b is None else t(b),
                     smooth=t(sm), eps=1e-5, bias=t(bias), residual=t(res), out=out)
    assert np.array_equal(n(out).view(np.uint16), ref.view(np.uint16))


@pytest.mark.parametrize("M", [2, 5, 8])
@pytest.mark.parametrize("N,K", [(47360, 256), (45058, 208), (37890, 1040)])
def test_fused_batch_stream_many_units_per_cta(M, N, K):
    """The token-batch (mma.sync) stream with 10-20 sixteen-row units per CTA: both partial-sum buffers, several
    full rounds of 8 parked units and a partial last round, a ragged last unit (N/2 not a multiple of 8), K not a
    multiple of 64 (a 64-byte piece that only some lanes load) and K spanning two batches per unit.  Three runs
    must agree bit for bit with the oracle (a shared-memory race would show as run-to-run differences)."""
    r = np.random.default_rng(M + N + K)
    x = (r.standard_normal((M, K)) * r.uniform(0.2, 3, (M, 1))).astype(np.float16)
    g = (1 + 0.1 * r.standard_normal(K)).astype(np.float16)
    sm = r.uniform(0.3, 3, K).astype(np.float32)
    w, ws = _w(r, N, K)
    q, s = O.norm_quant(x, g, None, sm, 1e-5)
    ref, _ = O.gemm_w8a8(q, w, s, ws)
    for _ in range(3):
        out = torch.zeros((M, N), dtype=torch.float16, device=DEV)
        ops.fused_linear(t(x), t(w), t(ws), x_kind=ops.X_RMS, gamma=t(g), smooth=t(sm), eps=1e-5, out=out)
        assert np.array_equal(n(out).view(np.uint16), ref.view(np.uint16))


@pytest.mark.parametrize("M,N,K", [(1, 4096, 4096), (5, 512, 1024), (8, 2048, 256)])
def test_fused_plain_linear_exact_and_zeroing(M, N, K):
    r = np.random.default_rng(M + N + K)
    x = (2 * r.standard_normal((M, K))).astype(np.float16)
    x[0] = 0                                                            # an all-zero row: scale 0
    sm = r.uniform(0.3, 3, K).astype(np.float32)
    w, ws = _w(r, N, K)
    q, s = O.smooth_quant(x, sm)
    ref, _ = O.gemm_w8a8(q, w, s, ws)
    out = torch.empty((M, N), dtype=torch.float16, device=DEV)
    dirty = torch.full((8,), 3.0, device=DEV)
    ops.fused_linear(t(x), t(w), t(ws), x_kind=ops.X_PLAIN, smooth=t(sm), out=out, zero_amax=dirty)
    assert np.array_equal(n(out).view(np.uint16), ref.view(np.uint16))
    assert n(dirty)[:M].tolist() == [0.0] * M and (n(dirty)[M:] == 3.0).all()


@pytest.mark.parametrize("M", [1, 3, 8])
@pytest.mark.parametrize("d,F", [(4096, 11008), (2048, 5632), (256, 512), (4096, 14336), (5120, 13824)])
def test_fused_gated_mlp_exact(M, d, F):
    """norm -> gate_up GEMV with act epilogue -> y32/amax -> down GEMV, against the oracle's
    norm_quant -> gemm -> act_mul_quant -> gemm chain."""
    r = np.random.default_rng(M + d + F)
    h = r.standard_normal((M, d)).astype(np.float16)
    g = (1 + 0.1 * r.standard_normal(d)).astype(np.float16)
    sm_gu = r.uniform(0.5, 2, d).astype(np.float32)
    sm_d = r.uniform(0.5, 2, F).astype(np.float32)
    wgu, sgu = _w(r, 2 * F, d)
    sgu *= 4                                                            # gate values of a few units
    wd, sd = _w(r, d, F)
    q3, s3 = O.norm_quant(h, g, None, sm_gu, 1e-5)
    gu, _ = O.gemm_w8a8(q3, wgu, s3, sgu)
    y_ref, amax_ref = O.act_mul_y(gu, sm_d, "silu")
    q4, s4 = O.quant_rows(y_ref)
    ref, _ = O.gemm_w8a8(q4, wd, s4, sd, None, h)
    y = torch.empty((M, F), dtype=torch.float32, device=DEV)
    amax = torch.zeros((8,), dtype=torch.float32, device=DEV)
    ops.fused_linear(t(h), t(wgu), t(sgu), x_kind=ops.X_RMS, gamma=t(g), smooth=t(sm_gu), eps=1e-5, out=y,
                     out_kind=ops.O_GATED, act=0, out_smooth=t(sm_d), out_amax=amax)
    assert np.array_equal(n(y), y_ref)
    assert np.array_equal(n(amax)[:M], amax_ref)
    out = torch.empty((M, d), dtype=torch.float16, device=DEV)
    ops.fused_linear(y, t(wd), t(sd), x_kind=ops.X_Y32, x_amax=amax, residual=t(h), out=out)
    assert np.array_equal(n(out).view(np.uint16), ref.view(np.uint16))


def test_fused_rejects_large_m():
    x = torch.zeros((9, 64), dtype=torch.float16, device=DEV)
    w = torch.zeros((16, 64), dtype=torch.int8, device=DEV)
    with pytest.raises(b200stage._abi.B2Error):
        ops.fused_linear(x, w, torch.ones(16, device=DEV), x_kind=ops.X_PLAIN, out=torch.empty((9, 16), dtype=torch.float16, device=DEV))


@pytest.mark.parametrize("model", ["tiny-llama-test", "tiny-gqa-test", "tiny-gpt2-test", "tiny-phi-test", "tiny-neox-test"])
def test_fused_decode_equals_unfused_decode(model):
    """Whole-stage check: greedy decode with the fused kernels == with the unfused chain, bit for
    bit (logits of every step), with and without CUDA graphs, single and batched decode."""
    cfg = b200stage.get_config(model)
    ids = np.random.default_rng(8).integers(0, cfg.vocab, 37)
    outs = {}
    for fuse in (False, True, "tail"):       # "tail": o_proj -> gate_up -> down -> next qkv in one persistent launch
        for graphs in (False, True):
            pipe = Pipeline.load(cfg, 2, seed=11)
            pipe.use_graphs = graphs
            for s in pipe.stages:
                s.fuse_decode = bool(fuse)
                s.fuse_tail = fuse == "tail"
            toks = pipe.generate(ids, max_new_tokens=10, chunk=16)
            caches = pipe.new_caches(3, 64)
            for b in range(3):
                pipe.prefill(torch.from_numpy(ids[: 20 + 5 * b]).to(DEV), caches, row=b)
            lg = pipe.decode_step(torch.tensor([1, 2, 3], device=DEV), caches, [20, 25, 30], rows=np.arange(3)).clone()
            outs[(fuse, graphs)] = (toks.cpu(), lg.cpu())
    base = outs[(False, False)]
    for k, v in outs.items():
        assert torch.equal(v[0], base[0]), k
        assert torch.equal(v[1], base[1]), k


@pytest.mark.parametrize("M", [1, 3, 8])
@pytest.mark.parametrize("H,KV,hd,d", [(4, 4, 64, 256), (8, 2, 64, 512), (32, 8, 128, 1024), (32, 8, 128, 4096),
                                       (40, 40, 128, 5120)])
def test_fused_qkv_rope_append_exact(M, H, KV, hd, d):
    """qkv GEMV with rotary + KV append in the epilogue == oracle norm_quant -> gemm -> apply_rope
    -> cache write, bit for bit (q columns of the output buffer, k/v rows of the caches)."""
    r = np.random.default_rng(M + H + hd)
    N = (H + 2 * KV) * hd
    Smax, rows = 96, 9
    x = r.standard_normal((M, d)).astype(np.float16)
    g = (1 + 0.1 * r.standard_normal(d)).astype(np.float16)
    sm = r.uniform(0.5, 2, d).astype(np.float32)
    w, ws = _w(r, N, d)
    ws *= 3
    bias = (0.05 * r.standard_normal(N)).astype(np.float16)
    pos = r.permutation(Smax)[:M].astype(np.int32)
    seq = r.permutation(rows)[:M].astype(np.int32)
    q8, s8 = O.norm_quant(x, g, None, sm, 1e-5)
    qkv_ref, _ = O.gemm_w8a8(q8, w, s8, ws, bias)
    cos, sin = O.rope_table(hd, Smax)
    q_ref = O.apply_rope(qkv_ref[:, :H * hd].reshape(M, H, hd), pos, cos, sin)
    k_ref = O.apply_rope(qkv_ref[:, H * hd:(H + KV) * hd].reshape(M, KV, hd), pos, cos, sin)
    v_ref = qkv_ref[:, (H + KV) * hd:].reshape(M, KV, hd)
    kc = torch.zeros((rows, KV, Smax, hd), dtype=torch.float16, device=DEV)
    vc = torch.zeros_like(kc)
    out = torch.zeros((M, N), dtype=torch.float16, device=DEV)
    ops.fused_qkv(t(x), t(w), t(ws), out, t(cos), t(sin), t(pos), kc, vc, H, KV, hd, x_kind=ops.X_RMS, gamma=t(g),
                  smooth=t(sm), eps=1e-5, bias=t(bias), seq_ids=t(seq))
    assert np.array_equal(n(out)[:, :H * hd].reshape(M, H, hd).view(np.uint16), q_ref.view(np.uint16))
    kcn, vcn = n(kc), n(vc)
    for i in range(M):
        assert np.array_equal(kcn[seq[i], :, pos[i]].view(np.uint16), k_ref[i].view(np.uint16))
        assert np.array_equal(vcn[seq[i], :, pos[i]].view(np.uint16), v_ref[i].view(np.uint16))
    assert int((kcn != 0).any(-1).sum()) == M * KV          # nothing else in the cache was touched


@pytest.mark.parametrize("M", [1, 3, 8])
@pytest.mark.parametrize("with_qkv", [False, True])
def test_fused_layer_tail_equals_separate_launches(M, with_qkv):
    """o_proj -> gate_up -> down [-> next qkv] as one persistent launch (grid barriers) == the same
    four fused launches one after the other, bit for bit; no barrier time-out flag."""
    r = np.random.default_rng(M + 7)
    d, F, H, KV, hd, Smax = 1024, 2816, 8, 2, 128, 64
    N = (H + 2 * KV) * hd
    attn = t(r.standard_normal((M, d)).astype(np.float16))
    h = t(r.standard_normal((M, d)).astype(np.float16))
    g2 = t((1 + 0.1 * r.standard_normal(d)).astype(np.float16))
    g1n = t((1 + 0.1 * r.standard_normal(d)).astype(np.float16))
    sm_o, sm_gu, sm_q = (t(r.uniform(0.5, 2, d).astype(np.float32)) for _ in range(3))
    sm_d = t(r.uniform(0.5, 2, F).astype(np.float32))
    wo, so = map(t, _w(r, d, d))
    wgu, sgu = _w(r, 2 * F, d)
    wgu, sgu = t(wgu), t(sgu * 4)
    wd, sd = map(t, _w(r, d, F))
    wq, sq = _w(r, N, d)
    wq, sq = t(wq), t(sq * 3)
    cos, sin = (t(x) for x in O.rope_table(hd, Smax))
    pos = t(r.permutation(Smax)[:M].astype(np.int32))
    seq = t(np.arange(M, dtype=np.int32))
    res = {}
    for mode in ("separate", "tail"):
        kc = torch.zeros((M, KV, Smax, hd), dtype=torch.float16, device=DEV)
        vc = torch.zeros_like(kc)
        h1 = torch.empty((M, d), dtype=torch.float16, device=DEV)
        h2 = torch.empty((M, d), dtype=torch.float16, device=DEV)
        y = torch.empty((M, F), dtype=torch.float32, device=DEV)
        amax = torch.full((8,), 5.0, device=DEV)
        qkv = torch.zeros((M, N), dtype=torch.float16, device=DEV)
        rope = dict(cos=cos, sin=sin, positions=pos, seq_ids=seq, kcache=kc, vcache=vc, H=H, KV=KV, hd=hd)
        if mode == "separate":
            ops.fused_linear(attn, wo, so, x_kind=ops.X_PLAIN, smooth=sm_o, residual=h, out=h1, zero_amax=amax)
            ops.fused_linear(h1, wgu, sgu, x_kind=ops.X_RMS, gamma=g2, smooth=sm_gu, out=y, out_kind=ops.O_GATED,
                             out_smooth=sm_d, out_amax=amax)
            ops.fused_linear(y, wd, sd, x_kind=ops.X_Y32, x_amax=amax, residual=h1, out=h2)
            if with_qkv:
                ops.fused_qkv(h2, wq, sq, qkv, cos, sin, pos, kc, vc, H, KV, hd, x_kind=ops.X_RMS, gamma=g1n, smooth=sm_q,
                              seq_ids=seq)
        else:
            ctr = torch.zeros(4, dtype=torch.int32, device=DEV)
            d_o = ops.fused_desc(attn, wo, so, x_kind=ops.X_PLAIN, smooth=sm_o, residual=h, out=h1, zero_amax=amax)
            d_g = ops.fused_desc(h1, wgu, sgu, x_kind=ops.X_RMS, gamma=g2, smooth=sm_gu, out=y, out_kind=ops.O_GATED,
                                 out_smooth=sm_d, out_amax=amax, prefetch=(wd,))
            d_d = ops.fused_desc(y, wd, sd, x_kind=ops.X_Y32, x_amax=amax, residual=h1, out=h2)
            d_q = ops.fused_desc(h2, wq, sq, x_kind=ops.X_RMS, gamma=g1n, smooth=sm_q, out=qkv, out_kind=ops.O_QKV,
                                 rope=rope) if with_qkv else None
            for _ in range(3):                      # replays must be idempotent given re-zeroed counters
                ctr.zero_()
                ops.fused_layer_tail(d_o, d_g, d_d, d_q, M, ctr)
            torch.cuda.synchronize()
            assert int(ctr[3]) == 0, "grid barrier timed out"
            assert ctr[:3].tolist() == ([ctr[0].item()] * 3 if with_qkv else [ctr[0].item(), ctr[0].item(), 0])
        res[mode] = [x.clone() for x in (h1, y, amax[:M], h2, qkv[:, :H * hd], kc, vc)]
    for a, b in zip(res["separate"], res["tail"]):
        assert torch.equal(a, b)
