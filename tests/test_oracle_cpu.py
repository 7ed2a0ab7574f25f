"""CPU tests of the Tier-S oracle (oracle/w8a8.py): its pinned arithmetic, and its agreement with
the golden vectors in tests/golden (made by tests/golden/make_golden.py from HF's own Llama code,
the module the reference's from_pretrained call instantiates).  No GPU, no /root/reference."""
import os

import numpy as np
import pytest
import torch

import b200stage
from b200stage import weights
from oracle import w8a8 as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
f32, f16 = np.float32, np.float16


# ---- pinned arithmetic -----------------------------------------------------------------------
def test_exp_det_accuracy_and_clamp():
    x = np.linspace(-79, 79, 100001).astype(f32)
    rel = np.abs(O.exp_det(x) - np.exp(x.astype(np.float64))) / np.exp(x.astype(np.float64))
    assert rel.max() < 2e-7
    assert O.exp_det(f32(-1000)) == O.exp_det(f32(-80)) and O.exp_det(f32(1000)) == O.exp_det(f32(80))
    assert O.exp_det(f32(0)) == f32(1)


def test_silu_gelu_close_to_reference_forms():
    g = np.linspace(-30, 30, 20001).astype(f32)
    gd = g.astype(np.float64)
    assert np.abs(O.silu_det(g) - gd / (1 + np.exp(-gd))).max() < 3e-6
    ref = 0.5 * gd * (1 + np.tanh(0.7978845608028654 * (gd + 0.044715 * gd ** 3)))
    assert np.abs(O.gelu_new_det(g) - ref).max() < 3e-6


def test_quant_rows_half_even_zero_and_clamp():
    y = np.array([[127.0, 0.5, 1.5, 2.5, -0.5, -1.5, -2.5, 126.5], [0] * 8, [-3.0, 3.0, 1.0, 0, 0, 0, 0, 0]], f32)
    q, s = O.quant_rows(y)
    assert q[0].tolist() == [127, 0, 2, 2, 0, -2, -2, 126]          # ties go to even
    assert s[0] == f32(1.0) and s[1] == 0 and not q[1].any()
    assert q[2].tolist()[:3] == [-127, 127, 42] and s[2] == f32(3.0) / f32(127.0)
    assert np.abs(q).max() <= 127


def test_int8_acc_exact_vs_integer_einsum():
    r = np.random.default_rng(0)
    for M, N, K in [(3, 5, 16), (17, 9, 1024), (2, 3, 14336)]:
        a = r.integers(-127, 128, (M, K), dtype=np.int8)
        w = r.integers(-127, 128, (N, K), dtype=np.int8)
        assert np.array_equal(O.int8_acc(a, w), O.int8_acc_slow(a, w))
    a = np.full((1, 14336), 127, np.int8)
    assert O.int8_acc(a, a)[0, 0] == 14336 * 127 * 127                # int32 headroom at max K


def test_epilogue_single_rounding():
    acc = np.array([[2 ** 24 + 1, -3, 7]], np.int32)                   # float(acc) rounds to even
    sa = f32(2.0 ** -12)
    out = O.epilogue(acc, np.array([sa], f32), np.array([0.25, 4096.0, 8192.0], f32),
                     np.array([1.0, 0.0, 0.5], f16), np.array([[0.0, 1.0, -0.5]], f16))
    assert out[0, 0] == f16(1025.0)          # (2^24 [+1 lost in float(acc)]) * 2^-12 * 0.25 + 1
    assert out[0, 1] == f16(-2.0) and out[0, 2] == f16(14.0)


def test_norm_quant_matches_float64_reference():
    r = np.random.default_rng(1)
    x = r.standard_normal((5, 512)).astype(f16)
    g = (1 + 0.1 * r.standard_normal(512)).astype(f16)
    y = O.rmsnorm_rows(x, g, 1e-5)
    xd = x.astype(np.float64)
    ref = xd / np.sqrt((xd ** 2).mean(-1, keepdims=True) + 1e-5) * g.astype(np.float64)
    assert np.abs(y - ref).max() < 1e-5
    q, s = O.norm_quant(x, g, None, None, 1e-5)
    assert np.abs(q.astype(np.float64) * s[:, None] - ref).max() <= s.max() * 0.5 + 1e-5
    b = (0.1 * r.standard_normal(512)).astype(f16)
    yl = O.layernorm_rows(x, g, b, 1e-5)
    refl = (xd - xd.mean(-1, keepdims=True)) / np.sqrt(xd.var(-1, keepdims=True) + 1e-5) * g.astype(np.float64) + b
    assert np.abs(yl - refl).max() < 1e-5


def test_rope_preserves_norm_and_identity_at_zero():
    r = np.random.default_rng(2)
    x = r.standard_normal((6, 4, 64)).astype(f16)
    cos, sin = O.rope_table(64, 128)
    assert np.array_equal(O.apply_rope(x, np.zeros(6, int), cos, sin), x)
    y = O.apply_rope(x, np.arange(6) * 7, cos, sin).astype(np.float64)
    assert np.abs(np.linalg.norm(y, axis=-1) - np.linalg.norm(x.astype(np.float64), axis=-1)).max() < 2e-2
    c2, s2 = weights.rope_table(64, 128)
    assert np.array_equal(c2.numpy(), cos) and np.array_equal(s2.numpy(), sin)


def test_llama3_rope_scaling_matches_hf_and_host_table():
    """The "llama3" frequency remap (Llama-3.2-1B is the refiner of the reference's pipeline,
    combiner_quant_updated.py:177-203): oracle == transformers' own `_compute_llama3_parameters`,
    and the host table the kernels read == the oracle's, bit for bit."""
    from transformers import LlamaConfig
    from transformers.modeling_rope_utils import ROPE_INIT_FUNCTIONS
    for hd, theta, sc in [(64, 500000.0, ("llama3", 32.0, 1.0, 4.0, 8192)), (64, 500000.0, ("llama3", 8.0, 1.0, 4.0, 64)),
                          (128, 10000.0, ("llama3", 8.0, 1.0, 4.0, 512))]:
        hc = LlamaConfig(hidden_size=hd * 4, num_attention_heads=4, head_dim=hd, rope_theta=theta,
                         max_position_embeddings=16384,
                         rope_scaling=dict(rope_type="llama3", factor=sc[1], low_freq_factor=sc[2], high_freq_factor=sc[3],
                                           original_max_position_embeddings=sc[4]))
        hf_inv, att = ROPE_INIT_FUNCTIONS["llama3"](hc, "cpu")
        base = (1.0 / (theta ** (np.arange(0, hd, 2, dtype=np.float64) / hd))).astype(np.float32)
        mine = O.llama3_inv_freq(base, *sc[1:])
        assert att == 1.0
        assert np.allclose(mine, hf_inv.numpy(), rtol=2e-6, atol=0)
        assert not np.allclose(mine, base)                      # the remap does something at these sizes
        cos, sin = O.rope_table(hd, 300, theta, sc)
        c2, s2 = weights.rope_table(hd, 300, theta, sc)
        assert np.array_equal(c2.numpy(), cos) and np.array_equal(s2.numpy(), sin)


def test_attention_causal_window_and_gqa():
    r = np.random.default_rng(3)
    S, H, KV, hd = 12, 4, 2, 16
    q = r.standard_normal((S, H, hd)).astype(f16)
    k = r.standard_normal((S, KV, hd)).astype(f16)
    v = r.standard_normal((S, KV, hd)).astype(f16)
    out = O.attention(q, k, v, np.arange(S), 0.25, 0)
    assert np.allclose(out[0, :, :], v[0, [0, 0, 1, 1], :], atol=1e-3)      # first token sees only itself
    for i in (5, 11):
        for h in range(H):
            s = (q[i, h].astype(np.float64) @ k[: i + 1, h // 2].astype(np.float64).T) * 0.25
            p = np.exp(s - s.max()); p /= p.sum()
            assert np.abs(out[i, h] - p @ v[: i + 1, h // 2].astype(np.float64)).max() < 2e-3
    w = O.attention(q, k, v, np.arange(S), 0.25, 3)
    s = (q[11, 0].astype(np.float64) @ k[9:12, 0].astype(np.float64).T) * 0.25
    p = np.exp(s - s.max()); p /= p.sum()
    assert np.abs(w[11, 0] - p @ v[9:12, 0].astype(np.float64)).max() < 2e-3


def test_weight_quantiser_matches_oracle_statement():
    import torch
    r = np.random.default_rng(4)
    W = (0.02 * r.standard_normal((48, 64))).astype(f32)
    W[3] = 0
    s = r.uniform(0.5, 2.0, 64).astype(f32)
    q_ref, sc_ref = O.quantize_weight(W, s)
    q, sc = weights.quantize_weight(torch.from_numpy(W), torch.from_numpy(s))
    assert np.array_equal(q.numpy(), q_ref) and np.array_equal(sc.numpy(), sc_ref)
    assert sc_ref[3] == 0 and not q_ref[3].any()


# ---- golden vectors --------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["tiny-llama-test", "tiny-gqa-test", "tiny-gpt2-test", "tiny-llama3-test"])
def test_oracle_against_golden(name):
    g = np.load(os.path.join(GOLD, f"{name}.npz"))
    cfg = b200stage.get_config(name)
    seed, new, ids = int(g["seed"]), int(g["new"]), g["ids"]
    T = len(ids)
    sw = weights.to_numpy(weights.build_stage_weights(cfg, 0, cfg.n_layers, True, True, seed, "cpu"))
    cos, sin = O.rope_table(cfg.head_dim, cfg.max_pos, cfg.rope_theta, cfg.rope_scaling)
    cache = O.KVCacheOracle(cfg.n_layers, 1, cfg.n_kv_heads, T + new, cfg.head_dim)
    logits = O.stage_forward(ids, sw, cfg.kernel_cfg(), cache, np.zeros(T, np.int64), np.arange(T), cos, sin)
    # regression anchor: the oracle reproduces its own committed output (BLAS order may differ
    # between hosts in attention/lm_head: 1e-3 absolute on logits of magnitude ~40)
    assert np.abs(logits - g["s_logits"]).max() < 1e-3
    # Tier R: HF's fp32 Llama on the same un-quantised weights.  W8A8 error budget: cosine >= 0.999,
    # max-abs <= 2 % of max|logit|, same argmax everywhere (margins here are > 20).
    hf = g["hf_logits"].astype(np.float64)
    a, b = logits.ravel().astype(np.float64), hf.ravel()
    assert a @ b / (np.linalg.norm(a) * np.linalg.norm(b)) >= 0.999
    assert np.abs(a - b).max() <= 0.02 * np.abs(b).max()
    assert np.array_equal(logits.argmax(-1), hf.argmax(-1).reshape(-1))
    # the oracle is at least as close to fp32 HF as the reference's own CPU int8 path is
    c = g["hf_dyn_logits"].ravel().astype(np.float64)
    assert np.linalg.norm(a - b) <= np.linalg.norm(c - b) * 1.05
    toks, margins = O.greedy_generate([sw], cfg.kernel_cfg(), ids, new, T + new)
    assert np.array_equal(toks, g["s_greedy"]) and np.array_equal(toks, g["hf_greedy"])


@pytest.mark.parametrize("name", ["tiny-phi-test", "tiny-neox-test"])
def test_oracle_against_golden_parallel_families(name):
    """Phi (the reference's first answerer, combiner_quant_updated.py:177): HF PhiForCausalLM holds the weights in the
    model's layout -- 80-wide heads, the first 32 dims rotary, one LayerNorm feeding attention and MLP in parallel,
    biases everywhere, lm_head bias -- the oracle sees them after weights.to_kernel_layout (96-wide heads, rotary pairs
    moved to (j, j + 48), identity rope entries elsewhere).  Agreement with HF therefore pins the layout transform, the
    partial rope table, the 1/sqrt(80) scale and the parallel residual at once.
    GPT-NeoX / Pythia (the second answerer, :178): 256-wide heads with 64 rotary dims (permuted, not padded), two
    LayerNorms feeding the parallel branches, exact-erf GELU, HF's head-interleaved query_key_value layout."""
    g = np.load(os.path.join(GOLD, f"{name}.npz"))
    cfg = b200stage.get_config(name)
    seed, new, ids = int(g["seed"]), int(g["new"]), g["ids"]
    T = len(ids)
    sw = weights.to_numpy(weights.build_stage_weights(cfg, 0, cfg.n_layers, True, True, seed, "cpu"))
    assert sw["layers"][0]["wqkv"].shape == (cfg.qkv_out, cfg.d_model)
    assert (cfg.head_dim, cfg.hd_model) == ((96, 80) if name == "tiny-phi-test" else (256, 256))
    cos, sin = O.rope_table(cfg.head_dim, cfg.max_pos, cfg.rope_theta, cfg.rope_scaling, cfg.rotary_dim)
    r2 = cfg.rotary_dim // 2
    assert (cos[:, r2:] == 1).all() and (sin[:, r2:] == 0).all() and not (cos[5, :r2] == 1).all()
    cache = O.KVCacheOracle(cfg.n_layers, 1, cfg.n_kv_heads, T + new, cfg.head_dim)
    logits = O.stage_forward(ids, sw, cfg.kernel_cfg(), cache, np.zeros(T, np.int64), np.arange(T), cos, sin)
    assert np.abs(logits - g["s_logits"]).max() < 1e-3
    hf = g["hf_logits"].astype(np.float64)
    a, b = logits.ravel().astype(np.float64), hf.ravel()
    assert a @ b / (np.linalg.norm(a) * np.linalg.norm(b)) >= 0.999
    budget = 0.02 * np.abs(b).max()
    assert np.abs(a - b).max() <= budget
    top2 = np.sort(hf, axis=-1)[:, -2:]
    clear = (top2[:, 1] - top2[:, 0]) > 2 * budget
    assert clear.sum() >= 3 and np.array_equal(logits.argmax(-1)[clear], hf.argmax(-1)[clear])
    c = g["hf_dyn_logits"].ravel().astype(np.float64)
    assert np.linalg.norm(a - b) <= np.linalg.norm(c - b) * 1.05
    toks, _ = O.greedy_generate([sw], cfg.kernel_cfg(), ids, new, T + new)
    assert np.array_equal(toks, g["s_greedy"]) and np.array_equal(toks, g["hf_greedy"])


def test_head_layout_transform_is_exact_on_qk_products():
    """to_kernel_layout permutes q and k alike and pads with zeros: every head's q.k product and v are unchanged."""
    import torch
    cfg = b200stage.get_config("tiny-phi-test")
    src = weights.head_slots(cfg)
    assert sorted(src[src >= 0].tolist()) == list(range(80)) and (src < 0).sum() == 16
    assert list(src[:16]) == list(range(16)) and list(src[48:64]) == list(range(16, 32))      # rotary pairs (j, j + 48)
    wm = weights.init_layer_fp32(cfg, 0, 1, layout="model")
    wk = weights.to_kernel_layout(cfg, wm)
    x = torch.randn(5, cfg.d_model)
    H = cfg.n_heads
    qm = (x @ wm["wqkv"][: H * 80].t() + wm["b_qkv"][: H * 80]).view(5, H, 80)
    km = (x @ wm["wqkv"][H * 80: 2 * H * 80].t() + wm["b_qkv"][H * 80: 2 * H * 80]).view(5, H, 80)
    qk = (x @ wk["wqkv"][: H * 96].t() + wk["b_qkv"][: H * 96]).view(5, H, 96)
    kk = (x @ wk["wqkv"][H * 96: 2 * H * 96].t() + wk["b_qkv"][H * 96: 2 * H * 96]).view(5, H, 96)
    assert torch.allclose(torch.einsum("thd,shd->hts", qm, km), torch.einsum("thd,shd->hts", qk, kk), atol=1e-5)
    a = torch.randn(5, H * 80)
    ak = torch.zeros(5, H, 96)
    ak[:, :, torch.from_numpy(np.nonzero(src >= 0)[0])] = a.view(5, H, 80)[:, :, torch.from_numpy(src[src >= 0])]
    assert torch.allclose(a @ wm["wo"].t(), ak.reshape(5, -1) @ wk["wo"].t(), atol=1e-5)
    llama = b200stage.get_config("tiny-gqa-test")
    wl = weights.init_layer_fp32(llama, 0, 1, layout="model")
    assert weights.to_kernel_layout(llama, wl) is wl                                          # identity for full-rotary heads


def test_two_stage_split_equals_one_stage():
    from b200stage.pipeline import split_layers
    cfg = b200stage.get_config("tiny-llama-test")
    ids = np.random.default_rng(6).integers(0, cfg.vocab, 12)
    one = [weights.to_numpy(weights.build_stage_weights(cfg, 0, cfg.n_layers, True, True, 1, "cpu"))]
    two = [weights.to_numpy(weights.build_stage_weights(cfg, b, e, i == 0, i == 1, 1, "cpu"))
           for i, (b, e) in enumerate(split_layers(cfg.n_layers, 2))]
    t1, _ = O.greedy_generate(one, cfg.kernel_cfg(), ids, 4, 16)
    t2, _ = O.greedy_generate(two, cfg.kernel_cfg(), ids, 4, 16, chunk=5)
    assert np.array_equal(t1, t2)


# ---- sampler oracle vs HF's own logits processors (Tier R for the generate kwargs) -------------
@pytest.mark.parametrize("params", [(0.7, 50, 0.9, 1.1), (0.5, 30, 0.85, 1.2), (1.0, 5, 0.5, 1.0), (0.5, 30, 0.9, 1.1)])
def test_sample_filter_keeps_what_hf_keeps(params):
    """The set of tokens that survives the oracle's chain == the set HF's processors leave finite,
    run here in HF's order (generation/utils.py:1089,1214,1219,1223) with the reference's own
    parameter values (config.yaml:3-7, try.py:308-314)."""
    import torch
    from transformers.generation.logits_process import (RepetitionPenaltyLogitsProcessor, TemperatureLogitsWarper,
                                                        TopKLogitsWarper, TopPLogitsWarper)
    T, k, p, pen = params
    r = np.random.default_rng(int(T * 100) + k)
    for trial in range(10):
        V = 3000
        lg = (r.standard_normal(V) * 3).astype(np.float32)
        hist = r.integers(0, V, 80)
        ids, sc = O.sample_filter(lg, hist, T, k, p, pen)
        s, h = torch.from_numpy(lg)[None].clone(), torch.from_numpy(hist)[None]
        for proc in (RepetitionPenaltyLogitsProcessor(pen), TemperatureLogitsWarper(T), TopKLogitsWarper(k), TopPLogitsWarper(p)):
            s = proc(h, s)
        kept = torch.nonzero(torch.isfinite(s[0])).flatten().numpy()
        assert set(kept.tolist()) == set(ids.tolist())
        assert np.allclose(s[0, ids].numpy(), sc, rtol=1e-6)          # same processed scores
        assert (np.diff(sc) <= 0).all()                               # ordered by score, descending


def test_sample_det_is_a_valid_reproducible_draw():
    r = np.random.default_rng(3)
    lg = (r.standard_normal(500) * 2).astype(np.float32)
    ids, _ = O.sample_filter(lg, None, 0.7, 50, 0.9, 1.0)
    draws = [O.sample_det(lg, None, 0.7, 50, 0.9, 1.0, 11, s, 0) for s in range(200)]
    assert set(draws) <= set(ids.tolist()) and len(set(draws)) > 3
    assert draws == [O.sample_det(lg, None, 0.7, 50, 0.9, 1.0, 11, s, 0) for s in range(200)]
    assert O.sample_det(lg, None, 1.0, 1, 1.0, 1.0, 0, 0, 0) == int(lg.argmax())
    u = [float(O.rng_uniform(5, s, 0)) for s in range(2000)]
    assert 0.0 <= min(u) and max(u) < 1.0 and abs(np.mean(u) - 0.5) < 0.03


# ---- LLM.int8() mixed decomposition ------------------------------------------------------------
def test_llm_int8_linear_oracle_properties():
    """No outlier -> exactly the plain W8A8 linear; with outliers -> the outlier columns leave the
    int8 product (so the row scales shrink and the result gets CLOSER to the fp32 product)."""
    r = np.random.default_rng(21)
    T, N, K = 6, 96, 256
    W = (0.05 * r.standard_normal((N, K))).astype(np.float32)
    wq, ws = weights.quantize_weight(torch.from_numpy(W), torch.ones(K))
    wq, ws = wq.numpy(), ws.numpy()
    y = np.clip(r.standard_normal((T, K)), -4, 4).astype(f16)
    out, acc, cols = O.llm_int8_linear(y, wq, ws)
    q, s = O.smooth_quant(y, None)
    ref, acc_ref = O.gemm_w8a8(q, wq, s, ws, None, None)
    assert len(cols) == 0 and np.array_equal(acc, acc_ref) and np.array_equal(out.view(np.uint16), ref.view(np.uint16))
    y2 = y.copy()
    y2[1, 10], y2[4, 10], y2[0, 77] = 30.0, -12.0, 6.0
    out2, _, cols2 = O.llm_int8_linear(y2, wq, ws)
    assert cols2.tolist() == [10, 77]
    exact = y2.astype(np.float64) @ (wq.astype(np.float64) * ws[:, None].astype(np.float64)).T
    q2, s2 = O.smooth_quant(y2, None)
    plain, _ = O.gemm_w8a8(q2, wq, s2, ws, None, None)
    assert np.abs(out2.astype(np.float64) - exact).max() < np.abs(plain.astype(np.float64) - exact).max()
    res = r.standard_normal((T, N)).astype(f16)
    out3, _, _ = O.llm_int8_linear(y2, wq, ws, None, res)
    assert np.array_equal(out3, (out2.astype(f32) + res.astype(f32)).astype(f16))
