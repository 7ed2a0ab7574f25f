"""LLM.int8() mixed decomposition (b2_outlier_scan / b2_smooth_quant / b2_gemm_w8a8 / b2_outlier_side)
against oracle/w8a8.py llm_int8_linear -- the restatement of what the reference's
`BitsAndBytesConfig(load_in_8bit=True)` (Code/Quantised Models/models_quant_updated.py:30-32,
bitsandbytes Linear8bitLt, threshold 6.0) does to every nn.Linear.  Integer parts (outlier column
list, int8 activations, row scales) and the fp16 result are bit-exact."""
import numpy as np
import pytest
import torch

import b200stage  # noqa: F401
from b200stage import ops, weights
from oracle import w8a8 as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def _weights(r, N, K):
    W = (0.05 * r.standard_normal((N, K))).astype(np.float32)
    q, s = weights.quantize_weight(torch.from_numpy(W), torch.ones(K))
    return q.numpy(), s.numpy()


def _acts(r, T, K, n_outlier_cols, rows_with_outliers=None):
    y = r.standard_normal((T, K)).astype(np.float32)
    y = np.clip(y, -5.5, 5.5)
    cols = np.sort(r.choice(K, n_outlier_cols, replace=False)) if n_outlier_cols else np.zeros((0,), int)
    for c in cols:
        rows = r.choice(T, max(1, T // 3), replace=False) if rows_with_outliers is None else rows_with_outliers
        y[rows, c] = r.choice([-1.0, 1.0], len(rows)) * r.uniform(6.0, 40.0, len(rows))
    return y.astype(np.float16), cols


@pytest.mark.parametrize("T,N,K,nout,bias,res", [
    (1, 256, 512, 0, False, False), (1, 256, 512, 3, True, True), (5, 384, 1024, 7, False, True),
    (8, 512, 2048, 1, True, False), (70, 768, 512, 12, True, True), (200, 1024, 4096, 40, False, True),
    (33, 256, 11008, 5, False, False)])
def test_llm_int8_linear_bit_exact(T, N, K, nout, bias, res):
    r = np.random.default_rng(T * 7 + N + K + nout)
    wq, ws = _weights(r, N, K)
    y16, cols = _acts(r, T, K, nout)
    b16 = (0.1 * r.standard_normal(N)).astype(np.float16) if bias else None
    r16 = r.standard_normal((T, N)).astype(np.float16) if res else None
    want, acc, ocols = O.llm_int8_linear(y16, wq, ws, b16, r16)
    assert np.array_equal(ocols, cols)
    sc = ops.OutlierScratch(K, DEV)
    q = torch.empty((T, K), dtype=torch.int8, device=DEV)
    rs = torch.empty((T,), dtype=torch.float32, device=DEV)
    got = ops.llm_int8_linear(t(y16), t(wq), t(ws), sc, bias=None if b16 is None else t(b16),
                              residual=None if r16 is None else t(r16), q_buf=q, scale_buf=rs)
    n = int(sc.count.item())
    assert n == len(cols) and np.array_equal(sc.cols[:n].cpu().numpy(), cols)
    mask = np.ones(K, np.float32)
    mask[cols] = 0
    assert np.array_equal(sc.mask.cpu().numpy(), mask)
    q_ref, s_ref = O.smooth_quant(y16, mask)
    assert np.array_equal(q.cpu().numpy(), q_ref) and np.array_equal(rs.cpu().numpy(), s_ref)
    assert np.array_equal(got.cpu().numpy().view(np.uint16), want.view(np.uint16))


@pytest.mark.parametrize("T,N,K,nout,bias,res", [(1, 512, 4096, 4, False, True), (3, 256, 1024, 0, True, False),
                                                  (8, 768, 2048, 9, True, True), (5, 1024, 11008, 2, False, True)])
def test_llm_int8_linear_decode_path_bit_exact(T, N, K, nout, bias, res):
    """Decode batches: one-launch scan + fused masked-quantise/GEMV kernel: the same bits as the oracle."""
    r = np.random.default_rng(T * 3 + N + K + nout)
    wq, ws = _weights(r, N, K)
    y16, cols = _acts(r, T, K, nout)
    b16 = (0.1 * r.standard_normal(N)).astype(np.float16) if bias else None
    r16 = r.standard_normal((T, N)).astype(np.float16) if res else None
    want, _, ocols = O.llm_int8_linear(y16, wq, ws, b16, r16)
    sc = ops.OutlierScratch(K, DEV)
    got = ops.llm_int8_linear(t(y16), t(wq), t(ws), sc, bias=None if b16 is None else t(b16),
                              residual=None if r16 is None else t(r16), fused_decode=True)
    n = int(sc.count.item())
    assert n == len(ocols) and np.array_equal(sc.cols[:n].cpu().numpy(), ocols)
    assert np.array_equal(got.cpu().numpy().view(np.uint16), want.view(np.uint16))


def test_threshold_is_inclusive_and_scratch_is_reusable():
    r = np.random.default_rng(3)
    K, N = 256, 128
    wq, ws = _weights(r, N, K)
    y = np.clip(r.standard_normal((4, K)), -3, 3).astype(np.float16)
    y[2, 17] = np.float16(6.0)            # |y| == threshold counts as an outlier (>=)
    y[1, 200] = np.float16(-5.996)        # just below: not
    sc = ops.OutlierScratch(K, DEV)
    ops.outlier_scan(t(y), sc)
    assert int(sc.count.item()) == 1 and int(sc.cols[0].item()) == 17
    y2 = np.clip(r.standard_normal((4, K)), -3, 3).astype(np.float16)      # second use: no outliers left over
    got = ops.llm_int8_linear(t(y2), t(wq), t(ws), sc)
    assert int(sc.count.item()) == 0
    want, _, _ = O.llm_int8_linear(y2, wq, ws)
    assert np.array_equal(got.cpu().numpy().view(np.uint16), want.view(np.uint16))


@pytest.mark.parametrize("act,gated", [(0, True), (1, True), (2, False), (3, False)])
def test_act_mul_f16_matches_oracle(act, gated):
    r = np.random.default_rng(act)
    T, F = 9, 1024
    x = (2 * r.standard_normal((T, 2 * F if gated else F))).astype(np.float16)
    got = ops.act_mul(t(x), F, act).cpu().numpy()
    if gated:
        want, _ = O.act_mul_y(x, None, "silu" if act == 0 else "gelu_new")
    else:
        xf = x.astype(np.float32)
        want = O.gelu_new_det(xf) if act == 2 else O.silu_det(xf)
    assert np.array_equal(got.view(np.uint16), want.astype(np.float16).view(np.uint16))


# ---- the whole stage in LLM.int8() mode ----------------------------------------------------------
@pytest.mark.parametrize("model,n_stages", [("tiny-llama-test", 1), ("tiny-gqa-test", 2), ("tiny-gpt2-test", 1)])
def test_stage_llm_int8_matches_oracle(model, n_stages):
    from b200stage.pipeline import Pipeline, split_layers
    cfg = b200stage.get_config(model)
    seed, T, new = 2, 40, 10
    ids = np.random.default_rng(8).integers(0, cfg.vocab, T)
    osw = [weights.to_numpy(weights.build_stage_weights(cfg, b, e, i == 0, i == n_stages - 1, seed, "cpu", quant="llm_int8"))
           for i, (b, e) in enumerate(split_layers(cfg.n_layers, n_stages))]
    kc = cfg.kernel_cfg()
    # one-stage trace of the oracle: outliers really occur, and where
    one = weights.to_numpy(weights.build_stage_weights(cfg, 0, cfg.n_layers, True, True, seed, "cpu", quant="llm_int8"))
    cos, sin = O.rope_table(cfg.head_dim, cfg.max_pos, cfg.rope_theta, cfg.rope_scaling)
    cache = O.KVCacheOracle(cfg.n_layers, 1, cfg.n_kv_heads, T + new, cfg.head_dim)
    otrace = []
    ref = O.stage_forward(ids, one, kc, cache, np.zeros(T, np.int64), np.arange(T), cos, sin, otrace)
    assert all(len(tr["cols"][0]) > 0 for tr in otrace)
    pipe = Pipeline.load(cfg, n_stages, seed=seed, quant="llm_int8")
    if n_stages == 1:
        st = pipe.stages[0]
        trace = []
        got = st.forward(torch.from_numpy(ids).to(DEV), st.new_kv_cache(1, T + new), np.arange(T), logits="all", trace=trace)
        # layer 0 up to the attention input: identical inputs -> bit-exact
        assert np.array_equal(trace[0]["qkv"].cpu().numpy().view(np.uint16), otrace[0]["qkv"].view(np.uint16))
        for li in range(cfg.n_layers):
            assert trace[li]["n_out"][0] == len(otrace[li]["cols"][2])           # scratch of width d: last used by gate_up
            a = trace[li]["h2"].cpu().numpy().astype(np.float64).ravel()
            b = otrace[li]["h2"].astype(np.float64).ravel()
            assert np.abs(a - b).max() <= 2.0 ** -5 * np.abs(b).max()
            assert a @ b / (np.linalg.norm(a) * np.linalg.norm(b)) >= 0.9995
        g = got.cpu().numpy()
        assert np.abs(g - ref).max() <= 0.02 * np.abs(ref).max()
        assert np.array_equal(g.argmax(-1), ref.argmax(-1))
    toks, margins = O.greedy_generate(osw, kc, ids, new, T + new)
    out = pipe.generate(torch.from_numpy(ids).to(DEV), max_new_tokens=new, chunk=16).cpu().numpy()
    assert np.array_equal(out, toks), (out[T:], toks[T:], min(margins))


@pytest.mark.parametrize("T,N,K,nout,res", [(96, 512, 1024, 700, True), (300, 768, 2048, 40, False), (130, 256, 512, 0, True)])
def test_dense_side_product_matches_oracle_within_tolerance(T, N, K, nout, res):
    """The tensor-core side product (long batches: hundreds of outlier columns) against the oracle: same
    mask / int8 part bit for bit, fp16 result within fp32-summation-order tolerance of the side term."""
    r = np.random.default_rng(T + N + nout)
    wq, ws = _weights(r, N, K)
    y16, cols = _acts(r, T, K, nout)
    r16 = r.standard_normal((T, N)).astype(np.float16) if res else None
    want, _, ocols = O.llm_int8_linear(y16, wq, ws, None, r16)
    assert len(ocols) == nout
    sc = ops.OutlierScratch(K, DEV)
    dense = ops.DenseSideScratch(T, K, N, N * K, DEV)
    got = ops.llm_int8_linear(t(y16), t(wq), t(ws), sc, residual=None if r16 is None else t(r16), dense=dense).cpu().numpy()
    assert int(sc.count.item()) == nout
    a, b = got.astype(np.float64), want.astype(np.float64)
    assert np.abs(a - b).max() <= 2e-3 * np.abs(b).max() + 2e-3, np.abs(a - b).max()
    if nout == 0:
        assert np.array_equal(got.view(np.uint16), want.view(np.uint16))       # zero side term: exact
