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
