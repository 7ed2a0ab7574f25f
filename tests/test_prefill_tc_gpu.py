"""b2_prefill_attn on the tcgen05/TMEM kernel (csrc/attn_tc.cu, head_dim 128, causal + optional sliding window): against the
oracle (oracle/w8a8.py attention, the statement of transformers modeling_llama.py:199-222 for
q_len > 1) within the attention tolerance, and against the mma.sync kernel it replaces on the same
inputs -- chunk lengths and start positions that are not multiples of the 128-row / 128-key tiles,
GQA, a second cache row, device-resident positions, and the Llama-2-7B head count at T = 1024."""
import numpy as np
import pytest
import torch

import b200stage  # noqa: F401
from b200stage import ops
from oracle import w8a8 as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _tol(got, ref, rel=2e-3):
    got = got.astype(np.float64).ravel()
    ref = ref.astype(np.float64).ravel()
    err = np.abs(got - ref).max()
    cos = float(got @ ref / (np.linalg.norm(got) * np.linalg.norm(ref) + 1e-30))
    assert err <= rel * np.abs(ref).max() + 1e-3, err
    assert cos >= 0.99999, cos


@pytest.fixture(autouse=True)
def _restore_impl():
    yield
    ops.set_prefill_attn_impl(0)


def _case(H, KV, T, pos0, row, rows, seed, slack=5):
    r = np.random.default_rng(seed)
    hd = 128
    S = pos0 + T
    Smax = S + slack
    kc = r.standard_normal((rows, KV, Smax, hd)).astype(np.float16)
    vc = r.standard_normal((rows, KV, Smax, hd)).astype(np.float16)
    qkv = r.standard_normal((T, (H + 2 * KV) * hd)).astype(np.float16)
    return kc, vc, qkv, S, 1.0 / np.sqrt(hd)


@pytest.mark.parametrize("H,KV,T,pos0,row,rows", [
    (2, 2, 64, 0, 0, 1), (4, 4, 128, 0, 0, 1), (4, 2, 129, 0, 1, 2), (8, 2, 300, 64, 1, 3), (2, 1, 200, 1000, 0, 1),
    (4, 4, 512, 128, 0, 1), (3, 1, 257, 333, 2, 3)])
def test_tc_matches_oracle_and_mma_kernel(H, KV, T, pos0, row, rows):
    kc, vc, qkv, S, scale = _case(H, KV, T, pos0, row, rows, H * 1000 + T + pos0)
    dq, dk, dv = (torch.from_numpy(x).to(DEV) for x in (qkv, kc, vc))
    ops.set_prefill_attn_impl(2)
    got = ops.prefill_attn(dq, dk, dv, pos0, row, H, KV, 128, scale, 0).cpu().numpy()
    ops.set_prefill_attn_impl(1)
    old = ops.prefill_attn(dq, dk, dv, pos0, row, H, KV, 128, scale, 0).cpu().numpy()
    ref = O.attention(qkv[:, :H * 128].reshape(T, H, 128), np.transpose(kc[row, :, :S], (1, 0, 2)),
                      np.transpose(vc[row, :, :S], (1, 0, 2)), np.arange(pos0, S), scale, 0)
    _tol(got.reshape(T, H, 128), ref)
    _tol(got, old, rel=1e-3)        # the two kernels differ only in summation order / P rounding point


@pytest.mark.parametrize("window", [16, 100, 128, 129, 300, 4096])
@pytest.mark.parametrize("H,KV,T,pos0,row,rows", [(4, 2, 300, 64, 1, 2), (2, 1, 200, 1000, 0, 1), (3, 1, 257, 333, 2, 3),
                                                  (2, 2, 256, 4200, 0, 1), (2, 2, 64, 0, 0, 1)])
def test_tc_sliding_window_matches_oracle_and_mma_kernel(H, KV, T, pos0, row, rows, window):
    """Mistral's attention (key > query - window): windows shorter than, equal to, one more than and longer than a
    128-key tile, starting inside the chunk and far before it; whole key tiles before the window are skipped."""
    kc, vc, qkv, S, scale = _case(H, KV, T, pos0, row, rows, H * 1000 + T + pos0 + window)
    dq, dk, dv = (torch.from_numpy(x).to(DEV) for x in (qkv, kc, vc))
    ops.set_prefill_attn_impl(2)
    got = ops.prefill_attn(dq, dk, dv, pos0, row, H, KV, 128, scale, window).cpu().numpy()
    ops.set_prefill_attn_impl(1)
    old = ops.prefill_attn(dq, dk, dv, pos0, row, H, KV, 128, scale, window).cpu().numpy()
    ref = O.attention(qkv[:, :H * 128].reshape(T, H, 128), np.transpose(kc[row, :, :S], (1, 0, 2)),
                      np.transpose(vc[row, :, :S], (1, 0, 2)), np.arange(pos0, S), scale, window)
    _tol(got.reshape(T, H, 128), ref)
    _tol(got, old, rel=1e-3)


def test_tc_device_positions_and_full_cache_row():
    """pos0 read from device memory (graph-replayable chunks); the chunk ends exactly at Smax, so the
    last key tile runs past the cache row and must come back zero-filled, not from the next row."""
    H, KV, T, pos0 = 4, 2, 150, 448
    kc, vc, qkv, S, scale = _case(H, KV, T, pos0, 1, 2, 77, slack=0)
    dq, dk, dv = (torch.from_numpy(x).to(DEV) for x in (qkv, kc, vc))
    pos = torch.arange(pos0, pos0 + T, dtype=torch.int32, device=DEV)
    ops.set_prefill_attn_impl(2)
    got = ops.prefill_attn(dq, dk, dv, 0, 1, H, KV, 128, scale, 0, positions=pos).cpu().numpy()
    ref = O.attention(qkv[:, :H * 128].reshape(T, H, 128), np.transpose(kc[1, :, :S], (1, 0, 2)),
                      np.transpose(vc[1, :, :S], (1, 0, 2)), np.arange(pos0, S), scale, 0)
    _tol(got.reshape(T, H, 128), ref)


def test_tc_llama7b_heads_vs_mma_kernel():
    H, KV, T, pos0 = 32, 32, 1024, 512
    kc, vc, qkv, S, scale = _case(H, KV, T, pos0, 0, 1, 5)
    dq, dk, dv = (torch.from_numpy(x).to(DEV) for x in (qkv, kc, vc))
    ops.set_prefill_attn_impl(2)
    got = ops.prefill_attn(dq, dk, dv, pos0, 0, H, KV, 128, scale, 0)
    ops.set_prefill_attn_impl(1)
    old = ops.prefill_attn(dq, dk, dv, pos0, 0, H, KV, 128, scale, 0)
    _tol(got.cpu().numpy(), old.cpu().numpy(), rel=1e-3)


def test_unsupported_shapes_fall_back_or_refuse():
    kc, vc, qkv, S, scale = _case(2, 2, 40, 0, 0, 1, 9)
    dq, dk, dv = (torch.from_numpy(x).to(DEV) for x in (qkv, kc, vc))
    ops.set_prefill_attn_impl(2)
    with pytest.raises(b200stage._abi.B2Error):
        ops.prefill_attn(dq, dk, dv, 0, 0, 2, 2, 128, scale, 16)       # 40 query rows: mma.sync kernel only
    ops.set_prefill_attn_impl(0)
    out = ops.prefill_attn(dq, dk, dv, 0, 0, 2, 2, 128, scale, 16)     # auto: falls back
    ref = O.attention(qkv[:, :256].reshape(40, 2, 128), np.transpose(kc[0, :, :S], (1, 0, 2)),
                      np.transpose(vc[0, :, :S], (1, 0, 2)), np.arange(0, S), scale, 16)
    _tol(out.cpu().numpy().reshape(40, 2, 128), ref)
