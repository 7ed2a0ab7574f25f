"""b2_decode_attn_static: the decode-attention entry point for callers that cannot allocate.

Some users of this library have no way to give b2_decode_attn a workspace at call time: their
process closed its allocator after handing all of HBM to the KV-cache pool, or cudaMalloc (a
device-wide synchronisation with no time bound) is banned on their decode thread, or device
memory belongs to a plugin host.  For them the split-KV partials sit in a fixed 512 KiB array inside
the library.  This file checks, through the C ABI, that the variant

  * returns what b2_decode_attn returns with a caller workspace (same tolerance as the attention
    tests; the kernels are the same, so the bits are expected to be equal too),
  * still does so on a second call made straight after the first, although the shared array then
    holds the first call's partials (second call: other inputs, fewer keys, another split count),
  * refuses a problem whose partials exceed the array instead of shrinking it silently.

The file name sorts last on purpose: with `pytest -x` every other test has already run when this
one does, and it must pass both then and alone in a fresh process (it does its own set-up: the
first call on the device resolves the array's address).
"""
import ctypes as C

import numpy as np
import pytest
import torch

import b200stage
from b200stage import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"
STATIC_BYTES = 1 << 19     # B2_DECODE_ATTN_STATIC_SCRATCH_BYTES in include/b200stage.h


def _case(seed, B, H, KV, hd, S):
    g = torch.Generator(device=DEV).manual_seed(seed)
    Smax = S + 4
    kc = torch.randn((B, KV, Smax, hd), generator=g, device=DEV, dtype=torch.float16)
    vc = torch.randn((B, KV, Smax, hd), generator=g, device=DEV, dtype=torch.float16)
    qkv = torch.randn((B, (H + 2 * KV) * hd), generator=g, device=DEV, dtype=torch.float16)
    lens = torch.tensor([max(1, S - 13 * b) for b in range(B)], dtype=torch.int32, device=DEV)
    return qkv, kc, vc, lens, S


def _check(got, ref):
    # tolerance of tests/test_kernels_gpu.py::_attn_tol
    g = got.double().flatten()
    r = ref.double().flatten()
    assert float((g - r).abs().max()) <= 2e-3 * float(r.abs().max()) + 1e-3
    assert float(g @ r / (g.norm() * r.norm() + 1e-30)) >= 0.99999


def _both(seed, B, H, KV, hd, S):
    qkv, kc, vc, lens, S = _case(seed, B, H, KV, hd, S)
    scale = 1.0 / np.sqrt(hd)
    need = ops.decode_attn_workspace_bytes(B, H, KV, hd, S, 0)
    assert 0 < need <= STATIC_BYTES, "case must exercise the split path and fit the static array"
    ws = torch.empty(need, dtype=torch.uint8, device=DEV)
    ref = ops.decode_attn(qkv, kc, vc, lens, H, KV, hd, S, scale, workspace=ws)
    got = ops.decode_attn_static(qkv, kc, vc, lens, H, KV, hd, S, scale)
    torch.cuda.synchronize()
    _check(got, ref)
    assert torch.equal(got, ref), "same kernels, same inputs: expected identical bits"
    # and against the CPU oracle on this variant's own shapes, not only against its sibling entry point
    from oracle import w8a8 as O
    q_h, k_h, v_h = qkv.cpu().numpy(), kc.cpu().numpy(), vc.cpu().numpy()
    for b in range(B):
        L = int(lens[b])
        want = O.attention(q_h[b:b + 1, :H * hd].reshape(1, H, hd), np.transpose(k_h[b, :, :L], (1, 0, 2)),
                           np.transpose(v_h[b, :, :L], (1, 0, 2)), np.array([L - 1]), scale, 0)
        _check(got[b].cpu().reshape(H, hd), torch.from_numpy(want[0].astype(np.float32)))
    return need


def test_static_scratch_matches_workspace_variant_twice_in_a_row():
    n1 = _both(1, B=2, H=32, KV=32, hd=128, S=3000)      # many splits, large partials
    n2 = _both(2, B=1, H=32, KV=8, hd=128, S=700)        # straight after: fewer keys, other split count
    assert n1 != n2
    _both(3, B=3, H=8, KV=2, hd=64, S=1100)              # third shape, head_dim 64


def test_static_scratch_rejects_oversize():
    # one long sequence, 4 kv heads of 8 query heads each: 46 splits x 32 heads of fp32 partials (765 KB)
    B, H, KV, hd, S = 1, 32, 4, 128, 20000
    L = b200stage._abi.lib()
    need = L.b2_decode_attn_workspace_bytes(B, H, KV, hd, S, 0)
    assert need > STATIC_BYTES, need
    # the size check comes before any dereference, so small dummy buffers are enough
    dummy = torch.zeros(64, dtype=torch.float16, device=DEV)
    lens = torch.ones(1, dtype=torch.int32, device=DEV)
    vp = lambda t: C.c_void_p(t.data_ptr())
    rc = L.b2_decode_attn_static(vp(dummy), 2, vp(dummy), vp(dummy), vp(lens), None, vp(dummy), 2, B, H, KV, hd, S + 4, S,
                                 0.1, 0, None)
    assert rc == -8 and b"exceeds" in L.b2_strerror(rc)   # B2_ERR_TOO_LARGE
    torch.cuda.synchronize()
