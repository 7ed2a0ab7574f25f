"""Torch-tensor front ends of the C-ABI entry points (one function per entry point).

Each wrapper only validates dtypes/devices, allocates the output with torch (device memory is
torch's job here) and forwards raw pointers.  All compute happens in libb200stage.so.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _abi
from ._abi import check, ptr, stream_handle

ACT_SILU_MUL, ACT_GELU_MUL, ACT_GELU, ACT_SILU, ACT_GELU_ERF = 0, 1, 2, 3, 4


def _cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise ValueError("b200stage ops need CUDA tensors (there is no CPU path)")
        if t is not None and not t.is_contiguous():
            raise ValueError("b200stage ops need contiguous tensors")


def norm_quant(x, gamma, beta=None, smooth=None, eps=1e-5, out_q=None, out_scale=None):
    """x [T,d] fp16 -> (q [T,d] int8, scale [T] fp32).  b2_norm_quant."""
    _cuda(x, gamma, beta, smooth)
    T, d = x.shape
    q = out_q if out_q is not None else torch.empty((T, d), dtype=torch.int8, device=x.device)
    s = out_scale if out_scale is not None else torch.empty((T,), dtype=torch.float32, device=x.device)
    check(_abi.lib().b2_norm_quant(ptr(x), ptr(gamma), ptr(beta), ptr(smooth), ptr(q), ptr(s), T, d, float(eps),
                                   stream_handle()), "b2_norm_quant")
    return q, s


def norm(x, gamma, beta=None, eps=1e-5, out=None):
    """x [T,d] fp16 -> normalised fp16 [T,d].  b2_norm."""
    _cuda(x, gamma, beta)
    T, d = x.shape
    y = out if out is not None else torch.empty_like(x)
    check(_abi.lib().b2_norm(ptr(x), ptr(gamma), ptr(beta), ptr(y), T, d, float(eps), stream_handle()), "b2_norm")
    return y


def smooth_quant(x, smooth=None, out_q=None, out_scale=None):
    """x [T,K] fp16 -> (q int8, scale fp32).  b2_smooth_quant."""
    _cuda(x, smooth)
    T, K = x.shape
    q = out_q if out_q is not None else torch.empty((T, K), dtype=torch.int8, device=x.device)
    s = out_scale if out_scale is not None else torch.empty((T,), dtype=torch.float32, device=x.device)
    check(_abi.lib().b2_smooth_quant(ptr(x), ptr(smooth), ptr(q), ptr(s), T, K, stream_handle()), "b2_smooth_quant")
    return q, s


def act_mul_quant(gate_up, smooth=None, act=ACT_SILU_MUL, out_q=None, out_scale=None):
    """gate_up [T,2F] (gated) or [T,F] (ungated) fp16 -> (q [T,F] int8, scale).  b2_act_mul_quant."""
    _cuda(gate_up, smooth)
    T = gate_up.shape[0]
    F = gate_up.shape[1] // 2 if act in (ACT_SILU_MUL, ACT_GELU_MUL) else gate_up.shape[1]
    q = out_q if out_q is not None else torch.empty((T, F), dtype=torch.int8, device=gate_up.device)
    s = out_scale if out_scale is not None else torch.empty((T,), dtype=torch.float32, device=gate_up.device)
    check(_abi.lib().b2_act_mul_quant(ptr(gate_up), ptr(smooth), ptr(q), ptr(s), T, F, int(act), stream_handle()),
          "b2_act_mul_quant")
    return q, s


def gemm_w8a8(a_q, w_q, a_scale, w_scale, bias=None, residual=None, out=None, acc_dump=None, schedule=0):
    """a_q [M,K] int8, w_q [N,K] int8 -> out [M,N] fp16.  b2_gemm_w8a8 (schedule 0) / _sched."""
    _cuda(a_q, w_q, a_scale, w_scale, bias, residual, out, acc_dump)
    M, K = a_q.shape
    N = w_q.shape[0]
    if w_q.shape[1] != K:
        raise ValueError("K mismatch")
    o = out if out is not None else torch.empty((M, N), dtype=torch.float16, device=a_q.device)
    L = _abi.lib()
    if schedule == 0:
        rc = L.b2_gemm_w8a8(ptr(a_q), ptr(w_q), ptr(a_scale), ptr(w_scale), ptr(bias), ptr(residual), ptr(o),
                            ptr(acc_dump), M, N, K, stream_handle())
    else:
        rc = L.b2_gemm_w8a8_sched(ptr(a_q), ptr(w_q), ptr(a_scale), ptr(w_scale), ptr(bias), ptr(residual), ptr(o),
                                  ptr(acc_dump), M, N, K, int(schedule), stream_handle())
    check(rc, "b2_gemm_w8a8")
    return o


def gemm_gated_act(a_q, w_q, a_scale, w_scale, *, y32, amax, bias=None, out_smooth=None, act=0):
    """a_q [M,K] int8, w_q [2F,K] = [gate; up] -> y32 [M,F] fp32 = act(gate)*up*out_smooth, amax [M].  b2_gemm_gated_act."""
    _cuda(a_q, w_q, a_scale, w_scale, bias, out_smooth, y32, amax)
    M, K = a_q.shape
    F = w_q.shape[0] // 2
    check(_abi.lib().b2_gemm_gated_act(ptr(a_q), ptr(w_q), ptr(a_scale), ptr(w_scale), ptr(bias), ptr(out_smooth), ptr(y32),
                                       ptr(amax), M, F, K, int(act), stream_handle()), "b2_gemm_gated_act")
    return y32, amax


def quant_f32(y32, amax, out_q=None, out_scale=None):
    """y32 [T,F] fp32 + row absmax -> (q [T,F] int8, scale [T]).  b2_quant_f32."""
    _cuda(y32, amax, out_q, out_scale)
    T, F = y32.shape
    q = out_q if out_q is not None else torch.empty((T, F), dtype=torch.int8, device=y32.device)
    s = out_scale if out_scale is not None else torch.empty((T,), dtype=torch.float32, device=y32.device)
    check(_abi.lib().b2_quant_f32(ptr(y32), ptr(amax), ptr(q), ptr(s), T, F, stream_handle()), "b2_quant_f32")
    return q, s


def gemm_qkv_rope(a_q, w_q, a_scale, w_scale, qkv_out, cos, sin, positions, kcache, vcache, H, KV, hd, *, bias=None,
                  seq_ids=None, seq0=0, use_rope=True):
    """Prefill qkv projection + RoPE + KV append in one launch.  b2_gemm_qkv_rope."""
    _cuda(a_q, w_q, a_scale, w_scale, qkv_out, cos, sin, positions, kcache, vcache, bias, seq_ids)
    M, K = a_q.shape
    Smax = kcache.shape[2]
    check(_abi.lib().b2_gemm_qkv_rope(ptr(a_q), ptr(w_q), ptr(a_scale), ptr(w_scale), ptr(bias), ptr(qkv_out),
                                      qkv_out.stride(0), ptr(cos), ptr(sin), ptr(positions), ptr(seq_ids), int(seq0),
                                      ptr(kcache), ptr(vcache), M, H, KV, hd, Smax, K, int(bool(use_rope)),
                                      stream_handle()), "b2_gemm_qkv_rope")
    return qkv_out


X_RMS, X_LN, X_PLAIN, X_Y32 = 0, 1, 2, 3
O_F16, O_GATED, O_QKV = 0, 1, 2


PREFETCH_BYTES = 24 << 20      # head of a weight matrix worth pulling into L2 ahead of its kernel


def _pf(prefetch, i):
    """(pointer, bytes) of the i-th prefetch hint: a tensor (its first PREFETCH_BYTES), a
    (tensor, byte_offset, bytes) slice of one, or None."""
    if prefetch is not None and i < len(prefetch) and isinstance(prefetch[i], tuple):
        t, off, nb = prefetch[i]
        total = t.numel() * t.element_size()
        off = min(int(off), total) & ~15
        nb = max(0, min(int(nb), total - off)) & ~15
        return (C.c_void_p(t.data_ptr() + off), nb) if nb else (None, 0)
    if prefetch is None or i >= len(prefetch) or prefetch[i] is None:
        return None, 0
    t = prefetch[i]
    return ptr(t), min(t.numel() * t.element_size(), PREFETCH_BYTES)


def fused_linear(x, w_q, w_scale, *, x_kind, out, gamma=None, beta=None, smooth=None, x_amax=None, eps=1e-5, bias=None,
                 residual=None, out_kind=O_F16, act=0, out_smooth=None, out_amax=None, zero_amax=None, prefetch=None):
    """Decode-time (M <= 8) prologue + quantise + W8A8 GEMV + epilogue in one launch.  b2_fused_linear.
    x: fp16 [M,K] (x_kind 0..2) or fp32 [M,K] (x_kind 3); out: fp16 [M,N] or fp32 [M,N/2] (gated).
    prefetch: up to two weight tensors of later launches whose heads are pulled into L2."""
    _cuda(w_q, w_scale, gamma, beta, smooth, x_amax, bias, residual, out, out_smooth, out_amax, zero_amax)
    if not x.is_cuda or x.stride(-1) != 1:
        raise ValueError("x must be a CUDA tensor with unit inner stride")
    M = x.shape[0]
    N, K = w_q.shape
    p0, n0 = _pf(prefetch, 0)
    p1, n1 = _pf(prefetch, 1)
    check(_abi.lib().b2_fused_linear(ptr(x), x.stride(0), int(x_kind), ptr(gamma), ptr(beta), ptr(smooth), ptr(x_amax),
                                     float(eps), ptr(w_q), ptr(w_scale), ptr(bias), ptr(residual), ptr(out), int(out_kind),
                                     int(act), ptr(out_smooth), ptr(out_amax), ptr(zero_amax), M, N, K, p0, n0, p1, n1,
                                     stream_handle()),
          "b2_fused_linear")
    return out


def fused_qkv(x, w_q, w_scale, qkv_out, cos, sin, positions, kcache, vcache, H, KV, hd, *, x_kind, gamma=None,
              beta=None, smooth=None, eps=1e-5, bias=None, seq_ids=None, seq0=0, use_rope=True, prefetch=None):
    """Decode-time qkv projection + RoPE + KV append in one launch (M <= 8).  b2_fused_qkv."""
    _cuda(x, w_q, w_scale, qkv_out, cos, sin, positions, kcache, vcache, gamma, beta, smooth, bias, seq_ids)
    M, K = x.shape
    Smax = kcache.shape[2]
    p0, n0 = _pf(prefetch, 0)
    p1, n1 = _pf(prefetch, 1)
    check(_abi.lib().b2_fused_qkv(ptr(x), x.stride(0), int(x_kind), ptr(gamma), ptr(beta), ptr(smooth), float(eps),
                                  ptr(w_q), ptr(w_scale), ptr(bias), ptr(qkv_out), ptr(cos), ptr(sin), ptr(positions),
                                  ptr(seq_ids), int(seq0), ptr(kcache), ptr(vcache), M, H, KV, hd, Smax, K,
                                  int(bool(use_rope)), p0, n0, p1, n1, stream_handle()), "b2_fused_qkv")
    return qkv_out


def fused_desc(x, w_q, w_scale, *, x_kind, out, out_kind=O_F16, gamma=None, beta=None, smooth=None, x_amax=None,
               eps=1e-5, bias=None, residual=None, act=0, out_smooth=None, out_amax=None, zero_amax=None,
               rope=None, prefetch=None):
    """b2_fused_desc for one phase of fused_layer_tail.  `rope` (out_kind 2 only) =
    dict(cos, sin, positions, seq_ids, seq0, kcache, vcache, H, KV, hd, use_rope).
    Returns (struct, keep-alive list)."""
    d = _abi.FusedDesc()
    keep = [x, w_q, w_scale, gamma, beta, smooth, x_amax, bias, residual, out, out_smooth, out_amax, zero_amax, prefetch]
    pv = lambda t: None if t is None else t.data_ptr()
    d.x, d.x_ld, d.x_kind = pv(x), x.stride(0), int(x_kind)
    d.gamma, d.beta, d.smooth, d.x_amax, d.eps = pv(gamma), pv(beta), pv(smooth), pv(x_amax), float(eps)
    d.W, d.w_scale, d.bias, d.residual = pv(w_q), pv(w_scale), pv(bias), pv(residual)
    d.out, d.out_kind, d.act = pv(out), int(out_kind), int(act)
    d.out_smooth, d.out_amax, d.zero_amax = pv(out_smooth), pv(out_amax), pv(zero_amax)
    d.N, d.K = int(w_q.shape[0]), int(w_q.shape[1])
    if rope is not None:
        keep.append(rope)
        d.cos_table, d.sin_table, d.positions = pv(rope["cos"]), pv(rope["sin"]), pv(rope["positions"])
        d.seq_ids, d.seq0 = pv(rope.get("seq_ids")), int(rope.get("seq0", 0))
        d.kcache, d.vcache = pv(rope["kcache"]), pv(rope["vcache"])
        d.H, d.KV, d.hd, d.Smax = int(rope["H"]), int(rope["KV"]), int(rope["hd"]), int(rope["kcache"].shape[2])
        d.use_rope = int(bool(rope.get("use_rope", True)))
    for i in (0, 1):
        p_, n_ = _pf(prefetch, i)
        setattr(d, f"prefetch{i}", None if p_ is None else p_.value)
        setattr(d, f"prefetch{i}_bytes", n_)
    return d, keep


def fused_layer_tail(o_proj, gate_up, down, next_qkv, M, counters):
    """o_proj -> gate_up -> down [-> next layer's qkv] as one persistent launch (M <= 8).
    Each argument is the (struct, keep) pair of fused_desc (next_qkv may be None); `counters` is a
    zeroed uint32[4] device tensor.  b2_fused_layer_tail."""
    import ctypes as C
    _cuda(counters)
    ref = lambda pair: None if pair is None else C.byref(pair[0])
    check(_abi.lib().b2_fused_layer_tail(ref(o_proj), ref(gate_up), ref(down), ref(next_qkv), int(M), ptr(counters),
                                         stream_handle()), "b2_fused_layer_tail")


def rope_kv_append(qkv, cos, sin, positions, kcache, vcache, H, KV, hd, seq_ids=None, seq0=0, use_rope=True):
    """qkv [T,(H+2KV)*hd] fp16 (q rotated in place); caches [rows,KV,Smax,hd].  b2_rope_kv_append."""
    _cuda(qkv, cos, sin, positions, kcache, vcache, seq_ids)
    T = qkv.shape[0]
    Smax = kcache.shape[2]
    check(_abi.lib().b2_rope_kv_append(ptr(qkv), qkv.stride(0), ptr(cos), ptr(sin), ptr(positions), ptr(seq_ids),
                                       int(seq0), ptr(kcache), ptr(vcache), T, H, KV, hd, Smax, int(bool(use_rope)),
                                       stream_handle()), "b2_rope_kv_append")


def decode_attn_workspace_bytes(B, H, KV, hd, max_seqlen, window=0):
    return int(_abi.lib().b2_decode_attn_workspace_bytes(B, H, KV, hd, int(max_seqlen), int(window)))


def decode_attn(qkv, kcache, vcache, seqlens, H, KV, hd, max_seqlen, scale, window=0, rows=None, out=None,
                workspace=None, prefetch=None):
    """One query token per sequence; q read in place from the qkv buffer.  b2_decode_attn.
    prefetch: up to two L2 hints (see _pf) for weights that later launches stream."""
    _cuda(qkv, kcache, vcache, seqlens, rows, out, workspace)
    B = qkv.shape[0]
    Smax = kcache.shape[2]
    o = out if out is not None else torch.empty((B, H * hd), dtype=torch.float16, device=qkv.device)
    wsb = 0 if workspace is None else workspace.numel() * workspace.element_size()
    p0, n0 = _pf(prefetch, 0)
    p1, n1 = _pf(prefetch, 1)
    check(_abi.lib().b2_decode_attn(ptr(qkv), qkv.stride(0), ptr(kcache), ptr(vcache), ptr(seqlens), ptr(rows), ptr(o),
                                    o.stride(0), ptr(workspace), wsb, B, H, KV, hd, Smax, int(max_seqlen),
                                    float(scale), int(window), p0, n0, p1, n1, stream_handle()), "b2_decode_attn")
    return o


def decode_attn_static(qkv, kcache, vcache, seqlens, H, KV, hd, max_seqlen, scale, window=0, rows=None, out=None):
    """decode_attn without a caller workspace, for hosts that cannot allocate device memory at
    call time (allocator closed after start-up, cudaMalloc banned on the decode thread, memory
    owned by a plugin host).  The split-KV partials use a fixed 512 KiB array inside the library;
    bigger problems raise B2Error (too large).  Calls share that array, so do not let two of them
    overlap on the GPU; make the first call on a device before any stream capture.
    b2_decode_attn_static.  (`out` is allocated here only when the caller passes none.)"""
    _cuda(qkv, kcache, vcache, seqlens, rows, out)
    B = qkv.shape[0]
    Smax = kcache.shape[2]
    o = out if out is not None else torch.empty((B, H * hd), dtype=torch.float16, device=qkv.device)
    check(_abi.lib().b2_decode_attn_static(ptr(qkv), qkv.stride(0), ptr(kcache), ptr(vcache), ptr(seqlens), ptr(rows),
                                           ptr(o), o.stride(0), B, H, KV, hd, Smax, int(max_seqlen), float(scale),
                                           int(window), stream_handle()), "b2_decode_attn_static")
    return o


# ---- LLM.int8() mixed decomposition ------------------------------------------------------------
LLM_INT8_THRESHOLD = 6.0      # bitsandbytes' default `llm_int8_threshold`


def act_mul(gate_up, F, act, out=None):
    """act(gate)*up (act 0/1: input [T,2F]) or act(x) (act 2/3: input [T,F]) as fp16 [T,F].  b2_act_mul."""
    _cuda(gate_up, out)
    T = gate_up.shape[0]
    o = out if out is not None else torch.empty((T, F), dtype=torch.float16, device=gate_up.device)
    check(_abi.lib().b2_act_mul(ptr(gate_up), ptr(o), T, F, int(act), stream_handle()), "b2_act_mul")
    return o


class OutlierScratch:
    """Device buffers of one LLM.int8() linear: mask [K] fp32, cols [K] int32, count [1] int32, flags [K/8] u32."""

    def __init__(self, K: int, device):
        self.K = K
        self.mask = torch.empty((K,), dtype=torch.float32, device=device)
        self.cols = torch.empty((K,), dtype=torch.int32, device=device)
        self.count = torch.zeros((1,), dtype=torch.int32, device=device)
        self.flags = torch.empty((K // 8,), dtype=torch.int32, device=device)


def outlier_scan(y, scratch: OutlierScratch, threshold=LLM_INT8_THRESHOLD):
    """Outlier columns of y [T,K] fp16 (any |y| >= threshold) into `scratch`.  b2_outlier_scan."""
    _cuda(y)
    T, K = y.shape
    check(_abi.lib().b2_outlier_scan(ptr(y), y.stride(0), T, K, float(threshold), ptr(scratch.mask), ptr(scratch.cols),
                                     ptr(scratch.count), ptr(scratch.flags), stream_handle()), "b2_outlier_scan")
    return scratch


def outlier_side(out, y, w_q, w_scale, scratch: OutlierScratch, residual=None):
    """out += fp16 side product over the outlier columns, then += residual (in place).  b2_outlier_side."""
    _cuda(out, y, w_q, w_scale, residual)
    T, N = out.shape
    check(_abi.lib().b2_outlier_side(ptr(out), ptr(y), y.stride(0), ptr(w_q), ptr(w_scale), ptr(scratch.cols),
                                     ptr(scratch.count), ptr(residual), T, N, w_q.shape[1], stream_handle()),
          "b2_outlier_side")
    return out


class DenseSideScratch:
    """Buffers of the tensor-core side product (b2_outlier_side_dense), sized for the largest linear of a model:
    ysub [T,Kmax] fp16, wdeq [max N*K] fp16, side [T, Nmax] fp32."""

    def __init__(self, T: int, kmax: int, nmax: int, nk_max: int, device):
        self.T = T
        self.ysub = torch.empty((T * kmax,), dtype=torch.float16, device=device)
        self.wdeq = torch.empty((nk_max,), dtype=torch.float16, device=device)
        self.side = torch.empty((T * nmax,), dtype=torch.float32, device=device)


def outlier_side_dense(out, y, w_q, w_scale, scratch: OutlierScratch, dense: DenseSideScratch, residual=None):
    """out += side product over the outlier columns as an fp16 tensor-core GEMM, then += residual."""
    _cuda(out, y, w_q, w_scale, residual)
    T, N = out.shape
    K = w_q.shape[1]
    check(_abi.lib().b2_outlier_side_dense(ptr(out), ptr(y), y.stride(0), ptr(w_q), ptr(w_scale), ptr(scratch.mask),
                                           ptr(residual), ptr(dense.ysub), ptr(dense.wdeq), ptr(dense.side), T, N, K,
                                           stream_handle()), "b2_outlier_side_dense")
    return out


def llm_int8_linear(y, w_q, w_scale, scratch: OutlierScratch, bias=None, residual=None, out=None, q_buf=None,
                    scale_buf=None, threshold=LLM_INT8_THRESHOLD, dense: DenseSideScratch | None = None,
                    fused_decode: bool = False):
    """One nn.Linear the way bitsandbytes' Linear8bitLt computes it (reference: load_in_8bit=True,
    models_quant_updated.py:30-32): outlier columns of the fp16 activation in fp16, the rest W8A8."""
    T, K = y.shape
    outlier_scan(y, scratch, threshold)
    if fused_decode and T <= 8 and dense is None:
        # decode: masked quantise + int8 GEMV + dequantising epilogue in one launch (bit-identical to the two below)
        o = out if out is not None else torch.empty((T, w_q.shape[0]), dtype=torch.float16, device=y.device)
        fused_linear(y, w_q, w_scale, x_kind=X_PLAIN, smooth=scratch.mask, bias=bias, out=o)
        return outlier_side(o, y, w_q, w_scale, scratch, residual)
    q = q_buf if q_buf is not None else torch.empty((T, K), dtype=torch.int8, device=y.device)
    sc = scale_buf if scale_buf is not None else torch.empty((T,), dtype=torch.float32, device=y.device)
    smooth_quant(y, scratch.mask, out_q=q, out_scale=sc)
    o = gemm_w8a8(q, w_q, sc, w_scale, bias, None, out=out)
    if dense is not None:       # long batches: thousands of outlier columns -> tensor-core side product
        return outlier_side_dense(o, y, w_q, w_scale, scratch, dense, residual)
    return outlier_side(o, y, w_q, w_scale, scratch, residual)


def set_prefill_attn_impl(impl: int):
    """Test hook: 0 = choose, 1 = always the mma.sync kernel, 2 = the tcgen05 kernel or an error."""
    check(_abi.lib().b2_set_prefill_attn_impl(int(impl)), "b2_set_prefill_attn_impl")


def prefill_attn(qkv, kcache, vcache, pos0, row, H, KV, hd, scale, window=0, out=None, positions=None):
    """Causal attention for one chunk of one sequence.  b2_prefill_attn."""
    _cuda(qkv, kcache, vcache, out)
    T = qkv.shape[0]
    Smax = kcache.shape[2]
    o = out if out is not None else torch.empty((T, H * hd), dtype=torch.float16, device=qkv.device)
    check(_abi.lib().b2_prefill_attn(ptr(qkv), qkv.stride(0), ptr(kcache), ptr(vcache), ptr(o), o.stride(0), ptr(positions), T,
                                     int(pos0), int(row), H, KV, hd, Smax, float(scale), int(window),
                                     stream_handle()), "b2_prefill_attn")
    return o


def embed(ids, table, positions=None, pos_table=None, out=None):
    """ids [T] int64 -> [T,d] fp16.  b2_embed."""
    _cuda(ids, table, positions, pos_table)
    T = ids.shape[0]
    V, d = table.shape
    o = out if out is not None else torch.empty((T, d), dtype=torch.float16, device=ids.device)
    check(_abi.lib().b2_embed(ptr(ids), ptr(positions), ptr(table), ptr(pos_table), ptr(o), T, d, V, stream_handle()),
          "b2_embed")
    return o


def lm_head(x, w, out=None):
    """x [M,d] fp16, w [V,d] fp16 -> logits [M,V] fp32.  b2_lm_head."""
    _cuda(x, w)
    M, d = x.shape
    V = w.shape[0]
    o = out if out is not None else torch.empty((M, V), dtype=torch.float32, device=x.device)
    check(_abi.lib().b2_lm_head(ptr(x), ptr(w), ptr(o), M, V, d, stream_handle()), "b2_lm_head")
    return o


def head_argmax_scratch(device):
    """The zeroed scratch b2_lm_head_argmax works in (it leaves it zeroed after every launch)."""
    return torch.zeros((int(_abi.lib().b2_lm_head_argmax_scratch_bytes()),), dtype=torch.uint8, device=device)


def lm_head_argmax(x, w, scratch, idx_offset=0, other=None, out_idx=None, out_idx2=None, out_val=None, logits=None,
                   wait=None, notify=None, err=None):
    """Greedy token of x [M<=4,d] fp16 over the head rows w [V,d] fp16 (global indices idx_offset + row) in one launch;
    `other` = (values [M] fp32, global indices [M] int64) of another vocabulary shard to merge with.
    wait = (flag, seq): x is a message in this GPU's inbox, wait for its sequence number first; notify = (flag, seq):
    publish seq there after storing the outputs (which may live in a peer's memory).  b2_lm_head_argmax."""
    _cuda(x, w, scratch, out_idx, out_idx2, out_val, logits)
    M, d = x.shape
    V = w.shape[0]
    oi = out_idx if out_idx is not None else torch.empty((M,), dtype=torch.int64, device=x.device)
    ov, ox = (other if other is not None else (None, None))
    wf, ws = wait if wait is not None else (None, 0)
    nf, ns = notify if notify is not None else (None, 0)
    check(_abi.lib().b2_lm_head_argmax(ptr(x), ptr(w), ptr(logits), M, V, d, int(idx_offset), ptr(ov), ptr(ox), ptr(oi), ptr(out_idx2),
                                       ptr(out_val), ptr(scratch), ptr(wf), int(ws) & 0xFFFFFFFF, ptr(err), ptr(nf),
                                       int(ns) & 0xFFFFFFFF, stream_handle()), "b2_lm_head_argmax")
    return oi


class MultiDst:
    """The destinations of a b2_p2p_send_multi, prepared once: (tensor, flag) pairs -> two host pointer arrays."""

    def __init__(self, pairs):
        n = len(pairs)
        self.n = n
        self.keep = list(pairs)
        self.dsts = (_abi.C.c_void_p * n)(*[t.data_ptr() for t, _ in pairs])
        self.flags = (_abi.C.c_void_p * n)(*[f.data_ptr() for _, f in pairs])


def p2p_send_multi(src, multi: MultiDst, seq: int):
    """Store the bytes of local `src` to every destination of `multi` and publish `seq` at each flag.  b2_p2p_send_multi."""
    nb = src.numel() * src.element_size()
    check(_abi.lib().b2_p2p_send_multi(ptr(src), nb, multi.dsts, multi.flags, multi.n, int(seq) & 0xFFFFFFFF, stream_handle()),
          "b2_p2p_send_multi")


def merge_winners(flags, seq: int, records, W: int, B: int, out_idx, out_idx2=None, err=None):
    """Greedy token from the winners of W vocabulary slices in this GPU's inbox (b2_merge_winners)."""
    check(_abi.lib().b2_merge_winners(ptr(flags), int(seq) & 0xFFFFFFFF, ptr(records), int(W), int(B), ptr(out_idx), ptr(out_idx2),
                                      ptr(err), stream_handle()), "b2_merge_winners")
    return out_idx


def argmax(logits, out=None):
    """logits [T,V] fp32 -> int64 [T].  b2_argmax."""
    _cuda(logits)
    T, V = logits.shape
    o = out if out is not None else torch.empty((T,), dtype=torch.int64, device=logits.device)
    check(_abi.lib().b2_argmax(ptr(logits), ptr(o), T, V, stream_handle()), "b2_argmax")
    return o


def p2p_send(src, dst_peer, flag_peer, seq: int):
    """Store the bytes of local `src` into the peer-mapped tensor `dst_peer`, then publish `seq` at `flag_peer`
    (a one-element uint32/int32 view of the receiver's memory).  b2_p2p_send."""
    nb = src.numel() * src.element_size()
    check(_abi.lib().b2_p2p_send(ptr(src), ptr(dst_peer), nb, ptr(flag_peer), int(seq) & 0xFFFFFFFF, stream_handle()),
          "b2_p2p_send")


def p2p_wait(flag, seq: int, src=None, dst=None, err=None):
    """Wait (on the device) until the local `flag` has reached `seq`; then copy inbox `src` to `dst` if given.
    b2_p2p_wait."""
    nb = 0 if dst is None else dst.numel() * dst.element_size()
    check(_abi.lib().b2_p2p_wait(ptr(flag), int(seq) & 0xFFFFFFFF, ptr(src) if dst is not None else None, ptr(dst), nb,
                                 ptr(err), stream_handle()), "b2_p2p_wait")


def argmax_shard(logits, idx_offset=0, out_idx=None, out_val=None, other=None):
    """Row argmax of a vocabulary shard [T,V] fp32 whose first column is global index `idx_offset`; `other` =
    (values [T] fp32, global indices [T] int64) of another shard to merge with.  b2_argmax_shard."""
    _cuda(logits, out_idx, out_val)
    T, V = logits.shape
    o = out_idx if out_idx is not None else torch.empty((T,), dtype=torch.int64, device=logits.device)
    ov, oi = other if other is not None else (None, None)
    check(_abi.lib().b2_argmax_shard(ptr(logits), logits.stride(0), ptr(o), ptr(out_val), int(idx_offset), ptr(ov), ptr(oi),
                                     T, V, stream_handle()), "b2_argmax_shard")
    return o


def bias_add_f32(logits, bias):
    """logits [T,V] fp32 += bias [V] fp32 (in place).  b2_bias_add_f32."""
    _cuda(logits, bias)
    T, V = logits.shape
    check(_abi.lib().b2_bias_add_f32(ptr(logits), ptr(bias), logits.stride(0), T, V, stream_handle()), "b2_bias_add_f32")
    return logits


def confidence(logits):
    """logits [T,V] fp32 -> (mean over rows of the largest softmax probability [1] fp32, per-row values [T]).
    b2_confidence."""
    _cuda(logits)
    if logits.dtype != torch.float32:
        raise ValueError("confidence needs fp32 logits")
    T, V = logits.shape
    pm = torch.empty((T,), dtype=torch.float32, device=logits.device)
    out = torch.empty((1,), dtype=torch.float32, device=logits.device)
    check(_abi.lib().b2_confidence(ptr(logits), logits.stride(0), ptr(pm), ptr(out), T, V, stream_handle()), "b2_confidence")
    return out, pm


def sample(logits, *, temperature=1.0, top_k=50, top_p=1.0, repetition_penalty=1.0, seed=0, step=0, history=None,
           hist_len=None, step_tensor=None, out=None):
    """logits [B,V] fp32 -> sampled ids int64 [B] with HF's processor chain (repetition penalty ->
    temperature -> top-k -> top-p) and a counter-based draw.  history [B,Hmax] int64 + hist_len [B]
    int32 = the ids seen so far per row; step_tensor: optional device uint64 counter read instead of
    `step`.  b2_sample."""
    _cuda(logits, history, hist_len, step_tensor, out)
    B, V = logits.shape
    o = out if out is not None else torch.empty((B,), dtype=torch.int64, device=logits.device)
    check(_abi.lib().b2_sample(ptr(logits), logits.stride(0), ptr(history), ptr(hist_len),
                               0 if history is None else history.stride(0), ptr(o), B, V, float(temperature), int(top_k),
                               float(top_p), float(repetition_penalty), int(seed) & 0xFFFFFFFFFFFFFFFF, ptr(step_tensor),
                               int(step), stream_handle()), "b2_sample")
    return o
