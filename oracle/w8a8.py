"""Tier-S oracle: CPU restatement (numpy) of the W8A8 decoder-block forward.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is on the product path: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline leg may import it.

PARITY UNPINNED.  The reference (parthabp55/LLM-for-Distributed-Egde-Devices) ships no
tests, golden vectors or fixtures, and the arithmetic of its quantised path
lives in unvendored, unpinned third-party code (``Code/others/requirements.txt:1-10``:
transformers, bitsandbytes).  This file therefore restates

  * the decoder-block *structure* the reference executes through HF ``generate``
    (call sites ``Code/Quantised Models/models_quant_updated.py:84-93`` and ``:66-67``;
    block = transformers 5.5.0 ``models/llama/modeling_llama.py:303-334`` (layer),
    ``:251-289`` (attention), ``:199-222`` (eager attention), ``:182-184`` (MLP),
    ``:62-67`` (RMSNorm), rotary ``:128-153``), and
  * the *W8A8 numerics* that ``BASELINE.json:north_star`` specifies (SmoothQuant,
    arXiv 2211.10438: per-output-channel symmetric int8 weights, per-token dynamic
    symmetric int8 activations, exact int32 accumulation, fp32 dequantising epilogue).
    bitsandbytes ``Linear8bitLt`` (``models_quant_updated.py:30-32``) with no activation
    outlier >= 6.0 is this scheme with smoothing alpha = 0.

Every floating-point step is written as individually rounded IEEE fp32 operations
(no fused multiply-add, no library transcendental) so the CUDA kernels can reproduce the
int8 activations, the int32 accumulators and the fp16 GEMM outputs bit for bit.  The one
exception is the softmax in attention (sum order and exp differ) which is compared within
a stated tolerance.

Pinned formulas (shared with the kernels in ``llm-for-distributed-egde-devices_b200/csrc``):

  quantise row y (fp32):  amax = max|y|;  scale = amax / 127;  inv = 127 / amax;
                          q = clamp(rint(y * inv), -127, 127)   (rint = half-to-even);
                          amax == 0  ->  q = 0, scale = 0.
  RMSNorm:   ss = sum_k double(x_k)^2  (fp64; order-insensitive to ~2^-53);
             var = float(ss / d);  r = 1 / sqrt(var + eps)  (IEEE fp32 div and sqrt);
             y_k = ((x_k * r) * gamma_k) [* smooth_k]
  epilogue:  v = (float(acc) * a_scale[m]) * w_scale[n] [+ bias[n]] [+ residual[m,n]]
             out = fp16(v)   (one rounding)
  exp_det:   Cody-Waite range reduction + degree-5 polynomial, all separately rounded.
"""
from __future__ import annotations

import numpy as np

f32 = np.float32
f16 = np.float16

# ----------------------------------------------------------------------------------------
# deterministic transcendental pieces
# ----------------------------------------------------------------------------------------
_LOG2E = f32(1.4426950408889634)
_LN2_HI = f32(0.693359375)            # 355/512: n*_LN2_HI is exact for |n| < 2^15
_LN2_LO = f32(-2.12194440e-4)
_EXP_C = [f32(1.9875691500e-4), f32(1.3981999507e-3), f32(8.3334519073e-3),
          f32(4.1665795894e-2), f32(1.6666665459e-1), f32(5.0000001201e-1)]
_EXP_CLAMP = f32(80.0)


def exp_det(t: np.ndarray) -> np.ndarray:
    """exp(t) for fp32 t, as a fixed sequence of individually rounded fp32 operations.

    Mirrors ``b2_exp_det`` in csrc/common.cuh operation for operation.
    """
    t = np.asarray(t, dtype=f32)
    t = np.minimum(np.maximum(t, -_EXP_CLAMP), _EXP_CLAMP)
    n = np.rint(t * _LOG2E).astype(f32)
    r = (t - n * _LN2_HI).astype(f32)
    r = (r - (n * _LN2_LO).astype(f32)).astype(f32)
    p = np.full_like(r, _EXP_C[0])
    for c in _EXP_C[1:]:
        p = ((p * r).astype(f32) + c).astype(f32)
    r2 = (r * r).astype(f32)
    res = ((p * r2).astype(f32) + r).astype(f32)
    res = (res + f32(1.0)).astype(f32)
    return np.ldexp(res, n.astype(np.int32)).astype(f32)


def silu_det(g: np.ndarray) -> np.ndarray:
    """SiLU g*sigmoid(g) (modeling_llama.py:182-184 act_fn) with the deterministic exp."""
    g = np.asarray(g, dtype=f32)
    e = exp_det(-g)
    den = (f32(1.0) + e).astype(f32)
    return (g / den).astype(f32)


def gelu_new_det(u: np.ndarray) -> np.ndarray:
    """GPT-2 ``gelu_new`` 0.5*u*(1+tanh(sqrt(2/pi)*(u+0.044715*u^3))) with tanh via exp_det."""
    u = np.asarray(u, dtype=f32)
    u3 = ((u * u).astype(f32) * u).astype(f32)
    inner = (u + (f32(0.044715) * u3).astype(f32)).astype(f32)
    z = (f32(0.7978845608028654) * inner).astype(f32)
    e = exp_det((f32(2.0) * z).astype(f32))
    th = (f32(1.0) - (f32(2.0) / (e + f32(1.0)).astype(f32)).astype(f32)).astype(f32)
    return ((f32(0.5) * u).astype(f32) * (f32(1.0) + th).astype(f32)).astype(f32)


def gelu_erf_det(u: np.ndarray) -> np.ndarray:
    """Exact GELU 0.5*u*(1+erf(u/sqrt 2)) (HF "gelu": GPT-NeoX / Pythia, modeling_gpt_neox.py:43) with erf from
    Abramowitz-Stegun 7.1.26 -- erf(a) = 1 - (a1 t + ... + a5 t^5) exp(-a^2), t = 1/(1 + p a), |error| <= 1.5e-7,
    i.e. at the resolution of fp32 -- on exp_det, every step individually rounded (kernel: gelu_erf_det)."""
    u = np.asarray(u, dtype=f32)
    z = (u * f32(0.7071067811865476)).astype(f32)
    a = np.abs(z).astype(f32)
    t = (f32(1.0) / (f32(1.0) + (f32(0.3275911) * a).astype(f32)).astype(f32)).astype(f32)
    p = f32(1.061405429)
    for c in (-1.453152027, 1.421413741, -0.284496736, 0.254829592):
        p = ((p * t).astype(f32) + f32(c)).astype(f32)
    p = (p * t).astype(f32)
    e = exp_det((-(a * a).astype(f32)).astype(f32))
    er = (f32(1.0) - (p * e).astype(f32)).astype(f32)
    er = np.where(z < 0, -er, er).astype(f32)
    return ((f32(0.5) * u).astype(f32) * (f32(1.0) + er).astype(f32)).astype(f32)


def act_fn(name: str):
    return {"silu": silu_det, "gelu_new": gelu_new_det, "gelu": gelu_erf_det}[name]


# ----------------------------------------------------------------------------------------
# quantisation
# ----------------------------------------------------------------------------------------
def quant_rows(y: np.ndarray):
    """Per-token (row) dynamic symmetric absmax int8 quantisation of fp32 y [T, K]."""
    y = np.asarray(y, dtype=f32)
    amax = np.max(np.abs(y), axis=-1).astype(f32)
    safe = np.where(amax > 0, amax, f32(1.0)).astype(f32)
    scale = np.where(amax > 0, (amax / f32(127.0)).astype(f32), f32(0.0)).astype(f32)
    inv = np.where(amax > 0, (f32(127.0) / safe).astype(f32), f32(0.0)).astype(f32)
    q = np.rint((y * inv[:, None]).astype(f32))
    q = np.clip(q, -127, 127).astype(np.int8)
    return q, scale


def rmsnorm_rows(x16: np.ndarray, gamma16: np.ndarray, eps: float) -> np.ndarray:
    """RMSNorm (modeling_llama.py:62-67) in the pinned fp64-sum / fp32-elementwise form."""
    x = np.asarray(x16, dtype=f16)
    xd = x.astype(np.float64)
    ss = np.sum(xd * xd, axis=-1)
    var = (ss / x.shape[-1]).astype(f32)
    r = (f32(1.0) / np.sqrt((var + f32(eps)).astype(f32)).astype(f32)).astype(f32)
    y = (x.astype(f32) * r[:, None]).astype(f32)
    return (y * np.asarray(gamma16, dtype=f16).astype(f32)[None, :]).astype(f32)


def layernorm_rows(x16, gamma16, beta16, eps: float) -> np.ndarray:
    """LayerNorm (GPT-2 ``ln_1``/``ln_2``): fp64 moments, fp32 elementwise."""
    x = np.asarray(x16, dtype=f16)
    xd = x.astype(np.float64)
    d = x.shape[-1]
    mean_d = np.sum(xd, axis=-1) / d
    var_d = np.sum(xd * xd, axis=-1) / d - mean_d * mean_d
    mean = mean_d.astype(f32)
    var = np.maximum(var_d, 0.0).astype(f32)
    r = (f32(1.0) / np.sqrt((var + f32(eps)).astype(f32)).astype(f32)).astype(f32)
    y = ((x.astype(f32) - mean[:, None]).astype(f32) * r[:, None]).astype(f32)
    y = (y * np.asarray(gamma16, dtype=f16).astype(f32)[None, :]).astype(f32)
    return (y + np.asarray(beta16, dtype=f16).astype(f32)[None, :]).astype(f32)


def norm_quant(x16, gamma16, beta16, smooth, eps):
    """Fused norm -> smoothing multiply -> per-token int8 quantise (kernel b2_norm_quant).

    ``beta16 is None`` selects RMSNorm.  ``smooth`` (fp32 [d], nullable) is the SmoothQuant
    multiplier 1/s applied to the activation (W carries s).
    """
    y = rmsnorm_rows(x16, gamma16, eps) if beta16 is None else layernorm_rows(x16, gamma16, beta16, eps)
    if smooth is not None:
        y = (y * np.asarray(smooth, dtype=f32)[None, :]).astype(f32)
    return quant_rows(y)


def smooth_quant(x16, smooth):
    """fp16 activation -> optional smoothing multiply -> per-token int8 (kernel b2_smooth_quant)."""
    y = np.asarray(x16, dtype=f16).astype(f32)
    if smooth is not None:
        y = (y * np.asarray(smooth, dtype=f32)[None, :]).astype(f32)
    return quant_rows(y)


def act_mul_y(gate_up16, smooth, act="silu"):
    """The fp32 values act(gate)*up*smooth that act_mul_quant quantises, and their row absmax
    (what the fused decode kernel b2_fused_linear writes with out_kind = 1)."""
    gu = np.asarray(gate_up16, dtype=f16).astype(f32)
    F = gu.shape[1] // 2
    g, u = gu[:, :F], gu[:, F:]
    a = act_fn(act)(g)
    y = (a * u).astype(f32)
    if smooth is not None:
        y = (y * np.asarray(smooth, dtype=f32)[None, :]).astype(f32)
    return y, np.max(np.abs(y), axis=-1).astype(f32)


def act_mul_quant(gate_up16, smooth, act="silu"):
    """act(gate)*up -> smoothing -> per-token int8 (kernel b2_act_mul_quant).

    gate_up16: fp16 [T, 2F], gate in columns [0,F), up in [F,2F)  (modeling_llama.py:182-184).
    """
    y, _ = act_mul_y(gate_up16, smooth, act)
    return quant_rows(y)


def act_quant(h16, smooth, act="gelu_new"):
    """act(h) -> smoothing -> per-token int8, for non-gated MLPs (GPT-2)."""
    h = np.asarray(h16, dtype=f16).astype(f32)
    y = act_fn(act)(h)
    if smooth is not None:
        y = (y * np.asarray(smooth, dtype=f32)[None, :]).astype(f32)
    return quant_rows(y)


def quantize_weight(W: np.ndarray, s: np.ndarray | None):
    """Offline weight quantisation: W [N,K] fp32, smoothing s [K] multiplied into columns;
    per-output-channel symmetric int8."""
    W = np.asarray(W, dtype=f32)
    if s is not None:
        W = (W * np.asarray(s, dtype=f32)[None, :]).astype(f32)
    amax = np.max(np.abs(W), axis=1).astype(f32)
    safe = np.where(amax > 0, amax, f32(1.0)).astype(f32)
    scale = np.where(amax > 0, (amax / f32(127.0)).astype(f32), f32(0.0)).astype(f32)
    Wq = np.clip(np.rint((W / (safe / f32(127.0)).astype(f32)[:, None]).astype(f32)), -127, 127).astype(np.int8)
    return Wq, scale


# ----------------------------------------------------------------------------------------
# W8A8 linear
# ----------------------------------------------------------------------------------------
def int8_acc(q: np.ndarray, wq: np.ndarray) -> np.ndarray:
    """Exact int32 accumulators q [M,K] x wq [N,K]^T.  fp64 BLAS is exact here:
    |acc| <= K*127^2 < 2^53 for any K this path sees (K <= 14336)."""
    acc = q.astype(np.float64) @ wq.astype(np.float64).T
    return acc.astype(np.int64).astype(np.int32)


def int8_acc_slow(q: np.ndarray, wq: np.ndarray) -> np.ndarray:
    """Pure integer restatement (int64 einsum) used to pin int8_acc on small shapes."""
    return np.einsum("mk,nk->mn", q.astype(np.int64), wq.astype(np.int64)).astype(np.int32)


def epilogue(acc, a_scale, w_scale, bias16=None, residual16=None) -> np.ndarray:
    v = (acc.astype(f32) * np.asarray(a_scale, f32)[:, None]).astype(f32)
    v = (v * np.asarray(w_scale, f32)[None, :]).astype(f32)
    if bias16 is not None:
        v = (v + np.asarray(bias16, f16).astype(f32)[None, :]).astype(f32)
    if residual16 is not None:
        v = (v + np.asarray(residual16, f16).astype(f32)).astype(f32)
    return v.astype(f16)


def gemm_w8a8(q, wq, a_scale, w_scale, bias16=None, residual16=None):
    """W8A8 linear with fused dequantising epilogue (kernel b2_gemm_w8a8).
    Returns (out fp16 [M,N], acc int32 [M,N])."""
    acc = int8_acc(q, wq)
    return epilogue(acc, a_scale, w_scale, bias16, residual16), acc


# ----------------------------------------------------------------------------------------
# rotary embedding, KV cache append, attention
# ----------------------------------------------------------------------------------------
def llama3_inv_freq(inv_freq: np.ndarray, factor, low_freq_factor, high_freq_factor, old_ctx) -> np.ndarray:
    """The "llama3" rope_scaling remap of Llama-3.1/3.2 (transformers modeling_rope_utils.py,
    `_compute_llama3_parameters`), fp32 like HF: wavelengths longer than old_ctx/low_freq_factor are
    stretched by `factor`, shorter than old_ctx/high_freq_factor kept, the band between blended."""
    inv = np.asarray(inv_freq, f32)
    wavelen = (f32(2.0 * np.pi) / inv).astype(f32)
    low_wl = f32(old_ctx / low_freq_factor)
    high_wl = f32(old_ctx / high_freq_factor)
    out = np.where(wavelen > low_wl, inv / f32(factor), inv).astype(f32)
    smooth = ((f32(old_ctx) / wavelen - f32(low_freq_factor)) / f32(high_freq_factor - low_freq_factor)).astype(f32)
    blended = ((f32(1.0) - smooth) * out / f32(factor) + smooth * out).astype(f32)
    medium = ~(wavelen < high_wl) & ~(wavelen > low_wl)
    return np.where(medium, blended, out).astype(f32)


# ----------------------------------------------------------------------------------------------
# LLM.int8() mixed decomposition
# ----------------------------------------------------------------------------------------------
LLM_INT8_THRESHOLD = 6.0


def outlier_columns(y16, threshold=LLM_INT8_THRESHOLD) -> np.ndarray:
    """Columns of the fp16 activation that hold any value of magnitude >= threshold, ascending
    (bitsandbytes int8_vectorwise_quant's `outlier_cols`; the reference asks for it with
    BitsAndBytesConfig(load_in_8bit=True), models_quant_updated.py:30-32, default threshold 6.0)."""
    y = np.asarray(y16, f16).astype(f32)
    if y.shape[0] == 0:
        return np.zeros((0,), np.int32)
    return np.nonzero((np.abs(y) >= f32(threshold)).any(axis=0))[0].astype(np.int32)


def llm_int8_linear(y16, w_q, w_scale, bias16=None, residual16=None, threshold=LLM_INT8_THRESHOLD):
    """bitsandbytes MatMul8bitLt.forward restated (the package is not installable here: PARITY UNPINNED):
    1. outlier columns O of the fp16 activation;  2. the other columns: per-row absmax int8 (vectorwise);
    3. int8 x int8 -> int32, dequantised with row scale x weight-row scale (+ bias) to fp16;
    4. + fp16( y[:,O] . fp16(W_q[:,O] * w_scale)^T ) accumulated in fp32 in ascending column order, fp16;
    5. + residual (fp16 add, as the decoder layer does outside the linear).
    Returns (out fp16 [T,N], acc int32 [T,N], O)."""
    y = np.asarray(y16, f16)
    cols = outlier_columns(y, threshold)
    mask = np.ones((y.shape[1],), f32)
    mask[cols] = 0.0
    q, s = smooth_quant(y, mask)
    out, acc = gemm_w8a8(q, w_q, s, w_scale, bias16, None)
    if len(cols):
        yf = y.astype(f32)
        side = np.zeros(out.shape, f32)
        for k in cols:
            w16 = (w_q[:, k].astype(f32) * np.asarray(w_scale, f32)).astype(f16).astype(f32)
            side = (side + (yf[:, k:k + 1] * w16[None, :]).astype(f32)).astype(f32)
        out = (out.astype(f32) + side.astype(f16).astype(f32)).astype(f16)
    if residual16 is not None:
        out = (out.astype(f32) + np.asarray(residual16, f16).astype(f32)).astype(f16)
    return out, acc, cols


def rope_table(head_dim: int, max_pos: int, theta: float = 10000.0, scaling=(), rotary_dim: int = 0):
    """cos/sin tables fp32 [max_pos, head_dim/2] (modeling_llama.py:100-153: inv_freq =
    theta^(-2i/hd), angle = pos*inv_freq; optional ("llama3", factor, low, high, old_ctx) remap).
    Computed in fp64, rounded once; the GPU reads this exact table, so no device-side sincos is
    involved."""
    rd = rotary_dim or head_dim
    i = np.arange(0, rd, 2, dtype=np.float64)
    inv_freq = (1.0 / (theta ** (i / rd))).astype(f32)
    if scaling:
        assert scaling[0] == "llama3", scaling
        inv_freq = llama3_inv_freq(inv_freq, *scaling[1:])
    inv_freq = inv_freq.astype(np.float64)
    ang = np.arange(max_pos, dtype=np.float64)[:, None] * inv_freq[None, :]
    ang = ang.astype(f32).astype(np.float64)   # HF forms the angle in fp32
    # partial rotary (Phi, GPT-NeoX: modeling_phi.py:216-231): in the kernel layout the rotary pairs are the first
    # rd/2 pairs (j, j + head_dim/2); every other pair gets cos = 1, sin = 0, which apply_rope passes through exactly
    cos, sin = np.ones((max_pos, head_dim // 2), f32), np.zeros((max_pos, head_dim // 2), f32)
    cos[:, : rd // 2], sin[:, : rd // 2] = np.cos(ang).astype(f32), np.sin(ang).astype(f32)
    return cos, sin


def apply_rope(x16: np.ndarray, positions: np.ndarray, cos: np.ndarray, sin: np.ndarray) -> np.ndarray:
    """x16 [T, nheads, hd] fp16 -> rotated fp16.  rotate_half convention
    (modeling_llama.py:128-153): out[:h] = x[:h]*c - x[h:]*s ; out[h:] = x[h:]*c + x[:h]*s."""
    x = np.asarray(x16, f16).astype(f32)
    h = x.shape[-1] // 2
    c = cos[positions][:, None, :]
    s = sin[positions][:, None, :]
    x1, x2 = x[..., :h], x[..., h:]
    o1 = ((x1 * c).astype(f32) - (x2 * s).astype(f32)).astype(f32)
    o2 = ((x2 * c).astype(f32) + (x1 * s).astype(f32)).astype(f32)
    return np.concatenate([o1, o2], axis=-1).astype(f16)


def attention(q16, k16, v16, q_pos, scale, window=0):
    """Causal (optionally sliding-window) attention, eager form (modeling_llama.py:199-222).

    q16 [Tq, H, hd]; k16/v16 [S, KV, hd] (the cache contents for this sequence, positions
    0..S-1); q_pos [Tq] absolute positions.  Key j is visible to query i iff
    j <= q_pos[i] and (window == 0 or j > q_pos[i] - window).  fp32 math, fp16 result.
    Compared against the kernels within tolerance, not bit-exactly.
    """
    q = np.asarray(q16, f16).astype(f32)
    k = np.asarray(k16, f16).astype(f32)
    v = np.asarray(v16, f16).astype(f32)
    Tq, H, hd = q.shape
    S, KV, _ = k.shape
    g = H // KV
    out = np.zeros((Tq, H, hd), dtype=f32)
    j = np.arange(S)[None, :]
    vis = j <= np.asarray(q_pos)[:, None]
    if window and window > 0:
        vis &= j > (np.asarray(q_pos)[:, None] - window)
    for h in range(H):
        kh = k[:, h // g, :]
        vh = v[:, h // g, :]
        s = (q[:, h, :] @ kh.T) * f32(scale)
        s = np.where(vis, s, -np.inf)
        m = np.max(s, axis=-1, keepdims=True)
        p = np.exp(s - m)
        p = p / np.sum(p, axis=-1, keepdims=True)
        out[:, h, :] = p @ vh
    return out.astype(f16)


# ----------------------------------------------------------------------------------------
# decoder block / stage (Llama, Mistral family)
# ----------------------------------------------------------------------------------------
class KVCacheOracle:
    """Per-stage KV cache, fp16, [layer][B][KV][Smax][hd] like the device layout."""

    def __init__(self, n_layers, batch, kv_heads, max_seq, head_dim):
        self.k = np.zeros((n_layers, batch, kv_heads, max_seq, head_dim), dtype=f16)
        self.v = np.zeros_like(self.k)
        self.seqlens = np.zeros((batch,), dtype=np.int32)


def block_forward(h16, lw, cfg, cache: KVCacheOracle, li, seq_ids, positions, cos, sin, trace=None):
    """One W8A8 decoder layer over T tokens.

    h16 [T,d] fp16; seq_ids [T] (which cache row each token belongs to); positions [T].
    lw: dict of numpy arrays as produced by weights.quantize_layer (see package).
    Follows modeling_llama.py:303-334 with every nn.Linear replaced by gemm_w8a8.
    """
    H, KV, hd = cfg["n_heads"], cfg["n_kv_heads"], cfg["head_dim"]
    eps = cfg["eps"]
    T = h16.shape[0]
    # --- attention half -------------------------------------------------------------
    q1, s1 = norm_quant(h16, lw["ln1_g"], lw.get("ln1_b"), lw.get("sm_qkv"), eps)
    qkv, acc_qkv = gemm_w8a8(q1, lw["wqkv"], s1, lw["s_qkv"], lw.get("b_qkv"), None)
    qh = qkv[:, : H * hd].reshape(T, H, hd)
    kh = qkv[:, H * hd: (H + KV) * hd].reshape(T, KV, hd)
    vh = qkv[:, (H + KV) * hd:].reshape(T, KV, hd)
    if cfg.get("rope", True):
        qh = apply_rope(qh, positions, cos, sin)
        kh = apply_rope(kh, positions, cos, sin)
    for t in range(T):
        cache.k[li, seq_ids[t], :, positions[t], :] = kh[t]
        cache.v[li, seq_ids[t], :, positions[t], :] = vh[t]
    attn = np.zeros((T, H, hd), dtype=f16)
    scale = 1.0 / np.sqrt(cfg.get("head_dim_model") or hd)      # heads padded to the kernel width keep the model's scale
    for b in np.unique(seq_ids):
        idx = np.nonzero(seq_ids == b)[0]
        S = int(positions[idx].max()) + 1
        kb = np.transpose(cache.k[li, b, :, :S, :], (1, 0, 2))
        vb = np.transpose(cache.v[li, b, :, :S, :], (1, 0, 2))
        attn[idx] = attention(qh[idx], kb, vb, positions[idx], scale, cfg.get("window", 0))
    q2, s2 = smooth_quant(attn.reshape(T, H * hd), lw.get("sm_o"))
    h1, acc_o = gemm_w8a8(q2, lw["wo"], s2, lw["s_o"], lw.get("b_o"), h16)
    # --- MLP half -------------------------------------------------------------------
    # parallel residual (Phi: modeling_phi.py:289-307; GPT-NeoX use_parallel_residual): the MLP reads the layer
    # INPUT, and the block returns h + attn + mlp = (h + attn) + mlp
    mlp_in = h16 if cfg.get("parallel", False) else h1
    q3, s3 = norm_quant(mlp_in, lw["ln2_g"], lw.get("ln2_b"), lw.get("sm_gu"), eps)
    gu, acc_gu = gemm_w8a8(q3, lw["wgu"], s3, lw["s_gu"], lw.get("b_gu"), None)
    if cfg.get("gated", True):
        q4, s4 = act_mul_quant(gu, lw.get("sm_d"), cfg.get("act", "silu"))
    else:
        q4, s4 = act_quant(gu, lw.get("sm_d"), cfg.get("act", "gelu_new"))
    h2, acc_d = gemm_w8a8(q4, lw["wd"], s4, lw["s_d"], lw.get("b_d"), h1)
    if trace is not None:
        trace.append(dict(q1=q1, s1=s1, acc_qkv=acc_qkv, qkv=qkv, attn=attn, q2=q2, s2=s2, acc_o=acc_o,
                          h1=h1, q3=q3, s3=s3, acc_gu=acc_gu, gu=gu, q4=q4, s4=s4, acc_d=acc_d, h2=h2))
    return h2


def _attend(qkv, cfg, cache, li, seq_ids, positions, cos, sin):
    """RoPE + KV append + attention over the cache for the fp16 qkv rows of one layer (shared by the modes)."""
    H, KV, hd = cfg["n_heads"], cfg["n_kv_heads"], cfg["head_dim"]
    T = qkv.shape[0]
    qh = qkv[:, : H * hd].reshape(T, H, hd)
    kh = qkv[:, H * hd: (H + KV) * hd].reshape(T, KV, hd)
    vh = qkv[:, (H + KV) * hd:].reshape(T, KV, hd)
    if cfg.get("rope", True):
        qh = apply_rope(qh, positions, cos, sin)
        kh = apply_rope(kh, positions, cos, sin)
    for t in range(T):
        cache.k[li, seq_ids[t], :, positions[t], :] = kh[t]
        cache.v[li, seq_ids[t], :, positions[t], :] = vh[t]
    attn = np.zeros((T, H, hd), dtype=f16)
    scale = 1.0 / np.sqrt(cfg.get("head_dim_model") or hd)
    for b in np.unique(seq_ids):
        idx = np.nonzero(seq_ids == b)[0]
        S = int(positions[idx].max()) + 1
        kb = np.transpose(cache.k[li, b, :, :S, :], (1, 0, 2))
        vb = np.transpose(cache.v[li, b, :, :S, :], (1, 0, 2))
        attn[idx] = attention(qh[idx], kb, vb, positions[idx], scale, cfg.get("window", 0))
    return attn.reshape(T, H * hd)


def block_forward_llm_int8(h16, lw, cfg, cache: KVCacheOracle, li, seq_ids, positions, cos, sin, trace=None):
    """One decoder layer with every nn.Linear computed the LLM.int8() way (llm_int8_linear): the
    layer structure of modeling_llama.py:303-334 as the reference runs it under
    BitsAndBytesConfig(load_in_8bit=True) (models_quant_updated.py:30-38).  Norms and act*up produce
    fp16 activations (the tensors bitsandbytes sees), residual adds are fp16 adds outside the linear."""
    eps = cfg["eps"]

    def norm16(x, g, b):
        y = rmsnorm_rows(x, g, eps) if b is None else layernorm_rows(x, g, b, eps)
        return y.astype(f16)

    y1 = norm16(h16, lw["ln1_g"], lw.get("ln1_b"))
    qkv, _, c1 = llm_int8_linear(y1, lw["wqkv"], lw["s_qkv"], lw.get("b_qkv"), None)
    attn = _attend(qkv, cfg, cache, li, seq_ids, positions, cos, sin)
    h1, _, c2 = llm_int8_linear(attn, lw["wo"], lw["s_o"], lw.get("b_o"), h16)
    y2 = norm16(h1, lw["ln2_g"], lw.get("ln2_b"))
    gu, _, c3 = llm_int8_linear(y2, lw["wgu"], lw["s_gu"], lw.get("b_gu"), None)
    if cfg.get("gated", True):
        hm = act_mul_y(gu, None, cfg.get("act", "silu"))[0].astype(f16)
    else:
        g = np.asarray(gu, f16).astype(f32)
        hm = act_fn(cfg.get("act", "gelu_new"))(g).astype(f16)
    h2, _, c4 = llm_int8_linear(hm, lw["wd"], lw["s_d"], lw.get("b_d"), h1)
    if trace is not None:
        trace.append(dict(y1=y1, qkv=qkv, attn=attn.reshape(h16.shape[0], cfg["n_heads"], cfg["head_dim"]), h1=h1, y2=y2,
                          gu=gu, hm=hm, h2=h2, cols=(c1, c2, c3, c4)))
    return h2


def final_norm(h16, g16, b16, eps):
    y = rmsnorm_rows(h16, g16, eps) if b16 is None else layernorm_rows(h16, g16, b16, eps)
    return y.astype(f16)


def lm_head(hn16, w16):
    """Unquantised fp16 lm_head with fp32 accumulation -> fp32 logits."""
    return (np.asarray(hn16, f16).astype(f32) @ np.asarray(w16, f16).astype(f32).T).astype(f32)


def confidence(logits) -> np.float32:
    """mean over positions of the largest softmax probability (the reference's confidence_score,
    models_quant_updated.py:63-70: softmax(logits) -> max over V -> mean over positions) as kernel b2_confidence
    computes it: exp_det of (x - rowmax) in fp32, row sums and the final mean in fp64."""
    x = np.asarray(logits, dtype=f32)
    m = x.max(axis=-1, keepdims=True)
    e = exp_det((x - m).astype(f32))
    p = (1.0 / e.astype(np.float64).sum(axis=-1)).astype(f32)
    return f32(p.astype(np.float64).sum() / p.shape[0])


def stage_forward(x, sw, cfg, cache, seq_ids, positions, cos, sin, trace=None, logits_rows=None):
    """Forward over a contiguous layer range (the build-defined Stage contract).

    x: int64 ids [T] if sw['first'] else fp16 hidden [T,d].  Returns fp16 hidden [T,d], or fp32
    logits for ``logits_rows`` (default: all rows) if sw['last'].
    """
    if sw["first"]:
        h = np.asarray(sw["embed"], f16)[np.asarray(x)]
        if "pos_embed" in sw:
            h = (h.astype(f32) + np.asarray(sw["pos_embed"], f16)[positions].astype(f32)).astype(f16)
    else:
        h = np.asarray(x, f16)
    block = block_forward_llm_int8 if sw.get("quant", "w8a8") == "llm_int8" else block_forward
    for li, lw in enumerate(sw["layers"]):
        h = block(h, lw, cfg, cache, li, seq_ids, positions, cos, sin, trace)
    if not sw["last"]:
        return h
    if logits_rows is not None:
        h = h[logits_rows]
    hn = final_norm(h, sw["lnf_g"], sw.get("lnf_b"), cfg["eps"])
    lg = lm_head(hn, sw["lm_head"])
    if "lm_head_b" in sw:                    # Phi's lm_head carries a bias: one fp32 add per logit
        lg = (lg + np.asarray(sw["lm_head_b"], f32)[None, :]).astype(f32)
    return lg


def greedy_generate(stages, cfg, ids, max_new_tokens, max_seq, chunk=None):
    """Greedy decode through a list of stage-weight dicts (prefill, then one token a step).
    Returns (tokens [len(ids)+max_new_tokens], top-1 margins per generated token)."""
    hd = cfg["head_dim"]
    cos, sin = rope_table(hd, max_seq, cfg.get("rope_theta", 10000.0), cfg.get("rope_scaling", ()), cfg.get("rotary_dim", 0))
    caches = [KVCacheOracle(len(s["layers"]), 1, cfg["n_kv_heads"], max_seq, hd) for s in stages]
    toks = list(int(t) for t in ids)
    margins = []

    def run(x, pos):
        seq = np.zeros(len(pos), dtype=np.int64)
        for s, c in zip(stages, caches):
            x = stage_forward(x, s, cfg, c, seq, pos, cos, sin,
                              logits_rows=np.array([len(pos) - 1]) if s["last"] else None)
        return x

    T = len(toks)
    chunk = chunk or T
    logits = None
    for c0 in range(0, T, chunk):
        pos = np.arange(c0, min(T, c0 + chunk))
        logits = run(np.array(toks[c0:c0 + chunk]), pos)
    for _ in range(max_new_tokens):
        row = logits[-1]
        order = np.argsort(row)
        margins.append(float(row[order[-1]] - row[order[-2]]))
        nxt = int(order[-1])
        toks.append(nxt)
        if len(toks) >= max_seq:
            break
        logits = run(np.array([nxt]), np.array([len(toks) - 1]))
    return np.array(toks, dtype=np.int64), margins


# ----------------------------------------------------------------------------------------
# sampling: the reference's real generate kwargs (do_sample=True, temperature, top_k, top_p,
# repetition_penalty -- Code/Quantised Models/models_quant_updated.py:84-93, Code/others/config.yaml:3-7)
# ----------------------------------------------------------------------------------------
_U64 = np.uint64


def rng_uniform(seed: int, step: int, row: int) -> np.float32:
    """Counter-based uniform in [0,1): splitmix64 finaliser of (seed, step, row), top 24 bits.
    torch.multinomial's generator cannot be reproduced on the device, so the draw is build-defined;
    everything before the draw follows HF."""
    m = (1 << 64) - 1
    x = (int(seed) ^ ((int(step) * 0xD1B54A32D192ED03) & m) ^ (((int(row) + 1) * 0x9E3779B97F4A7C15) & m)) & m
    x ^= x >> 30
    x = (x * 0xBF58476D1CE4E5B9) & m
    x ^= x >> 27
    x = (x * 0x94D049BB133111EB) & m
    x ^= x >> 31
    return f32((x >> 40) * (1.0 / 16777216.0))


def sample_filter(logits_row, history, temperature, top_k, top_p, repetition_penalty):
    """HF's processor chain on one row of fp32 logits, in HF's order
    (transformers/generation/utils.py:1089,1214,1219,1223; logits_process.py):
      RepetitionPenaltyLogitsProcessor (:371-392)  score<0 ? score*p : score/p for every id in history
      TemperatureLogitsWarper (:297-299)           score / temperature
      TopKLogitsWarper (:581-586)                  keep score >= k-th largest (ties kept)
      TopPLogitsWarper (:521-533)                  ascending cumsum of softmax; drop cum <= 1 - top_p; keep >= 1
    Returns (token ids kept, their scores), ordered by (score desc, id asc); fp32, individually
    rounded operations, exp via exp_det -- the order the kernel b2_sample reproduces bit for bit."""
    x = np.asarray(logits_row, dtype=f32).copy()
    if history is not None and len(history) and repetition_penalty != 1.0:
        ids = np.unique(np.asarray(history, dtype=np.int64))
        s = x[ids]
        x[ids] = np.where(s < 0, (s * f32(repetition_penalty)).astype(f32), (s / f32(repetition_penalty)).astype(f32))
    x = (x / f32(temperature)).astype(f32)
    k = min(int(top_k), x.shape[0])
    kth = np.sort(x)[-k]
    cand = np.nonzero(x >= kth)[0]
    order = np.lexsort((cand, -x[cand].astype(np.float64)))       # score desc, id asc
    cand = cand[order]
    sc = x[cand]
    # softmax over the candidates, summed in ascending-score order (sequential fp32)
    e = exp_det((sc - sc[0]).astype(f32))
    z = f32(0.0)
    for v in e[::-1]:
        z = f32(z + v)
    cum = f32(0.0)
    keep = np.ones(len(cand), dtype=bool)
    thr = f32(f32(1.0) - f32(top_p))
    for j in range(len(cand) - 1, -1, -1):                        # ascending score order
        cum = f32(cum + f32(e[j] / z))
        if j != 0 and cum <= thr:
            keep[j] = False
    return cand[keep], sc[keep]


def sample_det(logits_row, history, temperature, top_k, top_p, repetition_penalty, seed, step, row):
    """One sampled token id: sample_filter, then inverse-CDF over the kept tokens in
    (score desc, id asc) order with the counter-based uniform."""
    ids, sc = sample_filter(logits_row, history, temperature, top_k, top_p, repetition_penalty)
    e = exp_det((sc - sc[0]).astype(f32))
    z = f32(0.0)
    for v in e:
        z = f32(z + v)
    target = f32(rng_uniform(seed, step, row) * z)
    acc = f32(0.0)
    for j in range(len(ids)):
        acc = f32(acc + e[j])
        if acc > target:
            return int(ids[j])
    return int(ids[-1])
