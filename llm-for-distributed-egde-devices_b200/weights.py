"""Layer-range shard construction: fp32 layer weights -> W8A8 stage weights, and the seeded synthetic models.
(Real checkpoints are read by loader.py, which feeds the same quantiser.)

Counterpart of the reference's loaders ``load_int8_model`` (Code/Quantised Models/models_quant_updated.py:25-40)
and ``load_model_int8_cpu`` (Code/C-DAC Server/try.py:198-206), which load a *whole* model and let
bitsandbytes / quantize_dynamic quantise every nn.Linear on load.  Here a stage loads only its
contiguous layer range and quantises offline with the SmoothQuant recipe of SURVEY.md §8(c):

  smoothing   s[k] = amax_act[k]^alpha / amax_w[k]^(1-alpha)           (alpha = 0.5)
  weights     W'[n,k] = W[n,k]*s[k];  s_w[n] = max_k|W'[n,k]|/127;  Wq = clamp(rint(W'/s_w), +-127)
  activations multiplied by 1/s[k] (vector `sm_*`) inside the fused norm/act quantise kernels.

No checkpoints or network exist in this environment, so weights are random-init from the config
under a seed (HF default N(0, 0.02); embeddings x8 and tied to lm_head to widen greedy margins,
SURVEY.md §8d).  Per-layer generators make a stage able to build only its own layers.
This is host-side set-up code (numpy/torch CPU or torch CUDA), not the hot path.
"""
from __future__ import annotations

import numpy as np
import torch

from .config import ModelConfig

ALPHA = 0.5


def _gen(seed: int, *tags: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed((seed * 1000003 + sum((t + 1) * 7919 ** (i + 1) for i, t in enumerate(tags))) % (2 ** 63 - 1))
    return g


def head_slots(cfg: ModelConfig) -> np.ndarray:
    """Kernel layout of one attention head: slot j of the head_dim-wide head holds model dim head_slots[j], -1 = zero.

    The kernels rotate the pairs (j, j + head_dim/2) of a head and know nothing of partial rotary or odd head
    widths.  A model that rotates only its first rotary_dim dims (pairs (j, j + rotary_dim/2): Phi, GPT-NeoX) or
    whose heads are not a multiple of 32 wide (Phi-2: 80) is brought into that form when its weights are loaded:
    the rotary pairs go to slots (j, j + head_dim/2), j < rotary_dim/2, the other dims fill the remaining slots
    (their rope-table entries are cos = 1, sin = 0: an exact no-op), the rest is zero padding.  q and k are permuted
    alike, so every q.k product is unchanged; v and the o_proj columns take the same permutation."""
    hd, hm = cfg.head_dim, cfg.hd_model
    rd = cfg.rotary_dim or hm
    half, r2 = hd // 2, rd // 2
    if rd > hm or rd % 2 or hm > hd or r2 > half:
        raise ValueError("bad rotary_dim / head_dim_model")
    src = np.full((hd,), -1, dtype=np.int64)
    src[:r2] = np.arange(r2)
    src[half:half + r2] = np.arange(r2, rd)
    free = [j for j in range(hd) if not (j < r2 or half <= j < half + r2)]
    rest = np.arange(rd, hm)
    src[free[: len(rest)]] = rest
    return src


def to_kernel_layout(cfg: ModelConfig, w: dict) -> dict:
    """Layer weights in the MODEL's head layout (wqkv [(H+2KV)*hd_model, d], wo [d, H*hd_model], b_qkv) -> the
    kernels' layout (head_dim wide heads, see head_slots).  The identity for the Llama / GPT-2 families."""
    src = head_slots(cfg)
    hm, hd = cfg.hd_model, cfg.head_dim
    if hm == hd and np.array_equal(src, np.arange(hd)):
        return w
    H, KV = cfg.n_heads, cfg.n_kv_heads
    idx = torch.from_numpy(np.where(src >= 0, src, 0))
    keep = torch.from_numpy((src >= 0).astype(np.float32))

    def rows(t, nh):           # [nh*hm, ...] -> [nh*hd, ...]
        t = t.reshape(nh, hm, *t.shape[1:])[:, idx]
        return (t * keep.reshape(1, hd, *([1] * (t.dim() - 2)))).reshape(nh * hd, *t.shape[2:])
    out = dict(w)
    q, k, v = w["wqkv"][: H * hm], w["wqkv"][H * hm:(H + KV) * hm], w["wqkv"][(H + KV) * hm:]
    out["wqkv"] = torch.cat([rows(q, H), rows(k, KV), rows(v, KV)], 0).contiguous()
    if "b_qkv" in w:
        b = w["b_qkv"]
        out["b_qkv"] = torch.cat([rows(b[: H * hm], H), rows(b[H * hm:(H + KV) * hm], KV), rows(b[(H + KV) * hm:], KV)], 0).contiguous()
    out["wo"] = rows(w["wo"].t().contiguous(), H).t().contiguous()
    return out


def init_layer_fp32(cfg: ModelConfig, layer: int, seed: int = 0, layout: str = "kernel") -> dict:
    """fp32 weights of one decoder layer (HF naming collapsed: wqkv = [q;k;v], wgu = [gate;up]).
    Drawn in the model's own head layout; layout="kernel" (default) applies to_kernel_layout (the identity for
    Llama / GPT-2 shapes), layout="model" returns them as a HF module of that family would hold them."""
    g = _gen(seed, layer)
    d, F = cfg.d_model, cfg.ffn
    hm = cfg.hd_model
    std = 0.02

    def rn(*shape):
        return torch.randn(*shape, generator=g, dtype=torch.float32) * std

    qkv_rows = (cfg.n_heads + 2 * cfg.n_kv_heads) * hm
    w = dict(
        ln1_g=1.0 + 0.1 * torch.randn(d, generator=g), ln2_g=1.0 + 0.1 * torch.randn(d, generator=g),
        wqkv=rn(qkv_rows, d), wo=rn(d, cfg.n_heads * hm), wgu=rn(cfg.up_out, d), wd=rn(d, F))
    if cfg.norm == "layernorm":
        w["ln1_b"] = 0.02 * torch.randn(d, generator=g)
        w["ln2_b"] = 0.02 * torch.randn(d, generator=g)
    if cfg.bias:
        w["b_qkv"] = rn(qkv_rows)
        w["b_o"] = rn(d)
        w["b_gu"] = rn(cfg.up_out)
        w["b_d"] = rn(d)
    if cfg.shared_ln:          # one LayerNorm feeds attention and MLP: kept twice so that every code path stays uniform
        w["ln2_g"] = w["ln1_g"].clone()
        if "ln1_b" in w:
            w["ln2_b"] = w["ln1_b"].clone()
    return to_kernel_layout(cfg, w) if layout == "kernel" else w


def init_embed_fp32(cfg: ModelConfig, seed: int = 0) -> dict:
    g = _gen(seed, 100003)
    e = torch.randn(cfg.vocab, cfg.d_model, generator=g, dtype=torch.float32) * (0.02 * 8.0)
    out = dict(embed=e, lnf_g=1.0 + 0.1 * torch.randn(cfg.d_model, generator=g))
    if cfg.norm == "layernorm":
        out["lnf_b"] = 0.02 * torch.randn(cfg.d_model, generator=g)
    if not cfg.rope:
        out["pos_embed"] = torch.randn(cfg.max_pos, cfg.d_model, generator=g, dtype=torch.float32) * 0.02
    # untied heads (Phi): a separate matrix, but close to the embedding table, so that greedy margins stay as wide as
    # in the tied models (a purely random head leaves top-1 margins at the size of the W8A8 noise)
    out["lm_head"] = e if cfg.tie_embeddings else e + torch.randn(cfg.vocab, cfg.d_model, generator=g) * 0.02
    if cfg.head_bias:
        out["lm_head_b"] = torch.randn(cfg.vocab, generator=g) * 0.02
    return out


def _smooth_vec(W: torch.Tensor, act_amax: torch.Tensor | None, alpha: float) -> torch.Tensor:
    """s[k] for weight W [N,K]; act_amax None -> 1 (weight-only statistics)."""
    w_amax = W.abs().amax(dim=0).clamp_min(1e-5)
    a = torch.ones_like(w_amax) if act_amax is None else act_amax.to(w_amax).clamp_min(1e-5)
    s = a.pow(alpha) / w_amax.pow(1.0 - alpha)
    return s.clamp(1e-4, 1e4).to(torch.float32)


def quantize_weight(W: torch.Tensor, s: torch.Tensor | None):
    """W [N,K] fp32 (any device) -> (Wq int8 [N,K], s_w fp32 [N]).  Same formula as
    oracle/w8a8.py:quantize_weight (checked in tests/test_oracle_cpu.py)."""
    W = W.to(torch.float32)
    if s is not None:
        W = W * s[None, :]
    amax = W.abs().amax(dim=1)
    safe = torch.where(amax > 0, amax, torch.ones_like(amax))
    step = safe / 127.0
    scale = torch.where(amax > 0, amax / 127.0, torch.zeros_like(amax))
    Wq = torch.clamp(torch.round(W / step[:, None]), -127, 127).to(torch.int8)
    return Wq.contiguous(), scale.to(torch.float32).contiguous()


# LLM.int8() mode: real LLMs carry a few "outlier" feature channels; the synthetic models get them as
# large norm gains on OUTLIER_CHANNELS channels per norm, so the fp16 side path of the decomposition runs.
OUTLIER_CHANNELS = 3
OUTLIER_GAIN = 9.0


def outlier_gain_channels(d: int, layer: int, which: int):
    """The channels of norm `which` (0 = ln1, 1 = ln2) of `layer` that get OUTLIER_GAIN (deterministic)."""
    return [(37 * layer + 101 * which + 53 * j + 11) % d for j in range(OUTLIER_CHANNELS)]


def add_outlier_gains(w: dict, d: int, layer: int) -> dict:
    for which, key in enumerate(("ln1_g", "ln2_g")):
        for c in outlier_gain_channels(d, layer, which):
            w[key][c] = OUTLIER_GAIN
    return w


def quantize_layer(cfg: ModelConfig, w: dict, act_stats: dict | None = None, alpha: float = ALPHA,
                   smooth: bool = True) -> dict:
    """fp32 layer dict -> W8A8 layer dict (torch tensors on w's device).
    smooth = False: plain per-output-row absmax int8 (bitsandbytes' weight format), no sm_* vectors.

    act_stats (optional): per-input-channel activation absmax for 'qkv', 'o', 'gu', 'd' gathered by
    a calibration pass; absent -> weight-only smoothing statistics.
    """
    st = act_stats or {}
    out = {}
    for name, wk, smk, sk in (("qkv", "wqkv", "sm_qkv", "s_qkv"), ("o", "wo", "sm_o", "s_o"),
                              ("gu", "wgu", "sm_gu", "s_gu"), ("d", "wd", "sm_d", "s_d")):
        if not smooth:
            out[wk], out[sk] = quantize_weight(w[wk], None)
            continue
        s = _smooth_vec(w[wk], st.get(name), alpha)
        out[wk], out[sk] = quantize_weight(w[wk], s)
        out[smk] = (1.0 / s).to(torch.float32).contiguous()
    for k in ("ln1_g", "ln2_g", "ln1_b", "ln2_b", "b_qkv", "b_o", "b_gu", "b_d"):
        if k in w:
            out[k] = w[k].to(torch.float16).contiguous()
    return out


def build_stage_weights(cfg: ModelConfig, layer_begin: int, layer_end: int, first: bool, last: bool,
                        seed: int = 0, device="cpu", act_stats: list | None = None, quant: str = "w8a8") -> dict:
    """Seeded synthetic W8A8 shard for layers [layer_begin, layer_end) (CPU generation, exact
    and reproducible; used by the parity tests so the oracle sees identical weights)."""
    sw = dict(first=first, last=last, layers=[], quant=quant)
    for li in range(layer_begin, layer_end):
        w32 = init_layer_fp32(cfg, li, seed)
        if quant == "llm_int8":
            add_outlier_gains(w32, cfg.d_model, li)
        lw = quantize_layer(cfg, w32, None if act_stats is None else act_stats[li], smooth=(quant != "llm_int8"))
        sw["layers"].append({k: v.to(device) for k, v in lw.items()})
    if first or last:
        e = init_embed_fp32(cfg, seed)
        if first:
            sw["embed"] = e["embed"].to(torch.float16).to(device)
            if "pos_embed" in e:
                sw["pos_embed"] = e["pos_embed"].to(torch.float16).to(device)
        if last:
            sw["lnf_g"] = e["lnf_g"].to(torch.float16).to(device)
            if "lnf_b" in e:
                sw["lnf_b"] = e["lnf_b"].to(torch.float16).to(device)
            sw["lm_head"] = sw["embed"] if (first and cfg.tie_embeddings) else e["lm_head"].to(torch.float16).to(device)
            if "lm_head_b" in e:
                sw["lm_head_b"] = e["lm_head_b"].to(torch.float32).to(device)
    return sw


def build_stage_weights_fast(cfg: ModelConfig, layer_begin: int, layer_end: int, first: bool, last: bool,
                             seed: int = 0, device="cuda", quant: str = "w8a8", calib_ids=None) -> dict:
    """Same structure generated directly on `device` with torch's device RNG (benchmarks: the
    7B/13B shapes would take minutes through the CPU generator).  Statistically identical,
    not bit-identical to build_stage_weights.

    calib_ids (1-D token ids): SmoothQuant calibration on the way -- the fp32 layers are run over that prompt as
    they are drawn (loader._layer_fp32; layers below layer_begin are drawn too, forwarded and dropped), and every
    linear of this stage is smoothed with its input's per-channel activation maxima (alpha = 0.5) instead of
    weight-only statistics.  Every rank of a pipeline derives the same statistics for the layers it owns."""
    dev = torch.device(device)
    g = torch.Generator(device=dev)
    sw = dict(first=first, last=last, layers=[], quant=quant)
    d, F = cfg.d_model, cfg.ffn
    calib = calib_ids is not None and quant == "w8a8"
    hcal = None
    if calib and (cfg.bias or cfg.norm != "rms" or not cfg.gated):
        raise ValueError("calibration inside the fast builder covers the bias-free RMSNorm families (use loader.calibrate)")
    if calib:
        from . import loader
        g.manual_seed(seed * 1000003 + 100003)
        e0 = (torch.randn(cfg.vocab, cfg.d_model, generator=g, dtype=torch.float32, device=dev) * 0.16).to(torch.float16)
        hcal = e0[torch.as_tensor(np.asarray(calib_ids), dtype=torch.int64, device=dev)].to(torch.float32)
        del e0
        cos_c, sin_c = rope_table(cfg.head_dim, cfg.max_pos, cfg.rope_theta, cfg.rope_scaling, cfg.rotary_dim)
        cos_c, sin_c = cos_c.to(dev), sin_c.to(dev)
    for li in range(0 if calib else layer_begin, layer_end):
        g.manual_seed(seed * 1000003 + li)

        def rn(*shape, std=0.02):
            return torch.randn(*shape, generator=g, dtype=torch.float32, device=dev) * std

        w = dict(ln1_g=1.0 + rn(d, std=0.1), ln2_g=1.0 + rn(d, std=0.1))
        lw = {}
        if calib:
            # draw the four matrices in the usual order, run the fp32 layer over the calibration prompt, then quantise
            # with the statistics just gathered (gains rounded to fp16 first, as the kernels will read them)
            for wk, shape in (("wqkv", (cfg.qkv_out, d)), ("wo", (d, cfg.n_heads * cfg.head_dim)),
                              ("wgu", (cfg.up_out, d)), ("wd", (d, F))):
                w[wk] = rn(*shape)
            wf = dict(w, ln1_g=w["ln1_g"].to(torch.float16).float(), ln2_g=w["ln2_g"].to(torch.float16).float())
            st = {}
            hcal = loader._layer_fp32(cfg, wf, hcal, cos_c, sin_c, st)
            if li < layer_begin:
                continue
            for wk, name in (("wqkv", "qkv"), ("wo", "o"), ("wgu", "gu"), ("wd", "d")):
                sfx = wk[1:]
                s_ = _smooth_vec(w[wk], st[name], ALPHA)
                lw[wk], lw["s_" + sfx] = quantize_weight(w[wk], s_)
                lw["sm_" + sfx] = (1.0 / s_).contiguous()
                del w[wk]
        for wk, shape in (() if calib else (("wqkv", (cfg.qkv_out, d)), ("wo", (d, cfg.n_heads * cfg.head_dim)),
                                            ("wgu", (cfg.up_out, d)), ("wd", (d, F)))):
            W = rn(*shape)
            sfx = wk[1:]
            if quant == "llm_int8":
                lw[wk], lw["s_" + sfx] = quantize_weight(W, None)
            else:
                s = _smooth_vec(W, None, ALPHA)
                lw[wk], lw["s_" + sfx] = quantize_weight(W, s)
                lw["sm_" + sfx] = (1.0 / s).contiguous()
            del W
        if quant == "llm_int8":
            add_outlier_gains(w, d, li)
        lw["ln1_g"] = w["ln1_g"].to(torch.float16)
        lw["ln2_g"] = w["ln2_g"].to(torch.float16)
        if cfg.norm == "layernorm":
            lw["ln1_b"] = rn(d).to(torch.float16)
            lw["ln2_b"] = rn(d).to(torch.float16)
        if cfg.bias:
            lw["b_qkv"] = rn(cfg.qkv_out).to(torch.float16)
            lw["b_o"] = rn(d).to(torch.float16)
            lw["b_gu"] = rn(cfg.up_out).to(torch.float16)
            lw["b_d"] = rn(d).to(torch.float16)
        sw["layers"].append(lw)
    if first or last:
        g.manual_seed(seed * 1000003 + 100003)
        e = (torch.randn(cfg.vocab, cfg.d_model, generator=g, dtype=torch.float32, device=dev) * 0.16).to(torch.float16)
        if first:
            sw["embed"] = e
            if not cfg.rope:
                sw["pos_embed"] = (torch.randn(cfg.max_pos, d, generator=g, device=dev) * 0.02).to(torch.float16)
        if last:
            sw["lnf_g"] = (1.0 + 0.1 * torch.randn(d, generator=g, device=dev)).to(torch.float16)
            if cfg.norm == "layernorm":
                sw["lnf_b"] = (0.02 * torch.randn(d, generator=g, device=dev)).to(torch.float16)
            sw["lm_head"] = e
    return sw


def head_rows_fp16(cfg: ModelConfig, seed: int, lo: int, hi: int, device, fast: bool) -> torch.Tensor:
    """Rows [lo, hi) of the seeded model's lm_head, fp16 on `device`, exactly as build_stage_weights (fast = False)
    or build_stage_weights_fast gives them to the last stage: any stage of a pipeline can hold a slice of the head
    (DistPipeline's lm_head shared by all stages)."""
    if fast:
        dev = torch.device(device)
        g = torch.Generator(device=dev)
        g.manual_seed(seed * 1000003 + 100003)
        e = (torch.randn(cfg.vocab, cfg.d_model, generator=g, dtype=torch.float32, device=dev) * 0.16).to(torch.float16)
        return e[lo:hi].clone()
    return init_embed_fp32(cfg, seed)["lm_head"][lo:hi].to(torch.float16).to(device)


def to_numpy(sw: dict) -> dict:
    """Stage weights as plain numpy arrays (the form oracle/w8a8.py consumes)."""
    def cv(v):
        return v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v
    out = {k: cv(v) for k, v in sw.items() if k != "layers"}
    out["layers"] = [{k: cv(v) for k, v in lw.items()} for lw in sw["layers"]]
    return out


def rope_table(head_dim: int, max_pos: int, theta: float = 10000.0, scaling: tuple = (), rotary_dim: int = 0):
    """cos/sin fp32 [max_pos, head_dim/2]: angle formed in fp32 as HF does
    (modeling_llama.py:100-126), cos/sin evaluated in fp64 and rounded once.  Host-side table; the
    kernels only read it, so device and oracle use bit-identical values.
    scaling = ("llama3", factor, low_freq_factor, high_freq_factor, original_max_pos): the frequency
    remap of Llama-3.1/3.2 (transformers modeling_rope_utils.py `_compute_llama3_parameters`), in fp32 like HF."""
    rd = rotary_dim or head_dim          # partial rotary (kernel layout, head_slots): the first rd/2 pairs rotate
    i = np.arange(0, rd, 2, dtype=np.float64)
    inv_freq = (1.0 / (theta ** (i / rd))).astype(np.float32)
    if scaling:
        kind, factor, low_f, high_f, old_ctx = scaling
        if kind != "llama3":
            raise ValueError(f"rope scaling {kind!r} is not supported")
        f32 = np.float32
        wavelen = (f32(2.0 * np.pi) / inv_freq).astype(f32)
        low_wl, high_wl = f32(old_ctx / low_f), f32(old_ctx / high_f)
        scaled = np.where(wavelen > low_wl, inv_freq / f32(factor), inv_freq).astype(f32)
        smooth = ((f32(old_ctx) / wavelen - f32(low_f)) / f32(high_f - low_f)).astype(f32)
        mid = ((f32(1.0) - smooth) * scaled / f32(factor) + smooth * scaled).astype(f32)
        inv_freq = np.where((wavelen >= high_wl) & (wavelen <= low_wl), mid, scaled).astype(f32)
    inv_freq = inv_freq.astype(np.float64)
    ang = (np.arange(max_pos, dtype=np.float64)[:, None] * inv_freq[None, :]).astype(np.float32).astype(np.float64)
    cos, sin = np.ones((max_pos, head_dim // 2), np.float32), np.zeros((max_pos, head_dim // 2), np.float32)
    cos[:, : rd // 2], sin[:, : rd // 2] = np.cos(ang).astype(np.float32), np.sin(ang).astype(np.float32)
    return torch.from_numpy(cos), torch.from_numpy(sin)
