"""Checkpoint loading for a contiguous layer range, and the SmoothQuant calibration pass.

Counterpart of the reference's loaders -- ``AutoModelForCausalLM.from_pretrained(model_path, ...)`` in
``load_int8_model`` (Code/Quantised Models/models_quant_updated.py:25-40) and ``load_model_int8_cpu``
(Code/C-DAC Server/try.py:198-206) -- restricted to what a pipeline stage needs: it reads only the tensors of
its own layers (plus the embedding on the first stage, the final norm and lm_head on the last) from the
``*.safetensors`` shards (or ``pytorch_model*.bin``) of a Hugging Face model directory and hands them to the
same offline W8A8 quantiser the synthetic models go through (weights.quantize_layer).

A directory that has a config but no weight file is an error here, never a silently random model.

Calibration (alpha = 0.5 on activation maxima): ``calibrate`` runs the UNQUANTISED model
layer by layer in fp32 over a few prompts and records, per linear, the per-input-channel absmax of its input
(qkv: normed hidden state; o: attention output; gate/up: normed hidden state; down: act(gate)*up).  Set-up code in
plain torch -- it runs once per model, off the hot path, and never under a timed region.
"""
from __future__ import annotations

import glob
import json
import os

import numpy as np
import torch

from .config import ModelConfig, get_config
from . import weights

STATS_FILE = "b200stage_act_stats.npz"      # written next to a checkpoint by save_stats, picked up by Stage.load


# --------------------------------------------------------------------------------------------------
# tensor access
# --------------------------------------------------------------------------------------------------
def weight_files(model_dir: str) -> list[str]:
    fs = sorted(glob.glob(os.path.join(model_dir, "*.safetensors")))
    return fs or sorted(glob.glob(os.path.join(model_dir, "pytorch_model*.bin")))


def is_checkpoint_dir(spec) -> bool:
    return isinstance(spec, (str, os.PathLike)) and os.path.isdir(spec)


class TensorSource:
    """name -> fp32 CPU tensor, reading each shard lazily and only the tensors asked for (safetensors), or
    whole shards on first use (.bin)."""

    def __init__(self, model_dir: str):
        self.dir = model_dir
        files = weight_files(model_dir)
        if not files:
            raise FileNotFoundError(
                f"{model_dir} has no *.safetensors or pytorch_model*.bin: refusing to synthesise weights for a model "
                "directory (pass a preset name or a ModelConfig for a seeded synthetic model)")
        self.where = {}
        self._bin = {}
        self.safe = files[0].endswith(".safetensors")
        idx = glob.glob(os.path.join(model_dir, "*.index.json"))
        if idx:
            with open(idx[0]) as f:
                self.where = {k: os.path.join(model_dir, v) for k, v in json.load(f)["weight_map"].items()}
        elif self.safe:
            from safetensors import safe_open
            for fn in files:
                with safe_open(fn, framework="pt") as f:
                    for k in f.keys():
                        self.where[k] = fn
        else:
            for fn in files:
                self._bin[fn] = torch.load(fn, map_location="cpu", weights_only=True)
                for k in self._bin[fn]:
                    self.where[k] = fn

    def has(self, name: str) -> bool:
        return name in self.where

    def get(self, *names: str) -> torch.Tensor:
        """The first of `names` that exists (checkpoints differ in their prefixes)."""
        for name in names:
            fn = self.where.get(name)
            if fn is None:
                continue
            if fn.endswith(".safetensors"):
                from safetensors import safe_open
                with safe_open(fn, framework="pt") as f:
                    return f.get_tensor(name).to(torch.float32)
            if fn not in self._bin:
                self._bin[fn] = torch.load(fn, map_location="cpu", weights_only=True)
            return self._bin[fn][name].to(torch.float32)
        raise KeyError(f"none of {names} in {self.dir}")


def _family(model_dir: str) -> str:
    with open(os.path.join(model_dir, "config.json")) as f:
        mt = json.load(f).get("model_type", "llama")
    if mt == "gpt2":
        return "gpt2"
    if mt in ("llama", "mistral"):
        return "llama"
    if mt == "phi":
        return "phi"
    if mt == "gpt_neox":
        return "neox"
    raise ValueError(f"model_type {mt!r}: no tensor-name map for this family (llama, mistral, phi, gpt_neox and gpt2 are supported)")


# --------------------------------------------------------------------------------------------------
# HF names -> the stage's fp32 layer dict (the layout of weights.init_layer_fp32)
# --------------------------------------------------------------------------------------------------
def load_layer_fp32(src: TensorSource, cfg: ModelConfig, layer: int, family: str) -> dict:
    if family == "llama":
        p = f"model.layers.{layer}."
        g = lambda n: src.get(p + n, p.replace("model.", "", 1) + n)
        w = dict(ln1_g=g("input_layernorm.weight"), ln2_g=g("post_attention_layernorm.weight"),
                 wqkv=torch.cat([g("self_attn.q_proj.weight"), g("self_attn.k_proj.weight"), g("self_attn.v_proj.weight")], 0),
                 wo=g("self_attn.o_proj.weight"),
                 wgu=torch.cat([g("mlp.gate_proj.weight"), g("mlp.up_proj.weight")], 0), wd=g("mlp.down_proj.weight"))
        if src.has(p + "self_attn.q_proj.bias"):
            raise ValueError("attention biases are not supported for the llama family here")
    elif family == "phi":        # Phi-2: biases everywhere, ONE LayerNorm per layer, heads in the model's 80-wide layout
        p = f"model.layers.{layer}."
        g = lambda n: src.get(p + n, p.replace("model.", "", 1) + n)
        w = dict(ln1_g=g("input_layernorm.weight"), ln1_b=g("input_layernorm.bias"),
                 wqkv=torch.cat([g("self_attn.q_proj.weight"), g("self_attn.k_proj.weight"), g("self_attn.v_proj.weight")], 0),
                 b_qkv=torch.cat([g("self_attn.q_proj.bias"), g("self_attn.k_proj.bias"), g("self_attn.v_proj.bias")], 0),
                 wo=g("self_attn.dense.weight"), b_o=g("self_attn.dense.bias"),
                 wgu=g("mlp.fc1.weight"), b_gu=g("mlp.fc1.bias"), wd=g("mlp.fc2.weight"), b_d=g("mlp.fc2.bias"))
        w["ln2_g"], w["ln2_b"] = w["ln1_g"].clone(), w["ln1_b"].clone()
        w = weights.to_kernel_layout(cfg, w)
    elif family == "neox":       # GPT-NeoX / Pythia: query_key_value is stored head by head as [q_h | k_h | v_h]
        p = f"gpt_neox.layers.{layer}."
        g = lambda n: src.get(p + n, p.replace("gpt_neox.", "", 1) + n)
        H, hm = cfg.n_heads, cfg.hd_model
        qkv = g("attention.query_key_value.weight").view(H, 3, hm, cfg.d_model)
        bqkv = g("attention.query_key_value.bias").view(H, 3, hm)
        w = dict(ln1_g=g("input_layernorm.weight"), ln1_b=g("input_layernorm.bias"),
                 ln2_g=g("post_attention_layernorm.weight"), ln2_b=g("post_attention_layernorm.bias"),
                 wqkv=torch.cat([qkv[:, i].reshape(H * hm, cfg.d_model) for i in range(3)], 0).contiguous(),
                 b_qkv=torch.cat([bqkv[:, i].reshape(H * hm) for i in range(3)], 0).contiguous(),
                 wo=g("attention.dense.weight"), b_o=g("attention.dense.bias"),
                 wgu=g("mlp.dense_h_to_4h.weight"), b_gu=g("mlp.dense_h_to_4h.bias"),
                 wd=g("mlp.dense_4h_to_h.weight"), b_d=g("mlp.dense_4h_to_h.bias"))
        w = weights.to_kernel_layout(cfg, w)
    else:   # GPT-2: Conv1D weights are [in, out]
        p = f"transformer.h.{layer}."
        g = lambda n: src.get(p + n, p.replace("transformer.", "", 1) + n)
        w = dict(ln1_g=g("ln_1.weight"), ln1_b=g("ln_1.bias"), ln2_g=g("ln_2.weight"), ln2_b=g("ln_2.bias"),
                 wqkv=g("attn.c_attn.weight").t().contiguous(), b_qkv=g("attn.c_attn.bias"),
                 wo=g("attn.c_proj.weight").t().contiguous(), b_o=g("attn.c_proj.bias"),
                 wgu=g("mlp.c_fc.weight").t().contiguous(), b_gu=g("mlp.c_fc.bias"),
                 wd=g("mlp.c_proj.weight").t().contiguous(), b_d=g("mlp.c_proj.bias"))
    want = dict(wqkv=(cfg.qkv_out, cfg.d_model), wo=(cfg.d_model, cfg.n_heads * cfg.head_dim),
                wgu=(cfg.up_out, cfg.d_model), wd=(cfg.d_model, cfg.ffn))
    for k, shp in want.items():
        if tuple(w[k].shape) != shp:
            raise ValueError(f"layer {layer}: {k} has shape {tuple(w[k].shape)}, the config says {shp}")
    return w


def load_embed_fp32(src: TensorSource, cfg: ModelConfig, family: str, first: bool, last: bool) -> dict:
    out = {}
    if family == "llama":
        emb = ("model.embed_tokens.weight", "embed_tokens.weight")
        if first:
            out["embed"] = src.get(*emb)
        if last:
            out["lnf_g"] = src.get("model.norm.weight", "norm.weight")
            out["lm_head"] = src.get("lm_head.weight") if src.has("lm_head.weight") else src.get(*emb)
    elif family == "neox":
        emb = ("gpt_neox.embed_in.weight", "embed_in.weight")
        if first:
            out["embed"] = src.get(*emb)
        if last:
            out["lnf_g"] = src.get("gpt_neox.final_layer_norm.weight", "final_layer_norm.weight")
            out["lnf_b"] = src.get("gpt_neox.final_layer_norm.bias", "final_layer_norm.bias")
            out["lm_head"] = src.get("embed_out.weight") if src.has("embed_out.weight") else src.get(*emb)
    elif family == "phi":
        if first:
            out["embed"] = src.get("model.embed_tokens.weight", "embed_tokens.weight")
        if last:
            out["lnf_g"] = src.get("model.final_layernorm.weight", "final_layernorm.weight")
            out["lnf_b"] = src.get("model.final_layernorm.bias", "final_layernorm.bias")
            out["lm_head"] = src.get("lm_head.weight")
            out["lm_head_b"] = src.get("lm_head.bias")
    else:
        emb = ("transformer.wte.weight", "wte.weight")
        if first:
            out["embed"] = src.get(*emb)
            out["pos_embed"] = src.get("transformer.wpe.weight", "wpe.weight")
        if last:
            out["lnf_g"] = src.get("transformer.ln_f.weight", "ln_f.weight")
            out["lnf_b"] = src.get("transformer.ln_f.bias", "ln_f.bias")
            out["lm_head"] = src.get("lm_head.weight") if src.has("lm_head.weight") else src.get(*emb)
    return out


def build_stage_weights_from_checkpoint(model_dir: str, layer_begin: int, layer_end: int, first: bool, last: bool,
                                        device="cuda", act_stats: list | None = None, quant: str = "w8a8") -> dict:
    """W8A8 (or LLM.int8()) shard of layers [layer_begin, layer_end) of the checkpoint in `model_dir`.
    act_stats: per absolute layer index, the calibration dict of `calibrate` (None -> the file save_stats wrote next
    to the checkpoint if there is one, else weight-only smoothing statistics)."""
    cfg = get_config(model_dir)
    fam = _family(model_dir)
    src = TensorSource(model_dir)
    if act_stats is None:
        act_stats = load_stats(model_dir)
    sw = dict(first=first, last=last, layers=[], quant=quant, source=os.path.abspath(model_dir))
    for li in range(layer_begin, layer_end):
        w32 = load_layer_fp32(src, cfg, li, fam)
        lw = weights.quantize_layer(cfg, w32, None if act_stats is None else act_stats[li], smooth=(quant != "llm_int8"))
        sw["layers"].append({k: v.to(device) for k, v in lw.items()})
    if first or last:
        e = load_embed_fp32(src, cfg, fam, first, last)
        for k, v in e.items():
            sw[k] = v.to(torch.float32 if k == "lm_head_b" else torch.float16).to(device)
        if first and last and cfg.tie_embeddings:
            sw["lm_head"] = sw["embed"]
    return sw


def load_head_rows(model_dir: str, lo: int, hi: int, device="cuda") -> torch.Tensor:
    """Rows [lo, hi) of the checkpoint's lm_head (the embedding table where the model ties them), fp16 on `device`."""
    cfg = get_config(model_dir)
    e = load_embed_fp32(TensorSource(model_dir), cfg, _family(model_dir), False, True)
    return e["lm_head"][lo:hi].to(torch.float16).to(device)


# --------------------------------------------------------------------------------------------------
# calibration: per-input-channel activation maxima of the four linears of every layer
# --------------------------------------------------------------------------------------------------
def _norm(x, g, b, eps, kind):
    if kind == "layernorm":
        return torch.nn.functional.layer_norm(x, (x.shape[-1],), g, b, eps)
    return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps) * g


def _rope(x, cos, sin):          # x [T, heads, hd]; rotate_half convention (modeling_llama.py:128-153)
    h = x.shape[-1] // 2
    x1, x2 = x[..., :h], x[..., h:]
    c, s = cos[:, None, :], sin[:, None, :]
    return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], -1)


def _layer_fp32(cfg: ModelConfig, w: dict, h, cos, sin, stats: dict):
    """fp32 forward of one decoder layer over one prompt h [T,d] (full causal / sliding-window attention),
    updating `stats` with the channel maxima of the four linear inputs.  Returns the layer output."""
    T = h.shape[0]
    H, KV, hd, F = cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.ffn

    def seen(key, x):
        m = x.abs().amax(0)
        stats[key] = m if key not in stats else torch.maximum(stats[key], m)
    x = _norm(h, w["ln1_g"], w.get("ln1_b"), cfg.eps, cfg.norm)
    seen("qkv", x)
    qkv = x @ w["wqkv"].t()
    if "b_qkv" in w:
        qkv = qkv + w["b_qkv"]
    q = qkv[:, : H * hd].view(T, H, hd)
    k = qkv[:, H * hd:(H + KV) * hd].view(T, KV, hd)
    v = qkv[:, (H + KV) * hd:].view(T, KV, hd)
    if cfg.rope:
        q, k = _rope(q, cos[:T], sin[:T]), _rope(k, cos[:T], sin[:T])
    rep = H // KV
    k, v = k.repeat_interleave(rep, 1), v.repeat_interleave(rep, 1)
    sc = torch.einsum("thd,shd->hts", q, k) / float(np.sqrt(hd))
    i = torch.arange(T, device=h.device)
    mask = i[None, :] > i[:, None]
    if cfg.window:
        mask = mask | (i[None, :] <= i[:, None] - cfg.window)
    sc = sc.masked_fill(mask[None], float("-inf"))
    a = torch.einsum("hts,shd->thd", torch.softmax(sc, -1), v).reshape(T, H * hd)
    seen("o", a)
    h1 = h + a @ w["wo"].t() + (w["b_o"] if "b_o" in w else 0)
    x = _norm(h1, w["ln2_g"], w.get("ln2_b"), cfg.eps, cfg.norm)
    seen("gu", x)
    gu = x @ w["wgu"].t()
    if "b_gu" in w:
        gu = gu + w["b_gu"]
    act = {"silu": torch.nn.functional.silu, "gelu": torch.nn.functional.gelu,
           "gelu_new": lambda t: torch.nn.functional.gelu(t, approximate="tanh")}[cfg.act]
    y = act(gu[:, :F]) * gu[:, F:] if cfg.gated else act(gu)
    seen("d", y)
    return h1 + y @ w["wd"].t() + (w["b_d"] if "b_d" in w else 0)


@torch.no_grad()
def calibrate(model, prompts, seed: int = 0, device="cpu", layers=None, layer_fn=None) -> list:
    """Activation statistics for SmoothQuant: run the fp32 model over `prompts` (a list of 1-D id arrays) and return,
    for every layer, {'qkv','o','gu','d': per-input-channel absmax [K] fp32}.  `model` is a checkpoint directory or
    a preset / ModelConfig (the seeded synthetic weights of weights.init_layer_fp32).  One layer's fp32 weights are
    resident at a time.  `layers` limits the pass to the first `layers` layers (a stage needs no statistics beyond
    its own last layer).  layer_fn(li) -> fp32 layer dict overrides where the layers come from (tests)."""
    cfg = get_config(model)
    dev = torch.device(device)
    if layer_fn is not None:
        get_layer = layer_fn
        emb = weights.init_embed_fp32(cfg, seed)
    elif is_checkpoint_dir(model):
        fam, src = _family(model), TensorSource(model)
        get_layer = lambda li: load_layer_fp32(src, cfg, li, fam)
        emb = load_embed_fp32(src, cfg, fam, True, False)
    else:
        get_layer = lambda li: weights.init_layer_fp32(cfg, li, seed)
        emb = weights.init_embed_fp32(cfg, seed)
    cos, sin = weights.rope_table(cfg.head_dim, cfg.max_pos, cfg.rope_theta, cfg.rope_scaling)
    cos, sin = cos.to(dev), sin.to(dev)
    E = emb["embed"].to(torch.float16).to(torch.float32).to(dev)          # the stage stores the table in fp16
    hs = []
    for ids in prompts:
        ids = torch.as_tensor(np.asarray(ids), dtype=torch.int64, device=dev)
        h = E[ids]
        if "pos_embed" in emb:
            h = h + emb["pos_embed"].to(torch.float16).to(torch.float32).to(dev)[: ids.shape[0]]
        hs.append(h)
    out = []
    for li in range(cfg.n_layers if layers is None else layers):
        w = {k: v.to(torch.float32).to(dev) for k, v in get_layer(li).items()}
        for k in ("ln1_g", "ln2_g", "ln1_b", "ln2_b", "b_qkv", "b_o", "b_gu", "b_d"):
            if k in w:
                w[k] = w[k].to(torch.float16).to(torch.float32)           # as the kernels will read them
        st = {}
        hs = [_layer_fp32(cfg, w, h, cos, sin, st) for h in hs]
        out.append({k: v.to(torch.float32).cpu() for k, v in st.items()})
    return out


def save_stats(model_dir: str, stats: list) -> str:
    path = os.path.join(model_dir, STATS_FILE)
    np.savez_compressed(path, **{f"{li}.{k}": v.numpy() for li, st in enumerate(stats) for k, v in st.items()})
    return path


def load_stats(model_dir: str):
    path = os.path.join(model_dir, STATS_FILE)
    if not os.path.exists(path):
        return None
    z = np.load(path)
    n = 1 + max(int(k.split(".")[0]) for k in z.files)
    return [{k.split(".")[1]: torch.from_numpy(z[k]) for k in z.files if int(k.split(".")[0]) == li} for li in range(n)]
