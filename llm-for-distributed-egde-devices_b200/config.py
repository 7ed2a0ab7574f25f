"""Model shapes for the decoder families the stage path serves.

The dims are the public HF configs of the models BASELINE.json:configs names;
the reference itself only ever loads whole checkpoints by path
(Code/Quantised Models/models_quant_updated.py:34-38) and ships no config of its own.
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, asdict


@dataclass(frozen=True)
class ModelConfig:
    name: str
    n_layers: int
    d_model: int
    n_heads: int
    n_kv_heads: int
    head_dim: int
    ffn: int
    vocab: int
    eps: float = 1e-5
    rope_theta: float = 10000.0
    window: int = 0            # sliding-window size, 0 = full attention
    max_pos: int = 4096
    act: str = "silu"          # "silu" | "gelu_new"
    gated: bool = True         # gate/up MLP (Llama) vs single up projection (GPT-2)
    norm: str = "rms"          # "rms" | "layernorm"
    rope: bool = True          # False -> learned absolute positions (GPT-2)
    bias: bool = False         # linear layers carry a bias (GPT-2)
    tie_embeddings: bool = True
    # HF `rope_scaling` of type "llama3" (Llama-3.1/3.2): ("llama3", factor, low_freq_factor,
    # high_freq_factor, original_max_position_embeddings); () = plain RoPE
    rope_scaling: tuple = ()
    # --- the reference's own model families (Phi-2, Pythia / GPT-NeoX; Code/others/config.yaml:13-16) -----------
    # parallel residual: attention and MLP both read the layer input, h + attn(ln1 h) + mlp(ln2 h)
    parallel: bool = False
    shared_ln: bool = False    # Phi: ONE LayerNorm feeds both branches (stored twice, as ln1 and ln2)
    rotary_dim: int = 0        # partial rotary: only the first rotary_dim dims of a head are rotated (0 = all)
    # head_dim is the width the KERNELS see.  A model whose heads are not a multiple of 32 wide (Phi-2: 80) is
    # padded with zero rows / columns at load time (weights.to_kernel_layout); head_dim_model keeps the real width,
    # which the softmax scale uses (0 = head_dim).
    head_dim_model: int = 0
    head_bias: bool = False    # lm_head carries a bias (Phi)

    @property
    def qkv_out(self) -> int:
        return (self.n_heads + 2 * self.n_kv_heads) * self.head_dim

    @property
    def up_out(self) -> int:
        return (2 if self.gated else 1) * self.ffn

    def linear_params_per_layer(self) -> int:
        d = self.d_model
        return self.qkv_out * d + d * self.n_heads * self.head_dim + self.up_out * d + d * self.ffn

    def kernel_cfg(self) -> dict:
        """The plain-dict view the oracle consumes."""
        return dict(n_heads=self.n_heads, n_kv_heads=self.n_kv_heads, head_dim=self.head_dim, eps=self.eps,
                    rope_theta=self.rope_theta, window=self.window, act=self.act, gated=self.gated, rope=self.rope,
                    rope_scaling=self.rope_scaling, parallel=self.parallel, rotary_dim=self.rotary_dim,
                    head_dim_model=self.head_dim_model)

    @property
    def hd_model(self) -> int:
        return self.head_dim_model or self.head_dim

    @property
    def attn_scale(self) -> float:
        return 1.0 / float(self.hd_model) ** 0.5

    def to_dict(self) -> dict:
        return asdict(self)


PRESETS = {
    "gpt2": ModelConfig("gpt2", 12, 768, 12, 12, 64, 3072, 50257, eps=1e-5, max_pos=1024, act="gelu_new",
                        gated=False, norm="layernorm", rope=False, bias=True),
    "tinyllama-1.1b": ModelConfig("tinyllama-1.1b", 22, 2048, 32, 4, 64, 5632, 32000, max_pos=2048),
    "llama-2-7b": ModelConfig("llama-2-7b", 32, 4096, 32, 32, 128, 11008, 32000, max_pos=4096),
    "llama-2-13b": ModelConfig("llama-2-13b", 40, 5120, 40, 40, 128, 13824, 32000, max_pos=4096),
    # the refiner of the reference's three-model pipeline (combiner_quant_updated.py:177-203; try.py:303)
    "llama-3.2-1b": ModelConfig("llama-3.2-1b", 16, 2048, 32, 8, 64, 8192, 128256, rope_theta=500000.0, max_pos=8192,
                                rope_scaling=("llama3", 32.0, 1.0, 4.0, 8192)),
    # the two answerers of the reference's pipeline (combiner_quant_updated.py:177-178; config.yaml:13-14)
    "phi-2": ModelConfig("phi-2", 32, 2560, 32, 32, 96, 10240, 51200, eps=1e-5, max_pos=2048, act="gelu_new", gated=False,
                         norm="layernorm", bias=True, tie_embeddings=False, parallel=True, shared_ln=True, rotary_dim=32,
                         head_dim_model=80, head_bias=True),
    "pythia-1b": ModelConfig("pythia-1b", 16, 2048, 8, 8, 256, 8192, 50304, eps=1e-5, max_pos=2048, act="gelu", gated=False,
                             norm="layernorm", bias=True, tie_embeddings=False, parallel=True, rotary_dim=64),
    "mistral-7b": ModelConfig("mistral-7b", 32, 4096, 32, 8, 128, 14336, 32000, window=4096, max_pos=32768),
    # small shapes for tests (same structure, seconds on the CPU oracle)
    "tiny-llama-test": ModelConfig("tiny-llama-test", 4, 256, 4, 4, 64, 512, 512, max_pos=512),
    "tiny-gqa-test": ModelConfig("tiny-gqa-test", 4, 512, 8, 2, 64, 1024, 1024, window=0, max_pos=512),
    "tiny-window-test": ModelConfig("tiny-window-test", 2, 256, 4, 2, 64, 512, 512, window=48, max_pos=512),
    "tiny-llama3-test": ModelConfig("tiny-llama3-test", 3, 256, 4, 2, 64, 512, 640, rope_theta=500000.0, max_pos=512,
                                    rope_scaling=("llama3", 8.0, 1.0, 4.0, 64)),
    "tiny-hd128-test": ModelConfig("tiny-hd128-test", 2, 512, 4, 4, 128, 1024, 512, max_pos=512),
    "tiny-phi-test": ModelConfig("tiny-phi-test", 3, 320, 4, 4, 96, 640, 512, eps=1e-5, max_pos=256, act="gelu_new", gated=False,
                                 norm="layernorm", bias=True, tie_embeddings=False, parallel=True, shared_ln=True,
                                 rotary_dim=32, head_dim_model=80, head_bias=True),
    "tiny-neox-test": ModelConfig("tiny-neox-test", 3, 512, 2, 2, 256, 1024, 512, eps=1e-5, max_pos=256, act="gelu", gated=False,
                                  norm="layernorm", bias=True, tie_embeddings=False, parallel=True, rotary_dim=64),
    "tiny-gpt2-test": ModelConfig("tiny-gpt2-test", 3, 256, 4, 4, 64, 1024, 512, eps=1e-5, max_pos=256, act="gelu_new",
                                  gated=False, norm="layernorm", rope=False, bias=True),
}


def get_config(spec) -> ModelConfig:
    """Preset name, ModelConfig, dict, or a HF model dir / config.json (the reference's
    `from_pretrained(model_path)` convention, models_quant_updated.py:25-38)."""
    if isinstance(spec, ModelConfig):
        return spec
    if isinstance(spec, dict):
        return from_hf_dict(spec) if "hidden_size" in spec or "n_embd" in spec else ModelConfig(**spec)
    if spec in PRESETS:
        return PRESETS[spec]
    path = spec
    if os.path.isdir(path):
        path = os.path.join(path, "config.json")
    if os.path.exists(path):
        with open(path) as f:
            return from_hf_dict(json.load(f))
    raise ValueError(f"unknown model spec {spec!r}")


def from_hf_dict(c: dict) -> ModelConfig:
    if c.get("model_type") == "gpt2" or "n_embd" in c:
        d, h = c["n_embd"], c["n_head"]
        return ModelConfig(c.get("_name_or_path", "gpt2"), c["n_layer"], d, h, h, d // h, c.get("n_inner") or 4 * d,
                           c["vocab_size"], eps=c.get("layer_norm_epsilon", 1e-5), max_pos=c.get("n_positions", 1024),
                           act="gelu_new", gated=False, norm="layernorm", rope=False, bias=True)
    if c.get("model_type") == "gpt_neox":
        d, h = c["hidden_size"], c["num_attention_heads"]
        hd = d // h
        rp = c.get("rope_parameters") or {}
        pct = rp.get("partial_rotary_factor", c.get("rotary_pct", 0.25))
        if hd % 32 or not c.get("use_parallel_residual", True):
            raise ValueError("gpt_neox: head_dim must be a multiple of 32 and use_parallel_residual true")
        return ModelConfig(c.get("_name_or_path", "gpt_neox"), c["num_hidden_layers"], d, h, h, hd, c["intermediate_size"],
                           c["vocab_size"], eps=c.get("layer_norm_eps", 1e-5),
                           rope_theta=rp.get("rope_theta", c.get("rotary_emb_base", 10000.0)),
                           max_pos=c.get("max_position_embeddings", 2048), act=c.get("hidden_act", "gelu"), gated=False,
                           norm="layernorm", bias=bool(c.get("attention_bias", True)), tie_embeddings=bool(c.get("tie_word_embeddings", False)),
                           parallel=True, rotary_dim=int(hd * pct))
    if c.get("model_type") == "phi":
        d, h = c["hidden_size"], c["num_attention_heads"]
        hd = c.get("head_dim") or d // h
        rp = c.get("rope_parameters") or {}
        # transformers >= 5 keeps the factor the model uses in rope_parameters (a top-level copy defaults to 0.5)
        prf = rp.get("partial_rotary_factor", c.get("partial_rotary_factor", 0.5))
        if c.get("qk_layernorm"):
            raise ValueError("phi with qk_layernorm is not supported")
        return ModelConfig(c.get("_name_or_path", "phi"), c["num_hidden_layers"], d, h, c.get("num_key_value_heads") or h,
                           (hd + 31) // 32 * 32, c["intermediate_size"], c["vocab_size"], eps=c.get("layer_norm_eps", 1e-5),
                           rope_theta=rp.get("rope_theta", c.get("rope_theta", 10000.0)),
                           max_pos=c.get("max_position_embeddings", 2048), act=c.get("hidden_act", "gelu_new"), gated=False,
                           norm="layernorm", bias=True, tie_embeddings=False, parallel=True, shared_ln=True,
                           rotary_dim=int(hd * prf), head_dim_model=hd, head_bias=True)
    d, h = c["hidden_size"], c["num_attention_heads"]
    rs = c.get("rope_scaling") or {}
    kind = rs.get("rope_type", rs.get("type"))
    if kind not in (None, "default", "llama3"):
        raise ValueError(f"rope_scaling type {kind!r} is not supported")
    scaling = ("llama3", float(rs["factor"]), float(rs["low_freq_factor"]), float(rs["high_freq_factor"]),
               int(rs["original_max_position_embeddings"])) if kind == "llama3" else ()
    return ModelConfig(c.get("_name_or_path", c.get("model_type", "llama")), c["num_hidden_layers"], d, h,
                       c.get("num_key_value_heads", h), c.get("head_dim") or d // h, c["intermediate_size"],
                       c["vocab_size"], eps=c.get("rms_norm_eps", 1e-5), rope_theta=c.get("rope_theta", 10000.0),
                       window=c.get("sliding_window") or 0, max_pos=c.get("max_position_embeddings", 4096),
                       tie_embeddings=c.get("tie_word_embeddings", False), rope_scaling=scaling)
