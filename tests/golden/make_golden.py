"""Generate the golden fixtures in this directory (run in the build container, where transformers is
importable; the GPU box never needs to run this).

    python tests/golden/make_golden.py

For each small seeded config it writes `<name>.npz` with

  ids            prompt token ids
  hf_logits      Tier R: fp32 logits of HF `LlamaForCausalLM` -- the class the reference's
                 `AutoModelForCausalLM.from_pretrained` call instantiates for Llama-family
                 checkpoints (Code/Quantised Models/models_quant_updated.py:34-38) -- loaded with the
                 SAME synthetic fp32 weights the stage path quantises
  hf_greedy      Tier R: `generate(do_sample=False)` continuation (the greedy subset of the
                 reference's generate call, models_quant_updated.py:84-93)
  hf_dyn_logits  Tier R': the same model after `torch.quantization.quantize_dynamic(model,
                 {nn.Linear}, qint8)`, the reference's CPU int8 path (Code/C-DAC Server/try.py:204-205)
  s_logits, s_greedy, s_margins
                 Tier S: oracle/w8a8.py on the W8A8 weights (regression anchor for the oracle itself)

Nothing here reads /root/reference; the reference contributes only the call sequence cited above.
"""
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import b200stage  # noqa: E402
from b200stage import weights  # noqa: E402
from oracle import w8a8 as O  # noqa: E402

CASES = [("tiny-llama-test", 3, 24, 8), ("tiny-gqa-test", 5, 20, 8), ("tiny-llama3-test", 7, 200, 8)]


def hf_model(cfg, seed):
    from transformers import LlamaConfig, LlamaForCausalLM
    hc = LlamaConfig(vocab_size=cfg.vocab, hidden_size=cfg.d_model, intermediate_size=cfg.ffn,
                     num_hidden_layers=cfg.n_layers, num_attention_heads=cfg.n_heads,
                     num_key_value_heads=cfg.n_kv_heads, head_dim=cfg.head_dim, max_position_embeddings=cfg.max_pos,
                     rms_norm_eps=cfg.eps, rope_theta=cfg.rope_theta, tie_word_embeddings=False,
                     attention_bias=False, mlp_bias=False,
                     rope_scaling=dict(rope_type="llama3", factor=cfg.rope_scaling[1], low_freq_factor=cfg.rope_scaling[2],
                                       high_freq_factor=cfg.rope_scaling[3],
                                       original_max_position_embeddings=cfg.rope_scaling[4]) if cfg.rope_scaling else None)
    m = LlamaForCausalLM(hc).eval().float()
    sd = {}
    H, KV, hd, F = cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.ffn
    h16 = lambda t: t.to(torch.float16).to(torch.float32)     # the stage path stores these in fp16
    for li in range(cfg.n_layers):
        w = weights.init_layer_fp32(cfg, li, seed)
        p = f"model.layers.{li}."
        sd[p + "self_attn.q_proj.weight"] = w["wqkv"][: H * hd]
        sd[p + "self_attn.k_proj.weight"] = w["wqkv"][H * hd: (H + KV) * hd]
        sd[p + "self_attn.v_proj.weight"] = w["wqkv"][(H + KV) * hd:]
        sd[p + "self_attn.o_proj.weight"] = w["wo"]
        sd[p + "mlp.gate_proj.weight"] = w["wgu"][:F]
        sd[p + "mlp.up_proj.weight"] = w["wgu"][F:]
        sd[p + "mlp.down_proj.weight"] = w["wd"]
        sd[p + "input_layernorm.weight"] = h16(w["ln1_g"])
        sd[p + "post_attention_layernorm.weight"] = h16(w["ln2_g"])
    e = weights.init_embed_fp32(cfg, seed)
    sd["model.embed_tokens.weight"] = h16(e["embed"])
    sd["model.norm.weight"] = h16(e["lnf_g"])
    sd["lm_head.weight"] = h16(e["lm_head"])
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and all("rotary" in k or "inv_freq" in k for k in missing), (missing, unexpected)
    return m


def hf_gpt2(cfg, seed):
    """HF GPT2LMHeadModel (BASELINE.json configs[0] family) with the same synthetic weights.
    GPT-2 uses Conv1D: weight is [in, out] = our [out, in] transposed."""
    from transformers import GPT2Config, GPT2LMHeadModel
    hc = GPT2Config(vocab_size=cfg.vocab, n_positions=cfg.max_pos, n_embd=cfg.d_model, n_layer=cfg.n_layers,
                    n_head=cfg.n_heads, n_inner=cfg.ffn, activation_function="gelu_new", layer_norm_epsilon=cfg.eps,
                    resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0, tie_word_embeddings=True)
    m = GPT2LMHeadModel(hc).eval().float()
    h16 = lambda t: t.to(torch.float16).to(torch.float32)
    sd = {}
    for li in range(cfg.n_layers):
        w = weights.init_layer_fp32(cfg, li, seed)
        p = f"transformer.h.{li}."
        sd[p + "attn.c_attn.weight"] = w["wqkv"].t().contiguous()
        sd[p + "attn.c_attn.bias"] = h16(w["b_qkv"])
        sd[p + "attn.c_proj.weight"] = w["wo"].t().contiguous()
        sd[p + "attn.c_proj.bias"] = h16(w["b_o"])
        sd[p + "mlp.c_fc.weight"] = w["wgu"].t().contiguous()
        sd[p + "mlp.c_fc.bias"] = h16(w["b_gu"])
        sd[p + "mlp.c_proj.weight"] = w["wd"].t().contiguous()
        sd[p + "mlp.c_proj.bias"] = h16(w["b_d"])
        sd[p + "ln_1.weight"], sd[p + "ln_1.bias"] = h16(w["ln1_g"]), h16(w["ln1_b"])
        sd[p + "ln_2.weight"], sd[p + "ln_2.bias"] = h16(w["ln2_g"]), h16(w["ln2_b"])
    e = weights.init_embed_fp32(cfg, seed)
    sd["transformer.wte.weight"] = h16(e["embed"])
    sd["transformer.wpe.weight"] = h16(e["pos_embed"])
    sd["transformer.ln_f.weight"], sd["transformer.ln_f.bias"] = h16(e["lnf_g"]), h16(e["lnf_b"])
    sd["lm_head.weight"] = h16(e["embed"])
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and all(("attn.bias" in k or "masked_bias" in k) for k in missing), (missing, unexpected)
    return m


def hf_phi(cfg, seed):
    """HF PhiForCausalLM -- the class `AutoModelForCausalLM.from_pretrained("microsoft/phi-2")` instantiates, the first
    answerer of the reference's pipeline (combiner_quant_updated.py:177; config.yaml:13) -- with the synthetic weights
    in the MODEL's head layout (80-wide heads, first 32 dims rotary); the stage path loads the same weights through
    weights.to_kernel_layout."""
    from transformers import PhiConfig, PhiForCausalLM
    hm = cfg.hd_model
    hc = PhiConfig(vocab_size=cfg.vocab, hidden_size=cfg.d_model, intermediate_size=cfg.ffn, num_hidden_layers=cfg.n_layers,
                   num_attention_heads=cfg.n_heads, num_key_value_heads=cfg.n_kv_heads, hidden_act=cfg.act,
                   max_position_embeddings=cfg.max_pos, layer_norm_eps=cfg.eps, tie_word_embeddings=False,
                   resid_pdrop=0.0, embd_pdrop=0.0, attention_dropout=0.0,
                   rope_parameters=dict(rope_type="default", rope_theta=cfg.rope_theta, partial_rotary_factor=cfg.rotary_dim / hm))
    assert hc.hidden_size // hc.num_attention_heads == hm
    m = PhiForCausalLM(hc).eval().float()
    h16 = lambda t: t.to(torch.float16).to(torch.float32)
    H, KV = cfg.n_heads, cfg.n_kv_heads
    sd = {}
    for li in range(cfg.n_layers):
        w = weights.init_layer_fp32(cfg, li, seed, layout="model")
        p = f"model.layers.{li}."
        for name, lo, hi in (("q_proj", 0, H * hm), ("k_proj", H * hm, (H + KV) * hm), ("v_proj", (H + KV) * hm, (H + 2 * KV) * hm)):
            sd[p + f"self_attn.{name}.weight"] = w["wqkv"][lo:hi]
            sd[p + f"self_attn.{name}.bias"] = h16(w["b_qkv"][lo:hi])
        sd[p + "self_attn.dense.weight"], sd[p + "self_attn.dense.bias"] = w["wo"], h16(w["b_o"])
        sd[p + "mlp.fc1.weight"], sd[p + "mlp.fc1.bias"] = w["wgu"], h16(w["b_gu"])
        sd[p + "mlp.fc2.weight"], sd[p + "mlp.fc2.bias"] = w["wd"], h16(w["b_d"])
        sd[p + "input_layernorm.weight"], sd[p + "input_layernorm.bias"] = h16(w["ln1_g"]), h16(w["ln1_b"])
    e = weights.init_embed_fp32(cfg, seed)
    sd["model.embed_tokens.weight"] = h16(e["embed"])
    sd["model.final_layernorm.weight"], sd["model.final_layernorm.bias"] = h16(e["lnf_g"]), h16(e["lnf_b"])
    sd["lm_head.weight"], sd["lm_head.bias"] = h16(e["lm_head"]), e["lm_head_b"]
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and all("rotary" in k or "inv_freq" in k for k in missing), (missing, unexpected)
    return m


def hf_neox(cfg, seed):
    """HF GPTNeoXForCausalLM -- what `from_pretrained("EleutherAI/pythia-1b")` instantiates, the second answerer of the
    reference's pipeline (combiner_quant_updated.py:178; config.yaml:14): parallel residual with two LayerNorms, 25 %
    rotary, exact GELU, one fused query_key_value matrix stored head by head as [q_h | k_h | v_h]."""
    from transformers import GPTNeoXConfig, GPTNeoXForCausalLM
    hm, H = cfg.hd_model, cfg.n_heads
    hc = GPTNeoXConfig(vocab_size=cfg.vocab, hidden_size=cfg.d_model, intermediate_size=cfg.ffn, num_hidden_layers=cfg.n_layers,
                       num_attention_heads=H, hidden_act=cfg.act, max_position_embeddings=cfg.max_pos, layer_norm_eps=cfg.eps,
                       tie_word_embeddings=False, use_parallel_residual=True, attention_bias=True, hidden_dropout=0.0,
                       attention_dropout=0.0,
                       rope_parameters=dict(rope_type="default", rope_theta=cfg.rope_theta, partial_rotary_factor=cfg.rotary_dim / hm))
    m = GPTNeoXForCausalLM(hc).eval().float()
    h16 = lambda t: t.to(torch.float16).to(torch.float32)
    sd = {}
    for li in range(cfg.n_layers):
        w = weights.init_layer_fp32(cfg, li, seed, layout="model")
        p = f"gpt_neox.layers.{li}."
        q, k, v = (w["wqkv"][i * H * hm:(i + 1) * H * hm].view(H, hm, -1) for i in range(3))
        bq, bk, bv = (w["b_qkv"][i * H * hm:(i + 1) * H * hm].view(H, hm) for i in range(3))
        sd[p + "attention.query_key_value.weight"] = torch.cat([q, k, v], 1).reshape(3 * H * hm, -1).contiguous()
        sd[p + "attention.query_key_value.bias"] = h16(torch.cat([bq, bk, bv], 1).reshape(-1))
        sd[p + "attention.dense.weight"], sd[p + "attention.dense.bias"] = w["wo"], h16(w["b_o"])
        sd[p + "mlp.dense_h_to_4h.weight"], sd[p + "mlp.dense_h_to_4h.bias"] = w["wgu"], h16(w["b_gu"])
        sd[p + "mlp.dense_4h_to_h.weight"], sd[p + "mlp.dense_4h_to_h.bias"] = w["wd"], h16(w["b_d"])
        sd[p + "input_layernorm.weight"], sd[p + "input_layernorm.bias"] = h16(w["ln1_g"]), h16(w["ln1_b"])
        sd[p + "post_attention_layernorm.weight"], sd[p + "post_attention_layernorm.bias"] = h16(w["ln2_g"]), h16(w["ln2_b"])
    e = weights.init_embed_fp32(cfg, seed)
    sd["gpt_neox.embed_in.weight"] = h16(e["embed"])
    sd["gpt_neox.final_layer_norm.weight"], sd["gpt_neox.final_layer_norm.bias"] = h16(e["lnf_g"]), h16(e["lnf_b"])
    sd["embed_out.weight"] = h16(e["lm_head"])
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and all("rotary" in k or "inv_freq" in k or "bias" in k.split(".")[-1] and "attention.bias" in k
                                  or "masked_bias" in k for k in missing), (missing, unexpected)
    return m


def hf_for(cfg, seed):
    if cfg.parallel and cfg.shared_ln:
        return hf_phi(cfg, seed)
    if cfg.parallel:
        return hf_neox(cfg, seed)
    return hf_model(cfg, seed) if cfg.rope else hf_gpt2(cfg, seed)


def main():
    warnings.filterwarnings("ignore")
    torch.set_num_threads(4)
    only = sys.argv[1:]                       # optional: regenerate just the named cases
    for name, seed, T, new in CASES + [("tiny-gpt2-test", 9, 20, 8), ("tiny-phi-test", 13, 40, 8), ("tiny-neox-test", 17, 40, 8)]:
        if only and name not in only:
            continue
        cfg = b200stage.get_config(name)
        ids = np.random.default_rng(seed + 100).integers(0, cfg.vocab, T)
        m = hf_for(cfg, seed)
        tid = torch.from_numpy(ids)[None]
        with torch.no_grad():
            hf_logits = m(tid).logits[0].numpy()
            hf_greedy = m.generate(input_ids=tid, attention_mask=torch.ones_like(tid), max_new_tokens=new,
                                   min_new_tokens=new, do_sample=False, pad_token_id=0)[0].numpy()
            mq = torch.quantization.quantize_dynamic(m, {torch.nn.Linear}, dtype=torch.qint8)
            hf_dyn_logits = mq(tid).logits[0].numpy()
        sw = weights.to_numpy(weights.build_stage_weights(cfg, 0, cfg.n_layers, True, True, seed, "cpu"))
        cos, sin = O.rope_table(cfg.head_dim, cfg.max_pos, cfg.rope_theta, cfg.rope_scaling, cfg.rotary_dim)
        cache = O.KVCacheOracle(cfg.n_layers, 1, cfg.n_kv_heads, T + new, cfg.head_dim)
        s_logits = O.stage_forward(ids, sw, cfg.kernel_cfg(), cache, np.zeros(T, np.int64), np.arange(T), cos, sin)
        s_greedy, margins = O.greedy_generate([sw], cfg.kernel_cfg(), ids, new, T + new)
        out = os.path.join(HERE, f"{name}.npz")
        np.savez_compressed(out, ids=ids, hf_logits=hf_logits.astype(np.float32), hf_greedy=hf_greedy,
                            hf_dyn_logits=hf_dyn_logits.astype(np.float32), s_logits=s_logits.astype(np.float32),
                            s_greedy=s_greedy, s_margins=np.array(margins, np.float32), seed=seed, new=new)
        a, b = hf_logits.ravel().astype(np.float64), s_logits.ravel().astype(np.float64)
        c = hf_dyn_logits.ravel().astype(np.float64)
        print(f"{name}: cos(S,HF)={a @ b / np.linalg.norm(a) / np.linalg.norm(b):.6f} "
              f"cos(dyn,HF)={a @ c / np.linalg.norm(a) / np.linalg.norm(c):.6f} "
              f"maxabs(S-HF)={np.abs(a - b).max():.4f} max|HF|={np.abs(a).max():.2f} "
              f"argmax agree={np.mean(hf_logits.argmax(-1) == s_logits.argmax(-1)):.3f} "
              f"greedy equal={np.array_equal(hf_greedy, s_greedy)} min margin={min(margins):.3f} "
              f"size={os.path.getsize(out) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
