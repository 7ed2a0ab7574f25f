"""Small fixed workload for `ncu`: a few launches of the dominant kernels at Llama-2-7B shapes.
usage: python tools/ncu_target.py [gemm|gemm2|gemv|gemv8|attn|pattn|all]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200stage  # noqa: E402
from b200stage import ops, weights  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "all"
dev = "cuda"
cfg = b200stage.get_config("llama-2-7b")
ops._abi.init_device()
NL = 4
sw = weights.build_stage_weights_fast(cfg, 0, NL, False, False, 0, dev)
L = sw["layers"]
d, H, KV, hd, F = cfg.d_model, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.ffn
torch.cuda.synchronize()
if what in ("gemv", "all"):
    h = torch.randn((1, d), dtype=torch.float16, device=dev)
    y32 = torch.empty((1, F), dtype=torch.float32, device=dev)
    amax = torch.zeros((8,), dtype=torch.float32, device=dev)
    for lw in L:
        ops.fused_linear(h, lw["wgu"], lw["s_gu"], x_kind=ops.X_RMS, gamma=lw["ln2_g"], smooth=lw["sm_gu"], out=y32,
                         out_kind=ops.O_GATED, out_smooth=lw["sm_d"], out_amax=amax)
if what == "gemv8":       # the token-batch (mma.sync) variant of the same kernel, 8 tokens
    h = torch.randn((8, d), dtype=torch.float16, device=dev)
    y32 = torch.empty((8, F), dtype=torch.float32, device=dev)
    amax = torch.zeros((8,), dtype=torch.float32, device=dev)
    for lw in L:
        ops.fused_linear(h, lw["wgu"], lw["s_gu"], x_kind=ops.X_RMS, gamma=lw["ln2_g"], smooth=lw["sm_gu"], out=y32,
                         out_kind=ops.O_GATED, out_smooth=lw["sm_d"], out_amax=amax)
if what in ("gemm", "all"):
    T = 2048
    x8 = torch.randint(-127, 128, (T, d), dtype=torch.int8, device=dev)
    st = torch.full((T,), 0.01, device=dev)
    out = torch.empty((T, 2 * F), dtype=torch.float16, device=dev)
    for sched in (10, 14):
        for lw in L:
            ops.gemm_w8a8(x8, lw["wgu"], st, lw["s_gu"], out=out, schedule=sched)
if what == "gemm2":       # round 2: the CTA-pair (cta_group::2) kernel at T = 2048 -- gate_up plain, gate_up with the
    T = 2048              # act*up epilogue (+ the fp32 quantise pass), qkv with the RoPE + KV-append epilogue, down
    x8 = torch.randint(-127, 128, (T, max(d, F)), dtype=torch.int8, device=dev)
    xd, xf = x8[:, :d].contiguous(), x8[:, :F].contiguous()
    st = torch.full((T,), 0.01, device=dev)
    out = torch.empty((T, 2 * F), dtype=torch.float16, device=dev)
    y32 = torch.empty((T, F), dtype=torch.float32, device=dev)
    am = torch.empty((T,), dtype=torch.float32, device=dev)
    q8 = torch.empty((T, F), dtype=torch.int8, device=dev)
    sc = torch.empty((T,), dtype=torch.float32, device=dev)
    qkv = torch.empty((T, cfg.qkv_out), dtype=torch.float16, device=dev)
    hres = torch.randn((T, d), dtype=torch.float16, device=dev)
    hout = torch.empty((T, d), dtype=torch.float16, device=dev)
    cos, sin = (x.to(dev) for x in weights.rope_table(hd, cfg.max_pos))
    pos = torch.arange(T, dtype=torch.int32, device=dev)
    kc = torch.zeros((2, 1, KV, T + 8, hd), dtype=torch.float16, device=dev)
    vc = torch.zeros_like(kc)
    for i, lw in enumerate(L[:2]):
        ops.gemm_w8a8(xd, lw["wgu"], st, lw["s_gu"], out=out, schedule=20)
        ops.gemm_gated_act(xd, lw["wgu"], st, lw["s_gu"], y32=y32, amax=am, out_smooth=lw["sm_d"])
        ops.quant_f32(y32, am, out_q=q8, out_scale=sc)
        ops.gemm_qkv_rope(xd, lw["wqkv"], st, lw["s_qkv"], qkv, cos, sin, pos, kc[i], vc[i], H, KV, hd)
        ops.gemm_w8a8(xf, lw["wd"], st, lw["s_d"], None, hres, out=hout, schedule=20)
if what in ("attn", "all"):
    S = 2100
    kc = torch.randn((NL, 1, KV, S + 8, hd), dtype=torch.float16, device=dev)
    vc = torch.randn_like(kc)
    qkv = torch.randn((1, cfg.qkv_out), dtype=torch.float16, device=dev)
    lens = torch.full((1,), S, dtype=torch.int32, device=dev)
    attn = torch.empty((1, H * hd), dtype=torch.float16, device=dev)
    for i in range(NL):
        ops.decode_attn(qkv, kc[i], vc[i], lens, H, KV, hd, S, hd ** -0.5, 0, out=attn)
if what in ("pattn", "all"):
    T = 2048                                   # the whole bench prompt as one chunk (N = 1 bench)
    qkvT = torch.randn((T, cfg.qkv_out), dtype=torch.float16, device=dev)
    outT = torch.empty((T, H * hd), dtype=torch.float16, device=dev)
    kcp = torch.randn((2, 1, KV, T + 8, hd), dtype=torch.float16, device=dev)
    vcp = torch.randn_like(kcp)
    for i in range(2):
        ops.prefill_attn(qkvT, kcp[i], vcp[i], 0, 0, H, KV, hd, hd ** -0.5, 0, out=outT)
torch.cuda.synchronize()
print("done")
