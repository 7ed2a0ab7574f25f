"""Prefill GEMM timings at Llama-2-7B shapes (developer tool): the single-CTA tcgen05 kernel (schedules 10 / 14),
the CTA-pair kernel (schedule 20) and its fused epilogues, plus the row passes around them.

Each variant is captured in a CUDA graph over NL distinct weight copies (every launch streams cold weights:
NL x 202 MB >> L2) and timed with CUDA events around graph replays.

    python tools/gemm_bench.py [--model llama-2-7b] [--t 2048,512,256] [--nl 4]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200stage  # noqa: E402
from b200stage import ops, weights  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="llama-2-7b")
ap.add_argument("--t", default="2048,512,256")
ap.add_argument("--nl", type=int, default=4)
ap.add_argument("--reps", type=int, default=10)
a = ap.parse_args()
dev = "cuda"
cfg = b200stage.get_config(a.model)
ops._abi.init_device()
sw = weights.build_stage_weights_fast(cfg, 0, a.nl, True, False, 0, dev)
L = sw["layers"]
d, H, KV, hd, F = cfg.d_model, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.ffn


def timed(fn, reps=a.reps):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    g.replay()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps):
        g.replay()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / reps / len(L) * 1e3     # us per launch


def report(name, us, ops_=None, nbytes=None):
    extra = ""
    if ops_:
        extra = f"  {ops_ / us / 1e6:7.0f} TOP/s ({ops_ / us / 1e6 / 45:5.1f} % of nominal 4500)"
    if nbytes:
        extra = f"  {nbytes / us / 1e3:7.0f} GB/s"
    print(f"{name:58s} {us:8.2f} us{extra}", flush=True)


cos, sin = weights.rope_table(hd, cfg.max_pos)
cos, sin = cos.to(dev), sin.to(dev)
for T in [int(x) for x in a.t.split(",")]:
    print(f"== {cfg.name}, T = {T}")
    x8 = torch.randint(-127, 128, (T, max(d, F)), dtype=torch.int8, device=dev)
    st = torch.full((T,), 0.01, device=dev)
    hres = torch.randn((T, d), dtype=torch.float16, device=dev)
    tot = {}
    for name, wk, sk, K in (("qkv", "wqkv", "s_qkv", d), ("o", "wo", "s_o", H * hd), ("gate_up", "wgu", "s_gu", d),
                            ("down", "wd", "s_d", F)):
        N = L[0][wk].shape[0]
        xin = x8[:, :K].contiguous()
        out = torch.empty((T, N), dtype=torch.float16, device=dev)
        res = hres if N == d else None
        for sched, tag in ((10, "1-CTA 128x256"), (14, "1-CTA 128x128"), (20, "pair 256x256")):
            try:
                us = timed(lambda: [ops.gemm_w8a8(xin, lw[wk], st, lw[sk], None, res, out=out, schedule=sched) for lw in L])
                report(f"gemm {name} [{T}x{N}x{K}] {tag}", us, ops_=2.0 * T * N * K)
                tot.setdefault(tag, 0.0)
                tot[tag] += us
            except Exception as ex:  # noqa: BLE001
                print(f"gemm {name} {tag}: {ex}", flush=True)
    for tag, us in tot.items():
        report(f"  sum of the four linears, {tag}", us, ops_=2.0 * T * cfg.linear_params_per_layer())
    # fused epilogues
    xin = x8[:, :d].contiguous()
    y32 = torch.empty((T, F), dtype=torch.float32, device=dev)
    amax = torch.empty((T,), dtype=torch.float32, device=dev)
    q8 = torch.empty((T, F), dtype=torch.int8, device=dev)
    sc = torch.empty((T,), dtype=torch.float32, device=dev)
    if cfg.gated:
        us_g = timed(lambda: [ops.gemm_gated_act(xin, lw["wgu"], st, lw["s_gu"], y32=y32, amax=amax, out_smooth=lw["sm_d"]) for lw in L])
        report(f"pair gate_up + SiLU*up epilogue [{T}x{2 * F}x{d}]", us_g, ops_=2.0 * T * 2 * F * d)
        us_q = timed(lambda: [ops.quant_f32(y32, amax, out_q=q8, out_scale=sc) for _ in L])
        report(f"quant_f32 [{T}x{F}]", us_q, nbytes=T * F * 5)
        gu = torch.empty((T, 2 * F), dtype=torch.float16, device=dev)
        ops.gemm_w8a8(xin, L[0]["wgu"], st, L[0]["s_gu"], out=gu, schedule=10)
        us_a = timed(lambda: [ops.act_mul_quant(gu, lw["sm_d"], out_q=q8, out_scale=sc) for lw in L])
        report(f"act_mul_quant [{T}x{F}] (the unfused pass)", us_a, nbytes=T * F * 5)
    Smax = T + 64
    kc = torch.zeros((a.nl, 1, KV, Smax, hd), dtype=torch.float16, device=dev)
    vc = torch.zeros_like(kc)
    qkv = torch.empty((T, cfg.qkv_out), dtype=torch.float16, device=dev)
    pos = torch.arange(T, dtype=torch.int32, device=dev)
    us = timed(lambda: [ops.gemm_qkv_rope(xin, lw["wqkv"], st, lw["s_qkv"], qkv, cos, sin, pos, kc[i], vc[i], H, KV, hd)
                        for i, lw in enumerate(L)])
    report(f"pair qkv + RoPE + KV append epilogue [{T}x{cfg.qkv_out}x{d}]", us, ops_=2.0 * T * cfg.qkv_out * d)
    us = timed(lambda: [ops.rope_kv_append(qkv, cos, sin, pos, kc[i], vc[i], H, KV, hd) for i in range(a.nl)])
    report(f"rope_kv_append [{T}] (the unfused pass)", us, nbytes=T * cfg.qkv_out * 4)
    h16 = torch.randn((T, d), dtype=torch.float16, device=dev)
    q8d = torch.empty((T, d), dtype=torch.int8, device=dev)
    us = timed(lambda: [ops.norm_quant(h16, lw["ln1_g"], None, lw["sm_qkv"], 1e-5, out_q=q8d, out_scale=sc) for lw in L])
    report(f"norm_quant [{T}x{d}]", us, nbytes=T * d * 3)
    at = torch.randn((T, H * hd), dtype=torch.float16, device=dev)
    q8o = torch.empty((T, H * hd), dtype=torch.int8, device=dev)
    us = timed(lambda: [ops.smooth_quant(at, lw["sm_o"], out_q=q8o, out_scale=sc) for lw in L])
    report(f"smooth_quant [{T}x{H * hd}]", us, nbytes=T * H * hd * 3)
    kcp = torch.randn((a.nl, 1, KV, Smax, hd), dtype=torch.float16, device=dev)
    vcp = torch.randn_like(kcp)
    qq = torch.randn((T, cfg.qkv_out), dtype=torch.float16, device=dev)
    us = timed(lambda: [ops.prefill_attn(qq, kcp[i], vcp[i], 0, 0, H, KV, hd, hd ** -0.5, cfg.window, out=at) for i in range(a.nl)])
    report(f"prefill_attn T={T} from position 0", us, ops_=4.0 * T * (T / 2) * hd * H)
