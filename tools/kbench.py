"""Per-kernel device timings of the decode / prefill chain at Llama-2-7B shapes (developer tool).

Each kernel is captured in a CUDA graph over NL distinct weight copies (so every launch streams
cold weights: NL x 202 MB >> L2) and timed with CUDA events around graph replays.

    python tools/kbench.py [--model llama-2-7b] [--seq 2100] [--m 1] [--prefill 512]
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
ap.add_argument("--seq", type=int, default=2100)
ap.add_argument("--m", type=int, default=1)
ap.add_argument("--prefill", type=int, default=512)
ap.add_argument("--nl", type=int, default=8)
ap.add_argument("--chains", default="qaogd,qogd,qa,ao,og,gd,dq")
a = ap.parse_args()
dev = "cuda"
cfg = b200stage.get_config(a.model)
ops._abi.init_device()
sw = weights.build_stage_weights_fast(cfg, 0, a.nl, True, True, 0, dev)
L = sw["layers"]
d, H, KV, hd, F = cfg.d_model, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.ffn
M = a.m
HBM = 6550.4


def timed(fn, reps=20):
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


def report(name, us, nbytes=None, ops_=None):
    extra = ""
    if nbytes:
        extra = f"  {nbytes / us / 1e3:7.0f} GB/s ({nbytes / us / 1e3 / HBM * 100:5.1f} % HBM)"
    if ops_:
        extra = f"  {ops_ / us / 1e6:7.0f} TOP/s"
    print(f"{name:46s} {us:8.2f} us{extra}", flush=True)


h = torch.randn((M, d), dtype=torch.float16, device=dev)
h2 = torch.empty_like(h)
qkv = torch.empty((M, cfg.qkv_out), dtype=torch.float16, device=dev)
attn = torch.randn((M, H * hd), dtype=torch.float16, device=dev)
y32 = torch.empty((M, F), dtype=torch.float32, device=dev)
amax = torch.ones((8,), dtype=torch.float32, device=dev)
cos, sin = weights.rope_table(hd, 4096)
cos, sin = cos.to(dev), sin.to(dev)
Smax = a.seq + 8
kc = torch.randn((a.nl, M, KV, Smax, hd), dtype=torch.float16, device=dev)
vc = torch.randn_like(kc)
pos = torch.full((M,), a.seq - 1, dtype=torch.int32, device=dev)
lens = torch.full((M,), a.seq, dtype=torch.int32, device=dev)
seq = torch.arange(M, dtype=torch.int32, device=dev)
ws = torch.empty(max(16, ops.decode_attn_workspace_bytes(M, H, KV, hd, a.seq, cfg.window)), dtype=torch.uint8, device=dev)

print(f"== decode chain, M={M}, seq={a.seq}, {cfg.name}")
t = timed(lambda: [ops.fused_qkv(h, lw["wqkv"], lw["s_qkv"], qkv, cos, sin, pos, kc[i], vc[i], H, KV, hd, x_kind=ops.X_RMS,
                                 gamma=lw["ln1_g"], smooth=lw["sm_qkv"], seq_ids=seq) for i, lw in enumerate(L)])
report("fused_qkv (norm+quant+GEMV+rope+append)", t, cfg.qkv_out * d)
t = timed(lambda: [ops.decode_attn(qkv, kc[i], vc[i], lens, H, KV, hd, a.seq, hd ** -0.5, cfg.window, rows=seq, out=attn,
                                   workspace=ws) for i, lw in enumerate(L)])
report("decode_attn (+merge)", t, 2 * M * KV * min(a.seq, cfg.window or a.seq) * hd * 2)
t = timed(lambda: [ops.fused_linear(attn, lw["wo"], lw["s_o"], x_kind=ops.X_PLAIN, smooth=lw["sm_o"], residual=h, out=h2,
                                    zero_amax=amax) for lw in L])
report("fused o_proj (smooth+quant+GEMV+residual)", t, d * H * hd)
t = timed(lambda: [ops.fused_linear(h, lw["wgu"], lw["s_gu"], x_kind=ops.X_RMS, gamma=lw["ln2_g"], smooth=lw["sm_gu"], out=y32,
                                    out_kind=ops.O_GATED, out_smooth=lw["sm_d"], out_amax=amax) for lw in L])
report("fused gate_up (norm+quant+GEMV+SiLU*up)", t, 2 * F * d)
t = timed(lambda: [ops.fused_linear(y32, lw["wd"], lw["s_d"], x_kind=ops.X_Y32, x_amax=amax, residual=h, out=h2) for lw in L])
report("fused down (requant+GEMV+residual)", t, d * F)


def chain(parts):
    """The real per-layer launch order (Stage._core, fused branch, no L2 hints), restricted to `parts`."""
    for i, lw in enumerate(L):
        if "q" in parts:
            ops.fused_qkv(h, lw["wqkv"], lw["s_qkv"], qkv, cos, sin, pos, kc[i], vc[i], H, KV, hd, x_kind=ops.X_RMS,
                          gamma=lw["ln1_g"], smooth=lw["sm_qkv"], seq_ids=seq)
        if "a" in parts:
            ops.decode_attn(qkv, kc[i], vc[i], lens, H, KV, hd, a.seq, hd ** -0.5, cfg.window, rows=seq, out=attn, workspace=ws)
        if "o" in parts:
            ops.fused_linear(attn, lw["wo"], lw["s_o"], x_kind=ops.X_PLAIN, smooth=lw["sm_o"], residual=h, out=h2, zero_amax=amax)
        if "g" in parts:
            ops.fused_linear(h2, lw["wgu"], lw["s_gu"], x_kind=ops.X_RMS, gamma=lw["ln2_g"], smooth=lw["sm_gu"], out=y32,
                             out_kind=ops.O_GATED, out_smooth=lw["sm_d"], out_amax=amax)
        if "d" in parts:
            ops.fused_linear(y32, lw["wd"], lw["s_d"], x_kind=ops.X_Y32, x_amax=amax, residual=h2, out=h)


for parts in a.chains.split(","):
    report(f"chain [{parts}] per layer", timed(lambda: chain(parts)))
ctr = torch.zeros((a.nl, 4), dtype=torch.int32, device=dev)


def tail_all(with_qkv):
    ctr.zero_()
    for i, lw in enumerate(L):
        nl = L[(i + 1) % len(L)]
        d_o = ops.fused_desc(attn, lw["wo"], lw["s_o"], x_kind=ops.X_PLAIN, smooth=lw["sm_o"], residual=h, out=h2, zero_amax=amax)
        d_g = ops.fused_desc(h2, lw["wgu"], lw["s_gu"], x_kind=ops.X_RMS, gamma=lw["ln2_g"], smooth=lw["sm_gu"], out=y32,
                             out_kind=ops.O_GATED, out_smooth=lw["sm_d"], out_amax=amax)
        d_d = ops.fused_desc(y32, lw["wd"], lw["s_d"], x_kind=ops.X_Y32, x_amax=amax, residual=h2, out=h)
        d_q = ops.fused_desc(h, nl["wqkv"], nl["s_qkv"], x_kind=ops.X_RMS, gamma=nl["ln1_g"], smooth=nl["sm_qkv"], out=qkv,
                             out_kind=ops.O_QKV, rope=dict(cos=cos, sin=sin, positions=pos, seq_ids=seq, kcache=kc[i], vcache=vc[i],
                                                           H=H, KV=KV, hd=hd)) if with_qkv else None
        ops.fused_layer_tail(d_o, d_g, d_d, d_q, M, ctr[i])


t = timed(lambda: tail_all(False))
report("layer tail, 3 phases (o + gate_up + down)", t, d * H * hd + 2 * F * d + d * F)
t = timed(lambda: tail_all(True))
report("layer tail, 4 phases (+ next qkv)", t, d * H * hd + 2 * F * d + d * F + cfg.qkv_out * d)
q8 = torch.randint(-127, 128, (M, d), dtype=torch.int8, device=dev)
sc = torch.full((M,), 0.01, device=dev)
gu = torch.empty((M, 2 * F), dtype=torch.float16, device=dev)
t = timed(lambda: [ops.gemm_w8a8(q8, lw["wgu"], sc, lw["s_gu"], out=gu) for lw in L])
report("plain gemv_i8 gate_up (pre-quantised input)", t, 2 * F * d)
t = timed(lambda: [ops.norm_quant(h, lw["ln1_g"], None, lw["sm_qkv"], 1e-5) for lw in L])
report("norm_quant alone (M rows)", t)
logits = torch.empty((M, cfg.vocab), dtype=torch.float32, device=dev)
t = timed(lambda: [ops.lm_head(h, sw["lm_head"], out=logits) for _ in L])
report("lm_head gemv_f16 (warm after 1st)", t, cfg.vocab * d * 2)

if a.prefill:
    T = a.prefill
    print(f"== prefill chain, T={T}")
    x8 = torch.randint(-127, 128, (T, max(d, F)), dtype=torch.int8, device=dev)
    st = torch.full((T,), 0.01, device=dev)
    hT = torch.randn((T, d), dtype=torch.float16, device=dev)
    for name, wk, sk, K in (("qkv", "wqkv", "s_qkv", d), ("o", "wo", "s_o", H * hd), ("gate_up", "wgu", "s_gu", d), ("down", "wd", "s_d", F)):
        N = L[0][wk].shape[0]
        xin = x8[:, :K].contiguous()
        out = torch.empty((T, N), dtype=torch.float16, device=dev)
        for sched, tag in ((0, "auto"), (10, "BN256 CM1"), (11, "BN256 CM2"), (12, "BN256 CM4"), (14, "BN128 CM1"),
                           (16, "BN128 CM2"), (13, "BN128 CM4")):
            t = timed(lambda: [ops.gemm_w8a8(xin, lw[wk], st, lw[sk], out=out, schedule=sched) for lw in L], 10)
            report(f"gemm_tc {name} [{T}x{N}x{K}] {tag}", t, ops_=2.0 * T * N * K)
    qkvT = torch.randn((T, cfg.qkv_out), dtype=torch.float16, device=dev)
    outT = torch.empty((T, H * hd), dtype=torch.float16, device=dev)
    kcp = torch.randn((a.nl, 1, KV, a.seq + T, hd), dtype=torch.float16, device=dev)
    vcp = torch.randn_like(kcp)
    p0 = a.seq - (a.seq % 64)
    t = timed(lambda: [ops.prefill_attn(qkvT, kcp[i], vcp[i], p0, 0, H, KV, hd, hd ** -0.5, cfg.window, out=outT) for i in range(a.nl)], 10)
    report(f"prefill_attn T={T} pos0={p0}", t, ops_=4.0 * T * (p0 + T / 2) * hd * H)
    t = timed(lambda: [ops.norm_quant(hT, lw["ln1_g"], None, lw["sm_qkv"], 1e-5) for lw in L], 10)
    report(f"norm_quant T={T}", t, T * d * 3)
    guT = torch.randn((T, 2 * F), dtype=torch.float16, device=dev)
    t = timed(lambda: [ops.act_mul_quant(guT, lw["sm_d"]) for lw in L], 10)
    report(f"act_mul_quant T={T}", t, T * F * 5)
    t = timed(lambda: [ops.rope_kv_append(qkvT, cos, sin, torch.arange(T, dtype=torch.int32, device=dev), kcp[i, :], vcp[i, :], H, KV, hd) for i in range(a.nl)], 10)
    report(f"rope_kv_append T={T}", t, T * cfg.qkv_out * 4)
