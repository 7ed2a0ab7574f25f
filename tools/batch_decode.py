"""Batched decode step of a truncated model at a long context (developer tool; BASELINE.json configs[3]:
Llama-2-13B, batch 32, 4K KV cache): device time per layer against the HBM roofline of that step.

    python tools/batch_decode.py [--model llama-2-13b] [--batch 32] [--ctx 4000] [--layers 2]
"""
import argparse
import dataclasses
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200stage  # noqa: E402
from b200stage.stage import Stage  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="llama-2-13b")
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--ctx", type=int, default=4000)
ap.add_argument("--layers", type=int, default=2)
ap.add_argument("--fuse-tail", type=int, default=-1, help="override Stage.fuse_tail (0/1)")
a = ap.parse_args()
full = b200stage.get_config(a.model)
cfg = dataclasses.replace(full, n_layers=a.layers)
dev = "cuda"
st = Stage.load(cfg, 0, a.layers, first=False, last=False, device=dev, fast_init=True)
if a.fuse_tail >= 0:
    st.fuse_tail = bool(a.fuse_tail)
B = a.batch
cache = st.new_kv_cache(B, a.ctx + 64)
cache.k.normal_()
cache.v.normal_()
x = torch.randn((B, cfg.d_model), dtype=torch.float16, device=dev)
rows = np.arange(B)
ms = {}
for name, steps in (("warm", 3), ("timed", 20)):
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(steps):
        st.forward_auto(x, cache, np.full(B, a.ctx + (i % 32 if name == "timed" else 0)), rows, logits="none")
    e.record()
    torch.cuda.synchronize()
    ms[name] = s.elapsed_time(e) / steps
us_layer = ms["timed"] * 1e3 / a.layers
w_bytes = cfg.linear_params_per_layer()
kv_bytes = 2 * B * cfg.n_kv_heads * min(a.ctx, cfg.window or a.ctx) * cfg.head_dim * 2
gbs = (w_bytes + kv_bytes) / us_layer / 1e3
print(f"{full.name}: batch {B}, context {a.ctx}: {us_layer:.1f} us per layer; weights {w_bytes / 1e6:.0f} MB + KV {kv_bytes / 1e6:.0f} MB "
      f"-> {gbs:.0f} GB/s = {gbs / 6550.4 * 100:.0f} % of measured HBM; {B / (us_layer * full.n_layers * 1e-6):.0f} tok/s for {full.n_layers} layers")
