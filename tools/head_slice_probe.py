"""Developer probe: what one b2_lm_head_argmax launch costs against its bytes, by slice size (Llama-2 head, d = 4096).
Launches rotate over a 1 GB fp16 matrix so that the weights come from HBM, not from L2."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200stage  # noqa: E402,F401
from b200stage import ops  # noqa: E402

dev = torch.device("cuda:0")
ops._abi.init_device()
d, VT = 4096, 128000
W = (torch.randn((VT, d), dtype=torch.float16, device=dev) * 0.05)
x = torch.randn((1, d), dtype=torch.float16, device=dev)
scratch = ops.head_argmax_scratch(dev)
idx = torch.empty((1,), dtype=torch.int64, device=dev)
val = torch.empty((1,), dtype=torch.float32, device=dev)
for V in (4000, 5336, 8000, 16000, 32000):
    slices = [W[o: o + V] for o in range(0, VT - V + 1, V)]
    n = 400

    def go(k):
        for i in range(k):
            ops.lm_head_argmax(x, slices[i % len(slices)], scratch, idx_offset=0, out_idx=idx, out_val=val)
    go(20)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    go(n)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    mb = V * d * 2 / 1e6
    print(f"V {V:6d} rows = {mb:6.1f} MB: {us:6.2f} us per launch back to back = {mb / us:5.2f} TB/s "
          f"({us - mb / 6.55:5.2f} us over the bytes at 6.55 TB/s)", flush=True)
