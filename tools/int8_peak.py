"""On-box yardsticks for the prefill roofline denominator (never on the product path): a plain library int8 GEMM
(torch._int_mm -> cuBLASLt) and bf16 GEMM at 8192^3, burst (a few launches) and sustained (about 4 s), timed with
CUDA events like MEASURED_PEAKS.json was.  Prints one JSON line.

    python tools/int8_peak.py [--n 8192] [--sustain 4.0]
"""
import argparse
import json
import time

import torch

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=8192)
ap.add_argument("--sustain", type=float, default=4.0)
a = ap.parse_args()
dev = "cuda"
n = a.n


def run(fn, ops_):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 0.0
    for _ in range(5):                      # burst: best of five 10-launch bursts
        s.record()
        for _ in range(10):
            fn()
        e.record()
        torch.cuda.synchronize()
        best = max(best, ops_ * 10 / (s.elapsed_time(e) * 1e-3) / 1e12)
    t0 = time.time()
    launches = 0
    s.record()
    while time.time() - t0 < a.sustain:
        for _ in range(50):
            fn()
        launches += 50
        torch.cuda.synchronize()
    e.record()
    torch.cuda.synchronize()
    sus = ops_ * launches / (s.elapsed_time(e) * 1e-3) / 1e12
    return round(best, 1), round(sus, 1)


out = {"n": n}
try:
    A = torch.randint(-127, 128, (n, n), dtype=torch.int8, device=dev)
    B = torch.randint(-127, 128, (n, n), dtype=torch.int8, device=dev).t()      # column-major B, as cuBLASLt wants
    burst, sus = run(lambda: torch._int_mm(A, B), 2.0 * n ** 3)
    out["int8_tops_burst"], out["int8_tops_sustained"] = burst, sus
except Exception as ex:  # noqa: BLE001
    out["int8_error"] = repr(ex)[:200]
try:
    A = torch.randn((n, n), dtype=torch.bfloat16, device=dev)
    B = torch.randn((n, n), dtype=torch.bfloat16, device=dev)
    burst, sus = run(lambda: torch.mm(A, B), 2.0 * n ** 3)
    out["bf16_tflops_burst"], out["bf16_tflops_sustained"] = burst, sus
except Exception as ex:  # noqa: BLE001
    out["bf16_error"] = repr(ex)[:200]
try:
    A = torch.randn((n, n), device=dev).to(torch.float8_e4m3fn)
    B = torch.randn((n, n), device=dev).to(torch.float8_e4m3fn).t()
    one = torch.ones((), device=dev)
    burst, sus = run(lambda: torch._scaled_mm(A, B, scale_a=one, scale_b=one, out_dtype=torch.bfloat16), 2.0 * n ** 3)
    out["fp8_tflops_burst"], out["fp8_tflops_sustained"] = burst, sus
except Exception as ex:  # noqa: BLE001
    out["fp8_error"] = repr(ex)[:200]
print(json.dumps(out), flush=True)
