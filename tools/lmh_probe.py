"""lm_head (fp16 [32000, 4096]) device time per launch across batch sizes (developer tool): GEMV for 1-4 rows, tcgen05 tile above."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200stage
from b200stage import ops
ops._abi.init_device()
V, d = 32000, 4096
Ws = [torch.randn((V, d), dtype=torch.float16, device="cuda") * 0.02 for _ in range(3)]
for M in (1, 2, 4, 5, 8, 9, 16, 32):
    x = torch.randn((M, d), dtype=torch.float16, device="cuda")
    out = torch.empty((M, V), dtype=torch.float32, device="cuda")
    for W in Ws: ops.lm_head(x, W, out=out)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(10):
        for W in Ws: ops.lm_head(x, W, out=out)
    e.record(); torch.cuda.synchronize()
    print(M, f"{s.elapsed_time(e) / 30 * 1e3:.1f} us")
