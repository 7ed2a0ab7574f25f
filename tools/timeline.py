"""Where one decode step's time goes, CTA by CTA (developer tool).

Runs decode steps of a truncated Llama-2-7B stage through the -DB2_TIMELINE build of the library
(`make -C llm-for-distributed-egde-devices_b200/csrc timeline`), in which thread 0 of every CTA of
the fused decode kernels and of decode attention stamps %globaltimer at:
    0 kernel entry      1 own loads issued, about to wait for the predecessor      2 predecessor done
    3 activation prologue done     4 weight/KV stream done      5 epilogue done
and prints, per launch in time order, when the first/last CTA reached each point (us from the first
event of the step).  Timings under the probes are for the SHAPE of the timeline, not bench numbers.

    B2_LIBRARY=llm-for-distributed-egde-devices_b200/libb200stage_tl.so python tools/timeline.py --layers 4
"""
import argparse
import ctypes as C
import dataclasses
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("B2_LIBRARY", os.path.join(ROOT, "llm-for-distributed-egde-devices_b200", "libb200stage_tl.so"))
import b200stage  # noqa: E402
from b200stage import _abi  # noqa: E402
from b200stage.pipeline import DistPipeline  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="llama-2-7b")
ap.add_argument("--layers", type=int, default=4)
ap.add_argument("--ctx", type=int, default=2100)
ap.add_argument("--steps", type=int, default=3, help="decode steps recorded (the last one is printed)")
ap.add_argument("--batch", type=int, default=1, help="sequences decoded together (one group)")
ap.add_argument("--raw", default="", help="write the raw events of the printed step to this .npy")
a = ap.parse_args()

NAMES = {1: "decode_attn"}
for ok, okn in enumerate(["f16", "gated", "qkv"]):
    for xk, xkn in enumerate(["rms", "ln", "plain", "y32"]):
        NAMES[16 + ok * 4 + xk] = f"fused[{xkn}->{okn}]"

cfg = b200stage.get_config(a.model)
cfg = dataclasses.replace(cfg, n_layers=a.layers, name=f"{cfg.name}[{a.layers} layers]")
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
B = a.batch
pipe = DistPipeline.load(cfg, 0, 1, dev, max_rows=B, max_seq=a.ctx + 64)
lib = _abi.lib()
EV = np.dtype([("t", "<u8"), ("tag", "<u4"), ("cta", "<u4")])


def dump():
    out = []
    for fn in ("b2_tl_dump_fused", "b2_tl_dump_attn"):
        f = getattr(lib, fn)
        f.restype, f.argtypes = C.c_int, [C.c_void_p, C.c_int]
        buf = np.zeros(1 << 20, EV)
        n = f(buf.ctypes.data_as(C.c_void_p), len(buf))
        out.append(buf[:n].copy())
    return np.concatenate(out)


# fill the cache to ctx tokens (the values do not matter for timing), then decode
ids = torch.randint(0, cfg.vocab, (a.ctx,), device=dev)
pipe.prefill_stream([ids] * B, 512)
tok = [torch.zeros(B, dtype=torch.int64, device=dev)]
rows = [list(range(B))]
pipe.decode_stream(tok, [[a.ctx] * B], rows, 4)          # graph capture + warm
torch.cuda.synchronize()
dump()
pos = a.ctx + 4
for _ in range(a.steps):
    torch.cuda.synchronize()
    dump()
    pipe.decode_stream(tok, [[pos] * B], rows, 1)
    pos += 1
torch.cuda.synchronize()
ev = dump()
if a.raw:
    np.save(a.raw, ev)
ev = ev[np.argsort(ev["t"], kind="stable")]
t0 = int(ev["t"].min())
kid = ev["tag"] >> 8
pt = ev["tag"] & 0xFF
# split each kernel id's events into launches: a new launch starts when a CTA index repeats at point 0
launches = []
for k in np.unique(kid):
    e = ev[kid == k]
    p = pt[kid == k]
    cta = e["cta"] & 0xFFFFF
    starts = e[p == 0]
    # number of CTAs per launch = max cta index + 1
    ncta = int(cta.max()) + 1
    nl = len(starts) // ncta
    for li in range(nl):
        rec = {"k": int(k), "ncta": ncta}
        for q in range(6):
            eq = e[p == q]
            seg = eq[li * ncta:(li + 1) * ncta]
            if len(seg):
                tt = (seg["t"].astype(np.int64) - t0) / 1e3
                rec[q] = (tt.min(), np.median(tt), tt.max())
        launches.append(rec)
launches.sort(key=lambda r: r[0][0])
print(f"{'kernel':22s} {'ctas':>5s} | entry first..last | wait-released first..last | prologue done med..last | "
      f"stream done first..med..last | exit last   [us]")
prev_end = 0.0
for r in launches:
    g = lambda q, i: r[q][i] if q in r else float("nan")
    print(f"{NAMES.get(r['k'], r['k']):22s} {r['ncta']:5d} | {g(0,0):7.2f} {g(0,2):7.2f} | {g(2,0):7.2f} {g(2,2):7.2f} | "
          f"{g(3,1):7.2f} {g(3,2):7.2f} | {g(4,0):7.2f} {g(4,1):7.2f} {g(4,2):7.2f} | {g(5,2):7.2f}   "
          f"(+{g(5,2) - prev_end:5.2f} after previous exit)")
    prev_end = g(5, 2)
# per-CTA durations between consecutive points, per kernel type (all launches but the first layer's)
print("\nper-CTA time between points [us]: min / median / p90 / max")
for k in np.unique(kid):
    e = ev[kid == k]
    p = pt[kid == k]
    ncta = int((e["cta"] & 0xFFFFF).max()) + 1
    per = {}
    for q in range(8):
        eq = e[p == q]
        if len(eq):
            per[q] = eq
    # points 6 and 7 sit inside the activation prologue: 6 = x loaded and squared, 7 = norm statistics reduced
    qs = sorted(per, key=lambda q: {6: 0.3, 7: 0.6}.get(q, q))
    nl = len(per[qs[0]]) // ncta
    line = [f"{NAMES.get(int(k), k):22s}"]
    for qa, qb in zip(qs[:-1], qs[1:]):
        d = []
        for li in range(1, nl):
            sa = per[qa][li * ncta:(li + 1) * ncta]
            sb = per[qb][li * ncta:(li + 1) * ncta]
            ma = dict(zip((sa["cta"] & 0xFFFFF).tolist(), sa["t"].tolist()))
            d += [(int(t) - int(ma[c])) / 1e3 for c, t in zip((sb["cta"] & 0xFFFFF).tolist(), sb["t"].tolist())]
        d = np.array(d)
        line.append(f"p{qa}->p{qb} {d.min():5.2f}/{np.median(d):5.2f}/{np.percentile(d, 90):5.2f}/{d.max():5.2f}")
    print("  ".join(line))
res = np.diff(np.unique(ev["t"]))
print(f"globaltimer resolution seen: {res.min()} ns; step span {(int(ev['t'].max()) - t0) / 1e3:.1f} us for {a.layers} layers")
