"""Developer probe: what one decode tick of an 8-stage Llama-2-7B pipeline costs on ONE GPU, hand-off excluded.

A middle stage (4 layers) with 8 one-row decode runners, as DistPipeline._decode_stream_graphs drives them:
  a) the 8 graphs replayed in turn, nothing between them;
  b) a flag wait (flag already published: it never spins) + copy into the graph input, the replay, and a store + flag
     into this GPU's own memory -- the three launches of a real tick without the NVLink hop;
  c) one graph replayed back to back (same KV rows every time).
usage: python tools/tick_probe.py [layers=4] [ticks=1024]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200stage  # noqa: E402
from b200stage import ops  # noqa: E402
from b200stage.stage import Stage  # noqa: E402

NL = int(sys.argv[1]) if len(sys.argv) > 1 else 4
TICKS = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
dev = torch.device("cuda:0")
cfg = b200stage.get_config("llama-2-7b")
st = Stage.load(cfg, 8, 8 + NL, first=False, last=False, device=dev, seed=1, fast_init=True)
G = 8
cache = st.new_kv_cache(G, 2048 + 160)
runners = []
for g in range(G):
    r = st.runner(cache, 1, decode=True, rows=[g], logits="none")
    r.set_positions(np.array([2048]))
    r.x.normal_()
    runners.append(r)
flag = torch.full((64,), 1 << 30, dtype=torch.int32, device=dev)
inbox = torch.randn((G, cfg.d_model), dtype=torch.float16, device=dev)
outbox = torch.empty_like(inbox)
oflag = torch.zeros((64,), dtype=torch.int32, device=dev)
err = torch.zeros((1,), dtype=torch.int32, device=dev)


def reset():
    for r in runners:
        r.set_positions(np.array([2048]))


def timed(fn, ticks):
    reset()
    for i in range(16):
        fn(i)
    reset()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(ticks):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / ticks


def a(i):
    runners[i % G].replay()


def b(i):
    g = i % G
    r = runners[g]
    ops.p2p_wait(flag[g: g + 1], 1 + i // G, src=inbox[g: g + 1], dst=r.x, err=err)
    r.replay()
    ops.p2p_send(r.out, outbox[g: g + 1], oflag[g: g + 1], 1 + i // G)


def c(i):
    runners[0].replay()


def first_last():
    """The two end stages of the shared-lm_head ring as DistPipeline drives them, on this one GPU."""
    global runners
    from b200stage.stage import head_message
    out = {}
    for name, kw, lay in (("first", dict(first=True, last=False), (0, NL)), ("last", dict(first=False, last=True), (32 - NL, 32))):
        s2 = Stage.load(cfg, lay[0], lay[1], device=dev, seed=1, fast_init=True, **kw)
        c2 = s2.new_kv_cache(G, 2048 + 160)
        want = "split" if name == "last" else "none"
        runners = [s2.runner(c2, 1, decode=True, rows=[g], logits=want) for g in range(G)]
        for r in runners:
            r.set_positions(np.array([2048]))
            if name == "first":
                r.x.fill_(17)
            else:
                r.x.normal_()
        if name == "first":
            msgs = [head_message(1, cfg.d_model, dev) for _ in range(G)]
            for m in msgs:
                m[1].normal_()
            gen = torch.zeros((G, 1), dtype=torch.int64, device=dev)
            lg = torch.empty((1, s2.head_split()), dtype=torch.float32, device=dev)

            def full(i):
                g = i % G
                r = runners[g]
                ops.p2p_wait(flag[g: g + 1], 1 + i // G, err=err)
                s2.head_lower(msgs[g][1], msgs[g][2], msgs[g][3], gen[g], lg, out_tok2=r.x)
                r.replay()
                ops.p2p_send(r.out, outbox[g: g + 1], oflag[g: g + 1], 1 + i // G)

            def nohead(i):
                g = i % G
                r = runners[g]
                ops.p2p_wait(flag[g: g + 1], 1 + i // G, err=err)
                r.replay()
                ops.p2p_send(r.out, outbox[g: g + 1], oflag[g: g + 1], 1 + i // G)
            out[name] = (timed(full, n), timed(nohead, n), timed(a, n))
        else:
            mbox = torch.empty((G, runners[0].msg.numel()), dtype=torch.uint8, device=dev)

            def full(i):
                g = i % G
                r = runners[g]
                ops.p2p_wait(flag[g: g + 1], 1 + i // G, src=inbox[g: g + 1], dst=r.x, err=err)
                r.replay()
                ops.p2p_send(r.msg, mbox[g], oflag[g: g + 1], 1 + i // G)
            out[name] = (timed(full, n), timed(a, n))
        del runners, s2, c2
        torch.cuda.empty_cache()
    return out


n = min(TICKS, 128 * G)         # positions stay inside the cache: 128 steps per runner
ta, tb = timed(a, n), timed(b, n)
tc = timed(c, 128)
print(f"layers per stage {NL}: replay only {ta:.1f} us/tick; wait + replay + send {tb:.1f} us/tick; one runner {tc:.1f} us/tick; "
      f"kernels per graph {runners[0].n_launch}; per layer {ta / NL:.1f} us (N=1 bench: 56.7 us)", flush=True)
fl = first_last()
print(f"first stage (embedding + {NL} layers, lower lm_head half): whole tick {fl['first'][0]:.1f} us; without head_lower + token copy "
      f"{fl['first'][1]:.1f} us; replay only {fl['first'][2]:.1f} us")
print(f"last stage ({NL} layers + final norm + upper lm_head half + shard argmax): whole tick {fl['last'][0]:.1f} us; replay only {fl['last'][1]:.1f} us")
