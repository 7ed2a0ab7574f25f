"""Developer probe: does advancing two pipelines in turn (one host thread) change what either of them generates?"""
import sys
sys.path.insert(0, '.')
import numpy as np, torch
import b200stage
from b200stage.pipeline import Pipeline

ids = torch.from_numpy(np.random.default_rng(0).integers(0, 512, 48))
kw = dict(max_new_tokens=24, do_sample=True, temperature=0.7, top_k=50, top_p=0.9, repetition_penalty=1.1, seed=6)

def run_all(p, **k):
    return p.generate(ids, **{**kw, **k}).cpu().numpy()

def interleave(ps, sync=False, **k):
    its = [p.generate_steps(ids, **{**kw, **k}) for p in ps]
    out, live = [None] * len(ps), list(range(len(ps)))
    while live:
        for i in list(live):
            try:
                next(its[i])
                if sync:
                    torch.cuda.synchronize()
            except StopIteration as d:
                out[i] = d.value.cpu().numpy(); live.remove(i)
    return out

def case(name, ma, da, mb, db, **k):
    a = Pipeline.load(ma, 1, devices=[da], seed=1)
    b = Pipeline.load(mb, 1, devices=[db], seed=2)
    ra, rb = run_all(a, **k), run_all(b, **k)
    ra2, rb2 = run_all(a, **k), run_all(b, **k)
    ia, ib = interleave([a, b], **k)
    sa, sb = interleave([a, b], sync=True, **k)
    print(f"{name:34s} repeat a/b {np.array_equal(ra, ra2)}/{np.array_equal(rb, rb2)}  interleaved a/b "
          f"{np.array_equal(ia, ra)}/{np.array_equal(ib, rb)}  with sync a/b {np.array_equal(sa, ra)}/{np.array_equal(sb, rb)}", flush=True)
    if not np.array_equal(ib, rb):
        print("   b first diff at", int(np.nonzero(ib != rb)[0][0]), "of", len(rb), "(prompt 48)")
    if not np.array_equal(ia, ra):
        print("   a first diff at", int(np.nonzero(ia != ra)[0][0]), "of", len(ra), "(prompt 48)")

two = torch.cuda.device_count() >= 2
d1 = "cuda:1" if two else "cuda:0"
case("phi@0 + neox@1", "tiny-phi-test", "cuda:0", "tiny-neox-test", d1)
case("neox@0 + phi@1", "tiny-neox-test", "cuda:0", "tiny-phi-test", d1)
case("phi@0 + neox@0 (same GPU)", "tiny-phi-test", "cuda:0", "tiny-neox-test", "cuda:0")
case("gqa@0 + llama@1", "tiny-gqa-test", "cuda:0", "tiny-llama-test", d1)
case("gqa@0 + llama@0 (same GPU)", "tiny-gqa-test", "cuda:0", "tiny-llama-test", "cuda:0")
case("phi@0 + neox@1 greedy", "tiny-phi-test", "cuda:0", "tiny-neox-test", d1, do_sample=False)
