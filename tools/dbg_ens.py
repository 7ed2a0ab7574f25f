import sys, threading
sys.path.insert(0, '.')
import numpy as np, torch
import b200stage
from b200stage.pipeline import Pipeline
a = Pipeline.load("tiny-phi-test", 1, devices=["cuda:0"], seed=1)
b = Pipeline.load("tiny-neox-test", 1, devices=["cuda:1"], seed=2)
ids = torch.from_numpy(np.random.default_rng(0).integers(0, 512, 48))
kw = dict(max_new_tokens=24, do_sample=True, temperature=0.7, top_k=50, top_p=0.9, repetition_penalty=1.1, seed=6)
def gen(p): return p.generate(ids, **kw).cpu().numpy()
r = [gen(b) for _ in range(3)]
print("main x3 equal:", [np.array_equal(r[0], x) for x in r])
out = {}
def w(name, p): out[name] = gen(p)
t = threading.Thread(target=w, args=("b_alone", b)); t.start(); t.join()
print("worker alone == main:", np.array_equal(out["b_alone"], r[0]))
ra = gen(a)
ts = [threading.Thread(target=w, args=("a", a)), threading.Thread(target=w, args=("b", b))]
[t.start() for t in ts]; [t.join() for t in ts]
print("concurrent b == main:", np.array_equal(out["b"], r[0]), " a == main:", np.array_equal(out["a"], ra))
if not np.array_equal(out["b"], r[0]):
    d = np.nonzero(out["b"] != r[0])[0]; print("first diff at", d[0], "of", len(r[0]))
for k in range(3):
    ts = [threading.Thread(target=w, args=("a", a)), threading.Thread(target=w, args=("b", b))]
    [t.start() for t in ts]; [t.join() for t in ts]
    print("again concurrent b == main:", np.array_equal(out["b"], r[0]), " a == main:", np.array_equal(out["a"], ra))
# greedy too
kw2 = dict(max_new_tokens=24)
g0 = b.generate(ids, **kw2).cpu().numpy()
def w2(name, p): out[name] = p.generate(ids, **kw2).cpu().numpy()
ts = [threading.Thread(target=w2, args=("a", a)), threading.Thread(target=w2, args=("b", b))]
[t.start() for t in ts]; [t.join() for t in ts]
print("greedy concurrent b == main:", np.array_equal(out["b"], g0))
