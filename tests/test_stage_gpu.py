"""Stage / Pipeline parity against the Tier-S oracle on small seeded models (three levels,
SURVEY.md §8c): L1 int8 activations + GEMM outputs bit-exact wherever the inputs are identical
(layer 0 up to attention); L2 fp16 tensors per layer within max-abs <= 2^-5 * max|ref| and
cosine >= 0.9995; L3 greedy tokens identical for 32 steps (oracle top-1 margin reported).

Why L2 is not tighter: attention is the one step that is not bit-reproducible (softmax sum order,
exp2).  A last-bit fp16 difference in its output can move an int8 activation of the NEXT linear by
one LSB (1/127 of that row's absmax, about 0.8 %), and every later quantiser can do the same, so
differences of about 1 % of the tensor maximum are inherent to W8A8 and not a kernel defect."""
import numpy as np
import pytest
import torch

import b200stage
from b200stage import weights
from b200stage.pipeline import Pipeline, split_layers, confidence_score
from b200stage.stage import Stage
from oracle import w8a8 as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _oracle_stages(cfg, n_stages, seed):
    out = []
    for i, (b, e) in enumerate(split_layers(cfg.n_layers, n_stages)):
        sw = weights.build_stage_weights(cfg, b, e, i == 0, i == n_stages - 1, seed, "cpu")
        out.append(weights.to_numpy(sw))
    return out


def _close(got, ref, name):
    got = got.astype(np.float64).ravel()
    ref = ref.astype(np.float64).ravel()
    err = np.abs(got - ref).max()
    cos = float(got @ ref / (np.linalg.norm(got) * np.linalg.norm(ref) + 1e-30))
    assert err <= 2.0 ** -5 * np.abs(ref).max() + 1e-6, (name, err, np.abs(ref).max())
    assert cos >= 0.9995, (name, cos)


@pytest.mark.parametrize("model", ["tiny-llama-test", "tiny-gqa-test", "tiny-window-test", "tiny-hd128-test", "tiny-gpt2-test",
                                   "tiny-llama3-test", "tiny-phi-test", "tiny-neox-test"])
def test_stage_prefill_then_decode_matches_oracle(model):
    cfg = b200stage.get_config(model)
    seed, T, S = 3, 70, 96
    stage = Stage.load(cfg, 0, cfg.n_layers, first=True, last=True, device=DEV, seed=seed)
    osw = _oracle_stages(cfg, 1, seed)[0]
    ids = np.random.default_rng(5).integers(0, cfg.vocab, T)
    cos, sin = O.rope_table(cfg.head_dim, cfg.max_pos, cfg.rope_theta, cfg.rope_scaling, cfg.rotary_dim)
    ocache = O.KVCacheOracle(cfg.n_layers, 1, cfg.n_kv_heads, S, cfg.head_dim)
    otrace = []
    ref = O.stage_forward(ids, osw, cfg.kernel_cfg(), ocache, np.zeros(T, np.int64), np.arange(T), cos, sin, otrace)
    cache = stage.new_kv_cache(1, S)
    trace = []
    got = stage.forward(torch.from_numpy(ids).to(DEV), cache, np.arange(T), logits="all", trace=trace)
    # L1: layer 0, before attention, inputs identical -> bit-exact
    assert np.array_equal(trace[0]["q1"].cpu().numpy(), otrace[0]["q1"])
    assert np.array_equal(trace[0]["s1"].cpu().numpy(), otrace[0]["s1"])
    assert np.array_equal(trace[0]["qkv"].cpu().numpy().view(np.uint16), otrace[0]["qkv"].view(np.uint16))
    # L2: per layer
    for li in range(cfg.n_layers):
        for k in ("attn", "h1", "gu", "h2"):
            _close(trace[li][k].cpu().numpy(), otrace[li][k].reshape(trace[li][k].shape), f"L{li}.{k}")
    _close(got.cpu().numpy(), ref, "logits")
    # KV cache contents
    _close(cache.k[:, 0, :, :T].cpu().numpy(), ocache.k[:, 0, :, :T], "kcache")
    # one decode step on top
    nxt = np.array([int(ref[-1].argmax())])
    ref2 = O.stage_forward(nxt, osw, cfg.kernel_cfg(), ocache, np.zeros(1, np.int64), np.array([T]), cos, sin)
    got2 = stage.forward(torch.from_numpy(nxt).to(DEV), cache, [T])
    _close(got2.cpu().numpy(), ref2, "decode logits")
    assert int(got2.argmax()) == int(ref2.argmax())


def test_graph_replay_equals_eager():
    """The CUDA-graph runners (decode steps, full prefill chunks) must reproduce the eager launches
    bit for bit: same kernels, same inputs."""
    cfg = b200stage.get_config("tiny-gqa-test")
    ids = np.random.default_rng(4).integers(0, cfg.vocab, 3 * 64 + 10)
    outs = []
    for graphs in (False, True):
        pipe = Pipeline.load(cfg, 2, seed=5)
        pipe.use_graphs = graphs
        outs.append(pipe.generate(ids, max_new_tokens=12, chunk=64).cpu().numpy())
        if graphs:   # second call reuses the captured graphs and the pooled caches
            again = pipe.generate(ids, max_new_tokens=12, chunk=64).cpu().numpy()
            assert np.array_equal(again, outs[-1])
            lg_g = pipe.prefill(torch.from_numpy(ids).to(DEV), pipe.pooled_caches(1, 256), chunk=64)
        else:
            lg_e = pipe.prefill(torch.from_numpy(ids).to(DEV), pipe.new_caches(1, 256), chunk=64)
    assert np.array_equal(outs[0], outs[1])
    assert torch.equal(lg_e, lg_g)


def test_chunked_prefill_equals_one_shot():
    cfg = b200stage.get_config("tiny-gqa-test")
    pipe = Pipeline.load(cfg, 1, seed=1)
    ids = torch.from_numpy(np.random.default_rng(0).integers(0, cfg.vocab, 150)).to(DEV)
    a = pipe.prefill(ids, pipe.new_caches(1, 160), chunk=512)
    b = pipe.prefill(ids, pipe.new_caches(1, 160), chunk=64)
    _close(b.cpu().numpy(), a.cpu().numpy(), "chunked")
    assert int(a.argmax()) == int(b.argmax())


@pytest.mark.parametrize("model,n_stages", [("tiny-llama-test", 1), ("tiny-gqa-test", 2), ("tiny-window-test", 2), ("tiny-gpt2-test", 2),
                                            ("tiny-phi-test", 2), ("tiny-neox-test", 1)])
def test_greedy_tokens_identical(model, n_stages):
    cfg = b200stage.get_config(model)
    seed, T, NEW = 7, 40, 32
    pipe = Pipeline.load(cfg, n_stages, seed=seed)
    ids = np.random.default_rng(9).integers(0, cfg.vocab, T)
    got = pipe.generate(ids, max_new_tokens=NEW, chunk=16).cpu().numpy()
    ref, margins = O.greedy_generate(_oracle_stages(cfg, n_stages, seed), cfg.kernel_cfg(), ids, NEW, T + NEW)
    print(f"{model}: min oracle top-1 margin {min(margins):.4f}")
    assert np.array_equal(got, ref), (got[T:], ref[T:], min(margins))


@pytest.mark.parametrize("model", ["tiny-llama-test", "tiny-gqa-test", "tiny-hd128-test"])
def test_greedy_tokens_identical_at_low_margin(model):
    """The seeded models above are very peaked (top-1 margin 20-30 on logits of about 40): token identity there
    cannot tell a correct kernel from a slightly wrong one.  Here the lm_head is an independent random matrix: the
    oracle's margins are about 0.2 on logits of about 2.5 (8 %), many below 4 %.  Teacher-forced on the oracle's
    tokens, the kernels must pick the oracle's token at EVERY step whose margin exceeds 0.05 (2 % of the largest
    logit: a few attention roundings amplified by the int8 quantisers stay below that; a 5 %-wrong kernel does
    not), and may differ only at near-ties."""
    import dataclasses
    cfg = dataclasses.replace(b200stage.get_config(model), tie_embeddings=False)
    seed, T, NEW, TAU = 7, 40, 32, 0.05
    sw = weights.build_stage_weights(cfg, 0, cfg.n_layers, True, True, seed, "cpu")
    sw["lm_head"] = (torch.randn(sw["lm_head"].shape, generator=torch.Generator().manual_seed(5)) * 0.05).to(torch.float16)
    ids = np.random.default_rng(9).integers(0, cfg.vocab, T)
    ref, margins = O.greedy_generate([weights.to_numpy(sw)], cfg.kernel_cfg(), ids, NEW, T + NEW)
    dev_sw = weights.build_stage_weights(cfg, 0, cfg.n_layers, True, True, seed, DEV)
    dev_sw["lm_head"] = sw["lm_head"].to(DEV)
    pipe = Pipeline([Stage(cfg, dev_sw, 0, cfg.n_layers, DEV)])
    caches = pipe.new_caches(1, T + NEW + 8)
    lg = pipe.prefill(torch.from_numpy(ids).to(DEV), caches, 0, 0, 16)
    picks = []
    for i in range(NEW):
        picks.append(int(lg.view(-1).argmax()))
        if i + 1 < NEW:
            lg = pipe.decode_step(torch.tensor([int(ref[T + i])], device=DEV), caches, [T + i])
    picks, want, margins = np.array(picks), np.asarray(ref[T:]), np.array(margins)
    clear = margins > TAU
    print(f"{model}: margins min {margins.min():.4f} median {np.median(margins):.4f}; {int(clear.sum())} of {NEW} steps above {TAU}; "
          f"{int((picks != want).sum())} near-tie steps differ")
    assert np.median(margins) < 0.5 and clear.sum() >= NEW // 2
    assert np.array_equal(picks[clear], want[clear]), (picks, want, margins)
    assert (picks != want).sum() <= NEW // 8


def test_batched_decode_matches_single():
    cfg = b200stage.get_config("tiny-gqa-test")
    pipe = Pipeline.load(cfg, 1, seed=2)
    r = np.random.default_rng(1)
    B, S = 3, 64
    caches = pipe.new_caches(B, S)
    lens = [20, 33, 7]
    prompts = [torch.from_numpy(r.integers(0, cfg.vocab, L)).to(DEV) for L in lens]
    singles = []
    for b in range(B):
        singles.append(pipe.prefill(prompts[b], caches, row=b))
    toks = torch.stack([s.argmax(-1)[0] for s in singles])
    lg = pipe.decode_step(toks, caches, lens, rows=np.arange(B))
    for b in range(B):
        c1 = pipe.new_caches(1, S)
        pipe.prefill(prompts[b], c1)
        one = pipe.decode_step(toks[b:b + 1], c1, [lens[b]])
        _close(lg[b].cpu().numpy(), one[0].cpu().numpy(), f"row{b}")


def test_confidence_score_matches_oracle():
    """confidence_score (models_quant_updated.py:63-70): the [S,V] softmax -> max -> mean reduction runs in
    b2_confidence; its value is checked against the oracle's statement of the same reduction on the same logits
    (fp64 sums in a different order: 1e-6 relative, normally identical) and against torch's softmax."""
    from b200stage import ops
    cfg = b200stage.get_config("tiny-llama-test")
    pipe = Pipeline.load(cfg, 1, seed=0)
    ids = np.arange(30) % cfg.vocab
    c = confidence_score(pipe, ids)
    lg = pipe.logits(ids)
    ref = float(O.confidence(lg.cpu().numpy()))
    assert abs(c - ref) <= 1e-6 * ref, (c, ref)
    tref = float(torch.softmax(lg.float(), dim=-1).max(dim=-1).values.mean().item())
    assert abs(c - tref) <= 1e-5 * tref, (c, tref)
    assert 0.0 < c <= 1.0
    # wide rows, a padded row pitch and extreme logits
    r = np.random.default_rng(3)
    x = (r.standard_normal((37, 32000)) * 6).astype(np.float32)
    x[0, 5] = 80.0
    x[1] = -3.0
    xt = torch.from_numpy(np.pad(x, ((0, 0), (0, 16)))).to(DEV)[:, :32000]
    out, pm = ops.confidence(xt) if xt.is_contiguous() else ops.confidence(xt.contiguous())
    ref = float(O.confidence(x))
    assert abs(float(out) - ref) <= 1e-6 * ref
    assert abs(float(pm[1]) - 1.0 / 32000) <= 1e-6 / 32000 * 10


@pytest.mark.parametrize("model,rows", [("tiny-llama-test", 1), ("tiny-gqa-test", 2), ("tiny-gpt2-test", 1), ("tiny-phi-test", 1)])
def test_decode_stream_on_one_stage_matches_generate(model, rows):
    """The graph-driven decode stream of a one-stage DistPipeline (what bench.py times at N = 1): its greedy tokens
    come from b2_lm_head_argmax inside the graph (heads without a bias; Phi's biased head takes b2_lm_head +
    b2_argmax) and must equal Pipeline.generate's, which goes through logits + b2_argmax."""
    from b200stage.pipeline import DistPipeline
    cfg = b200stage.get_config(model)
    P, NEW, NP = 24, 12, 2 * rows
    pipe = DistPipeline.load(cfg, 0, 1, torch.device(DEV, 0), seed=3, fast_init=False, max_rows=NP, max_seq=P + NEW + 8)
    r = np.random.default_rng(5)
    prompts = [torch.from_numpy(r.integers(0, cfg.vocab, P)).to(DEV) for _ in range(NP)]
    ref = Pipeline.load(cfg, 1, seed=3)
    want = np.stack([ref.generate(p, max_new_tokens=NEW, chunk=16).cpu().numpy()[P:] for p in prompts])
    for rep in range(2):
        lgs = pipe.prefill_stream(prompts, 16)
        first = [lg.view(1, -1).argmax(-1) for lg in lgs]
        groups = [list(range(g * rows, (g + 1) * rows)) for g in range(NP // rows)]
        toks = [torch.cat([first[p] for p in grp]) for grp in groups]
        out = pipe.decode_stream(toks, [[P] * rows for _ in groups], groups, NEW - 1)
        got = np.zeros((NP, NEW), dtype=np.int64)
        for g, grp in enumerate(groups):
            seq = torch.stack([toks[g]] + out[g]).cpu().numpy()          # [NEW, rows]
            for j, p in enumerate(grp):
                got[p] = seq[:, j]
        assert np.array_equal(got, want), (rep, got, want)
    pipe.drain()


def test_stage_refuses_cpu():
    with pytest.raises(RuntimeError):
        Stage.load("tiny-llama-test", 0, 1, first=True, last=False, device="cpu")


def test_pipeline_from_a_checkpoint_directory(tmp_path):
    """Stage.load(model_dir): the reference's `from_pretrained(model_path)` convention (models_quant_updated.py:34-38).
    A tiny HF Llama checkpoint written here (the seeded synthetic weights in the HF module, `save_pretrained`) is
    read per layer range by two stages and must decode the tokens of the oracle on the same weights; a directory
    with a config but no weights must raise instead of yielding a random model."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import make_golden
    cfg = b200stage.get_config("tiny-gqa-test")
    seed, T, NEW = 6, 33, 16
    make_golden.hf_model(cfg, seed).save_pretrained(tmp_path)
    pipe = Pipeline.load(str(tmp_path), 2)
    assert pipe.stages[0].w["source"] == str(tmp_path) and pipe.stages[1].layer_begin == cfg.n_layers // 2
    ids = np.random.default_rng(2).integers(0, cfg.vocab, T)
    got = pipe.generate(ids, max_new_tokens=NEW, chunk=16).cpu().numpy()
    ref, margins = O.greedy_generate(_oracle_stages(cfg, 2, seed), cfg.kernel_cfg(), ids, NEW, T + NEW)
    assert np.array_equal(got, ref), (got[T:], ref[T:], min(margins))
    empty = tmp_path / "config_only"
    empty.mkdir()
    make_golden.hf_model(cfg, seed).config.save_pretrained(empty)
    with pytest.raises(FileNotFoundError):
        Stage.load(str(empty), 0, 1, first=True, last=False, device=DEV)


def test_ensemble_on_one_gpu():
    """answerer, answerer -> refiner (combiner_quant_updated.py:177-203) with the three pipelines on one GPU: the
    answerers then run one after the other, as in the reference; the result is reproducible under a seed."""
    from b200stage.ensemble import Ensemble

    class Tok:
        eos_token_id = None

        def __init__(self, vocab):
            self.vocab = vocab

        def encode(self, text):
            return [(sum(w.encode()) * 31 + i) % self.vocab for i, w in enumerate(text.split())][:48] or [1]

        def decode(self, ids):
            return " ".join(str(int(i)) for i in ids)
    pipes = [Pipeline.load(m, 1, seed=i + 1) for i, m in enumerate(("tiny-phi-test", "tiny-neox-test", "tiny-llama3-test"))]
    ens = Ensemble([(pipes[0], Tok(pipes[0].cfg.vocab)), (pipes[1], Tok(pipes[1].cfg.vocab))], (pipes[2], Tok(pipes[2].cfg.vocab)))
    r1 = ens("what is the capital of france", max_new_tokens=12, seed=3)
    r2 = ens("what is the capital of france", max_new_tokens=12, seed=3)
    assert r1["answers"] == r2["answers"] and r1["refined"] == r2["refined"] and r1["tps"] > 0
    assert all(len(a.split()) >= 7 + 12 for a in r1["answers"])      # prompt + new tokens, decoded
