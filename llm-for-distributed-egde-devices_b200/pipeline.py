"""Pipeline: stages chained over contiguous layer ranges, with greedy ``generate``.

Mirrors the subset of HF ``model.generate`` / ``model(**inputs).logits`` that the reference's
scripts call (Code/Quantised Models/models_quant_updated.py:84-93 and :66-67) on top of Stage.
Two forms:

* ``Pipeline(stages)``      -- all stages live in this process (one GPU, or several GPUs driven
  from one process; hidden states hop with a device-to-device copy).
* ``DistPipeline(stage)``   -- one process per GPU (torchrun); the hidden-state hand-off that the
  reference sketches as a gRPC hop between Jetsons (Code/gRPC/server.py:13-19, client.py:7-11;
  "Deploy models across Jetson ..." in Presentations/review 1.pptx) is an NCCL send/recv over
  NVLink for prefill chunks and a peer-memory store + flag (csrc/p2p.cu) for decode steps, with the
  lm_head shared between stages (two-way split, or one slice per middle stage on deep pipelines:
  ``ring_head_schedule``).  The KV cache is stage-local and never moves.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

from . import ops
from .config import ModelConfig, get_config
from .stage import Stage, KVCache, head_message, head_message_views, head_message_bytes


def split_layers(n_layers: int, n_stages: int):
    """Contiguous, near-equal layer ranges: stage i owns [b_i, e_i)."""
    base, rem = divmod(n_layers, n_stages)
    out, b = [], 0
    for i in range(n_stages):
        e = b + base + (1 if i < rem else 0)
        out.append((b, e))
        b = e
    return out


class Pipeline:
    def __init__(self, stages):
        if not stages or not stages[0].first or not stages[-1].last:
            raise ValueError("need stages[0].first and stages[-1].last")
        self.stages = list(stages)
        self.cfg: ModelConfig = stages[0].cfg
        self.use_graphs = True        # decode steps and full-size prefill chunks replay CUDA graphs
        self._cache_pool = {}

    @classmethod
    def load(cls, model, n_stages: int = 1, devices=None, seed: int = 0, fast_init: bool = False,
             quant: str = "w8a8", act_stats=None, calibrate_with=None) -> "Pipeline":
        """`model`: a Hugging Face model directory (its tensors are read per layer range), or a preset name /
        ModelConfig for the seeded synthetic model.  calibrate_with: prompts for the SmoothQuant statistics."""
        from . import loader
        cfg = get_config(model)
        devices = devices or ["cuda:0"] * n_stages
        rngs = split_layers(cfg.n_layers, n_stages)
        if calibrate_with is not None and act_stats is None:
            act_stats = loader.calibrate(model, calibrate_with, seed, devices[0])
        spec = model if loader.is_checkpoint_dir(model) else cfg
        return cls([Stage.load(spec, b, e, first=(i == 0), last=(i == n_stages - 1), device=devices[i], seed=seed,
                               fast_init=fast_init, quant=quant, act_stats=act_stats) for i, (b, e) in enumerate(rngs)])

    def new_caches(self, rows: int, max_seq: int):
        return [s.new_kv_cache(rows, max_seq) for s in self.stages]

    def pooled_caches(self, rows: int, max_seq: int):
        """Caches owned by the pipeline and reused across generate() calls, so the CUDA graphs
        captured against them are reused too."""
        key = (rows, max_seq)
        if key not in self._cache_pool:
            self._cache_pool[key] = self.new_caches(rows, max_seq)
        for c in self._cache_pool[key]:
            c.reset()
        return self._cache_pool[key]

    @property
    def launches(self) -> int:
        return sum(s.launches for s in self.stages)

    # ------------------------------------------------------------------------------------------
    def _run(self, x, caches, positions, seq_ids, logits, graph_chunk: int = 0):
        for s, c in zip(self.stages, caches):
            if x.device != s.device:
                x = x.to(s.device, non_blocking=True)
            want = logits if s.last else "none"
            if self.use_graphs:
                x = s.forward_auto(x, c, positions, seq_ids, logits=want, graph_chunk=graph_chunk)
            else:
                x = s.forward(x, c, positions, seq_ids, logits=want)
        return x

    def prefill(self, ids, caches, row: int = 0, pos0: int = 0, chunk: int = 512, logits: str = "last"):
        """Chunked prefill of one prompt into cache row `row`; returns the logits of the final
        position ("last") or of every position ("all")."""
        T = int(ids.shape[0])
        outs = []
        last = None
        for c0 in range(0, T, chunk):
            c1 = min(T, c0 + chunk)
            pos = np.arange(pos0 + c0, pos0 + c1)
            final = c1 == T
            want = "all" if logits == "all" else ("last" if final else "none")
            y = self._run(ids[c0:c1], caches, pos, np.full(c1 - c0, row), want, graph_chunk=chunk if T > chunk else 0)
            if logits == "all":
                outs.append(y.clone())
            elif final:
                last = y
        return torch.cat(outs, 0) if logits == "all" else last

    def decode_step(self, tokens, caches, positions, rows=None):
        """One token per sequence -> logits [B,V]."""
        B = int(tokens.shape[0])
        rows = np.arange(B) if rows is None else np.asarray(rows)
        return self._run(tokens, caches, np.asarray(positions), rows if B > 1 else np.asarray(rows[:1]), "last")

    def generate(self, input_ids, max_new_tokens: int = 32, do_sample: bool = False, chunk: int = 512,
                 max_seq: int | None = None, caches=None, eos_token_id=None, pad_token_id=None, **sample_kwargs):
        """Continuation of one prompt; returns prompt + new tokens (int64, on the device), like
        ``model.generate(**inputs, max_new_tokens=...)[0]`` (models_quant_updated.py:84-95): greedy, or with
        do_sample=True the reference's temperature / top_k / top_p / repetition_penalty chain.
        eos_token_id (int or list) stops the generation as HF does for one sequence: the result ends with the first
        EOS token produced.  The decode loop stays asynchronous; the host looks for an EOS every 8 steps.
        pad_token_id is accepted for call compatibility (one sequence is never padded).
        Sampling note: top_k <= 0 ("off" in HF) or > 1024 is capped at 1024 candidates (b2_sample)."""
        it = self.generate_steps(input_ids, max_new_tokens, do_sample, chunk, max_seq, caches, eos_token_id, **sample_kwargs)
        while True:
            try:
                next(it)
            except StopIteration as done:
                return done.value

    def generate_steps(self, input_ids, max_new_tokens: int = 32, do_sample: bool = False, chunk: int = 512,
                       max_seq: int | None = None, caches=None, eos_token_id=None, **sample_kwargs):
        """`generate` as a generator: yields after the prefill has been enqueued and after every decode step has been
        enqueued (nothing waits for the GPU in between unless an EOS check is due); the generator's return value is
        generate's result.  Lets ONE host thread keep several pipelines on several GPUs busy at the same time by
        advancing them in turn (ensemble.py)."""
        temperature = float(sample_kwargs.pop("temperature", 1.0))
        top_k = int(sample_kwargs.pop("top_k", 50))
        top_p = float(sample_kwargs.pop("top_p", 1.0))
        rep = float(sample_kwargs.pop("repetition_penalty", 1.0))
        seed = int(sample_kwargs.pop("seed", 0))
        if sample_kwargs:
            raise TypeError(f"unsupported generate() arguments: {sorted(sample_kwargs)}")
        with torch.no_grad():
            dev = self.stages[0].device
            ids = torch.as_tensor(input_ids, dtype=torch.int64).view(-1).to(dev)
            T = int(ids.shape[0])
            max_seq = max_seq or min(self.cfg.max_pos, T + max_new_tokens)
            if T + max_new_tokens > max_seq:
                raise ValueError("prompt + max_new_tokens exceeds max_seq")
            eos = None
            if eos_token_id is not None:
                eos = torch.as_tensor([eos_token_id] if isinstance(eos_token_id, int) else list(eos_token_id),
                                      dtype=torch.int64, device=dev)
            caches = caches or self.pooled_caches(1, max_seq)
            out = torch.empty((T + max_new_tokens,), dtype=torch.int64, device=dev)
            out[:T] = ids
            lg = self.prefill(ids, caches, 0, 0, chunk)
        yield 0
        hist2d = out.view(1, -1)
        hlen = torch.empty((1,), dtype=torch.int32, device=dev) if do_sample else None
        n_done = max_new_tokens
        for i in range(max_new_tokens):
            with torch.no_grad():
                tok = out[T + i: T + i + 1]
                with torch.cuda.device(dev):
                    lgd = lg.view(1, -1).to(dev) if lg.device != dev else lg.view(1, -1)
                    if do_sample:
                        # the reference's call: do_sample=True, temperature, top_k, top_p, repetition_penalty
                        # (models_quant_updated.py:84-93); the penalty sees prompt + everything generated so far
                        hlen.fill_(T + i)
                        ops.sample(lgd, temperature=temperature, top_k=top_k, top_p=top_p, repetition_penalty=rep, seed=seed,
                                   step=i, history=hist2d, hist_len=hlen, out=tok)
                    else:
                        ops.argmax(lgd, out=tok)
                self.stages[-1].launches += 1
                if eos is not None and (i % 8 == 7 or i + 1 == max_new_tokens):
                    hit = torch.isin(out[T: T + i + 1], eos).nonzero()
                    if hit.numel():
                        n_done = int(hit[0]) + 1
                        break
                if i + 1 < max_new_tokens:
                    lg = self.decode_step(tok, caches, [T + i])
            yield i + 1
        return out[: T + n_done]

    @torch.no_grad()
    def logits(self, input_ids, chunk: int = 512):
        """All-position logits of one prompt: the ``model(**inputs).logits`` form that the
        reference's confidence_score consumes (models_quant_updated.py:63-70)."""
        dev = self.stages[0].device
        ids = torch.as_tensor(input_ids, dtype=torch.int64).view(-1).to(dev)
        caches = self.new_caches(1, int(ids.shape[0]))
        return self.prefill(ids, caches, 0, 0, chunk, logits="all")


def confidence_score(pipe: Pipeline, input_ids) -> float:
    """mean over positions of max softmax probability (models_quant_updated.py:63-70)."""
    lg = pipe.logits(input_ids)
    with torch.cuda.device(lg.device):
        out, _ = ops.confidence(lg)
    pipe.stages[-1].launches += 2
    return float(out.item())


# ==============================================================================================
# one process per GPU
# ==============================================================================================
def ring_head_ranks(world: int):
    """The stages that hold a slice of the shared lm_head: from four stages up only the MIDDLE ones -- the first
    stage (embedding, merge) and the last (final norm, the send to all) are the busy ones of a tick as it is, and a
    slice's launch costs more than its bytes (profiles/r02_tick_trace_n8.txt) -- else all of them."""
    return list(range(1, world - 1)) if world >= 4 else list(range(world))


def ring_head_schedule(rank: int, world: int, groups: int, steps: int):
    """What stage `rank` of a `world`-stage pipeline enqueues, in order, for `steps` decode steps of `groups`
    sequence groups when the lm_head is shared by all stages (DistPipeline._decode_stream_ring).  Yields
    (op, group, step):
      "merge"   first stage: wait for every stage's winner of (group, step), write the token (the next input)
      "wait_h"  wait for the previous stage's hidden state of (group, step)
      "layers"  this stage's layer graph for (group, step)
      "send_h"  store the result into the next stage's inbox     "send_hn"  last stage: normed hidden state to ALL
      "slice"   this stage's slice of the head for (group, step): waits for "send_hn", stores its winner on stage 0
                (only on the stages of ring_head_ranks)
    Stage r runs world-1-r ticks ahead of the last stage, so its slice for the group that is leaving the last stage
    follows its own tick that much later.  Every wait is for something an EARLIER entry of another stage's list
    produces as long as groups >= world (tests/test_host_cpu.py runs the lists against each other)."""
    first, last = rank == 0, rank == world - 1
    lag = world - 1 - rank
    has_slice = rank in ring_head_ranks(world)
    for i in range(steps):
        for g in range(groups):
            if first:
                if i > 0:
                    yield ("merge", g, i - 1)
            else:
                yield ("wait_h", g, i)
            yield ("layers", g, i)
            yield ("send_hn" if last else "send_h", g, i)
            t = i * groups + g - lag
            if t >= 0 and has_slice:
                yield ("slice", t % groups, t // groups)
    for t in range(max(0, steps * groups - lag), steps * groups if has_slice else 0):
        yield ("slice", t % groups, t // groups)
    if first:
        for g in range(groups):
            yield ("merge", g, steps - 1)


class DistPipeline:
    """Rank r owns stage r of `world` stages.  Hidden tiles [T,d] fp16 move r -> r+1 with
    torch.distributed send/recv (NCCL on GPUs; gloo in the CPU schedule tests); the greedy token
    moves from the last rank back to rank 0.  All ranks make the same sequence of calls.

    `forward_fn(x, positions, seq_ids, logits)` abstracts the local stage so the scheduling and
    messaging logic can be exercised on CPU with a stand-in stage (tests/test_host_cpu.py).
    """

    def __init__(self, rank: int, world: int, d_model: int, device, forward_fn, vocab: int, group=None):
        import torch.distributed as dist
        self.dist, self.rank, self.world, self.d, self.device = dist, rank, world, d_model, device
        self.fwd, self.vocab, self.group = forward_fn, vocab, group
        self.first, self.last = rank == 0, rank == world - 1

    @classmethod
    def load(cls, model, rank: int, world: int, device, seed: int = 0, fast_init: bool = True, max_rows: int = 1,
             max_seq: int = 4096, quant: str = "w8a8", calibrate_with=None):
        from . import loader
        cfg = get_config(model)
        b, e = split_layers(cfg.n_layers, world)[rank]
        from_dir = loader.is_checkpoint_dir(model)
        stage = Stage.load(model if from_dir else cfg, b, e, first=(rank == 0), last=(rank == world - 1), device=device,
                           seed=seed, fast_init=fast_init and not from_dir, quant=quant, calibrate_with=calibrate_with)
        cache = stage.new_kv_cache(max_rows, max_seq)
        self = cls(rank, world, cfg.d_model, torch.device(device), None, cfg.vocab)
        self.stage, self.cache, self.cfg = stage, cache, cfg
        self.graph_chunk = 0          # set to the prefill chunk size to replay full chunks as graphs
        # Decode with two or more stages: the last stage streams its layers AND the fp16 lm_head (Llama-2-7B: 262 MB
        # on top of 4 x 202 MB at eight stages), so the pipeline ticks at its pace.  With tied embeddings the first
        # stage already holds the same matrix: it takes the lower half of the vocabulary, the last stage the upper
        # half, and instead of the token the last stage returns its normed hidden state plus its shard's winner.
        # Exact: a row's argmax is the better of the two shard winners under the lowest-index tie rule.
        self.split_head = world > 1 and cfg.tie_embeddings and os.environ.get("B2_SPLIT_HEAD", "1") != "0"
        # Deep pipelines: even a half of the head on each end stage keeps those two ~28 us per tick behind the
        # others, and with N sequences on N stages the ring ticks at the pace of its slowest stage
        # (tools/tick_probe.py).  ring_head: the MIDDLE stages (all of them below four stages) hold
        # vocab / n rows of the head each; the last stage sends its normed hidden state to them at once, each finds
        # its slice's winner, and the first stage merges the winners (_decode_stream_ring).  Measured on
        # Llama-2-7B, 8 stages: 59.0k tok/s (57.4k with a slice on all eight stages) vs 56.7k with the two-way
        # split; 4 stages, a slice on all four: 31.71k vs 31.74k, so it is the default from five stages up.
        # B2_RING_HEAD=0 turns it off, =2 uses it from two stages up (the tests do).
        rh = os.environ.get("B2_RING_HEAD", "1")
        self.ring_head = rh != "0" and world > (1 if rh == "2" else 4) and not cfg.head_bias
        self.fwd = lambda x, pos, seq, lg: stage.forward_auto(x, cache, pos, seq, logits=lg,
                                                              graph_chunk=self.graph_chunk)
        if world > 1:
            self.setup_groups()
            self.setup_p2p(max_rows)
        return self

    # -- peer-memory hand-off -----------------------------------------------------------------------
    def setup_p2p(self, max_rows: int):
        """Collective.  Give every stage an inbox in its own HBM that its neighbours can store into over NVLink
        (decode steps only; prefill chunks keep NCCL, they are bandwidth- not latency-bound): per cache row one slot
        for the incoming hidden state and, on the first stage, one for the last stage's head message or token, plus
        a sequence flag per group and edge.  The mapping comes from torch's symmetric memory, else from CUDA IPC
        handles exchanged through the process group; if neither works on this box the hand-off stays on NCCL
        send/recv (self.p2p is None).  B2_HANDOFF=nccl forces that."""
        self.p2p = None
        if self.world < 2 or self.device.type != "cuda" or os.environ.get("B2_HANDOFF", "p2p") == "nccl":
            return
        if torch.cuda.device_count() < self.world:
            return          # ranks sharing a GPU: their flag-polling kernels are not guaranteed to run at the same time
        dist, dev, R = self.dist, self.device, int(max_rows)
        slot = (2 * self.d + 16 + 255) & ~255
        off_head, off_flags = R * slot, 2 * R * slot
        nbytes = off_flags + 2 * R * 4 + 256
        # the lm_head shared by all stages (ring_head): per row a slot for the last stage's normed hidden state; per
        # group and stage a 64-byte winner record (read on the first stage only) and a flag; per group a flag for hn
        W = self.world
        off_hn = (nbytes + 255) & ~255
        off_win = off_hn + R * slot
        off_wflag = off_win + R * W * 64
        off_nflag = off_wflag + ((R * W * 4 + 255) & ~255)
        nbytes = off_nflag + R * 4 + 256

        def agree(ok: bool) -> bool:       # every rank takes the same branch
            t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            return bool(t.item())
        ring = bool(getattr(self, "ring_head", False))
        need = sorted(set(range(self.world)) - {self.rank}) if ring else sorted({0 if self.last else self.rank + 1})   # peers this rank stores into
        buf, peers, how = None, {}, None
        try:
            import torch.distributed._symmetric_memory as symm
            buf = symm.empty(nbytes, dtype=torch.uint8, device=dev)
            hdl = symm.rendezvous(buf, dist.group.WORLD)
            peers = {r: hdl.get_buffer(r, (nbytes,), torch.uint8) for r in need}
            ok, how = True, "torch symmetric memory"
        except Exception as e:          # noqa: BLE001
            ok, why = False, repr(e)[:200]
        if not agree(ok):
            buf, peers = None, {}
            try:
                from torch.multiprocessing.reductions import reduce_tensor
                buf = torch.zeros((nbytes,), dtype=torch.uint8, device=dev)
                objs = [None] * self.world
                dist.all_gather_object(objs, reduce_tensor(buf))
                for r in need:
                    fn, args = objs[r]
                    peers[r] = fn(*args)
                    torch.empty((16,), dtype=torch.uint8, device=dev).copy_(peers[r][:16])     # enables peer access
                ok, how = True, "CUDA IPC handles"
            except Exception as e:      # noqa: BLE001
                ok, why = False, repr(e)[:200]
            if not agree(ok):
                return
        buf.zero_()
        torch.cuda.synchronize(dev)
        dist.barrier()
        err = torch.zeros((1,), dtype=torch.int32, device=dev)
        nxt = 0 if self.last else self.rank + 1
        self.p2p = dict(buf=buf, peer=peers[nxt], slot=slot, off_head=off_head, off_flags=off_flags, rows=R, how=how,
                        err=err, base_h={}, base_t={}, base_n={}, peers={**peers, self.rank: buf} if ring else None,
                        off_hn=off_hn, off_win=off_win, off_wflag=off_wflag, off_nflag=off_nflag)

    def _p2p_views(self, buf, g: int, row0: int, B: int):
        """(hidden slot, head slot, hidden flag, head flag) of group g in `buf` (this rank's inbox or a peer's)."""
        P = self.p2p
        hs = buf[row0 * P["slot"]: row0 * P["slot"] + B * self.d * 2]
        ts = buf[P["off_head"] + row0 * P["slot"]: P["off_head"] + row0 * P["slot"] + head_message_bytes(B, self.d)]
        fl = buf[P["off_flags"]: P["off_flags"] + 2 * P["rows"] * 4].view(torch.int32)
        return hs, ts, fl[g: g + 1], fl[P["rows"] + g: P["rows"] + g + 1]

    # -- messaging --------------------------------------------------------------------------------
    # On GPUs every receive is posted from a side stream (so NCCL does not make it wait for compute
    # that is merely queued ahead of it) and every send runs on another side stream on a private
    # copy of the tile (so compute never waits for a send, and a graph's static output buffer can be
    # overwritten at once).  The compute stream waits only where it consumes received data.
    def setup_groups(self):
        """Collective: a separate communicator for the token return path, so that a token send
        queued on one rank can never sit in front of a hidden-state transfer (and vice versa)."""
        self.tok_group = self.dist.new_group(ranks=list(range(self.world))) if self.world > 1 else None
        # Two more communicators for the hidden-state edges, alternating along the pipeline (edge r -> r+1
        # uses number r % 2), so that a stage's receive from its predecessor and its send to its successor
        # never queue behind each other on one communicator's stream.
        self._edge_groups = [self.dist.new_group(ranks=list(range(self.world))) for _ in range(2)] if self.world > 1 else None
        if self._edge_groups is not None:
            self.group_in = self._edge_groups[(self.rank - 1) % 2]
            self.group_out = self._edge_groups[self.rank % 2]
        if self.device.type == "cuda":
            self._rs, self._ss = torch.cuda.Stream(self.device), torch.cuda.Stream(self.device)
            self._rs_tok, self._ss_tok = torch.cuda.Stream(self.device), torch.cuda.Stream(self.device)

    def _recv(self, shape, dtype, src, tok=False):
        buf = torch.empty(shape, dtype=dtype, device=self.device)
        group = getattr(self, "tok_group", None) if tok else getattr(self, "group_in", self.group)
        if self.device.type != "cuda" or not hasattr(self, "_rs"):
            self.dist.recv(buf, src=src, group=group)
            return buf
        side = self._rs_tok if tok else self._rs
        cur = torch.cuda.current_stream(self.device)
        side.wait_stream(cur)              # the allocation above is stream-ordered on `cur`
        with torch.cuda.stream(side):
            self.dist.irecv(buf, src=src, group=group).wait()
        cur.wait_stream(side)
        buf.record_stream(side)
        return buf

    def _send(self, t, dst, tok=False):
        group = getattr(self, "tok_group", None) if tok else getattr(self, "group_out", self.group)
        if self.device.type != "cuda" or not hasattr(self, "_ss"):
            # CPU (gloo tests): a blocking send could sit in front of the receive the peer needs
            # first; post it without waiting and reap finished handles later
            if not hasattr(self, "_pending"):
                self._pending = []
            self._pending = [(w, b) for w, b in self._pending if not w.is_completed()]
            c = t.clone()
            self._pending.append((self.dist.isend(c, dst=dst, group=group), c))
            return
        side = self._ss_tok if tok else self._ss
        cur = torch.cuda.current_stream(self.device)
        c = t.clone()                      # private copy: the caller may overwrite `t` immediately
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            self.dist.isend(c, dst=dst, group=group).wait()
        c.record_stream(side)

    def drain(self):
        """Wait for every message this rank has posted (call before tearing the group down)."""
        for w, _ in getattr(self, "_pending", []):
            w.wait()
        self._pending = []
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
            P = getattr(self, "p2p", None)
            if P is not None and int(P["err"].item()) != 0:
                raise RuntimeError("a peer-memory hand-off timed out: a neighbouring stage never published its step")

    # -- one hop --------------------------------------------------------------------------------
    def step(self, x, positions, seq_ids, want_logits: str):
        """x: ids on rank 0 (ignored elsewhere).  Returns logits on the last rank, else None."""
        T = len(positions)
        if not self.first:
            x = self._recv((T, self.d), torch.float16, self.rank - 1)
        y = self.fwd(x, positions, seq_ids, want_logits if self.last else "none")
        if not self.last:
            self._send(y, self.rank + 1)
            return None
        return y

    def return_tokens(self, tok):
        """Greedy ids [B] int64: last rank -> rank 0.  Returns them on rank 0 and on the last rank,
        None on the ranks in between."""
        if self.world == 1:
            return tok
        if self.last:
            self._send(tok, 0, tok=True)
            return tok
        if self.first:
            return self._recv(tuple(tok.shape), torch.int64, self.world - 1, tok=True)
        return None

    # -- schedules --------------------------------------------------------------------------------
    def prefill_stream(self, prompts, chunk: int, pos0: int = 0):
        """Pipelined chunked prefill of several prompts (cache row i for prompt i).  Chunk c of
        prompt p enters stage s as soon as stage s has finished the previous chunk; because send
        and recv are stream-ordered, issuing the chunks in order on every rank IS the pipeline
        schedule (chunk c on stage s at tick c+s).  Returns, on the last rank, the final-position
        logits per prompt."""
        outs = []
        for p, ids in enumerate(prompts):
            T = int(ids.shape[0]) if ids is not None else None
            T = self._bcast_len(T)
            for c0 in range(0, T, chunk):
                c1 = min(T, c0 + chunk)
                pos = np.arange(pos0 + c0, pos0 + c1)
                y = self.step(ids[c0:c1] if self.first else None, pos, np.full(c1 - c0, p),
                              "last" if c1 == T else "none")
                if self.last and c1 == T:
                    outs.append(y.clone())
        return outs

    def _bcast_len(self, T):
        # prompt lengths are part of the (synthetic) workload description known to every rank
        return int(T)

    def decode_stream(self, first_tokens, start_positions, rows_per_group, steps: int):
        """`steps` greedy decode steps for each of G independent groups as ONE continuous pipeline:
        group g's token of step i-1 is awaited only right before group g's step i, so with
        G >= world every stage is busy every tick (no round barrier).  first_tokens[g]: int64 [B_g]
        (needed on rank 0).  Returns, on rank 0 (and on the last rank unless the lm_head is shared with the first
        stage, `split_head`), tokens[g] = list of `steps` tensors [B_g]; None on the other ranks."""
        G = len(rows_per_group)
        if self.device.type == "cuda" and getattr(self, "stage", None) is not None and getattr(self, "use_graphs", True):
            return self._decode_stream_graphs(first_tokens, start_positions, rows_per_group, steps)
        gen = [[] for _ in range(G)]
        cur = list(first_tokens) if self.first else [None] * G
        for i in range(steps):
            for g in range(G):
                rows = rows_per_group[g]
                pos = [p + i for p in start_positions[g]]
                if i > 0 and self.first and not self.last:
                    cur[g] = self._recv((len(rows),), torch.int64, self.world - 1, tok=True)
                    gen[g].append(cur[g])
                lg = self.step(cur[g] if self.first else None, pos, rows, "last")
                if self.last:
                    t = torch.empty((len(rows),), dtype=torch.int64, device=self.device)
                    self._argmax(lg, t)
                    gen[g].append(t)
                    if self.first:
                        cur[g] = t
                    else:
                        self._send(t, 0, tok=True)
        if self.first and not self.last:
            for g in range(G):
                gen[g].append(self._recv((len(rows_per_group[g]),), torch.int64, self.world - 1, tok=True))
        return gen if (self.first or self.last) else None

    def _decode_stream_graphs(self, first_tokens, start_positions, rows_per_group, steps: int):
        """decode_stream on GPUs with as little host work per pipeline tick as possible (at 8 stages
        a tick is ~260 us of device time; the generic path spends ~250 us of Python on it).  Each
        group owns a captured graph (`GraphRunner`); hidden states are received straight into the
        graph's static input and sent straight from its static output, tokens are written into a
        preallocated [steps, B] tensor; sends are ordered with per-group events."""
        dev = self.device
        cur = torch.cuda.current_stream(dev)
        G = len(rows_per_group)
        multi = self.world > 1
        P0 = getattr(self, "p2p", None) if multi else None
        if (P0 is not None and getattr(self, "ring_head", False) and P0.get("peers") and self.world <= G <= P0["rows"]
                and sum(len(r_) for r_ in rows_per_group) <= P0["rows"] and all(len(r_) == 1 for r_ in rows_per_group)):
            # (one sequence per group: the case measured and tested on 2-8 GPUs; micro-batches keep the two-way split)
            return self._decode_stream_ring(first_tokens, start_positions, rows_per_group, steps)
        split = multi and getattr(self, "split_head", False)
        self.head_path = "first + last stage" if split else "last stage"
        want = ("split" if split else "last") if self.last else "none"
        # greedy token straight from the head's dot products (b2_lm_head_argmax) wherever the step is <= 4 rows
        tokmode = want == "last" and all(self.stage.can_fuse_head(len(r_)) for r_ in rows_per_group)
        if tokmode:
            want = "token"
        runners = []
        for g in range(G):
            r = self.stage.runner(self.cache, len(rows_per_group[g]), decode=True, rows=list(rows_per_group[g]), logits=want)
            pos = np.asarray(start_positions[g], dtype=np.int64).reshape(-1)
            if int(pos.max()) + steps > self.cache.max_seq:
                raise ValueError("decode would run past the KV cache capacity")
            if not np.array_equal(pos, r.host_pos):
                r.set_positions(pos)
            if self.first:
                r.x.copy_(first_tokens[g].view(r.x.shape))
            runners.append(r)
        keep = self.first or (self.last and not split)
        gen = [torch.empty((steps, len(rows_per_group[g])), dtype=torch.int64, device=dev) for g in range(G)] if keep else None
        P = getattr(self, "p2p", None) if multi else None
        if P is not None and (not split or sum(len(r_) for r_ in rows_per_group) > P["rows"] or G > P["rows"]):
            P = None            # (the token-only return path and oversize batches stay on NCCL)
        if P is not None:
            row0 = np.concatenate([[0], np.cumsum([len(r_) for r_ in rows_per_group])]).astype(int)
            mine = [self._p2p_views(P["buf"], g, int(row0[g]), len(rows_per_group[g])) for g in range(G)]
            peer = [self._p2p_views(P["peer"], g, int(row0[g]), len(rows_per_group[g])) for g in range(G)]
            bh = [P["base_h"].get(g, 0) for g in range(G)]
            bt = [P["base_t"].get(g, 0) for g in range(G)]
        if split and self.first:
            # per group: the last stage's message (hidden | shard maxima | indices) and the lower-shard logits
            if P is not None:       # the message lands in this GPU's inbox slot: read it in place
                inbox = [(mine[g][1],) + head_message_views(mine[g][1], len(rows_per_group[g]), self.d) for g in range(G)]
            else:
                inbox = [head_message(len(rows_per_group[g]), self.d, dev) for g in range(G)]
            lg_lo = [torch.empty((len(rows_per_group[g]), self.stage.head_split()), dtype=torch.float32, device=dev)
                     for g in range(G)]
        if multi:
            # the buffers used below were last touched by work queued on `cur` (earlier calls, the allocations above)
            for side in (self._rs, self._ss, self._rs_tok, self._ss_tok):
                side.wait_stream(cur)
        ev_sent = [None] * G        # on the send stream: runner g's output has left (the next replay may overwrite it)
        Event = torch.cuda.Event

        def recv_into(buf, src, group, side, g, then_copy_to=None):
            # Post the receive behind everything already queued on `cur`, not earlier: NCCL's receive
            # kernel polls on an SM until the data arrives, and the fused decode kernels put one
            # CTA on every SM -- a receive posted a whole tick ahead costs them a second wave
            # (measured at 2 stages: 946 vs 1045 tok/s).
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                self.dist.irecv(buf, src=src, group=group).wait()
                if then_copy_to is not None:
                    then_copy_to.copy_(buf.view(then_copy_to.shape))
                e = Event()
                e.record(side)
            cur.wait_event(e)

        def token_from_head(g, step, also_into=None):
            """first stage, shared head: receive (hidden, upper-shard winner) of `step`, finish the argmax here
            (`also_into`: the graph input of the next step)."""
            msg, hn, val, idx = inbox[g]
            if P is not None:
                ops.p2p_wait(mine[g][3], bt[g] + step + 1, err=P["err"])
                self.stage.launches += 1
            else:
                recv_into(msg, self.world - 1, self.tok_group, self._rs_tok, g)
            self.stage.head_lower(hn, val, idx, gen[g][step], lg_lo[g], out_tok2=also_into)

        for i in range(steps):
            for g in range(G):
                r = runners[g]
                if multi and self.first and i > 0:
                    if split:
                        token_from_head(g, i - 1, also_into=r.x)
                    else:
                        recv_into(gen[g][i - 1], self.world - 1, self.tok_group, self._rs_tok, g, then_copy_to=r.x)
                elif multi and not self.first:
                    if P is not None:       # wait for step i of group g in this GPU's inbox, then fetch it into the graph's input
                        ops.p2p_wait(mine[g][2], bh[g] + i + 1, src=mine[g][0], dst=r.x, err=P["err"])
                        self.stage.launches += 1
                    else:
                        recv_into(r.x, self.rank - 1, self.group_in, self._rs, g)
                if ev_sent[g] is not None:
                    cur.wait_event(ev_sent[g])
                r.replay()
                if self.last and tokmode:
                    gen[g][i].copy_(r.tok)            # (a one-stage pipeline's graph has already written the next input)
                elif self.last and not split:
                    ops.argmax(r.out, out=gen[g][i])
                    self.stage.launches += 1
                    if self.first:
                        r.x.copy_(gen[g][i].view(r.x.shape))
                if not multi:
                    continue
                if P is not None:
                    # store the result straight into the next stage's inbox (the first stage's, from the last one) and
                    # publish its sequence number: one small kernel behind the graph, on the same stream
                    if not self.last:
                        ops.p2p_send(r.out, peer[g][0], peer[g][2], bh[g] + i + 1)
                    else:
                        ops.p2p_send(r.msg, peer[g][1], peer[g][3], bt[g] + i + 1)
                    self.stage.launches += 1
                    continue
                ef = Event()
                ef.record(cur)
                if not self.last:
                    with torch.cuda.stream(self._ss):
                        self._ss.wait_event(ef)
                        self.dist.isend(r.out, dst=self.rank + 1, group=self.group_out).wait()
                        es = Event()
                        es.record(self._ss)
                    ev_sent[g] = es
                else:
                    with torch.cuda.stream(self._ss_tok):
                        self._ss_tok.wait_event(ef)
                        self.dist.isend(r.msg if split else gen[g][i], dst=0, group=self.tok_group).wait()
                        if split:               # r.msg is the runner's static buffer: the next replay must wait for the send
                            es = Event()
                            es.record(self._ss_tok)
                            ev_sent[g] = es
        if multi and self.first:
            for g in range(G):
                if split:
                    token_from_head(g, steps - 1)
                else:
                    recv_into(gen[g][steps - 1], self.world - 1, self.tok_group, self._rs_tok, g)
        if P is not None:
            for g in range(G):
                P["base_h"][g] = bh[g] + steps
                P["base_t"][g] = bt[g] + steps
        if multi:
            # side streams may still hold sends of this call: later work on `cur` (another call's buffers) follows them
            cur.wait_stream(self._ss)
            cur.wait_stream(self._ss_tok)
        return [list(gen[g].unbind(0)) for g in range(G)] if keep else None

    def head_bounds(self, rank: int):
        """Vocabulary rows [lo, hi) of the lm_head that stage `rank` (one of ring_head_ranks) holds when the stages
        share it."""
        hr = ring_head_ranks(self.world)
        V, n, k = self.vocab, len(hr), hr.index(rank)
        cut = lambda j: V if j >= n else ((V * j // n) & ~7)       # noqa: E731
        return cut(k), cut(k + 1)

    def _decode_stream_ring(self, first_tokens, start_positions, rows_per_group, steps: int):
        """decode_stream with the lm_head shared by the stages (ring_head; groups >= stages, one sequence per group --
        the code handles up to 4 rows per group, the dispatch in _decode_stream_graphs admits what has been measured).
        Per tick a stage enqueues: [wait for the hidden state | on the first stage: merge the winners of the
        previous step into the token] -> its layer graph -> [store the result into the next stage's inbox | on the
        last stage: store the normed hidden state into EVERY stage's inbox] -> its slice of the head for the group
        that is leaving the last stage at that moment (stage r runs W-1-r ticks ahead of the last one), as one
        launch that waits for the hidden state, finds the slice's winner and stores it into the first stage's
        inbox.  Every message is a peer-memory store + sequence flag; nothing here touches the host or NCCL."""
        dev, P, W, G, rank = self.device, self.p2p, self.world, len(rows_per_group), self.rank
        st = self.stage
        nh = len(ring_head_ranks(W))
        self.head_path = "all stages" if nh == W else f"the middle stages ({nh} of {W})"
        runners = []
        for g in range(G):
            r = st.runner(self.cache, len(rows_per_group[g]), decode=True, rows=list(rows_per_group[g]),
                          logits="norm" if self.last else "none")
            pos = np.asarray(start_positions[g], dtype=np.int64).reshape(-1)
            if int(pos.max()) + steps > self.cache.max_seq:
                raise ValueError("decode would run past the KV cache capacity")
            if not np.array_equal(pos, r.host_pos):
                r.set_positions(pos)
            if self.first:
                r.x.copy_(first_tokens[g].view(r.x.shape))
            runners.append(r)
        Bs = [len(r_) for r_ in rows_per_group]
        row0 = np.concatenate([[0], np.cumsum(Bs)]).astype(int)
        gen = [torch.empty((steps, Bs[g]), dtype=torch.int64, device=dev) for g in range(G)] if self.first else None
        slot, d, R = P["slot"], self.d, P["rows"]

        def views(buf, g):
            """group g in the inbox `buf` of some stage: hidden slot + flag, normed-hidden slot + flag"""
            hs = buf[int(row0[g]) * slot: int(row0[g]) * slot + Bs[g] * d * 2]
            hf = buf[P["off_flags"]: P["off_flags"] + R * 4].view(torch.int32)[g: g + 1]
            ns = buf[P["off_hn"] + int(row0[g]) * slot: P["off_hn"] + int(row0[g]) * slot + Bs[g] * d * 2]
            nf = buf[P["off_nflag"]: P["off_nflag"] + R * 4].view(torch.int32)[g: g + 1]
            return hs, hf, ns, nf

        hr = ring_head_ranks(W)                       # the stages holding a slice; record k belongs to stage hr[k]
        NH, mine_k = len(hr), (hr.index(rank) if rank in hr else None)

        def record(buf, g, k):
            """winner record k of group g in the first stage's inbox `buf`: indices [4], values [4], flag"""
            o = P["off_win"] + (g * W + k) * 64
            return (buf[o: o + 32].view(torch.int64), buf[o + 32: o + 48].view(torch.float32),
                    buf[P["off_wflag"]: P["off_wflag"] + R * W * 4].view(torch.int32)[g * W + k: g * W + k + 1])

        mine = [views(P["buf"], g) for g in range(G)]
        nxt = [views(P["peer"], g) for g in range(G)] if not self.last else None
        hn_all = [ops.MultiDst([views(P["peers"][r_], g)[2:4] for r_ in hr]) for g in range(G)] if self.last else None
        hn_in = [mine[g][2].view(torch.float16).view(Bs[g], d) for g in range(G)]
        if self.first:
            my_rec = [P["buf"][P["off_win"] + g * W * 64: P["off_win"] + (g * W + NH) * 64] for g in range(G)]
            my_wf = [P["buf"][P["off_wflag"]: P["off_wflag"] + R * W * 4].view(torch.int32)[g * W: g * W + NH] for g in range(G)]
        err = P["err"]
        if mine_k is not None:
            rec = [record(P["peers"][0], g, mine_k) for g in range(G)]
            lo, hi = self.head_bounds(rank)
            Wslice, scratch = st.head_slice(lo, hi), st.head_scratch()
        bh = [P["base_h"].get(g, 0) for g in range(G)]
        bn = [P["base_n"].get(g, 0) for g in range(G)]
        lag = W - 1 - rank

        def head_slice_for(t):
            """this stage's slice of the head for pipeline tick t = step * G + group (waits for the hidden state)"""
            g, step = t % G, t // G
            seq = bn[g] + step + 1
            ops.lm_head_argmax(hn_in[g], Wslice, scratch, idx_offset=lo, out_idx=rec[g][0][: Bs[g]], out_val=rec[g][1][: Bs[g]],
                               wait=(mine[g][3], seq), notify=(rec[g][2], seq), err=err)
            st.launches += 1

        def token(g, step, also_into):
            ops.merge_winners(my_wf[g], bn[g] + step + 1, my_rec[g], NH, Bs[g], gen[g][step], also_into, err)
            st.launches += 1

        trace = [] if os.environ.get("B2_TICK_TRACE") else None      # developer aid: where a tick's device time goes

        def mark(op):
            if trace is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                trace.append((op, e))

        for op, g, i in ring_head_schedule(rank, W, G, steps):
            r = runners[g]
            mark(op)
            if op == "layers":
                r.replay()
            elif op == "merge":
                token(g, i, r.x if i + 1 < steps else None)
            elif op == "wait_h":
                ops.p2p_wait(mine[g][1], bh[g] + i + 1, src=mine[g][0], dst=r.x, err=err)
                st.launches += 1
            elif op == "send_h":
                ops.p2p_send(r.out, nxt[g][0], nxt[g][1], bh[g] + i + 1)
                st.launches += 1
            elif op == "send_hn":
                ops.p2p_send_multi(r.hn_msg, hn_all[g], bn[g] + i + 1)
                st.launches += 1
            else:
                head_slice_for(i * G + g)
        if trace is not None:
            mark("end")
            torch.cuda.synchronize(dev)
            k0 = len(trace) // 4                      # steady state only
            per = {}
            for (op, e), (_, e_next) in zip(trace[k0:-1], trace[k0 + 1:]):
                per.setdefault(op, []).append(e.elapsed_time(e_next) * 1e3)
            n_ticks = max(1, len(per.get("layers", [])))
            print(f"[rank {rank}] tick {sum(sum(v) for v in per.values()) / n_ticks:.1f} us: " + "; ".join(
                f"{op} mean {np.mean(v):.1f} p90 {np.percentile(v, 90):.1f}" for op, v in per.items()), file=sys.stderr, flush=True)
        for g in range(G):
            P["base_h"][g] = bh[g] + steps
            P["base_n"][g] = bn[g] + steps
        return [list(gen[g].unbind(0)) for g in range(G)] if self.first else None

    def decode_round(self, tokens_per_group, positions_per_group, rows_per_group):
        """One decode step for each group (a decode_stream of a single step)."""
        gen = self.decode_stream(tokens_per_group, positions_per_group, rows_per_group, 1)
        return [x[0] for x in gen] if gen is not None else [None] * len(rows_per_group)

    def _argmax(self, lg, out):
        if lg.is_cuda:
            ops.argmax(lg, out=out)
            self.stage.launches += 1
        else:   # CPU stand-in used only by the gloo schedule tests
            out.copy_(lg.argmax(-1))
