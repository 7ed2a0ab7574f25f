#!/usr/bin/env python
"""bench.py -- tokens/sec (prefill + decode) of the W8A8 stage path on Llama-2-7B, 1..8 B200.

One step = one pass of the hot path over one batch of synthetic input: per GPU-stage pipeline,
N prompts (N = number of GPUs/stages; weak scaling: per-GPU layers x tokens is constant) of
PROMPT tokens are prefilled (chunked) and NEW tokens are greedily decoded for each.
value = (prompt + generated tokens) / second, the reference's single-model TPS definition
(Code/Quantised Models/models_quant_updated.py:95-96 counts len(output[0]) = prompt + new).

    python bench.py --gpus 1 --steps K --warmup W            # our arm
    python bench.py --impl reference ...                     # the reference's CPU int8 path
    torchrun --nproc-per-node N bench.py --gpus N ...        # one rank per GPU, NCCL send/recv

Prints ONE JSON line on rank 0 (see DESIGN.md "Measurement" for every key).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "tokens/sec prefill+decode, Llama-2-7B W8A8"
UNIT = "tokens/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="llama-2-7b")
    ap.add_argument("--prompt", type=int, default=2048)
    ap.add_argument("--new", type=int, default=128)
    ap.add_argument("--chunk", type=int, default=0, help="prefill chunk (0 = auto: whole prompt on one GPU, 1024 on several stages)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--layers", type=int, default=0, help="PROFILING ONLY: truncate the model to this many layers")
    ap.add_argument("--batch", type=int, default=0,
                    help="decode-only workload (BASELINE configs[3]): this many sequences, each with --kv tokens of context "
                         "already in the KV cache, decode --new tokens each; the sequences run as --batch/--micro groups")
    ap.add_argument("--kv", type=int, default=4096, help="with --batch: context length per sequence at the END of the run")
    ap.add_argument("--micro", type=int, default=0, help="with --batch: sequences per micro-batch (0 = batch / stages, at most 8)")
    ap.add_argument("--quant", default="w8a8", choices=["w8a8", "llm_int8"],
                    help="w8a8 = the headline SmoothQuant-style path; llm_int8 = bitsandbytes-style mixed decomposition "
                         "(SURVEY 8(f) rank 2; a side measurement, not the BASELINE metric)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=float(d["hbm_gbs"]), bf16=float(d["bf16_tflops"]), bf16_sus=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    which="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, which="fallback")


# ----------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi in the background; zero interference with the launching thread)
# ----------------------------------------------------------------------------------------------
class Clocks:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                       "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self) -> dict:
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own CPU int8 path (Code/C-DAC Server/try.py:198-206,
# :238-260): HF AutoModelForCausalLM -> quantize_dynamic({nn.Linear}, qint8) -> generate.
# ----------------------------------------------------------------------------------------------
def _cpu_llama_int8(cfg, n_layers: int):
    """HF LlamaForCausalLM at `n_layers` of the model's config -> quantize_dynamic({nn.Linear}, qint8) [try.py:204-205].
    No checkpoint exists offline, so the weights are random: one decoder layer is drawn (N(0, 0.02), HF's default
    init) and copied into the others -- the int8 kernels' time does not depend on the values, and drawing 6.5 G
    normals on the host would take longer than the measurement itself."""
    import torch
    from transformers import LlamaConfig, LlamaForCausalLM
    hc = LlamaConfig(vocab_size=cfg.vocab, hidden_size=cfg.d_model, intermediate_size=cfg.ffn, num_hidden_layers=n_layers,
                     num_attention_heads=cfg.n_heads, num_key_value_heads=cfg.n_kv_heads, head_dim=cfg.head_dim,
                     max_position_embeddings=max(4096, cfg.max_pos), rms_norm_eps=cfg.eps, rope_theta=cfg.rope_theta,
                     tie_word_embeddings=False)
    with torch.device("meta"):
        m = LlamaForCausalLM(hc)
    m = m.to_empty(device="cpu")
    g = torch.Generator().manual_seed(0)
    with torch.no_grad():
        first = m.model.layers[0]
        for name, p_ in first.named_parameters():
            if p_.dim() == 1:
                p_.fill_(1.0)
            else:
                p_.normal_(0.0, 0.02, generator=g)
        sd = first.state_dict()
        for layer in m.model.layers[1:]:
            layer.load_state_dict(sd)
        m.model.embed_tokens.weight.normal_(0.0, 0.02, generator=g)
        m.lm_head.weight.normal_(0.0, 0.02, generator=g)
        m.model.norm.weight.fill_(1.0)
        # buffers (rotary inv_freq) were created on the meta device and are garbage after to_empty: rebuild them
        hd = cfg.head_dim
        inv = 1.0 / (float(cfg.rope_theta) ** (torch.arange(0, hd, 2, dtype=torch.float32) / hd))
        for mod in m.modules():
            if "inv_freq" in dict(mod.named_buffers(recurse=False)):
                mod.inv_freq = inv.clone()
                if hasattr(mod, "original_inv_freq"):
                    mod.original_inv_freq = inv.clone()
                if not isinstance(getattr(mod, "attention_scaling", 1.0), float):
                    mod.attention_scaling = 1.0
        left = [n for n, b in m.named_buffers() if not torch.isfinite(b.float()).all()]
        assert not left, f"uninitialised buffers: {left}"
    m.eval()
    return torch.quantization.quantize_dynamic(m, {torch.nn.Linear}, dtype=torch.qint8, inplace=True)     # try.py:205


def cpu_reference_run(model: str, prompt: int, new: int, reps: int, warmup: int, depths=None):
    """The reference's CPU int8 path on the host cores: HF Llama -> torch.quantization.quantize_dynamic({nn.Linear},
    qint8) [try.py:204-205] -> greedy generate [try.py:247-256 with do_sample=False], on the bench's own shape:
    a `prompt`-token prompt and `new` generated tokens, B = 1, all host threads.

    depths = None: the full model, measured directly (`reps` generations after `warmup`, median).
    depths = (a, b): a bounded sample for the in-line cpu_baseline of the GPU arm -- the same generation at a and b
    of the model's layers, extended linearly in depth to the full model (stated in `sample`, `extrapolated` true)."""
    import warnings
    import torch
    warnings.filterwarnings("ignore")
    import importlib
    cfg = importlib.import_module("llm-for-distributed-egde-devices_b200").get_config(model)
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    ids = torch.randint(0, cfg.vocab, (1, prompt), generator=torch.Generator().manual_seed(1234))
    mask = torch.ones_like(ids)

    def measure(n_layers):
        m = _cpu_llama_int8(cfg, n_layers)

        def gen(n):
            t0 = time.perf_counter()
            m.generate(input_ids=ids, attention_mask=mask, max_new_tokens=n, min_new_tokens=n, do_sample=False, pad_token_id=0)
            return time.perf_counter() - t0
        pre, full = [], []
        with torch.no_grad():
            for i in range(warmup + reps):
                a1 = gen(1)              # prefill + first token
                a2 = gen(new)            # the whole generation
                if i >= warmup:
                    pre.append(a1)
                    full.append(a2)
        del m
        return statistics.median(pre), statistics.median(full), (max(full) - min(full)) / statistics.median(full)

    if depths is None:
        t_pre, t_full, spread = measure(cfg.n_layers)
        how = (f"all {cfg.n_layers} layers measured directly, median of {reps} generation(s) after {warmup} warm-up "
               f"(spread {100 * spread:.1f} %)")
        extrap = False
    else:
        a_, b_ = depths
        pa, fa, _ = measure(a_)
        pb, fb, _ = measure(b_)

        def ext(ta, tb):
            per_layer = (tb - ta) / (b_ - a_)
            if per_layer <= 0.0:                 # timing noise at toy sizes: fall back to proportional scaling
                return tb / b_ * cfg.n_layers
            return max(ta - a_ * per_layer, 0.0) + per_layer * cfg.n_layers
        t_pre, t_full = ext(pa, pb), ext(fa, fb)
        how = (f"bounded sample: the same generation at {a_} and {b_} of {cfg.n_layers} layers ({fa:.2f}s / {fb:.2f}s), "
               f"extended linearly in depth to {cfg.n_layers} layers")
        extrap = True
    t_dec = max(t_full - t_pre, 1e-9)
    out = dict(value=(prompt + new) / t_full, unit=UNIT, cores=threads, kind="reference",
               sample=(f"HF LlamaForCausalLM ({model} shape, random weights) + quantize_dynamic(qint8) [try.py:204-205], greedy "
                       f"generate of {new} tokens after a {prompt}-token prompt, B=1, {threads} threads; {how}: "
                       f"{prompt / t_pre:.1f} prefill tok/s, {(new - 1) / t_dec:.2f} decode tok/s, {t_full:.1f}s per generation"),
               ms_per_step=t_full * 1e3, prefill_tokens_per_s=prompt / t_pre, decode_tokens_per_s=(new - 1) / t_dec,
               extrapolated=extrap)
    if depths is not None:
        out["depths"] = list(depths)
    return out


def grpc_cpu_pipeline_run(steps: int = 3, warmup: int = 3) -> dict:
    """BASELINE.json configs[0], the CPU arm north_star asks to be timed beside the GPU pipeline: GPT-2 small
    fp32 in two layer-range stage servers over localhost gRPC (conventions of Code/gRPC/server.py:13-19 and
    client.py:7-11), 128-token synthetic prompt, 32 greedy tokens, on this box's host cores."""
    from oracle import grpc_stages
    r = grpc_stages.timed_run("gpt2-small", 2, 128, 32, steps=steps, warmup=warmup)
    return {k: r[k] for k in ("value", "unit", "cores", "kind", "prefill_tokens_per_s", "decode_tokens_per_s", "sample")}


def grpc_config_on_gpu(torch, b2, dev, steps: int = 3, warmup: int = 3) -> dict:
    """The same configs[0] workload (GPT-2 small, two layer-range stages, 128-token prompt + 32 greedy tokens)
    through this repo's public API on one GPU, W8A8, host prompt in / host tokens out every generation."""
    pipe = b2.Pipeline.load("gpt2", n_stages=2, devices=[dev, dev], seed=0, fast_init=True)
    ids = torch.randint(0, pipe.cfg.vocab, (128,), generator=torch.Generator().manual_seed(1234), dtype=torch.int64).pin_memory()
    ts = []
    for i in range(warmup + steps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = pipe.generate(ids, max_new_tokens=32, chunk=128).cpu()
        ts.append(time.perf_counter() - t0)
    t = statistics.median(ts[warmup:])
    return {"value": 160 / t, "unit": UNIT, "ms_per_generation": t * 1e3,
            "workload": "gpt2 (124M) W8A8, 2 layer-range stages on one GPU, 128-token prompt + 32 greedy tokens, host ids in / host tokens out"}


def workload_config(cfg_name: str, quant: str, n_prompts: int, P: int, NEW: int) -> dict:
    """The `config` object of the JSON line: the same for both arms (the GPU arm adds how it ran it)."""
    return {"workload": f"{cfg_name} {'W8A8' if quant == 'w8a8' else 'LLM.int8() mixed int8/fp16'} (random-init, seeded), "
                        f"{n_prompts} prompt(s) x ({P} prefill + {NEW} greedy decode), B=1 per prompt"}


def side_config_on_gpu(torch, b2, dev, model: str, prompt: int, new: int, pk, steps: int = 3, warmup: int = 2) -> dict:
    """Another BASELINE.json configuration through the public API on this GPU, reported inside the N = 1 line so that
    it is driver-visible: host prompt ids in, host tokens out, prefill and decode rates split by timing a one-token
    generation beside the full one (median of `steps` after `warmup`)."""
    pipe = b2.Pipeline.load(model, n_stages=1, devices=[dev], seed=0, fast_init=True)
    cfg = pipe.cfg
    ids = torch.randint(0, cfg.vocab, (prompt,), generator=torch.Generator().manual_seed(1234), dtype=torch.int64).pin_memory()
    t1, tn = [], []
    for i in range(warmup + steps):
        for n_new, acc in ((1, t1), (new, tn)):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            pipe.generate(ids, max_new_tokens=n_new, chunk=prompt).cpu()
            if i >= warmup:
                acc.append(time.perf_counter() - t0)
    pre, full = statistics.median(t1), statistics.median(tn)
    dec = max(full - pre, 1e-9)
    dec_tok_s = (new - 1) / dec
    dec_bytes = cfg.linear_params_per_layer() * cfg.n_layers + cfg.vocab * cfg.d_model * 2 \
        + 2 * cfg.n_layers * cfg.n_kv_heads * cfg.head_dim * 2 * (prompt + new // 2)
    out = {"workload": f"{cfg.name} W8A8 (random-init, seeded), 1 GPU, {prompt}-token prefill + {new} greedy tokens, batch 1, "
                       "host ids in / host tokens out", "value": (prompt + new) / full, "unit": UNIT,
           "prefill_tokens_per_s": prompt / pre, "decode_tokens_per_s": dec_tok_s,
           "decode_hbm_frac_whole_step": (dec_bytes / 1e9) * dec_tok_s / pk["hbm"], "ms_per_generation": full * 1e3}
    del pipe
    torch.cuda.empty_cache()
    return out


def run_reference(a):
    """The reference's own CPU implementation of the path, measured on the bench workload itself: the full-depth
    model, the full prompt and all generated tokens.  One generation takes seconds to tens of seconds on the host,
    so the repetitions actually timed are capped (and reported as `steps`); at N > 1 the workload is N independent
    prompts, which the single CPU model would run one after the other: tokens/s is that of one prompt."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    reps, warm = max(1, min(a.steps, 2)), 1      # one generation is ~40 s on 16 host cores; measured spread between generations: 2 %
    import importlib
    cfg = importlib.import_module("llm-for-distributed-egde-devices_b200").get_config(a.model)
    try:
        import psutil
        need = cfg.linear_params_per_layer() * cfg.n_layers * 5.5 + 4e9      # fp32 model + int8 copy + slack
        depths = None if psutil.virtual_memory().available > need else (cfg.n_layers // 4, cfg.n_layers // 2)
    except Exception:
        depths = None
    r = cpu_reference_run(a.model, a.prompt, a.new, reps, warm, depths)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": reps,
            "warmup": warm, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int8", "data": "synthetic",
            "config": workload_config(cfg.name, a.quant, a.gpus, a.prompt, a.new),
            "prefill_tokens_per_s": r["prefill_tokens_per_s"], "decode_tokens_per_s": r["decode_tokens_per_s"],
            "extrapolated": r["extrapolated"], "requested_steps": a.steps, "requested_warmup": a.warmup,
            "note": ("steps/warmup are the generations actually timed; ms_per_step is one prompt's generation -- with N > 1 "
                     "prompts in the workload the CPU model runs them back to back at this same tokens/s"),
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if r.get("depths"):
        line["depths"] = r["depths"]
    if not a.no_cpu_baseline and a.gpus == 1:
        try:
            line["cpu_grpc_pipeline"] = grpc_cpu_pipeline_run()
        except Exception as e:
            line["cpu_grpc_pipeline"] = {"value": None, "sample": f"failed: {type(e).__name__}: {e}"}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------------
def run_ours(a):
    import numpy as np
    import torch
    import importlib
    b2 = importlib.import_module("llm-for-distributed-egde-devices_b200")
    from b200stage import ops
    from b200stage.pipeline import DistPipeline, split_layers

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus:
        if world == 1 and a.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        # NCCL prints its version banner (NCCL_DEBUG >= VERSION) on stdout, ahead of the one JSON line:
        # send NCCL's debug output to stderr instead
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    cfg = b2.get_config(a.model)
    if a.layers:
        import dataclasses
        cfg = dataclasses.replace(cfg, n_layers=a.layers, name=f"{cfg.name}[TRUNCATED to {a.layers} layers: profiling only]")
    P, NEW = a.prompt, a.new
    # measured (value tok/s): N = 2: 512 -> 16 120, 1024 -> 16 267, 2048 -> 16 180; N = 4: 512 -> 29 632, 1024 -> 29 805
    # (bigger chunks run the GEMMs closer to peak, smaller ones fill the pipeline faster)
    # round 2, N = 8 with the CTA-pair GEMM: 512 -> 54 602, 1024 -> 55 526 (the pair kernel needs >= 48 tiles per launch)
    CH = a.chunk or (min(P, 2048) if world == 1 else min(P, 1024))
    max_seq = P + NEW
    n_prompts = world                       # weak scaling: one prompt stream per stage
    g = torch.Generator(device="cpu").manual_seed(1234)
    prompts_host = [torch.randint(0, cfg.vocab, (P,), generator=g, dtype=torch.int64).pin_memory() for _ in range(n_prompts)]
    # SmoothQuant (alpha = 0.5) calibrated on the synthetic prompts: the first 512 tokens of prompt 0, the
    # same at every N, run through the fp32 layers while the stage's weights are drawn
    calib = [prompts_host[0][:512].clone()] if (a.quant == "w8a8" and cfg.gated and not cfg.bias and cfg.norm == "rms") else None
    pipe = DistPipeline.load(cfg, rank, world, dev, seed=0, fast_init=True, max_rows=n_prompts, max_seq=max_seq,
                             quant=a.quant, calibrate_with=calib)
    stage = pipe.stage
    pipe.graph_chunk = CH
    prompts_dev = [p.to(dev) for p in prompts_host]

    def one_step(prompts, record=None):
        """prefill all prompts (pipelined chunks), then NEW-1 decode rounds (first token comes
        from the prefill logits).  Returns generated tokens [n_prompts, NEW] on rank 0 / last."""
        ev = lambda: torch.cuda.Event(enable_timing=True)
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        lgs = pipe.prefill_stream(prompts if pipe.first else [None] * n_prompts, CH)
        toks = []
        for p in range(n_prompts):
            t = None
            if pipe.last:
                t = torch.empty((1,), dtype=torch.int64, device=dev)
                ops.argmax(lgs[p].view(1, -1), out=t)
                stage.launches += 1
            if pipe.first and not pipe.last:
                t = torch.empty((1,), dtype=torch.int64, device=dev)
            toks.append(pipe.return_tokens(t))
        e1.record()
        # NEW-1 decode steps for the n_prompts sequences as one continuous pipeline (no round barrier)
        t_cpu0 = time.perf_counter()
        out = pipe.decode_stream(toks, [[P] for _ in range(n_prompts)], [[p] for p in range(n_prompts)], NEW - 1)
        if os.environ.get("B2_BENCH_DEBUG"):      # host enqueue time of the decode phase vs its device time
            print(f"[rank {rank}] decode enqueue {1e3 * (time.perf_counter() - t_cpu0):.1f} ms for {(NEW - 1) * n_prompts} ticks",
                  file=sys.stderr, flush=True)
        gen = [[toks[p]] + out[p] for p in range(n_prompts)] if out is not None else None
        e2.record()
        if record is not None:
            record.append((e0, e1, e2))
        if gen is not None:       # rank 0 (and the last rank unless the lm_head is shared with the first stage)
            return torch.stack([torch.cat(x) for x in gen])
        return None

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # prompt lengths are part of the workload description, known to every rank
    pipe._bcast_len = lambda T: P
    for _ in range(max(a.warmup, 3)):
        one_step(prompts_dev)
    barrier()
    clocks = Clocks(local) if rank == 0 else None
    l0 = stage.launches
    rec = []
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_start.record()
    for _ in range(a.steps):
        one_step(prompts_dev, rec)
    t_end.record()
    barrier()
    ms_total = t_start.elapsed_time(t_end)
    launches = stage.launches - l0
    pre_ms = sum(e0.elapsed_time(e1) for e0, e1, _ in rec) / len(rec)
    dec_ms = sum(e1.elapsed_time(e2) for _, e1, e2 in rec) / len(rec)
    clk = clocks.stop() if clocks else None
    # ---- e2e: host prompt (pinned) -> H2D -> generate -> D2H of the tokens, every step ---------
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        dp = [p.to(dev, non_blocking=True) for p in prompts_host] if pipe.first else prompts_dev
        out = one_step(dp)
        if pipe.first:
            host_out = out.cpu()          # D2H read of the step's result
        barrier()
    e2e_s = (time.perf_counter() - t0) / a.steps
    # max over ranks
    vals = torch.tensor([ms_total, pre_ms, dec_ms, e2e_s, float(launches)], dtype=torch.float64, device=dev)
    if dist is not None:
        mx = vals.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = vals.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms_total, pre_ms, dec_ms, e2e_s = (float(x) for x in mx[:4])
        launches = int(sm[4])
    ms_step = ms_total / a.steps
    tokens_step = n_prompts * (P + NEW)
    value = tokens_step / (ms_step / 1e3)

    # ---- roofline of the dominant kernel (decode weight stream, gemv_i8_kernel<1>) -------------
    pk = peaks()
    roof, roof_pre = None, None
    if rank == 0:
        roof, roof_pre = roofline_pass(torch, ops, stage, cfg, pk, dev, CH)
    line = None
    if rank == 0:
        n_layers_local = stage.layer_end - stage.layer_begin
        dec_bytes = cfg.linear_params_per_layer() * cfg.n_layers + cfg.vocab * cfg.d_model * 2 \
            + 2 * cfg.n_layers * cfg.n_kv_heads * cfg.head_dim * 2 * (P + NEW // 2)
        dec_tok_s = n_prompts * (NEW - 1) / (dec_ms / 1e3)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8",
            "data": "synthetic",
            "config": dict(workload_config(cfg.name, a.quant, n_prompts, P, NEW),
                           run=f"prefill in chunks of {CH}, {world} pipeline stage(s) x {n_layers_local} layers",
                           parallelism=f"pp{world}", l2="inputs larger than L2 (int8 weights 6.5 GB streamed per decode token)"),
            "prefill_tokens_per_s": n_prompts * P / (pre_ms / 1e3), "decode_tokens_per_s": dec_tok_s,
            "decode_hbm_frac_whole_step": (dec_bytes / 1e9) * dec_tok_s / (pk["hbm"] * world),
            "gpu_launches": launches,
            "e2e": {"value": tokens_step / e2e_s, "unit": UNIT, "h2d_bytes_per_step": n_prompts * P * 8,
                    "d2h_bytes_per_step": n_prompts * NEW * 8},
            "roofline": roof, "roofline_prefill": roof_pre, "clocks": clk, "peaks": pk["which"],
            "decode_handoff": (pipe.p2p["how"] + ": peer-memory store + flag kernels (b2_p2p_send / b2_p2p_send_multi / b2_p2p_wait)") if getattr(pipe, "p2p", None)
            else ("NCCL send/recv" if world > 1 else "none (one stage)"),
            "shared_lm_head": (getattr(pipe, "head_path", "last stage") if world > 1 else False),
            "smoothquant": "alpha 0.5, activation maxima calibrated on the first 512 tokens of prompt 0" if calib else "weight-only statistics",
        }
        line.update(token_check(host_out, cfg.name, a.quant, P, NEW, bool(a.layers)))
        if not a.no_cpu_baseline and world == 1:
            try:
                r = cpu_reference_run(a.model, P, NEW, 1, 1, depths=(max(1, cfg.n_layers // 8), max(2, cfg.n_layers // 4)))
                line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
            except Exception as e:   # the baseline must not take the GPU number down with it
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference",
                                        "sample": f"failed: {type(e).__name__}: {e}"}
            # configs[1]: TinyLlama-1.1B on one GPU, 512-token prefill and batch-1 decode
            if cfg.name == "llama-2-7b":
                try:
                    line["configs1_tinyllama"] = side_config_on_gpu(torch, b2, dev, "tinyllama-1.1b", 512, 128, pk)
                except Exception as e:
                    line["configs1_tinyllama"] = {"error": f"{type(e).__name__}: {e}"}
            # configs[0]: the CPU gRPC stage pipeline, and the same workload through the GPU path, side by side
            try:
                line["cpu_grpc_pipeline"] = grpc_cpu_pipeline_run()
                line["cpu_grpc_pipeline"]["same_workload_on_gpu"] = grpc_config_on_gpu(torch, b2, dev)
            except Exception as e:
                line.setdefault("cpu_grpc_pipeline", {})["error"] = f"{type(e).__name__}: {e}"
        print(json.dumps(line), flush=True)
    pipe.drain()            # raises if a peer-memory hand-off ever timed out
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None and line.get("tokens_match_n1") is False:
        raise SystemExit(f"generated tokens of prompt 0 differ from the stored single-GPU run ({line['token_key']})")


GOLDEN_TOKENS = os.path.join(ROOT, "tests", "golden", "bench_tokens.json")


def token_check(host_out, model: str, quant: str, P: int, NEW: int, truncated: bool) -> dict:
    """crc32 of every prompt's generated ids (the last end-to-end step's D2H result) and, for prompt 0 -- the same
    prompt at every N, since prompts are drawn in order from one seeded generator -- a comparison with the value a
    single-GPU run of the same workload stored in tests/golden/bench_tokens.json: the N-stage NCCL pipeline must
    decode exactly the tokens one GPU decodes."""
    import zlib
    sums = [zlib.crc32(row.numpy().tobytes()) for row in host_out]
    out = {"token_checksums": sums, "tokens_match_n1": None}
    key = f"{model}|{quant}|prompt{P}|new{NEW}|seed1234|calib512"
    if not truncated and os.path.exists(GOLDEN_TOKENS):
        with open(GOLDEN_TOKENS) as f:
            gold = json.load(f)
        if key in gold:
            out["tokens_match_n1"] = bool(gold[key]["crc32_prompt0"] == sums[0])
            out["tokens_first8_prompt0"] = [int(x) for x in host_out[0][:8]]
    out["token_key"] = key
    return out


def roofline_pass(torch, ops, stage, cfg, pk, dev, chunk):
    """CUDA-event timing of the dominant kernels on this stage's real weights, cycling through
    the layers so every launch streams cold weights (6.5 GB >> 126 MB L2)."""
    layers = stage.w["layers"]
    d = cfg.d_model

    def timed(fn, reps):
        """fn() is captured once in a CUDA graph (as the real decode/prefill steps are) so that the
        events bracket device time, not Python launch overhead."""
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        g.replay()
        torch.cuda.synchronize()
        s.record()
        for _ in range(reps):
            g.replay()
        e.record()
        torch.cuda.synchronize()
        return s.elapsed_time(e) / reps

    # decode: the M = 1 fused linear (RMSNorm + quantise + weight stream + SiLU*up epilogue) over the
    # biggest matrix (gate_up), all layers back to back -> every launch streams cold weights
    xh = torch.randn((1, d), dtype=torch.float16, device=dev)
    y32 = torch.empty((1, cfg.ffn), dtype=torch.float32, device=dev)
    amax = torch.zeros((8,), dtype=torch.float32, device=dev)
    xk = ops.X_LN if cfg.norm == "layernorm" else ops.X_RMS

    def dec():
        for lw in layers:
            ops.fused_linear(xh, lw["wgu"], lw["s_gu"], x_kind=xk, gamma=lw["ln2_g"], beta=lw.get("ln2_b"),
                             smooth=lw.get("sm_gu"), eps=cfg.eps, out=y32, out_kind=ops.O_GATED, out_smooth=lw.get("sm_d"),
                             out_amax=amax)
    ms = timed(dec, 5) / len(layers)
    # algorithmic bytes: int8 weights + fp16 x, gamma + fp32 smooth (in and out) + w_scale + fp32 y
    bytes_launch = cfg.up_out * d + 2 * d * 2 + d * 4 + cfg.ffn * 4 + cfg.up_out * 4 + cfg.ffn * 4
    achieved = bytes_launch / (ms * 1e-3) / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum per launch and the isolated duration of this kernel at this shape come
    # from the committed `ncu --set full` capture (profiles/ncu_traffic.json names the summary they were read from)
    ncu = {}
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            ncu = json.load(f).get(f"fused_gemv_kernel<1,RMS,GATED>|d{cfg.d_model}|F{cfg.ffn}", {})
    except Exception:
        pass
    roof = {"bound": "hbm", "kernel": "fused_gemv_kernel<1,RMS,GATED> (norm+quant+gate_up GEMV+SiLU*up, M=1)",
            "achieved": achieved, "peak": pk["hbm"], "unit": "GB/s",
            "frac": achieved / pk["hbm"],
            "traffic": ncu.get("dram_bytes_per_launch"), "traffic_unit": "bytes/launch", "traffic_source": ncu.get("source"),
            "algorithmic_bytes_per_launch": bytes_launch,
            # avg_launch_us: back-to-back launches inside a graph (programmatic dependent launch overlaps each launch's
            # ramp-up with its predecessor's tail); isolated_ncu_us: the same kernel alone, cold, under ncu
            "avg_launch_us": ms * 1e3, "isolated_ncu_us": ncu.get("isolated_ncu_us"),
            "frac_isolated": (bytes_launch / (ncu["isolated_ncu_us"] * 1e-6) / 1e9 / pk["hbm"]) if ncu.get("isolated_ncu_us") else None,
            "peak_source": pk["which"]}
    # prefill: the dominant tensor-core kernel = the gate/up projection of one chunk (CTA-pair tcgen05 GEMM with the
    # act(gate)*up epilogue when the chunk is long enough for it, else the plain W8A8 GEMM)
    M = chunk
    a8 = torch.randint(-127, 128, (M, d), dtype=torch.int8, device=dev)
    as_ = torch.full((M,), 0.01, device=dev)
    fused = stage.fuse_prefill and cfg.gated and M > 128 and cfg.ffn % 32 == 0
    if fused:
        y32 = torch.empty((M, cfg.ffn), dtype=torch.float32, device=dev)
        am = torch.empty((M,), dtype=torch.float32, device=dev)

        def pre():
            for lw in layers:
                ops.gemm_gated_act(a8, lw["wgu"], as_, lw["s_gu"], y32=y32, amax=am, out_smooth=lw.get("sm_d"),
                                   act=0 if cfg.act == "silu" else 1)
        kname = f"gemm2_kernel<GATED> (CTA-pair tcgen05 int8, gate_up + act*up epilogue, M={M})"
    else:
        outp = torch.empty((M, cfg.up_out), dtype=torch.float16, device=dev)

        def pre():
            for lw in layers:
                ops.gemm_w8a8(a8, lw["wgu"], as_, lw["s_gu"], out=outp)
        kname = f"W8A8 GEMM (gate_up, M={M})"
    msp = timed(pre, 3) / len(layers)
    ops_launch = 2.0 * M * cfg.up_out * d
    ach = ops_launch / (msp * 1e-3) / 1e12
    i8 = {}
    try:
        with open(os.path.join(ROOT, "profiles", "int8_peak.json")) as f:
            i8 = json.load(f)
    except Exception:
        pass
    # a kernel timed inside a seconds-long loop: the sustained figures are the right ones of each pair
    peak_i8 = float(i8.get("int8_tops_sustained") or 2.0 * pk["bf16_sus"])
    roof_pre = {"bound": "tensor", "kernel": kname, "achieved": ach, "peak": peak_i8, "unit": "TOP/s", "frac": ach / peak_i8,
                "avg_launch_us": msp * 1e3,
                "peak_source": ("library int8 GEMM (torch._int_mm, 8192^3) measured on this pool's B200, sustained -- profiles/int8_peak.json"
                                if i8 else f"2 x {pk['which']} bf16 sustained (no int8 measurement found)"),
                "other_denominators": {"int8_library_burst": i8.get("int8_tops_burst"), "two_x_bf16_sustained": 2.0 * pk["bf16_sus"],
                                       "two_x_bf16_burst": 2.0 * pk["bf16"], "nominal_dense_int8": 4500.0}}
    return roof, roof_pre


def run_decode_batch(a):
    """BASELINE configs[3]-style workload: batched decode over a long, already filled KV cache, the batch split into
    micro-batches that travel through the layer-range stages as one continuous pipeline (DistPipeline.decode_stream).
    One step = every sequence decodes --new tokens.  The KV cache is pre-filled with N(0,1) fp16 (seed 4321): decode
    attention streams it whatever it holds."""
    import torch
    import importlib
    b2 = importlib.import_module("llm-for-distributed-egde-devices_b200")
    from b200stage.pipeline import DistPipeline
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    cfg = b2.get_config(a.model)
    if a.layers:
        import dataclasses
        cfg = dataclasses.replace(cfg, n_layers=a.layers, name=f"{cfg.name}[TRUNCATED to {a.layers} layers: profiling only]")
    B, S, NEW = a.batch, a.kv, a.new
    micro = a.micro or max(1, min(8, B // max(world, 1)))
    if B % micro:
        raise SystemExit("--batch must be a multiple of --micro")
    G = B // micro
    pipe = DistPipeline.load(cfg, rank, world, dev, seed=0, fast_init=True, max_rows=B, max_seq=S, quant=a.quant)
    stage = pipe.stage
    g = torch.Generator(device=dev).manual_seed(4321 + rank)
    for t in (pipe.cache.k, pipe.cache.v):                 # [layers][rows][KV][Smax][hd]
        for li in range(t.shape[0]):
            for r0 in range(0, B, 8):
                t[li, r0:r0 + 8].normal_(generator=g)
    start = S - NEW                                         # the last decoded token lands at position S - 1
    rows = [list(range(i * micro, (i + 1) * micro)) for i in range(G)]
    first = [torch.randint(0, cfg.vocab, (micro,), device=dev, generator=g) for _ in range(G)]
    pos0 = [[start] * micro for _ in range(G)]

    def one_step():
        return pipe.decode_stream(first, pos0, rows, NEW)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
    for _ in range(max(a.warmup, 3)):
        one_step()
    barrier()
    clocks = Clocks(local) if rank == 0 else None
    l0 = stage.launches
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0.record()
    for _ in range(a.steps):
        out = one_step()
    t1.record()
    barrier()
    ms_total = t0.elapsed_time(t1)
    launches = stage.launches - l0
    clk = clocks.stop() if clocks else None
    # e2e: first tokens from pinned host memory in, generated ids out to the host, every step
    first_host = [f.cpu().pin_memory() for f in first]
    barrier()
    w0 = time.perf_counter()
    for _ in range(a.steps):
        if pipe.first:
            first = [f.to(dev, non_blocking=True) for f in first_host]
        out = one_step()
        if out is not None and pipe.first:
            host_out = torch.stack([torch.stack(o) for o in out]).cpu()
        barrier()
    e2e_s = (time.perf_counter() - w0) / a.steps
    vals = torch.tensor([ms_total, e2e_s, float(launches)], dtype=torch.float64, device=dev)
    if dist is not None:
        mx = vals.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = vals.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms_total, e2e_s = float(mx[0]), float(mx[1])
        launches = int(sm[2])
    if rank == 0:
        pk = peaks()
        ms_step = ms_total / a.steps
        tok_s = B * NEW / (ms_step / 1e3)
        w_bytes = cfg.linear_params_per_layer() * cfg.n_layers + cfg.vocab * cfg.d_model * 2
        kv_tok = 2 * cfg.n_layers * cfg.n_kv_heads * cfg.head_dim * 2 * (min(cfg.window, S) if cfg.window else (S - NEW // 2))
        alg = (w_bytes + kv_tok * B) / B                     # weights read once per step of the whole batch
        act = (w_bytes * G + kv_tok * B) / B                 # as run: every micro-batch re-reads the stage weights
        line = {"metric": f"decode tokens/sec, {cfg.name} W8A8, batch {B}, {S}-token KV cache", "value": tok_s, "unit": UNIT,
                "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
                "config": {"workload": f"{cfg.name} W8A8 (random-init, seeded), decode of {NEW} tokens for each of {B} sequences "
                                       f"ending at a {S}-token KV cache (pre-filled N(0,1) fp16)",
                           "run": f"{G} micro-batches x {micro} sequences through {world} pipeline stage(s) x "
                                  f"{stage.layer_end - stage.layer_begin} layers", "parallelism": f"pp{world}",
                           "l2": f"inputs larger than L2 (KV cache {kv_tok * B / 1e9:.1f} GB + weights {w_bytes / 1e9:.1f} GB per step)"},
                "gpu_launches": launches,
                "e2e": {"value": B * NEW / e2e_s, "unit": UNIT, "h2d_bytes_per_step": B * 8, "d2h_bytes_per_step": B * NEW * 8},
                "roofline": {"bound": "hbm", "kernel": "whole decode step (decode attention over the KV cache + weight streams)",
                             "achieved": alg * tok_s / 1e9 / world, "peak": pk["hbm"], "unit": "GB/s per GPU",
                             "frac": alg * tok_s / 1e9 / world / pk["hbm"], "algorithmic_bytes_per_token": alg,
                             "bytes_per_token_as_run": act, "frac_as_run": act * tok_s / 1e9 / world / pk["hbm"],
                             "traffic": None, "peak_source": pk["which"]},
                "clocks": clk, "peaks": pk["which"]}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.batch:
        run_decode_batch(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
