"""Stage: the quantised decoder-block forward over a contiguous layer range.

Build-defined drop-in boundary modelled on the reference's only conventions,
``load_int8_model(model_path) -> (model, tokenizer)`` and ``model(**inputs).logits`` /
``model.generate(...)`` (Code/Quantised Models/models_quant_updated.py:25-40, :66-67, :84-93):

    stage = Stage.load(model_dir_or_config, layer_begin, layer_end, first=..., last=..., device=...)
    y = stage.forward(x, kv_cache, positions)

``x`` is ``input_ids [T] int64`` on the first stage, else ``hidden [T,d] fp16``; the result is
``hidden [T,d] fp16``, or ``logits fp32`` on the last stage.  ``kv_cache`` is caller-owned and
updated in place.  Every arithmetic step is a call into libb200stage.so (ops.py); torch only owns
device memory, streams and CUDA graphs.  There is no CPU path: loading on a non-CUDA device raises.

Launch-bound shapes (decode steps, fixed-size prefill chunks) are captured once in a CUDA graph
(`Stage.runner`): the kernels read positions / sequence lengths from device memory and the graph
advances them itself, so one graph serves every step.
"""
from __future__ import annotations

import gc
import os

import numpy as np
import torch

from . import ops, weights
from .config import ModelConfig, get_config


# LLM.int8() mode: batches longer than this use the tensor-core side product (ops.outlier_side_dense)
LLM_INT8_DENSE_SIDE_ABOVE = 64

# Measured on B200 (tools/envsweep.sh): see DESIGN.md "L2 hint plan".
DEFAULT_PF_PLAN = "qkv=wo:0:24,wgu:0:24;gu=wd:0:24;down=nqkv:0:24"


def _parse_pf_plan(text: str):
    """'launch=weight:offMB:MB,weight:offMB:MB;launch=...' -> {launch: [(weight, offMB, MB)]}."""
    plan = {}
    for part in filter(None, (x.strip() for x in text.split(";"))):
        launch, _, items = part.partition("=")
        plan[launch.strip()] = []
        for item in filter(None, (x.strip() for x in items.split(","))):
            name, off, mb = item.split(":")
            plan[launch.strip()].append((name, float(off), float(mb)))
    return plan


class KVCache:
    """Stage-local KV cache: k, v fp16 [n_layers][rows][KV][Smax][hd] plus per-row lengths.

    Lengths live on the host (python ints drive launch geometry without device syncs); the
    kernels get them through small device vectors owned by the Stage scratch / graph runners.
    """

    def __init__(self, cfg: ModelConfig, n_layers: int, rows: int, max_seq: int, device):
        self.cfg, self.n_layers, self.rows, self.max_seq = cfg, n_layers, rows, max_seq
        shape = (n_layers, rows, cfg.n_kv_heads, max_seq, cfg.head_dim)
        self.k = torch.zeros(shape, dtype=torch.float16, device=device)
        self.v = torch.zeros(shape, dtype=torch.float16, device=device)
        self.lens = np.zeros((rows,), dtype=np.int64)

    def reset(self):
        self.lens[:] = 0

    def bytes(self) -> int:
        return 2 * self.k.numel() * 2


class _Scratch:
    """Activation buffers for up to `cap` tokens, reused by every layer of the stage."""

    def __init__(self, cfg: ModelConfig, cap: int, device, last: bool):
        d, F = cfg.d_model, cfg.ffn
        hk = cfg.n_heads * cfg.head_dim
        kmax = max(d, hk, F)
        i8, f16, f32 = torch.int8, torch.float16, torch.float32
        self.cap = cap
        self.h = [torch.empty((cap, d), dtype=f16, device=device) for _ in range(2)]
        self.q8 = torch.empty((cap * kmax,), dtype=i8, device=device)
        self.scale = torch.empty((cap,), dtype=f32, device=device)
        self.qkv = torch.empty((cap, cfg.qkv_out), dtype=f16, device=device)
        self.attn = torch.empty((cap, hk), dtype=f16, device=device)
        self.gu = torch.empty((cap, cfg.up_out), dtype=f16, device=device)
        self.pos = torch.empty((cap,), dtype=torch.int32, device=device)
        self.seq = torch.empty((cap,), dtype=torch.int32, device=device)
        self.lens = torch.empty((cap,), dtype=torch.int32, device=device)
        self.ids = torch.empty((cap,), dtype=torch.int64, device=device)
        self.hn = torch.empty((cap, d), dtype=f16, device=device) if last else None
        # fused decode path (T <= 8): act(gate)*up in fp32 and its running row absmax
        self.y32 = torch.empty((8, F), dtype=f32, device=device)
        self.amax = torch.zeros((8,), dtype=f32, device=device)
        # prefill: row absmax of act(gate)*up, raised atomically by the gate/up GEMM epilogue (the fp32 product itself
        # lives in `gu`'s storage: [cap,F] fp32 is as large as [cap,2F] fp16)
        self.amax_p = torch.zeros((cap,), dtype=f32, device=device)
        # grid-barrier counters of the persistent layer-tail kernel: 4 x uint32 per layer, zeroed per forward
        self.ctr = torch.zeros((256 * 4,), dtype=torch.int32, device=device)

    def q8view(self, T, K):
        return self.q8[: T * K].view(T, K)

    def llm_int8(self, cfg, device):
        """Extra buffers of the LLM.int8() path (made on first use): fp16 norm output, fp16 act*up, and the
        outlier scratch (mask / column list / count) per distinct input width."""
        if getattr(self, "_int8x", None) is None:
            f16 = torch.float16
            self._int8x = dict(
                y=torch.empty((self.cap, cfg.d_model), dtype=f16, device=device),
                hm=torch.empty((self.cap, cfg.ffn), dtype=f16, device=device),
                scr_d=ops.OutlierScratch(cfg.d_model, device),
                scr_o=ops.OutlierScratch(cfg.n_heads * cfg.head_dim, device),
                scr_f=ops.OutlierScratch(cfg.ffn, device), dense=None)
            if self.cap > LLM_INT8_DENSE_SIDE_ABOVE:
                hk = cfg.n_heads * cfg.head_dim
                shapes = [(cfg.qkv_out, cfg.d_model), (cfg.d_model, hk), (cfg.up_out, cfg.d_model), (cfg.d_model, cfg.ffn)]
                self._int8x["dense"] = ops.DenseSideScratch(self.cap, max(k for _, k in shapes), max(n for n, _ in shapes),
                                                            max(n * k for n, k in shapes), device)
        return self._int8x


class Stage:
    def __init__(self, cfg: ModelConfig, sw: dict, layer_begin: int, layer_end: int, device):
        if torch.device(device).type != "cuda":
            raise RuntimeError("Stage runs only on a CUDA device: the hot path has no CPU fallback")
        ops._abi.lib()   # fail loudly now if the extension is missing
        with torch.cuda.device(torch.device(device)):
            ops._abi.init_device()
        self.cfg, self.w = cfg, sw
        self.first, self.last = bool(sw["first"]), bool(sw["last"])
        self.layer_begin, self.layer_end = layer_begin, layer_end
        self.device = torch.device(device)
        cos, sin = weights.rope_table(cfg.head_dim, cfg.max_pos, cfg.rope_theta, cfg.rope_scaling, cfg.rotary_dim)
        self.cos, self.sin = cos.to(self.device), sin.to(self.device)
        self.scale = cfg.attn_scale        # 1/sqrt of the MODEL's head width (heads padded for the kernels keep it)
        self._scratch = None
        self._ws = None
        self._runners = {}
        self._head_scr = None
        self._head_slices = {}
        self._origin = None
        self.launches = 0          # kernels of ours launched (bench.py's gpu_launches claim)
        self.fuse_decode = True    # decode steps with <= 8 tokens use b2_fused_linear
        self.l2_prefetch = True    # fused decode kernels pull the head of the next weights into L2
        # which launch of a decoder layer hints which later weights into L2: launch -> [(weight, MB offset, MB)]
        # ("nqkv" = the next layer's wqkv, or the LM head after the last layer)
        self.pf_plan = _parse_pf_plan(os.environ.get("B2_PF_PLAN", DEFAULT_PF_PLAN))
        # o_proj -> gate_up -> down -> next qkv as one persistent launch with grid barriers: exact, but measured
        # SLOWER than four PDL-chained launches on B200 (38.0 vs 33.8 us per layer tail), so off by default
        # (also slower on the small models: TinyLlama 1271 vs 1584 tok/s, Llama-3.2-1B 1506 vs 1857)
        self.fuse_tail = False
        # prefill chunks of more than 128 tokens: RoPE + KV append and act(gate)*up run as epilogues of the
        # CTA-pair GEMM (b2_gemm_qkv_rope, b2_gemm_gated_act + b2_quant_f32); bit-identical to the unfused calls
        self.fuse_prefill = True
        # "w8a8" (SmoothQuant-style, the north-star path) or "llm_int8" (bitsandbytes' mixed decomposition:
        # outlier activation columns in fp16, the rest int8 -- what the reference's load_in_8bit=True computes)
        self.quant = sw.get("quant", "w8a8")

    # ------------------------------------------------------------------------------------------
    @classmethod
    def load(cls, model_dir_or_config, layer_begin: int, layer_end: int, *, first: bool, last: bool,
             device="cuda", quant: str = "w8a8", seed: int = 0, fast_init: bool = False, act_stats=None,
             calibrate_with=None) -> "Stage":
        """Load the W8A8 shard for layers [layer_begin, layer_end).

        model_dir_or_config: a Hugging Face model DIRECTORY (the reference's `from_pretrained(model_path)`
        convention, models_quant_updated.py:34-38): only this range's tensors are read from its *.safetensors /
        pytorch_model*.bin shards and quantised on load (loader.py); a directory without weight files is an error.
        A preset name, ModelConfig or config dict instead builds the seeded synthetic model of that shape
        (`seed`; `fast_init` generates on the device for the benchmarks).
        act_stats: per-layer activation maxima for SmoothQuant (loader.calibrate); calibrate_with: prompts (list of
        id arrays) to gather them from here, over the fp32 model up to layer_end.  Without either, a checkpoint
        directory's saved statistics (loader.save_stats) are used if present, else weight-only statistics."""
        from . import loader
        if quant not in ("w8a8", "llm_int8"):
            raise ValueError("quant must be 'w8a8' or 'llm_int8'")
        cfg = get_config(model_dir_or_config)
        if not (0 <= layer_begin < layer_end <= cfg.n_layers):
            raise ValueError(f"bad layer range [{layer_begin},{layer_end}) for {cfg.n_layers} layers")
        if torch.device(device).type != "cuda":
            raise RuntimeError("Stage runs only on a CUDA device: the hot path has no CPU fallback")
        from_dir = loader.is_checkpoint_dir(model_dir_or_config)
        calib_ids = None
        if calibrate_with is not None and act_stats is None:
            if fast_init and not from_dir:
                calib_ids = calibrate_with[0]        # the fast builder calibrates while it draws the layers
            else:
                act_stats = loader.calibrate(model_dir_or_config, calibrate_with, seed, device, layers=layer_end)
        if from_dir:
            sw = loader.build_stage_weights_from_checkpoint(str(model_dir_or_config), layer_begin, layer_end, first, last,
                                                            device, act_stats, quant)
        elif fast_init:
            sw = weights.build_stage_weights_fast(cfg, layer_begin, layer_end, first, last, seed, device, quant=quant,
                                                  calib_ids=calib_ids)
        else:
            sw = weights.build_stage_weights(cfg, layer_begin, layer_end, first, last, seed, device, act_stats, quant=quant)
        st = cls(cfg, sw, layer_begin, layer_end, device)
        st._origin = dict(model=model_dir_or_config, seed=seed, fast=bool(fast_init and not from_dir), from_dir=from_dir)
        return st

    def new_kv_cache(self, rows: int, max_seq: int) -> KVCache:
        return KVCache(self.cfg, self.layer_end - self.layer_begin, rows, max_seq, self.device)

    def weight_bytes(self) -> int:
        n = 0
        for lw in self.w["layers"]:
            n += sum(v.numel() * v.element_size() for v in lw.values())
        return n

    def _buffers(self, T: int) -> _Scratch:
        if self._scratch is None or self._scratch.cap < T:
            self._scratch = _Scratch(self.cfg, max(T, 8), self.device, self.last)
        return self._scratch

    def _workspace(self, nbytes: int):
        if nbytes == 0:
            return None
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = torch.empty((nbytes,), dtype=torch.uint8, device=self.device)
        return self._ws

    # ------------------------------------------------------------------------------------------
    # device-only core: every input already lives on the device, nothing here touches the host
    # except launch geometry that depends on T alone -> capturable in a CUDA graph.
    # ------------------------------------------------------------------------------------------
    def _core(self, buf: _Scratch, x, T: int, kv_cache: KVCache, decode: bool, dpos, dseq, dlens, row0: int,
              max_len: int, ws, logits: str, out, trace=None, hn_out=None, head_out=None):
        cfg = self.cfg
        d, H, KV, hd, F = cfg.d_model, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.ffn
        n = 0
        if self.first:
            h = ops.embed(x.view(-1), self.w["embed"], dpos if "pos_embed" in self.w else None,
                          self.w.get("pos_embed"), out=buf.h[0][:T])
            n += 1
        else:
            h = buf.h[0][:T]
            if x.data_ptr() != h.data_ptr():
                h.copy_(x.view(T, d))
        h_alt = buf.h[1][:T]
        # parallel residual (Phi, GPT-NeoX): the MLP branch reads the layer input h, not h + attention; the down
        # projection still adds h_alt = h + attention, so the layer returns h + attention + mlp
        par = cfg.parallel
        qkv, attn, gu = buf.qkv[:T], buf.attn[:T], buf.gu[:T]
        sc = buf.scale[:T]
        act = (ops.ACT_SILU_MUL if cfg.act == "silu" else ops.ACT_GELU_MUL) if cfg.gated else \
              {"gelu_new": ops.ACT_GELU, "silu": ops.ACT_SILU, "gelu": ops.ACT_GELU_ERF}[cfg.act]
        fused = self.fuse_decode and decode and T <= 8 and trace is None
        xk_norm = ops.X_LN if cfg.norm == "layernorm" else ops.X_RMS
        layers = self.w["layers"]
        for li, lw in enumerate(layers):
            kc, vc = kv_cache.k[li], kv_cache.v[li]
            if self.quant == "llm_int8":
                n += self._layer_llm_int8(buf, lw, h, h_alt, T, kc, vc, decode, dpos, dseq, dlens, row0, max_len, ws, act,
                                          trace)
                continue
            if fused and cfg.gated and self.fuse_tail:
                # 2 launches per layer: attention, then ONE persistent kernel for
                # o_proj -> gate_up -> down -> next layer's qkv (grid barriers between the phases)
                pf = self.l2_prefetch
                rope = lambda k_, v_: dict(cos=self.cos, sin=self.sin, positions=dpos, seq_ids=dseq, seq0=row0,
                                           kcache=k_, vcache=v_, H=H, KV=KV, hd=hd, use_rope=cfg.rope)
                if li == 0:
                    buf.ctr.zero_()
                    ops.fused_qkv(h, lw["wqkv"], lw["s_qkv"], qkv, self.cos, self.sin, dpos, kc, vc, H, KV, hd,
                                  x_kind=xk_norm, gamma=lw["ln1_g"], beta=lw.get("ln1_b"), smooth=lw["sm_qkv"],
                                  eps=cfg.eps, bias=lw.get("b_qkv"), seq_ids=dseq, seq0=row0, use_rope=cfg.rope,
                                  prefetch=(lw["wo"], lw["wgu"]) if pf else None)
                    n += 1
                ops.decode_attn(qkv, kc, vc, dlens, H, KV, hd, max_len, self.scale, cfg.window, rows=dseq, out=attn,
                                workspace=ws)
                nl = layers[li + 1] if li + 1 < len(layers) else None
                d_o = ops.fused_desc(attn, lw["wo"], lw["s_o"], x_kind=ops.X_PLAIN, smooth=lw["sm_o"], bias=lw.get("b_o"),
                                     residual=h, out=h_alt, zero_amax=buf.amax)
                d_g = ops.fused_desc(h if par else h_alt, lw["wgu"], lw["s_gu"], x_kind=xk_norm, gamma=lw["ln2_g"], beta=lw.get("ln2_b"),
                                     smooth=lw["sm_gu"], eps=cfg.eps, bias=lw.get("b_gu"), out=buf.y32[:T],
                                     out_kind=ops.O_GATED, act=0 if cfg.act == "silu" else 1, out_smooth=lw["sm_d"],
                                     out_amax=buf.amax, prefetch=(lw["wd"],) if pf else None)
                nxt_w = nl["wqkv"] if nl is not None else self.w.get("lm_head")
                d_d = ops.fused_desc(buf.y32[:T], lw["wd"], lw["s_d"], x_kind=ops.X_Y32, x_amax=buf.amax,
                                     bias=lw.get("b_d"), residual=h_alt, out=h, prefetch=(nxt_w,) if pf else None)
                d_q = None
                if nl is not None:
                    d_q = ops.fused_desc(h, nl["wqkv"], nl["s_qkv"], x_kind=xk_norm, gamma=nl["ln1_g"], beta=nl.get("ln1_b"),
                                         smooth=nl["sm_qkv"], eps=cfg.eps, bias=nl.get("b_qkv"), out=qkv, out_kind=ops.O_QKV,
                                         rope=rope(kv_cache.k[li + 1], kv_cache.v[li + 1]),
                                         prefetch=(nl["wo"], nl["wgu"]) if pf else None)
                ops.fused_layer_tail(d_o, d_g, d_d, d_q, T, buf.ctr[li * 4: li * 4 + 4])
                n += 2 if ws is None else 3
                continue
            if fused:
                # 6-7 launches per layer: every linear redoes its (tiny) activation prologue in-kernel
                ops.fused_qkv(h, lw["wqkv"], lw["s_qkv"], qkv, self.cos, self.sin, dpos, kc, vc, H, KV, hd,
                              x_kind=xk_norm, gamma=lw["ln1_g"], beta=lw.get("ln1_b"), smooth=lw["sm_qkv"],
                              eps=cfg.eps, bias=lw.get("b_qkv"), seq_ids=dseq, seq0=row0, use_rope=cfg.rope,
                              prefetch=self._hints("qkv", lw, li))
                ops.decode_attn(qkv, kc, vc, dlens, H, KV, hd, max_len, self.scale, cfg.window, rows=dseq, out=attn,
                                workspace=ws, prefetch=self._hints("attn", lw, li))
                ops.fused_linear(attn, lw["wo"], lw["s_o"], x_kind=ops.X_PLAIN, smooth=lw["sm_o"], bias=lw.get("b_o"),
                                 residual=h, out=h_alt, zero_amax=buf.amax, prefetch=self._hints("o", lw, li))
                n += 3 if ws is None else 4
                if cfg.gated:
                    ops.fused_linear(h if par else h_alt, lw["wgu"], lw["s_gu"], x_kind=xk_norm, gamma=lw["ln2_g"],
                                     beta=lw.get("ln2_b"), smooth=lw["sm_gu"], eps=cfg.eps, bias=lw.get("b_gu"),
                                     out=buf.y32[:T], out_kind=ops.O_GATED, act=0 if cfg.act == "silu" else 1,
                                     out_smooth=lw["sm_d"], out_amax=buf.amax,
                                     prefetch=self._hints("gu", lw, li))
                    # (also tried on B200 and rejected: prefetching the next layer's K/V block here, 442 vs
                    #  480 tok/s, and an L2 evict-first policy on the weight stream, 459 vs 480 tok/s)
                    ops.fused_linear(buf.y32[:T], lw["wd"], lw["s_d"], x_kind=ops.X_Y32, x_amax=buf.amax,
                                     bias=lw.get("b_d"), residual=h_alt, out=h,
                                     prefetch=self._hints("down", lw, li))
                    n += 2
                else:
                    ops.fused_linear(h if par else h_alt, lw["wgu"], lw["s_gu"], x_kind=xk_norm, gamma=lw["ln2_g"],
                                     beta=lw.get("ln2_b"), smooth=lw["sm_gu"], eps=cfg.eps, bias=lw.get("b_gu"), out=gu)
                    q8 = buf.q8view(T, F)
                    ops.act_mul_quant(gu, lw["sm_d"], act, out_q=q8, out_scale=sc)
                    ops.gemm_w8a8(q8, lw["wd"], sc, lw["s_d"], lw.get("b_d"), h_alt, out=h)
                    n += 3
                continue
            pair = self.fuse_prefill and not decode and trace is None and T > 128
            q8 = buf.q8view(T, d)
            ops.norm_quant(h, lw["ln1_g"], lw.get("ln1_b"), lw["sm_qkv"], cfg.eps, out_q=q8, out_scale=sc)
            tr = None
            if pair and hd in (64, 128):
                ops.gemm_qkv_rope(q8, lw["wqkv"], sc, lw["s_qkv"], qkv, self.cos, self.sin, dpos, kc, vc, H, KV, hd,
                                  bias=lw.get("b_qkv"), seq_ids=dseq, seq0=row0, use_rope=cfg.rope)
                n -= 1
            else:
                ops.gemm_w8a8(q8, lw["wqkv"], sc, lw["s_qkv"], lw.get("b_qkv"), None, out=qkv)
                if trace is not None:
                    tr = dict(q1=q8.clone(), s1=sc.clone(), qkv=qkv.clone())
                    trace.append(tr)
                ops.rope_kv_append(qkv, self.cos, self.sin, dpos, kc, vc, H, KV, hd, seq_ids=dseq, seq0=row0,
                                   use_rope=cfg.rope)
            if decode:
                ops.decode_attn(qkv, kc, vc, dlens, H, KV, hd, max_len, self.scale, cfg.window, rows=dseq, out=attn,
                                workspace=ws)
                n += 1 if ws is None else 2
            else:
                ops.prefill_attn(qkv, kc, vc, 0, row0, H, KV, hd, self.scale, cfg.window, out=attn, positions=dpos)
                n += 1
            q8 = buf.q8view(T, H * hd)
            ops.smooth_quant(attn, lw["sm_o"], out_q=q8, out_scale=sc)
            ops.gemm_w8a8(q8, lw["wo"], sc, lw["s_o"], lw.get("b_o"), h, out=h_alt)
            q8 = buf.q8view(T, d)
            ops.norm_quant(h if par else h_alt, lw["ln2_g"], lw.get("ln2_b"), lw["sm_gu"], cfg.eps, out_q=q8, out_scale=sc)
            if pair and cfg.gated and F % 32 == 0:
                y32 = gu.view(torch.float32)                     # [T,F] fp32 in the storage of the fp16 [T,2F] buffer
                am = buf.amax_p[:T]
                ops.gemm_gated_act(q8, lw["wgu"], sc, lw["s_gu"], y32=y32, amax=am, bias=lw.get("b_gu"),
                                   out_smooth=lw["sm_d"], act=0 if cfg.act == "silu" else 1)
                q8 = buf.q8view(T, F)
                ops.quant_f32(y32, am, out_q=q8, out_scale=sc)
            else:
                ops.gemm_w8a8(q8, lw["wgu"], sc, lw["s_gu"], lw.get("b_gu"), None, out=gu)
                q8 = buf.q8view(T, F)
                ops.act_mul_quant(gu, lw["sm_d"], act, out_q=q8, out_scale=sc)
            ops.gemm_w8a8(q8, lw["wd"], sc, lw["s_d"], lw.get("b_d"), h_alt, out=h)
            n += 9
            if tr is not None:
                tr.update(attn=attn.clone(), h1=h_alt.clone(), gu=gu.clone(), q4=q8.clone(), s4=sc.clone(), h2=h.clone())
        if not self.last or logits == "none":
            if out is not None and out.data_ptr() != h.data_ptr():
                out.copy_(h)
            return (h if out is None else out), n
        hsel = h if (logits in ("all", "split") or decode) else h[T - 1:T]
        hn = ops.norm(hsel, self.w["lnf_g"], self.w.get("lnf_b"), cfg.eps,
                      out=hn_out if hn_out is not None else buf.hn[: hsel.shape[0]])
        if logits == "norm":            # the head itself runs elsewhere (shared by all stages of the pipeline)
            return hn, n + 1
        if head_out is not None:
            # greedy decode: the token (or this vocabulary shard's winner) straight from the head's dot products, no
            # logits round trip and no second launch.  head_out = (values or None, indices, second copy or None)
            val, idx, idx2 = head_out
            lo = self.head_split() if logits == "split" else 0
            ops.lm_head_argmax(hn, self.w["lm_head"][lo:], self.head_scratch(), idx_offset=lo, out_idx=idx, out_idx2=idx2,
                               out_val=val)
            return None, n + 2
        if logits == "split":
            # the upper vocabulary shard only: the first stage computes the lower one from the normed hidden state
            lg = ops.lm_head(hn, self.w["lm_head"][self.head_split():], out=out)
        else:
            lg = ops.lm_head(hn, self.w["lm_head"], out=out)
            if "lm_head_b" in self.w:           # Phi: lm_head carries a bias
                ops.bias_add_f32(lg, self.w["lm_head_b"])
                n += 1
        return lg, n + 2

    def head_split(self) -> int:
        """First vocabulary row of the upper lm_head shard when the head is shared between the first and the last
        pipeline stage (DistPipeline.split_head): rows [0, split) on the first stage, [split, V) on the last."""
        return (self.cfg.vocab // 2 + 7) & ~7

    def head_scratch(self):
        """Work area of b2_lm_head_argmax for this stage's stream of launches (zeroed once; every launch re-zeroes it)."""
        if self._head_scr is None:
            self._head_scr = ops.head_argmax_scratch(self.device)
        return self._head_scr

    def head_slice(self, lo: int, hi: int):
        """Rows [lo, hi) of the model's lm_head on this stage's GPU -- every stage of a pipeline can hold one (the
        head shared by all stages, DistPipeline.ring_head).  The last stage (and the first one of a model with tied
        embeddings) already holds the matrix; the others read / draw their rows once."""
        key = (int(lo), int(hi))
        if key not in self._head_slices:
            if self.last:
                t = self.w["lm_head"][lo:hi]
            elif self.first and self.cfg.tie_embeddings:
                t = self.w["embed"][lo:hi]
            else:
                o = self._origin
                if o is None:
                    raise RuntimeError("this stage was not built by Stage.load: it cannot fetch a slice of the lm_head")
                if o["from_dir"]:
                    from . import loader
                    t = loader.load_head_rows(str(o["model"]), lo, hi, self.device)
                else:
                    t = weights.head_rows_fp16(self.cfg, o["seed"], lo, hi, self.device, o["fast"])
            self._head_slices[key] = t.contiguous()
        return self._head_slices[key]

    def can_fuse_head(self, rows: int) -> bool:
        """lm_head + greedy argmax as one launch: decode steps of up to 4 rows, heads without a bias."""
        return rows <= 4 and not self.cfg.head_bias

    def head_lower(self, hn, other_val, other_idx, out_tok, lg_buf, out_tok2=None):
        """First stage: logits of the lower vocabulary shard from the last stage's normed hidden state hn [B,d] (the
        embedding table doubles as lm_head when the model ties them), then the greedy token over the whole
        vocabulary by merging with the upper shard's winner (other_val, other_idx) -> out_tok [B] int64 (and a
        second copy in out_tok2, e.g. the next step's input)."""
        if "embed" not in self.w or not self.cfg.tie_embeddings:
            raise RuntimeError("head_lower needs the first stage of a model with tied embeddings")
        if self.can_fuse_head(hn.shape[0]):
            ops.lm_head_argmax(hn, self.w["embed"][: self.head_split()], self.head_scratch(), other=(other_val, other_idx),
                               out_idx=out_tok, out_idx2=out_tok2)
            self.launches += 1
            return out_tok
        lg = ops.lm_head(hn, self.w["embed"][: self.head_split()], out=lg_buf)
        ops.argmax_shard(lg, 0, out_idx=out_tok, other=(other_val, other_idx))
        if out_tok2 is not None:
            out_tok2.copy_(out_tok.view(out_tok2.shape))
        self.launches += 2
        return out_tok

    def _layer_llm_int8(self, buf, lw, h, h_alt, T, kc, vc, decode, dpos, dseq, dlens, row0, max_len, ws, act, trace):
        """One decoder layer with every linear computed as bitsandbytes' Linear8bitLt does (ops.llm_int8_linear;
        reference: BitsAndBytesConfig(load_in_8bit=True), models_quant_updated.py:30-38).  h -> h (h_alt = the
        hidden state after attention).  Returns the number of kernels launched."""
        cfg = self.cfg
        d, H, KV, hd, F = cfg.d_model, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.ffn
        x = buf.llm_int8(cfg, self.device)
        y, hm, sc = x["y"][:T], x["hm"][:T], buf.scale[:T]
        # long batches: most columns of a wide activation hold some value >= 6 -> side product on the tensor cores
        dense = x["dense"] if T > LLM_INT8_DENSE_SIDE_ABOVE else None
        fd = decode and T <= 8 and trace is None and self.fuse_decode
        qkv, attn, gu = buf.qkv[:T], buf.attn[:T], buf.gu[:T]
        ops.norm(h, lw["ln1_g"], lw.get("ln1_b"), cfg.eps, out=y)
        ops.llm_int8_linear(y, lw["wqkv"], lw["s_qkv"], x["scr_d"], bias=lw.get("b_qkv"), out=qkv,
                            q_buf=buf.q8view(T, d), scale_buf=sc, dense=dense, fused_decode=fd)
        qkv_pre = qkv.clone() if trace is not None else None          # q and k are rotated in place below
        ops.rope_kv_append(qkv, self.cos, self.sin, dpos, kc, vc, H, KV, hd, seq_ids=dseq, seq0=row0, use_rope=cfg.rope)
        if decode:
            ops.decode_attn(qkv, kc, vc, dlens, H, KV, hd, max_len, self.scale, cfg.window, rows=dseq, out=attn, workspace=ws)
        else:
            ops.prefill_attn(qkv, kc, vc, 0, row0, H, KV, hd, self.scale, cfg.window, out=attn, positions=dpos)
        ops.llm_int8_linear(attn, lw["wo"], lw["s_o"], x["scr_o"], bias=lw.get("b_o"), residual=h, out=h_alt,
                            q_buf=buf.q8view(T, H * hd), scale_buf=sc, dense=dense, fused_decode=fd)
        ops.norm(h if cfg.parallel else h_alt, lw["ln2_g"], lw.get("ln2_b"), cfg.eps, out=y)
        ops.llm_int8_linear(y, lw["wgu"], lw["s_gu"], x["scr_d"], bias=lw.get("b_gu"), out=gu,
                            q_buf=buf.q8view(T, d), scale_buf=sc, dense=dense, fused_decode=fd)
        ops.act_mul(gu, F, act, out=hm)
        ops.llm_int8_linear(hm, lw["wd"], lw["s_d"], x["scr_f"], bias=lw.get("b_d"), residual=h_alt, out=h,
                            q_buf=buf.q8view(T, F), scale_buf=sc, dense=dense, fused_decode=fd)
        if trace is not None:
            trace.append(dict(qkv=qkv_pre, attn=attn.clone(), h1=h_alt.clone(), gu=gu.clone(), hm=hm.clone(), h2=h.clone(),
                              n_out=[int(x[k].count.item()) for k in ("scr_d", "scr_o", "scr_f")]))
        # per linear: scan (1 launch for decode batches, else 2 + memset) + quantise + GEMM (one fused launch in
        # decode) + side; plus 2 norms, rope, attention (+ merge), act
        return 4 * (3 if fd else 5) + 5 + (1 if (decode and ws is not None) else 0)

    def _hints(self, launch: str, lw, li: int):
        """The (up to two) L2 hints the `launch` kernel of layer li issues, per self.pf_plan."""
        if not self.l2_prefetch:
            return None
        out = []
        layers = self.w["layers"]
        for name, off_mb, mb in self.pf_plan.get(launch, [])[:2]:
            if name == "nqkv":
                t = layers[li + 1]["wqkv"] if li + 1 < len(layers) else self.w.get("lm_head")
            else:
                t = lw.get(name)
            if t is not None and mb > 0:
                out.append((t, int(off_mb * (1 << 20)), int(mb * (1 << 20))))
        return tuple(out) if out else None

    # ------------------------------------------------------------------------------------------
    @staticmethod
    def _classify(positions, seq_ids):
        pos = np.asarray(positions, dtype=np.int64).reshape(-1)
        T = int(pos.shape[0])
        if T == 0:
            raise ValueError("empty input")
        rows = np.zeros((T,), dtype=np.int64) if seq_ids is None else np.asarray(seq_ids, dtype=np.int64).reshape(-1)
        single_seq = bool((rows == rows[0]).all())
        decode = (T == 1) or not single_seq
        if decode and len(set(rows.tolist())) != T:
            raise ValueError("batched decode needs one token per distinct cache row")
        if not decode and not (np.diff(pos) == 1).all():
            raise ValueError("a prefill chunk must cover consecutive positions")
        return pos, rows, T, decode

    def forward(self, x, kv_cache: KVCache, positions, seq_ids=None, logits: str = "last", out=None, trace=None):
        """Run this stage's layers over T tokens (eager launches), on this stage's device whatever device is
        current in the calling thread (one process may hold stages on several GPUs; gRPC handler threads start
        with device 0 current).  See `_forward` for the arguments."""
        with torch.cuda.device(self.device):
            return self._forward(x, kv_cache, positions, seq_ids, logits, out, trace)

    def _forward(self, x, kv_cache: KVCache, positions, seq_ids=None, logits: str = "last", out=None, trace=None):
        """Run this stage's layers over T tokens (eager launches).

        positions : host ints [T] (list / numpy / CPU tensor).
        seq_ids   : None -> all tokens belong to cache row 0 (T > 1 = one prefill chunk at
                    consecutive positions; T == 1 = one decode step); host ints [T] with distinct
                    rows -> batched decode, one token per sequence.
        trace     : optional list; receives per-layer clones of the intermediates (tests only).
        logits    : last stage only: "last" (final row(s): last row for a chunk, every row for
                    batched decode), "all" (every row; the `model(**inputs).logits` form used by
                    confidence_score, models_quant_updated.py:63-70) or "none" (hidden only).
        """
        pos, rows, T, decode = self._classify(positions, seq_ids)
        if int(pos.max()) >= kv_cache.max_seq:
            raise ValueError("position beyond the KV cache capacity")
        buf = self._buffers(T)
        dpos = buf.pos[:T]
        dpos.copy_(torch.from_numpy(pos.astype(np.int32)))
        dseq = dlens = ws = None
        max_len = 0
        if decode:
            dseq = buf.seq[:T]
            dseq.copy_(torch.from_numpy(rows.astype(np.int32)))
            lens_host = (pos + 1).astype(np.int32)
            dlens = buf.lens[:T]
            dlens.copy_(torch.from_numpy(lens_host))
            max_len = int(lens_host.max())
            ws = self._workspace(ops.decode_attn_workspace_bytes(T, self.cfg.n_heads, self.cfg.n_kv_heads,
                                                                 self.cfg.head_dim, max_len, self.cfg.window))
        if self.first and not x.is_cuda:
            x = x.to(self.device, non_blocking=True)
        y, n = self._core(buf, x, T, kv_cache, decode, dpos, dseq, dlens, int(rows[0]), max_len, ws, logits, out, trace)
        self.launches += n
        for r_, p_ in zip(rows.tolist(), pos.tolist()):
            kv_cache.lens[r_] = max(kv_cache.lens[r_], p_ + 1)
        if out is None and (not self.last or logits == "none"):
            return y.clone()      # y aliases the scratch buffer
        return y

    # ------------------------------------------------------------------------------------------
    def forward_auto(self, x, kv_cache: KVCache, positions, seq_ids=None, logits: str = "last", graph_chunk: int = 0):
        """forward() through a captured CUDA graph when the shape is a repeating one: every decode
        step, and prefill chunks of exactly `graph_chunk` tokens.  Anything else (ragged tails,
        one-off shapes) launches eagerly.  The returned tensor may be the runner's static output
        buffer: consume it before the next call with the same shape."""
        pos, rows, T, decode = self._classify(positions, seq_ids)
        if decode:
            r = self.runner(kv_cache, T, decode=True, rows=rows.tolist(), logits=logits)
            return r(x, pos)
        if graph_chunk and T == graph_chunk:
            r = self.runner(kv_cache, T, decode=False, rows=[int(rows[0])], logits=logits)
            return r(x, pos)
        return self.forward(x, kv_cache, positions, seq_ids, logits=logits)

    def drop_runners(self):
        """Release every captured graph (and the scratch buffers they pin)."""
        self._runners.clear()

    def runner(self, kv_cache: KVCache, T: int, *, decode: bool, rows=None, logits: str = "last") -> "GraphRunner":
        """CUDA-graph runner for a fixed token count: a decode step over `rows` (one token per row)
        or a prefill chunk of T consecutive positions of cache row rows[0].
        logits: "none" (hidden state out), "last" / "all" (logits), and for decode steps of the last stage "split"
        (upper vocabulary shard's winner + normed hidden state as one message), "token" (greedy token from
        b2_lm_head_argmax, no logits) and "norm" (stop at the final norm: the head runs on other stages)."""
        rows_t = tuple(int(r) for r in (rows if rows is not None else ([0] * (1 if decode else T))))
        if decode and len(rows_t) != T:
            raise ValueError("decode runner needs one row per token")
        key = (id(kv_cache), T, decode, rows_t if decode else rows_t[:1], logits)
        r = self._runners.get(key)
        if r is None:
            r = GraphRunner(self, kv_cache, T, decode, rows_t, logits)
            self._runners[key] = r
        return r


def head_message(T: int, d: int, device):
    """The buffer the last stage sends to the first one per decode step when they share the lm_head, and its three
    views: normed hidden [T,d] fp16, upper-shard maxima [T] fp32, their global indices [T] int64."""
    off_val = T * d * 2
    off_idx = off_val + ((T * 4 + 7) & ~7)
    total = (off_idx + T * 8 + 15) & ~15
    msg = torch.zeros((total,), dtype=torch.uint8, device=device)
    return (msg,) + head_message_views(msg, T, d)


def head_message_views(msg, T: int, d: int):
    """(hidden [T,d] fp16, maxima [T] fp32, indices [T] int64) views of a head message held in the uint8 buffer `msg`."""
    off_val = T * d * 2
    off_idx = off_val + ((T * 4 + 7) & ~7)
    return (msg[:off_val].view(torch.float16).view(T, d), msg[off_val: off_val + T * 4].view(torch.float32),
            msg[off_idx: off_idx + T * 8].view(torch.int64))


def head_message_bytes(T: int, d: int) -> int:
    return (T * d * 2 + ((T * 4 + 7) & ~7) + T * 8 + 15) & ~15


class GraphRunner:
    """One captured CUDA graph = one stage forward at a fixed shape.

    Static device state: input (ids or hidden), positions (and lengths for decode), output.  The
    graph ends by advancing positions/lengths by its own token stride, so consecutive decode steps
    (or consecutive chunks of one prompt) replay it with no host->device traffic besides the input.
    `set_positions` re-seeds that state (start of a new prompt / after a ragged tail).
    """

    def __init__(self, stage: Stage, kv_cache: KVCache, T: int, decode: bool, rows, logits: str):
        cfg, dev = stage.cfg, stage.device
        self.stage, self.cache, self.T, self.decode, self.rows, self.logits = stage, kv_cache, T, decode, rows, logits
        self.buf = _Scratch(cfg, max(T, 8), dev, stage.last)      # private: graphs pin their buffers
        self.x = self.buf.ids[:T] if stage.first else self.buf.h[0][:T]
        self.dpos = self.buf.pos[:T]
        self.dseq = self.dlens = self.ws = None
        self.max_len = 0
        if decode:
            self.dseq = self.buf.seq[:T]
            with torch.cuda.device(dev):
                self.dseq.copy_(torch.tensor(rows, dtype=torch.int32))
            self.dlens = self.buf.lens[:T]
            self.max_len = kv_cache.max_seq      # split count fixed for the whole graph lifetime
            nb = ops.decode_attn_workspace_bytes(T, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, self.max_len, cfg.window)
            self.ws = torch.empty((nb,), dtype=torch.uint8, device=dev) if nb else None
        n_out = T if (decode or logits in ("all", "split")) else 1
        self.msg = self.hn_msg = self.val_msg = self.idx_msg = self.tok = None
        self.fused_head = stage.last and decode and logits in ("split", "token") and stage.can_fuse_head(T)
        if logits == "norm":
            # the graph ends at the final norm: self.hn_msg [T,d] fp16 is what the last stage sends to every stage
            if not (stage.last and decode):
                raise ValueError("logits='norm' is a decode runner of the last stage")
            self.hn_msg = torch.empty((T, cfg.d_model), dtype=torch.float16, device=dev)
            self.out = None
        elif logits == "token":
            # greedy decode on the last stage: the graph ends in b2_lm_head_argmax -> self.tok [T] int64 (and, on a
            # one-stage pipeline, straight into the next step's input ids); no logits are produced
            if not self.fused_head:
                raise ValueError("logits='token' needs a decode runner of <= 4 rows on the last stage (no lm_head bias)")
            self.tok = torch.empty((T,), dtype=torch.int64, device=dev)
            self.out = None
        elif stage.last and logits == "split":
            # one message to the first stage per step: normed hidden [T,d] fp16 | shard maxima [T] fp32 | indices [T] int64
            self.out = torch.empty((n_out, cfg.vocab - stage.head_split()), dtype=torch.float32, device=dev)
            self.msg, self.hn_msg, self.val_msg, self.idx_msg = head_message(T, cfg.d_model, dev)
        elif stage.last and logits != "none":
            self.out = torch.empty((n_out, cfg.vocab), dtype=torch.float32, device=dev)
        else:
            self.out = torch.empty((T, cfg.d_model), dtype=torch.float16, device=dev)
        self.stride = 1 if decode else T
        self.host_pos = None
        self.set_positions(np.zeros(T, dtype=np.int64) if decode else np.arange(T))
        self.graph = None
        self.n_launch = 0
        self._capture()

    def _run_core(self):
        head_out = None
        if self.fused_head:
            head_out = (self.val_msg, self.idx_msg, None) if self.msg is not None else \
                       (None, self.tok, self.x if self.stage.first else None)
        _, n = self.stage._core(self.buf, self.x, self.T, self.cache, self.decode, self.dpos, self.dseq, self.dlens,
                                self.rows[0], self.max_len, self.ws, "last" if self.logits == "token" else self.logits,
                                self.out, hn_out=self.hn_msg, head_out=head_out)
        if self.msg is not None and not self.fused_head:
            ops.argmax_shard(self.out, self.stage.head_split(), out_idx=self.idx_msg, out_val=self.val_msg)
            n += 1
        self.dpos.add_(self.stride)
        if self.dlens is not None:
            self.dlens.add_(self.stride)
        return n

    def _capture(self):
        # No warm-up run: it would write K/V into a cache that may already hold a prompt.  All
        # one-time set-up was done by b2_init() (Stage.__init__), so the first launch of every
        # kernel can happen under capture.
        dev = self.stage.device
        saved_pos = self.host_pos.copy()
        torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        # Garbage collection must not run inside the capture: freeing an older runner there
        # destroys its CUDA graph, which is illegal while a stream is capturing.
        gc.collect()
        gc_was_on = gc.isenabled()
        gc.disable()
        try:
            # The capture stream must belong to THIS stage's device: torch's default capture stream is one process-wide
            # stream on whichever device captured first, and entering it switches the current device -- a stage on
            # another GPU would then record its kernels on the wrong device (they ran there through peer access when
            # that happened to be enabled, and faulted when it was not).
            # thread_local: another host thread may be driving another pipeline and allocate meanwhile.
            with torch.cuda.device(dev), torch.cuda.graph(g, stream=torch.cuda.Stream(dev), capture_error_mode="thread_local"):
                self.n_launch = self._run_core()
        finally:
            if gc_was_on:
                gc.enable()
        self.graph = g
        self.set_positions(saved_pos)

    def set_positions(self, positions):
        pos = np.asarray(positions, dtype=np.int64).reshape(-1)
        self.host_pos = pos.copy()
        with torch.cuda.device(self.stage.device):
            self.dpos.copy_(torch.from_numpy(pos.astype(np.int32)))
            if self.dlens is not None:
                self.dlens.copy_(torch.from_numpy((pos + 1).astype(np.int32)))

    def replay(self):
        """Replay on the state the graph itself maintains (input already in `self.x`, positions
        advanced by the previous replay): the per-tick path of DistPipeline's decode stream."""
        with torch.cuda.device(self.stage.device):      # the graph launches on the current device's current stream
            self.graph.replay()
        self.stage.launches += self.n_launch
        lens = self.cache.lens
        for r_, p_ in zip(self.rows, self.host_pos.tolist()):
            if lens[r_] < p_ + 1:
                lens[r_] = p_ + 1
        self.host_pos = self.host_pos + self.stride
        return self.out

    def __call__(self, x, positions=None):
        """Replay.  x: ids [T] int64 (first stage) or hidden [T,d] fp16, on the device, or None if
        the caller already wrote into `self.x`.  positions (host) only when they differ from the
        graph's own running state.  Returns the static output tensor (valid until the next replay)."""
        if positions is not None:
            p = np.asarray(positions, dtype=np.int64).reshape(-1)
            if not np.array_equal(p, self.host_pos):
                self.set_positions(p)
        if int(self.host_pos.max()) >= self.cache.max_seq:
            raise ValueError("position beyond the KV cache capacity")
        with torch.cuda.device(self.stage.device):
            if x is not None and x.data_ptr() != self.x.data_ptr():
                self.x.copy_(x.view(self.x.shape), non_blocking=True)
            self.graph.replay()
        self.stage.launches += self.n_launch
        rows = self.rows if self.decode else [self.rows[0]] * self.T
        for r_, p_ in zip(rows, self.host_pos.tolist()):
            self.cache.lens[r_] = max(self.cache.lens[r_], p_ + 1)
        self.host_pos = self.host_pos + self.stride
        return self.out
