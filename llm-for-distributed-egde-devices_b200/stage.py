"""Stage: the quantised decoder-block forward over a contiguous layer range.

Build-defined drop-in boundary modelled on the reference's only conventions,
``load_int8_model(model_path) -> (model, tokenizer)`` and ``model(**inputs).logits`` /
``model.generate(...)`` (Code/Quantised Models/models_quant_updated.py:25-40, :66-67, :84-93):

    stage = Stage.load(model_dir_or_config, layer_begin, layer_end, first=..., last=..., device=...)
    y = stage.forward(x, kv_cache, positions)

``x`` is ``input_ids [T] int64`` on the first stage, else ``hidden [T,d] fp16``; the result is
``hidden [T,d] fp16``, or ``logits fp32`` on the last stage.  ``kv_cache`` is caller-owned and
updated in place.  Every arithmetic step is a call into libb200stage.so (ops.py); torch only owns
device memory and streams.  There is no CPU path: loading on a non-CUDA device raises.
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops, weights
from .config import ModelConfig, get_config


class KVCache:
    """Stage-local KV cache: k, v fp16 [n_layers][rows][KV][Smax][hd] plus per-row lengths.

    Lengths live on the host (python ints drive launch geometry without device syncs) and are
    mirrored to a device int32 vector that the decode-attention kernel reads.
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
        self.hn = torch.empty((cap, d), dtype=f16, device=device) if last else None

    def q8view(self, T, K):
        return self.q8[: T * K].view(T, K)


class Stage:
    def __init__(self, cfg: ModelConfig, sw: dict, layer_begin: int, layer_end: int, device):
        if torch.device(device).type != "cuda":
            raise RuntimeError("Stage runs only on a CUDA device: the hot path has no CPU fallback")
        ops._abi.lib()   # fail loudly now if the extension is missing
        self.cfg, self.w = cfg, sw
        self.first, self.last = bool(sw["first"]), bool(sw["last"])
        self.layer_begin, self.layer_end = layer_begin, layer_end
        self.device = torch.device(device)
        cos, sin = weights.rope_table(cfg.head_dim, cfg.max_pos, cfg.rope_theta)
        self.cos, self.sin = cos.to(self.device), sin.to(self.device)
        self.scale = 1.0 / float(np.sqrt(cfg.head_dim))
        self._scratch = None
        self._ws = None
        self.launches = 0          # kernels of ours launched (bench.py's gpu_launches claim)

    # ------------------------------------------------------------------------------------------
    @classmethod
    def load(cls, model_dir_or_config, layer_begin: int, layer_end: int, *, first: bool, last: bool,
             device="cuda", quant: str = "w8a8", seed: int = 0, fast_init: bool = False, act_stats=None) -> "Stage":
        """Load (here: synthesise under `seed`, no checkpoints exist offline) the W8A8 shard for
        layers [layer_begin, layer_end).  `fast_init` generates on the device (benchmarks)."""
        if quant != "w8a8":
            raise ValueError("only quant='w8a8' is implemented")
        cfg = get_config(model_dir_or_config)
        if not (0 <= layer_begin < layer_end <= cfg.n_layers):
            raise ValueError(f"bad layer range [{layer_begin},{layer_end}) for {cfg.n_layers} layers")
        if fast_init:
            sw = weights.build_stage_weights_fast(cfg, layer_begin, layer_end, first, last, seed, device)
        else:
            sw = weights.build_stage_weights(cfg, layer_begin, layer_end, first, last, seed, device, act_stats)
        return cls(cfg, sw, layer_begin, layer_end, device)

    def new_kv_cache(self, rows: int, max_seq: int) -> KVCache:
        return KVCache(self.cfg, self.layer_end - self.layer_begin, rows, max_seq, self.device)

    def weight_bytes(self) -> int:
        n = 0
        for lw in self.w["layers"]:
            n += sum(v.numel() * v.element_size() for v in lw.values())
        return n

    def _buffers(self, T: int) -> _Scratch:
        if self._scratch is None or self._scratch.cap < T:
            cap = max(T, 8)
            self._scratch = _Scratch(self.cfg, cap, self.device, self.last)
        return self._scratch

    def _workspace(self, nbytes: int):
        if nbytes == 0:
            return None
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = torch.empty((nbytes,), dtype=torch.uint8, device=self.device)
        return self._ws

    # ------------------------------------------------------------------------------------------
    def forward(self, x, kv_cache: KVCache, positions, seq_ids=None, logits: str = "last", out=None, trace=None):
        """Run this stage's layers over T tokens.

        positions : host ints [T] (list / numpy / CPU tensor).
        seq_ids   : None -> all tokens belong to cache row 0 (T > 1 = one prefill chunk at
                    consecutive positions; T == 1 = one decode step); host ints [T] with distinct
                    rows -> batched decode, one token per sequence.
        trace     : optional list; receives per-layer clones of the intermediates (tests only).
        logits    : last stage only: "last" (final row(s): last row for a chunk, every row for
                    batched decode), "all" (every row; the `model(**inputs).logits` form used by
                    confidence_score, models_quant_updated.py:63-70) or "none" (hidden only).
        """
        cfg = self.cfg
        pos = np.asarray(positions, dtype=np.int64).reshape(-1)
        T = int(pos.shape[0])
        if T == 0:
            raise ValueError("empty input")
        if seq_ids is None:
            rows = np.zeros((T,), dtype=np.int64)
        else:
            rows = np.asarray(seq_ids, dtype=np.int64).reshape(-1)
        single_seq = bool((rows == rows[0]).all())
        decode = (T == 1) or not single_seq
        if decode and len(set(rows.tolist())) != T:
            raise ValueError("batched decode needs one token per distinct cache row")
        if not decode and not (np.diff(pos) == 1).all():
            raise ValueError("a prefill chunk must cover consecutive positions")
        if int(pos.max()) >= kv_cache.max_seq:
            raise ValueError("position beyond the KV cache capacity")
        buf = self._buffers(T)
        d, H, KV, hd, F = cfg.d_model, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.ffn
        dpos = buf.pos[:T]
        dpos.copy_(torch.from_numpy(pos.astype(np.int32)), non_blocking=True)
        dseq = None
        if decode:
            dseq = buf.seq[:T]
            dseq.copy_(torch.from_numpy(rows.astype(np.int32)), non_blocking=True)
            lens_host = (pos + 1).astype(np.int32)
            dlens = buf.lens[:T]
            dlens.copy_(torch.from_numpy(lens_host), non_blocking=True)
            max_len = int(lens_host.max())
            ws = self._workspace(ops.decode_attn_workspace_bytes(T, H, KV, hd, max_len, cfg.window))
        # ---- input -----------------------------------------------------------------------------
        if self.first:
            ids = x if x.is_cuda else x.to(self.device, non_blocking=True)
            h = ops.embed(ids.view(-1), self.w["embed"], dpos if "pos_embed" in self.w else None,
                          self.w.get("pos_embed"), out=buf.h[0][:T])
            self.launches += 1
        else:
            h = buf.h[0][:T]
            h.copy_(x.view(T, d))
        h_alt = buf.h[1][:T]
        qkv, attn, gu = buf.qkv[:T], buf.attn[:T], buf.gu[:T]
        sc = buf.scale[:T]
        act = (ops.ACT_SILU_MUL if cfg.act == "silu" else ops.ACT_GELU_MUL) if cfg.gated else \
              (ops.ACT_GELU if cfg.act == "gelu_new" else ops.ACT_SILU)
        # ---- layers ----------------------------------------------------------------------------
        for li, lw in enumerate(self.w["layers"]):
            kc, vc = kv_cache.k[li], kv_cache.v[li]
            q8 = buf.q8view(T, d)
            ops.norm_quant(h, lw["ln1_g"], lw.get("ln1_b"), lw["sm_qkv"], cfg.eps, out_q=q8, out_scale=sc)
            ops.gemm_w8a8(q8, lw["wqkv"], sc, lw["s_qkv"], lw.get("b_qkv"), None, out=qkv)
            tr = None
            if trace is not None:
                tr = dict(q1=q8.clone(), s1=sc.clone(), qkv=qkv.clone())
                trace.append(tr)
            ops.rope_kv_append(qkv, self.cos, self.sin, dpos, kc, vc, H, KV, hd, seq_ids=dseq, seq0=int(rows[0]),
                               use_rope=cfg.rope)
            if decode:
                ops.decode_attn(qkv, kc, vc, dlens, H, KV, hd, max_len, self.scale, cfg.window, rows=dseq, out=attn,
                                workspace=ws)
                self.launches += 1 if ws is None else 2
            else:
                ops.prefill_attn(qkv, kc, vc, int(pos[0]), int(rows[0]), H, KV, hd, self.scale, cfg.window, out=attn)
                self.launches += 1
            q8 = buf.q8view(T, H * hd)
            ops.smooth_quant(attn, lw["sm_o"], out_q=q8, out_scale=sc)
            ops.gemm_w8a8(q8, lw["wo"], sc, lw["s_o"], lw.get("b_o"), h, out=h_alt)
            q8 = buf.q8view(T, d)
            ops.norm_quant(h_alt, lw["ln2_g"], lw.get("ln2_b"), lw["sm_gu"], cfg.eps, out_q=q8, out_scale=sc)
            ops.gemm_w8a8(q8, lw["wgu"], sc, lw["s_gu"], lw.get("b_gu"), None, out=gu)
            q8 = buf.q8view(T, F)
            ops.act_mul_quant(gu, lw["sm_d"], act, out_q=q8, out_scale=sc)
            ops.gemm_w8a8(q8, lw["wd"], sc, lw["s_d"], lw.get("b_d"), h_alt, out=h)
            self.launches += 9
            if tr is not None:
                tr.update(attn=attn.clone(), h1=h_alt.clone(), gu=gu.clone(), q4=q8.clone(), s4=sc.clone(), h2=h.clone())
        for r_, p_ in zip(rows.tolist(), pos.tolist()):
            kv_cache.lens[r_] = max(kv_cache.lens[r_], p_ + 1)
        # ---- output ----------------------------------------------------------------------------
        if not self.last or logits == "none":
            if out is not None:
                out.copy_(h)
                return out
            return h.clone()
        hsel = h if (logits == "all" or decode) else h[T - 1:T]
        hn = ops.norm(hsel, self.w["lnf_g"], self.w.get("lnf_b"), cfg.eps, out=buf.hn[: hsel.shape[0]])
        lg = ops.lm_head(hn, self.w["lm_head"], out=out)
        self.launches += 2
        return lg
