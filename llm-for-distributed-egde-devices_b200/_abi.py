"""ctypes binding of libb200stage.so (the C ABI declared in include/b200stage.h).

The product path fails loudly when the CUDA library is missing: there is no CPU or PyTorch
fallback anywhere in this package.  Tensors are passed as raw device pointers
(``tensor.data_ptr()``) plus the current CUDA stream handle; torch is used for device memory and
streams only.
"""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# B2_LIBRARY: load another build of the same ABI (the -DB2_TIMELINE probe build of tools/timeline.py)
LIB_PATH = os.environ.get("B2_LIBRARY") or os.path.join(_HERE, "libb200stage.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "b200stage.h")

_vp, _i, _f, _ll, _sz = C.c_void_p, C.c_int, C.c_float, C.c_longlong, C.c_size_t

# name -> (restype, argtypes); must list every function include/b200stage.h declares
SIGNATURES = {
    "b2_version": (_i, []),
    "b2_strerror": (C.c_char_p, [_i]),
    "b2_init": (_i, []),
    "b2_norm_quant": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _f, _vp]),
    "b2_norm": (_i, [_vp, _vp, _vp, _vp, _i, _i, _f, _vp]),
    "b2_smooth_quant": (_i, [_vp, _vp, _vp, _vp, _i, _i, _vp]),
    "b2_act_mul_quant": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "b2_gemm_w8a8": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "b2_gemm_w8a8_sched": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "b2_gemm_gated_act": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "b2_quant_f32": (_i, [_vp, _vp, _vp, _vp, _i, _i, _vp]),
    "b2_gemm_qkv_rope": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _ll, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _i,
                              _i, _vp]),
    "b2_rope_kv_append": (_i, [_vp, _ll, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "b2_decode_attn": (_i, [_vp, _ll, _vp, _vp, _vp, _vp, _vp, _ll, _vp, _sz, _i, _i, _i, _i, _i, _i, _f, _i, _vp, _ll, _vp,
                             _ll, _vp]),
    "b2_decode_attn_static": (_i, [_vp, _ll, _vp, _vp, _vp, _vp, _vp, _ll, _i, _i, _i, _i, _i, _i, _f, _i, _vp]),
    "b2_decode_attn_workspace_bytes": (_sz, [_i, _i, _i, _i, _i, _i]),
    "b2_set_prefill_attn_impl": (_i, [_i]),
    "b2_act_mul": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "b2_outlier_scan": (_i, [_vp, _ll, _i, _i, _f, _vp, _vp, _vp, _vp, _vp]),
    "b2_outlier_side_dense": (_i, [_vp, _vp, _ll, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "b2_outlier_side": (_i, [_vp, _vp, _ll, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "b2_prefill_attn": (_i, [_vp, _ll, _vp, _vp, _vp, _ll, _vp, _i, _i, _i, _i, _i, _i, _i, _f, _i, _vp]),
    "b2_embed": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "b2_lm_head": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "b2_fused_linear": (_i, [_vp, _ll, _i, _vp, _vp, _vp, _vp, _f, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp,
                             _i, _i, _i, _vp, _ll, _vp, _ll, _vp]),
    "b2_fused_qkv": (_i, [_vp, _ll, _i, _vp, _vp, _vp, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp,
                          _i, _i, _i, _i, _i, _i, _i, _vp, _ll, _vp, _ll, _vp]),
    "b2_fused_layer_tail": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "b2_sample": (_i, [_vp, _ll, _vp, _vp, _ll, _vp, _i, _i, _f, _i, _f, _f, C.c_ulonglong, _vp, C.c_ulonglong, _vp]),
    "b2_argmax": (_i, [_vp, _vp, _i, _i, _vp]),
    "b2_p2p_send": (_i, [_vp, _vp, _ll, _vp, C.c_uint, _vp]),
    "b2_p2p_wait": (_i, [_vp, C.c_uint, _vp, _vp, _ll, _vp, _vp]),
    "b2_argmax_shard": (_i, [_vp, _ll, _vp, _vp, _ll, _vp, _vp, _i, _i, _vp]),
    "b2_lm_head_argmax_scratch_bytes": (_ll, []),
    "b2_lm_head_argmax": (_i, [_vp, _vp, _vp, _i, _i, _i, _ll, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_uint, _vp, _vp, C.c_uint, _vp]),
    "b2_p2p_send_multi": (_i, [_vp, _ll, _vp, _vp, _i, C.c_uint, _vp]),
    "b2_merge_winners": (_i, [_vp, C.c_uint, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "b2_bias_add_f32": (_i, [_vp, _vp, _ll, _i, _i, _vp]),
    "b2_confidence": (_i, [_vp, _ll, _vp, _vp, _i, _i, _vp]),
}



class FusedDesc(C.Structure):
    """struct b2_fused_desc (include/b200stage.h)."""
    _fields_ = [
        ("x", _vp), ("x_ld", _ll), ("x_kind", _i),
        ("gamma", _vp), ("beta", _vp), ("smooth", _vp), ("x_amax", _vp), ("eps", _f),
        ("W", _vp), ("w_scale", _vp), ("bias", _vp), ("residual", _vp),
        ("out", _vp), ("out_kind", _i), ("act", _i), ("out_smooth", _vp), ("out_amax", _vp), ("zero_amax", _vp),
        ("N", _i), ("K", _i),
        ("cos_table", _vp), ("sin_table", _vp), ("positions", _vp), ("seq_ids", _vp), ("seq0", _i),
        ("kcache", _vp), ("vcache", _vp), ("H", _i), ("KV", _i), ("hd", _i), ("Smax", _i), ("use_rope", _i),
        ("prefetch0", _vp), ("prefetch0_bytes", _ll), ("prefetch1", _vp), ("prefetch1_bytes", _ll),
    ]


_lib = None


def build(verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into libb200stage.so (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", os.path.join(_HERE, "csrc"), "-j8"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libb200stage.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    if verbose:
        print(r.stdout[-2000:])
    return LIB_PATH


def header_functions() -> list[str]:
    """Names of all functions declared in include/b200stage.h."""
    src = open(HEADER_PATH).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", src)))


def lib() -> C.CDLL:
    """Load the shared library (once).  Raises if it has not been built -- never falls back."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the CUDA extension has not been built "
                "(run __graft_entry__.build() or `make -C llm-for-distributed-egde-devices_b200/csrc`). "
                "There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


_inited = set()


def init_device() -> None:
    """b2_init() once per CUDA device (call with that device current)."""
    import torch
    dev = torch.cuda.current_device()
    if dev not in _inited:
        check(lib().b2_init(), "b2_init")
        _inited.add(dev)


class B2Error(RuntimeError):
    pass


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise B2Error(f"{what}: {lib().b2_strerror(rc).decode()} (code {rc})")


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_handle():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
