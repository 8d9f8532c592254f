"""ctypes binding of libfa100.so (include/fa100.h).  No CPU fallback: if the library is missing or a
CUDA device is absent the forward raises."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, os.environ.get("FA100_LIB", "libfa100.so"))  # FA100_LIB=libfa100_trace.so: timeline build

BF16, FP16, E4M3, FP32 = 0, 1, 2, 3
ERR_INVALID, ERR_UNSUPPORTED, ERR_CUDA = -1, -2, -3

i32, i64, f32, vp = C.c_int32, C.c_int64, C.c_float, C.c_void_p


class FwdParams(C.Structure):
    _fields_ = (
        [("q", vp), ("k", vp), ("v", vp), ("o", vp), ("lse", vp)]
        + [(n, i64) for n in (
            "q_batch_stride", "q_row_stride", "q_head_stride",
            "k_batch_stride", "k_row_stride", "k_head_stride",
            "v_batch_stride", "v_row_stride", "v_head_stride",
            "o_batch_stride", "o_row_stride", "o_head_stride")]
        + [("o_accum", vp), ("lse_accum", vp)]
        + [(n, i32) for n in ("b", "b_k", "seqlen_q", "seqlen_k", "total_q", "total_k", "h", "h_k", "d", "dv", "dtype")]
        + [(n, vp) for n in ("cu_seqlens_q", "cu_seqlens_k", "seqused_q", "seqused_k", "leftpad_k", "kv_batch_idx")]
        + [("page_table", vp), ("page_table_batch_stride", i64), ("page_size", i32), ("num_pages", i32)]
        + [("k_new", vp), ("v_new", vp)]
        + [(n, i64) for n in (
            "knew_batch_stride", "knew_row_stride", "knew_head_stride",
            "vnew_batch_stride", "vnew_row_stride", "vnew_head_stride")]
        + [("seqlen_knew", i32), ("total_knew", i32), ("cu_seqlens_knew", vp)]
        + [("rotary_cos", vp), ("rotary_sin", vp), ("seqlens_rotary", vp), ("rotary_dim", i32), ("is_rotary_interleaved", i32)]
        + [("q_descale", vp), ("k_descale", vp), ("v_descale", vp)]
        + [(n, i64) for n in (
            "q_descale_batch_stride", "q_descale_head_stride",
            "k_descale_batch_stride", "k_descale_head_stride",
            "v_descale_batch_stride", "v_descale_head_stride")]
        + [("softmax_scale", f32), ("softcap", f32)]
        + [(n, i32) for n in ("is_causal", "window_left", "window_right", "attention_chunk")]
        + [(n, i32) for n in ("num_splits", "pack_gqa", "num_sm")]
        + [("scheduler_metadata", vp), ("variant", i32), ("merge_partial", i32), ("lse_accum_prefilled", i32), ("reserved0", i32), ("tile_counter", vp), ("reserved", i32 * 2)]
        + [("qv", vp), ("qv_batch_stride", i64), ("qv_row_stride", i64), ("qv_head_stride", i64)]
    )


class BwdParams(C.Structure):
    _fields_ = (
        [(n, vp) for n in ("dout", "q", "k", "v", "out", "lse", "dq", "dk", "dv", "softmax_d")]
        + [(f"{t}_{k}_stride", i64) for t in ("q", "k", "v", "o", "do", "dq", "dk", "dv") for k in ("batch", "row", "head")]
        + [(n, i32) for n in ("b", "seqlen_q", "seqlen_k", "h", "h_k", "d", "dv_dim", "dtype", "is_causal", "window_left", "window_right")]
        + [("softmax_scale", f32), ("softcap", f32), ("deterministic", i32), ("num_sm", i32)]
        + [(n, vp) for n in ("cu_seqlens_q", "cu_seqlens_k", "seqused_q", "seqused_k")]
        + [("total_q", i32), ("total_k", i32), ("reserved", i32 * 6)]
    )


class CombineParams(C.Structure):
    _fields_ = (
        [("o_partial", vp), ("lse_partial", vp), ("o", vp), ("lse", vp)]
        + [(n, i64) for n in (
            "op_split_stride", "op_batch_stride", "op_row_stride", "op_head_stride",
            "lp_split_stride", "lp_batch_stride", "lp_row_stride", "lp_head_stride",
            "o_batch_stride", "o_row_stride", "o_head_stride",
            "lse_batch_stride", "lse_row_stride", "lse_head_stride")]
        + [(n, i32) for n in ("num_splits", "b", "s", "h", "d", "out_dtype")]
        + [("cu_seqlens", vp), ("seqused", vp), ("reserved", i32 * 8)]
    )


class Shape(C.Structure):
    _fields_ = [(n, i32) for n in (
        "b", "seqlen_q", "seqlen_k", "h", "h_k", "d", "dv", "dtype", "is_causal", "is_local", "is_varlen", "is_paged",
        "page_size", "has_new_kv", "num_splits", "pack_gqa", "num_sm")]


class Plan(C.Structure):
    _fields_ = [(n, i32) for n in ("block_m", "block_n", "num_splits", "pack_gqa", "num_sm", "num_work", "smem_bytes",
                                   "dual_stream")] + [("reserved", i32 * 8)]


EXPORTS = ("fa100_abi_version", "fa100_last_error", "fa100_plan", "fa100_fwd", "fa100_combine", "fa100_bwd", "fa100_append_kv", "fa100_prepare_scheduler",
           "fa100_scheduler_metadata_words", "fa100_fwd_splitkv_noalloc", "fa100_noalloc_scratch_bytes", "fa100_debug_hang_info",
           "fa100_debug_trace", "fa100_debug_prof")

_lib = None


def build(verbose: bool = False) -> str:
    """Compile libfa100.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", HERE, "-j4"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libfa100.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    if verbose:
        print(r.stdout[-2000:])
    return LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU or PyTorch fallback for the attention forward)")
        l = C.CDLL(LIB_PATH)
        l.fa100_abi_version.restype = C.c_int
        l.fa100_last_error.restype = C.c_char_p
        l.fa100_plan.argtypes = [C.POINTER(Shape), C.POINTER(Plan)]
        l.fa100_fwd.argtypes = [C.POINTER(FwdParams), vp]
        l.fa100_combine.argtypes = [C.POINTER(CombineParams), vp]
        l.fa100_bwd.argtypes = [C.POINTER(BwdParams), vp]
        l.fa100_bwd.restype = C.c_int
        l.fa100_append_kv.argtypes = [C.POINTER(FwdParams), vp, i64, i64, i64, vp, vp]
        l.fa100_prepare_scheduler.argtypes = [C.POINTER(FwdParams), vp, i64, vp]
        l.fa100_scheduler_metadata_words.argtypes = [i32]
        l.fa100_scheduler_metadata_words.restype = i64
        l.fa100_fwd_splitkv_noalloc.argtypes = [C.POINTER(FwdParams), vp]
        l.fa100_fwd_splitkv_noalloc.restype = C.c_int
        l.fa100_noalloc_scratch_bytes.restype = i64
        l.fa100_debug_hang_info.argtypes = [C.POINTER(C.c_uint32 * 4)]
        l.fa100_debug_trace.argtypes = [C.POINTER(C.c_uint32), C.c_int]
        l.fa100_debug_prof.argtypes = [C.POINTER(C.c_uint32), C.c_int]
        for n in ("fa100_plan", "fa100_fwd", "fa100_combine", "fa100_append_kv", "fa100_prepare_scheduler", "fa100_debug_hang_info",
           "fa100_debug_trace", "fa100_debug_prof"):
            getattr(l, n).restype = C.c_int
        if l.fa100_abi_version() != 2:
            raise RuntimeError("libfa100.so ABI version mismatch")
        _lib = l
    return _lib


def check(rc: int) -> None:
    if rc == 0:
        return
    msg = lib().fa100_last_error().decode()
    if rc == ERR_UNSUPPORTED:
        raise NotImplementedError("fa100: " + msg)
    raise RuntimeError("fa100: " + msg)


def hang_info():
    out = (C.c_uint32 * 4)()
    check(lib().fa100_debug_hang_info(C.byref(out)))
    return list(out)


def trace(max_events: int = 49152):
    """Events of the timeline build as a list of (tag, warp, begin, end); empty from the product library."""
    buf = (C.c_uint32 * (3 * max_events))()
    n = lib().fa100_debug_trace(buf, max_events)
    return [(buf[3 * i] & 0xFFFF, buf[3 * i] >> 16, buf[3 * i + 1], buf[3 * i + 2]) for i in range(max(n, 0))]


def prof(words: int = 160):
    """Per-warp cycle buckets of the timeline build: list of 8-tuples, one per warp; empty from the product library."""
    buf = (C.c_uint32 * words)()
    n = lib().fa100_debug_prof(buf, words)
    return [tuple(buf[w * 8 + i] for i in range(8)) for w in range(max(n, 0) // 8)]
