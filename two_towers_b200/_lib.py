"""ctypes binding of libtt_b200.so (C ABI: include/tt_b200.h).  There is NO fallback: if the library is
missing or fails to load, every op raises.  PyTorch is used only for device memory, streams and autograd
bookkeeping; all arithmetic happens in the library's kernels."""
import ctypes as C
import os

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libtt_b200.so")

c_f32p = C.c_void_p
u64 = C.c_uint64
i64 = C.c_int64
i32 = C.c_int32


class GruLayer(C.Structure):
    _fields_ = [
        ("B", i32), ("T", i32), ("In", i32), ("H", i32), ("ndir", i32), ("save_for_backward", i32),
        ("x", C.c_void_p), ("ids", C.c_void_p), ("ids_are_i64", i32), ("_pad0", i32),
        ("table", C.c_void_p), ("vocab", i64), ("pad_id", i64),
        ("dropout_p", C.c_float), ("_pad1", i32), ("seed", u64), ("offset", u64),
        ("w_ih", C.c_void_p * 2), ("w_hh", C.c_void_p * 2), ("b_ih", C.c_void_p * 2), ("b_hh", C.c_void_p * 2),
        ("lengths", C.c_void_p), ("out", C.c_void_p), ("h_n", C.c_void_p), ("gates", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", u64),
    ]


class GruLayerGrad(C.Structure):
    _fields_ = [
        ("fwd", GruLayer), ("d_out", C.c_void_p), ("d_h_n", C.c_void_p), ("d_x", C.c_void_p),
        ("d_w_ih", C.c_void_p * 2), ("d_w_hh", C.c_void_p * 2), ("d_b_ih", C.c_void_p * 2), ("d_b_hh", C.c_void_p * 2),
        ("workspace", C.c_void_p), ("workspace_bytes", u64),
    ]


class GruStack(C.Structure):
    _fields_ = [
        ("B", i32), ("T", i32), ("In", i32), ("H", i32), ("L", i32), ("ndir", i32),
        ("save_for_backward", i32), ("ids_are_i64", i32),
        ("x", C.c_void_p), ("ids", C.c_void_p), ("table", C.c_void_p), ("vocab", i64), ("pad_id", i64),
        ("dropout_p", C.c_float), ("_pad0", i32), ("seed", u64),
        ("weights", C.POINTER(C.c_void_p)), ("lengths", C.c_void_p), ("out", C.c_void_p), ("h_n", C.c_void_p),
        ("saved", C.c_void_p), ("saved_bytes", u64), ("workspace", C.c_void_p), ("workspace_bytes", u64),
        ("seed_dev", C.c_void_p),
    ]


class GruStackGrad(C.Structure):
    _fields_ = [
        ("fwd", GruStack), ("d_out", C.c_void_p), ("d_h_n", C.c_void_p), ("d_x", C.c_void_p),
        ("d_weights", C.POINTER(C.c_void_p)), ("workspace", C.c_void_p), ("workspace_bytes", u64),
    ]


class Head(C.Structure):
    _fields_ = [
        ("B", i32), ("In", i32), ("P1", i32), ("D2", i32), ("normalize", i32), ("dropout_p", C.c_float),
        ("seed", u64), ("offset", u64),
        ("v_lo", C.c_void_p), ("v_hi", C.c_void_p),
        ("w1", C.c_void_p), ("b1", C.c_void_p), ("ln_w", C.c_void_p), ("ln_b", C.c_void_p),
        ("w2", C.c_void_p), ("b2", C.c_void_p),
        ("y1", C.c_void_p), ("stats", C.c_void_p), ("y2", C.c_void_p), ("y3", C.c_void_p),
        ("out", C.c_void_p), ("norm", C.c_void_p), ("seed_dev", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", u64),
    ]


class HeadGrad(C.Structure):
    _fields_ = [
        ("fwd", Head), ("d_out", C.c_void_p), ("d_v_lo", C.c_void_p), ("d_v_hi", C.c_void_p),
        ("d_w1", C.c_void_p), ("d_b1", C.c_void_p), ("d_ln_w", C.c_void_p), ("d_ln_b", C.c_void_p),
        ("d_w2", C.c_void_p), ("d_b2", C.c_void_p), ("workspace", C.c_void_p), ("workspace_bytes", u64),
    ]


class ProfGemmRow(C.Structure):
    _fields_ = [("M", i32), ("N", i32), ("K", i32), ("nbatch", i32), ("npass", i32), ("flags", i32),
                ("total_ms", C.c_double), ("launches", u64)]


P = C.c_void_p
F = C.c_float
# name -> (restype, argtypes).  Must list EVERY symbol include/tt_b200.h declares (tests check this).
SIGNATURES = {
    "tt_abi_version": (C.c_int, []),
    "tt_last_error": (C.c_char_p, []),
    "tt_engine": (C.c_char_p, []),
    "tt_kernel_launches": (u64, []),
    "tt_prof_enable": (C.c_int, [i32]),
    "tt_prof_read": (C.c_int, [i32, C.POINTER(C.c_double), C.POINTER(u64)]),
    "tt_prof_family": (C.c_char_p, [i32]),
    "tt_prof_gemm_rows": (C.c_int, [C.POINTER(ProfGemmRow), i32]),
    "tt_set_engine": (C.c_int, [i32]),
    "tt_embed_gather": (C.c_int, [P, C.c_int, i64, P, i64, i32, i64, P, P]),
    "tt_gru_layer_fwd_workspace": (u64, [C.POINTER(GruLayer)]),
    "tt_gru_layer_fwd": (C.c_int, [C.POINTER(GruLayer), P]),
    "tt_gru_layer_bwd_workspace": (u64, [C.POINTER(GruLayerGrad)]),
    "tt_gru_layer_bwd": (C.c_int, [C.POINTER(GruLayerGrad), P]),
    "tt_gru_stack_saved_bytes": (u64, [C.POINTER(GruStack)]),
    "tt_gru_stack_fwd_workspace": (u64, [C.POINTER(GruStack)]),
    "tt_gru_stack_fwd": (C.c_int, [C.POINTER(GruStack), P]),
    "tt_gru_stack_bwd_workspace": (u64, [C.POINTER(GruStackGrad)]),
    "tt_gru_stack_bwd": (C.c_int, [C.POINTER(GruStackGrad), P]),
    "tt_head_fwd_workspace": (u64, [C.POINTER(Head)]),
    "tt_head_fwd": (C.c_int, [C.POINTER(Head), P]),
    "tt_head_bwd_workspace": (u64, [C.POINTER(HeadGrad)]),
    "tt_head_bwd": (C.c_int, [C.POINTER(HeadGrad), P]),
    "tt_l2norm_fwd": (C.c_int, [P, i32, i32, F, P, P, P]),
    "tt_l2norm_bwd": (C.c_int, [P, P, P, i32, i32, F, P, P]),
    "tt_score_ce_workspace": (u64, [i32, i32, i32]),
    "tt_score_ce_fwd": (C.c_int, [P, i32, P, i32, i32, i32, F, F, P, P, P, u64, P]),
    "tt_score_ce_bwd": (C.c_int, [P, i32, P, i32, i32, i32, F, F, P, F, P, P, P, u64, P]),
    "tt_score_ce_bwd_dev": (C.c_int, [P, i32, P, i32, i32, i32, F, F, P, F, P, P, P, P, u64, P]),
    "tt_hinge_bwd_dev": (C.c_int, [P, P, P, i32, i32, i32, F, F, P, P, P, P, P]),
    "tt_rows_topk_workspace": (u64, [i32, i64, i32, i32]),
    "tt_rows_topk": (C.c_int, [P, i32, P, i64, i32, i32, i32, i64, i64, P, P, P, u64, P]),
    "tt_corpus_index_bytes": (u64, [i64, i32]),
    "tt_corpus_index_build": (C.c_int, [P, i64, i32, i32, P, u64, P]),
    "tt_rows_topk_indexed_workspace": (u64, [i32, i64, i32, i32]),
    "tt_rows_topk_indexed": (C.c_int, [P, i32, P, i64, i32, i32, i32, i64, i64, P, P, P, u64, P, u64, P]),
    "tt_topk_merge": (C.c_int, [P, P, i32, i32, i32, P, P, P]),
    "tt_mrr_ranks": (C.c_int, [P, P, i32, i32, P, P, P]),
    "tt_hinge_fwd": (C.c_int, [P, P, P, i32, i32, i32, F, P, P]),
    "tt_hinge_bwd": (C.c_int, [P, P, P, i32, i32, i32, F, F, P, P, P, P]),
    "tt_gemm_nt": (C.c_int, [P, i64, P, i64, P, i64, P, i32, i32, i32, F, F, P]),
    "tt_dropout": (C.c_int, [P, P, i64, F, u64, u64, P]),
    "tt_gemm_x3_workspace": (u64, [i32, i32, i32]),
    "tt_gemm_x3": (C.c_int, [P, P, P, P, i32, i32, i32, F, F, i32, i32, P, u64, P]),
    "tt_debug_step_buffer": (C.c_int, [P]),
    "tt_debug_seq_mode": (C.c_int, [i32]),
    "tt_debug_bwd_seq_mode": (C.c_int, [i32]),
    "tt_debug_wgrad_overlap": (C.c_int, [i32]),
    "tt_debug_bwd_segments": (C.c_int, [i32]),
    "tt_debug_sel_resident": (C.c_int, [i32]),
    "tt_debug_sel_prefetch": (C.c_int, [i32]),
    "tt_debug_gemm_pairs": (C.c_int, [i32]),
    "tt_debug_sel_passes": (C.c_int, [i32]),
    "tt_debug_topk_fallback_flags": (i64, [i32, i64, i32, i32, i32, i32]),
    "tt_gemm_test": (C.c_int, [P, P, P, i32, i32, i32, F, F, i32, P, u64, P]),
    "tt_opt_step_workspace": (u64, [i32]),
    "tt_opt_step": (C.c_int, [P, P]),
    "tt_opt_step_noalloc": (C.c_int, [P, P]),
    "tt_vocab_create": (P, [P, P, i64, P]),
    "tt_vocab_destroy": (None, [P]),
    "tt_vocab_size": (i64, [P]),
    "tt_tokenize_batch": (C.c_int, [P, P, P, i64, i32, i64, i32, i32, P, P, i32]),
}

_lib = None


def lib():
    """Load (once) and return the CDLL.  Raises if the CUDA library is absent - no CPU path exists."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -m two_towers_b200.build` "
                "(two_towers_b200 has no CPU / eager fallback)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)        # AttributeError if the .so lacks a declared symbol
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {lib().tt_last_error().decode()}")


def ptr(t):
    """Device pointer of a contiguous fp32/int CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("two_towers_b200 kernels need CUDA tensors (no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("two_towers_b200 kernels need contiguous tensors")
    return t.data_ptr()


def stream(device=None):
    """Raw handle of torch's current stream ON `device` (None = the current device)."""
    return torch.cuda.current_stream(device).cuda_stream


def run(device, fn, what, *args):
    """Call a library entry point whose last argument is the stream, with `device` made current for the call (the C side
    launches on the current device and keeps per-device state) and torch's current stream of THAT device."""
    if device.type != "cuda":
        raise RuntimeError("two_towers_b200 kernels need CUDA tensors (no CPU fallback)")
    if torch.cuda.current_device() == device.index:
        check(fn(*args, torch.cuda.current_stream(device).cuda_stream), what)
    else:
        with torch.cuda.device(device):
            check(fn(*args, torch.cuda.current_stream(device).cuda_stream), what)


def f32c(t):
    """contiguous fp32 view/copy of a CUDA tensor."""
    if not t.is_cuda:
        raise RuntimeError("two_towers_b200 kernels need CUDA tensors (no CPU fallback)")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def workspace(nbytes, device):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
