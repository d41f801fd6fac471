"""ctypes binding of libgantron_b200.so (the C ABI declared in include/gantron_b200.h).

There is NO fallback: if the library is missing or a call fails, a RuntimeError is raised.
Tensors are passed as raw device pointers (`tensor.data_ptr()`), the stream as torch's current
CUDA stream handle.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgantron_b200.so")

_lib: Optional[C.CDLL] = None

c_p = C.c_void_p
c_i = C.c_int
c_ll = C.c_longlong
c_ull = C.c_ulonglong
c_f = C.c_float

# name -> (restype, argtypes).  Must list every symbol declared in include/gantron_b200.h
# (tests/test_abi.py cross-checks against the header).
SIGNATURES = {
    "gtn_last_error": (C.c_char_p, []),
    "gtn_version": (c_i, []),
    "gtn_launch_count": (c_ll, []),
    "gtn_reset_launch_count": (None, []),
    "gtn_device_info": (c_i, [C.POINTER(c_i)] * 4),
    "gtn_gemm_bf16": (c_i, [c_p, c_p, c_i, c_i, c_i, c_ll, c_ll, c_i, c_ll, c_ll,
                            c_p, c_i, c_ll, c_ll, c_i,
                            c_p, c_i, c_ll, c_ll, c_i,
                            c_p, c_p, c_i, c_i, c_f, c_ull, c_i, c_i, c_p, c_i, c_p]),
    "gtn_conv1d_bf16": (c_i, [c_p, c_i, c_i, c_i, c_i, c_i, c_i,
                              c_p, c_i, c_i, c_ll, c_ll,
                              c_p, c_i, c_ll, c_ll,
                              c_p, c_i, c_ll, c_ll,
                              c_p, c_i, c_f, c_ull, c_i, c_i, c_p, c_i, c_p]),
    "gtn_conv1d_wgrad_bf16": (c_i, [c_p, c_p, c_i, c_i, c_i, c_i, c_ll, c_ll, c_ll, c_ll,
                                    c_i, c_i, c_i, c_p, c_ll, c_i, c_i, c_p]),
    "gtn_gemm_tn_bf16": (c_i, [c_p, c_p, c_i, c_i, c_i, c_ll, c_ll, c_i, c_ll, c_ll,
                               c_p, c_i, c_ll, c_ll, c_i, c_p, c_p, c_i, c_i, c_p]),
    "gtn_bn_finalize": (c_i, [c_p, c_i, c_ll, c_f, c_f, c_p, c_p, c_p, c_p, c_p]),
    "gtn_bn_act_fwd": (c_i, [c_p, c_i, c_ll, c_ll, c_i, c_p, c_p, c_p, c_p, c_i, c_f, c_ull, c_i, c_p, c_i, c_ll, c_p]),
    "gtn_bn_act_bwd": (c_i, [c_p, c_i, c_ll, c_p, c_i, c_ll, c_ll, c_i, c_p, c_p, c_p, c_p, c_i, c_f, c_ull, c_i,
                             c_p, c_p, c_p, c_i, c_ll, c_p, c_p]),
    "gtn_bn_act_bwd_eval": (c_i, [c_p, c_i, c_ll, c_p, c_i, c_ll, c_ll, c_i, c_p, c_p, c_p, c_p, c_i, c_f, c_ull, c_i,
                                  c_p, c_p, c_p, c_i, c_ll, c_p, c_p]),
    "gtn_drop_tanh_bwd": (c_i, [c_p, c_i, c_ll, c_p, c_i, c_ll, c_ll, c_i, c_f, c_ull, c_i, c_p, c_i, c_ll, c_p, c_p]),
    "gtn_disc_head_fwd": (c_i, [c_p, c_i, c_ll, c_ll, c_i, c_p, c_p, c_p, c_p]),
    "gtn_disc_head_bwd": (c_i, [c_p, c_p, c_i, c_ll, c_ll, c_i, c_p, c_p, c_i, c_ll, c_p, c_p, c_p]),
    "gtn_window_mean_fwd": (c_i, [c_p, c_i, c_i, c_p, c_i, c_p, c_p]),
    "gtn_window_mean_bwd": (c_i, [c_p, c_i, c_i, c_p, c_i, c_p, c_p]),
    "gtn_taco_loss_fwd": (c_i, [c_p, c_p, c_p, C.POINTER(c_ll), c_i, c_i, c_i, c_p, c_ll, c_p, c_ll, c_p, c_p]),
    "gtn_taco_loss_bwd": (c_i, [c_p, c_p, c_p, C.POINTER(c_ll), c_i, c_i, c_i, c_p, c_ll, c_p, c_ll, c_p, c_p,
                                c_p, c_p, c_p, c_p]),
    "gtn_attn_guide_loss_fwd": (c_i, [c_p, c_i, c_i, c_i, c_p, c_p, c_f, c_p, c_p]),
    "gtn_attn_guide_loss_bwd": (c_i, [c_p, c_i, c_i, c_i, c_p, c_p, c_f, c_p, c_p, c_p]),
    "gtn_decoder_forward2": (c_i, [c_p, c_p]),
    "gtn_sizeof": (c_i, [c_i]),
    "gtn_bilstm_forward": (c_i, [c_p, c_p]),
    "gtn_bilstm_backward": (c_i, [c_p, c_p]),
    "gtn_decoder_infer1_ctas": (c_i, []),
    "gtn_decoder_infer1": (c_i, [c_p, c_p]),
    "gtn_decoder_infer1_static": (c_i, [c_p, c_p]),
    "gtn_decoder_bwd2_grid": (c_i, [c_i, c_i, c_i, c_i]),
    "gtn_decoder_backward2": (c_i, [c_p, c_p]),
}


c_u = C.c_uint


class DecoderFwdParams(C.Structure):
    """Mirror of `gtn_decoder_fwd_params` (include/gantron_b200.h) -- field order and types must match."""
    _fields_ = (
        [(n, c_i) for n in ("B", "Tin", "ldT", "n_steps", "E", "Cc", "D", "Dp", "Ha", "Hap", "Hd", "Hdp",
                            "P", "Pp", "A", "F", "KS", "NM", "NMp", "prec", "mult", "train", "use_dropout",
                            "mem_bf16", "ignore_stop")]
        + [(n, c_f) for n in ("p_att", "p_dec", "gate_threshold")]
        + [("thr_att", c_u), ("thr_dec", c_u), ("seed", c_ull)]
        + [(n, c_p) for n in ("w_att", "w_dec", "w_proj", "w_pre1", "w_pre2", "b_att", "b_dec", "b_proj",
                              "wq", "v", "wconv", "wdense", "mem_enc", "pm", "mem_const", "lengths", "pre_all",
                              "xh_a", "xh_d", "xctx", "xmel", "xp1", "xp2", "c_a", "c_d", "w_prev", "w_cum",
                              "qpart", "barrier", "done", "lengths_out", "stop", "steps_done",
                              "align_out", "mel_out", "gate_out",
                              "gates_a", "c_a_all", "h_a_all", "gates_d", "c_d_all", "h_d_all", "ctx_all",
                              "q_all", "th_all", "f_all", "cum_all",
                              "w_att2", "w_dec2", "u_loc", "wq_t", "prof", "gbias_a", "gbias_d", "pbias")]
        + [("Ep", c_i), ("reserved0", c_i)]
    )


class BiLstmParams(C.Structure):
    """Mirror of `gtn_bilstm_params` (include/gantron_b200.h)."""
    _fields_ = ([(n, c_i) for n in ("B", "T", "H", "reserved0")]
                + [(n, c_p) for n in ("xproj", "whh", "lengths", "out", "gates", "c_all", "dout", "dgates")])


class DecoderInfer1Params(C.Structure):
    """Mirror of `gtn_decoder_infer1_params` (include/gantron_b200.h)."""
    _fields_ = (
        [(n, c_i) for n in ("Tin", "Tp", "n_steps", "Ha", "Hd", "P", "A", "KS", "NM", "NMp", "nu", "ncta", "cols_a", "cols_d",
                            "ignore_stop", "reserved0")]
        + [("gate_threshold", c_f), ("reserved1", c_f), ("seed", c_ull)]
        + [(n, c_p) for n in ("wimg", "w1", "w2", "u", "wq", "v", "wp_h", "pc", "pbias", "b_att", "gbias_a", "gbias_d", "pm",
                              "h_att_g", "h_dec_g", "e_g", "qacc", "macc", "barrier", "mel_out", "gate_out", "align_out",
                              "steps_done", "done", "lengths_out", "prof")]
    )


class DecoderBwd2Params(C.Structure):
    """Mirror of `gtn_decoder_bwd2_params` (include/gantron_b200.h)."""
    _fields_ = (
        [(n, c_i) for n in ("B", "Tin", "ldT", "T", "E", "Ep", "D", "Ha", "Hd", "A", "KS", "prec", "mult",
                            "use_dropout", "mem_bf16", "grid")]
        + [("p_att", c_f), ("p_dec", c_f), ("thr_att", c_u), ("thr_dec", c_u), ("seed", c_ull)]
        + [(n, c_p) for n in ("wT_att2", "wT_dec2", "ut_loc", "wq_au", "v", "mem_enc", "lengths",
                              "gates_a", "c_a_all", "gates_d", "c_d_all", "align", "th_all", "cum_all",
                              "dproj_all", "dalign", "xdg_a", "xdg_d", "carry_ha", "dha_dec", "dc_a", "dc_d", "dhd_rec",
                              "dctx_att", "dctx_dec", "dwprev", "dcum", "barrier",
                              "dga_all", "dgd_all", "dq_all", "dctx_all", "dpm", "dv_part", "dU_part", "prof")]
        + [("flags", c_i)]
    )


def load() -> C.CDLL:
    """Load the shared library (raises if it has not been built)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m gantron_b200.build` "
            "(gantron_b200 has no CPU / eager fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().gtn_last_error().decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed (rc={rc}): {last_error()}")


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    if t is None:
        return None
    assert t.is_cuda, "gantron_b200 kernels take CUDA tensors only"
    return t.data_ptr()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def launch_count() -> int:
    return int(load().gtn_launch_count())


def reset_launch_count() -> None:
    load().gtn_reset_launch_count()


def device_info() -> dict:
    v = [c_i(0) for _ in range(4)]
    check(load().gtn_device_info(*[C.byref(x) for x in v]), "gtn_device_info")
    return dict(sm_count=v[0].value, cc_major=v[1].value, cc_minor=v[2].value, coop=v[3].value)
