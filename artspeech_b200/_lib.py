"""ctypes binding of libartspeech_b200.so (the C ABI in include/artspeech_b200.h).

There is no CPU fallback: if the library cannot be loaded the import of any compute
entry point raises, and every call needs CUDA tensors (or moves CPU tensors to the GPU).
"""
from __future__ import annotations

import ctypes
import os
import threading

import torch

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "libartspeech_b200.so")

ASB_OK, ASB_ERR_INVALID, ASB_ERR_UNSORTED, ASB_ERR_CUDA, ASB_ERR_WORKSPACE, ASB_ERR_UNSUPPORTED = range(6)
LAYOUT_PLANAR, LAYOUT_INTERLEAVED = 0, 1
CELL_GRU, CELL_LSTM = 3, 4
PRECISION_FAST, PRECISION_FP32 = 0, 1

_c_f32p = ctypes.c_void_p
_lock = threading.Lock()
_lib = None


class AsbWeights(ctypes.Structure):
    """Mirror of `struct asb_weights`."""
    _fields_ = (
        [("embedding", ctypes.c_void_p)]
        + [(n, (ctypes.c_void_p * 2) * 2) for n in ("rnn_w_ih", "rnn_w_hh", "rnn_b_ih", "rnn_b_hh")]
        + [("linear_w", ctypes.c_void_p), ("linear_b", ctypes.c_void_p)]
        + [(n, ctypes.POINTER(ctypes.c_void_p)) for n in (
            "ln0_w", "ln0_b", "fc1_w", "fc1_b", "ln1_w", "ln1_b", "fc2_w", "fc2_b",
            "ln2_w", "ln2_b", "xh_w", "xh_b", "yh_w", "yh_b")]
    )


class AsbPcWeights(ctypes.Structure):
    """Mirror of `struct asb_pc_weights`."""
    _fields_ = [(n, ctypes.c_void_p) for n in (
        "lin_w", "lin_b", "ln0_w", "ln0_b", "fc1_w", "fc1_b", "ln1_w", "ln1_b", "fc2_w", "fc2_b", "ln2_w", "ln2_b",
        "fc3_w", "fc3_b")]


class AsbEvalBuffers(ctypes.Structure):
    """Mirror of `struct asb_eval_buffers`."""
    _fields_ = [(n, ctypes.c_void_p) for n in (
        "outputs", "p2cp", "tv_pred", "tv_pred_idx", "tv_pred_poc", "tv_target", "tv_target_idx", "tv_target_poc",
        "scratch", "acc", "total", "workspace")] + [("workspace_bytes", ctypes.c_size_t)]


class AsbPredictorWeights(ctypes.Structure):
    """Mirror of `struct asb_predictor_weights`."""
    _fields_ = [(n, ctypes.c_void_p) for n in (
        "ln0_w", "ln0_b", "fc1_w", "fc1_b", "ln1_w", "ln1_b", "fc2_w", "fc2_b", "ln2_w", "ln2_b",
        "xh_w", "xh_b", "yh_w", "yh_b")]


# name -> (restype, argtypes); every symbol include/artspeech_b200.h declares
SIGNATURES = {
    "asb_abi_version": (ctypes.c_int, []),
    "asb_last_error_string": (ctypes.c_char_p, []),
    "asb_build_hash": (ctypes.c_char_p, []),
    "asb_launch_count": (ctypes.c_int64, []),
    "asb_p2cp": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                                ctypes.c_int, ctypes.c_void_p]),
    "asb_mean_euclidean": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, ctypes.c_int64, ctypes.c_int,
                                          ctypes.c_void_p]),
    "asb_point_distance": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p]),
    "asb_point_distance_backward": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, _c_f32p, ctypes.c_int64, ctypes.c_int,
                                                   ctypes.c_void_p]),
    "asb_pearson": (ctypes.c_int, [_c_f32p, _c_f32p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, _c_f32p,
                                   _c_f32p, ctypes.c_void_p]),
    "asb_min_distance": (ctypes.c_int, [_c_f32p, _c_f32p, ctypes.c_int64, ctypes.c_int, ctypes.c_int, _c_f32p,
                                        ctypes.c_void_p, ctypes.c_void_p]),
    "asb_tract_variables": (ctypes.c_int, [_c_f32p, ctypes.c_int64, ctypes.c_int,
                                           ctypes.POINTER(ctypes.c_int32), _c_f32p, ctypes.c_void_p, _c_f32p,
                                           ctypes.c_void_p]),
    "asb_tract_variables_ref": (ctypes.c_int, [_c_f32p, ctypes.c_int64, ctypes.c_int,
                                               ctypes.POINTER(ctypes.c_int32), _c_f32p, ctypes.c_int64, _c_f32p,
                                               ctypes.c_void_p, _c_f32p, ctypes.c_void_p]),
    "asb_tract_variables_ex": (ctypes.c_int, [_c_f32p, ctypes.c_int64, ctypes.c_int,
                                              ctypes.POINTER(ctypes.c_int32), _c_f32p, ctypes.c_int64, ctypes.c_int,
                                              _c_f32p, ctypes.c_void_p, _c_f32p, ctypes.c_void_p]),
    "asb_area_function": (ctypes.c_int, [_c_f32p, _c_f32p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p,
                                         ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_int, _c_f32p,
                                         ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    "asb_area_from_pairs": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int,
                                           ctypes.c_double, ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p,
                                           ctypes.c_void_p]),
    "asb_evenly_spaced_fx": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int,
                                            ctypes.c_int, _c_f32p, ctypes.c_void_p]),
    "asb_model_forward_resident": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                                  ctypes.POINTER(ctypes.c_int32), ctypes.c_void_p, ctypes.c_int,
                                                  _c_f32p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]),
    "asb_eval_accumulate": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                           ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    "asb_eval_accumulate_total": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                                 ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                                 ctypes.c_void_p]),
    "asb_eval_accumulate_static": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                                  ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]),
    "asb_eval_step": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int32),
                                     ctypes.c_void_p, ctypes.c_int, _c_f32p, ctypes.POINTER(ctypes.c_int32), _c_f32p,
                                     ctypes.c_int64, ctypes.POINTER(AsbEvalBuffers), ctypes.c_void_p,
                                     ctypes.c_void_p]),
    "asb_eval_sentence_metrics": (ctypes.c_int, [_c_f32p, _c_f32p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                                 ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                                 ctypes.c_void_p]),
    "asb_model_forward_hidden": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                                ctypes.POINTER(ctypes.c_int32), ctypes.c_void_p, ctypes.c_int,
                                                _c_f32p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]),
    "asb_pc_head_create": (ctypes.c_int, [ctypes.POINTER(AsbPcWeights), ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                          ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]),
    "asb_pc_head_destroy": (None, [ctypes.c_void_p]),
    "asb_pc_head_forward": (ctypes.c_int, [ctypes.c_void_p, _c_f32p, ctypes.c_int64, _c_f32p, ctypes.c_void_p]),
    "asb_predictor_create": (ctypes.c_int, [ctypes.POINTER(AsbPredictorWeights), ctypes.c_int, ctypes.c_int,
                                            ctypes.c_int, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]),
    "asb_predictor_destroy": (None, [ctypes.c_void_p]),
    "asb_predictor_forward": (ctypes.c_int, [ctypes.c_void_p, _c_f32p, ctypes.c_int64, _c_f32p, ctypes.c_void_p]),
    "asb_model_decode": (ctypes.c_int, [ctypes.c_void_p, _c_f32p, ctypes.c_int64, _c_f32p, ctypes.c_void_p]),
    "asb_gather_rows": (ctypes.c_int, [_c_f32p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                                       _c_f32p, ctypes.c_void_p]),
    "asb_gather_rows_flagged": (ctypes.c_int, [_c_f32p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                                               _c_f32p, ctypes.c_void_p, ctypes.c_void_p]),
    "asb_model_token_status": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int32),
                                              ctypes.POINTER(ctypes.c_int64), ctypes.c_int]),
    "asb_model_token_flag_ptr": (ctypes.c_void_p, [ctypes.c_void_p]),
    "asb_gemm": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int64,
                                ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, _c_f32p, ctypes.c_int,
                                ctypes.c_int, ctypes.c_void_p]),
    "asb_colsum": (ctypes.c_int, [_c_f32p, ctypes.c_int64, ctypes.c_int, ctypes.c_int64, _c_f32p, ctypes.c_int,
                                  ctypes.c_void_p]),
    "asb_layernorm_fwd": (ctypes.c_int, [_c_f32p, ctypes.c_int64, ctypes.c_int, _c_f32p, _c_f32p, _c_f32p, _c_f32p,
                                         _c_f32p, ctypes.c_void_p]),
    "asb_layernorm_bwd": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, _c_f32p, _c_f32p, ctypes.c_int64, ctypes.c_int,
                                         _c_f32p, ctypes.c_int, _c_f32p, _c_f32p, ctypes.c_int, ctypes.c_void_p]),
    "asb_sigmoid_bwd": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, ctypes.c_int64, ctypes.c_void_p]),
    "asb_relu_mask": (ctypes.c_int, [_c_f32p, _c_f32p, ctypes.c_int64, ctypes.c_void_p]),
    "asb_scatter_add_rows": (ctypes.c_int, [_c_f32p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                                            _c_f32p, ctypes.c_void_p]),
    "asb_dropout": (ctypes.c_int, [_c_f32p, _c_f32p, ctypes.c_int64, ctypes.c_float, ctypes.c_uint64, ctypes.c_void_p]),
    "asb_gru_train_forward": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, _c_f32p, _c_f32p, ctypes.c_void_p, ctypes.c_int,
                                             ctypes.c_int, ctypes.c_int, _c_f32p, _c_f32p, ctypes.c_void_p]),
    "asb_gru_train_backward": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, _c_f32p, _c_f32p, ctypes.c_void_p, ctypes.c_int,
                                              ctypes.c_int, ctypes.c_int, _c_f32p, _c_f32p, _c_f32p, ctypes.c_void_p]),
    "asb_lstm_train_forward": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, _c_f32p, _c_f32p, ctypes.c_void_p, ctypes.c_int,
                                              ctypes.c_int, ctypes.c_int, _c_f32p, _c_f32p, ctypes.c_void_p]),
    "asb_lstm_train_backward": (ctypes.c_int, [_c_f32p, _c_f32p, _c_f32p, _c_f32p, _c_f32p, ctypes.c_void_p, ctypes.c_int,
                                               ctypes.c_int, ctypes.c_int, _c_f32p, _c_f32p, ctypes.c_void_p]),
    "asb_model_set_decoder": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "asb_model_set_recurrent": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "asb_model_profile_enable": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "asb_model_profile_read": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_float),
                                              ctypes.POINTER(ctypes.c_int32), ctypes.c_int]),
    "asb_model_create": (ctypes.c_int, [ctypes.POINTER(AsbWeights), ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                        ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                        ctypes.POINTER(ctypes.c_void_p)]),
    "asb_model_destroy": (None, [ctypes.c_void_p]),
    "asb_model_workspace_bytes": (ctypes.c_size_t, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]),
    "asb_model_forward": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                         ctypes.POINTER(ctypes.c_int32), ctypes.c_int, _c_f32p, ctypes.c_void_p,
                                         ctypes.c_size_t, ctypes.c_void_p]),
}


def lib():
    """Load (once) and return the shared library; raise loudly if it is not there."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m artspeech_b200.build` "
                "(or __graft_entry__.build()). artspeech_b200 has no CPU fallback.")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (restype, argtypes) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the header and the library disagree
            fn.restype = restype
            fn.argtypes = argtypes
        if handle.asb_abi_version() != 1:
            raise RuntimeError("libartspeech_b200.so ABI version mismatch; rebuild the library")
        from . import build as _build
        if _build.sources():   # the sources travel with the library: a stale prebuilt .so is refused
            built, want = handle.asb_build_hash().decode(), _build.source_hash()
            if built != want:
                raise RuntimeError(f"{LIB_PATH} was built from other sources (hash {built}, sources {want}): "
                                   "rebuild it with `python -m artspeech_b200.build`")
        _lib = handle
    return _lib


def check(status):
    if status == ASB_OK:
        return
    msg = lib().asb_last_error_string().decode("utf-8", "replace")
    if status in (ASB_ERR_INVALID, ASB_ERR_WORKSPACE, ASB_ERR_UNSUPPORTED):
        raise ValueError(msg)
    raise RuntimeError(msg)


def launch_count():
    return int(lib().asb_launch_count())


def device_of(*tensors):
    """The CUDA device to run on: that of the first CUDA tensor, else the current device."""
    for t in tensors:
        if t.is_cuda:
            return t.device
    if not torch.cuda.is_available():
        raise RuntimeError("artspeech_b200 needs a CUDA device (B200); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
