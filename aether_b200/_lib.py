"""ctypes binding of libaether_b200.so (the C ABI declared in include/aether_b200.h).

The library is the product path: there is no Python/ATen fallback.  If the shared object is
missing it is built in-tree with nvcc (aether_b200.build); if that fails, importing the ops
raises.  Every call checks the integer return code and raises RuntimeError with the library's
message.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_uint32, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "lib", "libaether_b200.so")

F_SAVE = 1
F_BF16 = 2
NUM_PARAMS = 47


class AetherDims(ctypes.Structure):
    _fields_ = [("num_dims", c_int32), ("hidden", c_int32), ("n_nodes", c_int64), ("n_edges", c_int64)]


class AetherGraphC(ctypes.Structure):
    _fields_ = [("perm", c_void_p), ("rowptr", c_void_p), ("send", c_void_p), ("recv", c_void_p),
                ("sperm", c_void_p), ("colptr", c_void_p)]


SEQ_MAX_TYPES = 4


class AetherSeqDims(ctypes.Structure):
    _fields_ = [("num_dims", c_int32), ("num_vars", c_int32), ("hidden", c_int32), ("n_types", c_int32),
                ("skip_first", c_int32), ("field_hidden", c_int32), ("rnn_hidden", c_int32), ("prior_hidden", c_int32),
                ("flags", c_uint32), ("batch", c_int64)]


class AetherSeqEncParams(ctypes.Structure):
    _fields_ = [("filter_w0", c_void_p), ("filter_b0", c_void_p), ("filter_w2", c_void_p), ("filter_b2", c_void_p),
                ("res1_w", c_void_p), ("res1_b", c_void_p), ("mlp3_w", c_void_p * 2), ("mlp3_b", c_void_p * 2),
                ("mlp3_bn_scale", c_void_p), ("mlp3_bn_shift", c_void_p), ("mlp4_w", c_void_p * 2), ("mlp4_b", c_void_p * 2),
                ("mlp4_bn_scale", c_void_p), ("mlp4_bn_shift", c_void_p), ("lstm_w_ih", c_void_p), ("lstm_w_hh", c_void_p),
                ("lstm_b_ih", c_void_p), ("lstm_b_hh", c_void_p), ("prior_w", c_void_p * 4), ("prior_b", c_void_p * 4),
                ("n_prior_layers", c_int32), ("pos_offset", c_int32)]


class AetherSeqParams(ctypes.Structure):
    _T = c_void_p * SEQ_MAX_TYPES
    _fields_ = [("msg_fc1_w", _T), ("msg_fc1_b", _T), ("msg_fc2_w", _T), ("msg_fc2_b", _T),
                ("hidden_r_w", c_void_p), ("hidden_i_w", c_void_p), ("hidden_h_w", c_void_p),
                ("present_r_w", c_void_p), ("present_r_b", c_void_p), ("present_i_w", c_void_p), ("present_i_b", c_void_p),
                ("present_n_w", c_void_p), ("present_n_b", c_void_p),
                ("out_w", c_void_p * 3), ("out_b", c_void_p * 3),
                ("present_msg_fc1_w", _T), ("present_msg_fc1_b", _T), ("present_msg_fc2_w", _T), ("present_msg_fc2_b", _T),
                ("input_r_w", c_void_p), ("input_r_b", c_void_p), ("input_i_w", c_void_p), ("input_i_b", c_void_p),
                ("input_n_w", c_void_p), ("input_n_b", c_void_p),
                ("rff_B", c_void_p), ("field_w", c_void_p * 3), ("field_b", c_void_p * 3)]


# name -> (restype, argtypes); mirrors include/aether_b200.h one to one
SIGNATURES = {
    "aether_abi_version": (c_int, []),
    "aether_error_string": (c_char_p, [c_int]),
    "aether_graph_scratch_bytes": (c_size_t, [c_int64, c_int64]),
    "aether_graph_prepare": (c_int, [c_void_p, c_void_p, c_int64, c_int64, POINTER(AetherGraphC), c_void_p, c_void_p,
                                     c_size_t, c_void_p]),
    "aether_graph_prepare_static": (c_int, [c_void_p, c_void_p, c_int64, c_int64, POINTER(AetherGraphC), c_void_p, c_void_p]),
    "aether_fc_edges": (c_int, [c_int64, c_int64, c_void_p, c_void_p, c_void_p]),
    "aether_knn_edges": (c_int, [c_void_p, c_int32, c_void_p, c_void_p, c_int32, c_int32, c_void_p, c_void_p, c_void_p]),
    "aether_caller_prep": (c_int, [c_void_p, c_void_p, c_void_p, c_int32, c_void_p, c_void_p, c_int64, c_int64,
                                   c_void_p, c_void_p, c_void_p]),
    "aether_workspace_bytes": (c_size_t, [POINTER(AetherDims), c_uint32]),
    "aether_backward_scratch_bytes": (c_size_t, [POINTER(AetherDims), c_uint32]),
    "aether_step_forward": (c_int, [POINTER(AetherDims), POINTER(c_void_p), POINTER(AetherGraphC), c_void_p, c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_uint32, c_void_p]),
    "aether_step_backward": (c_int, [POINTER(AetherDims), POINTER(c_void_p), POINTER(AetherGraphC), c_void_p, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p, POINTER(c_void_p), c_void_p, c_void_p,
                                     c_void_p, c_size_t, c_void_p, c_size_t, c_uint32, c_void_p]),
    "aether_rollout_step": (c_int, [POINTER(AetherDims), POINTER(c_void_p), POINTER(AetherGraphC), c_void_p, c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                    c_uint32, c_void_p]),
    "aether_small_rollout_supported": (c_int, [c_int32, c_int32, c_int32, c_int32, c_uint32]),
    "aether_small_rollout_scratch_bytes": (c_size_t, []),
    "aether_small_rollout": (c_int, [POINTER(AetherDims), POINTER(c_void_p), c_int32, c_void_p, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_int32, c_int32, c_int32, c_void_p, c_void_p, c_void_p, c_int32, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_size_t, c_uint32, c_void_p]),
    "aether_part_forward_phase": (c_int, [POINTER(AetherDims), POINTER(c_void_p), POINTER(AetherGraphC), c_void_p, c_void_p,
                                          c_void_p, c_void_p, c_int64, c_int64, c_int32, c_void_p, c_void_p, c_void_p, c_void_p,
                                          c_size_t, c_uint32, c_void_p]),
    "aether_peer_allgather": (c_int, [POINTER(c_void_p), c_int32, c_int32, c_size_t, c_size_t, c_size_t, c_void_p, c_void_p]),
    "aether_radius_graph_scratch_bytes": (c_size_t, [c_int64, c_int32, POINTER(c_int32)]),
    "aether_radius_graph_neighbors": (c_int, [c_void_p, c_int32, c_int32, c_int64, c_float, c_float, c_int32,
                                              POINTER(c_float), POINTER(c_int32), c_void_p, c_void_p, c_void_p, c_size_t,
                                              c_void_p]),
    "aether_radius_graph_fill": (c_int, [c_void_p, c_void_p, c_int64, c_int32, c_void_p, c_void_p, c_void_p]),
    "aether_launch_count": (ctypes.c_uint64, []),
    "aether_profile_enable": (c_int, [c_int]),
    "aether_profile_collect": (c_int, [POINTER(ctypes.c_double), POINTER(c_int64), c_int32]),
    "aether_dynfield_workspace_bytes": (c_size_t, [c_int32, c_int64, c_int32]),
    "aether_dynfield_scratch_bytes": (c_size_t, [c_int32, c_int64, c_int32]),
    "aether_dynfield_forward": (c_int, [c_int32, c_int64, c_int32, POINTER(c_void_p), c_void_p, c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_size_t, c_void_p]),
    "aether_dynfield_backward": (c_int, [c_int32, c_int64, c_int32, POINTER(c_void_p), c_void_p, c_void_p, POINTER(c_void_p),
                                         c_void_p, c_void_p, c_void_p, c_size_t, c_void_p, c_size_t, c_void_p]),
    "aether_seq_workspace_bytes": (c_size_t, [POINTER(AetherSeqDims)]),
    "aether_seq_predict_field": (c_int, [POINTER(AetherSeqDims), POINTER(AetherSeqParams), c_void_p, c_void_p, c_void_p,
                                         c_size_t, c_void_p]),
    "aether_seq_decoder_step": (c_int, [POINTER(AetherSeqDims), POINTER(AetherSeqParams), c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "aether_seq_rollout_step": (c_int, [POINTER(AetherSeqDims), POINTER(AetherSeqParams), c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "aether_seq_encoder_step": (c_int, [POINTER(AetherSeqDims), POINTER(AetherSeqEncParams), c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "aether_seq_sample_hard": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_void_p]),
    "aether_seq_localizer": (c_int, [POINTER(AetherSeqDims), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "aether_adamw_step": (c_int, [c_int32, POINTER(c_void_p), POINTER(c_void_p), POINTER(c_void_p), POINTER(c_void_p),
                                  POINTER(c_int64), c_void_p, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_double,
                                  ctypes.c_double, c_void_p]),
    "aether_dbg_tc_gemm": (c_int, [c_void_p, c_int64, c_int32, c_int64, c_void_p, c_int64, c_int32, c_int32, c_int32,
                                   c_int32, c_int32, c_int32, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int32,
                                   c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "aether_dbg_tc_wgrad_partial_floats": (c_size_t, [c_int32, c_int32]),
    "aether_dbg_tc_wgrad": (c_int, [c_void_p, c_int32, c_void_p, c_int32, c_int32, c_int64, c_void_p, c_int32, c_int32,
                                    c_void_p, c_void_p, c_size_t, c_void_p]),
    "aether_dbg_tc_chain": (c_int, [c_void_p, c_int64, c_int32, c_int64, c_void_p, c_int64, c_int32, c_int32, c_void_p,
                                    c_int32, c_int32, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int32,
                                    c_void_p, c_void_p, c_void_p, c_void_p]),
    "aether_dbg_tc_chain_ex": (c_int, [c_void_p, c_int64, c_int32, c_int64, c_void_p, c_int64, c_int32, c_int32, c_void_p,
                                       c_int32, c_int32, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int32,
                                       c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_void_p]),
    "aether_dbg_tc_chain_bwd": (c_int, [c_void_p, c_int64, c_int32, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p,
                                        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "aether_dbg_gather4_probe": (c_int, [c_void_p, c_int64, c_int32, c_int64, c_int32, c_void_p, c_int32, c_void_p, c_void_p]),
    "aether_debug_stage": (c_int64, [POINTER(AetherDims), c_void_p, c_uint32, c_char_p, c_void_p, c_int64, c_void_p]),
}

_lib = None


def load(build_if_missing: bool = True) -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        if not build_if_missing:
            raise RuntimeError("aether_b200: %s is missing; run `python -m aether_b200.build`" % SO_PATH)
        from . import build as _build
        _build.build()
    lib = ctypes.CDLL(SO_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError here means the .so does not match the header
        fn.restype = res
        fn.argtypes = args
    if lib.aether_abi_version() != 1:
        raise RuntimeError("aether_b200: ABI version mismatch")
    _lib = lib
    return lib


def check(code: int, what: str) -> None:
    if code != 0:
        msg = load().aether_error_string(int(code))
        raise RuntimeError("%s failed (%d): %s" % (what, code, msg.decode() if msg else "?"))


def ptr(t) -> c_void_p:
    """Device pointer of a tensor (None -> NULL)."""
    return c_void_p(0) if t is None else c_void_p(t.data_ptr())


def ptr_array(tensors):
    arr = (c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr
