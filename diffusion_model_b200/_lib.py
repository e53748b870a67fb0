"""ctypes binding of libegnn_b200.so (the C ABI declared in include/egnn_b200.h).

There is no CPU fallback: if the library cannot be loaded, or a tensor is not a CUDA tensor, the call raises.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libegnn_b200.so")

PREC = {"fp32": 0, "bf16x3": 1, "bf16": 2}
NORM = {"call": 0, "graph": 1}


class NetDesc(C.Structure):
    _fields_ = [(k, C.c_int32) for k in ("L", "H", "Mh", "Ms", "Xh", "Hh")]


class GraphDesc(C.Structure):
    _fields_ = [("N", C.c_int32), ("E", C.c_int32), ("G", C.c_int32)] + [
        (k, C.c_void_p) for k in ("recv", "send", "row_ptr", "col_ptr", "col_edge", "node_ptr", "node_graph")]


class EgnnError(RuntimeError):
    pass


class UnsortedEdgeIndex(EgnnError):
    """edge_index[0] is not non-decreasing (found by the deferred validation of a training minibatch; direct calls sort)."""


_P, _I32, _I64, _F = C.c_void_p, C.c_int32, C.c_int64, C.c_float
_NET, _GR = C.POINTER(NetDesc), C.POINTER(GraphDesc)

# name -> (restype, argtypes); must list every symbol include/egnn_b200.h declares (tests check this)
SIGNATURES = {
    "egnn_b200_last_error": (C.c_char_p, []),
    "egnn_b200_version": (C.c_int, []),
    "egnn_launch_count": (_I64, []),
    "egnn_profile_enable": (C.c_int, [_I32]),
    "egnn_profile_read": (C.c_int, [_I32, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "egnn_tc_selftest": (C.c_int, [_P, _P, _P, _I32, _I32, _I32, _I32, _P, _P]),
    "egnn_param_count": (_I64, [_NET]),
    "egnn_packed_count": (_I64, [_NET]),
    "egnn_forward_workspace_floats": (_I64, [_NET, _I32, _I32, _I32]),
    "egnn_backward_workspace_floats": (_I64, [_NET, _I32, _I32, _I32]),
    "egnn_edge_tile": (_I32, [_NET]),
    "egnn_build_edges_fc": (C.c_int, [_P, _I32, _I32, _I64, _P, _P]),
    "egnn_cutoff_count": (C.c_int, [_P, _P, _P, _P, _F, _I32, _P, _P]),
    "egnn_cutoff_fill": (C.c_int, [_P, _P, _P, _P, _F, _I32, _P, _I64, _P, _P]),
    "egnn_exclusive_scan_i32": (C.c_int, [_P, _I32, _P, _P]),
    "egnn_cutoff_fallback": (C.c_int, [_P, _P, _P, _F, _I32, _P, _P, _P]),
    "egnn_cutoff_fallback_fill": (C.c_int, [_P, _P, _I32, _I64, _P, _P]),
    "egnn_graph_prepare": (C.c_int, [_P, _I32, _I64, _P, _P, _P, _P, _P, _P, _P, _P]),
    "egnn_graph_prepare_noalloc": (C.c_int, [_P, _I32, _I64, _P, _P, _P, _P, _P, _P, _P]),
    "egnn_node_ptr_from_batch": (C.c_int, [_P, _I32, _I32, _P, _P, _P]),
    "egnn_pack_weights": (C.c_int, [_NET, _P, _P, _P]),
    "egnn_net_forward": (C.c_int, [_NET, _GR, _P, _P, _P, _P, _P, _P, _P, _I64, _I32, _I32, _P]),
    "egnn_net_backward": (C.c_int, [_NET, _GR, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I64, _I32, _I32, _P]),
    "egnn_net_backward_events": (C.c_int, [_NET, _GR, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I64, _I32, _I32, _P, _P]),
    "egnn_remove_mean": (C.c_int, [_P, _I32, _P, _I32, _P]),
    "egnn_diffuse_batch": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _I32, _P, _I32, _I32, _I32, _P, _P, _P, _P]),
    "egnn_loss_fwd_bwd": (C.c_int, [_P, _P, _P, _P, _P, _P, _I32, _I32, _I32, _P, _P, _P, _P]),
    "egnn_sampler_step": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _I32, _I64, _I64, _I32, _F, _P, _I32, _I32, _I32,
                                    _I32, _P, _P, _P]),
    "egnn_counter_add": (C.c_int, [_P, _I32, _P]),
    "egnn_adam_step": (C.c_int, [_P, _P, _P, _P, _I64, _F, _F, _F, _F, _F, _I32, _F, _P]),
    "egnn_adam_step_dev": (C.c_int, [_P, _P, _P, _P, _I64, _F, _F, _F, _F, _F, _P, _F, _P]),
    "egnn_radam_sf_step": (C.c_int, [_P, _P, _P, _P, _I64, _F, _F, _F, _F, _F, _F, _F, _I32, _F, _P]),
    "egnn_lerp": (C.c_int, [_P, _P, _I64, _F, _P]),
    "egnn_rdf_local": (C.c_int, [_P, _P, _I32, _P, _P, _P, _I32, C.c_double, C.c_double, _P, _I32, _P, _P, _P, _P]),
    "egnn_cn2_sites": (C.c_int, [_P, _P, _I32, _F, _P, _P, _P, _P, _P]),
    "egnn_kabsch": (C.c_int, [_P, _P, _P, _I32, _P, _P, _P, _P]),
}

_lib = None


def load():
    """Load the shared library (once).  Raises if it is missing: the product has no other path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise EgnnError(
            f"{LIB_PATH} not found: build it with `python -m diffusion_model_b200.build` "
            "(or __graft_entry__.build()).  There is no CPU / PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here == header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().egnn_b200_last_error().decode("utf-8", "replace")
        raise EgnnError(f"{what or 'libegnn_b200'} failed (status {rc}): {msg}")


_arg_device = [None]  # CUDA device index of the last tensor handed to ptr(): the device the next call must run on


def call(name, *args):
    """Invoke a C-ABI entry point and raise on a non-zero status.  The library launches on the CURRENT device, so when
    the tensors of this call (seen by ptr() while the arguments were built) live on another GPU of the process, that
    device is made current for the duration of the call -- per-device kernel attributes and the static scratch address
    of egnn_graph_prepare_noalloc are resolved per device inside the library."""
    fn = getattr(load(), name)
    dev, _arg_device[0] = _arg_device[0], None
    if dev is not None and dev != torch.cuda.current_device():
        with torch.cuda.device(dev):
            rc = fn(*args)
    else:
        rc = fn(*args)
    check(rc, name)


def stream_ptr(device=None):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t, dtype=None, allow_none=False):
    """Device pointer of a contiguous CUDA tensor (raises otherwise)."""
    if t is None:
        if allow_none:
            return C.c_void_p(0)
        raise EgnnError("NULL tensor passed to libegnn_b200")
    if not t.is_cuda:
        raise EgnnError("libegnn_b200 needs CUDA tensors (got a CPU tensor); there is no CPU fallback")
    if not t.is_contiguous():
        raise EgnnError("libegnn_b200 needs contiguous tensors")
    if dtype is not None and t.dtype != dtype:
        raise EgnnError(f"expected dtype {dtype}, got {t.dtype}")
    _arg_device[0] = t.device.index  # call() makes this device current (ptr() may also be used for raw ctypes calls)
    return C.c_void_p(t.data_ptr())
