"""ctypes binding of libfm_b200.so (the C ABI in include/fm_b200.h).

There is no fallback of any kind: if the CUDA library is missing or a call
fails, this module raises.  torch is used for device memory and streams only.
"""
import ctypes as C
import os

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "lib", "libfm_b200.so")

FM_E = {-1: "invalid argument", -2: "workspace too small", -3: "CUDA error", -4: "size limit"}

PARSE_ERROR_PREFIX = {
    1: "Label could not be read in example: ",                               # cc/fm_parser_op.cc:70
    2: "Invalid format in example: ",                                        # :77, :88
    3: "Invalid feature id. Should be in range [0, vocabulary_size).",       # :90-92
    4: "Invalid feature value. ",                                            # :100
    5: "fm_parser: output capacity too small for example: ",
}
NO_ERROR = 0xFFFFFFFFFFFFFFFF
LOSS_KIND = {"logistic": 0, "mse": 1}
OPTIMIZER = {"adagrad": 0, "sgd": 1}

_i64, _i32, _f32, _vp, _sz = C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_size_t

_SIGNATURES = {
    "fm_abi_version": (C.c_int, []),
    "fm_last_error": (C.c_char_p, []),
    "fm_split_lines_workspace": (_sz, [_i64]),
    "fm_split_lines": (C.c_int, [_vp, _i64, _vp, _i64, _vp, _vp, _sz, _vp]),
    "fm_parse_workspace": (_sz, [_i64]),
    "fm_parse_count": (C.c_int, [_vp, _i64, _vp, _vp, _i64, _i64, _i64, _i32, _vp, _vp, _vp, _vp, _sz, _vp]),
    "fm_parse_fill": (C.c_int, [_vp, _i64, _vp, _vp, _i64, _i64, _i64, _i32, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _sz, _i32,
                                _vp]),
    "fm_unique_workspace": (_sz, [_i64]),
    "fm_unique": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "fm_lookup": (C.c_int, [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _vp, _vp]),
    "fm_score_workspace": (_sz, [_i64]),
    "fm_score": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _i64, _i64, _f32, _f32, _vp, _vp, _vp, _sz, _vp]),
    "fm_score_noalloc": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _i64, _i64, _f32, _f32, _vp, _vp, _vp]),
    "fm_loss_workspace": (_sz, [_i64]),
    "fm_loss_grad": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _vp, _vp, _vp, _sz, _vp]),
    "fm_grad": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _i64, _i64, _f32, _f32, _vp, _f32, _vp, _i64, _vp, _vp, _vp]),
    "fm_sparse_adagrad": (C.c_int, [_vp, _vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _f32, _vp, _vp]),
    "fm_fused_gacc_ld": (_i64, [_i64]),
    "fm_fused_workspace": (_sz, [_i64]),
    "fm_fused_score_scatter": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _vp, _i64, _f32, _f32, _i32, _i64,
                                         _vp, _vp, _vp, _vp, _i64, _vp, _vp, _sz, _vp]),
    "fm_fused_adagrad_claim": (C.c_int, [_vp, _vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _f32, _f32, _f32, _f32, _vp,
                                         _i32, _vp]),
    "fm_fused_adagrad_bitmap": (C.c_int, [_vp, _vp, _i64, _vp, _i64, _i64, _vp, _i64, _f32, _f32, _f32, _f32, _vp, _i32,
                                          _vp, _vp]),
    "fm_peer_score_scatter": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _vp, _i64, _vp, _i64, _vp, _f32, _f32,
                                        _i32, _i64, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "fm_peer_sync_bytes": (_sz, []),
    "fm_peer_sync": (C.c_int, [_vp, _i64, _i64, C.c_uint64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "fm_merge_tile_rows": (_i64, [_i64]),
    "fm_merge_inbox_bytes": (_sz, [_i64, _i64, _i64, _i64]),
    "fm_merge_push_workspace": (_sz, [_i64]),
    "fm_peer_sorted_score": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _vp, _i64, _f32, _f32, _i32, _i64, _vp,
                                       _vp, _vp, _vp, _i64, _vp, _vp, _sz, _vp]),
    "fm_peer_push_records": (C.c_int, [_vp, _vp, _vp, _i64, _vp, _i64, _i64, _i64, _i64, _i64, _i64, _vp, _i64, _vp, _vp,
                                       _sz, _vp]),
    "fm_peer_merge_update": (C.c_int, [_vp, _vp, _i64, _i64, _vp, _i64, _i64, _i64, _f32, _f32, _f32, _f32, _vp, _vp]),
    "fm_peer_touched_workspace": (_sz, [_i64]),
    "fm_peer_touched_rows": (C.c_int, [_vp, _i64, _vp, _i64, _vp, _vp, _sz, _vp]),
    "fm_shard_bucket_workspace": (_sz, [_i64, _i64]),
    "fm_shard_bucket": (C.c_int, [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _sz, _vp]),
    "fm_shard_remap": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp]),
    "fm_shard_local_rows": (C.c_int, [_vp, _i64, _i64, _vp, _vp]),
    "fm_shard_combine": (C.c_int, [_vp, _i64, _vp, _i64, _i64, _vp, _i64, _vp]),
    "fm_sort_workspace": (_sz, [_i64]),
    "fm_sort_by_row": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _sz, _i32, _vp]),
    "fm_sorted_record_ld": (_i64, [_i64]),
    "fm_sorted_score": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _vp, _i64, _f32, _f32, _i32, _i64, _vp, _vp, _vp,
                                  _vp, _i64, _vp, _vp, _sz, _vp]),
    "fm_sorted_workspace": (_sz, [_i64, _i64]),
    "fm_sorted_update": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _vp, _i64, _vp, _i64, _i64, _i32, _f32, _f32, _f32,
                                   _f32, _i64, _vp, _vp, _sz, _vp]),
    "fm_fused_adagrad": (C.c_int, [_vp, _vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _f32, _f32, _f32, _f32, _vp, _vp]),
}

_lib = None


class FmB200Error(RuntimeError):
    pass


def load():
    """Loads the CUDA library or raises; never degrades to a CPU path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FmB200Error(
            "fast_tffm_b200: %s is missing. Build it with `python -m fast_tffm_b200.build` "
            "(needs nvcc); there is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype, fn.argtypes = res, args
    if lib.fm_abi_version() != 1:
        raise FmB200Error("libfm_b200.so ABI version %d != 1" % lib.fm_abi_version())
    _lib = lib
    return lib


def exported_symbols():
    return sorted(_SIGNATURES)


def check(rc, what):
    if rc != 0:
        msg = load().fm_last_error().decode("utf-8", "replace")
        raise FmB200Error("%s failed (%s): %s" % (what, FM_E.get(rc, rc), msg))


def ptr(t):
    """device pointer of a tensor (None -> NULL)"""
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise FmB200Error("fast_tffm_b200 ops run on CUDA tensors only (got a %s tensor); "
                              "there is no CPU fallback" % t.device)


def workspace(nbytes, device):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
