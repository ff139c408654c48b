"""ctypes binding of libsrb200.so (include/srb200.h).  No CPU fallback: a missing library is an error."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "libsrb200.so")

c_void_p, c_int, c_i32, c_i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int32, ctypes.c_int64

# name -> (restype, argtypes); one entry per symbol declared in include/srb200.h
SIGNATURES = {
    "srb_version": (c_int, []),
    "srb_last_error_string": (ctypes.c_char_p, []),
    "srb_device_info": (c_int, [ctypes.POINTER(c_int)] * 4),
    "srb_packed_ab_floats": (c_i64, [c_int, c_int, c_int]),
    "srb_packed_c_floats": (c_i64, [c_int, c_int]),
    "srb_pack_model": (c_int, [c_void_p] * 4 + [c_int] * 3 + [c_void_p] * 4),
    "srb_gmm_score": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_i64,
                              c_void_p, c_void_p, c_void_p, c_void_p]),
    "srb_gmm_score_rows": (c_int, [c_void_p, c_i64, c_i64, c_i64, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                   c_void_p, c_i64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "srb_classify": (c_int, [c_void_p, c_void_p, c_i32, c_i64, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int,
                             c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "srb_tc_model_floats": (c_i64, [c_int, c_int, c_int]),
    "srb_pack_model_tc": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "srb_gmm_score_tc_supported": (c_int, [c_int, c_int, c_int]),
    "srb_gmm_score_tc": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_void_p, c_int, c_int, c_void_p, c_i32, c_void_p,
                                 c_i64, c_void_p]),
    "srb_merge_best": (c_int, [c_void_p, c_i64, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "srb_gmm_best_gemm_supported": (c_int, [c_int, c_int, c_int]),
    "srb_model_blocks_floats": (c_i64, [c_int, c_int, c_int]),
    "srb_pack_model_blocks": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "srb_gmm_best_gemm": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "srb_model_blocks_floats_sb": (c_i64, [c_int, c_int, c_int, c_int]),
    "srb_pack_model_blocks_sb": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "srb_gmm_score_gemm": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_i64,
                                   c_void_p, c_i64, c_void_p]),
    "srb_neg_lse_groups": (c_int, [c_void_p, c_i64, c_int, c_int, c_void_p, c_i64, c_void_p]),
    "srb_classify_tc_supported": (c_int, [c_int, c_int, c_int, c_int]),
    "srb_classify_tc": (c_int, [c_void_p, c_void_p, c_i32, c_i64, c_int, c_void_p, c_void_p, c_int, c_int,
                                c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "srb_classify_tc_timeline": (c_int, [c_void_p, c_void_p, c_i32, c_i64, c_int, c_void_p, c_void_p, c_int, c_int,
                                         c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "srb_viterbi_graph": (c_int, [c_void_p, c_i64, c_void_p, c_i32] + [c_void_p] * 11 + [c_i32] + [c_void_p] * 11),
    "srb_dtw_beam": (c_int, [c_void_p, c_i64, c_void_p, c_i32, c_void_p, c_i32, c_i32] + [c_void_p] * 8),
    "srb_viterbi_banded": (c_int, [c_void_p, c_i64, c_void_p, c_i32] + [c_void_p] * 16 + [c_i32] + [c_void_p] * 11),
    "srb_text_viterbi": (c_int, [c_void_p, c_void_p, c_i32] + [c_void_p] * 4 + [c_i32, c_i32, c_void_p, c_i32] +
                         [c_void_p] * 10),
    "srb_mfcc_num_frames": (c_i64, [c_i64, c_int]),
    "srb_mfcc": (c_int, [c_void_p, c_void_p, c_void_p, c_i32, c_i64, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                         c_void_p]),
    "srb_delta_standardize": (c_int, [c_void_p, c_void_p, c_i32, c_void_p, c_int, c_void_p]),
    "srb_path_words": (c_int, [c_void_p, c_void_p, c_void_p, c_i32, c_void_p, c_void_p, c_void_p, c_i32, c_void_p,
                               c_void_p, c_void_p]),
    "srb_fb_xi_parts": (c_int, []),
    "srb_forward_backward": (c_int, [c_void_p, c_i64, c_void_p, c_i32] + [c_void_p] * 12 + [c_i32, c_i32] +
                             [c_void_p] * 5 + [c_i32, c_void_p]),
    "srb_forward_backward_banded": (c_int, [c_void_p, c_i64, c_void_p, c_i32] + [c_void_p] * 9 + [c_i32] +
                                    [c_void_p] * 5 + [c_i32, c_void_p]),
    "srb_stats_row": (c_int, [c_int]),
    "srb_stats_parts": (c_int, []),
    "srb_group_stats": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "srb_pool_stats_parts": (c_int, []),
    "srb_pool_stats_partial_floats": (c_i64, [c_int, c_int, c_int]),
    "srb_pool_stats_supported": (c_int, [c_int, c_int, c_int]),
    "srb_pool_stats": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "srb_pool_stats_f64_supported": (c_int, [c_int, c_int, c_int]),
    "srb_pool_stats_f64_partial_doubles": (c_i64, [c_int, c_int, c_int]),
    "srb_pool_stats_f64": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "srb_fit_em_update": (c_int, [c_void_p, c_int, c_int, c_int, c_int] + [c_void_p] * 8 + [c_i64, c_void_p, c_void_p]),
    "srb_fit_kmeans_update": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_i64,
                                      c_void_p, c_void_p]),
    "srb_static_scratch_floats": (c_i64, []),
    "srb_group_stats_static": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "srb_gmm_stats": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                              c_void_p, c_void_p, c_int, c_void_p]),
    "srb_gmm_stats_soft": (c_int, [c_void_p, c_i64, c_i64, c_i64, c_int, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p,
                                   c_void_p, ctypes.c_double, c_void_p, c_void_p, c_int, c_void_p]),
    "srb_local_dist": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_i64,
                               c_void_p, c_void_p]),
    "srb_kmeans_assign": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "srb_mstep": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, ctypes.c_double,
                          c_void_p]),
    "srb_update_transitions": (c_int, [c_void_p] * 4 + [c_i32] + [c_void_p] * 4),
    "srb_gauss_consts": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "srb_em_tile_frames": (c_int, []),
    "srb_em_parts": (c_int, []),
    "srb_em_plan": (c_int, [c_void_p, c_void_p, c_i32, c_void_p, ctypes.POINTER(c_i32), ctypes.POINTER(c_i32)]),
    "srb_em_isolated": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_i32, c_void_p, c_i32, c_i32, c_void_p, c_void_p,
                                c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                                c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "srb_peer_alloc": (c_int, [c_i64, ctypes.POINTER(c_void_p)]),
    "srb_peer_free": (c_int, [c_void_p]),
    "srb_peer_export": (c_int, [c_void_p, c_void_p]),
    "srb_peer_open": (c_int, [c_void_p, ctypes.POINTER(c_void_p)]),
    "srb_peer_close": (c_int, [c_void_p]),
    "srb_em_sync_bytes": (c_i64, [c_int, c_i64]),
    "srb_em_push": (c_int, [c_void_p, c_i32, c_int, c_int, c_int, c_int, c_int, c_void_p, c_i32, c_void_p, c_int, c_int,
                            ctypes.c_uint64, c_void_p]),
    "srb_em_partial_floats": (c_i64, [c_int, c_int, c_int, c_int]),
    "srb_em_update": (c_int, [c_void_p, c_int, ctypes.c_uint64, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                              c_void_p, c_void_p, ctypes.c_double, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "srb_fma_peak": (c_int, [c_int, c_int, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_double), c_void_p]),
}


class SrbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libsrb200 error {code}: {msg}")
        self.code = code


_lib = None


def load():
    """Load the library (built by srb200/build.py).  Raises if it is absent -- there is no other path."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise ImportError(f"{SO_PATH} not built: run `python __graft_entry__.py build` (nvcc, sm_100a)")
        lib = ctypes.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc):
    if rc != 0:
        raise SrbError(rc, load().srb_last_error_string().decode())


def ptr(t):
    """device pointer of a torch tensor (or None)"""
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "libsrb200 takes contiguous CUDA tensors"
    return t.data_ptr()


def stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream
