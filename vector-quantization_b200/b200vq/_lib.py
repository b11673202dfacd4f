"""ctypes binding of libb200vq.so (C ABI declared in include/b200vq.h).

The library is built in-tree by ``__graft_entry__.build()`` (or ``make -C
vector-quantization_b200/csrc``). There is no CPU fallback: importing works
without a GPU (so the symbol table can be checked), but every compute call goes
to the CUDA kernels and raises ``B200VQError`` on any non-zero status.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200VQ_LIB: A/B a second build of the same C ABI (tools/ only); the in-tree library is the product
LIB_PATH = os.environ.get("B200VQ_LIB") or os.path.join(_HERE, "libb200vq.so")

OK = 0
METRIC_L2, METRIC_IP = 0, 1


class B200VQError(RuntimeError):
    pass


_p, _i, _i64, _sz, _u64 = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_uint64

# name -> (restype, argtypes); mirrors include/b200vq.h one to one.
SIGNATURES = {
    "b200vq_last_error": (C.c_char_p, []),
    "b200vq_abi_version": (_i, []),
    "b200vq_device_cc": (_i, []),
    "b200vq_pq_centroid_norms": (_i, [_p, _i, _i, _i, _p, _p]),
    "b200vq_pq_assign_f32": (_i, [_p, _i64, _i64, _p, _p, _i, _i, _i, _p, _p, _p]),
    "b200vq_pq_assign_tc_scratch_bytes": (_sz, [_i64, _i]),
    "b200vq_pq_assign_tc_scratch_bytes2": (_sz, [_i64, _i, _i]),
    "b200vq_pq_assign_tc": (_i, [_p, _i64, _i64, _p, _p, _i, _i, _i, _p, _p, _sz, _p, _p]),
    "b200vq_pq_decode": (_i, [_p, _i64, _p, _i, _i, _i, _p, _i64, _p]),
    "b200vq_pq_pack_bits": (_i, [_p, _i64, _i, _i, _p, _p]),
    "b200vq_pq_unpack_bits": (_i, [_p, _i64, _i, _i, _p, _p]),
    "b200vq_pq_decode_ids": (_i, [_p, _p, _i64, _p, _i, _i, _i, _p, _i64, _p]),
    "b200vq_pq_kmeans_accumulate": (_i, [_p, _i64, _i64, _p, _i, _i, _i, _p, _p, _p]),
    "b200vq_pq_kmeans_finalize": (_i, [_p, _p, _i, _i, _i, _p, _p, _p]),
    "b200vq_pq_sse": (_i, [_p, _i64, _i64, _p, _p, _i, _i, _i, _p, _p]),
    "b200vq_adc_lut": (_i, [_p, _i, _i64, _p, _i, _i, _i, _i, _p, _p]),
    "b200vq_adc_scan_scratch_bytes": (_sz, [_i64, _i, _i]),
    "b200vq_adc_scan_topk": (_i, [_p, _i64, _i, _i, _p, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "b200vq_adc_scan_topk_static": (_i, [_p, _i64, _i, _i, _p, _i, _i, _i, _p, _p, _p]),
    "b200vq_pq_codes_to_scan_layout": (_i, [_p, _i64, _i, _p, _p]),
    "b200vq_adc_scan_topk_rot": (_i, [_p, _i64, _i, _i, _p, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "b200vq_adc_scan_screen_scratch_bytes": (_sz, [_i64, _i, _i, _i]),
    "b200vq_adc_scan_topk_screen": (_i, [_p, _i64, _i, _i, _p, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "b200vq_flat_topk_scratch_bytes": (_sz, [_i64, _i, _i]),
    "b200vq_flat_topk": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _i, _i64, _i, _i, _p, _p, _p, _sz, _p]),
    "b200vq_flat_topk_after": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _i, _i64, _i, _i, _p, _p, _p, _p, _p, _sz, _p]),
    "b200vq_row_argmin": (_i, [_p, _i64, _i64, _i, _p, _p, _p]),
    "b200vq_ivf_search_tc_scratch_bytes": (_sz, [_i64, _i, _i, _i]),
    "b200vq_ivf_search_tc": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _i, _p, _i, _i64, _p, _i, _i, _i, _p, _p, _p, _p, _p, _sz, _p]),
    "b200vq_ivf_search_scratch_bytes": (_sz, [_i, _i, _i, _i]),
    "b200vq_ivf_search": (_i, [_p, _i64, _i, _p, _p, _p, _i, _p, _i, _i64, _p, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "b200vq_ivf_search_after": (_i, [_p, _i64, _i, _p, _p, _p, _i, _p, _i, _i64, _p, _i, _i, _i, _p, _p, _p, _p, _p, _sz,
                                     _p]),
    "b200vq_flat_topk_tc_scratch_bytes": (_sz, [_i64, _i, _i]),
    "b200vq_flat_topk_tc": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _i, _i64, _i, _i, _p, _p, _p, _sz, _p]),
    "b200vq_rotate_tc_applies": (_i, [_i64, _i64, _i64, _i64, _i64]),
    "b200vq_rotate_tc_scratch_bytes": (_sz, [_i64, _i64]),
    "b200vq_rotate_tc_prepare": (_i, [_p, _i64, _i64, _i64, _p, _sz, _p]),
    "b200vq_rotate_tc": (_i, [_p, _i64, _i64, _p, _i64, _i64, _p, _i64, _p]),
    "b200vq_rotate_tc_ex": (_i, [_p, _i64, _i64, _p, _i64, _i64, _p, _p, _p, _i, _p, _i64, _p]),
    "b200vq_row_sqnorms": (_i, [_p, _i64, _i64, _i, _p, _p]),
    "b200vq_topk_merge": (_i, [_p, _p, _i, _i, _i, _i, _p, _p, _p]),
    "b200vq_sq_minmax": (_i, [_p, _i64, _i64, _i, _p, _p, _p]),
    "b200vq_sq_encode": (_i, [_p, _i64, _i64, _i, _p, _p, _i, _p, _p]),
    "b200vq_sq_decode": (_i, [_p, _i64, _i, _p, _p, _i, _p, _i64, _p]),
    "b200vq_saq_encode": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _p, _i, _p, _p]),
    "b200vq_saq_decode": (_i, [_p, _i64, _i, _p, _p, _p, _p, _i, _p, _i64, _p]),
    "b200vq_erabitq_pack": (_i, [_p, _p, _i64, _i, _i, _p, _p, _p, _p]),
    "b200vq_resid_norms_f64": (_i, [_p, _i, _i64, _i64, _i, _p, _p, _p]),
    "b200vq_erabitq_pack_tc": (_i, [_p, _i64, _p, _i, _i64, _p, _p, _p, _i64, _i, _i, _p, _p, C.c_double, C.c_double,
                                    _p, _p, _p]),
    "b200vq_erabitq_pack_tc2_scratch_bytes": (_sz, [_i64, _i]),
    "b200vq_erabitq_pack_tc2": (_i, [_p, _i64, _p, _i, _i64, _p, _p, _p, _i64, _i, _i, _p, _p, C.c_double, C.c_double,
                                     _p, _p, _p, _sz, _p]),
    "b200vq_erabitq_rowscale": (_i, [_p, _i64, _i, _i, _p, _p]),
    "b200vq_erabitq_unpack": (_i, [_p, _i64, _i, _i, _p, _p, _p, _p]),
    "b200vq_erabitq_scan_scratch_bytes": (_sz, [_i64, _i, _i]),
    "b200vq_erabitq_decode_f32": (_i, [_p, _i64, _i, _i, _p, _p, _i64, _p, _p]),
    "b200vq_erabitq_scan_topk": (_i, [_p, _i64, _i, _i, _p, _p, _p, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "b200vq_est_row_norms": (_i, [_i, _p, _p, _p, _p, _p, _p, _p, _i64, _p, _p, _sz, _p]),
    "b200vq_est_scan_scratch_bytes": (_sz, [_i64, _i, _i]),
    "b200vq_est_scan_topk": (_i, [_i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i64, _p, _i, _i64, _p, _i, _i, _p, _p,
                                  _p, _sz, _p]),
    "b200vq_perdim_encode": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _p, _i, _p, _p]),
    "b200vq_perdim_decode": (_i, [_p, _i64, _i, _i, _p, _p, _p, _p, _p, _i64, _p]),
    "b200vq_perdim_encode_tc": (_i, [_p, _i64, _p, _i, _i64, _p, _p, _p, _i64, _i, _p, _p, _p, _p, _i, C.c_double,
                                     C.c_double, _p, _p, _p]),
    "b200vq_perdim_decode_f32": (_i, [_p, _i64, _i, _i, _p, _p, _p, _p, _p, _i64, _p]),
    "b200vq_codebook_lloyd_scratch_bytes": (_sz, [_i64, _i]),
    "b200vq_codebook_lloyd": (_i, [_p, _i64, _p, _p, _i64, _i, _p, _i, _i, _u64, _i, C.c_float, _p, _p, _p, _p, _sz, _p]),
    "b200vq_perdim_scan_scratch_bytes": (_sz, [_i64, _i, _i]),
    "b200vq_perdim_scan_topk": (_i, [_p, _i64, _i, _i, _p, _p, _p, _p, _p, _p, _i, _i64, _p, _i, _i, _p, _p, _p, _sz, _p]),
    "b200vq_ffd_pack": (_i, [_p, _i64, _i, _p, _p, _i, _p, _p]),
    "b200vq_ffd_unpack": (_i, [_p, _i64, _i, _p, _p, _p, _i, _p, _p]),
    "b200vq_lvq_encode": (_i, [_p, _i64, _i64, _i, _p, _i, _p, _p]),
    "b200vq_lvq_decode": (_i, [_p, _i64, _i, _p, _i, _p, _i64, _p]),
    "b200vq_col_mean_seq": (_i, [_p, _i64, _i64, _i, _p, _p]),
    "b200vq_rabitq1_encode": (_i, [_p, _i64, _i64, _i, _p, _i, _p, _p]),
    "b200vq_rabitq1_decode": (_i, [_p, _i64, _i, _p, _p, _i64, _p]),
    "b200vq_rabitq1_prepare_queries": (_i, [_p, _i, _i64, _i, _p, _i, _p, _p, _p]),
    "b200vq_rabitq1_scan_scratch_bytes": (_sz, [_i64, _i, _i]),
    "b200vq_rabitq1_scan_topk": (_i, [_p, _i64, _i, _p, _p, _i, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "b200vq_rabitq1_scan_topk_after": (_i, [_p, _i64, _i, _p, _p, _i, _i, _i, _i, _p, _p, _p, _p, _p, _sz, _p]),
    "b200vq_rabitq1_scan_tc_scratch_bytes": (_sz, [_i64, _i, _i, _i]),
    "b200vq_rabitq1_scan_topk_tc": (_i, [_p, _i64, _i, _p, _p, _i, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "b200vq_saq_plan_dp": (_i, [_p, _i, C.c_float, _p, _p, _p]),
    "b200vq_caq_encode_t": (_i, [_p, _i64, _i64, _i, _p, _i, _i, C.c_float, _p, _i64, _p, _p, _p, _p, _p, _p]),
    "b200vq_caq_encode_t_segments": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _i, C.c_float, _p, _i64, _p, _p, _p, _p, _p, _p]),
    "b200vq_caq_pack_t": (_i, [_p, _i64, _i64, _i, _i, _p, _p]),
    "b200vq_caq_decode": (_i, [_p, _i64, _p, _i64, _i, _i, _p, _i, _p, _i64, _p]),
    "b200vq_synth_rows": (_i, [_p, _i64, _i64, _i, _i64, _u64, _p, _i, _p]),
    "b200vq_sq_err_sum": (_i, [_p, _i64, _p, _i64, _i64, _i, _p, _p]),
}

_lib = None


def load() -> C.CDLL:
    """Load the shared library (once) and attach the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C vector-quantization_b200/csrc`. b200vq has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the build is stale
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


# number of compute entry points called (bench.py's gpu_launches claim; every call launches
# at least one kernel of libb200vq.so)
abi_calls = 0


def check(status: int, what: str) -> None:
    global abi_calls
    abi_calls += 1
    if status != OK:
        msg = load().b200vq_last_error()
        raise B200VQError(f"{what} failed (status {status}): {msg.decode(errors='replace') if msg else ''}")
