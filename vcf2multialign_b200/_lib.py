"""ctypes binding of libv2m_cuda.so (include/v2m_cuda.h). No fallback: a missing library or device is an error."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# V2M_CUDA_LIB: another build of the same library (kernel experiments); it must exist all the same.
LIB_PATH = os.environ.get("V2M_CUDA_LIB") or os.path.join(_HERE, "libv2m_cuda.so")

PLOIDY_MAX = 0xFFFFFFFF


class V2MError(RuntimeError):
    pass


class GraphView(C.Structure):
    """struct v2m_graph_view (include/v2m_cuda.h)."""
    _fields_ = [
        ("node_count", C.c_uint64),
        ("edge_count", C.c_uint64),
        ("reference_positions", C.c_void_p),
        ("aligned_positions", C.c_void_p),
        ("alt_edge_targets", C.c_void_p),
        ("alt_edge_count_csum", C.c_void_p),
        ("label_offsets", C.c_void_p),
        ("label_chars", C.c_void_p),
        ("paths_by_edge_and_chrom_copy", C.c_void_p),
        ("paths_by_chrom_copy_and_edge", C.c_void_p),
        ("h_pad", C.c_uint64),
        ("e_pad", C.c_uint64),
        ("total_chromosome_copies", C.c_uint32),
        ("ref_seq", C.c_void_p),
        ("ref_len", C.c_uint64),
        ("paths_on_device", C.c_int),
    ]


ROW_CALLBACK = C.CFUNCTYPE(None, C.c_uint32, C.c_void_p)
SINK = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint64, C.c_void_p)
ROW_SINK = C.CFUNCTYPE(C.c_int, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p)

# name → (restype, argtypes); every symbol include/v2m_cuda.h declares.
_vp, _u64, _u32, _i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
SYMBOLS = {
    "v2m_cuda_last_error": (C.c_char_p, []),
    "v2m_cuda_init": (_i32, [_i32]),
    "v2m_cuda_device_count": (_i32, [_vp]),
    "v2m_cuda_launch_count": (_u64, []),
    "v2m_cuda_host_alloc": (_vp, [_u64]),
    "v2m_cuda_host_free": (None, [_vp]),
    "v2m_cuda_device_alloc": (_vp, [_u64]),
    "v2m_cuda_device_free": (None, [_vp]),
    "v2m_cuda_memcpy_d2h": (_i32, [_vp, _vp, _u64]),
    "v2m_cuda_memcpy_h2d": (_i32, [_vp, _vp, _u64]),
    "v2m_cuda_synchronize": (_i32, []),
    "v2m_cuda_event_create": (_vp, []),
    "v2m_cuda_event_destroy": (None, [_vp]),
    "v2m_cuda_event_record": (_i32, [_vp, _vp]),
    "v2m_cuda_event_elapsed_ms": (_i32, [_vp, _vp, _vp]),
    "v2m_cuda_transpose_bits": (_i32, [_vp, _u64, _u64, _vp]),
    "v2m_cuda_transpose_bits_device": (_i32, [_vp, _u64, _u64, _vp, _vp]),
    "v2m_cuda_graph_upload": (_i32, [_vp, _vp]),
    "v2m_cuda_graph_attach_reference": (_i32, [_vp, _vp, _u64]),
    "v2m_cuda_graph_free": (_i32, [_vp]),
    "v2m_cuda_graph_aligned_length": (_i32, [_vp, _vp]),
    "v2m_cuda_graph_download_paths_by_chrom_copy_and_edge": (_i32, [_vp, _vp]),
    "v2m_cuda_graph_download_path_lines": (_i32, [_vp, _u32, _u32, _vp]),
    "v2m_cuda_graph_download_template": (_i32, [_vp, _vp]),
    "v2m_cuda_plan_haplotypes": (_i32, [_vp, _vp, _u32, _vp]),
    "v2m_cuda_plan_founders": (_i32, [_vp, _vp, _u64, _vp, _u32, _vp]),
    "v2m_cuda_plan_resolve": (_i32, [_vp, _vp]),
    "v2m_cuda_plan_free": (_i32, [_vp]),
    "v2m_cuda_plan_row_count": (_i32, [_vp, _vp]),
    "v2m_cuda_plan_download_taken": (_i32, [_vp, _vp]),
    "v2m_cuda_plan_row_lengths": (_i32, [_vp, _i32, _vp]),
    "v2m_cuda_emit_rows_device": (_i32, [_vp, _u32, _u32, _u64, _u64, _vp, _u64, _vp]),
    "v2m_cuda_emit_rows": (_i32, [_vp, _u32, _u32, _u64, _u64, _vp, _u64]),
    "v2m_cuda_release_staging": (None, []),
    "v2m_cuda_row_pitch": (_u64, [_u64]),
    "v2m_cuda_emit_rows_unaligned": (_i32, [_vp, _u32, _u32, _vp, _vp]),
    "v2m_cuda_write_a2m": (_i32, [_vp, C.c_char_p, _i32, _i32, ROW_CALLBACK, _vp]),
    "v2m_cuda_stream_a2m": (_i32, [_vp, C.c_char_p, _i32, SINK, _vp, ROW_CALLBACK, _vp]),
    "v2m_cuda_stream_rows": (_i32, [_vp, _i32, ROW_SINK, _vp]),
    "v2m_cuda_segment_classes": (_i32, [_vp, _vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _u64]),
    "v2m_cuda_find_matchings": (_i32, [_vp, _vp, _u64, _u32, _i32, _vp, _vp]),
    "v2m_cuda_last_segment_classes_timing": (_i32, [_vp, _vp]),
    "v2m_cuda_find_cut_positions": (_i32, [_vp, _u64, _vp, _u64, _vp, _vp, _vp]),
    "v2m_cuda_last_cut_positions_timing": (_i32, [_vp, _vp]),
    "v2m_cuda_paths_from_genotypes": (_i32, [_vp, _u64, _u32, _vp, _vp, _vp, _vp, _u64, _u64, _u64, _vp, _vp, _u64, _vp]),
    "v2m_cuda_paths_from_genotypes_device": (_i32, [_vp, _u64, _u32, _vp, _vp, _vp, _vp, _u64, _u64, _u64, _vp, _vp, _u64, _vp]),
    "v2m_cuda_paths_from_genotypes_both": (_i32, [_vp, _u64, _u32, _vp, _vp, _vp, _vp, _u64, _u64, _u64, _vp, _vp, _vp, _u64, _vp]),
    "v2m_cuda_ingest_genotypes_no_alloc": (_i32, [_vp, _u64, _u32, _vp, _vp, _vp, _vp, _u64, _u64, _u64, _vp, _vp, _u64, _vp, _vp]),
    "v2m_cuda_time_ingest_genotypes": (_i32, [_vp, _u64, _u32, _vp, _vp, _u64, _u64, _vp, _i32, _vp]),
    "v2m_cuda_synth_paths": (_i32, [_vp, _u64, _u64, _u32, _u64, _u32, _vp, _vp, _vp]),
    "v2m_cuda_time_emit_rows": (_i32, [_vp, _u32, _u32, _vp, _u64, _i32, _vp]),
    "v2m_cuda_time_transpose_bits": (_i32, [_vp, _u64, _u64, _vp, _i32, _vp]),
    "v2m_cuda_time_d2h_copies": (_i32, [_vp, _vp, _u64, _u64, _i32, _vp]),
}

_lib = None


def lib():
    """Loads libv2m_cuda.so; raises if it has not been built (python __graft_entry__.py build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise V2MError(f"{LIB_PATH} is missing: build it with `make -C vcf2multialign_b200/csrc` (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(status: int):
    if status != 0:
        raise V2MError(f"libv2m_cuda status {status}: {lib().v2m_cuda_last_error().decode()}")


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)
