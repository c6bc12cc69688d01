"""ctypes binding of libgg_b200.so (the C ABI declared in include/gg_b200.h).

The library is built in-tree by ``__graft_entry__.build()`` / ``genogrove_b200/csrc/Makefile``.  There is no
fallback: if the shared object is missing this module raises, and every compute entry point of the
library itself fails with GG_ERR_CUDA when no B200 is usable.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GG_B200_LIB") or os.path.join(_HERE, "libgg_b200.so")  # override: A/B builds of tools/

NO_CHROM = 0xFFFFFFFF
NO_RANK = 0xFFFFFFFF
STATIC_MAX_QUERIES = 1048576  # GG_STATIC_MAX_QUERIES: per-call limit of gg_intersect_fused_device_static
GG_OK = 0
STATUS_NAMES = {
    0: "GG_OK",
    1: "GG_ERR_INVALID_ARG",
    2: "GG_ERR_CUDA",
    3: "GG_ERR_OOM",
    4: "GG_ERR_INVALID_INTERVAL",
    5: "GG_ERR_NOT_BUILT",
    6: "GG_ERR_UNSORTED",
    7: "GG_ERR_CAPACITY",
}
TIE_INPUT_ORDER, TIE_REVERSE_INPUT, TIE_EXPLICIT = 0, 1, 2

# every symbol include/gg_b200.h declares (tests check the library exports exactly these)
EXPORTS = (
    "gg_version", "gg_device_count", "gg_index_create", "gg_index_destroy", "gg_last_error",
    "gg_index_build", "gg_index_build_device", "gg_index_upload_sorted", "gg_index_update_chrom", "gg_index_get_stats", "gg_index_chrom_counts",
    "gg_intersect_batch", "gg_intersect_batch_u32", "gg_intersect_batch_compact", "gg_compact_count", "gg_compact_rank",
    "gg_compact_expand", "gg_intersect_count", "gg_intersect_emit",
    "gg_intersect_count_device", "gg_intersect_emit_device", "gg_intersect_fused_device", "gg_intersect_fused_device_static", "gg_last_batch_stats",
    "gg_host_alloc", "gg_host_free", "gg_set_option",
    "gg_flanking_batch", "gg_flanking_device", "gg_flanking_batch_stranded", "gg_flanking_stranded_device",
    "gg_index_set_strands", "gg_intersect_batch_stranded", "gg_intersect_stranded_device",
    "gg_group_create", "gg_group_unique_id", "gg_group_create_rank", "gg_group_destroy", "gg_group_last_error",
    "gg_group_world", "gg_group_index", "gg_shard_assign", "gg_group_build", "gg_group_owners", "gg_group_get_stats",
    "gg_group_intersect_batch", "gg_group_intersect_device",
)


class IndexStats(C.Structure):
    _fields_ = [
        ("n_intervals", C.c_uint64), ("n_chrom", C.c_uint32), ("coord_bytes", C.c_uint32),
        ("n_levels", C.c_uint32), ("fanout", C.c_uint32), ("staged_levels", C.c_uint32),
        ("reserved", C.c_uint32), ("device_bytes", C.c_uint64), ("sort_ms", C.c_double),
        ("levels_ms", C.c_double), ("total_ms", C.c_double), ("h2d_bytes", C.c_uint64),
    ]


class BatchStats(C.Structure):
    _fields_ = [
        ("n_queries", C.c_uint64), ("n_hits", C.c_uint64), ("n_chunks", C.c_uint32),
        ("kernel_launches", C.c_uint32), ("count_ms", C.c_double), ("scan_ms", C.c_double),
        ("emit_ms", C.c_double), ("total_ms", C.c_double),
    ]


class CompactChunk(C.Structure):
    _fields_ = [("q_begin", C.c_uint64), ("n_queries", C.c_uint64), ("n_hits", C.c_uint64), ("hit_base", C.c_uint64),
                ("delta_byte_offset", C.c_uint64), ("delta_bytes", C.c_uint32), ("reserved", C.c_uint32)]


class CompactResult(C.Structure):
    _fields_ = [("n_queries", C.c_uint64), ("n_hits", C.c_uint64), ("chunk_queries", C.c_uint64),
                ("queries", C.c_void_p), ("deltas", C.c_void_p), ("delta_bytes_total", C.c_uint64),
                ("chunks", C.POINTER(CompactChunk)), ("n_chunks", C.c_uint32), ("n_chrom", C.c_uint32),
                ("chrom_slot_base", C.c_void_p)]


_u32p = C.POINTER(C.c_uint32)
_u64p = C.POINTER(C.c_uint64)
_lib = None


class GGError(RuntimeError):
    def __init__(self, status: int, message: str = ""):
        self.status = status
        super().__init__(f"{STATUS_NAMES.get(status, status)}: {message}")


def lib():
    """Load libgg_b200.so; raises if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). genogrove_b200 has no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.gg_version.restype = C.c_int
    L.gg_device_count.restype = C.c_int
    L.gg_index_create.argtypes = [C.POINTER(vp), C.c_int]
    L.gg_index_destroy.argtypes = [vp]
    L.gg_index_destroy.restype = None
    L.gg_last_error.argtypes = [vp]
    L.gg_last_error.restype = C.c_char_p
    L.gg_index_build.argtypes = [vp, vp, vp, vp, vp, C.c_int, C.c_size_t, C.c_uint32, vp]
    L.gg_index_upload_sorted.argtypes = [vp, vp, vp, vp, C.c_size_t, C.c_uint32]
    L.gg_index_update_chrom.argtypes = [vp, C.c_uint32, vp, vp, C.c_size_t]
    L.gg_index_get_stats.argtypes = [vp, C.POINTER(IndexStats)]
    L.gg_index_chrom_counts.argtypes = [vp, vp]
    L.gg_intersect_batch.argtypes = [vp, vp, vp, vp, C.c_size_t, vp, C.POINTER(_u32p), _u64p]
    L.gg_intersect_batch_u32.argtypes = [vp, vp, vp, vp, C.c_size_t, vp, C.POINTER(_u32p), _u64p]
    L.gg_intersect_batch_compact.argtypes = [vp, vp, vp, vp, C.c_size_t, C.POINTER(CompactResult)]
    L.gg_compact_count.argtypes = [C.POINTER(CompactResult), C.c_uint64]
    L.gg_compact_count.restype = C.c_uint64
    L.gg_compact_rank.argtypes = [C.POINTER(CompactResult), vp, C.c_uint64, C.c_uint64]
    L.gg_compact_rank.restype = C.c_uint32
    L.gg_compact_expand.argtypes = [C.POINTER(CompactResult), vp, vp, vp, C.c_int]
    L.gg_index_build_device.argtypes = [vp, vp, vp, vp, vp, C.c_int, C.c_size_t, C.c_uint32, vp]
    L.gg_intersect_count.argtypes = [vp, vp, vp, vp, C.c_size_t, vp]
    L.gg_intersect_emit.argtypes = [vp, vp, C.c_uint64]
    L.gg_intersect_count_device.argtypes = [vp, vp, vp, vp, C.c_int, C.c_size_t, vp, _u64p, vp]
    L.gg_intersect_emit_device.argtypes = [vp, vp, vp, vp, C.c_int, C.c_size_t, vp, vp, vp]
    L.gg_intersect_fused_device.argtypes = [vp, vp, vp, vp, C.c_int, C.c_size_t, vp, vp, C.c_uint64, _u64p, vp]
    L.gg_intersect_fused_device_static.argtypes = [vp, vp, vp, vp, C.c_int, C.c_size_t, vp, vp, C.c_uint64, _u64p, vp]
    L.gg_last_batch_stats.argtypes = [vp, C.POINTER(BatchStats)]
    L.gg_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
    L.gg_host_free.argtypes = [vp]
    L.gg_host_free.restype = None
    L.gg_set_option.argtypes = [vp, C.c_char_p, C.c_int64]
    L.gg_index_set_strands.argtypes = [vp, vp, C.c_size_t]
    L.gg_intersect_batch_stranded.argtypes = [vp, vp, vp, vp, vp, C.c_size_t, vp, C.POINTER(_u32p), _u64p]
    L.gg_intersect_stranded_device.argtypes = [vp, vp, vp, vp, vp, C.c_int, C.c_size_t, vp, vp, C.c_uint64, _u64p, vp]
    L.gg_flanking_batch.argtypes = [vp, vp, vp, vp, C.c_size_t, vp, vp]
    L.gg_flanking_device.argtypes = [vp, vp, vp, vp, C.c_int, C.c_size_t, vp, vp, C.c_int, vp]
    L.gg_flanking_batch_stranded.argtypes = [vp, vp, vp, vp, vp, C.c_size_t, vp, vp]
    L.gg_flanking_stranded_device.argtypes = [vp, vp, vp, vp, vp, C.c_int, C.c_size_t, vp, vp, C.c_int, vp]
    L.gg_group_create.argtypes = [C.POINTER(vp), C.POINTER(C.c_int), C.c_int]
    L.gg_group_unique_id.argtypes = [vp]
    L.gg_group_create_rank.argtypes = [C.POINTER(vp), C.c_int, C.c_int, C.c_int, vp]
    L.gg_group_destroy.argtypes = [vp]
    L.gg_group_destroy.restype = None
    L.gg_group_last_error.argtypes = [vp]
    L.gg_group_last_error.restype = C.c_char_p
    L.gg_group_world.argtypes = [vp]
    L.gg_group_world.restype = C.c_int
    L.gg_group_index.argtypes = [vp, C.c_int]
    L.gg_group_index.restype = vp
    L.gg_shard_assign.argtypes = [vp, C.c_uint32, C.c_int, vp]
    L.gg_group_build.argtypes = [vp, C.c_int, vp, vp, vp, vp, C.c_int, C.c_size_t, C.c_uint32, vp, C.c_size_t, C.POINTER(C.c_size_t)]
    L.gg_group_owners.argtypes = [vp, vp]
    L.gg_group_get_stats.argtypes = [vp, vp]
    L.gg_group_intersect_batch.argtypes = [vp, vp, vp, vp, C.c_int, C.c_size_t, vp, C.POINTER(_u32p), _u64p, _u64p]
    L.gg_group_intersect_device.argtypes = [vp, vp, vp, vp, C.c_int, C.c_size_t, C.c_int, vp]
    _lib = L
    return L
