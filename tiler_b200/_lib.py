"""ctypes loader for tiler_b200/libtmann.so (the C ABI in include/tmann.h).

The library is the product; this module only declares prototypes.  There is no
CPU fallback anywhere: if the shared object is missing the import fails loudly,
and every compute call fails with TM_ENODEVICE when no sm_100 GPU is present.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# TMANN_LIB: an alternative build of the same library (the debug-sweep build of scripts/dbg_sweep.sh), never a fallback
LIB_PATH = os.environ.get("TMANN_LIB") or os.path.join(_HERE, "libtmann.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `make -C tiler_b200/csrc` "
        "(or __graft_entry__.build()); tiler_b200 has no fallback path")

lib = C.CDLL(LIB_PATH)

vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_float


class TMStats(C.Structure):
    _fields_ = [("kernel_launches", i64), ("queries", i64), ("fallback_queries", i64),
                ("desc_ms", f32), ("search_ms", f32), ("coarse_ms", f32), ("rerank_ms", f32),
                ("coarse_launches", i64), ("max_error_ratio", f32), ("audited_candidates", i64)]


# name -> (restype, argtypes); must list every TM_API symbol of include/tmann.h
PROTOTYPES = {
    "ann_kdtree_create": (vp, [vp, i32, i32, i32, i32]),
    "ann_kdtree_destroy": (None, [vp]),
    "ann_kdtree_search": (i32, [vp, vp, f32, vp]),
    "ann_kdtree_pri_search": (i32, [vp, vp, f32, vp]),
    "ann_kdtree_search_multi": (i32, [vp, vp, vp, i32, vp, f32]),
    "tm_last_error": (C.c_char_p, []),
    "tm_version": (i32, []),
    "tm_set_device": (i32, [i32]),
    "tm_set_devices": (i32, [vp, i32]),
    "tm_get_devices": (i32, [vp, i32]),
    "tm_comm_init_ms": (C.c_double, []),
    "tm_dataset_replicas": (i32, [vp]),
    "tm_dataset_read_replica": (i32, [vp, i32, vp]),
    "tm_global_dataset_create_active": (vp, [vp, i64, vp, vp, vp]),
    "tm_set_engine": (i32, [i32]),
    "tm_set_stream": (i32, [vp]),
    "tm_device_sync": (i32, []),
    "tm_host_alloc": (vp, [C.c_size_t]),
    "tm_host_free": (None, [vp]),
    "tm_dataset_create": (vp, [vp, i64, i32, i32]),
    "tm_dataset_rows": (i64, [vp]),
    "tm_dataset_dim": (i32, [vp]),
    "ann_kdtree_search_batch": (i32, [vp, vp, i64, i32, vp, vp]),
    "tm_search_batch_dev": (i32, [vp, vp, i64, i32, vp, vp]),
    "tm_psyv_queries": (i32, [vp, i64, i32, i32, vp]),
    "tm_psyv_queries_dev": (i32, [vp, i64, i32, i32, vp]),
    "tm_tileset_dataset_create": (vp, [vp, i64, vp, i32, i32, vp, i32, i32]),
    "tm_tileset_dataset_create_dev": (vp, [vp, i64, vp, i32, i32, vp, i32, i32]),
    "tm_dataset_read": (i32, [vp, vp]),
    "tm_frametiling_search": (i32, [vp, vp, i64, i32, i32, vp, vp]),
    "tm_frametiling_search_dev": (i32, [vp, vp, i64, i32, i32, vp, vp]),
    "tm_global_dataset_create": (vp, [vp, i64, vp, vp]),
    "tm_keyframe_dataset_create": (vp, [vp, vp, i64, vp, vp, i32, i32, vp, vp, i64, i32, C.c_double, vp, i32, i32]),
    "tm_keyframe_tables": (i32, [vp, vp, vp, vp]),
    "tm_kmodes_assign": (i32, [vp, i64, vp, i64, i32, vp, vp]),
    "tm_kmodes_assign_dev": (i32, [vp, i64, vp, i64, i32, vp, vp]),
    "tm_kmodes_open": (vp, [vp, i64, i32]),
    "tm_kmodes_close": (None, [vp]),
    "tm_kmodes_set_centroids": (i32, [vp, vp, i64]),
    "tm_kmodes_update_centroids": (i32, [vp, vp, vp, i64]),
    "tm_kmodes_assign_range": (i32, [vp, i64, i64, vp, vp]),
    "tm_kmodes_read_centroids": (i32, [vp, vp]),
    "tm_kmodes_assign_noalloc": (i32, [vp, i64, vp, i64, i32, vp, vp, vp]),
    "tm_kmodes_update_min_distance": (i32, [vp, i64, vp, i32, vp]),
    "tm_kmodes_farthest_first": (i32, [vp, i64, i64, i64, i32, vp, vp]),
    "tm_kmodes_representatives": (i32, [vp, i64, vp, vp, i64, i32, vp, vp]),
    "tm_reindex": (i32, [vp, i64, i64, vp, vp, vp, vp]),
    "tm_reindex_dev": (i32, [vp, i64, i64, vp, vp, vp, vp]),
    "tm_kmodes_rows": (i32, [vp, i64, i32, vp, vp]),
    "tm_kmodes_assign_binned": (i32, [vp, i64, vp, vp, i64, vp, i32, i32, vp, vp]),
    "tm_temporal_smoothing": (i32, [vp, i64, vp, i32, i32, i32, vp, i64, i64, vp, vp, vp, C.c_double, vp, vp, vp, vp]),
    "tm_get_stats": (i32, [C.POINTER(TMStats)]),
    "tm_reset_stats": (i32, []),
    "tm_set_profiling": (i32, [i32]),
}

for _name, (_res, _args) in PROTOTYPES.items():
    _fn = getattr(lib, _name)  # AttributeError here = the .so does not export the header's symbol
    _fn.restype = _res
    _fn.argtypes = _args


class TMError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise TMError(f"libtmann error {rc}: {lib.tm_last_error().decode()}")


def check_handle(h):
    if not h:
        raise TMError(f"libtmann: {lib.tm_last_error().decode()}")
    return h
