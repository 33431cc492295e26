"""ctypes view of include/rela_b200.h (the C ABI of librela_b200.so).

The library is the product; this module only declares its signatures. There
is no fallback: if the library is missing or no CUDA device is present the
calls raise.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "librela_b200.so")

RR_OUT_RAW, RR_OUT_F32_DIV255, RR_OUT_BF16_DIV255 = 0, 1, 2
RR_OK, RR_ERR_ARG, RR_ERR_CUDA, RR_ERR_STATE, RR_ERR_EMPTY, RR_ERR_NOMEM, RR_ERR_OVERRUN = range(7)
RR_ROWS_DEVICE, RR_PRIO_DEVICE, RR_PRIO_IS_WEIGHT = 1, 2, 4
RR_NOALLOC_MAX_BATCH = 8192  # rr_update_priority_noalloc: elements of its static work area

SYMBOLS = [
    "rr_last_error", "rr_create", "rr_destroy", "rr_set_schema", "rr_add", "rr_add_rows", "rr_sample",
    "rr_update_priority", "rr_update_priority_noalloc", "rr_update_priority_sharded", "rr_size", "rr_num_add", "rr_sync", "rr_get_stats", "rr_last_sample_ids",
    "rr_debug_state", "rr_debug_weights", "rr_debug_tiles", "rr_debug_select_probe", "rr_debug_timeline", "rr_stream_fence",
    "rr_get_stream", "rr_stream_wait", "rr_save", "rr_load", "rr_restore_prefetched", "rr_prefetched_batch",
    "rr_record_prefetched",
    "rr_set_output_format", "rr_set_profiling", "rr_gather_bench", "rr_cycle_bench", "rr_nstep_return", "rr_td_priority",
    "rr_aggregate_priority", "rr_sample_select", "rr_sample_gather", "rr_prefetch_into", "rr_prefetch_routed",
    "rr_route_assignment",
    "rr_take_prefetched", "rr_prefetched", "rr_scatter_rows", "rr_peer_put", "rr_wait_flags", "rr_exchange",
    "rr_set_shard", "rr_shard_range", "rr_slab_ptrs", "rr_add_sharded_rows", "rr_add_sharded_commit",
    "rr_device_alloc", "rr_device_free", "rr_ipc_export", "rr_ipc_open", "rr_ipc_close",
    "rr_group_create", "rr_group_allgather", "rr_group_barrier", "rr_group_rank", "rr_group_world",
    "rr_group_destroy",
]


class RRField(C.Structure):
    _fields_ = [("name", C.c_char_p), ("row_bytes", C.c_uint64), ("chunks", C.c_uint32),
                ("reserved", C.c_uint32)]


class RRStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "kernel_launches", "samples", "samples_exact", "samples_ordered", "pops", "device_errors",
        "slab_bytes", "h2d_bytes")]


class RRRoute(C.Structure):
    _fields_ = [("n_learners", C.c_int), ("dst_tab", C.c_void_p), ("signal_ptrs", C.POINTER(C.c_void_p)),
                ("n_signals", C.c_int), ("wait_flags", C.c_void_p), ("wait_n", C.c_int),
                ("my_learner", C.c_int), ("out_weights", C.c_void_p), ("on_demand", C.c_int),
                ("my_chunk", C.POINTER(C.c_int))]


class ReplayError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"rela_b200 error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(rela_b200 has no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp, i, f, u = C.c_void_p, C.c_int, C.c_float, C.c_uint
        L.rr_last_error.restype = C.c_char_p
        L.rr_create.argtypes = [C.POINTER(vp), i, i, f, f, i, i]
        L.rr_destroy.argtypes = [vp]
        L.rr_set_schema.argtypes = [vp, i, C.POINTER(RRField)]
        L.rr_add.argtypes = [vp, i, vp, vp, vp, u, vp]
        L.rr_add_rows.argtypes = [vp, i, vp, vp, u, vp]
        L.rr_sample.argtypes = [vp, i, vp, vp, vp]
        L.rr_update_priority.argtypes = [vp, i, vp, u, vp]
        L.rr_update_priority_noalloc.argtypes = [vp, i, vp, u, vp]
        L.rr_update_priority_sharded.argtypes = [vp, i, vp, i, i, u, vp, i, C.c_uint32, vp, vp]
        L.rr_size.argtypes = [vp]
        L.rr_num_add.argtypes = [vp]
        L.rr_sync.argtypes = [vp]
        L.rr_get_stats.argtypes = [vp, C.POINTER(RRStats)]
        L.rr_last_sample_ids.argtypes = [vp, i, vp]
        L.rr_debug_state.argtypes = [vp, vp, vp, vp, vp, vp]
        L.rr_debug_weights.argtypes = [vp, vp, vp]
        L.rr_debug_tiles.argtypes = [vp, i, i, vp, vp, i, vp]
        L.rr_debug_select_probe.argtypes = [vp, vp]
        L.rr_debug_timeline.argtypes = [vp, i, vp, i]
        L.rr_stream_fence.argtypes = [vp, vp]
        L.rr_get_stream.argtypes = [vp, vp]
        L.rr_prefetch_into.argtypes = [vp, i, vp, vp, vp]
        L.rr_take_prefetched.argtypes = [vp, vp]
        L.rr_prefetched.argtypes = [vp]
        L.rr_stream_wait.argtypes = [vp, vp]
        L.rr_save.argtypes = [vp, C.c_char_p]
        L.rr_load.argtypes = [vp, C.c_char_p]
        L.rr_set_output_format.argtypes = [vp, i, i]
        L.rr_set_profiling.argtypes = [vp, i, i]
        L.rr_gather_bench.argtypes = [vp, i, vp, i, i, vp, vp, vp]
        L.rr_cycle_bench.argtypes = [vp, i, i, i, i, i, vp, vp, i, vp, u, vp, vp]
        L.rr_nstep_return.argtypes = [i, i, f, vp, vp, vp, vp, vp, i, vp]
        L.rr_td_priority.argtypes = [i, vp, vp, vp, vp, f, vp, vp]
        L.rr_aggregate_priority.argtypes = [i, i, vp, vp, f, i, vp, vp]
        L.rr_sample_select.argtypes = [vp, i, vp, vp]
        L.rr_sample_gather.argtypes = [vp, vp, vp]
        L.rr_scatter_rows.argtypes = [vp, i, i, vp, vp, vp, vp, vp]
        L.rr_peer_put.argtypes = [i, vp, C.c_uint64, vp, vp, i, C.c_uint64, vp]
        L.rr_wait_flags.argtypes = [i, vp, i, C.c_uint64, vp, vp]
        L.rr_exchange.argtypes = [vp, vp, C.c_uint64, vp, vp, i, C.c_uint64, vp, i, vp]
        L.rr_prefetch_routed.argtypes = [vp, i, C.POINTER(RRRoute), vp]
        L.rr_route_assignment.argtypes = [i, i, i, i, vp]
        L.rr_slab_ptrs.argtypes = [vp, vp, i]
        L.rr_add_sharded_rows.argtypes = [vp, i, i, i, vp, vp, vp]
        L.rr_add_sharded_commit.argtypes = [vp, vp, u]
        L.rr_restore_prefetched.argtypes = [vp, i, vp, vp, vp]
        L.rr_prefetched_batch.argtypes = [vp, i]
        L.rr_record_prefetched.argtypes = [vp, vp]
        L.rr_group_create.argtypes = [C.POINTER(vp), C.c_char_p, i, i]
        L.rr_group_allgather.argtypes = [vp, vp, C.c_uint64, vp]
        L.rr_group_barrier.argtypes = [vp]
        L.rr_group_rank.argtypes = [vp]
        L.rr_group_world.argtypes = [vp]
        L.rr_group_destroy.argtypes = [vp]
        L.rr_set_shard.argtypes = [vp, i, i]
        L.rr_shard_range.argtypes = [vp, vp, vp]
        L.rr_device_alloc.argtypes = [i, C.c_uint64, C.POINTER(vp)]
        L.rr_device_free.argtypes = [i, vp]
        L.rr_ipc_export.argtypes = [i, vp, vp]
        L.rr_ipc_open.argtypes = [i, vp, C.POINTER(vp)]
        L.rr_ipc_close.argtypes = [i, vp]
        _lib = L
    return _lib


def check(rc):
    if rc != RR_OK:
        raise ReplayError(rc, lib().rr_last_error().decode())
