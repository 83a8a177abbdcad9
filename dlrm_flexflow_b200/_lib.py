"""ctypes binding of include/ffemb_c.h (the C-ABI).  No torch types cross it.

The shared library is built in-tree (dlrm_flexflow_b200/libffemb.so).  If it
is missing it is compiled with nvcc; if that is impossible the import fails
loudly -- there is no CPU or PyTorch fallback for the compute path.
"""
import ctypes as C
import os

from . import build as _build

FFEMB_MAX_PARTS = 16

# status codes (include/ffemb_const.h)
FFEMB_OK = 0
FFEMB_ERR_INVALID_ARGUMENT = 1
FFEMB_ERR_CUDA = 2
FFEMB_ERR_INDEX_OUT_OF_RANGE = 3
FFEMB_ERR_UNSUPPORTED = 4
FFEMB_ERR_WORKSPACE_TOO_SMALL = 5
FFEMB_ERR_IO = 6
FFEMB_ERR_COMM = 7


class FFEmbError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("ffemb status %d: %s" % (status, message))
        self.status = status


class Lookup(C.Structure):
    """ffemb_lookup_t"""
    _fields_ = [("weight", C.c_void_p), ("indices", C.c_void_p), ("offsets", C.c_void_p),
                ("num_entries", C.c_int64), ("num_indices", C.c_int64), ("col", C.c_int64),
                ("bag_size", C.c_int32), ("reserved", C.c_int32)]


class Matrix(C.Structure):
    """ffemb_matrix_t"""
    _fields_ = [("base", C.c_void_p * FFEMB_MAX_PARTS), ("num_parts", C.c_int32),
                ("rows_per_part", C.c_int32), ("ld", C.c_int64)]


class Push(C.Structure):
    """ffemb_push_t"""
    _fields_ = [("dst", C.c_void_p), ("dst_ld", C.c_int64), ("src_col", C.c_int64)]


class Handle(C.Structure):
    """Every FFEMB_NEW_OPAQUE_TYPE: struct { void *impl; }"""
    _fields_ = [("impl", C.c_void_p)]


_H = Handle
_vp, _i, _i64, _sz, _f, _d, _b = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float, C.c_double, C.c_bool
_cp = C.c_char_p
_ip = C.POINTER(C.c_int)
_LP = C.POINTER(Lookup)
_MP = C.POINTER(Matrix)

# name: (restype, argtypes) -- one entry per declaration in include/ffemb_c.h
SIGNATURES = {
    "ffemb_last_error": (_cp, []),
    "ffemb_version": (_cp, []),
    "ffemb_compiled_sm": (_i, []),
    "ffemb_kernel_launch_count": (_i64, []),
    # kernel level
    "ffemb_embed_forward": (_i, [_LP, _i, _i, _i, _i, _MP, _vp, _vp]),
    "ffemb_embed_backward_workspace_bytes": (_sz, [_i64, _i64, _i]),
    "ffemb_embed_backward_sgd": (_i, [_LP, _i, _i, _i, _i, _MP, _f, _vp, _sz, _vp, _vp]),
    "ffemb_embed_backward_sort": (_i, [_LP, _i, _i, _i, _vp, _sz, _vp, _vp]),
    "ffemb_embed_sgd_update": (_i, [_LP, _i, _i, _i, _i, _MP, _f, _vp, _sz, _vp]),
    "ffemb_set_sort_impl": (_i, [_i]),
    "ffemb_get_sort_impl": (_i, []),
    "ffemb_static_scratch_bytes": (_sz, []),
    "ffemb_embed_backward_sgd_static": (_i, [_LP, _i, _i, _i, _i, _MP, _f, _vp, _vp]),
    "ffemb_uniform_init": (_i, [_vp, _i64, _i, _f, _f, _vp]),
    "ffemb_dense_sgd": (_i, [_vp, _vp, _i64, _f, _vp]),
    "ffemb_device_malloc": (_i, [C.POINTER(C.c_void_p), _sz]),
    "ffemb_device_free": (_i, [_vp]),
    "ffemb_ipc_get_handle": (_i, [_vp, _vp]),
    "ffemb_ipc_open_handle": (_i, [_vp, C.POINTER(C.c_void_p)]),
    "ffemb_ipc_close_handle": (_i, [_vp]),
    "ffemb_peer_barrier": (_i, [C.POINTER(C.c_void_p), _i, _i, _vp, _vp]),
    "ffemb_push_columns": (_i, [C.POINTER(Push), _i, _i, _i, _vp, _i64, _vp]),
    "ffemb_replica_reduce_update": (_i, [C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), _i64, _i, _i, _f, _vp]),
    # config
    "ffemb_config_create": (_H, []),
    "ffemb_config_destroy": (None, [_H]),
    "ffemb_config_set_batch_size": (_i, [_H, _i]),
    "ffemb_config_get_batch_size": (_i, [_H]),
    "ffemb_config_set_workers_per_node": (_i, [_H, _i]),
    "ffemb_config_get_workers_per_node": (_i, [_H]),
    "ffemb_config_set_num_nodes": (_i, [_H, _i]),
    "ffemb_config_get_num_nodes": (_i, [_H]),
    "ffemb_config_set_rank": (_i, [_H, _i, _i]),
    "ffemb_config_get_rank": (_i, [_H]),
    "ffemb_config_import_strategy_file": (_i, [_H, _cp]),
    "ffemb_config_set_replicate_below": (_i, [_H, _i64]),
    "ffemb_config_set_exchange_engine": (_i, [_H, _i]),
    "ffemb_config_set_gradient_exchange": (_i, [_H, _i]),
    # model
    "ffemb_model_create": (_H, [_H]),
    "ffemb_model_destroy": (None, [_H]),
    "ffemb_tensor_create": (_H, [_H, _i, _ip, _cp, _i, _b]),
    "ffemb_tensor_destroy": (None, [_H]),
    "ffemb_tensor_get_num_dims": (_i, [_H]),
    "ffemb_tensor_get_dim": (_i, [_H, _i]),
    "ffemb_tensor_get_data_type": (_i, [_H]),
    "ffemb_model_add_embedding": (_H, [_H, _H, _i, _i, _i, _H, _H]),
    "ffemb_model_set_sgd_optimizer": (_i, [_H, _H]),
    "ffemb_model_compile": (_i, [_H]),
    "ffemb_model_init_layers": (_i, [_H]),
    "ffemb_model_forward": (_i, [_H]),
    "ffemb_model_zero_gradients": (_i, [_H]),
    "ffemb_model_backward": (_i, [_H]),
    "ffemb_model_update": (_i, [_H]),
    "ffemb_model_sync": (_i, [_H]),
    "ffemb_model_set_stream": (_i, [_H, _vp]),
    "ffemb_model_set_presort": (_i, [_H, _b]),
    "ffemb_model_get_num_layers": (_i, [_H]),
    "ffemb_model_get_layer_by_id": (_H, [_H, _i]),
    "ffemb_model_set_strategy": (_i, [_H, _H]),
    "ffemb_model_is_sharded": (_i, [_H]),
    "ffemb_model_trace": (_i, [_H, _i]),
    "ffemb_model_trace_report": (_sz, [_H, _vp, _sz]),
    "ffemb_model_get_exchange_handle": (_i, [_H, _vp]),
    "ffemb_model_connect_peers": (_i, [_H, _vp, _i]),
    "ffemb_model_get_exchange_matrix": (_i, [_H, _b, C.POINTER(C.c_void_p), C.POINTER(_i64), _ip]),
    "ffemb_model_load_batch_async": (_i, [_H, C.POINTER(C.c_void_p), _vp]),
    "ffemb_model_begin_batch": (_i, [_H]),
    "ffemb_model_fetch_outputs_async": (_i, [_H, _vp]),
    "ffemb_model_wait_outputs": (_i, [_H, _b]),
    # op
    "ffemb_op_get_parameter_by_id": (_H, [_H, _i]),
    "ffemb_op_get_input_by_id": (_H, [_H, _i]),
    "ffemb_op_get_output_by_id": (_H, [_H, _i]),
    "ffemb_op_get_name": (_cp, [_H]),
    "ffemb_op_get_device": (_i, [_H]),
    # tensor data
    "ffemb_tensor_set_int64": (_i, [_H, _H, _vp, _i64]),
    "ffemb_tensor_set_grad_float": (_i, [_H, _H, _vp, _i64]),
    "ffemb_tensor_get_float": (_i, [_H, _H, _vp, _i64]),
    "ffemb_tensor_attach_device_ptr": (_i, [_H, _vp, _i64, _b]),
    "ffemb_tensor_get_device_ptr": (_vp, [_H, _b]),
    "ffemb_tensor_get_ld": (_i64, [_H, _b]),
    # parameter
    "ffemb_parameter_set_weights_float": (_b, [_H, _H, _i, _ip, _vp]),
    "ffemb_parameter_get_weights_float": (_b, [_H, _H, _vp]),
    "ffemb_parameter_get_device_ptr": (_vp, [_H]),
    "ffemb_parameter_get_num_entries": (_i64, [_H]),
    "ffemb_parameter_get_out_dim": (_i, [_H]),
    # optimizer / initialisers
    "ffemb_sgd_optimizer_create": (_H, [_H, _d, _d, _b, _d]),
    "ffemb_sgd_optimizer_destroy": (None, [_H]),
    "ffemb_sgd_optimizer_set_lr": (None, [_H, _d]),
    "ffemb_uniform_initializer_create": (_H, [_i, _f, _f]),
    "ffemb_uniform_initializer_destroy": (None, [_H]),
    "ffemb_zero_initializer_create": (_H, []),
    "ffemb_zero_initializer_destroy": (None, [_H]),
    # placement
    "ffemb_strategy_create": (_H, []),
    "ffemb_strategy_destroy": (None, [_H]),
    "ffemb_strategy_load_file": (_i, [_H, _cp]),
    "ffemb_strategy_save_file": (_i, [_H, _cp]),
    "ffemb_strategy_parse": (_i, [_H, _vp, _sz]),
    "ffemb_strategy_serialize": (_sz, [_H, _vp, _sz]),
    "ffemb_strategy_num_ops": (_i, [_H]),
    "ffemb_strategy_op_name": (_cp, [_H, _i]),
    "ffemb_strategy_op_device_type": (_i, [_H, _i]),
    "ffemb_strategy_op_num_dims": (_i, [_H, _i]),
    "ffemb_strategy_op_dim": (_i, [_H, _i, _i]),
    "ffemb_strategy_op_num_device_ids": (_i, [_H, _i]),
    "ffemb_strategy_op_device_id": (_i, [_H, _i, _i]),
    "ffemb_strategy_op_num_memory_types": (_i, [_H, _i]),
    "ffemb_strategy_op_memory_type": (_i, [_H, _i, _i]),
    "ffemb_strategy_add_op": (_i, [_H, _cp, _i, _ip, _i, _ip, _i, _ip, _i]),
    "ffemb_strategy_find": (_i, [_H, _cp]),
    "ffemb_strategy_generate_dlrm": (_i, [_H, _i, _i, _i]),
    "ffemb_strategy_generate_dlrm_simple": (_i, [_H, _i, _i]),
    "ffemb_strategy_device_for_table": (_i, [_H, _i, _cp, _i]),
    "ffemb_placement_balance": (_i, [C.POINTER(C.c_int64), C.POINTER(C.c_int64), _i, _i, _i, _i64, _ip]),
}

class CostModel(C.Structure):
    """ffemb_cost_model_t"""
    _fields_ = [("hbm_gbs", C.c_double), ("nvlink_gbs", C.c_double), ("hbm_passes", C.c_double),
                ("replica_latency_us", C.c_double)]


SIGNATURES["ffemb_placement_plan"] = (_i, [C.POINTER(C.c_int64), C.POINTER(C.c_int64), _i, _i, _i, _i, _i64,
                                           C.POINTER(CostModel), _ip, C.POINTER(C.c_double)])

_lib = None


def lib():
    """The loaded C-ABI library (built on first use)."""
    global _lib
    if _lib is None:
        path = _build.build()
        L = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError = header/library mismatch: fail loudly
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def last_error():
    return lib().ffemb_last_error().decode()


def check(status):
    """Raise FFEmbError unless status == FFEMB_OK."""
    if status != FFEMB_OK:
        raise FFEmbError(status, last_error())


def library_path():
    return os.path.abspath(_build.LIB)
