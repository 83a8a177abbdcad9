"""Host-side mirror of the reference's Python binding, for the embedding path.

Same class and method names, argument order and meaning as
python/flexflow/core/flexflow_cbinding.py (FFConfig :346-372, Tensor :380-527,
Parameter :529-562, FFModel :564-875, SGDOptimizer :877-886, initialisers
:904-950, Op/Embedding :52-84,191-193), so that the DLRM example's

    t = ffmodel.embedding(input, num_entries, out_dim, AggrMode.AGGR_MODE_SUM,
                          None, UniformInitializer(seed, -r, r))
    ffmodel.compile(optimizer=SGDOptimizer(ffmodel, 0.01), ...)
    ffmodel.forward(); ffmodel.zero_gradients(); ffmodel.backward(); ffmodel.update()

runs unchanged on top of libffemb.so.  Everything goes through the C-ABI of
include/ffemb_c.h via ctypes; torch/numpy appear only as buffer handles.
"""
import ctypes as C
from enum import Enum

import numpy as np

from . import _lib
from ._lib import FFEmbError, Handle, check, lib


# ---- python/flexflow/core/flexflow_type.py:1-27 ------------------------------
class AggrMode(Enum):
    AGGR_MODE_NONE = 20
    AGGR_MODE_SUM = 21
    AGGR_MODE_AVG = 22


class DataType(Enum):
    DT_FLOAT = 40
    DT_DOUBLE = 41
    DT_INT32 = 42
    DT_INT64 = 43
    DT_BOOLEAN = 44


class OpType(Enum):
    EMBEDDING = 2012


def enum_to_int(enum, enum_item):
    for item in enum:
        if enum_item == item:
            return item.value
    if isinstance(enum_item, int) and any(enum_item == item.value for item in enum):
        return enum_item
    assert 0, "unknow enum type " + str(enum_item) + " " + str(enum)
    return -1


def int_to_enum(enum, value):
    for item in enum:
        if item.value == value:
            return item
    assert 0, "unknow enum value " + str(value) + " " + str(enum)


_NULL = Handle(None)


class _Owned(object):
    """Calls the C destroy function when the Python wrapper dies (ffi.gc in the reference)."""

    def __init__(self, handle, destroy):
        self.handle = handle
        self._destroy = destroy

    def __del__(self):
        try:
            if self.handle.impl and self._destroy is not None:
                self._destroy(self.handle)
                self.handle = _NULL
        except Exception:
            pass


# ---- FFConfig ------------------------------------------------------------------
class FFConfig(_Owned):
    def __init__(self):
        super().__init__(lib().ffemb_config_create(), lib().ffemb_config_destroy)

    def parse_args(self, argv=None):
        """The flags of FFConfig::parse_args that matter here (model.cc:1313-1381):
        -b/--batch-size, -ll:gpu, --nodes, --import/--import-strategy."""
        import sys
        argv = list(sys.argv[1:] if argv is None else argv)
        i = 0
        while i < len(argv):
            a = argv[i]
            if a in ("-b", "--batch-size") and i + 1 < len(argv):
                self.set_batch_size(int(argv[i + 1])); i += 1
            elif a == "-ll:gpu" and i + 1 < len(argv):
                check(lib().ffemb_config_set_workers_per_node(self.handle, int(argv[i + 1]))); i += 1
            elif a == "--nodes" and i + 1 < len(argv):
                check(lib().ffemb_config_set_num_nodes(self.handle, int(argv[i + 1]))); i += 1
            elif a in ("--import", "--import-strategy") and i + 1 < len(argv):
                self.import_strategy_file(argv[i + 1]); i += 1
            elif a == "--replicate-below" and i + 1 < len(argv):
                self.set_replicate_below(argv[i + 1] if argv[i + 1] == "auto" else int(argv[i + 1])); i += 1
            i += 1

    def get_batch_size(self):
        return lib().ffemb_config_get_batch_size(self.handle)

    def set_batch_size(self, batch_size):
        check(lib().ffemb_config_set_batch_size(self.handle, batch_size))

    def get_workers_per_node(self):
        return lib().ffemb_config_get_workers_per_node(self.handle)

    def set_workers_per_node(self, n):
        check(lib().ffemb_config_set_workers_per_node(self.handle, n))

    def get_num_nodes(self):
        return lib().ffemb_config_get_num_nodes(self.handle)

    def set_num_nodes(self, n):
        check(lib().ffemb_config_set_num_nodes(self.handle, n))

    def set_rank(self, rank, cuda_device=-1):
        check(lib().ffemb_config_set_rank(self.handle, rank, cuda_device))

    def get_rank(self):
        return lib().ffemb_config_get_rank(self.handle)

    def import_strategy_file(self, path):
        check(lib().ffemb_config_import_strategy_file(self.handle, path.encode()))

    def set_replicate_below(self, rows):
        """More than one rank: tables with fewer rows than this stay on every rank
        (data-parallel, src/runtime/model.cc:697-724) instead of being sharded.
        0: shard everything (the reference's strategy files); -1 or "auto":
        batch / ranks, where a replica becomes cheaper than a shard."""
        check(lib().ffemb_config_set_replicate_below(self.handle, -1 if rows == "auto" else int(rows)))

    def set_gradient_exchange(self, mode):
        """'auto' (pull with two ranks and one-hot bags, push otherwise), 'push' or 'pull'."""
        check(lib().ffemb_config_set_gradient_exchange(self.handle, {"auto": 0, "push": 1, "pull": 2}[mode]))

    def set_exchange_engine(self, force=True):
        """Run the sharded engine even with one rank (single-GPU tests)."""
        check(lib().ffemb_config_set_exchange_engine(self.handle, int(bool(force))))

    def begin_trace(self, trace_id):
        """Legion tracing (flexflow_begin_trace) has no counterpart: steps are
        plain stream work and can be captured in a CUDA graph by the caller."""

    def end_trace(self, trace_id):
        pass


# ---- Tensor / Parameter --------------------------------------------------------
class Tensor(object):
    def __init__(self, handle, model=None, owner_op_type=None):
        assert handle.impl, "Tensor handle is wrong: " + _lib.last_error()
        self.handle = handle
        self._model = model
        self.owner_op_type = owner_op_type
        self.num_dims = lib().ffemb_tensor_get_num_dims(handle)
        self.dims = [lib().ffemb_tensor_get_dim(handle, i) for i in range(self.num_dims)]
        self.data_type = int_to_enum(DataType, lib().ffemb_tensor_get_data_type(handle))
        self._keep = []  # borrowed buffers stay alive as long as they are attached

    # -- explicit data movement (the reference uses dataloader tasks / inline maps)
    def set_array(self, ffmodel, np_array):
        """Host -> device copy into an int64 input tensor (enqueued)."""
        a = np.ascontiguousarray(np_array, dtype=np.int64)
        self._keep = [a]
        check(lib().ffemb_tensor_set_int64(self.handle, ffmodel.handle, a.ctypes.data, a.size))

    def set_pinned(self, ffmodel, torch_tensor):
        """Same, from a (pinned) torch CPU tensor without a numpy round trip."""
        self._keep = [torch_tensor]
        check(lib().ffemb_tensor_set_int64(self.handle, ffmodel.handle, torch_tensor.data_ptr(),
                                           torch_tensor.numel()))

    def set_grad(self, ffmodel, array):
        """Host -> device copy of d(loss)/d(output) (numpy or torch CPU, float32)."""
        if isinstance(array, np.ndarray):
            array = np.ascontiguousarray(array, dtype=np.float32)
            ptr, n = array.ctypes.data, array.size
        else:
            ptr, n = array.data_ptr(), array.numel()
        self._keep_grad = array
        check(lib().ffemb_tensor_set_grad_float(self.handle, ffmodel.handle, ptr, n))

    def get_array(self, ffmodel, data_type=None):
        """Device -> host copy of an embedding output (blocks)."""
        out = np.empty(self.dims, dtype=np.float32)
        check(lib().ffemb_tensor_get_float(self.handle, ffmodel.handle, out.ctypes.data, out.size))
        return out

    def get_array_slice(self, ffmodel):
        """Sharded mode: this rank's samples of an embedding output, [batch/ranks, out_dim] (blocks)."""
        world = ffmodel._ffconfig.get_workers_per_node() * ffmodel._ffconfig.get_num_nodes()
        out = np.empty((self.dims[0] // world, self.dims[1]), dtype=np.float32)
        check(lib().ffemb_tensor_get_float(self.handle, ffmodel.handle, out.ctypes.data, out.size))
        return out

    def get_into(self, ffmodel, torch_tensor):
        """Device -> host copy into a (pinned) torch CPU tensor (blocks)."""
        check(lib().ffemb_tensor_get_float(self.handle, ffmodel.handle, torch_tensor.data_ptr(),
                                           torch_tensor.numel()))

    def attach_torch(self, tensor, grad=False):
        """Borrow a CUDA torch tensor as this tensor's data (or gradient)."""
        if not tensor.is_cuda:
            raise ValueError("attach_torch needs a CUDA tensor")
        if tensor.dim() != 2 or tensor.stride(1) != 1:
            raise ValueError("attach_torch needs a 2-D tensor with unit column stride")
        if grad:
            self._keep_grad = tensor
        else:
            self._keep = [tensor]
        check(lib().ffemb_tensor_attach_device_ptr(self.handle, tensor.data_ptr(), tensor.stride(0),
                                                   grad))

    def device_ptr(self, grad=False):
        return lib().ffemb_tensor_get_device_ptr(self.handle, grad)

    def ld(self, grad=False):
        return lib().ffemb_tensor_get_ld(self.handle, grad)


class Parameter(object):
    def __init__(self, handle):
        assert handle.impl, "Parameter handle is wrong"
        self.parameter_handle = handle
        self.dims = [lib().ffemb_parameter_get_num_entries(handle),
                     lib().ffemb_parameter_get_out_dim(handle)]
        self.num_dims = 2

    def set_weights(self, ffmodel, np_array):
        assert np_array.__array_interface__['strides'] is None, \
            "Parameter set_weights, numpy array strides is not None"
        assert np_array.dtype == np.float32, "Parameter set_weights needs float32"
        np_shape = np_array.shape
        assert len(np_shape) == self.num_dims, "please check dims (%d == %d)" % (len(np_shape), self.num_dims)
        for i in range(self.num_dims):
            assert np_shape[i] == self.dims[i], \
                "please check shape dim %d (%d == %d)" % (i, np_shape[i], self.dims[i])
        c_dims = (C.c_int * self.num_dims)(*self.dims)
        ret_val = lib().ffemb_parameter_set_weights_float(self.parameter_handle, ffmodel.handle,
                                                          self.num_dims, c_dims, np_array.ctypes.data)
        assert ret_val is True, _lib.last_error()

    def get_weights(self, ffmodel):
        np_array = np.empty(self.dims, dtype=np.float32)
        ret_val = lib().ffemb_parameter_get_weights_float(self.parameter_handle, ffmodel.handle,
                                                          np_array.ctypes.data)
        assert ret_val is True, _lib.last_error()
        return np_array

    def device_ptr(self):
        return lib().ffemb_parameter_get_device_ptr(self.parameter_handle)

    def as_torch(self, device=None):
        """Zero-copy torch view of the table (owned by the library)."""
        import torch
        ptr = self.device_ptr()
        if not ptr:
            raise FFEmbError(_lib.FFEMB_ERR_INVALID_ARGUMENT, "table is not owned by this rank")

        class _Mem(object):
            pass
        n, d = self.dims
        holder = _Mem()
        holder.__cuda_array_interface__ = {"shape": (n, d), "typestr": "<f4", "data": (ptr, False),
                                           "version": 2, "strides": None}
        return torch.as_tensor(holder, device=device or "cuda")


# ---- Op / Embedding ------------------------------------------------------------
class Op(object):
    def __init__(self, handle, idx=None, name=None):
        assert handle.impl, "Op handle is wrong"
        self.handle = handle
        self.idx = idx
        self.name = name if name is not None else lib().ffemb_op_get_name(handle).decode()

    def _get_parameter_tensor_by_id(self, id):
        return Parameter(lib().ffemb_op_get_parameter_by_id(self.handle, id))

    def _get_input_tensor_by_id(self, id):
        return Tensor(lib().ffemb_op_get_input_by_id(self.handle, id))

    def _get_output_tensor_by_id(self, id):
        return Tensor(lib().ffemb_op_get_output_by_id(self.handle, id), owner_op_type=OpType.EMBEDDING)

    def get_device(self):
        return lib().ffemb_op_get_device(self.handle)


class Embedding(Op):
    def get_weight_tensor(self):
        return self._get_parameter_tensor_by_id(0)

    def get_input_tensor(self):
        return self._get_input_tensor_by_id(0)

    def get_output_tensor(self):
        return self._get_output_tensor_by_id(0)


# ---- optimiser / initialisers ----------------------------------------------------
class SGDOptimizer(_Owned):
    def __init__(self, ffmodel, lr=0.01, momentum=0.0, nesterov=False, weight_decay=0.0):
        super().__init__(lib().ffemb_sgd_optimizer_create(ffmodel.handle, lr, momentum, nesterov,
                                                          weight_decay),
                         lib().ffemb_sgd_optimizer_destroy)

    def set_learning_rate(self, learning_rate):
        lib().ffemb_sgd_optimizer_set_lr(self.handle, learning_rate)


class Initializer(_Owned):
    pass


class UniformInitializer(Initializer):
    def __init__(self, seed, minv, maxv):
        super().__init__(lib().ffemb_uniform_initializer_create(seed, minv, maxv),
                         lib().ffemb_uniform_initializer_destroy)


class ZeroInitializer(Initializer):
    def __init__(self):
        super().__init__(lib().ffemb_zero_initializer_create(), lib().ffemb_zero_initializer_destroy)


# ---- strategy / placement --------------------------------------------------------
class Strategy(_Owned):
    """FFProtoBuf.Strategy (src/runtime/strategy.proto) and the mapper's
    table -> device rule (src/mapper/mapper.cc:33-97)."""

    def __init__(self, path=None, data=None):
        super().__init__(lib().ffemb_strategy_create(), lib().ffemb_strategy_destroy)
        if path is not None:
            check(lib().ffemb_strategy_load_file(self.handle, path.encode()))
        if data is not None:
            buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
            check(lib().ffemb_strategy_parse(self.handle, buf, len(data)))

    @classmethod
    def dlrm(cls, gpus_per_node, embs_per_node, num_nodes):
        s = cls()
        check(lib().ffemb_strategy_generate_dlrm(s.handle, gpus_per_node, embs_per_node, num_nodes))
        return s

    @classmethod
    def dlrm_simple(cls, num_gpu, num_emb):
        s = cls()
        check(lib().ffemb_strategy_generate_dlrm_simple(s.handle, num_gpu, num_emb))
        return s

    def serialize(self):
        n = lib().ffemb_strategy_serialize(self.handle, None, 0)
        buf = (C.c_uint8 * max(n, 1))()
        lib().ffemb_strategy_serialize(self.handle, buf, n)
        return bytes(buf[:n])

    def save(self, path):
        check(lib().ffemb_strategy_save_file(self.handle, path.encode()))

    def add_op(self, name, device_type=0, dims=(), device_ids=(), memory_types=()):
        def arr(v):
            return (C.c_int * max(len(v), 1))(*v)
        check(lib().ffemb_strategy_add_op(self.handle, name.encode(), device_type, arr(dims), len(dims),
                                          arr(device_ids), len(device_ids), arr(memory_types),
                                          len(memory_types)))

    def ops(self):
        L, h = lib(), self.handle
        out = []
        for i in range(L.ffemb_strategy_num_ops(h)):
            out.append({
                "name": L.ffemb_strategy_op_name(h, i).decode(),
                "device_type": L.ffemb_strategy_op_device_type(h, i),
                "dims": [L.ffemb_strategy_op_dim(h, i, k) for k in range(L.ffemb_strategy_op_num_dims(h, i))],
                "device_ids": [L.ffemb_strategy_op_device_id(h, i, k)
                               for k in range(L.ffemb_strategy_op_num_device_ids(h, i))],
                "memory_types": [L.ffemb_strategy_op_memory_type(h, i, k)
                                 for k in range(L.ffemb_strategy_op_num_memory_types(h, i))],
            })
        return out

    def device_for_table(self, table_index, num_devices, op_name=None):
        return lib().ffemb_strategy_device_for_table(self.handle, table_index,
                                                     op_name.encode() if op_name else None, num_devices)


def round_robin_device(table_index, num_devices):
    """Default placement (src/runtime/dlrm_strategy.cc:252-256)."""
    return lib().ffemb_strategy_device_for_table(_NULL, table_index, None, num_devices)


def balance_placement(num_entries, lookups, out_dim, num_devices, bytes_capacity=0):
    T = len(num_entries)
    ne = (C.c_int64 * T)(*num_entries)
    lk = (C.c_int64 * T)(*lookups)
    owner = (C.c_int * T)()
    check(lib().ffemb_placement_balance(ne, lk, T, out_dim, num_devices, bytes_capacity, owner))
    return list(owner)


def plan_placement(num_entries, lookups, out_dim, batch, num_devices, bytes_capacity=0, cost=None):
    """Shard / replicate decision and balanced owners from the step cost model
    (ffemb_placement_plan).  Returns (owner list with -1 = replicated, modelled step in us)."""
    T = len(num_entries)
    ne = (C.c_int64 * T)(*num_entries)
    lk = (C.c_int64 * T)(*lookups)
    owner = (C.c_int * T)()
    us = C.c_double()
    cm = None
    if cost is not None:
        cm = _lib.CostModel(*cost)
    check(lib().ffemb_placement_plan(ne, lk, T, out_dim, batch, num_devices, bytes_capacity,
                                     C.byref(cm) if cm is not None else None, owner, C.byref(us)))
    return list(owner), us.value


# ---- FFModel ---------------------------------------------------------------------
class FFModel(_Owned):
    def __init__(self, ffconfig):
        h = lib().ffemb_model_create(ffconfig.handle)
        if not h.impl:
            raise FFEmbError(_lib.FFEMB_ERR_INVALID_ARGUMENT, _lib.last_error())
        super().__init__(h, lib().ffemb_model_destroy)
        self._layers = dict()
        self._nb_layers = 0
        self._ffconfig = ffconfig
        self._keep = []  # optimiser / initialisers / strategy must outlive the model

    def get_layers(self):
        return self._layers

    def add_layer(self, op_type, name):
        layer_id = self._nb_layers
        op_handle = lib().ffemb_model_get_layer_by_id(self.handle, layer_id)
        self._layers[layer_id] = Embedding(op_handle, idx=layer_id, name=name)
        self._nb_layers += 1

    def create_tensor(self, dims, data_type, create_grad=True, name=None):
        c_dims = (C.c_int * len(dims))(*dims)
        c_data_type = enum_to_int(DataType, data_type)
        handle = lib().ffemb_tensor_create(self.handle, len(dims), c_dims,
                                           name.encode() if name else None, c_data_type, create_grad)
        if not handle.impl:
            raise FFEmbError(_lib.FFEMB_ERR_INVALID_ARGUMENT, _lib.last_error())
        return Tensor(handle, self)

    def embedding(self, input, num_entires, out_dim, aggr, shared_op=None, kernel_initializer=None,
                  name=None):
        shared_op_handle = shared_op.handle if shared_op is not None else _NULL
        c_aggr = enum_to_int(AggrMode, aggr)
        assert kernel_initializer is None or isinstance(kernel_initializer, Initializer), \
            "unknow initializer type"
        init_handle = kernel_initializer.handle if kernel_initializer is not None else _NULL
        if kernel_initializer is not None:
            self._keep.append(kernel_initializer)
        handle = lib().ffemb_model_add_embedding(self.handle, input.handle, num_entires, out_dim,
                                                 c_aggr, shared_op_handle, init_handle)
        if not handle.impl:
            raise FFEmbError(_lib.FFEMB_ERR_INVALID_ARGUMENT, _lib.last_error())
        self.add_layer(OpType.EMBEDDING, name)
        return Tensor(handle, self, owner_op_type=OpType.EMBEDDING)

    def set_sgd_optimizer(self, optimizer):
        self._keep.append(optimizer)
        check(lib().ffemb_model_set_sgd_optimizer(self.handle, optimizer.handle))

    def set_strategy(self, strategy):
        self._keep.append(strategy)
        check(lib().ffemb_model_set_strategy(self.handle, strategy.handle))

    def set_presort(self, enable=True):
        check(lib().ffemb_model_set_presort(self.handle, bool(enable)))

    def set_stream(self, cuda_stream_ptr):
        check(lib().ffemb_model_set_stream(self.handle, cuda_stream_ptr))

    def compile(self, optimizer=None, loss_type=None, metrics=None, group=None):
        """loss_type / metrics belong to the dense part of the model and are ignored.

        With more than one rank (FFConfig workers_per_node * num_nodes, one
        process per GPU) the ranks' exchange buffers are connected here: the
        64-byte IPC handles travel through torch.distributed (`group`, default
        group if None), which must be initialised."""
        if optimizer is not None:
            assert isinstance(optimizer, SGDOptimizer), "unknow optimizer type (this path: SGD)"
            self.set_sgd_optimizer(optimizer)
        check(lib().ffemb_model_compile(self.handle))
        world = self._ffconfig.get_workers_per_node() * self._ffconfig.get_num_nodes()
        self._group = group
        self._world = world
        if world > 1:
            import torch.distributed as dist
            if not dist.is_initialized():
                raise FFEmbError(_lib.FFEMB_ERR_COMM, "more than one rank: initialise torch.distributed "
                                 "before FFModel.compile (it carries the ranks' IPC handles)")
            mine = (C.c_uint8 * 64)()
            check(lib().ffemb_model_get_exchange_handle(self.handle, mine))
            handles = [None] * world
            dist.all_gather_object(handles, bytes(mine), group=group)
            buf = (C.c_uint8 * (64 * world)).from_buffer_copy(b"".join(handles))
            check(lib().ffemb_model_connect_peers(self.handle, buf, world))
            dist.barrier(group=group)      # every rank is connected before anyone stores into a peer

    # -- double-buffered host ingest / egress (the reference: DataLoader::next_batch,
    #    examples/cpp/DLRM/dlrm.cc:486-524, and zero-copy outputs, embedding.cu:280-284)
    def load_batch_async(self, indices, out_grad=None):
        """indices: per layer a pinned torch int64 [batch, bag] tensor (or None for a
        table this rank does not read); out_grad: pinned torch float32 [rows, sum D]."""
        n = len(indices)
        arr = (C.c_void_p * n)(*[None if i is None else i.data_ptr() for i in indices])
        self._pipe_keep = (indices, out_grad)
        check(lib().ffemb_model_load_batch_async(self.handle, arr, None if out_grad is None else out_grad.data_ptr()))

    def begin_batch(self):
        check(lib().ffemb_model_begin_batch(self.handle))

    def fetch_outputs_async(self, out_host):
        self._pipe_out = out_host
        check(lib().ffemb_model_fetch_outputs_async(self.handle, out_host.data_ptr()))

    def wait_outputs(self, all=False):
        check(lib().ffemb_model_wait_outputs(self.handle, bool(all)))

    def trace(self, enable=True):
        check(lib().ffemb_model_trace(self.handle, int(enable)))

    def trace_report(self):
        n = lib().ffemb_model_trace_report(self.handle, None, 0)
        buf = C.create_string_buffer(n + 16)
        lib().ffemb_model_trace_report(self.handle, buf, n + 16)
        return buf.value.decode()

    def is_sharded(self):
        return bool(lib().ffemb_model_is_sharded(self.handle))

    def exchange_matrix(self, grad=False):
        """Sharded mode: zero-copy torch view of this rank's [batch/ranks, sum D]
        output (or gradient) matrix, every table's block side by side."""
        import torch
        ptr, ld, rows = C.c_void_p(), C.c_int64(), C.c_int()
        check(lib().ffemb_model_get_exchange_matrix(self.handle, grad, C.byref(ptr), C.byref(ld), C.byref(rows)))

        class _Mem(object):
            pass
        holder = _Mem()
        holder.__cuda_array_interface__ = {"shape": (rows.value, ld.value), "typestr": "<f4",
                                           "data": (ptr.value, False), "version": 2, "strides": None}
        return torch.as_tensor(holder, device="cuda")

    def close(self):
        """Sharded mode: no rank unmaps its peers' memory while one of them may
        still be storing into it.  Collective; call it on every rank."""
        if self.handle.impl:
            if getattr(self, "_world", 1) > 1:
                import torch
                import torch.distributed as dist
                torch.cuda.synchronize()
                dist.barrier(group=self._group)
            lib().ffemb_model_destroy(self.handle)
            self.handle = _NULL

    def init_layers(self):
        check(lib().ffemb_model_init_layers(self.handle))

    def forward(self):
        check(lib().ffemb_model_forward(self.handle))

    def zero_gradients(self):
        check(lib().ffemb_model_zero_gradients(self.handle))

    def backward(self):
        check(lib().ffemb_model_backward(self.handle))

    def update(self):
        check(lib().ffemb_model_update(self.handle))

    def sync(self):
        check(lib().ffemb_model_sync(self.handle))

    def get_layer_by_id(self, layer_id):
        return self._layers[layer_id]

    def get_layer_by_name(self, layer_name):
        for layer_id in self._layers:
            layer = self._layers[layer_id]
            if layer.name == layer_name:
                return layer
        assert 0, "Can not find the layer with the name"
        return None
