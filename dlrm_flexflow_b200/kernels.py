"""Kernel-level entry points over torch tensors (handles only: data_ptr()).

Thin argument marshalling for the (K) layer of include/ffemb_c.h.  PyTorch
supplies device memory and the stream; all arithmetic happens in libffemb.so.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import Lookup, Matrix, check

AGGR_MODE_SUM = 21
AGGR_MODE_AVG = 22


def _stream_ptr(stream=None):
    if stream is None:
        stream = torch.cuda.current_stream()
    return C.c_void_p(stream.cuda_stream)


def _require_cuda(t, dtype, what):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise ValueError("%s must be a CUDA tensor: the embedding path has no CPU fallback" % what)
    if t.dtype != dtype:
        raise TypeError("%s must be %s, got %s" % (what, dtype, t.dtype))


def make_lookups(weights, indices, offsets=None, cols=None):
    """Build the ffemb_lookup_t array for T tables.

    weights[t]  [N_t, D] float32 cuda, contiguous
    indices[t]  [B, L] int64 cuda (fixed bags) or flat [n_t] with offsets[t] [B+1]
    cols[t]     column of table t's block in the out / grad matrix (default t*D)
    """
    T = len(weights)
    arr = (Lookup * T)()
    for t in range(T):
        w, idx = weights[t], indices[t]
        _require_cuda(w, torch.float32, "weights[%d]" % t)
        _require_cuda(idx, torch.int64, "indices[%d]" % t)
        if not w.is_contiguous() or not idx.is_contiguous():
            raise ValueError("tables and index tensors must be contiguous")
        off = offsets[t] if offsets is not None else None
        l = arr[t]
        l.weight = w.data_ptr()
        l.indices = idx.data_ptr() if idx.numel() else None
        l.num_entries = w.shape[0]
        l.num_indices = idx.numel()
        l.col = cols[t] if cols is not None else t * w.shape[1]
        if off is not None:
            _require_cuda(off, torch.int64, "offsets[%d]" % t)
            l.offsets = off.data_ptr()
            l.bag_size = 0
        else:
            if idx.dim() != 2:
                raise ValueError("fixed-size bags need a [B, L] index tensor")
            l.offsets = None
            l.bag_size = idx.shape[1]
    return arr


def make_matrix(parts, rows_per_part=None):
    """ffemb_matrix_t from one [B, ld] tensor or a list of per-part base pointers/tensors."""
    m = Matrix()
    if isinstance(parts, torch.Tensor):
        parts = [parts]
    first = parts[0]
    if isinstance(first, torch.Tensor):
        for p in parts:
            _require_cuda(p, torch.float32, "matrix part")
            if p.dim() != 2 or p.stride(1) != 1:
                raise ValueError("matrix parts must be 2-D with unit column stride")
        m.ld = first.stride(0)
        ptrs = [p.data_ptr() for p in parts]
        rows = first.shape[0]
    else:  # (ptr list, ld) given explicitly
        ptrs, m.ld = parts[0], parts[1]
        rows = rows_per_part
    for r, p in enumerate(ptrs):
        m.base[r] = p
    m.num_parts = len(ptrs)
    m.rows_per_part = rows_per_part if rows_per_part is not None else rows
    return m


def embed_forward(lookups, batch, out_dim, aggr, out, err_flag=None, stream=None):
    """ffemb_embed_forward; `out` is a Matrix or a [B, ld] float32 cuda tensor."""
    mat = out if isinstance(out, Matrix) else make_matrix(out)
    err = C.c_void_p(err_flag.data_ptr()) if err_flag is not None else None
    check(_lib.lib().ffemb_embed_forward(lookups, len(lookups), batch, out_dim, int(aggr),
                                         C.byref(mat), err, _stream_ptr(stream)))


def backward_workspace_bytes(total_indices, total_entries, out_dim):
    return _lib.lib().ffemb_embed_backward_workspace_bytes(total_indices, total_entries, out_dim)


def alloc_workspace(total_indices, total_entries, out_dim, device=None):
    n = backward_workspace_bytes(total_indices, total_entries, out_dim)
    return torch.empty(n + 256, dtype=torch.uint8, device=device or "cuda")


def _ws_ptr(ws):
    p = (ws.data_ptr() + 255) & ~255
    return C.c_void_p(p), ws.numel() - (p - ws.data_ptr())


def embed_backward_sgd(lookups, batch, out_dim, aggr, grad, lr, workspace, err_flag=None,
                       stream=None):
    """ffemb_embed_backward_sgd: deterministic backward + sparse SGD on touched rows."""
    mat = grad if isinstance(grad, Matrix) else make_matrix(grad)
    err = C.c_void_p(err_flag.data_ptr()) if err_flag is not None else None
    p, n = _ws_ptr(workspace)
    check(_lib.lib().ffemb_embed_backward_sgd(lookups, len(lookups), batch, out_dim, int(aggr),
                                              C.byref(mat), lr, p, n, err, _stream_ptr(stream)))


def embed_backward_sgd_static(lookups, batch, out_dim, aggr, grad, lr, err_flag=None, stream=None):
    """ffemb_embed_backward_sgd_static: the same step without a workspace argument.

    For hosts that are not allowed to allocate device memory (no cudaMalloc, no
    torch.empty): the scratch is a fixed 8 MiB array inside libffemb.so.  Larger
    batches raise FFEmbError(status 5).  Calls on one device share that array,
    so run them one after the other on one stream; make the first call on a
    device outside stream capture.
    """
    mat = grad if isinstance(grad, Matrix) else make_matrix(grad)
    err = C.c_void_p(err_flag.data_ptr()) if err_flag is not None else None
    check(_lib.lib().ffemb_embed_backward_sgd_static(lookups, len(lookups), batch, out_dim, int(aggr),
                                                     C.byref(mat), lr, err, _stream_ptr(stream)))


def static_scratch_bytes():
    return _lib.lib().ffemb_static_scratch_bytes()


def embed_backward_sort(lookups, batch, out_dim, workspace, err_flag=None, stream=None):
    err = C.c_void_p(err_flag.data_ptr()) if err_flag is not None else None
    p, n = _ws_ptr(workspace)
    check(_lib.lib().ffemb_embed_backward_sort(lookups, len(lookups), batch, out_dim, p, n, err,
                                               _stream_ptr(stream)))


def embed_sgd_update(lookups, batch, out_dim, aggr, grad, lr, workspace, stream=None):
    mat = grad if isinstance(grad, Matrix) else make_matrix(grad)
    p, n = _ws_ptr(workspace)
    check(_lib.lib().ffemb_embed_sgd_update(lookups, len(lookups), batch, out_dim, int(aggr),
                                            C.byref(mat), lr, p, n, _stream_ptr(stream)))


def uniform_init(weight, seed, min_val, max_val, stream=None):
    _require_cuda(weight, torch.float32, "weight")
    check(_lib.lib().ffemb_uniform_init(C.c_void_p(weight.data_ptr()), weight.numel(), seed,
                                        min_val, max_val, _stream_ptr(stream)))


def set_sort_impl(impl):
    """0: CUDA toolkit DeviceRadixSort; 1: the library's own segmented per-table sort."""
    check(_lib.lib().ffemb_set_sort_impl(int(impl)))


def get_sort_impl():
    return _lib.lib().ffemb_get_sort_impl()


def kernel_launch_count():
    return _lib.lib().ffemb_kernel_launch_count()
