"""ctypes front-end to the CPU oracle.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this module.  The shipped package
(dlrm_flexflow_b200) never does: it has no CPU fallback.

Two shared objects are wrapped:

* ``oracle/libffemb_oracle.so``  -- our C restatement (oracle/ffemb_oracle.c).
* ``oracle/_ref/libembedding_avx2.so`` -- the reference's own
  ``EmbeddingLookup_int64_t_float_float__avx2_fma``
  (src/ops/embedding_avx2.cc:5-296) compiled in place from /root/reference by
  oracle/Makefile.  Present whenever it was built in the authoring container;
  it travels to the GPU box with the snapshot (git-ignored, not gpurun-ignored).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libffemb_oracle.so")
_REF = os.path.join(_HERE, "_ref", "libembedding_avx2.so")
_REF_BWD = os.path.join(_HERE, "_ref", "libembed_backward_generic.so")
_REF_SYM = "_Z45EmbeddingLookup_int64_t_float_float__avx2_fmaiiiiPKfPKlPKiS0_bPf"

AGGR_MODE_SUM = 21  # include/ffconst.h:11-15
AGGR_MODE_AVG = 22

_f32p = C.POINTER(C.c_float)
_i64p = C.POINTER(C.c_int64)
_i32p = C.POINTER(C.c_int)


def build(force=False):
    """Compile the oracle (and oracle/_ref when /root/reference is present)."""
    if force or not os.path.exists(_LIB) or (
            os.path.getmtime(_LIB) < os.path.getmtime(os.path.join(_HERE, "ffemb_oracle.c"))):
        subprocess.check_call(["make", "-s", "-C", _HERE])
    elif (not os.path.exists(_REF) or not os.path.exists(_REF_BWD)) and \
            os.path.exists("/root/reference/src/ops/embedding_avx2.cc"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "ref"])


def _fp(a):
    return a.ctypes.data_as(_f32p)


def _ip(a):
    return a.ctypes.data_as(_i64p)


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB)
        L.ffo_embed_forward.restype = C.c_int
        L.ffo_embed_forward.argtypes = [C.c_int, C.c_int64, C.c_int64, _f32p, _i64p, _i32p,
                                        C.c_int, C.c_int, _f32p, C.c_int64]
        L.ffo_embed_forward_mt.restype = C.c_int
        L.ffo_embed_forward_mt.argtypes = [C.c_int, C.c_int64, C.c_int64, _f32p, _i64p, C.c_int,
                                           C.c_int, _f32p, C.c_int64, C.c_int]
        L.ffo_embed_backward_dense.restype = C.c_int
        L.ffo_embed_backward_dense.argtypes = [C.c_int, C.c_int64, C.c_int64, _i64p, _i32p,
                                               C.c_int, C.c_int, _f32p, C.c_int64, _f32p]
        L.ffo_sgd_update_dense.restype = None
        L.ffo_sgd_update_dense.argtypes = [C.c_int64, C.c_float, C.c_float, C.c_float, C.c_int,
                                           _f32p, _f32p, _f32p]
        L.ffo_embed_backward_sgd_touched.restype = C.c_int
        L.ffo_embed_backward_sgd_touched.argtypes = [C.c_int, C.c_int64, C.c_int64, _i64p,
                                                     C.c_int, C.c_int, _f32p, C.c_int64,
                                                     C.c_float, _f32p, _f32p]
        L.ffo_step_tables_mt.restype = C.c_int
        L.ffo_step_tables_mt.argtypes = [C.c_int, C.c_int, C.c_int64, _i64p,
                                         C.POINTER(_f32p), C.POINTER(_i64p), C.c_int, C.c_int,
                                         C.POINTER(_f32p), C.POINTER(_f32p), C.c_float,
                                         C.POINTER(_f32p), C.c_int]
        L.ffo_uniform_init.restype = None
        L.ffo_uniform_init.argtypes = [_f32p, C.c_int64, C.c_int, C.c_float, C.c_float]
        L.ffo_max_threads.restype = C.c_int
        _lib = L
    return _lib


def have_ref():
    build()
    return os.path.exists(_REF)


def ref():
    """The reference's own AVX2 kernel (raises if oracle/_ref was never built)."""
    global _ref
    if _ref is None:
        if not have_ref():
            raise RuntimeError("oracle/_ref/libembedding_avx2.so absent: run `make -C oracle` "
                               "in a container that has /root/reference")
        R = C.CDLL(_REF)
        fn = getattr(R, _REF_SYM)
        fn.restype = None
        fn.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, _f32p, _i64p, _i32p, _f32p,
                       C.c_bool, _f32p]
        _ref = fn
    return _ref


_ref_bwd = None


def have_ref_bwd():
    build()
    return os.path.exists(_REF_BWD)


def backward_ref(N, D, indices, g, gradW=None):
    """The reference's own CPU backward, embed_backward_generic
    (src/ops/embedding.cc:54-72), compiled in place by oracle/Makefile: one
    index per sample (bag size 1), SUM, accumulating into gradW [N, D]."""
    global _ref_bwd
    if _ref_bwd is None:
        if not have_ref_bwd():
            raise RuntimeError("oracle/_ref/libembed_backward_generic.so absent: run `make -C oracle` "
                               "in a container that has /root/reference")
        fn = C.CDLL(_REF_BWD).embed_backward_generic
        fn.restype = None
        fn.argtypes = [_i64p, _i32p, _f32p, _f32p, C.c_int, C.c_int, C.c_int, C.c_int]
        _ref_bwd = fn
    idx = np.ascontiguousarray(indices, dtype=np.int64).reshape(-1)
    g = np.ascontiguousarray(g, dtype=np.float32)
    B = g.shape[0]
    assert idx.size == B and g.shape[1] == D and B * D < 2 ** 31
    assert idx.size == 0 or (idx.min() >= 0 and idx.max() < N), "the reference does not check"
    if gradW is None:
        gradW = np.zeros((N, D), dtype=np.float32)
    assert gradW.dtype == np.float32 and gradW.flags.c_contiguous
    lengths = np.ones(B, dtype=np.int32)
    _ref_bwd(_ip(idx), lengths.ctypes.data_as(_i32p), _fp(g), _fp(gradW), D, B, B, N)
    return gradW


def _lengths_arg(lengths):
    if lengths is None:
        return None, None
    l32 = np.ascontiguousarray(lengths, dtype=np.int32)
    return l32, l32.ctypes.data_as(_i32p)


def forward(W, indices, lengths=None, aggr=AGGR_MODE_SUM):
    """Oracle forward.  W [N,D] f32; indices int64 [B,L], or flat with `lengths`."""
    W = np.ascontiguousarray(W, dtype=np.float32)
    idx = np.ascontiguousarray(indices, dtype=np.int64)
    N, D = W.shape
    if lengths is None:
        assert idx.ndim == 2, "fixed-size bags need a [B, L] index array"
        B, L = idx.shape
    else:
        B, L = len(lengths), 0
        assert int(np.sum(lengths)) == idx.size
    l32, lp = _lengths_arg(lengths)
    out = np.empty((B, D), dtype=np.float32)
    rc = lib().ffo_embed_forward(D, B, N, _fp(W), _ip(idx), lp, int(L),
                                 int(aggr == AGGR_MODE_AVG), _fp(out), D)
    if rc:
        raise IndexError("embedding index out of range")
    return out


def forward_ref(W, indices, lengths, aggr=AGGR_MODE_SUM):
    """The reference's AVX2 routine, called directly with the true data_size=N
    and real lengths (not through the Legion wrapper, which hard-codes
    data_size=1000000 and lengths=1: src/ops/embedding.cc:114-119)."""
    W = np.ascontiguousarray(W, dtype=np.float32)
    idx = np.ascontiguousarray(indices, dtype=np.int64).reshape(-1)
    l32 = np.ascontiguousarray(lengths, dtype=np.int32)
    N, D = W.shape
    B = l32.size
    assert int(l32.sum()) == idx.size
    assert idx.size == 0 or (idx.min() >= 0 and idx.max() < N), "reference would assert"
    out = np.zeros((B, D), dtype=np.float32)
    ref()(D, B, idx.size, N, _fp(W), _ip(idx), l32.ctypes.data_as(_i32p), None,
          bool(aggr == AGGR_MODE_AVG), _fp(out))
    return out


def backward_dense(N, D, indices, g, lengths=None, aggr=AGGR_MODE_SUM, gradW=None):
    """Dense table gradient, accumulating (reference semantics)."""
    idx = np.ascontiguousarray(indices, dtype=np.int64)
    g = np.ascontiguousarray(g, dtype=np.float32)
    B = g.shape[0]
    L = 0
    if lengths is None:
        assert idx.ndim == 2 and idx.shape[0] == B
        L = idx.shape[1]
    l32, lp = _lengths_arg(lengths)
    if gradW is None:
        gradW = np.zeros((N, D), dtype=np.float32)
    rc = lib().ffo_embed_backward_dense(D, B, N, _ip(idx), lp, int(L), int(aggr), _fp(g),
                                        g.shape[1], _fp(gradW))
    if rc:
        raise IndexError("embedding index out of range")
    return gradW


def sgd_update_dense(W, G, lr, weight_decay=0.0, momentum=0.0, nesterov=False, V=None):
    """In-place dense SGD on W (reference semantics)."""
    assert W.dtype == np.float32 and W.flags.c_contiguous
    G = np.ascontiguousarray(G, dtype=np.float32)
    if V is None:
        V = np.zeros(1, dtype=np.float32) if momentum <= 0 else np.zeros_like(W)
    lib().ffo_sgd_update_dense(W.size, lr, weight_decay, momentum, int(nesterov), _fp(G), _fp(V),
                               _fp(W))
    return W


def step(W, indices, g, lr, lengths=None, aggr=AGGR_MODE_SUM):
    """Reference training step on one table: forward, zero grad, backward,
    dense SGD (wd = momentum = 0).  Returns (out, W_new); W is not modified."""
    out = forward(W, indices, lengths=lengths, aggr=aggr)
    N, D = W.shape
    G = backward_dense(N, D, indices, g, lengths=lengths, aggr=aggr)
    Wn = np.array(W, dtype=np.float32, copy=True)
    sgd_update_dense(Wn, G, lr)
    return out, Wn


def uniform_init(count, seed, min_val, max_val):
    W = np.empty(count, dtype=np.float32)
    lib().ffo_uniform_init(_fp(W), count, seed, min_val, max_val)
    return W


def step_tables_mt(Ws, idxs, gs, L, aggr, lr, gradWs, outs, nthreads):
    """Timed CPU baseline: forward + touched-rows backward/SGD over T tables,
    in place on Ws, OpenMP over tables."""
    T = len(Ws)
    D = Ws[0].shape[1]
    B = gs[0].shape[0]
    Ns = np.array([w.shape[0] for w in Ws], dtype=np.int64)
    FP = _f32p * T
    IP = _i64p * T
    rc = lib().ffo_step_tables_mt(T, D, B, _ip(Ns), FP(*[_fp(w) for w in Ws]),
                                  IP(*[_ip(i) for i in idxs]), L, aggr,
                                  FP(*[_fp(o) for o in outs]), FP(*[_fp(x) for x in gs]), lr,
                                  FP(*[_fp(x) for x in gradWs]), nthreads)
    if rc:
        raise IndexError("embedding index out of range")


def max_threads():
    return lib().ffo_max_threads()
