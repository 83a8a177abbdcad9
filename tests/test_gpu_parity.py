"""GPU parity tests: the CUDA path, called through the C-ABI, against the CPU
oracle (which is pinned to the reference's own CPU kernel, tests/test_oracle.py).

Bars (BASELINE.json north_star):
  * index handling and forward: BIT-EXACT against the reference's CPU path;
  * tables after backward + SGD: within 1e-5 (rtol = atol = 1e-5, the
    reference test-suite's tolerance, src/ops/tests/test_harness.py:44-60);
    bit-exact whenever every row's lookups fall inside one 32-position chunk
    of the sorted order (always true for all-distinct rows).
"""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden", "embedding_avx2_golden.npz")
SUM, AVG = 21, 22
TOL = dict(rtol=1e-5, atol=1e-5)


@pytest.fixture(scope="module")
def K():
    if not torch.cuda.is_available():
        pytest.fail("these tests need a CUDA device; there is no CPU fallback to test")
    from dlrm_flexflow_b200 import kernels
    return kernels


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def run_forward(K, Ws, idxs, aggr, offsets=None, B=None, concat=True):
    T, D = len(Ws), Ws[0].shape[1]
    dW = [dev(w) for w in Ws]
    dI = [dev(i) for i in idxs]
    dO = [dev(o) for o in offsets] if offsets is not None else None
    B = B if B is not None else idxs[0].shape[0]
    out = torch.full((B, T * D), float("nan"), device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    lk = K.make_lookups(dW, dI, dO)
    K.embed_forward(lk, B, D, aggr, out, err)
    torch.cuda.synchronize()
    return out.cpu().numpy(), int(err.item())


def run_step(K, Ws, idxs, g, aggr, lr, offsets=None, B=None, repeat=1):
    """backward + sparse SGD on copies of Ws; g is [B, T*D]."""
    T, D = len(Ws), Ws[0].shape[1]
    B = B if B is not None else idxs[0].shape[0]
    dI = [dev(i) for i in idxs]
    dO = [dev(o) for o in offsets] if offsets is not None else None
    dg = dev(g)
    n = sum(i.size for i in idxs)
    ws = K.alloc_workspace(n, sum(w.shape[0] for w in Ws), D)
    results = []
    for _ in range(repeat):
        dW = [dev(w) for w in Ws]
        err = torch.zeros(1, dtype=torch.int32, device="cuda")
        lk = K.make_lookups(dW, dI, dO)
        K.embed_backward_sgd(lk, B, D, aggr, dg, lr, ws, err)
        torch.cuda.synchronize()
        results.append(([w.cpu().numpy() for w in dW], int(err.item())))
    return results if repeat > 1 else results[0]


def oracle_step(O, Ws, idxs, g, aggr, lr, lengths=None):
    D = Ws[0].shape[1]
    outs = []
    for t, (W, idx) in enumerate(zip(Ws, idxs)):
        gt = np.ascontiguousarray(g[:, t * D:(t + 1) * D])
        _, Wn = O.step(W, idx, gt, lr, lengths=None if lengths is None else lengths[t], aggr=aggr)
        outs.append(Wn)
    return outs


# ---------------------------------------------------------------------------
# forward
# ---------------------------------------------------------------------------
def _gold():
    z = np.load(GOLD)
    return [(str(n), z[str(n) + "_W"], z[str(n) + "_idx"], z[str(n) + "_len"], z[str(n) + "_sum"],
             z[str(n) + "_avg"]) for n in z["names"]]


@pytest.mark.parametrize("case", _gold(), ids=lambda c: c[0])
def test_forward_bit_exact_vs_reference_golden(K, case):
    """Vectors computed by the reference's embedding_avx2.cc."""
    name, W, idx, lengths, want_sum, want_avg = case
    B = len(lengths)
    off = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int64)
    for aggr, want in ((SUM, want_sum), (AVG, want_avg)):
        got, err = run_forward(K, [W], [idx], aggr, offsets=[off], B=B)   # CSR path
        assert err == 0 and got.tobytes() == want.tobytes(), (name, aggr, "csr")
        if not name.endswith("ragged"):                                    # fixed-bag path
            L = int(lengths[0])
            got, err = run_forward(K, [W], [idx.reshape(B, L)], aggr)
            assert err == 0 and got.tobytes() == want.tobytes(), (name, aggr, "fixed")


@pytest.mark.parametrize("D", [4, 8, 16, 32, 64, 128, 20, 7, 100, 256])
@pytest.mark.parametrize("L", [1, 2, 5, 32, 37])
def test_forward_bit_exact_vs_oracle(K, oracle, D, L):
    rng = np.random.default_rng(100 * D + L)
    T, B = 3, 301                       # B not a multiple of 32: ragged last warp
    Ns = [997, 5, 40000]
    Ws = [rng.normal(size=(n, D)).astype(np.float32) for n in Ns]
    idxs = [rng.integers(0, n, size=(B, L)) for n in Ns]
    for aggr in (SUM, AVG):
        got, err = run_forward(K, Ws, idxs, aggr)
        assert err == 0
        for t in range(T):
            want = oracle.forward(Ws[t], idxs[t], aggr=aggr)
            assert got[:, t * D:(t + 1) * D].tobytes() == want.tobytes(), (D, L, aggr, t)


def test_forward_many_tables_crosses_launch_groups(K, oracle):
    """More than 64 tables: several batched launches."""
    rng = np.random.default_rng(11)
    T, B, D = 70, 64, 16
    Ws = [rng.normal(size=(50 + t, D)).astype(np.float32) for t in range(T)]
    idxs = [rng.integers(0, 50 + t, size=(B, 1)) for t in range(T)]
    got, err = run_forward(K, Ws, idxs, SUM)
    assert err == 0
    for t in range(T):
        assert got[:, t * D:(t + 1) * D].tobytes() == oracle.forward(Ws[t], idxs[t]).tobytes()


def test_forward_split_matrix_parts(K, oracle):
    """The output matrix split into sample blocks with separate base pointers
    (what the sharded path uses to write straight into each rank's buffer)."""
    rng = np.random.default_rng(12)
    T, B, D, P = 2, 256, 64, 4
    Ws = [rng.normal(size=(300, D)).astype(np.float32) for _ in range(T)]
    idxs = [rng.integers(0, 300, size=(B, 3)) for _ in range(T)]
    parts = [torch.zeros(B // P, T * D, device="cuda") for _ in range(P)]
    lk = K.make_lookups([dev(w) for w in Ws], [dev(i) for i in idxs])
    K.embed_forward(lk, B, D, AVG, K.make_matrix(parts, rows_per_part=B // P))
    torch.cuda.synchronize()
    got = torch.cat(parts).cpu().numpy()
    for t in range(T):
        assert got[:, t * D:(t + 1) * D].tobytes() == oracle.forward(Ws[t], idxs[t], aggr=AVG).tobytes()


def test_forward_empty_and_out_of_range(K, oracle):
    W = np.arange(40, dtype=np.float32).reshape(10, 4)
    # empty bags -> zeros, also with AVG
    off = np.array([0, 0, 2, 2], dtype=np.int64)
    got, err = run_forward(K, [W], [np.array([1, 2], dtype=np.int64)], AVG, offsets=[off], B=3)
    assert err == 0
    assert got.tobytes() == oracle.forward(W, np.array([1, 2]), lengths=[0, 2, 0], aggr=AVG).tobytes()
    # out-of-range and negative indices: flagged, contribute nothing, no fault
    idx = np.array([[0, 10], [-1, 3], [2, 2]], dtype=np.int64)
    got, err = run_forward(K, [W], [idx], SUM)
    assert err == 3                                   # FFEMB_ERR_INDEX_OUT_OF_RANGE
    np.testing.assert_array_equal(got, np.stack([W[0], W[3], W[2] + W[2]]))
    got, err = run_forward(K, [W], [np.array([[11], [1]], dtype=np.int64)], SUM)
    assert err == 3
    np.testing.assert_array_equal(got, np.stack([np.zeros(4, np.float32), W[1]]))


# ---------------------------------------------------------------------------
# backward + fused sparse SGD
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("D", [4, 16, 64, 128, 20, 200])
@pytest.mark.parametrize("L,aggr", [(1, SUM), (3, AVG), (32, AVG)])
def test_step_unique_rows_bit_exact(K, oracle, D, L, aggr):
    """All lookups distinct: every run is one position long."""
    rng = np.random.default_rng(D + L)
    T, B = 2, 130
    Ns = [B * L + 7, 3 * B * L]
    Ws = [rng.normal(size=(n, D)).astype(np.float32) for n in Ns]
    idxs = [rng.permutation(n)[:B * L].reshape(B, L).astype(np.int64) for n in Ns]
    g = rng.normal(size=(B, T * D)).astype(np.float32)
    got, err = run_step(K, Ws, idxs, g, aggr, 0.01)
    want = oracle_step(oracle, Ws, idxs, g, aggr, 0.01)
    assert err == 0
    for t in range(T):
        assert got[t].tobytes() == want[t].tobytes(), (D, L, t)


@pytest.mark.parametrize("D", [16, 64, 128, 24])
@pytest.mark.parametrize("N,L,aggr", [(5000, 1, SUM), (700, 4, AVG), (7, 1, SUM), (3, 8, AVG), (1, 2, SUM)])
def test_step_with_duplicates_within_tolerance(K, oracle, D, N, L, aggr):
    """Duplicate rows, from a few per row to runs thousands long that span
    many chunks (N = 1, 3, 7 -- MLPerf-Terabyte really has 3- and 4-row tables)."""
    rng = np.random.default_rng(N * 7 + D)
    T, B = 3, 2048
    Ws = [rng.normal(size=(N + t, D)).astype(np.float32) for t in range(T)]
    idxs = [rng.integers(0, N + t, size=(B, L)) for t in range(T)]
    g = rng.normal(size=(B, T * D)).astype(np.float32)
    got, err = run_step(K, Ws, idxs, g, aggr, 0.01)
    want = oracle_step(oracle, Ws, idxs, g, aggr, 0.01)
    assert err == 0
    for t in range(T):
        np.testing.assert_allclose(got[t], want[t], **TOL)


@pytest.mark.parametrize("D,L,aggr,B", [(128, 1, SUM, 40001), (64, 3, AVG, 40001), (16, 1, SUM, 40001),
                                        (20, 1, SUM, 40001), (128, 1, SUM, 160001), (64, 4, AVG, 100001)])
def test_step_very_long_runs_cross_many_blocks(K, oracle, D, L, aggr, B):
    """Two-row tables: each row collects tens of thousands of lookups, i.e. a run that spans many level-1
    blocks (like the 3- and 4-row tables of the MLPerf-Terabyte shape).  Chains of more than 32 partials
    go to the block-level kernel reduce_l3: ~80 partials in the 40001-bag cases at D = 128, 300-400 in
    the large ones; the D = 16 case (2048-position blocks) stays a short chain for reduce_l2."""
    rng = np.random.default_rng(D * 10 + L)
    T = 2
    Ws = [rng.normal(size=(2, D)).astype(np.float32) for _ in range(T)]
    idxs = [rng.integers(0, 2, size=(B, L)) for _ in range(T)]
    g = (0.05 * rng.normal(size=(B, T * D))).astype(np.float32)
    got, err = run_step(K, Ws, idxs, g, aggr, 0.01)
    want = oracle_step(oracle, Ws, idxs, g, aggr, 0.01)
    assert err == 0
    for t in range(T):
        assert not np.array_equal(got[t], Ws[t])
        np.testing.assert_allclose(got[t], want[t], **TOL)


def test_step_short_runs_inside_one_chunk_are_bit_exact(K, oracle):
    """Each touched row has exactly 4 lookups, and runs are aligned so none
    crosses a 32-position chunk: sequential sum == CPU order, bit for bit."""
    rng = np.random.default_rng(21)
    D, B = 64, 1024
    rows = np.repeat(rng.permutation(5000)[:B // 4], 4)     # 256 rows x 4
    idx = rows[rng.permutation(B)].reshape(B, 1).astype(np.int64)
    W = rng.normal(size=(5000, D)).astype(np.float32)
    g = rng.normal(size=(B, D)).astype(np.float32)
    (got,), err = run_step(K, [W], [idx], g, SUM, 0.01)
    (want,) = oracle_step(oracle, [W], [idx], g, SUM, 0.01)
    assert err == 0 and got.tobytes() == want.tobytes()


def test_step_is_deterministic(K):
    """Two runs on identical inputs: bit-identical tables (no atomics)."""
    rng = np.random.default_rng(22)
    D, B, L, T = 64, 4096, 4, 4
    Ws = [rng.normal(size=(50 * (t + 1), D)).astype(np.float32) for t in range(T)]
    idxs = [rng.zipf(1.3, size=(B, L)) % (50 * (t + 1)) for t in range(T)]
    g = rng.normal(size=(B, T * D)).astype(np.float32)
    (a, _), (b, _), (c, _) = run_step(K, Ws, idxs, g, AVG, 0.05, repeat=3)
    for t in range(T):
        assert a[t].tobytes() == b[t].tobytes() == c[t].tobytes()


def test_step_ragged_bags(K, oracle):
    rng = np.random.default_rng(23)
    D, B, N = 32, 500, 300
    lengths = rng.integers(0, 9, size=B)
    lengths[:3] = 0
    off = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int64)
    idx = rng.integers(0, N, size=int(off[-1])).astype(np.int64)
    W = rng.normal(size=(N, D)).astype(np.float32)
    g = rng.normal(size=(B, D)).astype(np.float32)
    for aggr in (SUM, AVG):
        (got,), err = run_step(K, [W], [idx], g, aggr, 0.01, offsets=[off], B=B)
        (want,) = oracle_step(oracle, [W], [idx], g, aggr, 0.01, lengths=[lengths])
        assert err == 0
        np.testing.assert_allclose(got, want, **TOL)


def test_step_out_of_range_is_flagged_and_skipped(K, oracle):
    rng = np.random.default_rng(24)
    D, B, N = 16, 64, 100
    W = rng.normal(size=(N, D)).astype(np.float32)
    idx = rng.permutation(N)[:B].reshape(B, 1).astype(np.int64)
    g = rng.normal(size=(B, D)).astype(np.float32)
    bad = idx.copy()
    bad[5, 0] = N
    bad[9, 0] = -3
    (got,), err = run_step(K, [W], [bad], g, SUM, 0.01)
    assert err == 3
    g2 = g.copy()
    g2[[5, 9]] = 0                       # as if those two lookups did not exist
    ok = idx.copy()
    (want,) = oracle_step(oracle, [W], [ok], g2, SUM, 0.01)
    want[idx[5, 0]] = W[idx[5, 0]]
    want[idx[9, 0]] = W[idx[9, 0]]
    assert got.tobytes() == want.tobytes()


def test_sort_update_split_equals_fused(K):
    rng = np.random.default_rng(25)
    D, B, L = 64, 512, 2
    W = rng.normal(size=(100, D)).astype(np.float32)
    idx = rng.integers(0, 100, size=(B, L))
    g = dev(rng.normal(size=(B, D)).astype(np.float32))
    ws = K.alloc_workspace(B * L, 100, D)
    w1, w2 = dev(W), dev(W)
    di = dev(idx)
    K.embed_backward_sgd(K.make_lookups([w1], [di]), B, D, AVG, g, 0.1, ws)
    lk = K.make_lookups([w2], [di])
    K.embed_backward_sort(lk, B, D, ws)
    K.embed_sgd_update(lk, B, D, AVG, g, 0.1, ws)
    torch.cuda.synchronize()
    assert torch.equal(w1, w2) and not torch.equal(w1, dev(W))


def test_workspace_too_small_is_an_error(K):
    from dlrm_flexflow_b200 import FFEmbError
    w = torch.zeros(10, 16, device="cuda")
    idx = torch.zeros(64, 1, dtype=torch.int64, device="cuda")
    g = torch.zeros(64, 16, device="cuda")
    small = torch.empty(512, dtype=torch.uint8, device="cuda")
    with pytest.raises(FFEmbError) as e:
        K.embed_backward_sgd(K.make_lookups([w], [idx]), 64, 16, SUM, g, 0.1, small)
    assert e.value.status == 5


def test_uniform_init_bit_exact_vs_oracle(K, oracle):
    w = torch.empty(100003, device="cuda")
    K.uniform_init(w, 1234, -0.003, 0.003)
    torch.cuda.synchronize()
    assert w.cpu().numpy().tobytes() == oracle.uniform_init(100003, 1234, -0.003, 0.003).tobytes()


# ---------------------------------------------------------------------------
# model-level API (the reference's Python surface)
# ---------------------------------------------------------------------------
def test_model_api_small_dlrm_config(oracle):
    """BASELINE config C1: 8 tables x 100k x 64, bag 1, SUM, batch 256, driven
    the way examples/cpp/DLRM/dlrm.cc:166-187 drives the reference."""
    import dlrm_flexflow_b200 as ff
    rng = np.random.default_rng(31)
    T, N, D, B, lr = 8, 100000, 64, 256, 0.01
    cfg = ff.FFConfig()
    cfg.set_batch_size(B)
    m = ff.FFModel(cfg)
    ins, outs = [], []
    for t in range(T):
        x = m.create_tensor([B, 1], ff.DataType.DT_INT64)
        r = (1.0 / N) ** 0.5                                        # dlrm.cc:44
        outs.append(m.embedding(x, N, D, ff.AggrMode.AGGR_MODE_SUM, None,
                                ff.UniformInitializer(1000 + t, -r, r)))
        ins.append(x)
    m.compile(optimizer=ff.SGDOptimizer(m, lr))
    m.init_layers()
    # device-side initialiser == oracle's restatement, bit for bit
    W0 = [m.get_layer_by_id(t).get_weight_tensor().get_weights(m) for t in range(T)]
    for t in range(T):
        r = np.float32((1.0 / N) ** 0.5)
        assert W0[t].tobytes() == oracle.uniform_init(N * D, 1000 + t, -r, r).tobytes()
    for step in range(2):
        idx = [rng.integers(0, N, size=(B, 1)) for _ in range(T)]
        g = [rng.normal(size=(B, D)).astype(np.float32) for _ in range(T)]
        for t in range(T):
            ins[t].set_array(m, idx[t])
        m.forward()
        m.zero_gradients()
        for t in range(T):
            outs[t].set_grad(m, g[t])
        m.backward()
        m.update()
        m.sync()
        for t in range(T):
            out = outs[t].get_array(m)
            want_out, want_W = oracle.step(W0[t], idx[t], g[t], lr)
            assert out.tobytes() == want_out.tobytes()
            got_W = m.get_layer_by_id(t).get_weight_tensor().get_weights(m)
            np.testing.assert_allclose(got_W, want_W, **TOL)
            W0[t] = got_W


def test_model_api_set_weights_attach_and_errors():
    import dlrm_flexflow_b200 as ff
    B, N, D = 64, 500, 16
    cfg = ff.FFConfig()
    cfg.set_batch_size(B)
    m = ff.FFModel(cfg)
    x = m.create_tensor([B, 2], ff.DataType.DT_INT64)
    y = m.embedding(x, N, D, ff.AggrMode.AGGR_MODE_AVG, None, ff.ZeroInitializer())
    m.compile(optimizer=ff.SGDOptimizer(m, 0.5))
    p = m.get_layer_by_id(0).get_weight_tensor()
    assert not p.get_weights(m).any()                               # ZeroInitializer
    W = np.random.default_rng(1).normal(size=(N, D)).astype(np.float32)
    p.set_weights(m, W)
    assert p.get_weights(m).tobytes() == W.tobytes()
    with pytest.raises(AssertionError):
        p.set_weights(m, W[:10])                                    # shape check like the reference
    # zero-copy torch view of the table, attached torch buffers for in/out/grad
    idx = torch.randint(0, N, (B, 2), device="cuda")
    out = torch.zeros(B, D, device="cuda")
    grad = torch.ones(B, D, device="cuda")
    x.attach_torch(idx)
    y.attach_torch(out)
    y.attach_torch(grad, grad=True)
    m.forward()
    m.sync()
    Wt = p.as_torch()
    assert torch.equal(out, Wt[idx].sum(1) * (1.0 / 2))
    before = Wt.clone()
    m.update()                                                      # no backward yet: no-op
    m.sync()
    assert torch.equal(Wt, before)
    m.backward()
    m.update()
    m.sync()
    ref = before.clone()
    ref.index_put_((idx.reshape(-1),), torch.full((B * 2, D), -0.5 * 0.5, device="cuda"), accumulate=True)
    torch.testing.assert_close(Wt, ref, rtol=1e-5, atol=1e-5)
    # out-of-range index is reported by sync, then cleared
    idx[0, 0] = N
    m.forward()
    with pytest.raises(ff.FFEmbError) as e:
        m.sync()
    assert e.value.status == 3
    idx[0, 0] = 0
    m.forward()
    m.sync()
    # momentum / weight decay: refused, not silently wrong
    m2 = ff.FFModel(cfg)
    x2 = m2.create_tensor([B, 1], ff.DataType.DT_INT64)
    m2.embedding(x2, N, D, ff.AggrMode.AGGR_MODE_SUM)
    m2.compile(optimizer=ff.SGDOptimizer(m2, 0.01, momentum=0.9))
    with pytest.raises(ff.FFEmbError) as e:
        m2.update()
    assert e.value.status == 4


def test_native_library_is_the_one_loaded():
    """The extension in the tree is what these tests exercised."""
    import dlrm_flexflow_b200 as ff
    path = ff.library_path()
    assert os.path.exists(path)
    maps = open("/proc/self/maps").read()
    assert path in maps
    from dlrm_flexflow_b200 import kernels
    assert kernels.kernel_launch_count() > 0
