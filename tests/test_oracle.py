"""CPU tests of the oracle itself: it must reproduce the reference's own CPU
kernel (golden vectors made from it, and the kernel itself when oracle/_ref is
present) before anything is allowed to be compared against it."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden", "embedding_avx2_golden.npz")


def _cases():
    z = np.load(GOLD)
    for name in z["names"]:
        name = str(name)
        yield name, z[name + "_W"], z[name + "_idx"], z[name + "_len"], z[name + "_sum"], z[name + "_avg"]


def test_golden_file_covers_all_paths():
    names = [c[0] for c in _cases()]
    for d in (128, 64, 32, 16, 20):           # three AVX2 specialisations + generic path
        for kind in ("L1", "L4", "L32", "ragged"):
            assert "D%d_%s" % (d, kind) in names


@pytest.mark.parametrize("case", list(_cases()), ids=lambda c: c[0])
def test_oracle_forward_matches_reference_golden(oracle, case):
    """Bit-exact against vectors produced by embedding_avx2.cc:5-296."""
    _, W, idx, lengths, want_sum, want_avg = case
    got = oracle.forward(W, idx, lengths=lengths, aggr=oracle.AGGR_MODE_SUM)
    assert got.tobytes() == want_sum.tobytes()
    got = oracle.forward(W, idx, lengths=lengths, aggr=oracle.AGGR_MODE_AVG)
    assert got.tobytes() == want_avg.tobytes()


def test_oracle_forward_matches_reference_kernel_live(oracle):
    """Same, against the reference kernel itself on fresh random shapes."""
    if not oracle.have_ref():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    rng = np.random.default_rng(7)
    for D in (128, 64, 32, 16, 8, 4, 20, 7, 100):
        for L in (1, 2, 5, 32, 100):
            N, B = 503, 67
            W = rng.normal(size=(N, D)).astype(np.float32)
            idx = rng.integers(0, N, size=(B, L))
            for aggr in (oracle.AGGR_MODE_SUM, oracle.AGGR_MODE_AVG):
                a = oracle.forward(W, idx, aggr=aggr)
                b = oracle.forward_ref(W, idx, np.full(B, L, np.int32), aggr)
                assert a.tobytes() == b.tobytes(), (D, L, aggr)


BWD_GOLD = os.path.join(os.path.dirname(__file__), "golden", "embedding_bwd_golden.npz")


def _bwd_cases():
    z = np.load(BWD_GOLD)
    for name in z["names"]:
        name = str(name)
        yield (name,) + tuple(z[name + s] for s in ("_idx", "_g1", "_g2", "_G1", "_G2"))


@pytest.mark.parametrize("case", list(_bwd_cases()), ids=lambda c: c[0])
def test_oracle_backward_matches_reference_golden(oracle, case):
    """Bit-exact against vectors produced by the reference's own CPU backward,
    embed_backward_generic (src/ops/embedding.cc:54-72): bag size 1, SUM, a
    zeroed gradient and a second call accumulating on top of the first."""
    name, idx, g1, g2, G1, G2 = case
    N, D = G1.shape
    got = oracle.backward_dense(N, D, idx.reshape(-1, 1), g1)
    assert got.tobytes() == G1.tobytes()
    got = oracle.backward_dense(N, D, idx[::-1].reshape(-1, 1), g2, gradW=got)
    assert got.tobytes() == G2.tobytes()


def test_oracle_backward_matches_reference_function_live(oracle):
    """Same, against the compiled reference function itself on fresh shapes."""
    if not oracle.have_ref_bwd():
        pytest.skip("oracle/_ref/libembed_backward_generic.so not built (no /root/reference here)")
    rng = np.random.default_rng(11)
    for D in (128, 64, 32, 16, 4, 20, 7):
        for N, B in ((3, 500), (97, 300), (4099, 128)):
            idx = rng.integers(0, N, size=B)
            g = rng.normal(size=(B, D)).astype(np.float32)
            a = oracle.backward_dense(N, D, idx.reshape(-1, 1), g)
            b = oracle.backward_ref(N, D, idx, g)
            assert a.tobytes() == b.tobytes(), (D, N, B)


def test_oracle_forward_fixed_equals_lengths(oracle):
    rng = np.random.default_rng(1)
    W = rng.normal(size=(50, 16)).astype(np.float32)
    idx = rng.integers(0, 50, size=(9, 3))
    a = oracle.forward(W, idx)
    b = oracle.forward(W, idx.reshape(-1), lengths=np.full(9, 3))
    assert a.tobytes() == b.tobytes()


def test_oracle_rejects_out_of_range(oracle):
    W = np.zeros((4, 4), np.float32)
    with pytest.raises(IndexError):
        oracle.forward(W, np.array([[4]]))
    with pytest.raises(IndexError):
        oracle.forward(W, np.array([[-1]]))
    with pytest.raises(IndexError):
        oracle.backward_dense(4, 4, np.array([[7]]), np.zeros((1, 4), np.float32))


def test_reference_gpu_avg_differs_from_cpu_avg(oracle):
    """the CUDA kernel divides inside the bag loop
    (embedding.cu:187-195); the CPU path multiplies the sum by 1/len.  For
    {1,2,4}: GPU 1.5926, CPU 2.3333.  We follow the CPU path."""
    W = np.array([[1.0], [2.0], [4.0]], np.float32)
    out = oracle.forward(W, np.array([[0, 1, 2]]), aggr=oracle.AGGR_MODE_AVG)
    assert out[0, 0] == np.float32(7.0) * (np.float32(1.0) / np.float32(3.0))
    gpu = np.float32(0)
    for v in (1.0, 2.0, 4.0):
        gpu = np.float32((gpu + np.float32(v)) / np.float32(3))
    assert abs(float(gpu) - 1.5926) < 1e-4 and abs(float(out[0, 0]) - 2.3333) < 1e-4


@pytest.mark.parametrize("aggr,mode", [(21, "sum"), (22, "mean")])
def test_oracle_step_matches_torch(oracle, aggr, mode):
    """Secondary cross-check: torch EmbeddingBag + SGD(lr=0.01) at
    the reference test-suite's tolerance rtol=atol=1e-5."""
    import torch
    rng = np.random.default_rng(3)
    N, D, B, L = 97, 24, 40, 6
    W = rng.normal(size=(N, D)).astype(np.float32)
    idx = rng.integers(0, 11, size=(B, L))   # few rows: heavy duplication
    g = rng.normal(size=(B, D)).astype(np.float32)
    out, Wn = oracle.step(W, idx, g, 0.01, aggr=aggr)
    eb = torch.nn.EmbeddingBag(N, D, mode=mode)
    eb.weight.data = torch.from_numpy(W.copy())
    opt = torch.optim.SGD(eb.parameters(), lr=0.01)
    o = eb(torch.from_numpy(idx))
    o.backward(torch.from_numpy(g))
    opt.step()
    np.testing.assert_allclose(out, o.detach().numpy(), rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(Wn, eb.weight.detach().numpy(), rtol=1e-5, atol=1e-5)


def test_oracle_touched_rows_equals_dense(oracle):
    """The touched-rows CPU step used for the timed baseline is bit-identical
    to the reference's dense zero + backward + sgd_update (wd = m = 0)."""
    rng = np.random.default_rng(5)
    N, D, B, L = 64, 16, 128, 4
    W = rng.normal(size=(N, D)).astype(np.float32)
    idx = rng.integers(0, N, size=(B, L))
    g = rng.normal(size=(B, D)).astype(np.float32)
    _, dense = oracle.step(W, idx, g, 0.01, aggr=oracle.AGGR_MODE_AVG)
    W2 = W.copy()
    gradW = np.zeros_like(W)
    out = np.empty((B, D), np.float32)
    oracle.step_tables_mt([W2], [idx], [g], L, oracle.AGGR_MODE_AVG, 0.01, [gradW], [out], 2)
    assert W2.tobytes() == dense.tobytes()
    assert not gradW.any()
    assert out.tobytes() == oracle.forward(W, idx, aggr=oracle.AGGR_MODE_AVG).tobytes()


def test_oracle_sgd_momentum_follows_reference_formula(oracle):
    """optimizer_kernel.cu:23-41 with momentum/nesterov/weight decay."""
    rng = np.random.default_rng(9)
    W = rng.normal(size=100).astype(np.float32)
    G = rng.normal(size=100).astype(np.float32)
    V = rng.normal(size=100).astype(np.float32)
    lr, wd, m = np.float32(0.1), np.float32(0.01), np.float32(0.9)
    W0, V0 = W.copy(), V.copy()
    oracle.sgd_update_dense(W, G, lr, wd, m, True, V)
    gt = G.astype(np.float64) + wd * W0
    v = V0 * np.float64(m) + gt
    want = W0 - lr * (gt + m * v)
    np.testing.assert_allclose(W, want, rtol=1e-5, atol=1e-6)


def test_oracle_uniform_init_range_and_determinism(oracle):
    a = oracle.uniform_init(100000, 42, -0.25, 0.5)
    b = oracle.uniform_init(100000, 42, -0.25, 0.5)
    c = oracle.uniform_init(100000, 43, -0.25, 0.5)
    assert a.tobytes() == b.tobytes() and a.tobytes() != c.tobytes()
    assert a.min() > -0.25 and a.max() <= 0.5          # U(min, max]
    assert abs(a.mean() - 0.125) < 5e-3 and abs(a.std() - 0.75 / 12 ** 0.5) < 5e-3
