"""Full-size checks at BASELINE.json's configurations, through properties that
do not need the (seconds-per-step) CPU oracle at that size:

  * one-hot forward is a pure row gather: compare bit-for-bit with an
    independent gather (torch index_select) -- C3, 96 GB of tables;
  * multi-hot AVG forward against torch's embedding_bag at 1e-5 (its
    summation order is unspecified, so not bitwise) -- C4;
  * the update is linear in the gradient: a step with +g followed by a step
    with -g on the same indices restores every table (checksums), and the
    column sums of a step's table change equal -lr * column sums of the
    scattered gradient -- C5 Zipf;
  * determinism at full size: the same step from the same (re-seeded) state
    gives the same fp64 checksum twice.

Indices/gradients are generated on the GPU (torch is plumbing here); tables
come from the library's own seeded initialiser.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu
SUM, AVG = 21, 22


@pytest.fixture(scope="module")
def K():
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device")
    from dlrm_flexflow_b200 import kernels
    return kernels


def _free_gb():
    free, _ = torch.cuda.mem_get_info()
    return free / 2 ** 30


def _tables(K, rows, D, seed=1234):
    Ws = []
    for t, n in enumerate(rows):
        w = torch.empty((n, D), device="cuda")
        r = (1.0 / n) ** 0.5
        K.uniform_init(w, seed + t, -r, r)
        Ws.append(w)
    return Ws


def _checksum(Ws):
    return [float(w.double().sum().item()) for w in Ws]


def test_c3_terabyte_forward_is_an_exact_gather(K):
    from dlrm_flexflow_b200.workloads import WORKLOADS, gen_indices
    w = WORKLOADS["C3"]
    if _free_gb() < 110:
        pytest.skip("needs ~100 GB of free HBM")
    rows, D, B = w["rows"], w["dim"], w["batch"]
    T = len(rows)
    Ws = _tables(K, rows, D)
    gen = torch.Generator(device="cuda")
    gen.manual_seed(1)
    idx = [gen_indices(torch, rows[t], B, 1, "uniform", 0.0, gen) for t in range(T)]
    out = torch.empty((B, T * D), device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    K.embed_forward(K.make_lookups(Ws, idx), B, D, SUM, out, err)
    torch.cuda.synchronize()
    assert int(err.item()) == 0
    for t in range(T):
        want = Ws[t].index_select(0, idx[t].reshape(-1)) + 0.0      # 0 + w, as the CPU path
        assert torch.equal(out[:, t * D:(t + 1) * D], want), "table %d" % t
    del Ws, out
    torch.cuda.empty_cache()


def test_c4_multihot_avg_forward_vs_torch(K):
    from dlrm_flexflow_b200.workloads import WORKLOADS, gen_indices
    w = WORKLOADS["C4"]
    if _free_gb() < 50:
        pytest.skip("needs ~45 GB of free HBM")
    rows, D, B, L = w["rows"], w["dim"], w["batch"], w["bag"]
    T = len(rows)
    Ws = _tables(K, rows, D)
    gen = torch.Generator(device="cuda")
    gen.manual_seed(2)
    idx = [gen_indices(torch, rows[t], B, L, "uniform", 0.0, gen) for t in range(T)]
    out = torch.empty((B, T * D), device="cuda")
    K.embed_forward(K.make_lookups(Ws, idx), B, D, AVG, out)
    torch.cuda.synchronize()
    for t in range(T):
        want = torch.nn.functional.embedding_bag(idx[t], Ws[t], mode="mean")
        torch.testing.assert_close(out[:, t * D:(t + 1) * D], want, rtol=1e-5, atol=1e-6)
        # and exactly: the first bag summed in bag order in fp32, times 1/L
        acc = torch.zeros(D, device="cuda")
        for j in range(L):
            acc = acc + Ws[t][idx[t][0, j]]
        assert torch.equal(out[0, t * D:(t + 1) * D], acc * (1.0 / L))
    del Ws, out
    torch.cuda.empty_cache()


@pytest.mark.parametrize("alpha", [0.0, 1.05])
def test_c5_update_linearity_and_determinism(K, alpha):
    from dlrm_flexflow_b200.workloads import WORKLOADS, gen_indices
    w = WORKLOADS["C5"]
    if _free_gb() < 24:
        pytest.skip("needs ~20 GB of free HBM")
    rows, D, B = w["rows"], w["dim"], w["batch"]
    T, lr = len(rows), 0.01
    gen = torch.Generator(device="cuda")
    gen.manual_seed(3)
    idx = [gen_indices(torch, rows[t], B, 1, "zipf" if alpha else "uniform", alpha, gen) for t in range(T)]
    g = torch.randn((B, T * D), device="cuda", generator=gen)
    ws = K.alloc_workspace(T * B, sum(rows), D)
    err = torch.zeros(1, dtype=torch.int32, device="cuda")

    def one_step(Ws, grad):
        K.embed_backward_sgd(K.make_lookups(Ws, idx), B, D, SUM, grad, lr, ws, err)
        torch.cuda.synchronize()

    Ws = _tables(K, rows, D)
    before = _checksum(Ws)
    col_before = [w.double().sum(0) for w in Ws[:4]]
    one_step(Ws, g)
    first = _checksum(Ws)
    # column sums move by -lr * (sum of the gradients scattered into the table)
    for t in range(4):
        delta = Ws[t].double().sum(0) - col_before[t]
        want = -lr * g[:, t * D:(t + 1) * D].double().sum(0)
        torch.testing.assert_close(delta, want, rtol=1e-3, atol=2e-4)
    neg = -g
    one_step(Ws, neg)                                        # undo
    after = _checksum(Ws)
    for b, a in zip(before, after):
        assert abs(a - b) <= 1e-3 * max(1.0, abs(b)), (a, b)
    assert any(abs(f - b) > 1e-6 for f, b in zip(first, before))          # the step did something
    # determinism: same state, same inputs -> same bits
    del Ws
    Ws2 = _tables(K, rows, D)
    one_step(Ws2, g)
    assert _checksum(Ws2) == first
    assert int(err.item()) == 0
    del Ws2
    torch.cuda.empty_cache()
