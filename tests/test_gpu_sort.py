"""The library's own key sort (csrc/ffemb_sort.cuh: segmented per table, only as
many radix passes as each table's size needs) against the toolkit sort that is
the default.  Both are stable sorts of the same keys, so the backward + SGD
results must be bit-identical; with an out-of-range index in the batch the
invalid key sits at the end of its table's segment instead of the end of the
whole array, which shifts later tables by one position against the 32-position
chunks -- still deterministic and within the usual 1e-5."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture()
def K():
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device")
    from dlrm_flexflow_b200 import kernels
    default = kernels.get_sort_impl()
    yield kernels
    kernels.set_sort_impl(default)


def _step(K, impl, Ws, idxs, g, aggr, B, offsets=None):
    K.set_sort_impl(impl)
    assert K.get_sort_impl() == impl
    D = Ws[0].shape[1]
    dW = [w.clone() for w in Ws]
    ws = K.alloc_workspace(sum(i.numel() for i in idxs), sum(w.shape[0] for w in Ws), D)
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    K.embed_backward_sgd(K.make_lookups(dW, idxs, offsets), B, D, aggr, g, 0.05, ws, err)
    torch.cuda.synchronize()
    return dW, int(err.item())


@pytest.mark.parametrize("rows,D,B,L,aggr", [
    ([5000], 64, 512, 1, 21),
    ([5000, 300], 64, 512, 1, 21),
    ([300, 65536, 9], 32, 1000, 5, 22),          # 1, 2 and 1 radix passes in one launch
    ([3, 1], 64, 4096, 1, 21),                    # everything is one long run
    ([3, 1, 700, 5000, 100000, 70000, 2, 513, 511, 512], 64, 4096, 1, 21),
    ([1000000, 17, 5000000], 128, 8192, 1, 21),   # 3 passes for the big tables
    ([2000] * 70, 16, 333, 2, 21),                # more than 64 tables: two launch groups
])
def test_own_sort_gives_the_same_tables_bit_for_bit(K, rows, D, B, L, aggr):
    gen = torch.Generator(device="cuda")
    gen.manual_seed(len(rows) * 1000 + D)
    Ws = [torch.randn(n, D, device="cuda", generator=gen) for n in rows]
    idxs = [torch.randint(0, n, (B, L), device="cuda", generator=gen) for n in rows]
    g = torch.randn(B, len(rows) * D, device="cuda", generator=gen)
    a, ea = _step(K, 0, Ws, idxs, g, aggr, B)
    b, eb = _step(K, 1, Ws, idxs, g, aggr, B)
    assert ea == eb == 0
    assert any(not torch.equal(x, w) for x, w in zip(a, Ws))
    for t, (x, y) in enumerate(zip(a, b)):
        assert torch.equal(x, y), "table %d" % t


def test_own_sort_ragged_bags_and_invalid_indices(K):
    gen = torch.Generator(device="cuda")
    gen.manual_seed(5)
    B, D, N = 777, 64, 5000
    lengths = torch.randint(0, 9, (B,), device="cuda", generator=gen)
    off = torch.cat([torch.zeros(1, dtype=torch.int64, device="cuda"), lengths.cumsum(0)])
    idx = torch.randint(0, N, (int(off[-1].item()),), device="cuda", generator=gen)
    W = torch.randn(N, D, device="cuda", generator=gen)
    g = torch.randn(B, D, device="cuda", generator=gen)
    a, _ = _step(K, 0, [W], [idx], g, 22, B, offsets=[off])
    b, _ = _step(K, 1, [W], [idx], g, 22, B, offsets=[off])
    assert torch.equal(a[0], b[0])
    # invalid indices: flagged by both, skipped by both; later tables may differ in the last bit
    rows = [5000, 300]
    Ws = [torch.randn(n, D, device="cuda", generator=gen) for n in rows]
    idxs = [torch.randint(0, n, (512, 1), device="cuda", generator=gen) for n in rows]
    idxs[0][0, 0] = rows[0]
    idxs[1][3, 0] = -7
    g = torch.randn(512, 2 * D, device="cuda", generator=gen)
    a, ea = _step(K, 0, Ws, idxs, g, 21, 512)
    b, eb = _step(K, 1, Ws, idxs, g, 21, 512)
    assert ea == eb == 3
    for x, y in zip(a, b):
        np.testing.assert_allclose(x.cpu().numpy(), y.cpu().numpy(), rtol=1e-5, atol=1e-5)
    c, _ = _step(K, 1, Ws, idxs, g, 21, 512)
    assert all(torch.equal(x, y) for x, y in zip(b, c))       # and it repeats itself exactly
