"""The library's own key sorts -- 2: the one-launch on-chip cluster sort
(csrc/ffemb_csort.cuh, the default), 1: the segmented sort through global
memory (csrc/ffemb_sort.cuh) -- against the toolkit sort (0).  All are stable
sorts of the same keys, so the backward + SGD results must be bit-identical;
with an out-of-range index in the batch the
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
    ([40000000, 7, 30000], 16, 65536, 1, 21),     # full 4-CTA clusters, 3 / 1 / 2 passes
    ([700, 200000], 32, 16385, 1, 21),            # one lookup more than a CTA's slab: 2-CTA clusters
    ([5000, 123], 16, 8192, 8, 22),               # bags of 8: 65536 lookups per table
    ([150000000], 4, 4096, 1, 21),                # 28 key bits: 4 passes
    ([900, 70000], 16, 8192, 9, 21),              # 73728 lookups per table: too big for a cluster -> falls back
])
@pytest.mark.parametrize("impl", [1, 2])
def test_own_sort_gives_the_same_tables_bit_for_bit(K, rows, D, B, L, aggr, impl):
    gen = torch.Generator(device="cuda")
    gen.manual_seed(len(rows) * 1000 + D)
    Ws = [torch.randn(n, D, device="cuda", generator=gen) for n in rows]
    idxs = [torch.randint(0, n, (B, L), device="cuda", generator=gen) for n in rows]
    g = torch.randn(B, len(rows) * D, device="cuda", generator=gen)
    a, ea = _step(K, 0, Ws, idxs, g, aggr, B)
    b, eb = _step(K, impl, Ws, idxs, g, aggr, B)
    assert ea == eb == 0
    assert any(not torch.equal(x, w) for x, w in zip(a, Ws))
    for t, (x, y) in enumerate(zip(a, b)):
        assert torch.equal(x, y), "table %d" % t


@pytest.mark.parametrize("impl", [1, 2])
def test_own_sort_ragged_bags_and_invalid_indices(K, impl):
    gen = torch.Generator(device="cuda")
    gen.manual_seed(5)
    B, D, N = 777, 64, 5000
    lengths = torch.randint(0, 9, (B,), device="cuda", generator=gen)
    off = torch.cat([torch.zeros(1, dtype=torch.int64, device="cuda"), lengths.cumsum(0)])
    idx = torch.randint(0, N, (int(off[-1].item()),), device="cuda", generator=gen)
    W = torch.randn(N, D, device="cuda", generator=gen)
    g = torch.randn(B, D, device="cuda", generator=gen)
    a, _ = _step(K, 0, [W], [idx], g, 22, B, offsets=[off])
    b, _ = _step(K, impl, [W], [idx], g, 22, B, offsets=[off])
    assert torch.equal(a[0], b[0])
    # ragged bags spread over a 4-CTA cluster (about 40k lookups), two tables
    B2 = 9000
    offs, idxs2, Ws2 = [], [], []
    for n_rows in (33, 2000000):
        ln = torch.randint(0, 10, (B2,), device="cuda", generator=gen)
        ln[0] = 0
        ln[B2 - 1] = 0
        offs.append(torch.cat([torch.zeros(1, dtype=torch.int64, device="cuda"), ln.cumsum(0)]))
        idxs2.append(torch.randint(0, n_rows, (int(offs[-1][-1].item()),), device="cuda", generator=gen))
        Ws2.append(torch.randn(n_rows, 16, device="cuda", generator=gen))
    g2 = torch.randn(B2, 32, device="cuda", generator=gen)
    a2, _ = _step(K, 0, Ws2, idxs2, g2, 22, B2, offsets=offs)
    b2, _ = _step(K, impl, Ws2, idxs2, g2, 22, B2, offsets=offs)
    assert all(torch.equal(x, y) for x, y in zip(a2, b2))
    # invalid indices: flagged by both, skipped by both; later tables may differ in the last bit
    rows = [5000, 300]
    Ws = [torch.randn(n, D, device="cuda", generator=gen) for n in rows]
    idxs = [torch.randint(0, n, (512, 1), device="cuda", generator=gen) for n in rows]
    idxs[0][0, 0] = rows[0]
    idxs[1][3, 0] = -7
    g = torch.randn(512, 2 * D, device="cuda", generator=gen)
    a, ea = _step(K, 0, Ws, idxs, g, 21, 512)
    b, eb = _step(K, impl, Ws, idxs, g, 21, 512)
    assert ea == eb == 3
    for x, y in zip(a, b):
        np.testing.assert_allclose(x.cpu().numpy(), y.cpu().numpy(), rtol=1e-5, atol=1e-5)
    c, _ = _step(K, impl, Ws, idxs, g, 21, 512)
    assert all(torch.equal(x, y) for x, y in zip(b, c))       # and it repeats itself exactly


def test_wide_key_path_matches_the_oracle(K, oracle):
    """Launch groups whose tables hold 2^32-1 rows or more in all use 64-bit keys
    (make_keys<u64>, the library sort on u64, reduce_kernel<u64>).  Two tables
    DECLARED with 2^31 + k rows (backed by small allocations; the indices stay
    inside them) take that path; one index is out of range so that the
    all-ones invalid key is sorted and skipped too."""
    gen = torch.Generator(device="cuda")
    gen.manual_seed(21)
    D, B, L, real = 32, 700, 2, 3000
    declared = [(1 << 31) + 5, (1 << 31) + 9]
    Ws = [torch.randn(real, D, device="cuda", generator=gen) for _ in declared]
    idx = [torch.randint(0, real, (B, L), device="cuda", generator=gen) for _ in declared]
    idx[0][:50] = 7                                   # a long run
    idx[1][3, 1] = declared[1]                        # out of range even for the declared size
    g = torch.randn(B, 2 * D, device="cuda", generator=gen)
    W0 = [w.cpu().numpy().copy() for w in Ws]
    lk = K.make_lookups(Ws, idx)
    for t, n in enumerate(declared):
        lk[t].num_entries = n
    ws = K.alloc_workspace(2 * B * L, sum(declared), D)
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    K.embed_backward_sgd(lk, B, D, 21, g, 0.05, ws, err)
    torch.cuda.synchronize()
    assert int(err.item()) == 3                       # FFEMB_ERR_INDEX_OUT_OF_RANGE, flagged not applied
    g_h = g.cpu().numpy()
    for t in range(2):
        i_h = idx[t].cpu().numpy().copy()
        gt = np.ascontiguousarray(g_h[:, t * D:(t + 1) * D]).copy()
        if t == 1:                                    # the oracle refuses bad indices: drop that lookup's share
            i_h[3, 1] = 0
            want_fix = gt[3].copy()
        _, want = oracle.step(W0[t], i_h, gt, 0.05)
        if t == 1:
            want[0] += np.float32(0.05) * want_fix    # undo the stand-in lookup of row 0
        np.testing.assert_allclose(Ws[t].cpu().numpy(), want, rtol=1e-5, atol=2e-5)
