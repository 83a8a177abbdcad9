"""The allocation-free backward entry point (ffemb_embed_backward_sgd_static).

Some users of the library run where device allocation is off limits -- a
plugin inside a framework that owns all GPU memory, a serving loop that must
never touch the allocator -- so they cannot hand ffemb_embed_backward_sgd a
workspace.  The static variant uses an 8 MiB array built into libffemb.so
instead.  These tests call it through the C-ABI, twice in a row on one device
(the second call finds the first call's leftovers in the shared array), and
require the tables to match what the workspace-taking entry point produces
from the same inputs, at the tolerance the other backward tests use
(rtol = atol = 1e-5), and what the CPU oracle computes.  They also check the
property the entry point exists for -- repeated calls leave the amount of
device memory in use exactly where it was -- and both sides of its size
limit.  The file name keeps it last in collection order so a
`pytest -x` run has exercised everything else before it gets here; it also
passes on its own in a fresh process.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

SUM, AVG = 21, 22
TOL = dict(rtol=1e-5, atol=1e-5)


@pytest.fixture(scope="module")
def K():
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device; there is no CPU fallback to test")
    from dlrm_flexflow_b200 import kernels
    return kernels


def _inputs(seed, T, D, B, L, hot):
    """hot=True: tiny tables, so rows repeat thousands of times and the chain
    lists / partial sums in the scratch array are really used."""
    rng = np.random.default_rng(seed)
    rows = [(5 + t) if hot else (B * L * 3 + 17 * t) for t in range(T)]
    Ws = [rng.normal(size=(n, D)).astype(np.float32) for n in rows]
    idx = [rng.integers(0, n, size=(B, L)).astype(np.int64) for n in rows]
    g = rng.normal(size=(B, T * D)).astype(np.float32)
    return Ws, idx, g


def _both(K, Ws, idx, g, aggr, lr):
    """Tables after one step: (workspace entry point, static entry point)."""
    T, D = len(Ws), Ws[0].shape[1]
    B = idx[0].shape[0]
    di = [torch.from_numpy(i).cuda() for i in idx]
    dg = torch.from_numpy(g).cuda()
    n = sum(i.size for i in idx)
    ws = K.alloc_workspace(n, sum(w.shape[0] for w in Ws), D)
    wa = [torch.from_numpy(w).cuda() for w in Ws]
    wb = [torch.from_numpy(w).cuda() for w in Ws]
    err = torch.zeros(2, dtype=torch.int32, device="cuda")
    K.embed_backward_sgd(K.make_lookups(wa, di), B, D, aggr, dg, lr, ws, err[0:1])
    K.embed_backward_sgd_static(K.make_lookups(wb, di), B, D, aggr, dg, lr, err[1:2])
    torch.cuda.synchronize()
    assert err.tolist() == [0, 0]
    return [w.cpu().numpy() for w in wa], [w.cpu().numpy() for w in wb]


def test_static_entry_matches_workspace_entry_twice_in_a_row(K):
    # call 1: hot rows (long chains), AVG; call 2 right after it: different
    # shape, distinct rows, SUM -- it must not be disturbed by call 1's leftovers
    for seed, (T, D, B, L, hot, aggr) in enumerate([(3, 64, 4096, 2, True, AVG),
                                                     (2, 128, 1024, 1, False, SUM)]):
        Ws, idx, g = _inputs(100 + seed, T, D, B, L, hot)
        ref, got = _both(K, Ws, idx, g, aggr, 0.05)
        for t in range(T):
            assert not np.array_equal(got[t], Ws[t])            # the update happened
            np.testing.assert_allclose(got[t], ref[t], **TOL)


def test_static_entry_matches_the_oracle(K, oracle):
    T, D, B, L = 3, 64, 2048, 2
    Ws, idx, g = _inputs(31, T, D, B, L, True)
    _, got = _both(K, Ws, idx, g, AVG, 0.05)
    for t in range(T):
        _, want = oracle.step(Ws[t], idx[t], g[:, t * D:(t + 1) * D], 0.05, aggr=oracle.AGGR_MODE_AVG)
        np.testing.assert_allclose(got[t], want, **TOL)


def test_static_entry_allocates_no_device_memory(K):
    """What the entry point is for: after the first call on the device (which
    looks the array's address up) further calls must not move the driver's
    free-memory figure nor torch's allocator."""
    T, D, B, L = 2, 128, 8192, 1
    Ws, idx, g = _inputs(5, T, D, B, L, False)
    di = [torch.from_numpy(i).cuda() for i in idx]
    dg = torch.from_numpy(g).cuda()
    w = [torch.from_numpy(x).cuda() for x in Ws]
    lk = K.make_lookups(w, di)
    K.embed_backward_sgd_static(lk, B, D, SUM, dg, 0.01)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    alloc0, reserved0 = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    for _ in range(5):
        K.embed_backward_sgd_static(lk, B, D, SUM, dg, 0.01)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == (alloc0, reserved0)


def test_static_entry_size_limit_both_sides(K):
    """The limit is the exact scratch layout: workspace_bytes() minus the 1 MiB
    of slack that function adds.  The largest batch that fits runs and matches
    the workspace entry point; one past the limit is refused."""
    from dlrm_flexflow_b200 import FFEmbError
    D, rows = 16, 1000
    need = lambda b: K.backward_workspace_bytes(b, rows, D) - (1 << 20)
    lo, hi = 1024, 4 << 20
    assert need(lo) <= K.static_scratch_bytes() < need(hi)
    while hi - lo > 1:
        mid = (lo + hi) // 2
        lo, hi = (mid, hi) if need(mid) <= K.static_scratch_bytes() else (lo, mid)
    gen = torch.Generator(device="cuda")
    gen.manual_seed(3)
    for B, fits in ((lo, True), (hi, False)):
        idx = torch.randint(0, rows, (B, 1), device="cuda", generator=gen)
        g = torch.randn(B, D, device="cuda", generator=gen)
        w0 = torch.randn(rows, D, device="cuda", generator=gen)
        wa, wb = w0.clone(), w0.clone()
        if fits:
            K.embed_backward_sgd(K.make_lookups([wa], [idx]), B, D, SUM, g, 0.01, K.alloc_workspace(B, rows, D))
            K.embed_backward_sgd_static(K.make_lookups([wb], [idx]), B, D, SUM, g, 0.01)
            torch.cuda.synchronize()
            np.testing.assert_allclose(wb.cpu().numpy(), wa.cpu().numpy(), **TOL)
        else:
            with pytest.raises(FFEmbError) as e:
                K.embed_backward_sgd_static(K.make_lookups([wb], [idx]), B, D, SUM, g, 0.01)
            assert e.value.status == 5
            assert torch.equal(wb, w0)


def test_static_entry_repeats_itself_bit_for_bit(K):
    Ws, idx, g = _inputs(7, 2, 32, 2048, 4, True)
    _, a = _both(K, Ws, idx, g, SUM, 0.01)
    _, b = _both(K, Ws, idx, g, SUM, 0.01)
    for x, y in zip(a, b):
        assert x.tobytes() == y.tobytes()


def test_static_entry_refuses_batches_beyond_its_array(K):
    from dlrm_flexflow_b200 import FFEmbError
    assert K.static_scratch_bytes() == 8 << 20
    D, B = 128, 65536 * 8                   # needs far more than 8 MiB of scratch
    assert K.backward_workspace_bytes(B, 1000, D) > K.static_scratch_bytes()
    w = torch.zeros(1000, D, device="cuda")
    idx = torch.zeros(B, 1, dtype=torch.int64, device="cuda")
    g = torch.zeros(B, D, device="cuda")
    with pytest.raises(FFEmbError) as e:
        K.embed_backward_sgd_static(K.make_lookups([w], [idx]), B, D, SUM, g, 0.1)
    assert e.value.status == 5 and "built-in array" in str(e.value)
    assert not w.any()                      # nothing ran
