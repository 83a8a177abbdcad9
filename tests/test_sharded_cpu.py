"""CPU tests of the sharded path's host logic, including a world_size-2 run over
the gloo backend: placement -> staging layout -> all-to-all -> concat layout,
forward and backward, against the single-process oracle.  The compute on each
rank is done by the oracle here (this is a test of the exchange plumbing; the
CUDA kernels are covered by the -m gpu tests)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from dlrm_flexflow_b200.sharded import ShardPlan  # noqa: E402
import dlrm_flexflow_b200 as ff  # noqa: E402


def test_plan_round_robin_and_layout():
    plan = ShardPlan([10] * 26, 128, 65536, 8)
    assert plan.owner == [t % 8 for t in range(26)]                 # dlrm_strategy.cc:252-256
    assert [len(plan.tables_of(r)) for r in range(8)] == [4, 4, 3, 3, 3, 3, 3, 3]
    assert abs(plan.imbalance_bound() - 26 / 32.0) < 1e-12            # 81 %
    assert plan.Bs == 8192 and plan.global_col(5) == 5 * 128
    assert plan.local_col(9) == 1 * 128 and plan.local_col(1) == 0    # rank 1 owns tables 1, 9, 17, 25
    assert plan.forward_send_splits(0) == [8192 * 4 * 128] * 8
    assert plan.forward_recv_splits(3) == [8192 * len(plan.tables_of(r)) * 128 for r in range(8)]
    assert plan.backward_send_splits(2) == plan.forward_recv_splits(2)
    # C3, N=8: 95.4 MB leave each 3-table rank... (B*T_r*D*4*(N-1)/N)
    assert plan.exchange_bytes_per_rank(0) == 65536 * 4 * 128 * 4 * 7 // 8
    with pytest.raises(ValueError):
        ShardPlan([10] * 4, 16, 30, 4)                                # batch not divisible
    with pytest.raises(ValueError):
        ShardPlan([10] * 4, 16, 32, 4, owners=[0, 1, 2, 7])


def test_plan_replicates_small_tables():
    from dlrm_flexflow_b200.workloads import TERABYTE_ROWS
    plan = ShardPlan(TERABYTE_ROWS, 128, 65536, 8, replicate_below="auto")     # rule: N_t < B/N = 8192
    assert plan.replicate_below == 8192
    assert plan.replicated == [t for t, n in enumerate(TERABYTE_ROWS) if n < 8192] and len(plan.replicated) == 13
    assert all(plan.owner[t] == -1 for t in plan.replicated)
    assert all(plan.owner[t] == t % 8 for t in range(26) if t not in plan.replicated)
    assert max(len(plan.tables_of(r)) for r in range(8)) == 2           # exchange volume halves (4 -> 2 tables)
    assert plan.replica_allreduce_bytes() == sum(TERABYTE_ROWS[t] for t in plan.replicated) * 128 * 4 < 11e6
    assert ShardPlan(TERABYTE_ROWS, 128, 65536, 1, replicate_below="auto").replicated == []   # one GPU: nothing
    assert ShardPlan(TERABYTE_ROWS, 128, 65536, 8).replicated == []                           # default: off


def test_plan_follows_a_strategy_file():
    s = ff.Strategy.dlrm_simple(8, 16)           # the reference's 16embs_8gpus fixture
    plan = ShardPlan([100] * 16, 64, 1024, 8, strategy=s)
    assert plan.owner == [i % 8 for i in range(16)]
    plan4 = ShardPlan([100] * 16, 64, 1024, 4, strategy=s)            # folded: % devices.size()
    assert plan4.owner == [(i % 8) % 4 for i in range(16)]
    s2 = ff.Strategy()
    for i, d in enumerate([3, 3, 0, 1]):
        s2.add_op("embedding%d" % i, 0, [1, 1], [d])
    assert ShardPlan([5] * 4, 8, 8, 4, strategy=s2).owner == [3, 3, 0, 1]


def test_pack_unpack_roundtrip_numpy():
    plan = ShardPlan([7] * 5, 4, 12, 3, owners=[2, 0, 2, 1, 0])
    rng = np.random.default_rng(0)
    concat = rng.normal(size=(plan.Bs, plan.T * plan.D)).astype(np.float32)
    send = np.zeros(plan.Bs * plan.T * plan.D, np.float32)
    plan.pack_backward(concat, send)
    back = np.zeros_like(concat)
    plan.unpack_forward(send, back)             # same block layout in both directions
    assert back.tobytes() == concat.tobytes()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, owners, seed, result_dir):
    import torch
    import torch.distributed as dist
    from oracle import oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        T, D, B, L, lr, aggr = len(owners), 8, 16, 3, 0.05, O.AGGR_MODE_AVG
        rows = [11 + 3 * t for t in range(T)]
        plan = ShardPlan(rows, D, B, world, owners=owners)
        rng = np.random.default_rng(seed)                        # same stream on every rank
        Ws = [rng.normal(size=(rows[t], D)).astype(np.float32) for t in range(T)]
        idx = [rng.integers(0, rows[t], size=(B, L)) for t in range(T)]
        g_full = rng.normal(size=(B, T * D)).astype(np.float32)
        mine = plan.tables_of(rank)
        w, Bs = plan.width(rank), plan.Bs
        # forward on owned tables, whole batch, into the staging layout [q][Bs][T_r*D]
        send = np.zeros(world * Bs * w, np.float32)
        for t in mine:
            out_t = O.forward(Ws[t], idx[t], aggr=aggr)         # [B, D]
            c = plan.local_col(t)
            for q in range(world):
                blk = send[q * Bs * w:(q + 1) * Bs * w].reshape(Bs, w)
                blk[:, c:c + D] = out_t[q * Bs:(q + 1) * Bs]
        recv = torch.zeros(Bs * T * D)
        dist.all_to_all_single(recv, torch.from_numpy(send), plan.forward_recv_splits(rank),
                               plan.forward_send_splits(rank))
        out_local = np.zeros((Bs, T * D), np.float32)
        plan.unpack_forward(recv.numpy(), out_local)
        want = np.concatenate([O.forward(Ws[t], idx[t], aggr=aggr) for t in range(T)], axis=1)
        assert out_local.tobytes() == want[rank * Bs:(rank + 1) * Bs].tobytes()
        # backward: this rank holds the gradient of its samples, owners need all samples
        bsend = np.zeros(Bs * T * D, np.float32)
        plan.pack_backward(g_full[rank * Bs:(rank + 1) * Bs], bsend)
        brecv = torch.zeros(world * Bs * w)
        dist.all_to_all_single(brecv, torch.from_numpy(bsend), plan.backward_recv_splits(rank),
                               plan.backward_send_splits(rank))
        brecv = brecv.numpy()
        for t in mine:
            c = plan.local_col(t)
            g_t = np.concatenate([brecv[r * Bs * w:(r + 1) * Bs * w].reshape(Bs, w)[:, c:c + D]
                                  for r in range(world)])
            assert g_t.tobytes() == np.ascontiguousarray(g_full[:, t * D:(t + 1) * D]).tobytes()
            _, Wn = O.step(Ws[t], idx[t], g_t, lr, aggr=aggr)
            np.save(os.path.join(result_dir, "W%d.npy" % t), Wn)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("owners", [[0, 1, 0, 1, 0], [1, 1, 1, 0, 1], [0, 0, 0, 0, 0]])
def test_world2_gloo_exchange_matches_single_process(tmp_path, owners):
    import torch.multiprocessing as mp
    from oracle import oracle as O
    O.build()
    world, seed = 2, 77
    mp.spawn(_worker, args=(world, _free_port(), owners, seed, str(tmp_path)), nprocs=world, join=True)
    # single-process reference of the same step
    T, D, B, L, lr = len(owners), 8, 16, 3, 0.05
    rows = [11 + 3 * t for t in range(T)]
    rng = np.random.default_rng(seed)
    Ws = [rng.normal(size=(rows[t], D)).astype(np.float32) for t in range(T)]
    idx = [rng.integers(0, rows[t], size=(B, L)) for t in range(T)]
    g_full = rng.normal(size=(B, T * D)).astype(np.float32)
    for t in range(T):
        _, Wn = O.step(Ws[t], idx[t], np.ascontiguousarray(g_full[:, t * D:(t + 1) * D]), lr,
                       aggr=O.AGGR_MODE_AVG)
        got = np.load(os.path.join(str(tmp_path), "W%d.npy" % t))
        assert got.tobytes() == Wn.tobytes()
