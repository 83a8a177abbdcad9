"""GPU tests of the NCCL all-to-all baseline of the table-wise sharded path
(dlrm_flexflow_b200/sharded.py; the peer-memory engine behind the operator API
is covered by test_gpu_model_sharded.py).  World size 1 runs on any GPU box;
the 2-rank runs need two GPUs (skipped otherwise) and check, on every rank,
bit-identical forward outputs to the single-GPU / oracle result (SURVEY 8c rule
4) and tables within 1e-5."""
import os
import socket
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _problem(seed, T=5, D=64, B=256, L=2):
    rows = [40 + 300 * t for t in range(T)]
    rng = np.random.default_rng(seed)
    idx = [rng.integers(0, rows[t], size=(B, L)) for t in range(T)]
    g = rng.normal(size=(B, T * D)).astype(np.float32)
    return rows, D, B, L, idx, g


def _run_rank(rank, world, exchange, owners, seed, lr=0.05, aggr=22, replicate_below=0):
    from oracle import oracle as O
    from dlrm_flexflow_b200.sharded import ShardPlan, ShardedEmbeddingBags
    rows, D, B, L, idx, g = _problem(seed)
    plan = ShardPlan(rows, D, B, world, owners=owners, replicate_below=replicate_below)
    sh = ShardedEmbeddingBags(plan, rank, L, aggr, lr, exchange=exchange)
    mine = plan.tables_of(rank)
    W0 = {t: sh.tables[i].cpu().numpy() for i, t in enumerate(mine)}
    rep = plan.replicated
    W0.update({t: sh.rep_tables[i].cpu().numpy() for i, t in enumerate(rep)})
    for t in mine:                                   # initialiser parity
        r = np.float32((1.0 / rows[t]) ** 0.5)
        assert W0[t].tobytes() == O.uniform_init(rows[t] * D, 1234 + t, -r, r).tobytes()
    didx = [torch.from_numpy(idx[t]).cuda() for t in mine]
    Bs = plan.Bs
    ridx = [torch.from_numpy(np.ascontiguousarray(idx[t][rank * Bs:(rank + 1) * Bs])).cuda() for t in rep] or None
    for step in range(2):
        out = sh.forward(didx, ridx)
        sh.check_errors()
        # every table's initial weights are reproducible from the seed on any rank
        Wall = []
        for t in range(len(rows)):
            r = np.float32((1.0 / rows[t]) ** 0.5)
            Wall.append(O.uniform_init(rows[t] * D, 1234 + t, -r, r).reshape(rows[t], D))
        if step == 0:
            want = np.concatenate([O.forward(Wall[t], idx[t], aggr=aggr) for t in range(len(rows))], axis=1)
            assert out.cpu().numpy().tobytes() == want[rank * Bs:(rank + 1) * Bs].tobytes()
        sh.grad.copy_(torch.from_numpy(g[rank * Bs:(rank + 1) * Bs]).cuda())
        sh.backward_update(didx, ridx)
        sh.check_errors()
        for tabs, ids in ((sh.tables, mine), (sh.rep_tables, rep)):
            for i, t in enumerate(ids):
                _, Wn = O.step(W0[t], idx[t], np.ascontiguousarray(g[:, t * D:(t + 1) * D]), lr, aggr=aggr)
                np.testing.assert_allclose(tabs[i].cpu().numpy(), Wn, rtol=1e-5, atol=1e-5)
                W0[t] = tabs[i].cpu().numpy()
        if rep and world > 1:      # replicas stay bit-identical across ranks
            import torch.distributed as dist
            mine_sum = torch.stack([w.double().sum() for w in sh.rep_tables])
            all_sums = [torch.empty_like(mine_sum) for _ in range(world)]
            dist.all_gather(all_sums, mine_sum)
            for other in all_sums:
                assert torch.equal(other, mine_sum)
    sh.close()


def test_world1_sharded_wrapper():
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device")
    _run_rank(0, 1, "nccl", [0] * 5, seed=5)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, exchange, owners, seed, replicate_below=0):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        _run_rank(rank, world, exchange, owners, seed, replicate_below=replicate_below)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("owners", [[0, 1, 0, 1, 0], [1, 1, 1, 1, 0]])
def test_world2_nccl(owners):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), "nccl", owners, 9), nprocs=2, join=True)


def test_peer_memory_exchange_is_behind_the_operator_api():
    from dlrm_flexflow_b200.sharded import ShardPlan, ShardedEmbeddingBags
    with pytest.raises(ValueError):
        ShardedEmbeddingBags(ShardPlan([10, 10], 16, 8, 1), 0, 1, 21, 0.1, exchange="p2p")


def test_world2_with_replicated_small_tables():
    """Tables 0 and 1 (40 and 340 rows) are kept on both ranks (data-parallel,
    gradient replicas summed by all-reduce), the rest are sharded."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), "nccl", [0, 1, 0, 1, 0], 11, 500), nprocs=2, join=True)
