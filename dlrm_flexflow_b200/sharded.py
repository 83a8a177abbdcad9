"""Table-wise sharding of the embedding tables over the GPUs of one box.

One process per GPU (torch.distributed supplies rendezvous and NCCL).  Table t
lives wholly on rank owner(t) -- the reference's DLRM strategy (embedding<i>
-> device i % ngpu, dims [1,1]: src/runtime/dlrm_strategy.cc:252-256) -- and
that rank processes the WHOLE batch for its tables, while the dense part of
the model is sample-parallel: rank q consumes rows [q*B/N, (q+1)*B/N) of every
table's output, side by side as [B/N][T*D] (what the reference's Concat builds,
src/ops/concat.cu:159-215), and produces the gradient in the same layout.

Two exchanges per step (the reference does them implicitly by Legion DMA
through zero-copy host memory, src/ops/embedding.cu:280-284,343-347).  The
production implementation is the sharded engine inside libffemb.so, behind the
operator API (FFModel with workers_per_node > 1; csrc/ffemb_model.cu): both
directions over peer memory, fused into the kernels.  This module holds

  ShardPlan              the pure host logic (placement, column layout, split
                         sizes, pack/unpack of the NCCL staging layout); runs on
                         CPU and is covered by the world_size-2 gloo tests;
  ShardedEmbeddingBags   the unfused NCCL all-to-all formulation of the same
                         step, kept as the baseline (`bench.py --exchange nccl`).

"""
import ctypes as C

from . import _lib
from ._lib import check, lib


class ShardPlan(object):
    """Which rank owns which table, and which small tables are replicated.

    replicate_below: tables with fewer rows than this are kept on EVERY rank
    (the reference's data-parallel embedding: replicated table, gradient
    replicas summed, src/runtime/model.cc:697-724) instead of sharded.  A
    replica costs a dense all-reduce of N_t*D*4 bytes per step, a shard costs
    two exchanges of B*D*4*(N-1)/N^2 per GPU: replicate iff N_t < B/N
    (hard part 1).  "auto" applies that rule; 0 replicates nothing.
    owner[t] is -1 for a replicated table.
    """

    def __init__(self, rows, dim, global_batch, world, owners=None, strategy=None, replicate_below=0):
        if global_batch % world:
            raise ValueError("global batch %d is not divisible by world size %d" % (global_batch, world))
        self.rows = list(rows)
        self.T, self.D, self.B, self.N = len(rows), dim, global_batch, world
        self.Bs = global_batch // world
        if replicate_below == "auto":
            replicate_below = self.Bs if world > 1 else 0
        self.replicate_below = int(replicate_below) if world > 1 else 0
        self.replicated = [t for t in range(self.T) if self.rows[t] < self.replicate_below]
        if owners is None:
            handle = strategy.handle if strategy is not None else _lib.Handle(None)
            owners = [lib().ffemb_strategy_device_for_table(handle, t, None, world) for t in range(self.T)]
        owners = [-1 if t in self.replicated else o for t, o in enumerate(owners)]
        if len(owners) != self.T or any(not 0 <= o < world for t, o in enumerate(owners)
                                         if t not in self.replicated):
            raise ValueError("owners must name a rank in [0, world) for every table")
        self.owner = list(owners)
        self.local = [[t for t in range(self.T) if self.owner[t] == r] for r in range(world)]

    # ---- layout ---------------------------------------------------------------
    def tables_of(self, rank):
        return self.local[rank]

    def global_col(self, t):
        """First column of table t in the [B/N][T*D] concat matrix."""
        return t * self.D

    def local_col(self, t):
        """First column of table t in its owner's [..][T_r*D] staging block."""
        return self.local[self.owner[t]].index(t) * self.D

    def width(self, rank):
        return len(self.local[rank]) * self.D

    def forward_send_splits(self, rank):
        """Elements rank sends to each peer in the forward exchange (NCCL path)."""
        return [self.Bs * self.width(rank)] * self.N

    def forward_recv_splits(self, rank):
        return [self.Bs * self.width(r) for r in range(self.N)]

    def lookups_per_rank(self, bag):
        return [len(self.local[r]) * self.B * bag for r in range(self.N)]

    def imbalance_bound(self):
        """Best possible scaling efficiency from the integer split of the sharded tables alone."""
        most = max(len(l) for l in self.local)
        sharded = self.T - len(self.replicated)
        return sharded / float(self.N * most) if most else 1.0

    def replica_allreduce_bytes(self):
        """Dense gradient bytes all-reduced per step for the replicated tables."""
        return sum(self.rows[t] for t in self.replicated) * self.D * 4

    def exchange_bytes_per_rank(self, rank):
        """Bytes that leave `rank` over NVLink per direction per step."""
        return self.B * self.width(rank) * 4 * (self.N - 1) // self.N

    # ---- NCCL staging layout <-> concat layout (works on torch and numpy) -----
    def unpack_forward(self, recv_flat, out_concat):
        """recv_flat: blocks [r][B/N][T_r*D] in rank order -> out_concat [B/N][T*D]."""
        pos = 0
        for r in range(self.N):
            w = self.width(r)
            if w:
                block = recv_flat[pos:pos + self.Bs * w].reshape(self.Bs, w)
                for i, t in enumerate(self.local[r]):
                    out_concat[:, t * self.D:(t + 1) * self.D] = block[:, i * self.D:(i + 1) * self.D]
            pos += self.Bs * w
        return out_concat

    def pack_backward(self, grad_concat, send_flat):
        """grad_concat [B/N][T*D] -> blocks [q][B/N][T_q*D] for the owners q."""
        pos = 0
        for q in range(self.N):
            w = self.width(q)
            if w:
                block = send_flat[pos:pos + self.Bs * w].reshape(self.Bs, w)
                for i, t in enumerate(self.local[q]):
                    block[:, i * self.D:(i + 1) * self.D] = grad_concat[:, t * self.D:(t + 1) * self.D]
            pos += self.Bs * w
        return send_flat

    def backward_send_splits(self, rank):
        return [self.Bs * self.width(q) for q in range(self.N)]

    def backward_recv_splits(self, rank):
        return [self.Bs * self.width(rank)] * self.N


class ShardedEmbeddingBags(object):
    """Table-wise sharding with an NCCL all-to-all: the unfused baseline.

    The production path is the sharded engine inside libffemb.so, reached
    through the operator API (FFModel with workers_per_node > 1: forward stores
    and gradient pushes over peer memory, peer-memory barriers, replicas reduced
    by one kernel).  This class keeps the textbook formulation next to it for
    comparison (`bench.py --exchange nccl`): the kernels write / read local
    staging buffers, `all_to_all_single` with uneven splits moves them, and a
    strided copy builds / splits the concat layout.

    forward(indices)         indices[i]: [B, bag] int64 cuda for this rank's
                             i-th table, WHOLE batch.  Returns this rank's
                             [B/N][T*D] slice of all tables' pooled outputs.
    backward_update(indices) consumes self.grad ([B/N][T*D], filled by the
                             caller) on every rank and applies the sparse SGD
                             update to this rank's tables.
    rep_indices              [B/N, bag] indices of this rank's own samples for
                             the replicated tables (plan.replicated), if any.

    With one rank, forward() already sorts the keys of the coming backward; the
    sort is reused when backward_update() gets the same index tensors (compared
    by address), so do not modify them in between -- or pass presort=False.
    """

    def __init__(self, plan, rank, bag, aggr, lr, exchange="nccl", group=None, init_seed=1234, presort=True):
        import torch
        import torch.distributed as dist
        from . import kernels
        if not torch.cuda.is_available():
            raise RuntimeError("ShardedEmbeddingBags needs a CUDA device: there is no CPU fallback")
        if exchange != "nccl":
            raise ValueError("the peer-memory exchange lives behind the operator API (FFModel with "
                             "workers_per_node > 1); this class is the NCCL all-to-all baseline")
        self.plan, self.rank, self.bag, self.aggr, self.lr = plan, rank, bag, aggr, lr
        self.exchange, self.group, self.K, self.torch, self.dist = exchange, group, kernels, torch, dist
        self.tables_idx = plan.tables_of(rank)
        D, Bs, T, N, B = plan.D, plan.Bs, plan.T, plan.N, plan.B
        self.tables = []
        for t in self.tables_idx:
            w = torch.empty((plan.rows[t], D), device="cuda")
            r = (1.0 / plan.rows[t]) ** 0.5
            kernels.uniform_init(w, init_seed + t, -r, r)
            self.tables.append(w)
        n_lookups = len(self.tables_idx) * B * bag
        self.workspace = kernels.alloc_workspace(max(n_lookups, 1), sum(plan.rows[t] for t in self.tables_idx) + 1, D)
        self.err = torch.zeros(1, dtype=torch.int32, device="cuda")
        # Small tables kept on every rank: all replicas in one flat buffer (so one
        # all-reduce and one dense step cover them), same seeds => identical copies.
        self.rep_idx = list(plan.replicated)
        self.rep_tables, self._rep_grads = [], []
        if self.rep_idx:
            total = sum(plan.rows[t] for t in self.rep_idx) * D
            self._rep_W = torch.empty(total, device="cuda")
            self._rep_G = torch.zeros(total, device="cuda")
            pos = 0
            for t in self.rep_idx:
                n = plan.rows[t] * D
                w = self._rep_W[pos:pos + n].view(plan.rows[t], D)
                r = (1.0 / plan.rows[t]) ** 0.5
                kernels.uniform_init(w, init_seed + t, -r, r)
                self.rep_tables.append(w)
                self._rep_grads.append(self._rep_G[pos:pos + n].view(plan.rows[t], D))
                pos += n
            self._rep_cols = [plan.global_col(t) for t in self.rep_idx]
            self._rep_ws = kernels.alloc_workspace(len(self.rep_idx) * Bs * bag,
                                                   sum(plan.rows[t] for t in self.rep_idx) + 1, D)
        self._presort = presort and N == 1
        self._sorted_for = None
        self._side = torch.cuda.Stream(priority=-1) if self._presort else None
        w_me = plan.width(rank)
        self.out = torch.zeros((Bs, T * D), device="cuda")
        self.grad = torch.zeros((Bs, T * D), device="cuda")
        if N == 1:
            self._out_mat = kernels.make_matrix(self.out)
            self._grad_mat = kernels.make_matrix(self.grad)
            self._out_cols = self._grad_cols = [plan.global_col(t) for t in self.tables_idx]
        else:
            w = w_me
            self._fwd_send = torch.empty(N * Bs * max(w, 1), device="cuda")
            self._fwd_recv = torch.empty(Bs * sum(plan.width(r) for r in range(N)), device="cuda")
            self._bwd_send = torch.empty(Bs * sum(plan.width(r) for r in range(N)), device="cuda")
            self._bwd_recv = torch.empty(N * Bs * max(w, 1), device="cuda")
            ptr = self._fwd_send.data_ptr()
            self._out_mat = kernels.make_matrix(([ptr + q * Bs * w * 4 for q in range(N)], max(w, 4)),
                                                rows_per_part=Bs)
            ptr = self._bwd_recv.data_ptr()
            self._grad_mat = kernels.make_matrix(([ptr + q * Bs * w * 4 for q in range(N)], max(w, 4)),
                                                 rows_per_part=Bs)
            self._out_cols = self._grad_cols = [plan.local_col(t) for t in self.tables_idx]

    def _lookups(self, indices, cols):
        if len(indices) != len(self.tables):
            raise ValueError("expected indices for %d local tables" % len(self.tables))
        return self.K.make_lookups(self.tables, indices, cols=cols)

    def _rep_lookups(self, weights, rep_indices):
        if rep_indices is None or len(rep_indices) != len(self.rep_idx):
            raise ValueError("expected [B/N, bag] indices of this rank's own samples for the %d replicated "
                             "tables %s" % (len(self.rep_idx), self.rep_idx))
        return self.K.make_lookups(weights, rep_indices, cols=self._rep_cols)

    def forward(self, indices, rep_indices=None):
        plan, K, torch = self.plan, self.K, self.torch
        self._sorted_for = None
        if self.rep_idx:    # replicated tables: this rank's own samples, straight into its own slice
            K.embed_forward(self._rep_lookups(self.rep_tables, rep_indices), plan.Bs, plan.D, self.aggr,
                            K.make_matrix(self.out), self.err)
        if self.tables:
            lk = self._lookups(indices, self._out_cols)
            done = None
            if self._presort:       # one GPU: the backward's key sort runs under the gather
                main = torch.cuda.current_stream()
                ready = torch.cuda.Event()
                ready.record(main)
                self._side.wait_event(ready)
                K.embed_backward_sort(lk, plan.B, plan.D, self.workspace, self.err, stream=self._side)
                done = torch.cuda.Event()
                done.record(self._side)
            K.embed_forward(lk, plan.B, plan.D, self.aggr, self._out_mat, self.err)
            if done is not None:
                torch.cuda.current_stream().wait_event(done)
                self._sorted_for = [i.data_ptr() for i in indices]
        if plan.N > 1:
            self.dist.all_to_all_single(self._fwd_recv, self._fwd_send[:plan.N * plan.Bs * plan.width(self.rank)],
                                        plan.forward_recv_splits(self.rank),
                                        plan.forward_send_splits(self.rank), group=self.group)
            plan.unpack_forward(self._fwd_recv, self.out)
        return self.out

    def backward_update(self, indices, rep_indices=None):
        plan, K = self.plan, self.K
        if self.rep_idx:
            # Replicas (the reference's data-parallel embedding): dense gradient of this rank's own
            # samples (the sort/segment-reduce kernels with lr = -1 into a zeroed G), replicas summed
            # by an all-reduce, the same dense SGD step on every rank.
            rlk = self._rep_lookups(self._rep_grads, rep_indices)
            K.embed_backward_sgd(rlk, plan.Bs, plan.D, self.aggr, K.make_matrix(self.grad), -1.0,
                                 self._rep_ws, self.err)
            if plan.N > 1:
                self.dist.all_reduce(self._rep_G, group=self.group)
            check(lib().ffemb_dense_sgd(C.c_void_p(self._rep_W.data_ptr()), C.c_void_p(self._rep_G.data_ptr()),
                                        self._rep_W.numel(), self.lr,
                                        C.c_void_p(self.torch.cuda.current_stream().cuda_stream)))
        if plan.N > 1:
            plan.pack_backward(self.grad, self._bwd_send)
            self.dist.all_to_all_single(self._bwd_recv[:plan.N * plan.Bs * plan.width(self.rank)], self._bwd_send,
                                        plan.backward_recv_splits(self.rank),
                                        plan.backward_send_splits(self.rank), group=self.group)
        if self.tables:
            lk = self._lookups(indices, self._grad_cols)
            if self._sorted_for == [i.data_ptr() for i in indices]:
                K.embed_sgd_update(lk, plan.B, plan.D, self.aggr, self._grad_mat, self.lr, self.workspace)
            else:
                K.embed_backward_sgd(lk, plan.B, plan.D, self.aggr, self._grad_mat, self.lr, self.workspace,
                                     self.err)
            self._sorted_for = None

    def check_errors(self):
        self.torch.cuda.synchronize()
        e = int(self.err.item())
        if e:
            self.err.zero_()
            raise _lib.FFEmbError(e, "an embedding index was outside [0, num_entries)")

    def close(self):
        self.torch.cuda.synchronize()
