"""Table-wise sharding of the embedding tables over the GPUs of one box.

One process per GPU (torch.distributed supplies rendezvous and NCCL).  Table t
lives wholly on rank owner(t) -- the reference's DLRM strategy (embedding<i>
-> device i % ngpu, dims [1,1]: src/runtime/dlrm_strategy.cc:252-256) -- and
that rank processes the WHOLE batch for its tables, while the dense part of
the model is sample-parallel: rank q consumes rows [q*B/N, (q+1)*B/N) of every
table's output, side by side as [B/N][T*D] (what the reference's Concat builds,
src/ops/concat.cu:159-215), and produces the gradient in the same layout.

Two exchanges per step (the reference does them implicitly by Legion DMA
through zero-copy host memory, src/ops/embedding.cu:280-284,343-347):

  exchange="p2p"   fused: the forward kernel stores each pooled row straight
                   into the consuming rank's [B/N][T*D] matrix over NVLink and
                   the update kernel loads each gradient row straight from the
                   producing rank's (peer mappings by CUDA IPC).  No staging
                   buffer, no separate exchange pass, concat layout for free.
  exchange="nccl"  baseline: kernels write/read local staging buffers and
                   NCCL all-to-all (all_to_all_single with uneven splits)
                   moves them; a strided copy builds / splits the concat
                   layout.

`ShardPlan` is the pure host logic (placement, column layout, split sizes,
pack/unpack of the NCCL staging layout); it runs on CPU and is covered by the
world_size-2 gloo tests.
"""
import ctypes as C

from . import _lib
from ._lib import check, lib


class ShardPlan(object):
    def __init__(self, rows, dim, global_batch, world, owners=None, strategy=None):
        if global_batch % world:
            raise ValueError("global batch %d is not divisible by world size %d" % (global_batch, world))
        self.rows = list(rows)
        self.T, self.D, self.B, self.N = len(rows), dim, global_batch, world
        self.Bs = global_batch // world
        if owners is None:
            handle = strategy.handle if strategy is not None else _lib.Handle(None)
            owners = [lib().ffemb_strategy_device_for_table(handle, t, None, world) for t in range(self.T)]
        if len(owners) != self.T or any(not 0 <= o < world for o in owners):
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
        """Best possible scaling efficiency from the integer table split alone."""
        most = max(len(l) for l in self.local)
        return self.T / float(self.N * most) if most else 1.0

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


class _DeviceBuffer(object):
    """cudaMalloc'ed (IPC-exportable) float buffer with a torch view."""

    def __init__(self, numel):
        import torch
        ptr = C.c_void_p()
        check(lib().ffemb_device_malloc(C.byref(ptr), numel * 4))
        self.ptr, self.numel = ptr.value, numel
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": "<f4", "data": (self.ptr, False),
                                         "version": 2, "strides": None}
        self.tensor = torch.as_tensor(self, device="cuda")

    def handle(self):
        h = (C.c_uint8 * 64)()
        check(lib().ffemb_ipc_get_handle(self.ptr, h))
        return bytes(h)

    def free(self):
        if self.ptr:
            self.tensor = None
            lib().ffemb_device_free(self.ptr)
            self.ptr = None


class ShardedEmbeddingBags(object):
    """The embedding tables of one DLRM, sharded table-wise; this rank's part.

    forward(indices)         indices[i]: [B, bag] int64 cuda for this rank's
                             i-th table, WHOLE batch.  Returns this rank's
                             [B/N][T*D] slice of all tables' pooled outputs.
    backward_update(indices) consumes self.grad ([B/N][T*D], filled by the
                             caller) on every rank and applies the sparse SGD
                             update to this rank's tables.
    """

    def __init__(self, plan, rank, bag, aggr, lr, exchange="p2p", group=None, init_seed=1234, presort=True):
        import torch
        import torch.distributed as dist
        from . import kernels
        if not torch.cuda.is_available():
            raise RuntimeError("ShardedEmbeddingBags needs a CUDA device: there is no CPU fallback")
        if exchange not in ("p2p", "nccl"):
            raise ValueError("exchange must be 'p2p' or 'nccl'")
        if plan.N > _lib.FFEMB_MAX_PARTS:
            raise ValueError("at most %d ranks" % _lib.FFEMB_MAX_PARTS)
        self.plan, self.rank, self.bag, self.aggr, self.lr = plan, rank, bag, aggr, lr
        self.exchange, self.group, self.K, self.torch, self.dist = exchange, group, kernels, torch, dist
        self.tables_idx = plan.tables_of(rank)
        D, Bs, T, N = plan.D, plan.Bs, plan.T, plan.N
        self.tables = []
        for t in self.tables_idx:
            w = torch.empty((plan.rows[t], D), device="cuda")
            r = (1.0 / plan.rows[t]) ** 0.5
            kernels.uniform_init(w, init_seed + t, -r, r)
            self.tables.append(w)
        n_lookups = len(self.tables_idx) * plan.B * bag
        self.workspace = kernels.alloc_workspace(max(n_lookups, 1), sum(plan.rows[t] for t in self.tables_idx) + 1, D)
        self.err = torch.zeros(1, dtype=torch.int32, device="cuda")
        self._flag = torch.zeros(1, dtype=torch.int32, device="cuda")
        self._peers = []
        # the backward's key sort needs only the indices: forward() runs it on a
        # side stream underneath the gather (and its exchange)
        self._presort = presort
        self._sorted_for = None
        # high priority: the sort's small grids take SM slots as they free up
        # instead of queueing behind the gather's thousands of blocks
        self._side = torch.cuda.Stream(priority=-1) if presort else None
        if exchange == "p2p" and N > 1:
            self._buf = _DeviceBuffer(2 * Bs * T * D)
            handles = [None] * N
            dist.all_gather_object(handles, self._buf.handle(), group=group)
            base = []
            for q in range(N):
                if q == rank:
                    base.append(self._buf.ptr)
                else:
                    p = C.c_void_p()
                    hb = (C.c_uint8 * 64).from_buffer_copy(handles[q])
                    check(lib().ffemb_ipc_open_handle(hb, C.byref(p)))
                    self._peers.append(p.value)
                    base.append(p.value)
            self.out = self._buf.tensor[:Bs * T * D].view(Bs, T * D)
            self.grad = self._buf.tensor[Bs * T * D:].view(Bs, T * D)
            self._out_mat = kernels.make_matrix(([b for b in base], T * D), rows_per_part=Bs)
            self._grad_mat = kernels.make_matrix(([b + Bs * T * D * 4 for b in base], T * D), rows_per_part=Bs)
            self._cols = [plan.global_col(t) for t in self.tables_idx]
        else:
            self._buf = None
            self.out = torch.zeros((Bs, T * D), device="cuda")
            self.grad = torch.zeros((Bs, T * D), device="cuda")
            if N == 1:
                self._out_mat = kernels.make_matrix(self.out)
                self._grad_mat = kernels.make_matrix(self.grad)
                self._cols = [plan.global_col(t) for t in self.tables_idx]
            else:
                w = plan.width(rank)
                self._fwd_send = torch.empty(N * Bs * max(w, 1), device="cuda")
                self._fwd_recv = torch.empty(Bs * T * D, device="cuda")
                self._bwd_send = torch.empty(Bs * T * D, device="cuda")
                self._bwd_recv = torch.empty(N * Bs * max(w, 1), device="cuda")
                ptr = self._fwd_send.data_ptr()
                self._out_mat = kernels.make_matrix(([ptr + q * Bs * w * 4 for q in range(N)], max(w, 4)),
                                                    rows_per_part=Bs)
                ptr = self._bwd_recv.data_ptr()
                self._grad_mat = kernels.make_matrix(([ptr + q * Bs * w * 4 for q in range(N)], max(w, 4)),
                                                     rows_per_part=Bs)
                self._cols = [plan.local_col(t) for t in self.tables_idx]
        self.barrier()

    # -- stream-ordered cross-rank barrier (no host sync) ---------------------------
    def barrier(self):
        if self.plan.N > 1:
            self.dist.all_reduce(self._flag, group=self.group)

    def _lookups(self, indices):
        if len(indices) != len(self.tables):
            raise ValueError("expected indices for %d local tables" % len(self.tables))
        return self.K.make_lookups(self.tables, indices, cols=self._cols)

    def forward(self, indices):
        plan, K, torch = self.plan, self.K, self.torch
        self._sorted_for = None
        if self.tables:
            lk = self._lookups(indices)
            if self._presort:
                main = torch.cuda.current_stream()
                ready = torch.cuda.Event()
                ready.record(main)
                self._side.wait_event(ready)
                K.embed_backward_sort(lk, plan.B, plan.D, self.workspace, self.err, stream=self._side)
                done = torch.cuda.Event()
                done.record(self._side)
            K.embed_forward(lk, plan.B, plan.D, self.aggr, self._out_mat, self.err)
            if self._presort:
                main.wait_event(done)
                self._sorted_for = [i.data_ptr() for i in indices]
        if plan.N > 1 and self.exchange == "nccl":
            self.dist.all_to_all_single(self._fwd_recv, self._fwd_send[:plan.N * plan.Bs * plan.width(self.rank)],
                                        plan.forward_recv_splits(self.rank),
                                        plan.forward_send_splits(self.rank), group=self.group)
            plan.unpack_forward(self._fwd_recv, self.out)
        else:
            self.barrier()      # every rank's rows have landed in self.out
        return self.out

    def backward_update(self, indices):
        plan, K = self.plan, self.K
        if plan.N > 1 and self.exchange == "nccl":
            plan.pack_backward(self.grad, self._bwd_send)
            self.dist.all_to_all_single(self._bwd_recv[:plan.N * plan.Bs * plan.width(self.rank)], self._bwd_send,
                                        plan.backward_recv_splits(self.rank),
                                        plan.backward_send_splits(self.rank), group=self.group)
        else:
            self.barrier()      # every rank's self.grad is complete before peers read it
        if self.tables:
            lk = self._lookups(indices)
            if self._sorted_for == [i.data_ptr() for i in indices]:     # forward() sorted these already
                K.embed_sgd_update(lk, plan.B, plan.D, self.aggr, self._grad_mat, self.lr, self.workspace)
            else:
                K.embed_backward_sgd(lk, plan.B, plan.D, self.aggr, self._grad_mat, self.lr, self.workspace,
                                     self.err)
            self._sorted_for = None
        if not (plan.N > 1 and self.exchange == "nccl"):
            self.barrier()      # peers are done reading self.grad / may overwrite self.out

    def check_errors(self):
        self.torch.cuda.synchronize()
        e = int(self.err.item())
        if e:
            self.err.zero_()
            raise _lib.FFEmbError(e, "an embedding index was outside [0, num_entries)")

    def close(self):
        self.torch.cuda.synchronize()
        for p in self._peers:
            lib().ffemb_ipc_close_handle(p)
        self._peers = []
        if self._buf is not None:
            self.out = self.grad = None
            self._buf.free()
            self._buf = None
