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

  exchange="p2p"   over peer memory (CUDA IPC mappings), both directions as
                   NVLink STORES (fire-and-forget; remote loads cost the SM a
                   round trip each and measured about half the bandwidth):
                   forward -- the gather kernel stores each pooled row straight
                   into the consuming rank's [B/N][T*D] matrix (fused, concat
                   layout for free); backward -- one push kernel copies every
                   table's gradient block into its owner's [B][T_r*D] staging
                   matrix, under which the key sort runs on a side stream; the
                   update then reads local memory only.  Ordering between
                   ranks: two ffemb_peer_barrier kernels per step.
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

    Hazards between steps are covered by the two barriers of a step: peers
    overwrite this rank's `out` only in their next forward, which they reach
    after the backward barrier that this rank passes only after it has used
    `out`; peers overwrite the gradient staging only in their next push, after
    the forward barrier that this rank passes only after its update.
    """

    def __init__(self, plan, rank, bag, aggr, lr, exchange="p2p", group=None, init_seed=1234, presort=True,
                 barrier="peer", backward_exchange="auto"):
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
        self._flag = torch.zeros(1, dtype=torch.int32, device="cuda")
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
            self._rep_stream = torch.cuda.Stream()
        self._peers = []
        self._peer_flags = None
        self._push = None
        self.barrier_impl = barrier      # "peer": ffemb_peer_barrier kernel; "nccl": 4-byte all-reduce
        # p2p backward exchange.  "push": gradient blocks are stored into the owners' staging, the
        # update reads local memory.  "pull": the update kernel loads gradient rows from the peers
        # while it works (no staging pass, but remote loads reach only about half the NVLink rate).
        # Measured on C3: pull wins at N = 2 (0.80 vs 0.88 ms/step), push from N = 4 up.
        if backward_exchange not in ("auto", "push", "pull"):
            raise ValueError("backward_exchange must be 'auto', 'push' or 'pull'")
        self.bwd_mode = backward_exchange if backward_exchange != "auto" else ("push" if plan.N >= 4 else "pull")
        self.trace = None                # set to [] to collect per-call CUDA events (development aid)
        # The backward's key sort needs only the indices.  It runs on a
        # high-priority side stream (its small grids then take SM slots as they
        # free up): under the gather when there is one GPU, under the
        # NVLink-bound gradient push when there are several.
        self._presort = presort
        self._sorted_for = None
        self._rep_sorted_for = None
        self._side = torch.cuda.Stream(priority=-1) if presort else None
        self.grad = None
        w_me = plan.width(rank)
        if exchange == "p2p" and N > 1:
            # one IPC-exportable allocation:
            # [ out Bs x T*D | 64 barrier words | grad Bs x T*D | staging B x T_r*D (push mode) ]
            off_flags = Bs * T * D
            off_grad = off_flags + 64
            off_stage = off_grad + Bs * T * D
            self._buf = _DeviceBuffer(off_stage + (B * max(w_me, 4) if self.bwd_mode == "push" else 0))
            self._buf.tensor[off_flags:off_stage].zero_()
            self.grad = self._buf.tensor[off_grad:off_stage].view(Bs, T * D)
            torch.cuda.synchronize()
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
            self._peer_flags = (C.c_void_p * N)(*[b + off_flags * 4 for b in base])
            self._out_mat = kernels.make_matrix(([b for b in base], T * D), rows_per_part=Bs)
            self._out_cols = [plan.global_col(t) for t in self.tables_idx]
            # backward: table t's block of my gradient goes to rows [rank*Bs, ..) of owner(t)'s staging
            sharded = [t for t in range(T) if plan.owner[t] >= 0]
            items = (_lib.Push * max(len(sharded), 1))()
            for i, t in enumerate(sharded):
                q = plan.owner[t]
                wq = plan.width(q)
                items[i].dst = base[q] + (off_stage + rank * Bs * wq + plan.local_col(t)) * 4
                items[i].dst_ld = wq
                items[i].src_col = plan.global_col(t)
            self._n_push = len(sharded)
            if self.bwd_mode == "push":
                self._push = items
                self._grad_mat = kernels.make_matrix(([self._buf.ptr + off_stage * 4], max(w_me, 4)),
                                                     rows_per_part=B)
                self._grad_cols = [plan.local_col(t) for t in self.tables_idx]
            else:
                self._grad_mat = kernels.make_matrix(([b + off_grad * 4 for b in base], T * D), rows_per_part=Bs)
                self._grad_cols = self._out_cols
        else:
            self._buf = None
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
        if N > 1:   # every rank's buffers and zeroed flags exist before any peer touches them
            torch.cuda.synchronize()
            dist.all_reduce(self._flag, group=group)
            torch.cuda.synchronize()

    # -- stream-ordered cross-rank barrier (no host sync) ---------------------------
    def barrier(self):
        if self.plan.N <= 1:
            return
        if self._peer_flags is not None and self.barrier_impl == "peer":   # own NVLink barrier kernel
            check(lib().ffemb_peer_barrier(self._peer_flags, self.rank, self.plan.N,
                                           C.c_void_p(self.err.data_ptr()),
                                           C.c_void_p(self.torch.cuda.current_stream().cuda_stream)))
        else:
            self.dist.all_reduce(self._flag, group=self.group)

    def _mark(self, name):
        if self.trace is not None:
            ev = self.torch.cuda.Event(enable_timing=True)
            ev.record()
            self.trace.append((name, ev))

    def trace_summary(self):
        """Mean device time between consecutive marks, by (from, to) pair, in us."""
        self.torch.cuda.synchronize()
        acc = {}
        for (a, ea), (b, eb) in zip(self.trace[:-1], self.trace[1:]):
            k = a + " -> " + b
            tot, n = acc.get(k, (0.0, 0))
            acc[k] = (tot + ea.elapsed_time(eb) * 1e3, n + 1)
        return {k: round(t / n, 1) for k, (t, n) in acc.items()}

    def _lookups(self, indices, cols):
        if len(indices) != len(self.tables):
            raise ValueError("expected indices for %d local tables" % len(self.tables))
        return self.K.make_lookups(self.tables, indices, cols=cols)

    def _fork_sort(self, lk):
        """Start the key sort on the side stream; returns the event to join on."""
        torch, plan = self.torch, self.plan
        main = torch.cuda.current_stream()
        ready = torch.cuda.Event()
        ready.record(main)
        self._side.wait_event(ready)
        self.K.embed_backward_sort(lk, plan.B, plan.D, self.workspace, self.err, stream=self._side)
        done = torch.cuda.Event()
        done.record(self._side)
        return done

    def _rep_lookups(self, weights, rep_indices):
        if rep_indices is None or len(rep_indices) != len(self.rep_idx):
            raise ValueError("expected [B/N, bag] indices of this rank's own samples for the %d replicated "
                             "tables %s" % (len(self.rep_idx), self.rep_idx))
        return self.K.make_lookups(weights, rep_indices, cols=self._rep_cols)

    def forward(self, indices, rep_indices=None):
        plan, K, torch = self.plan, self.K, self.torch
        self._sorted_for = None
        self._mark("fwd_begin")
        self._rep_sorted_for = None
        if self.rep_idx:    # replicated tables: this rank's own samples, straight into its own slice
            if self._presort:
                # their backward is a chain of small latency-bound kernels that would otherwise end up
                # on the critical path of backward_update: sort the keys now, under the gather
                main = torch.cuda.current_stream()
                ready = torch.cuda.Event()
                ready.record(main)
                self._rep_stream.wait_event(ready)
                K.embed_backward_sort(self._rep_lookups(self._rep_grads, rep_indices), plan.Bs, plan.D,
                                      self._rep_ws, self.err, stream=self._rep_stream)
                rep_sorted = torch.cuda.Event()
                rep_sorted.record(self._rep_stream)
            K.embed_forward(self._rep_lookups(self.rep_tables, rep_indices), plan.Bs, plan.D, self.aggr,
                            K.make_matrix(self.out), self.err)
            if self._presort:
                main.wait_event(rep_sorted)      # join: one timeline for callers and graph capture
                self._rep_sorted_for = [i.data_ptr() for i in rep_indices]
        if self.tables:
            lk = self._lookups(indices, self._out_cols)
            done = None
            if self._presort and plan.N == 1:
                done = self._fork_sort(lk)
            K.embed_forward(lk, plan.B, plan.D, self.aggr, self._out_mat, self.err)
            self._mark("fwd_kernel_done")
            if done is not None:
                torch.cuda.current_stream().wait_event(done)
                self._sorted_for = [i.data_ptr() for i in indices]
        if plan.N > 1 and self.exchange == "nccl":
            self.dist.all_to_all_single(self._fwd_recv, self._fwd_send[:plan.N * plan.Bs * plan.width(self.rank)],
                                        plan.forward_recv_splits(self.rank),
                                        plan.forward_send_splits(self.rank), group=self.group)
            plan.unpack_forward(self._fwd_recv, self.out)
        else:
            self.barrier()      # every rank's rows have landed in self.out
        self._mark("fwd_end")
        return self.out

    def backward_update(self, indices, rep_indices=None):
        plan, K, torch = self.plan, self.K, self.torch
        rep_done = None
        if self.rep_idx:
            # Replicas (the reference's data-parallel embedding): dense gradient of this rank's own
            # samples (the sort/segment-reduce kernels with lr = -1 into a zeroed G), replicas summed
            # by an all-reduce, the same dense SGD step on every rank.  On its own stream, under the
            # exchange of the sharded tables.
            main = torch.cuda.current_stream()
            ready = torch.cuda.Event()
            ready.record(main)
            with torch.cuda.stream(self._rep_stream):
                self._rep_stream.wait_event(ready)
                rlk = self._rep_lookups(self._rep_grads, rep_indices)
                if self._rep_sorted_for == [i.data_ptr() for i in rep_indices]:     # forward() sorted these
                    K.embed_sgd_update(rlk, plan.Bs, plan.D, self.aggr, K.make_matrix(self.grad), -1.0,
                                       self._rep_ws, stream=self._rep_stream)
                else:
                    K.embed_backward_sgd(rlk, plan.Bs, plan.D, self.aggr, K.make_matrix(self.grad), -1.0,
                                         self._rep_ws, self.err, stream=self._rep_stream)
                self._rep_sorted_for = None
                self.dist.all_reduce(self._rep_G, group=self.group)
                check(lib().ffemb_dense_sgd(C.c_void_p(self._rep_W.data_ptr()), C.c_void_p(self._rep_G.data_ptr()),
                                            self._rep_W.numel(), self.lr,
                                            C.c_void_p(self._rep_stream.cuda_stream)))
                rep_done = torch.cuda.Event()
                rep_done.record(self._rep_stream)
        lk = self._lookups(indices, self._grad_cols) if self.tables else None
        done = None
        presorted = lk is not None and self._sorted_for == [i.data_ptr() for i in indices]
        if lk is not None and self._presort and not presorted and plan.N > 1:
            done = self._fork_sort(lk)              # runs under the exchange below
        if plan.N > 1 and self.exchange == "nccl":
            plan.pack_backward(self.grad, self._bwd_send)
            self.dist.all_to_all_single(self._bwd_recv[:plan.N * plan.Bs * plan.width(self.rank)], self._bwd_send,
                                        plan.backward_recv_splits(self.rank),
                                        plan.backward_send_splits(self.rank), group=self.group)
        elif plan.N > 1:
            if self._push is not None:
                check(lib().ffemb_push_columns(self._push, self._n_push, plan.Bs, plan.D,
                                               C.c_void_p(self.grad.data_ptr()), self.grad.stride(0),
                                               C.c_void_p(torch.cuda.current_stream().cuda_stream)))
                self._mark("push_done")
            self.barrier()      # push: blocks have landed in the owners' staging; pull: every grad is complete
        self._mark("bwd_exchange_done")
        if lk is not None:
            if done is not None:
                torch.cuda.current_stream().wait_event(done)
                presorted = True
            if presorted:
                K.embed_sgd_update(lk, plan.B, plan.D, self.aggr, self._grad_mat, self.lr, self.workspace)
            else:
                K.embed_backward_sgd(lk, plan.B, plan.D, self.aggr, self._grad_mat, self.lr, self.workspace,
                                     self.err)
            self._sorted_for = None
        if rep_done is not None:
            torch.cuda.current_stream().wait_event(rep_done)
        self._mark("bwd_end")

    def check_errors(self):
        self.torch.cuda.synchronize()
        e = int(self.err.item())
        if e:
            self.err.zero_()
            raise _lib.FFEmbError(e, "a peer rank did not reach a barrier within 5 s" if e == _lib.FFEMB_ERR_COMM
                                  else "an embedding index was outside [0, num_entries)")

    def close(self):
        self.torch.cuda.synchronize()
        if self.plan.N > 1:
            self.dist.barrier(group=self.group)      # nobody unmaps memory a peer may still write
        for p in self._peers:
            lib().ffemb_ipc_close_handle(p)
        self._peers = []
        if self._buf is not None:
            self.out = self.grad = None
            self._buf.free()
            self._buf = None
