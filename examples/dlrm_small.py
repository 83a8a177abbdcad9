#!/usr/bin/env python
"""DLRM-small end to end: the reference's DLRM example with its embedding
operator replaced by this library.

Follows examples/cpp/DLRM/dlrm.cc of the reference: sparse inputs
[batch, bag] int64 per table (:100-108), bottom MLP on the dense input (:117),
`ff.embedding(input, num_entries, out_dim, AGGR_MODE_SUM, NULL,
UniformInitializer(seed, -sqrt(1/n), sqrt(1/n)))` per table (:41-47,120-128),
interaction "cat" (:49-65), top MLP (:130), MSE-avg loss, SGD lr 0.01 (:134),
then the loop forward / zero_gradients / backward / update (:166-187) and the
THROUGHPUT line (:195-198).  Defaults are run_random.sh's shape scaled to
BASELINE config C1 (8 tables x 100k x 64, batch 256, MLP 64-512-512-64).

The dense layers are plain torch modules -- they are outside the path this
repository replaces; the embedding tables, their forward, backward and SGD
update are libffemb.so, driven through the reference's own call sequence.
The embedding outputs and their gradients live in one [batch, T*D] torch
tensor that the library reads and writes in place (the Concat layout), so the
"cat" interaction costs no copy.

Under `torchrun --nproc-per-node N examples/dlrm_small.py -ll:gpu N [--import-strategy
<file.pb>]` the same program runs sharded, as the reference's does under
Legion with a strategy file (src/mapper/mapper.cc:282-286): the tables are
placed table-wise by the library, every rank keeps batch / N samples of the
dense part (data parallel, gradients averaged with an all-reduce) and reads
its slice of all tables' outputs from the library's exchange matrix.
"""
import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dlrm_flexflow_b200 as ff  # noqa: E402


def mlp(sizes, sigmoid_layer=-1):
    """create_mlp, dlrm.cc:26-39: dense + relu, sigmoid on `sigmoid_layer`."""
    layers = []
    for i in range(len(sizes) - 1):
        lin = torch.nn.Linear(sizes[i], sizes[i + 1])
        std = (2.0 / (sizes[i] + sizes[i + 1])) ** 0.5
        torch.nn.init.normal_(lin.weight, 0.0, std)
        torch.nn.init.normal_(lin.bias, 0.0, (2.0 / sizes[i + 1]) ** 0.5)
        layers += [lin, torch.nn.Sigmoid() if i == sigmoid_layer else torch.nn.ReLU()]
    return torch.nn.Sequential(*layers)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("-b", "--batch-size", type=int, default=256)
    ap.add_argument("-e", "--epochs", type=int, default=2)
    ap.add_argument("--iterations", type=int, default=20, help="iterations per epoch (synthetic data)")
    ap.add_argument("--arch-sparse-feature-size", type=int, default=64)
    ap.add_argument("--arch-embedding-size", default="-".join(["100000"] * 8))
    ap.add_argument("--embedding-bag-size", type=int, default=1)
    ap.add_argument("--arch-mlp-bot", default="64-512-512-64")
    ap.add_argument("--arch-mlp-top", default="576-1024-1024-1024-1")
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("-ll:gpu", dest="gpus", type=int, default=1)
    ap.add_argument("--import-strategy", "--import", dest="strategy", default=None)
    args = ap.parse_args(argv)
    world, rank = args.gpus, int(os.environ.get("RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))

    assert torch.cuda.is_available(), "needs a CUDA device: the embedding path has no CPU fallback"
    dev = torch.device("cuda")
    B, D, L = args.batch_size, args.arch_sparse_feature_size, args.embedding_bag_size
    rows = [int(x) for x in args.arch_embedding_size.split("-")]
    bot = [int(x) for x in args.arch_mlp_bot.split("-")]
    top = [int(x) for x in args.arch_mlp_top.split("-")]
    T = len(rows)
    assert bot[-1] == D, "bottom MLP must end at the sparse feature size"
    assert top[0] == (T + 1) * D, "top MLP input must be (tables + 1) * sparse feature size for 'cat'"

    # ---- the model, built as the reference example builds it -------------------
    ffconfig = ff.FFConfig()
    ffconfig.set_batch_size(B)
    ffconfig.set_workers_per_node(world)
    ffconfig.set_rank(rank, torch.cuda.current_device())
    if args.strategy:
        ffconfig.import_strategy_file(args.strategy)
    ffmodel = ff.FFModel(ffconfig)
    Bs = B // world                                        # samples of the dense part on this rank
    sparse_inputs = [ffmodel.create_tensor([B, L], ff.DataType.DT_INT64) for _ in range(T)]
    ly = []
    for i in range(T):
        r = (1.0 / rows[i]) ** 0.5
        ly.append(ffmodel.embedding(sparse_inputs[i], rows[i], D, ff.AggrMode.AGGR_MODE_SUM, None,
                                    ff.UniformInitializer(1000 + i, -r, r)))
    ffmodel.compile(optimizer=ff.SGDOptimizer(ffmodel, args.lr))
    ffmodel.init_layers()
    ffmodel.set_stream(torch.cuda.current_stream().cuda_stream)

    torch.manual_seed(0)
    bot_mlp, top_mlp = mlp(bot).to(dev), mlp(top, sigmoid_layer=len(top) - 2).to(dev)
    dense_opt = torch.optim.SGD(list(bot_mlp.parameters()) + list(top_mlp.parameters()), lr=args.lr)

    if world == 1:
        # z = cat(x, ly...): one [B, (T+1)*D] buffer; the library writes the T embedding blocks in place
        z = torch.zeros((B, (T + 1) * D), device=dev)
        gz = torch.zeros_like(z)
        for i in range(T):
            ly[i].attach_torch(z[:, (i + 1) * D:(i + 2) * D])
            ly[i].attach_torch(gz[:, (i + 1) * D:(i + 2) * D], grad=True)
        emb_out, emb_grad = z[:, D:], gz[:, D:]
    else:
        # sharded: this rank's [B/N, T*D] slices of all tables' outputs / gradients (peer-mapped, library-owned)
        emb_out, emb_grad = ffmodel.exchange_matrix(False), ffmodel.exchange_matrix(True)

    gen = torch.Generator(device="cuda")
    gen.manual_seed(1)                                     # the same data on every rank; each takes its samples
    dense_in = torch.rand((B, bot[0]), device=dev, generator=gen)[rank * Bs:(rank + 1) * Bs]
    label = torch.rand((B, 1), device=dev, generator=gen)[rank * Bs:(rank + 1) * Bs]
    idx = [torch.randint(0, rows[i], (B, L), device=dev, generator=gen) for i in range(T)]   # rand() % N
    for i in range(T):
        sparse_inputs[i].attach_torch(idx[i])              # a rank reads only what its tables need
    dense_params = list(bot_mlp.parameters()) + list(top_mlp.parameters())

    def iteration():
        ffmodel.forward()                                  # embeddings -> this rank's samples of every table
        x = bot_mlp(dense_in)
        emb = emb_out.detach().requires_grad_(True)        # the T embedding blocks, already side by side
        p = top_mlp(torch.cat([x, emb], dim=1))            # interact_features "cat" (dlrm.cc:49-65)
        loss = torch.nn.functional.mse_loss(p, label) / world   # LOSS_MEAN_SQUARED_ERROR_AVG_REDUCE over the batch
        dense_opt.zero_grad(set_to_none=True)
        ffmodel.zero_gradients()
        loss.backward()
        emb_grad.copy_(emb.grad)                           # d loss / d embedding outputs
        ffmodel.backward()
        if world > 1:                                      # data-parallel dense part: sum of the ranks' gradients
            for q in dense_params:
                dist.all_reduce(q.grad)
        dense_opt.step()
        ffmodel.update()
        total = loss.detach().clone()
        if world > 1:
            dist.all_reduce(total)
        return total

    loss0 = float(iteration())                           # warm-up iteration (dlrm.cc:143-152)
    ffmodel.sync()
    torch.cuda.synchronize()
    ts = time.perf_counter()
    loss = None
    for _ in range(args.epochs):
        for _ in range(args.iterations):
            loss = iteration()
    ffmodel.sync()
    torch.cuda.synchronize()
    run_time = time.perf_counter() - ts
    n = B * args.iterations * args.epochs
    if rank == 0:
        print("loss: first %.6f last %.6f" % (loss0, float(loss)))
        print("ELAPSED TIME = %.4fs, THROUGHPUT = %.2f samples/s" % (run_time, n / run_time))
    if world > 1:
        ffmodel.close()
    return loss0, float(loss)


if __name__ == "__main__":
    main()
