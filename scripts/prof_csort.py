#!/usr/bin/env python
"""Development aid: per-phase clock counts of cluster_sort_kernel (one CTA).
Build the profiling variant first:
  FFEMB_VARIANT=_prof FFEMB_NVCC_EXTRA="-DFFEMB_CS_PROFILE=<block index>" python -m dlrm_flexflow_b200.build
and run with FFEMB_VARIANT=_prof on a GPU box."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlrm_flexflow_b200 import _lib, kernels as K, workloads  # noqa: E402

NAMES = ["start+csr", "load", "count", "warp prefix", "cluster.sync 1", "bases", "reorder", "copy out",
         "cluster.sync 2"]


def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 else "C3"
    gen = torch.Generator(device="cuda")
    gen.manual_seed(0)
    w = workloads.WORKLOADS[cfg]
    rows, D, B, L = w["rows"], 4, w["batch"], w["bag"]
    Ws = [torch.zeros(n, D, device="cuda") for n in rows]
    idxs = [workloads.gen_indices(torch, n, B, L, "uniform", 0.0, gen) for n in rows]
    ws = K.alloc_workspace(len(rows) * B * L, sum(rows), D)
    lk = K.make_lookups(Ws, idxs)
    K.set_sort_impl(2)
    out = (C.c_longlong * 16)()
    fn = _lib.lib().ffemb_debug_cs_prof
    K.embed_backward_sort(lk, B, D, ws)
    fn(out, 1)
    iters = 20
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        K.embed_backward_sort(lk, B, D, ws)
    e.record()
    fn(out, 0)
    print("%s: sort phase %.1f us per call" % (cfg, s.elapsed_time(e) / iters * 1e3))
    tot = sum(out[:9]) / iters
    for i, nm in enumerate(NAMES):
        print("  %-16s %9.0f clk  %5.1f %%" % (nm, out[i] / iters, 100.0 * out[i] / iters / max(tot, 1)))
    print("  total            %9.0f clk = %.1f us at 1.965 GHz" % (tot, tot / 1965.0))


if __name__ == "__main__":
    main()
