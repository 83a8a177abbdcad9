#!/usr/bin/env python
"""Development probe: pinned-host <-> device copy rates on this box, one and both
directions at once, with the process bound to each NUMA node in turn."""
import os
import subprocess
import sys
import torch


def rates(tag):
    n = 512 << 20
    h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_in.fill_(1); h_out.fill_(2)
    d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(h2d, d2h, it=6):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(it):
            if h2d:
                with torch.cuda.stream(s1):
                    d_in.copy_(h_in, non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    h_out.copy_(d_out, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s1)
        torch.cuda.current_stream().wait_stream(s2)
        b.record()
        torch.cuda.synchronize()
        return n * it / (a.elapsed_time(b) * 1e-3) / 1e9
    run(True, True, 2)
    print("%-28s H2D %.1f GB/s   D2H %.1f GB/s   both: %.1f GB/s each way" % (tag, run(True, False), run(False, True), run(True, True)))


if __name__ == "__main__":
    if len(sys.argv) > 1:
        cpus = [int(c) for c in sys.argv[1].split(",")]
        os.sched_setaffinity(0, cpus)
        rates("cpus %s..%s" % (cpus[0], cpus[-1]))
    else:
        print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout[:1500])
        print(subprocess.run("lscpu | grep -i -E 'numa|^cpu\\(s\\)|model name'", shell=True, capture_output=True, text=True).stdout)
        print("affinity now:", len(os.sched_getaffinity(0)), "cpus", sorted(os.sched_getaffinity(0))[:4], "...")
        rates("default")
        nodes = {}
        for d in sorted(os.listdir("/sys/devices/system/node")):
            if d.startswith("node"):
                lst = open("/sys/devices/system/node/%s/cpulist" % d).read().strip()
                nodes[d] = lst
        print(nodes)
        for d, lst in nodes.items():
            cpus = []
            for part in lst.split(","):
                if "-" in part:
                    a, b = part.split("-"); cpus += list(range(int(a), int(b) + 1))
                elif part:
                    cpus.append(int(part))
            cpus = [c for c in cpus if c in os.sched_getaffinity(0)]
            if cpus:
                subprocess.run([sys.executable, __file__, ",".join(map(str, cpus))])
