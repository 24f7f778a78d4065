#!/usr/bin/env python
"""Where does pinned host memory have to live for a GPU's H2D copies to run at full speed?
Prints the NUMA layout the box exposes, then times pinned H2D copies with the allocating process
bound to each NUMA node in turn (first-touch places the ring's pages on that node).
Single process:  python tools/numa_probe.py            (GPU 0 against every node)
All GPUs:        torchrun --nproc-per-node N tools/numa_probe.py --mode local|node0|spread"""
import argparse
import glob
import os
import re
import subprocess
import time

import torch


def cpulist(s):
    out = []
    for part in s.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        out.extend(range(int(a), int(b or a) + 1))
    return out


def nodes():
    res = {}
    for d in sorted(glob.glob("/sys/devices/system/node/node[0-9]*")):
        with open(os.path.join(d, "cpulist")) as f:
            res[int(re.search(r"node(\d+)$", d).group(1))] = cpulist(f.read())
    return res


def gpu_node(idx):
    bdf = torch.cuda.get_device_properties(idx).pci_bus_id if hasattr(torch.cuda.get_device_properties(idx), "pci_bus_id") else None
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(idx)],
                             capture_output=True, text=True, timeout=20).stdout.strip()
        bdf = out.lower()
        if bdf.startswith("00000000:"):
            bdf = bdf[4:]
        with open("/sys/bus/pci/devices/%s/numa_node" % bdf) as f:
            return int(f.read()), bdf
    except Exception as e:                                  # noqa: BLE001
        return -1, "%s (%s)" % (bdf, e)


def h2d_gbs(n_bytes, reps=8):
    h = torch.empty(n_bytes, dtype=torch.uint8).pin_memory()
    h.fill_(1)
    d = torch.empty(n_bytes, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    return n_bytes * reps / (time.perf_counter() - t0) / 1e9


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="sweep", choices=["sweep", "local", "node0", "spread", "none"])
    ap.add_argument("--mb", type=int, default=656)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    lr = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    nd = nodes()
    all_cpus = sorted(os.sched_getaffinity(0))
    if world == 1:
        print("cpus allowed: %d  numa nodes: %s" % (len(all_cpus), {k: "%d cpus" % len(v) for k, v in nd.items()}))
        for cmd in (["nvidia-smi", "topo", "-m"], ["lscpu"]):
            try:
                txt = subprocess.run(cmd, capture_output=True, text=True, timeout=30).stdout
                keep = [l for l in txt.splitlines() if cmd[0] != "lscpu" or re.search(r"NUMA|Model name|Socket|^CPU\(s\)|Thread|Hypervisor", l)]
                print("\n".join(keep))
            except Exception as e:                          # noqa: BLE001
                print(cmd, "failed:", e)
        gn, bdf = gpu_node(lr)
        print("GPU %d pci %s numa_node %d" % (lr, bdf, gn))
        print("unbound: %.1f GB/s" % h2d_gbs(a.mb << 20))
        for k, cpus in nd.items():
            use = [c for c in cpus if c in all_cpus]
            if not use:
                continue
            os.sched_setaffinity(0, use)
            print("bound to node %d (%d cpus): %.1f GB/s" % (k, len(use), h2d_gbs(a.mb << 20)))
        os.sched_setaffinity(0, all_cpus)
        return
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    gn, bdf = gpu_node(lr)
    keys = sorted(nd)
    target = {"local": gn if gn >= 0 else None, "node0": keys[0], "spread": keys[lr % len(keys)], "none": None}[a.mode]
    if target is not None:
        use = [c for c in nd[target] if c in all_cpus]
        if use:
            os.sched_setaffinity(0, use)
    h = torch.empty(a.mb << 20, dtype=torch.uint8).pin_memory()
    h.fill_(1)
    d = torch.empty(a.mb << 20, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(8):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    g = torch.tensor([(a.mb << 20) * 8 / (time.perf_counter() - t0) / 1e9, float(gn), float(-1 if target is None else target)], device="cuda")
    out = [torch.zeros_like(g) for _ in range(world)]
    dist.all_gather(out, g)
    if rank == 0:
        print("mode %-6s N=%d  GB/s per rank: %s  sum %.1f  | gpu numa nodes %s  ring on %s" % (
            a.mode, world, " ".join("%.1f" % float(x[0]) for x in out), sum(float(x[0]) for x in out),
            [int(x[1]) for x in out], [int(x[2]) for x in out]))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
