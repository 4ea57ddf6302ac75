#!/usr/bin/env python
"""Concurrent pinned device->host copy bandwidth of N ranks (one per GPU), the platform bound of the e2e metric for small trees
(844 MB of posteriors per 1M ALARM cases). Prints per-rank and aggregate GB/s, the GPUs' NUMA nodes, and repeats the
measurement with the pinned buffer bound (set_mempolicy) to the GPU's NUMA node when the container allows it.
  python -m torch.distributed.run --nproc-per-node N tools/pcie_probe.py"""
import ctypes
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def gpu_numa_node(i):
    try:
        bus = torch.cuda.get_device_properties(i)
        pci = "%04x:%02x:%02x.0" % (bus.pci_domain_id, bus.pci_bus_id, bus.pci_device_id)
        return int(open("/sys/bus/pci/devices/%s/numa_node" % pci).read()), pci
    except Exception as e:  # noqa: BLE001
        return -1, str(e)


def measure(tag):
    n = 840_000_000 // 8
    d = torch.empty(n, dtype=torch.float64, device="cuda")
    h = torch.empty(n, dtype=torch.float64).pin_memory()
    for _ in range(2):
        h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(5):
        h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 5
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    g = [torch.zeros_like(t) for _ in range(world)]
    if world > 1:
        dist.all_gather(g, t)
    else:
        g = [t]
    if rank == 0:
        ts = [float(x.item()) for x in g]
        print("%s: per-rank GB/s %s  aggregate %.1f GB/s" % (tag, " ".join("%.1f" % (n * 8 / x / 1e9) for x in ts), world * n * 8 / max(ts) / 1e9),
              flush=True)
    del h, d


node, pci = gpu_numa_node(local)
nodes = sorted(x for x in os.listdir("/sys/devices/system/node") if x.startswith("node")) if os.path.isdir("/sys/devices/system/node") else []
print("rank %d gpu %d pci %s numa_node %d, host nodes %s, cpus %s" % (rank, local, pci, node, nodes, sorted(os.sched_getaffinity(0))[:4]), flush=True)
measure("default placement")
if node >= 0:
    libc = ctypes.CDLL("libc.so.6", use_errno=True)
    mask = ctypes.c_ulong(1 << node)
    rc = libc.syscall(238, 2, ctypes.byref(mask), 64)  # set_mempolicy(MPOL_BIND, mask, maxnode) on x86_64
    print("rank %d set_mempolicy(bind node %d) rc %d errno %d" % (rank, node, rc, ctypes.get_errno()), flush=True)
    measure("bound to the GPU's node")
if world > 1:
    dist.destroy_process_group()
