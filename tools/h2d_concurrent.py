"""Aggregate pinned-memory H2D bandwidth of the box when 1 / 2 / 4 / 8 ranks copy at the same time.

Run under torchrun with N ranks: every rank copies a 15.36 MB page-locked block (one bench step's input) to its GPU `reps` times back to
back, all ranks start together (barrier); prints per-rank and aggregate GB/s for (a) cudaHostAlloc default, (b) write-combined,
(c) default memory with the rank pinned to the CPUs of its GPU's NUMA node before the allocation.  Rank 0 prints one JSON line."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("gloo")
rt = C.CDLL("libcudart.so.12") if os.path.exists("/usr/local/cuda/lib64/libcudart.so.12") else C.CDLL("libcudart.so")
NBYTES = 15_360_000
REPS = 200


def numa_node_of_gpu(idx):
    try:
        bus = torch.cuda.get_device_properties(idx).pci_bus_id if hasattr(torch.cuda.get_device_properties(idx), "pci_bus_id") else None
        import subprocess
        out = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(idx)], capture_output=True, text=True).stdout.strip()
        dev = out.lower()[4:] if out.lower().startswith("0000") else out.lower()
        p = f"/sys/bus/pci/devices/{dev}/numa_node"
        return int(open(p).read()) if os.path.exists(p) else -1
    except Exception:  # noqa: BLE001
        return -1


def host_alloc(nbytes, flags):
    p = C.c_void_p()
    e = rt.cudaHostAlloc(C.byref(p), C.c_size_t(nbytes), C.c_uint(flags))
    assert e == 0, e
    a = np.frombuffer((C.c_char * nbytes).from_address(p.value), dtype=np.uint8)
    a[:] = 7     # first touch by this rank's thread
    return p, a


def measure(flags):
    p, a = host_alloc(NBYTES, flags)
    d = torch.empty(NBYTES, dtype=torch.uint8, device="cuda")
    st = torch.cuda.Stream()
    rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    for _ in range(5):
        rt.cudaMemcpyAsync(d.data_ptr(), p, NBYTES, 1, st.cuda_stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(REPS):
        rt.cudaMemcpyAsync(d.data_ptr(), p, NBYTES, 1, st.cuda_stream)
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    rt.cudaFreeHost(p)
    return NBYTES * REPS / ms / 1e6


def gather(x):
    if world == 1:
        return [x]
    out = [None] * world
    dist.all_gather_object(out, x)
    return out


res = {"world": world, "bytes_per_copy": NBYTES, "copies": REPS}
res["default"] = gather(measure(0))
res["write_combined"] = gather(measure(4))     # cudaHostAllocWriteCombined
node = numa_node_of_gpu(local)
cpus = None
try:
    if node >= 0:
        txt = open(f"/sys/devices/system/node/node{node}/cpulist").read().strip()
        cpus = set()
        for part in txt.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        os.sched_setaffinity(0, cpus)
except Exception:  # noqa: BLE001
    cpus = None
res["numa_bound"] = gather(measure(0))
res["numa_node_of_gpu"] = gather(node)
if rank == 0:
    for k in ("default", "write_combined", "numa_bound"):
        res[k + "_aggregate_GBps"] = float(sum(res[k]))
    try:
        res["numa_nodes_online"] = open("/sys/devices/system/node/online").read().strip()
        res["cpus"] = os.cpu_count()
    except Exception:  # noqa: BLE001
        pass
    print(json.dumps(res), flush=True)
if world > 1:
    dist.destroy_process_group()
