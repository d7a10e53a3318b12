"""Multi-GPU sharding of a request stream: shard = request id % n_gpus, no exchange step.

Requests are independent (the reference's MockService is stateless, internal/service/mock.go:14),
so every GPU routes its own requests with its own handle and a replica of the (tiny, read-only)
routing table; nothing crosses NVLink and no collective is needed on the data path.
"""
from __future__ import annotations

import numpy as np


def shard_ids(total: int, world: int, rank: int) -> np.ndarray:
    """Request ids owned by `rank`: rank, rank + world, rank + 2*world, ..."""
    return np.arange(rank, total, world, dtype=np.int64)


def select_requests(batch, ids):
    """The flat batch (include/pk_route.h layout) holding requests `ids` of `batch`, in that order.
    Payload starts stay 16-byte aligned."""
    ids = np.asarray(ids, dtype=np.int64)
    n = int(ids.size)
    pay_len = batch["pay_len"][ids].astype(np.uint32)
    pay_off = np.zeros(n + 1, dtype=np.uint64)
    if n:
        np.cumsum((pay_len.astype(np.uint64) + np.uint64(15)) & np.uint64(0xFFFFFFFFFFFFFFF0), out=pay_off[1:])
    payload = np.zeros(max(int(pay_off[-1]), 16), dtype=np.uint8)[:int(pay_off[-1])]
    nlen = (batch["name_off"][1:] - batch["name_off"][:-1])[ids].astype(np.uint32)
    name_off = np.zeros(n + 1, dtype=np.uint32)
    if n:
        np.cumsum(nlen, out=name_off[1:])
    names = np.empty(int(name_off[-1]), dtype=np.uint8)
    src_off, src_noff = batch["pay_off"], batch["name_off"]
    for j, i in enumerate(ids):
        o, l = int(src_off[i]), int(pay_len[j])
        payload[int(pay_off[j]):int(pay_off[j]) + l] = batch["payload"][o:o + l]
        a, b = int(src_noff[i]), int(src_noff[i + 1])
        names[int(name_off[j]):int(name_off[j + 1])] = batch["names"][a:b]
    return {"payload": payload, "pay_off": pay_off, "pay_len": pay_len, "names": names, "name_off": name_off}


def merge_shards(total: int, world: int, shard_records):
    """Inverse of sharding: shard_records[rank] holds the records of shard_ids(total, world, rank)."""
    out = np.empty(total, dtype=shard_records[0].dtype)
    for rank, rec in enumerate(shard_records):
        out[shard_ids(total, world, rank)] = rec
    return out


def parse_cpulist(text: str) -> list[int]:
    """'0-15,32-47' -> [0..15, 32..47] (the kernel's cpulist format)."""
    cpus: list[int] = []
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.extend(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_local_cpus(pci_bus_id: str, sysfs: str = "/sys/bus/pci/devices") -> list[int]:
    """CPUs on the NUMA node a GPU hangs off, from sysfs (<bdf>/local_cpulist); [] if unknown.
    pci_bus_id as CUDA/NVML print it: '00000000:1B:00.0' or '0000:1b:00.0'."""
    dom, _, rest = pci_bus_id.strip().lower().partition(":")
    bdf = f"{dom[-4:]:0>4}:{rest}"
    try:
        with open(f"{sysfs}/{bdf}/local_cpulist") as f:
            return parse_cpulist(f.read())
    except (OSError, ValueError):
        return []


def bind_to_gpu_numa(pci_bus_id: str, sysfs: str = "/sys/bus/pci/devices") -> list[int]:
    """Pin this process to the CPUs local to its GPU BEFORE it allocates page-locked arenas, so
    that first-touch places them on the GPU's NUMA node and every rank's copies stay off the
    inter-socket link (the multi-GPU limiter is host-side).  Returns the CPU set now
    in force; leaves the affinity alone (and returns []) when sysfs has no answer or the local
    set does not intersect what the process may use (containers with a cpuset)."""
    import os
    local = set(gpu_local_cpus(pci_bus_id, sysfs))
    allowed = os.sched_getaffinity(0)
    chosen = sorted(local & allowed)
    if not chosen or len(chosen) == len(allowed):
        return []
    os.sched_setaffinity(0, chosen)
    return chosen
