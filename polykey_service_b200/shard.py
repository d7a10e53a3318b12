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
