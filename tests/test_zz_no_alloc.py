"""GPU: pk_route_device_noalloc — the entry point for callers that cannot allocate GPU memory.

Some hosts of the routing library do not own their GPU's memory: an inference engine that
reserved all of HBM at start-up, a fixed-budget pool, or a flusher thread that must never run
into the device-wide synchronisation cudaMalloc implies.  They hand the library device
pointers carved out of their own reservation; the scratch the kernel needs for long requests
is a static array inside the library.  This test plays such a caller (the buffers come from
torch, standing in for "somebody else's allocator"), calls the entry point through the C ABI
twice in a row on different batches, and requires its records to be bit-identical both to the
CPU oracle's and to those of the ordinary entry point, pk_route_batch, on the same batches
(integer work: tolerance 0).  Two more cases cover what the shared static scratch makes special:
a call that leaves the scratch dirty must not disturb the next one, and a caller's own
(non-default) stream works like the default one.

Kept in a file that pytest collects last, so a failure here cannot hide any other test; it
needs nothing from the other tests and also passes when run alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

from helpers import assert_records_equal

pytestmark = pytest.mark.gpu

SEG = 16384
MAX_REQUESTS = 65536          # PK_NOALLOC_MAX_REQUESTS in include/pk_route.h


def extra_segments(pay_len):
    """(request, segment) pairs for every segment >= 1 of every request longer than 16 KiB."""
    pairs = []
    for i, l in enumerate(pay_len.tolist()):
        for s in range(1, -(-l // SEG)):
            pairs.append((i, s))
    return np.array(pairs, dtype=np.uint32).reshape(-1, 2)


def to_device(torch, a, pad=0):
    raw = np.frombuffer(np.ascontiguousarray(a).tobytes() + b"\0" * pad, dtype=np.uint8)
    t = torch.empty(max(raw.size, 16), dtype=torch.uint8, device="cuda")
    if raw.size:
        t[:raw.size] = torch.from_numpy(raw.copy()).cuda()
    return t


def route_noalloc(pk, torch, batch, extra=None, stream=None):
    """Plays the caller: places the batch in 'its own' device memory (torch's allocator stands in
    for it), calls the entry point on `stream` (default: the current stream) and reads the records
    back.  `extra` overrides the (request, segment) list — used to leave the scratch dirty."""
    n = int(batch["pay_len"].size)
    if extra is None:
        extra = extra_segments(batch["pay_len"])
    d = {k: to_device(torch, batch[k], pad=32 if k == "payload" else 0) for k in batch}
    d_extra = to_device(torch, extra)
    d_out = torch.zeros(24 * n, dtype=torch.uint8, device="cuda")
    table = pk.default_table()
    torch.cuda.synchronize()
    rc = pk.lib().pk_route_device_noalloc(
        C.byref(table), d["payload"].data_ptr(), d["pay_off"].data_ptr(), d["pay_len"].data_ptr(),
        d["names"].data_ptr(), d["name_off"].data_ptr(), d_extra.data_ptr() if len(extra) else None,
        len(extra), n, int(batch["pay_len"].max()), d_out.data_ptr(),
        (stream or torch.cuda.current_stream()).cuda_stream)
    assert rc == 0, pk.STATUS.get(rc, rc)
    torch.cuda.synchronize()
    return np.frombuffer(d_out.cpu().numpy().tobytes(), dtype=pk.DECISION_DTYPE), len(extra)


def test_noalloc_entry_matches_route_batch_twice(pk, oracle):
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    router = pk.Router(device=0, max_requests=4096, max_payload_bytes=64 << 20, n_slots=1)
    try:
        # two different batches back to back; both hold requests of several segments, so the
        # static scratch is used, and the second call reuses entries the first one touched
        first = oracle.gen_batch(8801, 700, 0, 90000, mode=0)
        second = oracle.gen_batch(8802, 1500, 0, 40000, mode=3)
        table = oracle.default_table()
        for which, batch in (("first call", first), ("second call", second)):
            rc, want = oracle.route_batch(table, batch, threads=4)        # the checker: the CPU oracle
            assert rc == 0
            got, n_extra = route_noalloc(pk, torch, batch)
            assert n_extra > 0
            assert_records_equal(got, want, f"{which} vs oracle")
            assert_records_equal(got, router.route_batch(batch), f"{which} vs pk_route_batch")
        # a batch of short requests takes the path that needs no scratch at all
        short = oracle.gen_batch(8803, 3000, 0, 200, mode=0)
        got, n_extra = route_noalloc(pk, torch, short)
        assert n_extra == 0
        assert_records_equal(got, oracle.route_batch(table, short, threads=4)[1], "short batch vs oracle")
        assert_records_equal(got, router.route_batch(short), "short batch vs pk_route_batch")
    finally:
        router.close()


def test_noalloc_entry_survives_dirty_scratch_and_own_stream(pk, oracle):
    """Every call shares one static scratch array, and it keeps what the last call left in it.
    A call whose segment list is incomplete never finishes its long requests, so their scratch
    entries stay non-zero — the state a failed or interrupted caller leaves behind.  The next call
    must clear what it is going to use and produce exact records all the same; here it also runs on
    a stream the caller created, not the default one."""
    import torch
    torch.cuda.set_device(0)
    table = oracle.default_table()
    dirty = oracle.gen_batch(8811, 400, 40000, 90000, mode=0)           # every request has >= 3 segments
    full = extra_segments(dirty["pay_len"])
    # drop the LAST segment of every request: partial sums and arrival counts stay in the scratch
    keep = np.ones(len(full), dtype=bool)
    keep[np.r_[np.nonzero(np.diff(full[:, 0]))[0], len(full) - 1]] = False
    assert 0 < keep.sum() < len(full)
    route_noalloc(pk, torch, dirty, extra=full[keep])                   # records of this call are unfinished: ignored

    own = torch.cuda.Stream()
    batch = oracle.gen_batch(8812, 400, 0, 90000, mode=3)               # same request indices, correct list
    rc, want = oracle.route_batch(table, batch, threads=4)
    assert rc == 0
    got, n_extra = route_noalloc(pk, torch, batch, stream=own)
    assert n_extra > 0
    assert_records_equal(got, want, "after a call that left the scratch dirty, on the caller's own stream")
    # and once more on the default stream, now over a scratch the previous launch left clean
    got, _ = route_noalloc(pk, torch, batch)
    assert_records_equal(got, want, "repeat on the default stream")


def test_noalloc_entry_rejects_oversized_batches(pk):
    """More requests than the static scratch holds: a clear error, and no GPU work."""
    table = pk.default_table()
    dummy = np.zeros(16, dtype=np.uint8).ctypes.data        # never dereferenced: refused first
    rc = pk.lib().pk_route_device_noalloc(C.byref(table), dummy, dummy, dummy, dummy, dummy, None, 0,
                                          MAX_REQUESTS + 1, 16, dummy, None)
    assert rc == -4 and b"capacity" in pk.lib().pk_strerror(rc)          # PK_ERR_CAPACITY
    assert pk.lib().pk_route_device_noalloc(None, dummy, dummy, dummy, dummy, dummy, None, 0, 1, 16,
                                            dummy, None) == -1           # PK_ERR_ARG
