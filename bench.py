#!/usr/bin/env python
"""bench.py — routed requests/s of the batched routing stage on B200, beside the CPU path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c1|c2|c3]

A "step" is one pass of the hot path over one batch of synthetic requests.  The default
workload is BASELINE.json configs[1] (C1): 4,096 tool-execution requests x 2 KB payloads,
8 providers.  One JSON line is printed by rank 0:

  value      whole-job requests/s with the batches already resident in HBM: K launches of the
             fused kernel issued back to back (one CUDA graph of K kernel nodes) over a ring of
             resident batches larger than L2, timed with CUDA events on the launching stream
  e2e        the same metric through the C-ABI call a caller makes (pk_submit/pk_wait with HOST
             buffers: H2D copies, kernel, D2H copy inside the timed region)
  roofline   algorithmic bytes per launch / average launch duration, against the measured HBM
             peak in MEASURED_PEAKS.json
  cpu_baseline  the CPU oracle (work-equivalent port) timed on this box's host cores

`--impl reference` times the CPU path alone (the oracle port — the Go reference cannot be
built here: no Go toolchain, un-vendored spounge-proto) on the same workload.

Multi-GPU (torchrun, one rank per GPU): requests shard by request id, every rank routes its
own batches, no collective on the data path ("scaling": "weak").
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (requests per batch, payload bytes, description)
    "c1": (4096, 2048, "C1: 4096 requests x 2048 B payloads, 8 providers (BASELINE configs[1])"),
    "c2": (65536, 256, "C2: 65536 requests x 256 B payloads, 8 providers (BASELINE configs[2])"),
    "c3": (1024, 1 << 20, "C3: 1024 requests x 1 MiB payloads, 8 providers (BASELINE configs[3])"),
}
L2_BYTES = 126 << 20
METRIC, UNIT = "routed_requests_per_sec", "requests/s"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# Libraries (NCCL's version banner, for one) write to fd 1.  The contract is ONE JSON line on
# stdout, so everything but that line is sent to stderr: fd 1 is pointed at fd 2 for the whole
# run and the JSON line is written to the saved, real stdout.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line: dict):
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


def algorithmic_bytes(batch) -> int:
    """payload_len + name_len + 12 (offsets read) + 24 (record written) per request,
    the payload counted once."""
    n = int(batch["pay_len"].size)
    return int(batch["pay_len"].sum()) + int(batch["name_off"][-1] - batch["name_off"][0]) + 36 * n


class ClockSampler(threading.Thread):
    """Samples SM clock and clock-event reasons through NVML while the GPU is under load."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake", 0x2: "applications_clocks_setting"}

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.samples, self.stop_flag, self.ok, self.max_mhz = [], False, False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:                      # NVML missing: report that, do not invent clocks
            self.err = repr(e)

    def _reasons(self):
        for fn in ("nvmlDeviceGetCurrentClocksEventReasons", "nvmlDeviceGetCurrentClocksThrottleReasons"):
            if hasattr(self.nv, fn):
                try:
                    return int(getattr(self.nv, fn)(self.h))
                except Exception:
                    continue
        return 0

    def run(self):
        if not self.ok:
            return
        while not self.stop_flag:
            try:
                sm = int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                self.samples.append((time.perf_counter(), sm, self._reasons()))
            except Exception:
                pass
            time.sleep(0.001)

    def summary(self, windows):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "NVML unavailable: " + self.err}
        inside = [(sm, r) for (t, sm, r) in self.samples if any(a <= t <= b for a, b in windows)]
        src = "timed regions"
        if len(inside) < 3:
            inside, src = [(sm, r) for (_, sm, r) in self.samples], "whole measurement phase (timed regions too short to sample)"
        mask = 0
        for _, r in inside:
            mask |= r
        return {"sm_mhz": statistics.median([sm for sm, _ in inside]) if inside else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(v for k, v in self.REASONS.items() if mask & k),
                "samples": len(inside), "window": src}


def dist_setup(n_gpus: int):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        # keep this rank's threads and (first-touch) its page-locked arenas on the GPU's NUMA node
        try:
            from polykey_service_b200 import shard
            pr = torch.cuda.get_device_properties(local)
            bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
            cpus = shard.bind_to_gpu_numa(bdf)
            print(f"[rank {rank}] gpu {bdf}: bound to {len(cpus)} local cpus" if cpus else
                  f"[rank {rank}] gpu {bdf}: affinity unchanged", file=sys.stderr)
        except (AttributeError, OSError) as e:
            print(f"[rank {rank}] NUMA binding skipped: {e}", file=sys.stderr)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        return world, rank, local, dist
    return 1, 0, 0, None


def reduce_max(dist, local, x: float) -> float:
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local}")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier_sync(dist, local):
    import torch
    if dist is not None:
        dist.barrier()
    if torch.cuda.is_available():
        torch.cuda.synchronize(local)


# ------------------------------------------------------------------ CPU arms ------
def cpu_arm(O, table, batch, threads: int, budget_s: float, switch_only: bool = False, min_reps: int = 2):
    """Times the oracle over `batch` repeatedly for about budget_s; returns (requests/s, reps)."""
    n = int(batch["pay_len"].size)
    O.route_batch(table, batch, threads=threads, switch_only=switch_only, fast=True)      # warm
    reps, t0 = 0, time.perf_counter()
    while True:
        rc, _ = O.route_batch(table, batch, threads=threads, switch_only=switch_only, fast=True)
        assert rc == 0
        reps += 1
        dt = time.perf_counter() - t0
        if (dt >= budget_s and reps >= min_reps) or reps >= 100000:
            break
    return n * reps / dt, reps


def run_reference(args):
    """--impl reference: the CPU path on this box's host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pk_oracle as O
    O.build()
    n, plen, desc = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    table = O.default_table()
    # bounded sample: for C3 a step is 64 of the 1,024 requests (64 MiB), otherwise the full batch
    n_step = n if n * plen <= (64 << 20) else max(1, (64 << 20) // plen)
    batches = [O.gen_batch(O.SEED_BASE + 1 + i, n_step, plen, mode=0, threads=min(cores, 16)) for i in range(4)]
    # the efficient CPU port (oracle/pk_cpu_fast.c) is what is timed; check it against the plain
    # oracle once, outside the timed region
    assert (O.route_batch(table, batches[0], threads=cores, fast=True)[1] ==
            O.route_batch(table, batches[0], threads=cores)[1]).all()
    for w in range(args.warmup):
        O.route_batch(table, batches[w % 4], threads=cores, fast=True)
    t0 = time.perf_counter()
    for k in range(args.steps):
        rc, _ = O.route_batch(table, batches[k % 4], threads=cores, fast=True)
        assert rc == 0
    dt = time.perf_counter() - t0
    value = n_step * args.steps / dt
    # CPU-A: the reference-faithful work; the threaded form spawns its threads per call, so for a
    # 4,096-request batch one core is usually faster — report the better of the two
    sw_all, _ = cpu_arm(O, table, batches[0], cores, 1.0, switch_only=True)
    sw_one, _ = cpu_arm(O, table, batches[0], 1, 1.0, switch_only=True)
    sw, sw_cores = (sw_all, cores) if sw_all >= sw_one else (sw_one, 1)
    sample = f"{args.steps} steps x {n_step} requests x {plen} B (oracle/pk_cpu_fast.c port, {cores} threads)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/u32/u64 integer",
        "data": "synthetic", "config": {"workload": desc, "batch_requests": n_step, "payload_bytes": plen,
                                        "providers": 8, "parallelism": "host threads, static partition"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "reference_switch_only": {"value": sw, "unit": UNIT, "cores": sw_cores,
                                  "note": "CPU-A: only the work the Go reference does per request "
                                          "(mock.go:32-64 switch; payload never read)"},
        "note": "Go reference cannot be built on this box (no Go toolchain, un-vendored spounge-proto); "
                "this is the C oracle port doing the work-equivalent scan+hash+select",
    }
    emit(line)


# ------------------------------------------------------------------ GPU arm ------
def run_gpu(args):
    import torch
    import polykey_service_b200 as pk
    from oracle import pk_oracle as O            # cpu_baseline leg + input generator only

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    world, rank, local, dist = dist_setup(args.gpus)
    O.build()
    n, plen, desc = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    batch_bytes = n * ((plen + 15) & ~15)

    # ring of resident batches larger than L2 so every launch streams from HBM
    ring = max(2, min(64, -(-(2 * L2_BYTES + batch_bytes) // batch_bytes)))
    if batch_bytes > L2_BYTES:
        ring = 2
    depth = 3                                      # e2e pipeline depth (slots in flight)
    router = pk.Router(device=local, max_requests=n, max_payload_bytes=batch_bytes,
                       n_slots=ring + depth)
    log(f"[rank {rank}] workload {args.workload}: n={n} payload={plen} ring={ring} "
        f"({ring * batch_bytes / 2**20:.0f} MiB resident)")

    # request ids: rank r owns ids r, r+world, ... (req_id % G sharding); batch j of rank r starts at
    # a disjoint id range, so every rank routes different requests
    def make_batch(j, payload_out=None):
        first_id = (rank * 1_000_003 + j) * n
        return O.gen_batch(O.SEED_BASE + 1, n, plen, mode=0, first_id=first_id, threads=min(cores, 16),
                           payload_out=payload_out)

    table = O.default_table()
    check_batch = None
    for j in range(ring):
        b = make_batch(j)
        router.stage(j, b)
        if j == 0:
            check_batch = b
    alg_bytes = algorithmic_bytes(check_batch)

    # correctness gate before timing anything: slot 0 against the oracle, bit-exact
    router.launch(0)
    got = router.fetch(0, n)
    rc, want = O.route_batch(table, check_batch, threads=cores)
    assert rc == 0 and (got == want).all(), "GPU records differ from the oracle — refusing to time"

    sampler = ClockSampler(local)
    sampler.start()
    slots = list(range(ring))
    PDL = 0x100                                                   # PK_LAUNCH_PDL, include/pk_route.h
    issue = "plain" if args.no_graph else args.issue
    ISSUE = {"pdl": (1 | PDL, "one CUDA graph of K kernel nodes in a chain, programmatic dependent launch "
                              "(a kernel may start while its predecessor drains; batches are independent)"),
             "serial": (1, "one CUDA graph of K kernel nodes in a chain, each starting after its predecessor completed"),
             "plain": (0, "K plain launches")}
    use_graph, issue_desc = ISSUE[issue]

    # bring clocks up, then the W warm-up steps
    t_end = time.perf_counter() + (0.0 if args.no_prewarm else 0.4)
    while time.perf_counter() < t_end:
        router.launch_many(slots, 512, use_graph=False)
    router.launch_many(slots, max(3, args.warmup), use_graph=use_graph)

    # ---- timed: K steps, device-resident ----
    launches_before = router.launch_count
    barrier_sync(dist, local)
    w0 = time.perf_counter()
    ms = router.launch_many(slots, args.steps, use_graph=use_graph)
    torch.cuda.synchronize(local)
    w1 = time.perf_counter()
    barrier_sync(dist, local)
    launches = router.launch_count - launches_before
    ms = reduce_max(dist, local, ms)
    value = world * args.steps * n / (ms / 1e3)
    per_launch_s = ms / 1e3 / args.steps
    # the other issue modes, for the record (same K launches, same stream)
    other = {}
    for name, mode in (("pdl", 1 | PDL), ("serial", 1), ("plain", 0), ("plain_pdl", PDL), ("four_chains", 4)):
        if mode != use_graph:
            other[name] = reduce_max(dist, local, router.launch_many(slots, args.steps, use_graph=mode)) / args.steps
    ms_serial = other.get("serial", ms / args.steps)

    # ---- e2e: host buffers through pk_submit / pk_wait, copies inside the timed region ----
    e2e_batches, host_bufs = [], []
    for j in range(depth + 1):
        buf = router.host_alloc(batch_bytes + 64)                 # pinned arena owned by the C side
        e2e_batches.append(make_batch(100 + j, payload_out=buf))
        host_bufs.append(buf)
    outs = [np.frombuffer(router.host_alloc(24 * n), dtype=pk.DECISION_DTYPE) for _ in range(depth)]
    e2e_slots = [ring + d for d in range(depth)]
    meta_bytes = int(check_batch["pay_off"].nbytes + check_batch["pay_len"].nbytes +
                     check_batch["name_off"].nbytes + check_batch["names"].nbytes)
    h2d = int(check_batch["pay_off"][n - 1]) + int(check_batch["pay_len"][n - 1]) + meta_bytes
    d2h = 24 * n
    k_e2e = max(20, min(args.steps, int(3.0 / max(1e-5, (h2d / 20e9)))))   # ~3 s at a pessimistic 20 GB/s

    def e2e_pipelined(k):
        inflight = [False] * depth
        t0 = time.perf_counter()
        for s in range(k):
            d = s % depth
            if inflight[d]:
                router.wait(e2e_slots[d])
            router.submit(e2e_slots[d], e2e_batches[s % len(e2e_batches)], outs[d])
            inflight[d] = True
        for d in range(depth):
            if inflight[d]:
                router.wait(e2e_slots[d])
        return time.perf_counter() - t0

    e2e_pipelined(max(3, min(10, k_e2e)))                          # warm
    barrier_sync(dist, local)
    direct_before = router.direct_count
    e0 = time.perf_counter()
    e2e_s = e2e_pipelined(k_e2e)
    e1 = time.perf_counter()
    barrier_sync(dist, local)
    used_direct = router.direct_count - direct_before == k_e2e       # the library's own choice for this batch size
    e2e_s = reduce_max(dist, local, e2e_s)
    e2e_value = world * k_e2e * n / e2e_s
    rc, want = O.route_batch(table, e2e_batches[(k_e2e - 1) % len(e2e_batches)], threads=cores)
    assert (outs[(k_e2e - 1) % depth] == want).all(), "e2e records differ from the oracle"
    # the other transfer path, for the record (same call, same buffers)
    router.set_direct_bytes(0 if used_direct else 1 << 40)
    k_alt = max(10, k_e2e // 4)
    e2e_pipelined(3)
    barrier_sync(dist, local)
    e2e_alt_s = reduce_max(dist, local, e2e_pipelined(k_alt)) / k_alt
    router.set_direct_bytes(16 << 20)
    transfer = {True: "none issued: the kernel reads the page-locked arenas over PCIe and writes the records into the "
                      "caller's page-locked buffer (pk_set_direct_bytes, default for batches up to 16 MiB)",
                False: "cudaMemcpyAsync H2D (payload, metadata) and D2H (records) around the kernel"}

    # synchronous single-call latency (what one flush costs a waiting request)
    lat = []
    for s in range(min(k_e2e, 300)):
        t0 = time.perf_counter()
        router.route_batch(e2e_batches[s % len(e2e_batches)], out=outs[0])
        lat.append((time.perf_counter() - t0) * 1e6)
    lat.sort()
    sampler.stop_flag = True
    sampler.join(timeout=2)
    clocks = sampler.summary([(w0, w1), (e0, e1)])

    if rank != 0:
        router.close()
        return

    # ---- CPU baseline on this box's host cores (rank 0, N=1 only) ----
    cpu_baseline = None
    extra_cpu = {}
    if world == 1 and not args.no_cpu:
        sample_batch = check_batch if batch_bytes <= (64 << 20) else make_sample(O, plen)
        v_all, reps = cpu_arm(O, table, sample_batch, cores, 8.0)
        v_one, _ = cpu_arm(O, table, sample_batch, 1, 4.0)
        v_sw, _ = cpu_arm(O, table, sample_batch, cores, 1.5, switch_only=True)
        v_sw1, _ = cpu_arm(O, table, sample_batch, 1, 1.0, switch_only=True)
        cpu_baseline = {"value": v_all, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"{reps} passes over one batch of {int(sample_batch['pay_len'].size)} requests x {plen} B "
                                  f"(about 8 s), oracle/pk_cpu_fast.c scan+hash+select, {cores} threads"}
        extra_cpu = {"cpu_baseline_1core": {"value": v_one, "unit": UNIT, "cores": 1},
                     "cpu_switch_only": {"value": v_sw, "unit": UNIT, "cores": cores,
                                         "note": "CPU-A: what the Go reference actually does per request "
                                                 "(mock.go:32-64 switch; never reads the payload)"},
                     "cpu_switch_only_1core": {"value": v_sw1, "unit": UNIT, "cores": 1}}

    peak, peak_src = measured_peak()
    achieved = alg_bytes / per_launch_s / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(args.workload)
        except Exception:
            traffic = None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8/u32/u64 integer", "data": "synthetic",
        "config": {"workload": desc, "batch_requests": n, "payload_bytes": plen, "providers": 8,
                   "parallelism": f"dp{world}: requests sharded by id, no collective",
                   "l2": f"ring of {ring} resident batches = {ring * batch_bytes / 2**20:.0f} MiB > 126 MiB L2, "
                         "launch k reads batch k mod ring",
                   "issue": issue_desc + "; one stream, CUDA events on that stream"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "kernel": "route_kernel (fused scan+hash+select)",
                     "algorithmic_bytes_per_launch": alg_bytes, "avg_launch_us": per_launch_s * 1e6},
        "cpu_baseline": cpu_baseline,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": k_e2e, "pipeline_depth": depth, "ms_per_step": e2e_s / k_e2e * 1e3,
                "api": "pk_submit/pk_wait, pinned host arenas", "copies": transfer[used_direct],
                "other_path": {"copies": transfer[not used_direct], "ms_per_step": e2e_alt_s * 1e3,
                               "value": world * n / e2e_alt_s, "unit": UNIT},
                "sync_call_latency_us": {"p50": lat[len(lat) // 2], "p99": lat[min(len(lat) - 1, int(len(lat) * 0.99))],
                                         "api": "pk_route_batch, one batch per call"}},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "issue_modes_ms_per_step": dict(other, **{issue: ms / args.steps}),
        "serial": {"ms_per_step": ms_serial, "value": world * n / (ms_serial / 1e3), "unit": UNIT,
                   "frac": alg_bytes / (ms_serial / 1e3) / 1e9 / peak,
                   "note": "the same K launches with every kernel waiting for its predecessor to complete: what ONE "
                           "batch costs on an idle GPU (launch + two dependent HBM round trips + drain)"},
    }
    line.update(extra_cpu)
    emit(line)
    router.close()


def make_sample(O, plen):
    n_s = max(1, (64 << 20) // plen)
    return O.gen_batch(O.SEED_BASE + 1, n_s, plen, mode=0, threads=16)



# ------------------------------------------------------------------ C4: strong scaling ------
C4_TOTAL = 1_000_000


def run_gpu_c4(args):
    """BASELINE configs[4]: ONE stream of 1,000,000 requests x 2 KB sharded by request id over the
    G ranks (rank r owns r, r+G, ...; no exchange), each rank working through its shard in batches
    of --c4-batch requests.  Strong scaling: the work is fixed, `value` = 1e6 / the slowest rank's
    time for one pass over its shard.  A step is one batch on every rank."""
    import torch
    import polykey_service_b200 as pk
    from oracle import pk_oracle as O

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    world, rank, local, dist = dist_setup(args.gpus)
    O.build()
    n, plen = args.c4_batch, 2048
    cores = os.cpu_count() or 1
    mine = len(range(rank, C4_TOTAL, world))                  # requests of this rank's shard
    nb = -(-mine // n)                                        # its batches; the last one is ragged
    batch_bytes = n * plen
    depth = 3
    router = pk.Router(device=local, max_requests=n, max_payload_bytes=batch_bytes, n_slots=nb + depth)
    table = O.default_table()
    log(f"[rank {rank}] C4: shard of {mine} requests = {nb} batches of {n} ({nb * batch_bytes / 2**20:.0f} MiB resident)")

    def shard_batch(j, payload_out=None):
        # rank-local numbering: the shard's k-th request is generated from id rank * 2^32 + k, so
        # shards are disjoint and the same whatever the batch size
        cnt = min(n, mine - j * n)
        return O.gen_batch(O.SEED_BASE + 4, cnt, plen, mode=0, first_id=(rank << 32) + j * n,
                           threads=min(cores, 16), payload_out=payload_out)

    first = last = None
    for j in range(nb):
        b = shard_batch(j)
        router.stage(j, b)
        if j == 0:
            first = b
        if j == nb - 1:
            last = b
    for slot, b in ((0, first), (nb - 1, last)):              # correctness gate: full and ragged batch
        router.launch(slot)
        got = router.fetch(slot, int(b["pay_len"].size))
        assert (got == O.route_batch(table, b, threads=cores)[1]).all(), "GPU records differ from the oracle"
    alg_bytes = algorithmic_bytes(first) * (mine / n)        # per pass over the shard

    sampler = ClockSampler(local)
    sampler.start()
    slots = list(range(nb))
    PDL = 0x100
    reps = max(3, args.steps // nb)
    t_end = time.perf_counter() + (0.0 if args.no_prewarm else 0.4)
    while time.perf_counter() < t_end:
        router.launch_many(slots, nb, use_graph=False)
    router.launch_many(slots, nb * max(1, args.warmup // nb + 1), use_graph=1 | PDL)
    launches_before = router.launch_count
    barrier_sync(dist, local)
    w0 = time.perf_counter()
    ms = router.launch_many(slots, nb * reps, use_graph=1 | PDL)
    torch.cuda.synchronize(local)
    w1 = time.perf_counter()
    barrier_sync(dist, local)
    launches = router.launch_count - launches_before
    pass_s = reduce_max(dist, local, ms) / 1e3 / reps          # slowest rank's time for ONE pass over its shard
    ms_serial = reduce_max(dist, local, router.launch_many(slots, nb * reps, use_graph=1)) / reps

    # ---- e2e: the shard streamed from page-locked host arenas through pk_submit / pk_wait ----
    host = [shard_batch(j % nb, payload_out=router.host_alloc(batch_bytes + 64)) for j in range(depth + 1)]
    ragged = shard_batch(nb - 1, payload_out=router.host_alloc(batch_bytes + 64))
    outs = [np.frombuffer(router.host_alloc(24 * n), dtype=pk.DECISION_DTYPE) for _ in range(depth)]
    e2e_slots = [nb + d for d in range(depth)]

    def e2e_pass():
        inflight = [None] * depth
        t0 = time.perf_counter()
        for j in range(nb):
            d = j % depth
            if inflight[d] is not None:
                router.wait(e2e_slots[d])
            b = ragged if j == nb - 1 else host[j % len(host)]
            cnt = int(b["pay_len"].size)
            router.submit(e2e_slots[d], b, outs[d][:cnt])
            inflight[d] = cnt
        for d in range(depth):
            if inflight[d] is not None:
                router.wait(e2e_slots[d])
        return time.perf_counter() - t0

    e2e_pass()
    barrier_sync(dist, local)
    e0 = time.perf_counter()
    e2e_reps = 3
    e2e_s = sum(e2e_pass() for _ in range(e2e_reps)) / e2e_reps
    e1 = time.perf_counter()
    barrier_sync(dist, local)
    e2e_s = reduce_max(dist, local, e2e_s)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    clocks = sampler.summary([(w0, w1), (e0, e1)])
    if rank != 0:
        router.close()
        return
    meta = int(first["pay_off"].nbytes + first["pay_len"].nbytes + first["name_off"].nbytes + first["names"].nbytes)
    peak, peak_src = measured_peak()
    achieved = alg_bytes / pass_s / 1e9
    cpu_baseline = None
    if world == 1 and not args.no_cpu:
        v_all, reps_cpu = cpu_arm(O, table, first, cores, 6.0)
        cpu_baseline = {"value": v_all, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"{reps_cpu} passes over one batch of {n} requests x {plen} B, "
                                  f"oracle/pk_cpu_fast.c scan+hash+select, {cores} threads"}
    emit({
        "metric": METRIC, "value": C4_TOTAL / pass_s, "unit": UNIT, "n_gpus": world, "steps": nb * reps,
        "warmup": max(3, args.warmup), "ms_per_step": pass_s * 1e3 / nb, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u8/u32/u64 integer", "data": "synthetic",
        "config": {"workload": f"C4: one stream of {C4_TOTAL} requests x {plen} B sharded by request id over {world} GPU(s), "
                               f"per-GPU batches of {n} (BASELINE configs[4])",
                   "shard_requests_rank0": mine, "batches_per_shard": nb, "batch_requests": n, "payload_bytes": plen,
                   "parallelism": f"dp{world}: requests sharded by id, no collective",
                   "l2": f"the shard is resident ({nb * batch_bytes / 2**20:.0f} MiB on rank 0) and each pass walks all of it",
                   "issue": "one CUDA graph of K kernel nodes in a chain, programmatic dependent launch; one stream, "
                            "CUDA events on that stream"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "peak_source": peak_src, "kernel": "route_kernel (fused scan+hash+select)",
                     "algorithmic_bytes_per_launch": algorithmic_bytes(first), "avg_launch_us": pass_s * 1e6 / nb},
        "cpu_baseline": cpu_baseline,
        "e2e": {"value": C4_TOTAL / e2e_s, "unit": UNIT, "h2d_bytes_per_step": batch_bytes + meta, "d2h_bytes_per_step": 24 * n,
                "steps": nb * e2e_reps, "pipeline_depth": depth, "ms_per_step": e2e_s * 1e3 / nb,
                "api": "pk_submit/pk_wait, pinned host arenas", "pass_ms": e2e_s * 1e3},
        "gpu_launches": int(launches), "clocks": clocks,
        "pass_ms": pass_s * 1e3, "serial": {"pass_ms": ms_serial, "value": C4_TOTAL / (ms_serial / 1e3), "unit": UNIT},
    })
    router.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c1", choices=sorted(WORKLOADS) + ["c4"])
    ap.add_argument("--c4-batch", type=int, default=4096, help="requests per batch of the C4 stream (4096 / 16384 / 65536)")
    ap.add_argument("--issue", default="pdl", choices=["pdl", "serial", "plain"],
                    help="how the K timed launches are issued on their one stream (default: graph chain + programmatic dependent launch)")
    ap.add_argument("--no-graph", action="store_true", help="same as --issue plain")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-prewarm", action="store_true",
                    help="skip the 0.4 s clock ramp-up before the W warm-up steps (for ncu launch lists)")
    args = ap.parse_args()
    if args.steps < 1:
        raise SystemExit("--steps must be >= 1")
    if args.impl == "reference":
        if args.workload == "c4":
            args.workload = "c1"          # the same 4,096 x 2 KB batches, timed on the host cores
        run_reference(args)
    elif args.workload == "c4":
        run_gpu_c4(args)
    else:
        run_gpu(args)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        try:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.destroy_process_group()
        except Exception:
            pass


if __name__ == "__main__":
    main()
