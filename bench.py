#!/usr/bin/env python
"""bench.py — routed requests/s of the batched routing stage on B200, beside the CPU path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c1|c2|c3|c4]

A "step" is one pass of the hot path over one batch of synthetic requests.  The default
workload is BASELINE.json configs[1] (C1): 4,096 tool-execution requests x 2 KB payloads,
8 providers.  One JSON line is printed by rank 0:

  value         whole-job requests/s with the batches already resident in HBM: K launches of the
                fused kernel issued back to back (one CUDA graph of K kernel nodes in four parallel
                chains, programmatic dependent launch inside each) over a ring of
                resident batches larger than L2, timed with CUDA events on the launching stream
  steady_state  the same issue mode over >= 2,000 launches (fill and drain of the overlapped
                pipeline amortised) — `value` at the driver's K, this one K-independent
  e2e           the same metric through the C-ABI call a caller makes (pk_submit/pk_wait with HOST
                buffers: H2D copies, kernel, D2H copy inside the timed region); e2e.with_pack
                starts one step earlier, from per-request heap buffers, through the batching
                front-end at the service.Service seam (producers copy into the pinned arena)
  roofline      algorithmic bytes per launch / average launch duration, against the measured HBM
                peak in MEASURED_PEAKS.json
  workloads     (N=1) compact records for BASELINE configs[2] and [3]: C2 65,536 x 256 B and
                C3 1,024 x 1 MiB — ms per launch, roofline fraction, e2e, parity gate
  c4            BASELINE configs[4]: ONE stream of 1,000,000 requests x 2 KB sharded by request id
                over the N ranks (strong scaling): time of a pass, kernel-only and end to end
  latency       p50/p99 per-request routing latency through the batching front-end, open loop
  cpu_baseline  the CPU port (work-equivalent scan+hash+select) timed on this box's host cores

`--impl reference` times the CPU path alone (the oracle port — the Go reference cannot be
built here: no Go toolchain, un-vendored spounge-proto) on the same workload.

Multi-GPU (torchrun, one rank per GPU): requests shard by request id, every rank routes its
own batches, no collective on the data path ("scaling": "weak").
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import statistics
import subprocess
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
DTYPE = "u8/u32/u64 integer"
PDL = 0x100                                                        # PK_LAUNCH_PDL, include/pk_route.h


def FUSE(m):                                                       # PK_LAUNCH_FUSE(m): m staged batches per kernel launch
    return m << 12


ISSUE = {"pdl4": (4 | PDL, "one CUDA graph of K kernel nodes in four parallel chains (a batch's chain = its ring position mod 4), "
                            "programmatic dependent launch inside each chain (a kernel may start while its predecessor "
                            "drains; batches are independent)"),
         "pdl": (1 | PDL, "one CUDA graph of K kernel nodes in a chain, programmatic dependent launch "
                          "(a kernel may start while its predecessor drains; batches are independent)"),
         "fused": (1 | PDL | FUSE(32), "up to 32 staged batches per kernel launch (grid.y = batch; PK_LAUNCH_FUSE), the launches "
                                       "in one CUDA graph chain with programmatic dependent launch — K = 20 is ONE launch, "
                                       "whose duration a profiler can check directly"),
         "serial": (1, "one CUDA graph of K kernel nodes in a chain, each starting after its predecessor completed"),
         "plain": (0, "K plain launches")}
C4_TOTAL = 1_000_000


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


def config_of(desc: str, n: int, plen: int) -> dict:
    """The `config` object — the same keys in the product arm and the reference arm."""
    return {"workload": desc, "batch_requests": n, "payload_bytes": plen, "providers": 8}


def algorithmic_bytes(batch) -> int:
    """payload_len + name_len + 12 (offsets read) + 24 (record written) per request,
    the payload counted once."""
    n = int(batch["pay_len"].size)
    return int(batch["pay_len"].sum()) + int(batch["name_off"][-1] - batch["name_off"][0]) + 36 * n


def kernel_sources_sha() -> str:
    """Identifies the kernel code a DRAM-traffic capture belongs to."""
    h = hashlib.sha256()
    for f in ("pk_kernels.cu", "pk_kernels.cuh", "pk_spec.cuh"):
        with open(os.path.join(ROOT, "polykey_service_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def ncu_traffic(workload: str):
    """dram__bytes_read + write per launch from the committed ncu capture — only while that capture
    was taken from the kernel sources this run uses; otherwise null (never a stale constant)."""
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        t = json.load(open(tpath))
    except Exception:
        return None, "no ncu capture on file (profiles/ncu_traffic.json)"
    if t.get("kernel_sources_sha") != kernel_sources_sha():
        return None, (f"the ncu capture on file ({t.get('captured_at', 'round 1')}) is of older kernel sources "
                      f"(sha {t.get('kernel_sources_sha')}): not quoted")
    return t.get(workload), f"ncu --set full capture, {t.get('captured_at')}, same kernel sources (sha {t['kernel_sources_sha']})"


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


def reduce_sum(dist, local, x: float) -> float:
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local}")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def barrier_sync(dist, local):
    import torch
    if dist is not None:
        dist.barrier()
    if torch.cuda.is_available():
        torch.cuda.synchronize(local)


class Ctx:
    """What every leg of the GPU arm needs."""

    def __init__(self, args):
        import torch
        import polykey_service_b200 as pk
        from oracle import pk_oracle as O            # cpu_baseline leg, parity gates and input generator only
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback "
                             "(use --impl reference for the CPU arm)")
        self.args, self.torch, self.pk, self.O = args, torch, pk, O
        self.world, self.rank, self.local, self.dist = dist_setup(args.gpus)
        O.build()
        self.cores = os.cpu_count() or 1
        self.gen_threads = min(self.cores, 16)
        self.table = O.default_table()
        self.peak, self.peak_src = measured_peak()
        self.windows = []                             # (t0, t1) wall-clock windows of the timed regions

    def max(self, x):
        return reduce_max(self.dist, self.local, x)

    def barrier(self):
        barrier_sync(self.dist, self.local)


# ------------------------------------------------------------------ CPU arms ------
def llc_bytes() -> int:
    """Size of the last-level cache of cpu0 (sysfs), 0 if unknown."""
    best = 0
    base = "/sys/devices/system/cpu/cpu0/cache"
    try:
        for d in os.listdir(base):
            try:
                txt = open(os.path.join(base, d, "size")).read().strip()
                mult = {"K": 1 << 10, "M": 1 << 20, "G": 1 << 30}.get(txt[-1], 1)
                best = max(best, int(txt.rstrip("KMG")) * mult)
            except (OSError, ValueError):
                continue
    except OSError:
        pass
    return best


def cpu_ring(O, n: int, plen: int, threads: int, seed_shift: int = 0):
    """The CPU arm's inputs: a ring of distinct batches larger than the last-level cache, so that —
    like the GPU arm, whose ring is larger than L2 and whose end-to-end leg pulls every byte out of
    host DRAM — each pass streams its payload from memory instead of re-reading one cached batch."""
    batch_bytes = n * ((plen + 15) & ~15)
    target = max(256 << 20, 2 * llc_bytes())
    k = max(2, min(64, -(-target // batch_bytes)))
    return [O.gen_batch(O.SEED_BASE + 1 + seed_shift, n, plen, mode=0, first_id=(7_000_003 + j) * n, threads=min(threads, 16))
            for j in range(k)]


def cpu_arm(O, table, batches, threads: int, budget_s: float, switch_only: bool = False, min_reps: int = 2):
    """Times the CPU port over the ring `batches` (a list; one batch = one step) for about budget_s;
    returns (requests/s, steps)."""
    if isinstance(batches, dict):
        batches = [batches]
    n = int(batches[0]["pay_len"].size)
    out = np.zeros(n, dtype=O.DECISION_DTYPE)                       # reused: no allocation in the timed loop
    O.route_batch(table, batches[-1], threads=threads, switch_only=switch_only, fast=True, out=out)      # warm
    reps, t0 = 0, time.perf_counter()
    while True:
        rc, _ = O.route_batch(table, batches[reps % len(batches)], threads=threads, switch_only=switch_only, fast=True, out=out)
        assert rc == 0
        reps += 1
        dt = time.perf_counter() - t0
        if (dt >= budget_s and reps >= min_reps) or reps >= 1000000:
            break
    return n * reps / dt, reps


def cpu_port_note(O, cores):
    return (f"oracle/pk_cpu_fast.c: one fused pass (scan+hash+select), {O.fast_impl()} inner loop, "
            f"persistent pool of {cores} threads, dynamic chunks")


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
    batches = cpu_ring(O, n_step, plen, cores)                   # larger than the last-level cache
    nr = len(batches)
    out = np.zeros(n_step, dtype=O.DECISION_DTYPE)
    # the efficient CPU port (oracle/pk_cpu_fast.c) is what is timed; check it against the plain
    # oracle once, outside the timed region
    assert (O.route_batch(table, batches[0], threads=cores, fast=True)[1] ==
            O.route_batch(table, batches[0], threads=cores)[1]).all()
    for w in range(args.warmup):
        O.route_batch(table, batches[(nr - 1 - w) % nr], threads=cores, fast=True, out=out)
    t0 = time.perf_counter()
    for k in range(args.steps):
        rc, _ = O.route_batch(table, batches[k % nr], threads=cores, fast=True, out=out)
        assert rc == 0
    dt = time.perf_counter() - t0
    value = n_step * args.steps / dt
    v_one, _ = cpu_arm(O, table, batches, 1, 1.5)
    v_hot, _ = cpu_arm(O, table, batches[:1], cores, 1.0)            # one batch over and over: cache-resident
    # CPU-A: the reference-faithful work (the switch only) — report the better of all cores / one
    sw_all, _ = cpu_arm(O, table, batches, cores, 1.0, switch_only=True)
    sw_one, _ = cpu_arm(O, table, batches, 1, 1.0, switch_only=True)
    sw, sw_cores = (sw_all, cores) if sw_all >= sw_one else (sw_one, 1)
    sample = (f"{args.steps} steps x {n_step} requests x {plen} B over a ring of {nr} distinct batches "
              f"({nr * n_step * plen >> 20} MiB > last-level cache {llc_bytes() >> 20} MiB: payloads stream from DRAM) "
              f"({cpu_port_note(O, cores)})")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": DTYPE,
        "data": "synthetic", "config": config_of(desc, n, plen),
        "parallelism": "host threads (persistent pool), requests in dynamic chunks",
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "one_core": v_one, "scaling_efficiency": value / (v_one * cores), "impl": O.fast_impl(),
                         "sample_requests_per_step": n_step, "cache_resident_value": v_hot},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "reference_switch_only": {"value": sw, "unit": UNIT, "cores": sw_cores,
                                  "note": "CPU-A: only the work the Go reference does per request "
                                          "(mock.go:32-64 switch; payload never read)"},
        "note": "Go reference cannot be built on this box (no Go toolchain, un-vendored spounge-proto); "
                "this is the C port doing the work-equivalent scan+hash+select",
    }
    emit(line)


# ------------------------------------------------------------------ GPU legs ------
def resident_leg(ctx: Ctx, workload: str, steps: int, warmup: int, issue: str, all_modes: bool,
                 steady_steps: int, depth: int, prewarm: bool = True):
    """Stages a ring of batches larger than L2, checks slot 0 bit-exactly against the oracle, then
    times `steps` launches issued as `issue` (device time, max over ranks).  Returns the router
    (slots [ring, ring+depth) are left for the e2e leg) and the numbers."""
    pk, O = ctx.pk, ctx.O
    n, plen, desc = WORKLOADS[workload]
    batch_bytes = n * ((plen + 15) & ~15)
    ring = max(2, min(64, -(-(2 * L2_BYTES + batch_bytes) // batch_bytes)))
    if batch_bytes > L2_BYTES:
        ring = 2
    router = pk.Router(device=ctx.local, max_requests=n, max_payload_bytes=batch_bytes, n_slots=ring + depth)
    log(f"[rank {ctx.rank}] workload {workload}: n={n} payload={plen} ring={ring} "
        f"({ring * batch_bytes / 2**20:.0f} MiB resident)")

    # request ids: rank r owns ids r, r+world, ... (req_id % G sharding); batch j of rank r starts at
    # a disjoint id range, so every rank routes different requests
    def make_batch(j, payload_out=None):
        first_id = (ctx.rank * 1_000_003 + j) * n
        return O.gen_batch(O.SEED_BASE + 1, n, plen, mode=0, first_id=first_id, threads=ctx.gen_threads,
                           payload_out=payload_out)

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
    rc, want = O.route_batch(ctx.table, check_batch, threads=ctx.cores)
    assert rc == 0 and (got == want).all(), "GPU records differ from the oracle — refusing to time"

    slots = list(range(ring))
    use_graph, issue_desc = ISSUE[issue]
    if issue == "fused":                                  # (a batch may appear once per launch: no more than the ring holds)
        use_graph = 1 | PDL | FUSE(min(32, ring))
    # bring clocks up, then the W warm-up steps
    t_end = time.perf_counter() + (0.4 if prewarm else 0.0)
    while time.perf_counter() < t_end:
        router.launch_many(slots, 512, use_graph=False)
    router.launch_many(slots, max(3, warmup), use_graph=use_graph)

    # ---- timed: K steps, device-resident ----
    launches_before = router.launch_count
    ctx.barrier()
    w0 = time.perf_counter()
    ms = router.launch_many(slots, steps, use_graph=use_graph)
    ctx.torch.cuda.synchronize(ctx.local)
    w1 = time.perf_counter()
    ctx.barrier()
    ctx.windows.append((w0, w1))
    launches = router.launch_count - launches_before
    ms = ctx.max(ms)
    res = {"n": n, "plen": plen, "desc": desc, "batch_bytes": batch_bytes, "ring": ring, "alg_bytes": alg_bytes,
           "ms": ms, "steps": steps, "launches": int(launches), "issue": issue, "issue_desc": issue_desc,
           "check_batch": check_batch, "make_batch": make_batch, "parity": True}
    # the same issue mode over enough launches that fill and drain of the overlapped pipeline vanish
    if steady_steps:
        router.launch_many(slots, 64, use_graph=use_graph)
        t0 = time.perf_counter()
        ms_ss = ctx.max(router.launch_many(slots, steady_steps, use_graph=use_graph))
        ctx.windows.append((t0, time.perf_counter()))
        res["steady"] = (steady_steps, ms_ss)
    other = {}
    modes = (("pdl4", 4 | PDL), ("pdl", 1 | PDL), ("serial", 1), ("plain", 0), ("plain_pdl", PDL), ("four_chains", 4),
             ("two_chains_pdl", 2 | PDL), ("fused", 1 | PDL | FUSE(min(32, ring)))) if all_modes \
        else (("serial", 1),)
    for name, mode in modes:
        if name != issue:
            other[name] = ctx.max(router.launch_many(slots, steps, use_graph=mode)) / steps
    res["other"] = other
    # parity gate on the timed path itself: one more pass of the SAME issue mode over the whole ring
    # (so slot 0 is written by the kernel that was timed, not by the single launch above), then slot
    # 0's records against the oracle's again
    router.launch_many(slots, ring, use_graph=use_graph)
    assert (router.fetch(0, n) == want).all(), f"records written under issue '{issue}' differ from the oracle"
    return router, res


def e2e_leg(ctx: Ctx, router, res, depth: int, k_e2e: int, n_arenas: int, both_paths: bool, sync_latency: bool):
    """pk_submit / pk_wait on HOST buffers, `depth` batches in flight: every step's payload and
    metadata cross PCIe inside the timed region and its records come back."""
    pk, O = ctx.pk, ctx.O
    n, batch_bytes, ring = res["n"], res["batch_bytes"], res["ring"]
    e2e_batches = []
    for j in range(n_arenas):
        buf = router.host_alloc(batch_bytes + 64)                 # pinned arena owned by the C side
        e2e_batches.append(res["make_batch"](100 + j, payload_out=buf))
    outs = [np.frombuffer(router.host_alloc(24 * n), dtype=pk.DECISION_DTYPE) for _ in range(depth)]
    e2e_slots = [ring + d for d in range(depth)]
    cb = res["check_batch"]
    meta_bytes = int(cb["pay_off"].nbytes + cb["pay_len"].nbytes + cb["name_off"].nbytes + cb["names"].nbytes)
    h2d = int(cb["pay_off"][n - 1]) + int(cb["pay_len"][n - 1]) + meta_bytes
    d2h = 24 * n

    def pipelined(k):
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

    pipelined(max(3, min(10, k_e2e)))                               # warm
    ctx.barrier()
    direct_before = router.direct_count
    e0 = time.perf_counter()
    e2e_s = pipelined(k_e2e)
    e1 = time.perf_counter()
    ctx.barrier()
    ctx.windows.append((e0, e1))
    used_direct = router.direct_count - direct_before == k_e2e      # the library's own choice for this batch size
    e2e_s = ctx.max(e2e_s)
    rc, want = O.route_batch(ctx.table, e2e_batches[(k_e2e - 1) % len(e2e_batches)], threads=ctx.cores)
    assert (outs[(k_e2e - 1) % depth] == want).all(), "e2e records differ from the oracle"
    transfer = {True: "none issued: the kernel reads the page-locked arenas over PCIe and writes the records into the "
                      "caller's page-locked buffer (pk_set_direct_bytes, default for batches up to 16 MiB)",
                False: "cudaMemcpyAsync H2D (payload, metadata) and D2H (records) around the kernel"}
    e2e = {"value": ctx.world * k_e2e * n / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "steps": k_e2e, "pipeline_depth": depth, "ms_per_step": e2e_s / k_e2e * 1e3,
           "host_gbs_per_gpu": (h2d + d2h) / (e2e_s / k_e2e) / 1e9,
           "api": "pk_submit/pk_wait, pinned host arenas", "copies": transfer[used_direct]}
    if both_paths:
        # the other transfer path, for the record (same call, same buffers)
        router.set_direct_bytes(0 if used_direct else 1 << 40)
        k_alt = max(10, k_e2e // 4)
        pipelined(3)
        ctx.barrier()
        alt_s = ctx.max(pipelined(k_alt)) / k_alt
        router.set_direct_bytes(16 << 20)
        e2e["other_path"] = {"copies": transfer[not used_direct], "ms_per_step": alt_s * 1e3,
                             "value": ctx.world * n / alt_s, "unit": UNIT}
    if sync_latency:
        # synchronous single-call latency (what one flush costs a waiting request)
        lat = []
        for s in range(min(k_e2e, 300)):
            t0 = time.perf_counter()
            router.route_batch(e2e_batches[s % len(e2e_batches)], out=outs[0])
            lat.append((time.perf_counter() - t0) * 1e6)
        lat.sort()
        e2e["sync_call_latency_us"] = {"p50": lat[len(lat) // 2], "p99": lat[min(len(lat) - 1, int(len(lat) * 0.99))],
                                       "api": "pk_route_batch, one batch per call"}
    return e2e


def pcie_probe(ctx: Ctx, router, nbytes: int, reps: int = 40):
    """The host ceiling an end-to-end step runs under: all ranks copy `nbytes` of page-locked host
    memory to their GPU at the same time with plain cudaMemcpyAsync (and, separately, both
    directions at once).  Device time, max over ranks."""
    torch = ctx.torch
    dev = torch.device("cuda", ctx.local)
    h_in = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    h_out = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    d_in = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    d_out = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
    h_in.fill_(7)
    s2 = torch.cuda.Stream(device=dev)

    def timed(duplex):
        for _ in range(3):
            d_in.copy_(h_in, non_blocking=True)
        ctx.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        s2.wait_stream(torch.cuda.current_stream(dev))
        for _ in range(reps):
            d_in.copy_(h_in, non_blocking=True)
            if duplex:
                with torch.cuda.stream(s2):
                    h_out.copy_(d_out, non_blocking=True)
        torch.cuda.current_stream(dev).wait_stream(s2)
        b.record()
        torch.cuda.synchronize(dev)
        return ctx.max(a.elapsed_time(b)) / 1e3

    t_h2d = timed(False)
    t_dup = timed(True)
    per_rank = nbytes * reps / t_h2d / 1e9
    # the same bytes pulled by a plain read kernel (128-bit loads of mapped host memory): what the
    # in-place path of pk_submit runs under — on these boxes faster than one copy engine
    arena = router.host_alloc((nbytes + 15) & ~15)
    arena[:] = 7
    router.probe_host_read(arena, 3)
    ctx.barrier()
    t0 = time.perf_counter()
    kern = router.probe_host_read(arena, reps)
    kern_min = -ctx.max(-kern)                               # slowest rank
    ctx.windows.append((t0, time.perf_counter()))
    return {"memcpy_h2d_gbs_per_gpu": per_rank, "kernel_read_gbs_per_gpu": kern_min,
            "peak_gbs_per_gpu": max(per_rank, kern_min), "peak_gbs_aggregate": max(per_rank, kern_min) * ctx.world,
            "memcpy_duplex_gbs_per_gpu_each_way": nbytes * reps / t_dup / 1e9,
            "what": f"{ctx.world} rank(s) at once, {reps} passes over {nbytes} B of page-locked host memory: "
                    "cudaMemcpyAsync H2D (one copy engine), the same with a D2H copy of equal size on a second stream, "
                    "and a plain read kernel (pk_probe_host_read); device time, slowest rank"}


def c4_leg(ctx: Ctx, batch: int, steps: int, warmup: int, e2e_reps: int = 3, prewarm: bool = True):
    """BASELINE configs[4]: ONE stream of 1,000,000 requests x 2 KB sharded by request id over the
    G ranks (rank r owns r, r+G, ...; no exchange), each rank working through its shard in batches
    of `batch` requests.  Strong scaling: the work is fixed, value = 1e6 / the slowest rank's time
    for one pass over its shard."""
    pk, O = ctx.pk, ctx.O
    n, plen = batch, 2048
    mine = len(range(ctx.rank, C4_TOTAL, ctx.world))           # requests of this rank's shard
    nb = -(-mine // n)                                         # its batches; the last one is ragged
    batch_bytes = n * plen
    depth = 3
    router = pk.Router(device=ctx.local, max_requests=n, max_payload_bytes=batch_bytes, n_slots=nb + depth)
    log(f"[rank {ctx.rank}] C4: shard of {mine} requests = {nb} batches of {n} ({nb * batch_bytes / 2**20:.0f} MiB resident)")

    def shard_batch(j, payload_out=None):
        # rank-local numbering: the shard's k-th request is generated from id rank * 2^32 + k, so
        # shards are disjoint and the same whatever the batch size
        cnt = min(n, mine - j * n)
        return O.gen_batch(O.SEED_BASE + 4, cnt, plen, mode=0, first_id=(ctx.rank << 32) + j * n,
                           threads=ctx.gen_threads, payload_out=payload_out)

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
        assert (got == O.route_batch(ctx.table, b, threads=ctx.cores)[1]).all(), "GPU records differ from the oracle"
    alg_pass = algorithmic_bytes(first) * (mine / n)         # per pass over the shard

    slots = list(range(nb))
    reps = max(3, steps // nb)
    t_end = time.perf_counter() + (0.4 if prewarm else 0.0)
    while time.perf_counter() < t_end:
        router.launch_many(slots, nb, use_graph=False)
    router.launch_many(slots, nb * max(1, warmup // nb + 1), use_graph=4 | PDL)
    launches_before = router.launch_count
    ctx.barrier()
    w0 = time.perf_counter()
    ms = router.launch_many(slots, nb * reps, use_graph=4 | PDL)
    ctx.torch.cuda.synchronize(ctx.local)
    w1 = time.perf_counter()
    ctx.barrier()
    ctx.windows.append((w0, w1))
    launches = router.launch_count - launches_before
    pass_s = ctx.max(ms) / 1e3 / reps                          # slowest rank's time for ONE pass over its shard
    ms_serial = ctx.max(router.launch_many(slots, nb * reps, use_graph=1)) / reps

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
    ctx.barrier()
    e0 = time.perf_counter()
    e2e_s = sum(e2e_pass() for _ in range(e2e_reps)) / e2e_reps
    e1 = time.perf_counter()
    ctx.barrier()
    ctx.windows.append((e0, e1))
    e2e_s = ctx.max(e2e_s)
    meta = int(first["pay_off"].nbytes + first["pay_len"].nbytes + first["name_off"].nbytes + first["names"].nbytes)
    router.close()
    return {"first": first, "n": n, "plen": plen, "mine": mine, "nb": nb, "reps": reps, "batch_bytes": batch_bytes,
            "pass_s": pass_s, "serial_pass_ms": ms_serial, "launches": int(launches), "alg_pass": alg_pass,
            "e2e_s": e2e_s, "e2e_reps": e2e_reps, "meta": meta, "depth": depth}


def c4_record(ctx: Ctx, c):
    """The compact `c4` sub-record of the default line."""
    return {"workload": f"C4: one stream of {C4_TOTAL} requests x {c['plen']} B sharded by request id over {ctx.world} GPU(s), "
                        f"per-GPU batches of {c['n']} (BASELINE configs[4])",
            "scaling": "strong", "value": C4_TOTAL / c["pass_s"], "unit": UNIT, "pass_ms": c["pass_s"] * 1e3,
            "batches_per_shard": c["nb"], "shard_requests_rank0": c["mine"],
            "frac": c["alg_pass"] / c["pass_s"] / 1e9 / ctx.peak,
            "serial_pass_ms": c["serial_pass_ms"],
            "e2e": {"value": C4_TOTAL / c["e2e_s"], "unit": UNIT, "pass_ms": c["e2e_s"] * 1e3,
                    "h2d_bytes_per_step": c["batch_bytes"] + c["meta"], "d2h_bytes_per_step": 24 * c["n"],
                    "pipeline_depth": c["depth"], "api": "pk_submit/pk_wait, pinned host arenas"},
            "gpu_launches": c["launches"], "parity_gate": "first and last (ragged) batch of the shard bit-exact vs the oracle"}


def frontend_leg(ctx: Ctx, seconds: float = 1.5):
    """Through the service.Service seam: the C++ batching front-end (host/batching_service.cpp),
    driven by host/latency_bench.cpp in ONE process with one backend (flusher, table replica) per
    GPU — the shape of the reference's server (cmd/polykey/main.go:85-89).  Rank 0 runs it while the
    other ranks wait; open loop, Poisson arrivals of 2 KB prompts that start on the callers' heap
    (each producer copies its request into the pinned arena inside the measured time)."""
    if ctx.rank != 0:
        return None
    exe = os.path.join(ROOT, "build", "latency_bench")
    try:
        from polykey_service_b200 import build as pkbuild
        exe = pkbuild.build_latency_bench()
    except Exception as e:                                   # no compiler on this box: say so, invent nothing
        if not os.path.exists(exe):
            return {"unavailable": f"host/latency_bench could not be built: {e!r}"}
    G = ctx.world
    env = dict(os.environ, PK_BACKENDS=str(G))
    gens = max(2, min(ctx.cores - G - 3, 24))              # leave a core per flusher, and some for this process
    out = {"gpus": G, "generators": gens, "payload_bytes": 2048, "max_batch": 4096, "max_wait_us": 200,
           "api": "BatchingService::EnqueueAsync -> decision (C++ mirror of the Go batcher), open loop, Poisson arrivals",
           "rates": []}
    # offered rates per GPU: light load (latency = service time), moderate, heavy, very heavy, and flat out
    # (more than one process of producers can deliver: its completed rate is the front-end's ceiling)
    for per_gpu in (1e5, 1e6, 5e6, 1e7, 2e7):
        rate = per_gpu * G
        try:
            p = subprocess.run([exe, "open", str(int(rate)), str(seconds), "4096", "200", "2048", str(gens)],
                               capture_output=True, text=True, timeout=120, env=env)
            r = json.loads(p.stdout.strip().splitlines()[-1])
        except Exception as e:
            out["rates"].append({"offered_per_s": rate, "error": repr(e)[:200]})
            continue
        out["rates"].append({"offered_per_s": rate, "completed_per_s": r["completed_per_s"], "dropped": r["dropped"],
                             "errors": r["errors"], "p50_us": r["latency_us"]["p50"], "p99_us": r["latency_us"]["p99"],
                             "mean_batch": r["mean_batch"]})
    # the same flat out with TWO shards (handle, flusher, open batch) per GPU: one open batch means one
    # reservation counter every producer of the GPU meets on, and that counter is the single-shard ceiling
    try:
        p = subprocess.run([exe, "open", str(int(2e7 * G)), str(seconds), "4096", "200", "2048", str(gens)],
                           capture_output=True, text=True, timeout=120, env=dict(env, PK_SHARDS_PER_GPU="2"))
        r = json.loads(p.stdout.strip().splitlines()[-1])
        out["two_shards_per_gpu"] = {"offered_per_s": 2e7 * G, "completed_per_s": r["completed_per_s"], "mean_batch": r["mean_batch"]}
    except Exception as e:
        out["two_shards_per_gpu"] = {"error": repr(e)[:200]}
    ok = [r for r in out["rates"] if "error" not in r]
    if ok:
        light = ok[0]
        out.update({"p50_us": light["p50_us"], "p99_us": light["p99_us"], "rate": light["offered_per_s"]})
        out["sustained_per_s"] = max(r["completed_per_s"] for r in ok)
    return out


def gcm_leg(ctx: Ctx):
    """SURVEY §8(f) rank 1 — batched AES-256-GCM (internal/adapters/security/cipher.go:110-141): two
    batches sealed on the GPU through its C ABI (include/pk_gcm.h) from page-locked arenas, checked
    byte for byte against OpenSSL, timed beside OpenSSL on the host cores with the key set once per
    thread (the reference's own loop rebuilds the cipher per message: cpu_reference_shape_ms)."""
    import ctypes as C
    from polykey_service_b200 import gcm
    ref = C.CDLL(os.path.join(ROOT, "oracle", "libpkaesref.so"))
    ref.pk_aes_ref_seal_batch.argtypes = [C.c_char_p] + [C.c_void_p] * 4 + [C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int]
    rng = np.random.default_rng(7)
    key = rng.bytes(32)
    rows = {}

    def med(f, reps):
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            f()
            ts.append(time.perf_counter() - t0)
        return sorted(ts)[len(ts) // 2] * 1e3

    for name, n, length in (("65536x256B", 65536, 256), ("32768x4KiB", 32768, 4096)):
        plen = np.full(n, length, dtype=np.uint32)
        poff, ptotal = gcm.slots(plen)
        soff, stotal = gcm.slots(plen.astype(np.uint64) + gcm.OVERHEAD)
        with gcm.Gcm(key, device=ctx.local, max_messages=n, max_bytes=stotal + 64) as g:
            arrs = {"plain": rng.integers(0, 256, ptotal, dtype=np.uint8), "poff": poff, "plen": plen,
                    "nonces": rng.integers(0, 256, 12 * n, dtype=np.uint8), "sealed": np.zeros(stotal, dtype=np.uint8), "soff": soff}
            pin = {}
            for k, a in arrs.items():                          # page-locked, as the host mirror packs them
                pin[k] = g.pinned(a.nbytes, a.dtype)[:a.size]
                pin[k][:] = a
            call = lambda: g.seal_arena(pin["plain"], pin["poff"], pin["plen"], pin["nonces"], pin["sealed"], pin["soff"])  # noqa: E731
            call()
            want = np.zeros(stotal, dtype=np.uint8)
            cpu = lambda threads, per_item: ref.pk_aes_ref_seal_batch(  # noqa: E731
                key, arrs["plain"].ctypes.data, poff.ctypes.data, plen.ctypes.data, arrs["nonces"].ctypes.data, n,
                want.ctypes.data, soff.ctypes.data, threads, per_item)
            cpu(ctx.cores, 0)
            sealed = np.asarray(pin["sealed"])
            ok = all((sealed[int(o):int(o) + length + gcm.OVERHEAD] == want[int(o):int(o) + length + gcm.OVERHEAD]).all()
                     for o in soff[:: max(1, n // 512)]) and (sealed[int(soff[-1]):] == want[int(soff[-1]):]).all()
            assert ok, f"AES-GCM: GPU and OpenSSL differ ({name})"
            for _ in range(5):
                call()
            t0 = time.perf_counter()
            call_ms = med(call, 15)
            ctx.windows.append((t0, time.perf_counter()))
            g.set_chunks(1)                                    # kernels alone: no copy overlapping them
            ks = []
            for _ in range(7):
                call()
                ks.append(g.last_kernel_ms)
            rows[name] = {"messages": n, "bytes": length, "kernel_ms": sorted(ks)[len(ks) // 2], "call_ms": call_ms,
                          "openssl_key_once_ms": med(lambda: cpu(ctx.cores, 0), 5),
                          "cpu_reference_shape_ms": med(lambda: cpu(ctx.cores, 1), 3), "cores": ctx.cores,
                          "kernel_gbs": n * length / (sorted(ks)[len(ks) // 2] / 1e3) / 1e9,
                          "parity_gate": "sealed bytes equal OpenSSL's (every 128th message and the last, all bytes)"}
    rows["what"] = ("pk_gcm_seal_batch, page-locked arenas; kernel_ms = the kernels alone (CUDA events); call_ms = one call, "
                    "copies included; OpenSSL EVP on all host cores, key set once per thread / fresh cipher per message")
    return rows


def run_gpu(args):
    ctx = Ctx(args)
    O, pk = ctx.O, ctx.pk
    workload = args.workload
    depth = args.depth
    issue = "plain" if args.no_graph else args.issue
    sampler = ClockSampler(ctx.local)
    sampler.start()

    steady = 0 if args.steps >= 2000 or args.lean else 4000
    router, res = resident_leg(ctx, workload, args.steps, args.warmup, issue, all_modes=True, steady_steps=steady,
                               depth=depth, prewarm=not args.no_prewarm)
    n, plen, desc, alg_bytes = res["n"], res["plen"], res["desc"], res["alg_bytes"]
    ms = res["ms"]
    h2d_guess = res["batch_bytes"]
    k_e2e = max(20, min(max(args.steps, 200), int(3.0 / max(1e-5, (h2d_guess / 20e9)))))   # <= ~3 s at a pessimistic 20 GB/s
    e2e = e2e_leg(ctx, router, res, depth, k_e2e, n_arenas=depth + 1 if res["batch_bytes"] <= (64 << 20) else 2,
                  both_paths=True, sync_latency=True)
    if not args.lean:
        probe = pcie_probe(ctx, router, e2e["h2d_bytes_per_step"])
        e2e["pcie_probe"] = probe
        e2e["pcie_peak_gbs"] = probe["peak_gbs_per_gpu"]
        e2e["frac_of_pcie"] = (e2e["h2d_bytes_per_step"] / (e2e["ms_per_step"] / 1e3) / 1e9) / probe["peak_gbs_per_gpu"]
    router.close()

    per_launch_s = ms / 1e3 / args.steps                 # per step; a launch covers batches_per_launch steps (1 unless fused)
    batches_per_launch = args.steps / max(1, res["launches"])
    value = ctx.world * args.steps * n / (ms / 1e3)
    ms_serial = res["other"].get("serial", ms / args.steps)

    # ---- the other BASELINE configurations, compact ----
    workloads, c4 = {}, None
    if not args.lean:
        if ctx.world == 1 and workload == "c1":
            for w, k, k_e in (("c2", 2000, 60), ("c3", 60, 8)):
                r2, s2 = resident_leg(ctx, w, k, 10, issue, all_modes=False, steady_steps=0, depth=2)
                e2 = e2e_leg(ctx, r2, s2, 2, k_e, n_arenas=2, both_paths=False, sync_latency=False)
                r2.close()
                workloads[w] = {"workload": s2["desc"], "steps": k, "ms_per_step": s2["ms"] / k,
                                "value": s2["n"] * k / (s2["ms"] / 1e3), "unit": UNIT,
                                "frac": s2["alg_bytes"] / (s2["ms"] / 1e3 / k) / 1e9 / ctx.peak,
                                "algorithmic_bytes_per_launch": s2["alg_bytes"],
                                "serial_ms_per_step": s2["other"].get("serial"),
                                "e2e": {k_: e2[k_] for k_ in ("value", "unit", "ms_per_step", "steps", "h2d_bytes_per_step",
                                                              "d2h_bytes_per_step", "pipeline_depth", "copies")},
                                "parity_gate": "slot 0 (single launch before timing, and again as written by the timed issue "
                                               "mode) and the last e2e batch bit-exact vs the oracle"}
        c4 = c4_record(ctx, c4_leg(ctx, args.c4_batch, 245 * 4, 5))
    gcm_rec = None
    if not args.lean and ctx.world == 1 and workload == "c1":
        try:
            gcm_rec = gcm_leg(ctx)
        except (OSError, ImportError) as e:                  # no libcrypto / cryptography on this box: say so
            gcm_rec = {"unavailable": repr(e)[:200]}
    latency = None
    if not (args.lean or args.no_frontend):
        # rank 0 drives every GPU from one server-shaped process; the other ranks wait on the
        # rendezvous store — on the CPU, not in an NCCL barrier, whose kernel would spin on their GPUs
        ctx.barrier()
        store = None
        if ctx.dist is not None:
            from torch.distributed.distributed_c10d import _get_default_store
            store = _get_default_store()
        if ctx.rank == 0:
            latency = frontend_leg(ctx)
            if store is not None:
                store.set("pk_frontend_done", "1")
        elif store is not None:
            import datetime
            store.wait(["pk_frontend_done"], datetime.timedelta(seconds=900))
    ctx.barrier()

    sampler.stop_flag = True
    sampler.join(timeout=2)
    clocks = sampler.summary(ctx.windows)
    if ctx.rank != 0:
        return

    # ---- CPU baseline on this box's host cores (rank 0, N=1 only) ----
    cpu_baseline, extra_cpu = None, {}
    if ctx.world == 1 and not args.no_cpu:
        n_s = n if res["batch_bytes"] <= (64 << 20) else max(1, (64 << 20) // plen)
        ring_cpu = cpu_ring(O, n_s, plen, ctx.cores)
        v_all, reps = cpu_arm(O, ctx.table, ring_cpu, ctx.cores, 6.0)
        v_one, _ = cpu_arm(O, ctx.table, ring_cpu, 1, 3.0)
        v_hot, _ = cpu_arm(O, ctx.table, ring_cpu[:1], ctx.cores, 1.0)
        v_sw, _ = cpu_arm(O, ctx.table, ring_cpu, ctx.cores, 1.0, switch_only=True)
        v_sw1, _ = cpu_arm(O, ctx.table, ring_cpu, 1, 1.0, switch_only=True)
        cpu_baseline = {"value": v_all, "unit": UNIT, "cores": ctx.cores, "kind": "port",
                        "sample": f"{reps} steps of {n_s} requests x {plen} B (about 6 s) over a ring of {len(ring_cpu)} distinct "
                                  f"batches ({len(ring_cpu) * n_s * plen >> 20} MiB > last-level cache {llc_bytes() >> 20} MiB: "
                                  f"payloads stream from DRAM), {cpu_port_note(O, ctx.cores)}",
                        "one_core": v_one, "scaling_efficiency": v_all / (v_one * ctx.cores), "impl": O.fast_impl(),
                        "cache_resident_value": v_hot}
        extra_cpu = {"cpu_baseline_1core": {"value": v_one, "unit": UNIT, "cores": 1},
                     "cpu_switch_only": {"value": v_sw, "unit": UNIT, "cores": ctx.cores,
                                         "note": "CPU-A: what the Go reference actually does per request "
                                                 "(mock.go:32-64 switch; never reads the payload)"},
                     "cpu_switch_only_1core": {"value": v_sw1, "unit": UNIT, "cores": 1}}

    achieved = alg_bytes / per_launch_s / 1e9
    traffic, traffic_src = ncu_traffic(workload)       # per batch on file; a launch covers batches_per_launch of them
    if traffic is not None:
        traffic = int(round(traffic * batches_per_launch))
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": ctx.world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
        "config": config_of(desc, n, plen),
        "parallelism": f"dp{ctx.world}: requests sharded by id, no collective",
        "l2": f"ring of {res['ring']} resident batches = {res['ring'] * res['batch_bytes'] / 2**20:.0f} MiB > 126 MiB L2; launch k "
              "reads the batch after the one launch k-1 read, and the walk carries on across calls (warm-up launches "
              "never pre-load the timed ones)",
        "issue": res["issue_desc"] + "; CUDA events on the stream the graph is launched on",
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": ctx.peak, "unit": "GB/s",
                     "frac": achieved / ctx.peak, "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": ctx.peak_src,
                     "kernel": ("route_fused_kernel" if batches_per_launch > 1 else "route_kernel") + " (fused scan+hash+select)",
                     "algorithmic_bytes_per_launch": alg_bytes * batches_per_launch,
                     "avg_launch_us": per_launch_s * batches_per_launch * 1e6, "batches_per_launch": batches_per_launch},
        "cpu_baseline": cpu_baseline,
        "e2e": e2e,
        "gpu_launches": res["launches"],
        "parity_gate": "before timing: slot 0 routed by a single launch, bit-exact vs the oracle; after timing: slot 0 as "
                       "written by the timed issue mode, again; the last e2e batch likewise",
        "clocks": clocks,
        "issue_modes_ms_per_step": dict(res["other"], **{issue: ms / args.steps}),
        "serial": {"ms_per_step": ms_serial, "value": ctx.world * n / (ms_serial / 1e3), "unit": UNIT,
                   "frac": alg_bytes / (ms_serial / 1e3) / 1e9 / ctx.peak,
                   "note": "the same K launches with every kernel waiting for its predecessor to complete: what ONE "
                           "batch costs on an idle GPU (launch + two dependent HBM round trips + drain)"},
    }
    if "steady" in res or args.steps >= 2000:
        k_ss, ms_ss = res.get("steady", (args.steps, ms))
        line["steady_state"] = {"steps": k_ss, "ms_per_step": ms_ss / k_ss, "value": ctx.world * k_ss * n / (ms_ss / 1e3),
                                "unit": UNIT, "frac": alg_bytes / (ms_ss / 1e3 / k_ss) / 1e9 / ctx.peak,
                                "note": "same issue mode over enough steps that the ramp and the tail of the first and "
                                        "last launch are amortised (fused: 32 batches per launch, consecutive launches "
                                        "overlapping through programmatic dependent launch)"}
    if workloads:
        line["workloads"] = workloads
    if c4:
        line["c4"] = c4
    if gcm_rec:
        line["gcm"] = gcm_rec
    if latency:
        line["latency"] = latency
        if "sustained_per_s" in latency:
            e2e["with_pack"] = {"value": latency["sustained_per_s"], "unit": UNIT,
                                "what": "open-loop arrivals through the batching front-end at the service.Service seam: "
                                        "per-request heap buffers -> producers copy into the pinned arena -> pk_submit/pk_wait "
                                        "-> decision read by the caller (see `latency`)"}
    line.update(extra_cpu)
    emit(line)


def make_sample(O, plen):
    n_s = max(1, (64 << 20) // plen)
    return O.gen_batch(O.SEED_BASE + 1, n_s, plen, mode=0, threads=16)


def run_gpu_c4(args):
    """`--workload c4`: BASELINE configs[4] as the headline of its own line (strong scaling)."""
    ctx = Ctx(args)
    O = ctx.O
    sampler = ClockSampler(ctx.local)
    sampler.start()
    c = c4_leg(ctx, args.c4_batch, args.steps, args.warmup, prewarm=not args.no_prewarm)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    clocks = sampler.summary(ctx.windows)
    if ctx.rank != 0:
        return
    n, plen, nb, reps, pass_s, first = c["n"], c["plen"], c["nb"], c["reps"], c["pass_s"], c["first"]
    achieved = c["alg_pass"] / pass_s / 1e9
    cpu_baseline = None
    if ctx.world == 1 and not args.no_cpu:
        v_all, reps_cpu = cpu_arm(O, ctx.table, first, ctx.cores, 6.0)
        cpu_baseline = {"value": v_all, "unit": UNIT, "cores": ctx.cores, "kind": "port",
                        "sample": f"{reps_cpu} passes over one batch of {n} requests x {plen} B, {cpu_port_note(O, ctx.cores)}"}
    emit({
        "metric": METRIC, "value": C4_TOTAL / pass_s, "unit": UNIT, "n_gpus": ctx.world, "steps": nb * reps,
        "warmup": max(3, args.warmup), "ms_per_step": pass_s * 1e3 / nb, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
        "config": {"workload": f"C4: one stream of {C4_TOTAL} requests x {plen} B sharded by request id over {ctx.world} GPU(s), "
                               f"per-GPU batches of {n} (BASELINE configs[4])",
                   "shard_requests_rank0": c["mine"], "batches_per_shard": nb, "batch_requests": n, "payload_bytes": plen,
                   "providers": 8},
        "parallelism": f"dp{ctx.world}: requests sharded by id, no collective",
        "l2": f"the shard is resident ({nb * c['batch_bytes'] / 2**20:.0f} MiB on rank 0) and each pass walks all of it",
        "issue": ISSUE["pdl4"][1] + "; CUDA events on the stream the graph is launched on",
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": ctx.peak, "unit": "GB/s", "frac": achieved / ctx.peak,
                     "traffic": None, "peak_source": ctx.peak_src, "kernel": "route_kernel (fused scan+hash+select)",
                     "algorithmic_bytes_per_launch": algorithmic_bytes(first), "avg_launch_us": pass_s * 1e6 / nb},
        "cpu_baseline": cpu_baseline,
        "e2e": {"value": C4_TOTAL / c["e2e_s"], "unit": UNIT, "h2d_bytes_per_step": c["batch_bytes"] + c["meta"],
                "d2h_bytes_per_step": 24 * n, "steps": nb * c["e2e_reps"], "pipeline_depth": c["depth"],
                "ms_per_step": c["e2e_s"] * 1e3 / nb, "api": "pk_submit/pk_wait, pinned host arenas", "pass_ms": c["e2e_s"] * 1e3},
        "gpu_launches": c["launches"], "clocks": clocks,
        "pass_ms": pass_s * 1e3, "serial": {"pass_ms": c["serial_pass_ms"], "value": C4_TOTAL / (c["serial_pass_ms"] / 1e3), "unit": UNIT},
    })


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c1", choices=sorted(WORKLOADS) + ["c4"])
    ap.add_argument("--c4-batch", type=int, default=4096, help="requests per batch of the C4 stream (4096 / 16384 / 65536)")
    ap.add_argument("--issue", default="fused", choices=["fused", "pdl4", "pdl", "serial", "plain"],
                    help="how the K timed steps are issued (default fused: up to 32 staged batches per kernel launch, the "
                         "launches in a graph chain with programmatic dependent launch; pdl4: one launch per batch, a graph of "
                         "four parallel chains; pdl: one such chain)")
    ap.add_argument("--depth", type=int, default=3, help="e2e pipeline depth: batches in flight through pk_submit/pk_wait")
    ap.add_argument("--no-graph", action="store_true", help="same as --issue plain")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-frontend", action="store_true", help="skip the latency / with_pack leg (C++ front-end)")
    ap.add_argument("--lean", action="store_true",
                    help="headline workload only: no steady_state / workloads / c4 / latency / PCIe probe (for ncu launch lists)")
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
