#!/usr/bin/env python
"""Print gpurun_out/latency.jsonl and latency_open.jsonl (tools/gpu_latency.sh) as tables."""
import json
import sys

d = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
for l in open(f"{d}/latency.jsonl"):
    r = json.loads(l)
    print("closed %5d callers  B=%-5d eager=%d  %9.3g req/s  p50 %7.1f  p90 %7.1f  p99 %7.1f  batch %6.1f" % (
        r["callers"], r["payload_bytes"], r.get("eager", -1), r["requests_per_s"], r["latency_us"]["p50"],
        r["latency_us"]["p90"], r["latency_us"]["p99"], r["mean_batch"]))
for l in open(f"{d}/latency_open.jsonl"):
    r = json.loads(l)
    print("open %9.0f/s  gens %d  B=%-5d eager=%d wait %d idle %d  done %9.3g/s drop %d  p50 %8.1f  p90 %8.1f  p99 %8.1f  p999 %8.1f  batch %6.1f" % (
        r["offered_per_s"], r["generators"], r["payload_bytes"], r.get("eager", -1), r["max_wait_us"], r["idle_flush_us"],
        r["completed_per_s"], r["dropped"], r["latency_us"]["p50"], r["latency_us"]["p90"], r["latency_us"]["p99"],
        r["latency_us"]["p999"], r["mean_batch"]))
