#!/usr/bin/env python
"""Warp instructions of route_kernel by code region, from the source page of an .ncu-rep captured
with --import-source on:  python tools/ncu_groups.py prof.ncu-rep <requests per launch> <smsp__inst_executed.sum>
(regions are found by marker comments, so the script follows the sources as they move)."""
import csv
import io
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def marker_lines(path, markers):
    lines = open(path).read().split("\n")
    out = {}
    for name, text in markers.items():
        out[name] = next(i + 1 for i, l in enumerate(lines) if text in l)
    return out, len(lines)


def main(rep, nreq, real):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source=cuda,sass"],
                         capture_output=True, text=True).stdout
    cur, hdr, agg = None, None, {}
    for r in csv.reader(io.StringIO(out)):
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if r and r[0] == "Line No":
            hdr = r
            continue
        if hdr and len(r) == len(hdr) and r[0] != "":
            try:
                inst = int(r[hdr.index("Instructions Executed")])
            except ValueError:
                continue
            agg[(cur, int(r[0]))] = agg.get((cur, int(r[0])), 0) + inst
    tot = sum(agg.values())
    sp, _ = marker_lines(os.path.join(ROOT, "polykey_service_b200/csrc/pk_spec.cuh"), {
        "fmix": "PK_HD uint64_t fmix64", "classifier": "struct WordScan", "stripe": "struct Stripe { uint32_t a[4]; };",
        "name": "PK_HD uint64_t name_term", "table": "struct alignas(16) DevTable", "decide": "PK_HD void decide(",
        "final": "const uint64_t h = route_hash_final("})
    kn, _ = marker_lines(os.path.join(ROOT, "polykey_service_b200/csrc/pk_kernels.cu"), {
        "scan_tiny": "__device__ __forceinline__ void scan_tiny(", "body": "__device__ __forceinline__ void route_body(",
        "phase1": "---- phase 1:", "staging": "---- staging for phase 2", "phase2": "---- phase 2:", "wrapper": "route_kernel(const RouteParams P"})
    groups = [
        ("classifier (SPEC B)", [("pk_spec.cuh", 41, sp["fmix"] - 1), ("pk_spec.cuh", sp["classifier"], sp["stripe"] - 1)]),
        ("hash rounds + digests (SPEC C)", [("pk_spec.cuh", 30, 40), ("pk_spec.cuh", sp["fmix"], sp["classifier"] - 1), ("pk_spec.cuh", sp["stripe"], sp["name"] - 1)]),
        ("name terms", [("pk_spec.cuh", sp["name"], sp["table"] - 1)]),
        ("decide: names + tool switch", [("pk_spec.cuh", sp["decide"], sp["final"] - 1)]),
        ("decide: final hash + providers + store", [("pk_spec.cuh", sp["final"], 9999), ("pk_spec.cuh", sp["table"], sp["decide"] - 1)]),
        ("kernel: prologue", [("pk_kernels.cu", kn["body"], kn["phase1"] - 1)]),
        ("kernel: scan loop", [("pk_kernels.cu", kn["phase1"], kn["staging"] - 1), ("pk_kernels.cu", kn["scan_tiny"], kn["scan_tiny"] + 28), ("pk_kernels.cu", 55, 75)]),
        ("kernel: names staging", [("pk_kernels.cu", kn["staging"], kn["phase2"] - 1)]),
        ("kernel: phase 2 (partials)", [("pk_kernels.cu", kn["phase2"], kn["wrapper"] - 12)]),
    ]
    used = set()
    print(f"{rep}: {real:.0f} warp instructions per launch = {real / nreq:.1f} per request")
    for name, rs in groups:
        v = 0
        for (f, a, b) in rs:
            for (ff, l), x in agg.items():
                if ff == f and a <= l <= b and (ff, l) not in used:
                    v += x
                    used.add((ff, l))
        print("  %-42s %5.1f %%  %5.1f / request" % (name, 100 * v / tot, v / tot * real / nreq))
    rest = sum(x for k, x in agg.items() if k not in used)
    print("  %-42s %5.1f %%  %5.1f / request" % ("other", 100 * rest / tot, rest / tot * real / nreq))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]), float(sys.argv[3]))
