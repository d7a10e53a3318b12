#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
# one sanitizer tool, on a small subset (edge lengths, golden vectors, long requests, no-alloc entry)
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 99 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q \
   -k "golden or length_edges or long_requests or empty_batch or noalloc_entry_matches" > gpurun_out/sanitizer_memcheck.log 2>&1; echo "memcheck rc=$?"
grep -E "ERROR SUMMARY|passed|failed|Invalid|out of bounds" gpurun_out/sanitizer_memcheck.log | tail -8
