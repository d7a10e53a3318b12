#!/bin/bash
set -u
mkdir -p gpurun_out
echo "== alone, fresh process"
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -15
echo "== whole gpu suite (no-alloc file collected last)"
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
