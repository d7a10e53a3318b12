#!/bin/bash
# Requests per CTA (PK_REQS_PER_CTA) x finalisation mode (PK_LATENCY_MODE) under every issue mode
# of bench.py; prints µs per step.  Tuning aid for launch_route's choice of R.
for w in ${WORKLOADS:-c1 c2}; do
for R in ${RS:-0 8 16 32 64 128 256}; do for M in ${MODES:-x}; do
  export PK_REQS_PER_CTA=$R; if [ "$M" = "x" ]; then unset PK_LATENCY_MODE; else export PK_LATENCY_MODE=$M; fi
  [ "$R" = "0" ] && unset PK_REQS_PER_CTA
  python bench.py --workload $w --no-cpu --steps 2000 2>/dev/null | python -c "
import json,sys; d=json.load(sys.stdin); m=d['issue_modes_ms_per_step']; print('$w R=$R mode=$M', ' '.join('%s=%.2f'%(k,v*1e3) for k,v in sorted(m.items())))"
done; done; done
