#!/bin/bash
# C2 (65,536 x 256 B) under the bench's PDL issue: requests per CTA x resident-CTA bound.
set -u
run() { python bench.py --workload c2 --steps 2000 --warmup 10 --no-cpu --lean 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1', 'pdl %.3f us frac %.3f serial %.3f us'%(d['ms_per_step']*1e3, d['roofline']['frac'], d['serial']['ms_per_step']*1e3))"; }
for r in 32 64 128 256; do PK_REQS_PER_CTA=$r run "default R=$r"; done
for v in tiny6 tiny5 tiny4; do for r in 64 128 256; do POLYKEY_PKROUTE_LIB=$PWD/build/pk_$v.so PK_REQS_PER_CTA=$r run "$v R=$r"; done; done
PK_TINY_SHAPE=8 PK_REQS_PER_CTA=64 run "8x2 R=64"
PK_TINY_SHAPE=8 PK_REQS_PER_CTA=128 run "8x2 R=128"
# (the 16 lanes x 1 block shape of round 1 went with SPEC v3: a stripe would span two lanes)
