#!/bin/bash
# A/B of library variants on the GPU box: tools/ab.sh <bench args> -- lib1.so lib2.so ...   (kernel ms of each)
args=()
while [ "$1" != "--" ]; do args+=("$1"); shift; done; shift
for lib in "$@"; do
  OHCU_LIB=$(realpath $lib) python bench.py "${args[@]}" --no-cpu --no-chain 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); k=d['kernels']
print('$lib', 'frames/s', d['frames_per_s'], 'e2e', d['e2e']['frames_per_s'], 'mc', k['mc']['ms'], 'res', k['residual']['ms'], 'lf', k['loopfilter']['ms'], 'pad', k['pad']['ms'], d['parity_check'])"
done
