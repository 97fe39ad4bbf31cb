#!/bin/bash
# Bench lines and ncu evidence of one round: usage tools/final_profile.sh <tag>  (run under gpurun, one GPU)
# Writes gpurun_out/<tag>_*.  Every ncu pass follows a plain run of the same command that exited 0.
tag=${1:-vX}
o=gpurun_out
mkdir -p $o
line() { grep '^{' | tail -1; }
python bench.py 2> $o/${tag}_4k.err | line > $o/${tag}_bench_4k.json
python bench.py --impl reference 2>/dev/null | line > $o/${tag}_bench_reference.json
python bench.py --workload 1080p --no-cpu 2>/dev/null | line > $o/${tag}_bench_1080p.json
python bench.py --workload 8k --no-cpu 2>/dev/null | line > $o/${tag}_bench_8k.json
python bench.py --coeffs levels --bs derived --copy-out packed --no-cpu 2>/dev/null | line > $o/${tag}_bench_4k_levels_derived_packed.json
cmd="python bench.py --steps 2 --warmup 1 --no-cpu --no-check --frames 4 --bs derived"
$cmd > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches.csv $cmd > $o/ncu1.log 2>&1
for k in mc_kernel residual_kernel loopfilter_kernel bs_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 6 -c 1 -f -o $o/${tag}_$k $cmd > $o/ncu_$k.log 2>&1
done
ls -la $o | tail -20
