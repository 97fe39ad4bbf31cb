#!/bin/bash
# Bench lines and ncu evidence of one round: usage tools/final_profile.sh <tag>  (run under gpurun, one GPU)
# Writes gpurun_out/<tag>_*.  Every ncu pass follows a plain run of the same command that exited 0.
tag=${1:-vX}
o=gpurun_out
mkdir -p $o
line() { grep '^{' | tail -1; }
python bench.py 2> $o/${tag}_4k.err | line > $o/${tag}_bench_4k.json
python bench.py --impl reference 2>/dev/null | line > $o/${tag}_bench_reference.json
for w in 1080p 1080p422 1080p444; do python bench.py --workload $w --no-cpu 2>/dev/null | line > $o/${tag}_bench_$w.json; done
python bench.py --workload 8k --no-cpu --steps 5 2>/dev/null | line > $o/${tag}_bench_8k.json
python bench.py --coeffs levels --bs derived --copy-out packed --no-cpu 2>/dev/null | line > $o/${tag}_bench_4k_levels_derived_packed.json
python bench.py --fill append --no-cpu 2>/dev/null | line > $o/${tag}_bench_4k_append.json
# pictures with intra CUs: 5 % intra CUs in every picture; the same with every 8th picture all intra
python bench.py --intra-pct 5 --no-cpu 2>/dev/null | line > $o/${tag}_bench_4k_intra5.json
python bench.py --intra-pct 5 --gop 8 --no-cpu 2>/dev/null | line > $o/${tag}_bench_4k_intra5_gop8.json
python tools/intra_cost.py > $o/${tag}_intra_cost.txt 2>&1
cmd="python tools/one_frame.py"
$cmd > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-check --no-chain --frames 8 > $o/ncu1.log 2>&1
for k in mc_kernel residual_kernel loopfilter_kernel prepare_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -f -o $o/${tag}_$k $cmd > $o/ncu_$k.log 2>&1
done
python tools/intra_one_frame.py --4k > /dev/null 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:intra_kernel -s 1 -c 1 -f -o $o/${tag}_intra_kernel python tools/intra_one_frame.py --4k --pct 25 > $o/ncu_intra.log 2>&1
ls -la $o | tail -20
