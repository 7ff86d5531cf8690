#!/bin/bash
# Collect the round's evidence on the GPU box (run under gpurun from the repo root):
#   gpurun --timeout 1500 -- 'bash profiles/collect.sh r01 launches'   # bench lines + ncu launch list
#   gpurun --timeout 1500 -- 'bash profiles/collect.sh r01 full'       # one --set full capture
# Outputs land in gpurun_out/; copy what should be judged into profiles/ afterwards.
set -u
tag=${1:-r01}; what=${2:-launches}
out=gpurun_out; mkdir -p $out
if [ "$what" = launches ]; then
  python bench.py > $out/${tag}_bench_ur5e_1m.json 2> $out/${tag}_bench.err || exit 1
  python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference.json 2>> $out/${tag}_bench.err
  for wl in panda_4096 panda_64m h1_16m rrr_float fr3_edges; do
    python bench.py --workload $wl --no-cpu-baseline > $out/${tag}_bench_${wl}.json 2>> $out/${tag}_bench.err
  done
  python bench.py --workload panda_64m --kernel fk --no-cpu-baseline > $out/${tag}_bench_fk_panda.json 2>> $out/${tag}_bench.err
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $out/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_ur5e.csv \
      python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $out/ncu_launch.log 2>&1
else
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"check_kernel|gjk_queue" -s 6 -c 4 -f \
      -o $out/${tag}_prof_check_ur5e python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/ncu_full.log 2>&1
fi
tail -n 3 $out/*.err $out/ncu_*.log 2>/dev/null | tail -n 20
