#!/bin/bash
# Collect the round's evidence on the GPU box (run under gpurun from the repo root):
#   gpurun --timeout 1500 -- 'bash profiles/collect.sh r02 bench'      # bench lines (default run, reference arm, single workloads)
#   gpurun --timeout 1500 -- 'bash profiles/collect.sh r02 launches'   # ncu launch list of the headline run
#   gpurun --timeout 1500 -- 'bash profiles/collect.sh r02 full'       # ncu --set full captures of the two kernels, 3 workloads
# Outputs land in gpurun_out/; summarise here (no GPU) and copy what should be judged into profiles/.
set -u
tag=${1:-r02}; what=${2:-bench}
out=gpurun_out; mkdir -p $out
short="--no-subs --no-cpu-baseline --steps 2 --warmup 3 --passes 1"
if [ "$what" = bench ]; then
  python bench.py > $out/${tag}_bench_default.json 2> $out/${tag}_bench.err || exit 1
  python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference.json 2>> $out/${tag}_bench.err
elif [ "$what" = launches ]; then
  python bench.py $short --rows 4194304 > $out/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"check_kernel|gjk_queue|pack_flags" -c 200 --csv \
      --log-file $out/${tag}_launches_panda.csv python bench.py $short --rows 4194304 > $out/ncu_launch.log 2>&1
else
  for spec in panda_64m:4194304 ur5e_1m:1048576 panda_boxy_64m:4194304; do
    wl=${spec%%:*}; rows=${spec##*:}
    python bench.py --workload $wl --rows $rows $short > $out/plain_$wl.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k regex:"check_kernel|gjk_queue" -s 6 -c 4 -f \
        -o $out/${tag}_prof_$wl python bench.py --workload $wl --rows $rows $short > $out/ncu_full_$wl.log 2>&1
  done
fi
tail -n 3 $out/*.err $out/ncu_*.log 2>/dev/null | tail -n 20
