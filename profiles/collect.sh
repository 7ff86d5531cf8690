#!/bin/bash
# Collect the round's evidence on the GPU box (run under gpurun from the repo root):
#   gpurun --timeout 1500 -- 'bash profiles/collect.sh r02 bench'      # bench lines (default run, reference arm, single workloads)
#   gpurun --timeout 1500 -- 'bash profiles/collect.sh r02 launches'   # ncu launch list of the headline run
#   gpurun --timeout 1500 -- 'bash profiles/collect.sh r02 full panda_64m:4194304'   # ONE ncu --set full capture of the two kernels
#                                                                  (one workload:rows per call; also ur5e_1m:1048576,
#                                                                   panda_boxy_64m:4194304, h1_16m:2097152)
#   gpurun --timeout 1500 -- 'bash profiles/collect.sh r02 fk'         # the same for fk_kernel<1> (poses stored)
#   gpurun --timeout 1500 -- 'bash profiles/collect.sh r02 edges'      # launch list of the planner-edge path, 2^18 edges
# then here (no GPU): summarize_ncu.py / regions.py / summarize_launches.py per capture, make_constants.py for the two JSON files.
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
elif [ "$what" = fk ]; then
  python bench.py --kernel fk --steps 3 --warmup 3 > $out/plain_fk.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:fk_kernel -s 3 -c 1 -f -o $out/${tag}_prof_fk \
      python bench.py --kernel fk --steps 3 --warmup 3 > $out/ncu_full_fk.log 2>&1
elif [ "$what" = edges ]; then
  e="--workload fr3_edges --rows 262144 --plan-seeds 0 --steps 2 --warmup 3"
  python bench.py $e > $out/plain_edges.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"edge_|gjk_queue" --csv \
      --log-file $out/${tag}_launches_edges.csv python bench.py $e > $out/ncu_launch_edges.log 2>&1
else
  spec=${3:-panda_64m:4194304}
  wl=${spec%%:*}; rows=${spec##*:}
  python bench.py --workload $wl --rows $rows $short > $out/plain_$wl.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"^(void )?check_kernel|gjk_queue" -s 6 -c 4 -f \
      -o $out/${tag}_prof_$wl python bench.py --workload $wl --rows $rows $short > $out/ncu_full_$wl.log 2>&1
fi
tail -n 3 $out/*.err $out/ncu_*.log 2>/dev/null | tail -n 20
