#!/bin/bash
# The two ncu recipes behind the summaries under profiles/ (B200_PROFILING.md), to be run on the GPU box through gpurun.
# (Round 1 kept them as scratch scripts tools/run38.sh / run41.sh / run49.sh / run56.sh, which the r1 summaries name.)
#   tools/ncu_recipes.sh full  TAG     one `--set full` capture of sweep_kernel / init_kernel on a 1600x1200 x 10-source view
#   tools/ncu_recipes.sh list  TAG     per-launch durations of one whole scan (the bench command, 1 warm-up + 1 step)
# Each recipe first runs its command WITHOUT ncu and only profiles if that exits 0; numbers printed under ncu are not bench values.
set -u
mode=${1:-full}; tag=${2:-r2}
mkdir -p gpurun_out
KERNELS='regex:(sweep|init|rng_init|setup_pairs|depth_normal|median_filter|merge_color|split_maps|pack_planes|jbu_depth|resize_linear|pad_replicate|support_cells|tri_raster|big_raster|prior_finalize|check_u8|to_u8_replicate|u8_to_float)_kernel'
if [ "$mode" = full ]; then
  CMD="python tools/run_one_view.py --runs 2 --tex auto"
  timeout 300 $CMD > gpurun_out/${tag}_plain.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:sweep_kernel|init_kernel' -s 5 -c 5 -f -o gpurun_out/${tag}_sweep_full $CMD > gpurun_out/${tag}_ncu.log 2>&1
  echo "ncu rc=$?"; tail -n 2 gpurun_out/${tag}_ncu.log
  ncu -i gpurun_out/${tag}_sweep_full.ncu-rep --page raw --csv > gpurun_out/${tag}_sweep_full_raw.csv 2>/dev/null
else
  CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
  timeout 600 $CMD > gpurun_out/${tag}_bench_short.json 2>&1 && \
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KERNELS" -c 6000 --csv --log-file gpurun_out/${tag}_launches_one_scan.csv $CMD > gpurun_out/${tag}_ncu_list.log 2>&1
  echo "rc=$?"; tail -n 2 gpurun_out/${tag}_ncu_list.log | cut -c1-200; wc -l gpurun_out/${tag}_launches_one_scan.csv
fi
