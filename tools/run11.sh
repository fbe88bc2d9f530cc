set -x
CMD="python tools/scan_check.py --views 12 --width 1600 --height 1200 --nsrc 10"
$CMD > gpurun_out/plain12.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:^(sweep|init|rng_init|setup_pairs|depth_normal|median_filter|split_maps|pack_planes|jbu_depth|resize_linear|pad_replicate|support_cells|tri_raster|prior_finalize)_kernel' -c 4000 --csv --log-file gpurun_out/launches_scan12.csv $CMD > gpurun_out/ncu12.log 2>&1
echo "rc=$?"; tail -n 3 gpurun_out/ncu12.log; wc -l gpurun_out/launches_scan12.csv
