set -x
CMD="python tools/scan_check.py --views 49 --width 1600 --height 1200 --nsrc 10"
timeout 600 $CMD > gpurun_out/plain49_13.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:^(setup_pairs|split_maps|pack_planes|jbu_depth|resize_linear|pad_replicate|support_cells|tri_raster|big_raster|prior_finalize|depth_normal|median_filter|rng_init)_kernel' -c 6000 --csv --log-file gpurun_out/launches_aux49.csv $CMD > gpurun_out/ncu49_13.log 2>&1
echo "rc=$?"; tail -n 3 gpurun_out/ncu49_13.log; wc -l gpurun_out/launches_aux49.csv
