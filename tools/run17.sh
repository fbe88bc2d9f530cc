set -x
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
timeout 900 python bench.py > gpurun_out/bench17.json 2> gpurun_out/bench17.err; echo "bench rc=$?"; cat gpurun_out/bench17.json | cut -c1-3000; tail -n 3 gpurun_out/bench17.err
timeout 900 $CMD > gpurun_out/bench17_short.json 2>&1 && timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:^(sweep|init|rng_init|setup_pairs|depth_normal|median_filter|split_maps|pack_planes|jbu_depth|resize_linear|pad_replicate|support_cells|tri_raster|big_raster|prior_finalize|fp32_peak|mufu_peak|tex_peak)_kernel' -c 30000 --csv --log-file gpurun_out/launches_bench17.csv $CMD > gpurun_out/ncu17.log 2>&1
echo "rc=$?"; tail -n 2 gpurun_out/ncu17.log | cut -c1-300; wc -l gpurun_out/launches_bench17.csv
