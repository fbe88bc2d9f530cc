set -x
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
timeout 900 python bench.py > gpurun_out/r41_bench.json 2> gpurun_out/r41_bench.err; echo "bench rc=$?"; tail -n 3 gpurun_out/r41_bench.err
timeout 900 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/r41_bench_ref.json 2> gpurun_out/r41_bench_ref.err; echo "ref rc=$?"; tail -n 3 gpurun_out/r41_bench_ref.err
timeout 900 $CMD > gpurun_out/r41_bench_short.json 2>&1 && timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:(sweep|init|rng_init|setup_pairs|depth_normal|median_filter|split_maps|pack_planes|jbu_depth|resize_linear|pad_replicate|support_cells|tri_raster|big_raster|prior_finalize|fp32_peak|mufu_peak|tex_peak|check_u8|to_u8_replicate)_kernel' -c 40000 --csv --log-file gpurun_out/r41_launches_bench.csv $CMD > gpurun_out/r41_ncu.log 2>&1
echo "rc=$?"; tail -n 2 gpurun_out/r41_ncu.log | cut -c1-300; wc -l gpurun_out/r41_launches_bench.csv
