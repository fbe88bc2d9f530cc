timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_scale_n1.json 2> gpurun_out/bench_scale_n1.err; echo "bench1 rc=$?"; cut -c1-400 gpurun_out/bench_scale_n1.json
bash tools/run_scale.sh 2
bash tools/run_scale.sh 4
