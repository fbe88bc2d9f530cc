nproc
for t in 1 0; do
ACMMP_DELAUNAY_THREADS=$t timeout 600 python tools/scan_check.py --views 5 --width 4032 --height 3024 --nsrc 4 > gpurun_out/r52_scan_c4_t$t.log 2>&1; echo "threads $t rc=$?"; grep -E "photo|total" gpurun_out/r52_scan_c4_t$t.log | cut -c1-200
done
timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r52_bench.json 2> gpurun_out/r52_bench.err; echo "bench rc=$?"
grep "^{" gpurun_out/r52_bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'], d['result_checksum'], d['kernel_ms_per_step'])"; tail -n 2 gpurun_out/r52_bench.err
