timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r50_tests.log 2>&1; echo "tests rc=$?"; tail -n 6 gpurun_out/r50_tests.log
timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r50_bench.json 2> gpurun_out/r50_bench.err; echo "bench rc=$?"
grep "^{" gpurun_out/r50_bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e'], d['result_checksum'], d['fusion']['points'])"; tail -n 3 gpurun_out/r50_bench.err
