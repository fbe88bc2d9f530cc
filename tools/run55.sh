timeout 900 python bench.py > gpurun_out/r55_bench.json 2> gpurun_out/r55_bench.err; echo "bench rc=$?"; tail -n 3 gpurun_out/r55_bench.err
grep "^{" gpurun_out/r55_bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'], d['result_checksum'], d['planar_prior'], d['fusion'])"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r55_tests.log 2>&1; echo "tests rc=$?"; tail -n 4 gpurun_out/r55_tests.log
