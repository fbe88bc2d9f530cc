timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests24.log 2>&1; echo "tests rc=$?"; tail -n 4 gpurun_out/gpu_tests24.log
timeout 1200 python bench.py > gpurun_out/bench24.json 2> gpurun_out/bench24.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench24.json')); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','fusion','kernel_ms_per_step')}); print(d['e2e']); print(d['roofline']['frac'], d['roofline']['achieved'], d['roofline']['peak'])"; tail -n 3 gpurun_out/bench24.err
