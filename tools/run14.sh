set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests14.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/gpu_tests14.log
timeout 600 python tools/scan_check.py --views 49 --width 1600 --height 1200 --nsrc 10 --passes 2 > gpurun_out/scan49_14.log 2>&1; echo "scan rc=$?"; grep -E "photo|total" gpurun_out/scan49_14.log | cut -c1-250
timeout 900 python bench.py > gpurun_out/bench14.json 2> gpurun_out/bench14.err; echo "bench rc=$?"; cat gpurun_out/bench14.json; tail -n 5 gpurun_out/bench14.err
