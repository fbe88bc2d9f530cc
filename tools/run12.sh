set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests12.log 2>&1; echo "tests rc=$?"; tail -n 5 gpurun_out/gpu_tests12.log
timeout 600 python tools/scan_check.py --views 49 --width 1600 --height 1200 --nsrc 10 --passes 2 > gpurun_out/scan49_12.log 2>&1; echo "scan rc=$?"; tail -n 40 gpurun_out/scan49_12.log
