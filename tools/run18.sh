set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/gpu_tests18.log 2>&1; echo "tests rc=$?"; tail -n 15 gpurun_out/gpu_tests18.log
timeout 600 python tools/scan_check.py --views 49 --width 1600 --height 1200 --nsrc 10 --passes 2 > gpurun_out/scan49_18.log 2>&1; echo "scan rc=$?"; grep -E "stage|total" gpurun_out/scan49_18.log | cut -c1-230
