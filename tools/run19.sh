timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests19.log 2>&1; echo "tests rc=$?"; tail -n 8 gpurun_out/gpu_tests19.log
timeout 600 python tools/scan_check.py --views 12 --width 1600 --height 1200 --nsrc 10 > gpurun_out/scan12_19.log 2>&1; echo "scan rc=$?"; grep -E "stage|total" gpurun_out/scan12_19.log | cut -c1-230
