timeout 900 python -m pytest tests/test_gpu_cli.py -m gpu -q -x > gpurun_out/gpu_tests25.log 2>&1; echo "tests rc=$?"; tail -n 25 gpurun_out/gpu_tests25.log
