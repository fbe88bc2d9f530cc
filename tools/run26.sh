timeout 900 python -m pytest tests/test_gpu_jpeg.py tests/test_gpu_cli.py -m gpu -q > gpurun_out/gpu_tests26.log 2>&1; echo "tests rc=$?"; tail -n 30 gpurun_out/gpu_tests26.log
