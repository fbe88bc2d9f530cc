timeout 1200 python -m pytest tests/test_gpu_cli.py tests/test_gpu_fusion.py -m gpu -q > gpurun_out/gpu_tests31.log 2>&1; echo "tests rc=$?"; tail -n 4 gpurun_out/gpu_tests31.log
