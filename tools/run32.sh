timeout 1200 python -m pytest tests/test_gpu_pipeline.py -m gpu -q -k "ragged" > gpurun_out/gpu_tests32.log 2>&1; echo "tests rc=$?"; tail -n 25 gpurun_out/gpu_tests32.log
