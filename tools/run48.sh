bash tools/run47.sh
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r48_tests.log 2>&1; echo "tests rc=$?"; tail -n 6 gpurun_out/r48_tests.log
