timeout 600 python tools/prior_ub_probe.py 2>&1 | tee gpurun_out/r44_prior_ub.txt | cut -c1-600
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r44_tests.log 2>&1; echo "tests rc=$?"; tail -n 15 gpurun_out/r44_tests.log
