timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/gpu_tests30.log 2>&1; echo "tests rc=$?"; tail -n 6 gpurun_out/gpu_tests30.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -n 2
