set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python tests/golden/make_ref_golden.py > gpurun_out/golden.log 2>&1; echo "golden rc=$?"
python tools/gpu_parity.py --tag r1c --big > gpurun_out/parity_r1c.log 2>&1; echo "parity rc=$?"
python tools/run_one_view.py --runs 3 > gpurun_out/one_view.log 2>&1; echo "oneview rc=$?"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.log 2>&1
tail -3 gpurun_out/smoke.log gpurun_out/golden.log gpurun_out/one_view.log
