timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r49_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/r49_smoke.log
timeout 900 python bench.py > gpurun_out/r49_bench.json 2> gpurun_out/r49_bench.err; echo "bench rc=$?"; tail -n 2 gpurun_out/r49_bench.err
grep "^{" gpurun_out/r49_bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'], d['result_checksum'], d['fusion']['points'], d['roofline']['achieved'], d['roofline']['frac'], d['kernel_ms_per_step'])"
CMD="python tools/run_one_view.py --runs 2 --tex auto"
timeout 300 $CMD > gpurun_out/r49_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:sweep_kernel|init_kernel' -s 5 -c 5 -f -o gpurun_out/r1r_sweep_u8_full $CMD > gpurun_out/r49_ncu.log 2>&1
echo "ncu rc=$?"; tail -n 2 gpurun_out/r49_ncu.log
