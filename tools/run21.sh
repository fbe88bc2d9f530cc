set -x
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke21.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke21.log
CMD="python tools/scan_check.py --views 12 --width 1600 --height 1200 --nsrc 10"
timeout 600 $CMD > gpurun_out/plain12_21.log 2>&1 && timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:^(sweep|init)_kernel' -s 240 -c 10 -f -o gpurun_out/r1g_sweep_init_full $CMD > gpurun_out/ncu21.log 2>&1
echo "rc=$?"; tail -n 3 gpurun_out/ncu21.log; ls -la gpurun_out/*.ncu-rep
