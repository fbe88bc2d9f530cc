set -x
CMD="python tools/scan_check.py --views 12 --width 1600 --height 1200 --nsrc 10"
timeout 600 $CMD > gpurun_out/plain12_15.log 2>&1 && timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:^sweep_kernel' -s 200 -c 6 -f -o gpurun_out/r1c_sweep_full $CMD > gpurun_out/ncu15.log 2>&1
echo "rc=$?"; tail -n 5 gpurun_out/ncu15.log; ls -la gpurun_out/*.ncu-rep
