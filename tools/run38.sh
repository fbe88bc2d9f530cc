CMD="python tools/run_one_view.py --runs 2 --tex auto"
timeout 300 $CMD > gpurun_out/r38_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:sweep_kernel|init_kernel' -s 5 -c 5 -f -o gpurun_out/r1o_sweep_u8_full $CMD > gpurun_out/r38_ncu.log 2>&1
echo "rc=$?"; tail -n 3 gpurun_out/r38_ncu.log; ls -la gpurun_out/*.ncu-rep
