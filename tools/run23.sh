timeout 600 python tools/fusion_check.py --views 12 --oracle > gpurun_out/fusion12.log 2>&1; echo "rc=$?"; tail -n 8 gpurun_out/fusion12.log
timeout 600 python tools/fusion_check.py --views 49 > gpurun_out/fusion49.log 2>&1; echo "rc=$?"; tail -n 6 gpurun_out/fusion49.log
