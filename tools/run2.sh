set -x
python tools/scan_check.py > gpurun_out/scan5.log 2>&1; echo "scan5 rc=$?"; tail -n 12 gpurun_out/scan5.log
python tools/scan_check.py --views 49 --width 1600 --height 1200 --nsrc 10 > gpurun_out/scan49.log 2>&1; echo "scan49 rc=$?"; tail -n 10 gpurun_out/scan49.log
