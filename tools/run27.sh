echo "== 3 levels: 5 views 4032x3024 (cap 3200 -> 3200x2400, 1600x1200, 800x600), 4 sources"
timeout 600 python tools/scan_check.py --views 5 --width 4032 --height 3024 --nsrc 4 > gpurun_out/scan_c4.log 2>&1; echo "rc=$?"; grep -E "levels|stage|total|Error|error" gpurun_out/scan_c4.log | cut -c1-220
echo "== 20 sources: 24 views 1920x1080"
timeout 600 python tools/scan_check.py --views 24 --width 1920 --height 1080 --nsrc 20 > gpurun_out/scan_c5.log 2>&1; echo "rc=$?"; grep -E "levels|stage|total|Error|error" gpurun_out/scan_c5.log | cut -c1-220
