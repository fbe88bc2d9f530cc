for t in unset 2 4 8; do
if [ $t = unset ]; then unset ACMMP_DELAUNAY_THREADS; else export ACMMP_DELAUNAY_THREADS=$t; fi
timeout 600 python tools/scan_check.py --views 5 --width 4032 --height 3024 --nsrc 4 > gpurun_out/r53_scan_c4_t$t.log 2>&1; echo "threads $t rc=$?"; grep -E "photo|total" gpurun_out/r53_scan_c4_t$t.log | cut -c1-200 | tail -2
done
