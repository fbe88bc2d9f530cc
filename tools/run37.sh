# times the kernel variants under build/variants on one full-resolution view, R8 and R32F texels
for f in build/variants/lib_*.so; do
  for t in auto r32f; do
    echo "== $f $t $(ACMMP_B200_LIB=$PWD/$f timeout 300 python tools/run_one_view.py --runs 3 --tex $t 2>&1 | tail -n 2 | head -n 1)"
  done
done 2>&1 | tee gpurun_out/r37_variants.txt
