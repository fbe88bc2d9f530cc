# usage: tools/build_variant.sh NAME [nvcc -D flags...]   -> build/variants/lib_NAME.so (+ register report)
name=$1; shift
mkdir -p build/variants
nvcc -O3 --use_fast_math -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared "$@" -Xptxas -v \
  -o build/variants/lib_$name.so acmmp_b200/csrc/kernels.cu acmmp_b200/csrc/capi.cu acmmp_b200/csrc/noalloc.cu acmmp_b200/csrc/prior.cpp acmmp_b200/csrc/jpeg.cpp build/fusion.o build/pairs.o build/seeded.o -lnvjpeg_static -lculibos -lcudart -ldl 2>&1 \
  | grep -A2 "Compiling entry function '_ZN5acmmp1[12]\(sweep\|init\)_kernel" | grep -E "Used|spill" | tr '\n' ' ' | sed "s/^/$name: /"; echo
