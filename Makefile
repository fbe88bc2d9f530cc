# Build of the product library (libacmmp_b200.so, sm_100a only) and of the test
# infrastructure under oracle/.  `python -c "import __graft_entry__ as g; g.build()"`
# drives the same targets.
NVCC      ?= nvcc
CXX       ?= g++
CC        ?= gcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
# the reference's arithmetic flags (src/CMakeLists.txt:19) define its results
NVFLAGS   := -O3 --use_fast_math -std=c++17 $(ARCH) -lineinfo -Xcompiler -fPIC
CSRC      := acmmp_b200/csrc
LIB       := acmmp_b200/libacmmp_b200.so
REF       := /root/reference/src
CUDA_INC  ?= /usr/local/cuda/include
CUDA_LIB  ?= /usr/local/cuda/lib64
# the g++ on PATH: the image's $$CXX wrapper links libstdc++ statically, which breaks iostreams
# inside a library that is dlopen()ed into python
REF_CXX   ?= $(shell command -v g++)

all: $(LIB) oracle cli

# command line with the reference's flags and file layout (cli/acmmp_main.cpp) over the C ABI
cli: build/acmmp_main build/fuse_data build/io_selftest
build/io_selftest: cli/io_selftest.cpp $(wildcard acmmp_b200/host/*.hpp) include/acmmp_b200.h
	@mkdir -p build
	$(REF_CXX) -O2 -std=c++17 -Wall -o $@ cli/io_selftest.cpp -lz
build/fuse_data: cli/fuse_data.cpp cli/fuse_common.hpp $(wildcard acmmp_b200/host/*.hpp) include/acmmp_b200.h $(LIB)
	@mkdir -p build
	$(REF_CXX) -O2 -std=c++17 -Wall -o $@ cli/fuse_data.cpp -Lacmmp_b200 -lacmmp_b200 -Wl,-rpath,'$$ORIGIN/../acmmp_b200' -L$(CUDA_LIB) -lcudart -lz
build/acmmp_main: cli/acmmp_main.cpp cli/fuse_common.hpp $(wildcard acmmp_b200/host/*.hpp) include/acmmp_b200.h $(LIB)
	@mkdir -p build
	$(REF_CXX) -O2 -std=c++17 -Wall -pthread -o $@ cli/acmmp_main.cpp -Lacmmp_b200 -lacmmp_b200 -Wl,-rpath,'$$ORIGIN/../acmmp_b200' -L$(CUDA_LIB) -lcudart -lz

# fusion.cu restates host code of the reference (IEEE float / double): no fast math, no contraction
FUSIONFLAGS := -O3 -std=c++17 $(ARCH) -lineinfo -fmad=false -Xcompiler -fPIC

build/fusion.o: $(CSRC)/fusion.cu include/acmmp_b200.h
	@mkdir -p build
	$(NVCC) $(FUSIONFLAGS) -c -o $@ $(CSRC)/fusion.cu

build/pairs.o: $(CSRC)/pairs.cu include/acmmp_b200.h
	@mkdir -p build
	$(NVCC) $(FUSIONFLAGS) -c -o $@ $(CSRC)/pairs.cu

build/seeded.o: $(CSRC)/seeded.cu $(CSRC)/kernels.h include/acmmp_b200.h
	@mkdir -p build
	$(NVCC) $(FUSIONFLAGS) -c -o $@ $(CSRC)/seeded.cu

$(LIB): $(CSRC)/kernels.cu $(CSRC)/capi.cu $(CSRC)/noalloc.cu $(CSRC)/prior.cpp $(CSRC)/jpeg.cpp build/fusion.o build/pairs.o build/seeded.o $(wildcard $(CSRC)/*.cuh) $(wildcard $(CSRC)/*.inc) $(wildcard $(CSRC)/*.h) include/acmmp_b200.h
	$(NVCC) $(NVFLAGS) -shared -o $@ $(CSRC)/kernels.cu $(CSRC)/capi.cu $(CSRC)/noalloc.cu $(CSRC)/prior.cpp $(CSRC)/jpeg.cpp build/fusion.o build/pairs.o build/seeded.o -lnvjpeg_static -lculibos -lcudart -ldl

oracle: oracle/liboracle.so

oracle/liboracle.so: $(wildcard oracle/*.c) $(wildcard oracle/*.h)
	$(CC) -O2 -ffp-contract=off -fPIC -shared -o $@ $(wildcard oracle/*.c) -lm

# The reference's own kernels, compiled from where they lie (never copied into the
# repo), sm_100 + the reference's flags; only built where /root/reference exists.
ref: oracle/_ref/libacmmp_ref.so oracle/_ref/libacmmp_ref_host.so oracle/_ref/libacmmp_ref_host_ieee.so oracle/_ref/libacmmp_ref_full.so oracle/_ref/libacmmp_ref_full_p2.so oracle/_ref/libacmmp_dropin.so
oracle/_ref/libacmmp_ref.so: oracle/ref_harness.cu $(wildcard oracle/ref_stub/opencv2/*.hpp)
	@if [ -d $(REF) ]; then mkdir -p oracle/_ref && \
	$(NVCC) -O3 --use_fast_math --maxrregcount=128 -std=c++14 -gencode arch=compute_100,code=sm_100 \
	  -Ioracle/ref_stub -I$(REF) -Iinclude -Xcompiler -fPIC -shared -o $@ oracle/ref_harness.cu -lcudart ; \
	else echo "reference sources absent: keeping prebuilt oracle/_ref"; fi

# The reference PROGRAM (kernels + host class + ProcessProblem + JBU + RunFusion) in the order of its
# main(); cv::Subdiv2D is served by the product's Delaunay (acmmp_b200/csrc/prior.cpp) through a hook.
oracle/_ref/libacmmp_ref_full.so: oracle/ref_full_harness.cu acmmp_b200/csrc/prior.cpp $(wildcard oracle/ref_stub/opencv2/*.hpp)
	@if [ -d $(REF) ]; then mkdir -p oracle/_ref && \
	$(NVCC) -O3 --use_fast_math --maxrregcount=128 -std=c++14 -gencode arch=compute_100,code=sm_100 -w \
	  -Ioracle/ref_stub -I$(REF) -Iinclude -Xcompiler -fPIC -shared -o $@ oracle/ref_full_harness.cu acmmp_b200/csrc/prior.cpp -lcudart ; \
	else echo "reference sources absent: keeping prebuilt oracle/_ref"; fi

# The same program with patch P2 of BASELINE.md: `#define DEBUG` (ACMMP.cu:3) removed.  The patched translation unit is
# GENERATED at build time into the git-ignored oracle/_ref/gen_p2/ (sed deletes that one line; removed again after the
# compile); it is found first on the
# include path, everything else still comes from $(REF).
oracle/_ref/libacmmp_ref_full_p2.so: oracle/ref_full_harness.cu acmmp_b200/csrc/prior.cpp $(wildcard oracle/ref_stub/opencv2/*.hpp)
	@if [ -d $(REF) ]; then mkdir -p oracle/_ref/gen_p2 && \
	sed '/^#define DEBUG/d' $(REF)/ACMMP.cu > oracle/_ref/gen_p2/ACMMP.cu && ! grep -q '^#define DEBUG' oracle/_ref/gen_p2/ACMMP.cu && \
	$(NVCC) -O3 --use_fast_math --maxrregcount=128 -std=c++14 -gencode arch=compute_100,code=sm_100 -w \
	  -Ioracle/_ref/gen_p2 -Ioracle/ref_stub -I$(REF) -Iinclude -Xcompiler -fPIC -shared -o $@ oracle/ref_full_harness.cu acmmp_b200/csrc/prior.cpp -lcudart ; \
	rc=$$?; rm -rf oracle/_ref/gen_p2; exit $$rc; \
	else echo "reference sources absent: keeping prebuilt oracle/_ref"; fi

# The drop-in proof: the reference's UNCHANGED ACMMP.h + acmmp_definitions.cpp on top of
# oracle/dropin/ACMMP_dropin.cpp (class ACMMP / RunJBU / helpers over the C ABI) instead of its ACMMP.cpp / ACMMP.cu.
dropin: oracle/_ref/libacmmp_dropin.so
oracle/_ref/libacmmp_dropin.so: oracle/dropin/ACMMP_dropin.cpp oracle/dropin/dropin_main.cpp include/acmmp_b200.h $(LIB) $(wildcard oracle/ref_stub/opencv2/*.hpp)
	@if [ -d $(REF) ]; then mkdir -p oracle/_ref && \
	$(REF_CXX) -O2 -std=c++14 -fPIC -shared -w -Ioracle/ref_stub -I$(REF) -Iinclude -I$(CUDA_INC) -o $@ oracle/dropin/ACMMP_dropin.cpp oracle/dropin/dropin_main.cpp \
	  -Lacmmp_b200 -lacmmp_b200 -Wl,-rpath,'$$ORIGIN/../../acmmp_b200' -L$(CUDA_LIB) -lcudart ; \
	else echo "reference sources absent: keeping prebuilt oracle/_ref"; fi

# The reference's HOST code (RunFusion + helpers: ACMMP.cpp, acmmp_definitions.cpp) against the
# stand-in OpenCV.  Two builds: the reference's own flags (src/CMakeLists.txt:28, minus
# -march=native so that the library also runs on the GPU box's CPU) and plain IEEE arithmetic
# (-O2), against which the C restatement under oracle/ is bit-exact.
REF_HOST_SRC := oracle/ref_host_harness.cpp $(wildcard oracle/ref_stub/opencv2/*.hpp)
oracle/_ref/libacmmp_ref_host.so: $(REF_HOST_SRC)
	@if [ -d $(REF) ]; then mkdir -p oracle/_ref && \
	$(REF_CXX) -O3 -ffast-math -std=c++14 -fPIC -shared -w -Ioracle/ref_stub -I$(REF) -I$(CUDA_INC) -o $@ oracle/ref_host_harness.cpp -L$(CUDA_LIB) -lcudart ; \
	else echo "reference sources absent: keeping prebuilt oracle/_ref"; fi
oracle/_ref/libacmmp_ref_host_ieee.so: $(REF_HOST_SRC)
	@if [ -d $(REF) ]; then mkdir -p oracle/_ref && \
	$(REF_CXX) -O2 -std=c++14 -fPIC -shared -w -Ioracle/ref_stub -I$(REF) -I$(CUDA_INC) -o $@ oracle/ref_host_harness.cpp -L$(CUDA_LIB) -lcudart ; \
	else echo "reference sources absent: keeping prebuilt oracle/_ref"; fi

clean:
	rm -f $(LIB) build/fusion.o build/pairs.o build/seeded.o oracle/liboracle.so oracle/_ref/libacmmp_ref.so oracle/_ref/libacmmp_ref_host.so oracle/_ref/libacmmp_ref_host_ieee.so oracle/_ref/libacmmp_ref_full.so oracle/_ref/libacmmp_ref_full_p2.so
.PHONY: all oracle ref cli clean dropin
