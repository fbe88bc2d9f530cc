./build/tex_lds_probe 2>&1 | tee gpurun_out/tex_lds_probe.txt
