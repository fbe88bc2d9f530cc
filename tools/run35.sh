./build/tex_u8_probe > gpurun_out/tex_u8_probe.txt 2>&1; cat gpurun_out/tex_u8_probe.txt
