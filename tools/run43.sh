bash tools/run42.sh
timeout 600 python tools/prior_ub_probe.py 2>&1 | tee gpurun_out/r43_prior_ub.txt | cut -c1-1500
