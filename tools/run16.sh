set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests16.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/gpu_tests16.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench16_n1.json 2> gpurun_out/bench16_n1.err; echo "bench1 rc=$?"; cut -c1-900 gpurun_out/bench16_n1.json; tail -n 3 gpurun_out/bench16_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/bench16_n2.json 2> gpurun_out/bench16_n2.err; echo "bench2 rc=$?"; cut -c1-900 gpurun_out/bench16_n2.json; tail -n 12 gpurun_out/bench16_n2.err
