# usage: tools/run_scale.sh N  -> gpurun_out/bench_scale_nN.json
N=$1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --no-cpu-baseline > gpurun_out/bench_scale_n$N.json 2> gpurun_out/bench_scale_n$N.err; echo "bench$N rc=$?"; grep "^{" gpurun_out/bench_scale_n$N.json | cut -c1-700; tail -n 5 gpurun_out/bench_scale_n$N.err
