set -x
python bench.py --steps 1 --warmup 1 > gpurun_out/bench_ours.log 2>&1; echo "ours rc=$?"; tail -n 3 gpurun_out/bench_ours.log
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "ref rc=$?"; tail -n 3 gpurun_out/bench_ref.log
