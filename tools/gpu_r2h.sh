# round 2, 8-GPU lines (driver-style launch)
N=${1:-8}
set -x
mkdir -p gpurun_out/r2h
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus $N --steps 2 --warmup 3 > gpurun_out/r2h/bench_default_n$N.json 2> gpurun_out/r2h/bench_default_n$N.err; cut -c1-300 gpurun_out/r2h/bench_default_n$N.json; tail -2 gpurun_out/r2h/bench_default_n$N.err
for w in clt_stream lutopt_stream prbs_detect; do
  timeout 300 $TR bench.py --gpus $N --steps 3 --warmup 3 --workload $w > gpurun_out/r2h/bench_${w}_n$N.json 2> gpurun_out/r2h/bench_${w}_n$N.err; cut -c1-400 gpurun_out/r2h/bench_${w}_n$N.json; tail -2 gpurun_out/r2h/bench_${w}_n$N.err
done
