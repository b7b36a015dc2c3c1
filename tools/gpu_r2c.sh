# round 2, call C (2 GPUs): NCCL multirank test, default bench and sharded legs under torchrun, multi-GPU CLI
set -x
mkdir -p gpurun_out/r2c
nvidia-smi -L
timeout 900 python -m pytest tests/test_gpu_multirank.py tests/test_gpu_cli.py -m gpu -x -q --durations=5 2>&1 | tail -12
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/r2c/bench_default_n2.json 2> gpurun_out/r2c/bench_default_n2.err; cat gpurun_out/r2c/bench_default_n2.json; tail -3 gpurun_out/r2c/bench_default_n2.err
for w in prbs_detect prbs_slicer; do
  timeout 400 $TR bench.py --gpus 2 --steps 3 --warmup 3 --workload $w > gpurun_out/r2c/bench_${w}_n2.json 2> gpurun_out/r2c/bench_${w}_n2.err; cut -c1-1500 gpurun_out/r2c/bench_${w}_n2.json; tail -2 gpurun_out/r2c/bench_${w}_n2.err
done
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2c/bench_reference.json 2>/dev/null; cut -c1-600 gpurun_out/r2c/bench_reference.json
