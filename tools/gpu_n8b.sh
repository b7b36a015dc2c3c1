set -x
mkdir -p gpurun_out/n8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517"
timeout 300 $TR bench.py --gpus 8 --steps 3 --warmup 3 --workload lutopt_stream > gpurun_out/n8/bench_lutopt_stream.json 2> gpurun_out/n8/bench_lutopt_stream.err; cut -c1-160 gpurun_out/n8/bench_lutopt_stream.json
