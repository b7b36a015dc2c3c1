# 8-GPU lines, launched the way the driver does
set -x
mkdir -p gpurun_out/n8
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/n8/bench_default.json 2> gpurun_out/n8/bench_default.err; python -c "
import json; d=json.load(open('gpurun_out/n8/bench_default.json')); h=d['secondary']['rnghunt']; print(d['value'], d['result_check']['reduced_hist_total']==d['result_check']['reduced_expected'], h['value'], h['ms_per_step'], h['result_check']['best'][:4], h['result_check']['stage_counts'])"
for w in lutopt_stream prbs_slicer clt_stream; do
  timeout 300 $TR bench.py --gpus 8 --steps 3 --warmup 3 --workload $w > gpurun_out/n8/bench_$w.json 2> gpurun_out/n8/bench_$w.err; cut -c1-140 gpurun_out/n8/bench_$w.json
done
timeout 600 python -m pytest tests/test_gpu_multirank.py -x -q -m gpu 2>&1 | tail -2
