set -x
mkdir -p gpurun_out/n4
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus 4 --steps 3 --warmup 3 > gpurun_out/n4/bench_default_n4.json 2> gpurun_out/n4/bench_default_n4.err; python -c "
import json; d=json.load(open('gpurun_out/n4/bench_default_n4.json')); h=d['secondary']['rnghunt']; print(d['value'], d['result_check']['reduced_hist_total']==d['result_check']['reduced_expected'], h['value'], h['ms_per_step'], h['result_check']['best'][:4], h['result_check']['stage_counts'])"
timeout 300 $TR bench.py --gpus 4 --steps 3 --warmup 3 --workload prbs_detect > gpurun_out/n4/bench_prbs_detect_n4.json 2>/dev/null; cut -c1-120 gpurun_out/n4/bench_prbs_detect_n4.json
