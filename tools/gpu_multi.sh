# N-GPU pass of bench.py (launched the way the driver does); N = $1
N=${1:-2}
WL=${2:-"clt_fused rnghunt prbs_detect prbs_stream"}
set -x
nvidia-smi -L
for w in $WL; do
  EXTRA=""; [ $w = clt_fused ] && EXTRA="--log2-samples 38 --no-cpu-baseline"
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 --workload $w $EXTRA > gpurun_out/multi${N}_$w.json 2> gpurun_out/multi${N}_$w.err
  cat gpurun_out/multi${N}_$w.json; tail -3 gpurun_out/multi${N}_$w.err
done
