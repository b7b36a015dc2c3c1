# end-of-session evidence pass on one GPU: tests, smoke, every bench workload, launch lists
mkdir -p gpurun_out/fin
python -m pytest tests -m gpu -q > gpurun_out/fin/pytest_gpu.log 2>&1; tail -2 gpurun_out/fin/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/fin/smoke.log 2>&1; echo smoke rc=$?
python bench.py > gpurun_out/fin/bench_default.json 2> gpurun_out/fin/bench_default.err; echo default rc=$?
for w in prbs_stream prbs_detect prbs_slicer prbs_shaped_slicer tx_stream rnghunt clt_stream; do
  python bench.py --workload $w > gpurun_out/fin/bench_$w.json 2> gpurun_out/fin/bench_$w.err; echo $w rc=$?
done
python bench.py --workload clt_stream --noise-var-stream 8 > gpurun_out/fin/bench_clt_stream_i16.json 2>/dev/null; echo i16 rc=$?
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/fin/bench_reference.json 2>/dev/null; echo ref rc=$?
for f in gpurun_out/fin/bench_*.json; do echo $f; cut -c1-110 $f; done
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/fin/tx_launches.csv python bench.py --workload tx_stream --steps 2 --warmup 1 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:"prbs_stream_kernel|shaper_stream_kernel|shaped_slicer_err_kernel" -c 4 -o gpurun_out/fin/prof_tx python tools/ncu_targets_dec.py > gpurun_out/fin/ncu_tx.log 2>&1; tail -1 gpurun_out/fin/ncu_tx.log
