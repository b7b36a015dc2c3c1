set -x
mkdir -p gpurun_out/r2g
for i in 1 2; do
  timeout 900 python -X faulthandler -m pytest tests -m gpu -q -x --deselect tests/test_gpu_config_scale.py > gpurun_out/r2g/pytest_$i.log 2>&1; echo rc=$?; tail -3 gpurun_out/r2g/pytest_$i.log | cut -c1-300
done
timeout 600 python -m pytest tests/test_gpu_clt.py -m gpu -q -x > gpurun_out/r2g/pytest_clt.log 2>&1; echo rc=$?; tail -3 gpurun_out/r2g/pytest_clt.log | cut -c1-300
