python -m pytest tests/test_gpu_dropin.py -m gpu -x -q > gpurun_out/r2o_dropin.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2o_dropin.log
