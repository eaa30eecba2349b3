set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2n_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2n_gputests.log
true
