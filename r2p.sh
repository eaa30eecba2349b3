set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2p_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2p_gputests.log
timeout 600 python bench.py --workload c3 --no-extra --batch 1208 --steps 2 --warmup 1 > gpurun_out/r2p_c3_small.json 2> gpurun_out/r2p_c3_small.err; echo "rc=$?" >> gpurun_out/r2p_c3_small.err
free -g | head -2 >> gpurun_out/r2p_c3_small.err
timeout 1200 python bench.py --workload c3 --no-extra --steps 2 --warmup 1 > gpurun_out/r2p_c3_full.json 2> gpurun_out/r2p_c3_full.err; echo "rc=$?" >> gpurun_out/r2p_c3_full.err
true
