set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_gputests.log
python bench.py --workload c4 --no-cpu-baseline --steps 1 --warmup 1 --batch 2048 > gpurun_out/r2a_c4_plain.json 2> gpurun_out/r2a_c4_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:route_kernel -s 2 -c 1 -f -o gpurun_out/r2a_c4_before python bench.py --workload c4 --no-cpu-baseline --steps 1 --warmup 1 --batch 2048 > gpurun_out/r2a_c4_ncu.log 2>&1
python bench.py --workload c2 --no-cpu-baseline --no-extra --steps 2 --warmup 1 > gpurun_out/r2a_c2_plain.json 2> gpurun_out/r2a_c2_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:route_kernel -s 2 -c 1 -f -o gpurun_out/r2a_c2_before python bench.py --workload c2 --no-cpu-baseline --no-extra --steps 2 --warmup 1 > gpurun_out/r2a_c2_ncu.log 2>&1
ls -la gpurun_out | tail -20
