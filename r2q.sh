set -x
cp mqt-qmap_b200/csrc/obj/route_h0.o gpurun_out/r2q_route_h0.o; cp mqt-qmap_b200/csrc/obj/route_h4.o gpurun_out/r2q_route_h4.o; cp mqt-qmap_b200/csrc/route_core.hpp gpurun_out/r2q_route_core.hpp
python -m pytest tests -m gpu -x -q > gpurun_out/r2q_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2q_gputests.log
timeout 600 python bench.py --no-extra --no-cpu-baseline --steps 2 --warmup 1 --batch 32768 > gpurun_out/r2q_c4_32k.json 2> gpurun_out/r2q_c4_32k.err
python bench.py --workload c4 --no-cpu-baseline --no-extra --steps 1 --warmup 1 --batch 2048 > gpurun_out/r2q_c4_plain.json 2> gpurun_out/r2q_c4_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:route_kernel -s 2 -c 1 -f -o gpurun_out/r2q_c4 python bench.py --workload c4 --no-cpu-baseline --no-extra --steps 1 --warmup 1 --batch 2048 > gpurun_out/r2q_c4_ncu.log 2>&1
python bench.py --workload c2 --no-cpu-baseline --no-extra --steps 2 --warmup 1 > gpurun_out/r2q_c2_plain.json 2> gpurun_out/r2q_c2_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:route_kernel -s 2 -c 1 -f -o gpurun_out/r2q_c2 python bench.py --workload c2 --no-cpu-baseline --no-extra --steps 2 --warmup 1 > gpurun_out/r2q_c2_ncu.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2q_launches_c2.csv python bench.py --workload c2 --no-cpu-baseline --no-extra --steps 2 --warmup 1 > gpurun_out/r2q_launches_c2.log 2>&1
true
