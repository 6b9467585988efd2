set -x
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1h_gputests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1h_gputests.log
timeout 600 python bench.py > gpurun_out/r1h_bench.json 2> gpurun_out/r1h_bench.err; echo "bench exit $?"
MPCB_LIB=mpc_motion_planner_b200/libmpcb_prof.so timeout 300 python tools/phase_profile.py 148 > gpurun_out/r1h_phase_cycles.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1h_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1; echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:admm -c 1 -o gpurun_out/r1h_admm python tools/run_small.py > gpurun_out/ncu2.log 2>&1; echo "ncu full exit $?"
tail -3 gpurun_out/r1h_gputests.log
cat gpurun_out/r1h_bench.json
