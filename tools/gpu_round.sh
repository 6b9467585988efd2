# full evidence run: gpu tests, bench (with cpu baseline), phase profile, ncu launch list, ncu full capture of admm_kernel
TAG=${1:-round}
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_gputests.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${TAG}_gputests.log
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
MPCB_LIB=mpc_motion_planner_b200/libmpcb_prof.so timeout 300 python tools/phase_profile.py 148 > gpurun_out/${TAG}_phase_cycles.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_ncu1.log 2>&1; echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:admm -c 1 -o gpurun_out/${TAG}_admm python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_ncu2.log 2>&1; echo "ncu full exit $?"
cat gpurun_out/${TAG}_bench.json
