# quick GPU check: parity tests, bench without the CPU leg, per-phase cycle profile (outputs under gpurun_out/)
TAG=${1:-quick}
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_gputests.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/${TAG}_gputests.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench.json'));print(d['value'],d['ms_per_step'],d['roofline']['frac'],d['roofline']['step'])"
MPCB_LIB=mpc_motion_planner_b200/libmpcb_prof.so timeout 300 python tools/phase_profile.py 148 > gpurun_out/${TAG}_phase_cycles.txt 2>&1; cat gpurun_out/${TAG}_phase_cycles.txt
if [ "$2" = "ncu" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:admm -c 1 -o gpurun_out/${TAG}_admm python tools/run_small.py > gpurun_out/${TAG}_ncu.log 2>&1; echo "ncu full exit $?"
fi
