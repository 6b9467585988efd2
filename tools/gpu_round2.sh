# Round-2 evidence run (one gpurun call): GPU tests, bench with the CPU baseline, the reference arm, BASELINE config 4 shapes,
# the phase profile (libmpcb_prof.so from `make prof`), the ncu launch list and one ncu --set full capture each of admm_kernel
# and of the linearise+factorise stage launch.  Summaries are derived on the CPU side afterwards (tools/ncu_hotlines.py, ...).
TAG=${1:-r2}
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_gputests.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${TAG}_gputests.log
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
timeout 900 python bench.py --impl reference > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "reference arm exit $?"
for sh in 3,12 5,6 7,4; do
  timeout 600 python bench.py --shape $sh --steps 2 --warmup 3 --batch 296 --no-cpu-baseline > gpurun_out/${TAG}_bench_shape_${sh/,/_}.json 2> gpurun_out/${TAG}_bench_shape.err; echo "bench $sh exit $?"
done
if [ -f mpc_motion_planner_b200/libmpcb_prof.so ]; then
  MPCB_LIB=mpc_motion_planner_b200/libmpcb_prof.so timeout 300 python tools/phase_profile.py 148 > gpurun_out/${TAG}_phase_cycles.txt 2>&1; echo "phase profile exit $?"
fi
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_ncu1.log 2>&1; echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:admm -c 1 -f -o gpurun_out/${TAG}_admm python tools/run_small.py 148 > gpurun_out/${TAG}_ncu_admm.log 2>&1; echo "ncu admm exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sqp_stage -s 1 -c 1 -f -o gpurun_out/${TAG}_stage python tools/run_small.py 148 > gpurun_out/${TAG}_ncu_stage.log 2>&1; echo "ncu stage exit $?"
python - <<EOF
import json
for f in ["bench", "bench_reference", "bench_shape_3_12", "bench_shape_5_6", "bench_shape_7_4"]:
    try:
        d = json.load(open("gpurun_out/${TAG}_%s.json" % f))
        print(f, d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("value"), (d.get("roofline") or {}).get("frac"), (d.get("clocks") or {}).get("reasons"))
    except Exception as e:
        print(f, "unreadable:", e)
EOF
