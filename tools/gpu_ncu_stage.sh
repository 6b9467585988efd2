# ncu --set full capture of the linearise+factorise stage launch (kind 1 = 2nd sqp_stage_kernel launch) of a small solve
TAG=${1:-ncu}
cd $GRAFT_REPO_ROOT
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sqp_stage -s 1 -c 1 -f -o gpurun_out/${TAG}_stage python tools/run_small.py 148 > gpurun_out/${TAG}_ncu_stage.log 2>&1; echo "ncu stage exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:admm -c 1 -f -o gpurun_out/${TAG}_admm python tools/run_small.py 148 > gpurun_out/${TAG}_ncu_admm.log 2>&1; echo "ncu admm exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_ncu1.log 2>&1; echo "ncu list exit $?"
