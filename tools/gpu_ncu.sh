# ncu --set full capture of the (second) admm_kernel launch of a small solve; summaries are produced on the CPU side afterwards
TAG=${1:-ncu}
cd $GRAFT_REPO_ROOT
python tools/run_small.py 148 > gpurun_out/${TAG}_plain.log 2>&1; echo "plain exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:admm -c 1 -f -o gpurun_out/${TAG}_admm python tools/run_small.py 148 > gpurun_out/${TAG}_ncu.log 2>&1; echo "ncu full exit $?"
