# round-2 first GPU check: all GPU tests (incl. shapes / adaptive rho), quick bench of the base shape
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_gputests.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/r2a_gputests.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench exit $?"; cat gpurun_out/r2a_bench.json | head -c 600
