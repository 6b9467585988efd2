cd $GRAFT_REPO_ROOT
python tools/dbg_shape.py 3 12 2>&1 | cut -c1-250 | head
python tools/dbg_shape.py 5 6 2>&1 | cut -c1-250 | head
python tools/dbg_shape.py 7 4 2>&1 | cut -c1-250 | head
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_gputests.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/r2b_gputests.log
