# the no-allocation entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last)
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/${1:-na}_alone.log 2>&1; echo "alone exit $?"; tail -15 gpurun_out/${1:-na}_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${1:-na}_all.log 2>&1; echo "all exit $?"; tail -15 gpurun_out/${1:-na}_all.log
python -m pytest tests -m gpu --collect-only -q 2>/dev/null | tail -8
