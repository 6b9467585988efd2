cd $GRAFT_REPO_ROOT
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/${1:-t}_gputests.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/${1:-t}_gputests.log
