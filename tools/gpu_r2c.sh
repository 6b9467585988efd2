cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2c_gputests.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/r2c_gputests.log
timeout 900 python bench.py --global-batch 20000 > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench exit $?"; tail -3 gpurun_out/r2c_bench.err
for sh in 3,12 5,6 7,4; do timeout 900 python bench.py --shape $sh --steps 2 --warmup 1 --batch 296 > gpurun_out/r2c_bench_shape_${sh/,/_}.json 2> gpurun_out/r2c_bench_shape.err; echo "bench $sh exit $?"; tail -3 gpurun_out/r2c_bench_shape.err; done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2c_bench_reference.json 2>/dev/null; echo "ref exit $?"
