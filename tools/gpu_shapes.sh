cd $GRAFT_REPO_ROOT
TAG=${1:-shapes}
python tools/dbg_shape.py 3 12; python tools/dbg_shape.py 5 6; python tools/dbg_shape.py 7 4; python tools/dbg_shape.py 3 6 qp_adaptive_rho=1
timeout 900 python -m pytest tests/test_gpu_shapes.py -x -q 2>&1 | tail -3
for sh in 3,12 5,6 7,4; do timeout 900 python bench.py --shape $sh --steps 2 --warmup 1 --batch 296 --no-cpu-baseline > gpurun_out/${TAG}_bench_shape_${sh/,/_}.json 2> gpurun_out/${TAG}_bench_shape.err; echo "bench $sh exit $?"; python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench_shape_${sh/,/_}.json'));print(d['value'],d['ms_per_step'],d['roofline']['frac'],d['roofline']['operator_stream']['achieved_gbs'])"; done
