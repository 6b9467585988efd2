# multi-part splitting inside mpcb_solve_batch: GPU tests with the default (4 parts), bench with MPCB_PARTS = 1 / 2 / 4
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/parts_gputests.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/parts_gputests.log
for P in 1 2 4 1 4; do echo "== parts $P"; MPCB_PARTS=$P timeout 300 python bench.py --no-cpu-baseline --steps 5 --warmup 3 2>gpurun_out/parts.err | python -c "
import json,sys;d=json.loads(sys.stdin.read());print(d['value'],d['ms_per_step'],d['e2e']['value'],d['roofline']['frac'],d['roofline']['kernel_ms'],d['gpu_launches'],d['roofline']['step'])"; done
