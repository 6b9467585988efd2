cd $GRAFT_REPO_ROOT
for v in "$@"; do echo "== $v"; MPCB_LIB=mpc_motion_planner_b200/libmpcb_$v.so timeout 300 python bench.py --no-cpu-baseline --steps 3 --warmup 2 2>gpurun_out/variant.err | python -c "
import json,sys;d=json.loads(sys.stdin.read());print(d['value'],d['ms_per_step'],d['roofline']['step'])"; done
