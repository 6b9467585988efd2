# bench a list of library variants: tools/gpu_variants.sh name1 name2 ...   (libmpcb_<name>.so)
cd $GRAFT_REPO_ROOT
for v in "$@"; do
  MPCB_LIB=mpc_motion_planner_b200/libmpcb_$v.so timeout 300 python bench.py --no-cpu-baseline --steps 3 > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err
  python -c "
import json;d=json.load(open('gpurun_out/var_$v.json'));print('$v',round(d['value']),round(d['ms_per_step'],2),d['roofline']['step']['admm_ms'])" || tail -3 gpurun_out/var_$v.err
done
