# per-phase cycles of profiling builds with one phase's arithmetic stubbed out (libmpcb_abl<N>.so, -DMPCB_PROFILE -DMPCB_ABLATE=N)
cd $GRAFT_REPO_ROOT
for v in "$@"; do echo "== abl$v"; MPCB_LIB=mpc_motion_planner_b200/libmpcb_abl$v.so timeout 300 python tools/phase_profile.py 148 2>&1 | sed -n 1,12p; done
