# multi-GPU bench lines (weak scaling at 1000 OCP/GPU + BASELINE config 3 as stats.config3); N = $1
N=${1:-2}
cd $GRAFT_REPO_ROOT
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2_bench_${N}gpu.json 2> gpurun_out/r2_bench_${N}gpu.err; echo "bench $N exit $?"; tail -2 gpurun_out/r2_bench_${N}gpu.err
python -c "
import json;d=json.load(open('gpurun_out/r2_bench_${N}gpu.json'));print(d['n_gpus'],d['value'],d['ms_per_step'],d['e2e']['value'],d['stats'].get('config3'))"
