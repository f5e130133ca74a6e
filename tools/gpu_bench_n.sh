# usage: bash tools/gpu_bench_n.sh N   — the driver's bench command on N GPUs (both arms)
N=$1
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 500 $TR --master-port 29621 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n${N}_final.log 2>&1; echo "rc=$?"; grep '^{' gpurun_out/bench_n${N}_final.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print({k:d.get(k) for k in ('value','n_gpus','verified','ms_per_step')}, d['config']['parallelism'], 'fp32', d['fp32_mode']['value'], 'e2e', d['e2e']['value'], d['e2e'].get('e2e_roofline'), d['e2e']['how'][:60])"
