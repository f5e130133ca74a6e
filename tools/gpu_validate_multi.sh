# usage: bash tools/gpu_multi_c.sh N "<layouts>" "<pytest -k>"   — sweep + tests + driver-style bench on N GPUs
N=$1; LAYOUTS=$2
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
rm -f gpurun_out/scaling_sweep_n${N}_packed.jsonl
timeout 600 $TR --master-port 29612 tools/scaling_sweep.py --grid 32768 --steps 400 --layouts $LAYOUTS --min-waves 2 --ariths fp32,strict-fma --tag _packed > gpurun_out/sweep${N}_packed.log 2>&1; grep '^{' gpurun_out/sweep${N}_packed.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['arith'], d['px'],'x',d['py'], round(d.get('glups',0)), 'waits', d.get('halo_waits_rank0'), d.get('verified'), d.get('error',''))"
timeout 400 $TR --master-port 29613 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n${N}_packed.log 2>&1; grep '^{' gpurun_out/bench_n${N}_packed.log | tail -1
timeout 500 python -m pytest tests/test_dist_gpu.py -m gpu -q --timeout=400 -k "$3" > gpurun_out/gputest_multi_n${N}_packed.log 2>&1; echo "rc=$?" >> gpurun_out/gputest_multi_n${N}_packed.log; tail -3 gpurun_out/gputest_multi_n${N}_packed.log
