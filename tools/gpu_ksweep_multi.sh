# gpurun payload: strict-fma, sweep depth 8 vs 4 on N GPUs.  usage: bash tools/gpu_ksweep_multi.sh N PxQ
N=$1; LAYOUTS=$2
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
rm -f gpurun_out/scaling_sweep_n${N}_ks.jsonl
timeout 600 $TR --master-port 29612 tools/scaling_sweep.py --grid 32768 --steps 400 --layouts $LAYOUTS --min-waves 2 --ariths strict-fma --ks 8,4 --tag _ks > gpurun_out/sweep${N}_ks.log 2>&1; grep '^{' gpurun_out/sweep${N}_ks.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['arith'], 'k', d['k'], d['px'],'x',d['py'], round(d.get('glups',0)), 'waits', d.get('halo_waits_rank0'), d.get('verified'), d.get('error',''))"
