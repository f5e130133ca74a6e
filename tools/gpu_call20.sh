set -x
timeout 900 python -m pytest tests -m gpu -q --timeout=120 -x > gpurun_out/gputest20.log 2>&1; echo "rc=$?" >> gpurun_out/gputest20.log; tail -8 gpurun_out/gputest20.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
rm -f gpurun_out/scaling_sweep_n2*.jsonl
timeout 600 $TR --master-port 29612 tools/scaling_sweep.py --grid 32768 --steps 400 --layouts 2x1,1x2 --min-waves 2 --ariths fp32,strict-fma 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['arith'], d['px'],'x',d['py'], round(d.get('glups',0)), 'waits', d.get('halo_waits_rank0'), d.get('verified'), d.get('error',''))"
timeout 300 $TR --master-port 29650 tools/nvlink_bytes.py --grid 32768 --steps 400 --px 2 --py 1 2>&1 | tail -1 | cut -c1-900
