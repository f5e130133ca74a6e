set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -q --timeout=400 2>&1 | tail -3
rm -f gpurun_out/scaling_sweep_n2*.jsonl
for cfg in "1 0 80" "0 0 80" "1 1 80" "1 0 100000"; do set -- $cfg; H2D_FWD_BULK=$1 H2D_FWD_LINES=$2 H2D_EDGE_CHUNK=$3 timeout 300 $TR --master-port 29612 tools/scaling_sweep.py --grid 32768 --steps 400 --layouts 2x1,1x2 --min-waves 4 --ariths fp32 --tag _b$1l$2c$3 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('bulk $1 lines $2 chunk $3', d['px'],'x',d['py'], round(d['glups']), 'waits', d['halo_waits_rank0'], d['verified'])"; done
