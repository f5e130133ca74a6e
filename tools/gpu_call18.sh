set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for cfg in "0 100000" "1 100000" "0 512" "0 256" "1 256" "0 128"; do set -- $cfg; H2D_FWD_LINES=$1 H2D_EDGE_CHUNK=$2 timeout 200 $TR --master-port 29612 tools/scaling_sweep.py --grid 32768 --steps 400 --layouts 1x2 --min-waves 2 --ariths fp32 --tag _l$1c$2 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('lines $1 chunk $2', d['px'],'x',d['py'], round(d['glups']), 'waits', d['halo_waits_rank0'], round(d['halo_wait_us_rank0']/max(1,d['halo_waits_rank0']),1), d['verified'])"; done
