set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 200 python tools/perf_probe.py --nx 32768 --ny 16384 --steps 200 --ks 8 --layouts v4 --tag _tall 2>&1 | tail -1
timeout 200 python tools/perf_probe.py --nx 16384 --ny 32768 --steps 200 --ks 8 --layouts v4 --tag _wide 2>&1 | tail -1
for cfg in "1 100000" "1 80"; do set -- $cfg; H2D_DEBUG_NOFWD=$1 H2D_EDGE_CHUNK=$2 timeout 300 $TR --master-port 29612 tools/scaling_sweep.py --grid 32768 --steps 400 --layouts 2x1,1x2 --min-waves 4 --ariths fp32 --tag _nofwd$2 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('nofwd $1 chunk $2', d['px'],'x',d['py'], round(d['glups']), 'waits', d['halo_waits_rank0'], d['verified'])"; done
