# usage: bash tools/gpu_multi.sh N "<layouts>" "<test filter>"   (N GPUs of one box)
N=$1; LAYOUTS=$2; KFILT=$3
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
timeout 900 python -m pytest tests/test_dist_gpu.py -m gpu -q --timeout=400 -k "$KFILT" > gpurun_out/gputest_multi_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/gputest_multi_n$N.log; tail -4 gpurun_out/gputest_multi_n$N.log
timeout 200 $TR --master-port 29611 tools/pcie_probe.py > gpurun_out/pcie$N.log 2>&1; tail -1 gpurun_out/pcie$N.log
rm -f gpurun_out/scaling_sweep_n$N.jsonl
timeout 900 $TR --master-port 29612 tools/scaling_sweep.py --grid 32768 --steps 400 --layouts $LAYOUTS --min-waves 1,2,4 --bulk 1,0 --ariths fp32,strict-fma > gpurun_out/sweep$N.log 2>&1; grep '^{' gpurun_out/sweep$N.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['arith'], d['px'],'x',d['py'], 'mw',d['min_waves'],'bulk',d['bulk'], round(d.get('glups',0)), 'waits', d.get('halo_waits_rank0'), d.get('verified'), d.get('error',''))"
timeout 300 $TR --master-port 29613 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.log 2>&1; tail -1 gpurun_out/bench_n$N.log | cut -c1-300
