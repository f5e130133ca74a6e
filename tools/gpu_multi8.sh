N=8
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -q --timeout=400 -k "eight" > gpurun_out/gputest_multi_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/gputest_multi_n$N.log; tail -3 gpurun_out/gputest_multi_n$N.log
timeout 200 $TR --master-port 29611 tools/pcie_probe.py > gpurun_out/pcie$N.log 2>&1; tail -1 gpurun_out/pcie$N.log | cut -c1-700
rm -f gpurun_out/scaling_sweep_n$N.jsonl
timeout 900 $TR --master-port 29612 tools/scaling_sweep.py --grid 32768 --steps 400 --layouts 8x1,4x2,2x4 --min-waves 1,2,4 --bulk 1 --ariths fp32,strict-fma > gpurun_out/sweep$N.log 2>&1; grep '^{' gpurun_out/sweep$N.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['arith'], d['px'],'x',d['py'], 'mw',d['min_waves'],'bulk',d['bulk'], round(d.get('glups',0)), 'waits', d.get('halo_waits_rank0'), d.get('verified'), d.get('error',''))"
timeout 300 $TR --master-port 29613 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.log 2>&1; tail -1 gpurun_out/bench_n$N.log | cut -c1-300
timeout 300 $TR --master-port 29614 bench.py --gpus $N --steps 1000 --warmup 24 --arith fp32 --no-extra --no-e2e --impl nccl-halo --k 8 > gpurun_out/bench_n${N}_nccl_k8.log 2>&1; tail -1 gpurun_out/bench_n${N}_nccl_k8.log | cut -c1-200
timeout 300 $TR --master-port 29615 bench.py --gpus $N --steps 1000 --warmup 24 --no-e2e --conv > gpurun_out/bench_n${N}_conv.log 2>&1; tail -1 gpurun_out/bench_n${N}_conv.log | cut -c1-200
timeout 400 $TR --master-port 29616 bench.py --gpus $N --steps 400 --warmup 24 --no-e2e --weak --grid 65536 --no-verify > gpurun_out/bench_n${N}_weak.log 2>&1; tail -1 gpurun_out/bench_n${N}_weak.log | cut -c1-200
