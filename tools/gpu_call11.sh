set -x
timeout 900 python -m pytest tests -m gpu -q --timeout=120 -x > gpurun_out/gputest11.log 2>&1; echo "rc=$?" >> gpurun_out/gputest11.log; tail -4 gpurun_out/gputest11.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
rm -f gpurun_out/scaling_sweep_n2.jsonl
timeout 600 $TR --master-port 29612 tools/scaling_sweep.py --grid 32768 --steps 400 --layouts 2x1,1x2 --min-waves 4 --ariths fp32,strict-fma > gpurun_out/sweep2b.log 2>&1; grep '^{' gpurun_out/sweep2b.log | cut -c1-330
