# gpurun payload (2 GPUs): the whole -m gpu suite, then the driver's bench command on 2 GPUs and on 1
set -x
timeout 1500 python -m pytest tests -m gpu -q --timeout=300 -x > gpurun_out/gputest_2gpu_final.log 2>&1; echo "rc=$?" >> gpurun_out/gputest_2gpu_final.log; tail -6 gpurun_out/gputest_2gpu_final.log
bash tools/gpu_bench_n.sh 2
CUDA_VISIBLE_DEVICES=0 timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_n1_final.log 2>&1; grep '^{' gpurun_out/bench_n1_final.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print({k:d.get(k) for k in ('value','verified','gpu_launches')}, 'k', d['config']['k'], 'fp32', d['fp32_mode']['value'], d['fp32_mode']['k'], 'team', d.get('team_mode',{}).get('value'), 'e2e', d['e2e']['value'], d['e2e']['e2e_roofline'])"
