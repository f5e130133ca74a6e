# gpurun payload (4 GPUs): the driver's bench command on 4 GPUs, on 1 GPU (both arms), and the team tests
set -x
bash tools/gpu_bench_n.sh 4
CUDA_VISIBLE_DEVICES=0 timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_ref_final.log 2>&1; grep '^{' gpurun_out/bench_ref_final.log | tail -1 | cut -c1-330
CUDA_VISIBLE_DEVICES=0 timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_n1_final.log 2>&1; grep '^{' gpurun_out/bench_n1_final.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print({k:d.get(k) for k in ('value','verified','gpu_launches')}, 'k', d['config']['k'], 'fp32', d['fp32_mode']['value'], d['fp32_mode']['k'], 'team', d.get('team_mode',{}).get('value'), 'e2e', d['e2e']['value'], d['e2e']['e2e_roofline'])"
CUDA_VISIBLE_DEVICES=0 timeout 600 python -m pytest tests/test_gpu_team.py tests/test_gpu_kernels.py -m gpu -q --timeout=120 -x 2>&1 | tail -3
