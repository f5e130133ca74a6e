# gpurun payload: the whole -m gpu suite, smoke() and the bench (own arm) on ONE GPU.
set -x
timeout 600 python -m pytest tests -m gpu -q --timeout=120 -x > gpurun_out/gputest_final.log 2>&1; echo "rc=$?" >> gpurun_out/gputest_final.log; tail -6 gpurun_out/gputest_final.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_n1_final.log 2>&1; grep '^{' gpurun_out/bench_n1_final.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print({k:d.get(k) for k in ('value','verified','gpu_launches')}, 'k', d['config']['k'], 'fp32', d['fp32_mode']['value'], 'team', d.get('team_mode',{}).get('value'), 'e2e', d['e2e']['value'], d['e2e']['e2e_roofline'])"
