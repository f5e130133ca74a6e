# gpurun payload: the whole -m gpu suite, smoke(), both bench arms and an ncu capture of the team kernels on ONE GPU.
set -x
timeout 1200 python -m pytest tests -m gpu -q --timeout=120 -x > gpurun_out/gputest_final.log 2>&1; echo "rc=$?" >> gpurun_out/gputest_final.log; tail -6 gpurun_out/gputest_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_ref_final.log 2>&1; grep '^{' gpurun_out/bench_ref_final.log | tail -1 | cut -c1-400
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_n1_final.log 2>&1; grep '^{' gpurun_out/bench_n1_final.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print({k:d.get(k) for k in ('value','verified','gpu_launches')}, 'fp32', d['fp32_mode']['value'], 'e2e', {k:v for k,v in d['e2e'].items() if k not in ('what','how')})"
timeout 200 python tools/ncu_team_target.py 16384 > gpurun_out/ncu_team_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tk_kernel|tk_team_kernel" -o gpurun_out/prof_team python tools/ncu_team_target.py 16384 > gpurun_out/ncu_team_run.log 2>&1
tail -2 gpurun_out/ncu_team_plain.log; tail -2 gpurun_out/ncu_team_run.log
