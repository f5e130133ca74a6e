# gpurun payload: tests of the team launches + solver-level timing with team on / off.
set -x
timeout 900 python -m pytest tests/test_gpu_team.py -m gpu -x -q --timeout=120 > gpurun_out/gputest_team.log 2>&1; echo "rc=$?" >> gpurun_out/gputest_team.log; tail -15 gpurun_out/gputest_team.log
for t in off on; do
timeout 300 python - <<PY 2>&1 | grep -v Warning
import torch, json
from heat2d_b200 import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver
for arith in ("strict-fma",):
    for n, steps in ((32768, 20), (32768, 200), (16384, 200)):
        s = Heat2DSolver(HeatConfig(nx=n, ny=n, steps=steps, k=8, arith=arith, team="$t"), "cuda:0").initialize()
        s.run(16)
        best = min(s.run(steps).elapsed_s for _ in range(3))
        print(json.dumps({"team": "$t", "arith": arith, "n": n, "steps": steps, "glups": round(n * n * steps / best / 1e9, 1)}), flush=True)
        s.close(); del s; torch.cuda.empty_cache()
PY
done
