#!/bin/bash
# Second sweep: strong scaling (config 3), convergence (config 4), NCCL-halo baseline,
# weak scaling 65536^2 per GPU with k = 1/4/8 (config 5), p2p and fp32 peaks.
set -u
OUT=gpurun_out/scaling2_$(date +%H%M%S).jsonl
tr() { local n=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) "$@"; }
b() { local n=$1; shift; if [ "$n" = 1 ]; then timeout 900 python bench.py --gpus 1 "$@" 2>>gpurun_out/err2.log | tail -1 >> $OUT; else tr $n bench.py --gpus $n "$@" 2>>gpurun_out/err2.log | tail -1 >> $OUT; fi; }
./tools/microbench/fp32_pipes | head -1 | tee gpurun_out/fp32_peak.log
tr 8 tools/p2p_probe.py 2>>gpurun_out/err2.log | tail -1 | tee gpurun_out/p2p8.log
tr 2 tools/p2p_probe.py 2>>gpurun_out/err2.log | tail -1 | tee gpurun_out/p2p2.log
for n in 1 2 4 8; do b $n --steps 1000 --warmup 24; done
b 4 --steps 1000 --warmup 24 --px 4 --py 1 --no-e2e
b 4 --steps 1000 --warmup 24 --px 1 --py 4 --no-e2e
for n in 1 8; do b $n --steps 1000 --warmup 24 --conv --no-e2e; done
b 8 --steps 200 --warmup 24 --impl nccl-halo
b 8 --steps 200 --warmup 24 --impl nccl-halo --conv
for k in 1 4 8; do b 8 --steps 96 --warmup 24 --weak --grid 65536 --k $k --no-e2e; done
b 1 --steps 96 --warmup 24 --weak --grid 65536 --k 8 --no-e2e
python - <<PY
import json
for l in open("$OUT"):
    try: d=json.loads(l)
    except Exception: print("BAD", l[:200]); continue
    c=d["config"]
    print(d.get("impl"), d["n_gpus"], d["scaling"], c["grid"], c.get("parallelism","")[:9], "k", c.get("k"), "conv" if c.get("conv") else "", round(d["value"],1), "e2e", round(d["e2e"]["value"],1) if "e2e" in d else None, d["clocks"].get("sm_mhz"), d["clocks"].get("reasons"))
PY
