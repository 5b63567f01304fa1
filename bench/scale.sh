#!/bin/sh
# Scaling run on one box: bench.py at N = 1, 2, 4, 8 (ours), device-timed, max over ranks.
# Usage: bench/scale.sh [steps] [warmup]   -> gpurun_out/scale_n{N}.json
STEPS=${1:-20000}; WARM=${2:-200}
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for N in 1 2 4 8; do
  [ "$N" -gt "$NG" ] && break
  if [ "$N" -eq 1 ]; then
    timeout 300 python bench.py --gpus 1 --steps $STEPS --warmup $WARM 2>gpurun_out/scale_n$N.err | tail -1 > gpurun_out/scale_n$N.json
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $((29600 + N)) bench.py --gpus $N --steps $STEPS --warmup $WARM 2>gpurun_out/scale_n$N.err | tail -1 > gpurun_out/scale_n$N.json
  fi
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/scale_n$N.json"))
    print("N=$N value", d["value"], "img/s  ms/step", d["ms_per_step"], " e2e", d["e2e"]["value"], " clocks", d["clocks"].get("sm_mhz"), d["clocks"].get("reasons"), "timeouts", d["peer_timeouts"])
except Exception as e:
    print("N=$N failed:", e); print(open("gpurun_out/scale_n$N.err").read()[-1500:])
PY
done
