#!/bin/sh
# bench.py at the given world sizes only: bench/scale_some.sh "4 8" [steps] [warmup] -> gpurun_out/scale_n{N}.json
STEPS=${2:-10000}; WARM=${3:-200}
mkdir -p gpurun_out
for N in $1; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port $((29600 + N)) bench.py --gpus $N --steps $STEPS --warmup $WARM 2>gpurun_out/scale_n$N.err | tail -1 > gpurun_out/scale_n$N.json
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/scale_n$N.json"))
    print("N=$N value", d["value"], "img/s  ms/step", d["ms_per_step"], " e2e", d["e2e"]["value"], " clocks", d["clocks"].get("sm_mhz"), d["clocks"].get("reasons"), "timeouts", d["peer_timeouts"], d["config"].get("exchange"))
except Exception as e:
    print("N=$N failed:", e); print(open("gpurun_out/scale_n$N.err").read()[-1500:])
PY
done
