#!/bin/bash
# what the seam guard and the longer warm-up cost per launch (resident leg only)
cd /root/repo; mkdir -p gpurun_out/abw
B="python bench.py --no-cpu-baseline --no-fast-mode --polish-synth 0 --no-sustained --steps 12"
$B > gpurun_out/abw/default.json 2>/dev/null
MARGIN_B200_WARMUP=128 $B > gpurun_out/abw/w128.json 2>/dev/null
MARGIN_B200_WARMUP=128 MARGIN_B200_GUARD=0 $B > gpurun_out/abw/w128_noguard.json 2>/dev/null
MARGIN_B200_GUARD=0 $B > gpurun_out/abw/w192_noguard.json 2>/dev/null
python - <<'PY'
import json
for n in ("default", "w128", "w128_noguard", "w192_noguard"):
    d = json.load(open("gpurun_out/abw/%s.json" % n))
    k = d["kernel_ms_per_step"]
    print("%-14s value %.2f  ms/step %.3f  fwd %.3f bwd %.3f guard %.3f tiles %.0f fail %s" % (n, d["value"], d["ms_per_step"], k["forward"], k["backward"], k["guard"], d["config"]["tiles_per_step"], d["config"]["guard_failures"]))
PY
