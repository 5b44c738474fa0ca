mkdir -p gpurun_out
run() { # name, env...
  name=$1; shift
  env "$@" timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu --no-verify > gpurun_out/v_$name.json 2> gpurun_out/v_$name.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/v_$name.json")); s=d["stage_ms"]; k=d["kernel_share"]
    print("$name", "total=%.1f a=%.2f b1=%.2f c=%.2f b2=%.2f h=%.2f" % (s["total"], s["msm_a"], s["msm_b1"], s["msm_c"], s["msm_b2"], s["msm_h"]), {x:round(v*s["total"],1) for x,v in k.items() if "acc" in x or "sum" in x})
except Exception as e:
    print("$name FAILED", e, open("gpurun_out/v_$name.err").read()[-300:])
PY
}
run batch UGB_MSM_BATCH=1
run nobatch UGB_MSM_BATCH=0
