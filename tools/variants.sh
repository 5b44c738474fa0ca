mkdir -p gpurun_out
run() { # name, env...
  name=$1; shift
  env "$@" timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/v_$name.json 2> gpurun_out/v_$name.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/v_$name.json")); s=d["stage_ms"]; k=d["kernel_share"]
    print("$name", "verified", d["verified"], "total=%.2f a=%.2f b2=%.2f h=%.2f" % (s["total"], s["msm_a"], s["msm_b2"], s["msm_h"]), {x:round(v*s["total"],1) for x,v in k.items() if "sum" in x})
except Exception as e:
    print("$name FAILED", e, open("gpurun_out/v_$name.err").read()[-300:])
PY
}
run mu5 UGB_RS_MU=5
run mu4 UGB_RS_MU=4
run mu3 UGB_RS_MU=3
run mu6 UGB_RS_MU=6
