mkdir -p gpurun_out
run() { # name, env...
  name=$1; shift
  env "$@" timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu --no-verify > gpurun_out/v_$name.json 2> gpurun_out/v_$name.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/v_$name.json")); s=d["stage_ms"]; k=d["kernel_share"]
    print("$name", "total=%.1f a=%.2f b2=%.2f h=%.2f" % (s["total"], s["msm_a"], s["msm_b2"], s["msm_h"]), {x:round(v*s["total"],1) for x,v in k.items() if "acc" in x or "sum" in x})
except Exception as e:
    print("$name FAILED", e, open("gpurun_out/v_$name.err").read()[-300:])
PY
}
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run base X=1
run rsg2_0 UGB_RS_G2=0
run rsg2_13 UGB_RS_G2=13
run rsg1_13 UGB_RS_G1=13
run rsg1_3 UGB_RS_G1=3
run rsg1_4 UGB_RS_G1=4
