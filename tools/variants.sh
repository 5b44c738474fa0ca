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
UGB_ACC_G1=23 UGB_ACC_G2=32 timeout 600 python -m pytest tests/test_gpu_msm.py tests/test_gpu_prover.py tests/test_gpu_synth.py -m gpu -x -q 2>&1 | tail -4
run ba_g1_22 UGB_ACC_G1=22
run ba_g1_23 UGB_ACC_G1=23
run ba_g1_24 UGB_ACC_G1=24
run ba_g1_33 UGB_ACC_G1=33
run ba_g2_32 UGB_ACC_G2=32
run ba_g2_33 UGB_ACC_G2=33
