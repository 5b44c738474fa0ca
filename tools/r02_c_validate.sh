# the recalibrated planner (resident tables: 54 per bucket, smallest c per W; chunk rule by bucket count) against its neighbours
mkdir -p gpurun_out
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu $BARGS > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "$name rc=$?"
  python -c "
import json;d=json.load(open('gpurun_out/bench_$name.json'));print(d['value'],d['e2e']['value'],d['stage_ms'])"
}
BARGS=""
run v_default X=1
run v_mu3 UGB_RS_MU=3
run v_c19 UGB_MSM_PRE_C=19
BARGS="--workload groth16 --logn 20"
run v_g16_default X=1
run v_g16_c20 UGB_MSM_PRE_C=20
run v_g16_c19 UGB_MSM_PRE_C=19
for w in 8 4 2; do
  timeout 300 python tools/shard_phase_times.py --world $w --logn 22 --reps 3 > gpurun_out/shard_v_w$w.txt 2>&1
  echo "world=$w rc=$? $(grep 'shard 0/' gpurun_out/shard_v_w$w.txt)"
done
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu8.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu8.log
