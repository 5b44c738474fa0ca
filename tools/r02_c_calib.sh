# calibration of the window planner for resident tables: per-MSM stage times of the single-GPU proof at c = 20, 21 (22 is the default),
# and the witness-MSM phase of 1/4 and 1/2 shards over c
mkdir -p gpurun_out
for c in 21 20; do
  UGB_MSM_PRE_C=$c timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_c$c.json 2> gpurun_out/bench_c$c.err; echo "c=$c rc=$?"
  python -c "
import json;d=json.load(open('gpurun_out/bench_c$c.json'));print(d['value'],d['e2e']['value'],d['stage_ms'])"
done
for c in 17 18 19 20; do
  UGB_MSM_PRE_C=$c timeout 300 python tools/shard_phase_times.py --world 4 --logn 22 --reps 3 > gpurun_out/shard4_c$c.txt 2>&1
  echo "world=4 c=$c rc=$? $(grep 'shard 0/' gpurun_out/shard4_c$c.txt)"
done
for c in 18 19 20 21; do
  UGB_MSM_PRE_C=$c timeout 300 python tools/shard_phase_times.py --world 2 --logn 22 --reps 3 > gpurun_out/shard2_c$c.txt 2>&1
  echo "world=2 c=$c rc=$? $(grep 'shard 0/' gpurun_out/shard2_c$c.txt)"
done
