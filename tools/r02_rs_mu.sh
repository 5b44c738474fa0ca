# chunk size of the running-sum level (2^mu buckets per thread; the bit sums shrink with it) against the window size, on a 1/8 shard
# and on the single-GPU proof
mkdir -p gpurun_out
for c in 0 17 16; do
  for mu in 3 4 5 6; do
    if [ $c = 0 ]; then unset UGB_MSM_PRE_C; else export UGB_MSM_PRE_C=$c; fi
    UGB_RS_MU=$mu timeout 300 python tools/shard_phase_times.py --world 8 --logn 22 --reps 3 > gpurun_out/shard_c${c}_mu$mu.txt 2>&1
    echo "c=$c mu=$mu rc=$? $(grep 'shard 0/' gpurun_out/shard_c${c}_mu$mu.txt)"
  done
done
unset UGB_MSM_PRE_C
for mu in 6 7; do
  UGB_RS_MU=$mu timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_mu$mu.json 2> gpurun_out/bench_mu$mu.err; echo "mu=$mu rc=$?"
  python -c "
import json;d=json.load(open('gpurun_out/bench_mu$mu.json'));print(d['value'],d['e2e']['value'],d['stage_ms'])"
done
