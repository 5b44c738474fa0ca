# where does the 9.9 ms witness-MSM phase of a 1/8 shard go, and is the planner's window the best one at that size?
mkdir -p gpurun_out
timeout 400 python tools/shard_phase_times.py --world 8 --logn 22 --kernels --reps 3 > gpurun_out/shard_kernels_w8.txt 2>&1; echo "rc=$?"; tail -32 gpurun_out/shard_kernels_w8.txt
for c in 17 18 20 21; do
  UGB_MSM_PRE_C=$c timeout 400 python tools/shard_phase_times.py --world 8 --logn 22 --reps 3 > gpurun_out/shard_c$c.txt 2>&1; echo "c=$c rc=$?"; grep "shard [07]/" gpurun_out/shard_c$c.txt
done
