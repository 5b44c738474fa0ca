# is the 147 B of DRAM traffic per accumulated entry (68 B algorithmic) the 128-byte L2 line fill of a 64-byte gather?
mkdir -p gpurun_out
for g in 0 32; do
  if [ $g = 0 ]; then unset UGB_L2_FETCH; else export UGB_L2_FETCH=$g; fi
  timeout 100 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --profile-from-start off -k regex:"k_accumulate" -c 4 --csv --log-file gpurun_out/l2fetch_$g.csv python bench.py --workload groth16 --logn 20 --steps 1 --warmup 3 --no-cpu --no-verify > gpurun_out/l2fetch_$g.log 2>&1; echo "g=$g rc=$?"
  grep -c k_accumulate gpurun_out/l2fetch_$g.csv
done
