# launch list of the timed proofs + ncu --set full of the NTT passes (after the tile-shape change)
mkdir -p gpurun_out /tmp/p
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-verify > gpurun_out/pre.json 2> gpurun_out/pre.err; echo "bench rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_v4.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-verify > gpurun_out/ncu_launches.log 2>&1; echo "launches rc=$?"
timeout 400 ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:"k_ntt_pass" -c 6 -o /tmp/p/ntt python bench.py --steps 1 --warmup 3 --no-cpu --no-verify > gpurun_out/ncu_ntt.log 2>&1; echo "ntt rc=$?"
ncu -i /tmp/p/ntt.ncu-rep --page raw --csv > gpurun_out/ntt_raw_v4.csv
ls -la gpurun_out | tail -8
