# final evidence pass: tests, bench (both arms), launch list and ncu --set full summaries of the timed proofs
mkdir -p gpurun_out /tmp/p
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-verify > gpurun_out/ncu_launches.log 2>&1; echo "launches rc=$?"
timeout 400 ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:"k_accumulate|k_running_sum|k_bit_sums|k_sum_points" --launch-skip 6 -c 12 -o /tmp/p/msm python bench.py --steps 1 --warmup 3 --no-cpu --no-verify > gpurun_out/ncu_msm.log 2>&1
ncu -i /tmp/p/msm.ncu-rep --page raw --csv > gpurun_out/msm_raw.csv
timeout 400 ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:"k_ntt_pass|k_digits|k_h_final|k_spmv_abc|k_seg_reduce" -c 10 -o /tmp/p/ntt python bench.py --steps 1 --warmup 3 --no-cpu --no-verify > gpurun_out/ncu_ntt.log 2>&1
ncu -i /tmp/p/ntt.ncu-rep --page raw --csv > gpurun_out/ntt_raw.csv
tail -c 300 gpurun_out/bench_full.err; cut -c1-200 gpurun_out/bench_full.json
