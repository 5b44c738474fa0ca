mkdir -p gpurun_out /tmp/p
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-verify > gpurun_out/ncu_launches.log 2>&1; echo "launches rc=$?"
timeout 400 ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:"k_ntt_pass|k_digits_scatter|k_h_final|k_spmv_abc" -c 6 -o /tmp/p/ntt python bench.py --steps 1 --warmup 3 --no-cpu --no-verify > gpurun_out/ncu_ntt.log 2>&1
ncu -i /tmp/p/ntt.ncu-rep --page raw --csv > gpurun_out/ntt_raw.csv
timeout 300 python sweep.py --msm 16,18,20,22,24 --ntt 16,18,20,22,24 > gpurun_out/sweep_g1.jsonl 2> gpurun_out/sweep.err
timeout 300 python sweep.py --msm 16,18,20,22,24 --ntt "" --dist witness > gpurun_out/sweep_g1_witness.jsonl 2>> gpurun_out/sweep.err
timeout 200 python sweep.py --msm 16,18,20,22 --ntt "" --g2 > gpurun_out/sweep_g2.jsonl 2>> gpurun_out/sweep.err
tail -c 300 gpurun_out/bench_full.err; cut -c1-200 gpurun_out/bench_full.json
