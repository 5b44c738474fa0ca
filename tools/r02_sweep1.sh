mkdir -p gpurun_out
timeout 1500 python sweep.py --msm 16,18,20,22,24,26 --ntt 16,18,20,22,24,26 --cpu --cpu-max 24 --reps 5 > gpurun_out/sweep_n1.jsonl 2> gpurun_out/sweep_n1.err; echo "sweep rc=$?"; tail -c 500 gpurun_out/sweep_n1.err
cut -c1-330 gpurun_out/sweep_n1.jsonl
timeout 900 python bench.py --steps 3 --warmup 3 --workload groth16 --logn 24 --no-cpu > gpurun_out/bench_g16_24_n1.json 2> gpurun_out/bench_g16_24_n1.err; echo "bench rc=$?"; tail -c 500 gpurun_out/bench_g16_24_n1.err
python -c "
import json;d=json.load(open('gpurun_out/bench_g16_24_n1.json'));print(d['value'],d['e2e'],d['stage_ms'],d['create_ms'],d['setup_s'])"
