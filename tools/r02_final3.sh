# round 2 closing evidence, one GPU, final build (recalibrated window planner): smoke, both bench arms, launch list, ncu of the accumulation
mkdir -p gpurun_out /tmp/p
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 500 python bench.py > gpurun_out/bench_final3.json 2> gpurun_out/bench_final3.err; echo "bench rc=$?"; tail -c 300 gpurun_out/bench_final3.err
timeout 500 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_final3.json 2> gpurun_out/bench_ref_final3.err; echo "ref rc=$?"
timeout 500 python bench.py --workload groth16 --logn 20 --no-cpu > gpurun_out/bench_g16_20_final3.json 2> gpurun_out/bench_g16_20_final3.err; echo "g16 rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_final3.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-verify > gpurun_out/ncu_launches_final3.log 2>&1; echo "launches rc=$?"
timeout 500 ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:"k_accumulate|k_running_sum" --launch-skip 5 -c 7 -o /tmp/p/msm python bench.py --steps 1 --warmup 3 --no-cpu --no-verify > gpurun_out/ncu_msm_final3.log 2>&1; echo "ncu msm rc=$?"
ncu -i /tmp/p/msm.ncu-rep --page raw --csv > gpurun_out/msm_raw_final3.csv
python -c "
import json;d=json.load(open('gpurun_out/bench_final3.json'));print(d['value'],d['e2e'],d['stage_ms'],d['create_ms'],d['cpu_baseline']['value'],d['roofline']['frac'])"
cut -c1-400 gpurun_out/bench_ref_final3.json
python -c "
import json;d=json.load(open('gpurun_out/bench_g16_20_final3.json'));print(d['value'],d['e2e']['value'],d['stage_ms'])"
