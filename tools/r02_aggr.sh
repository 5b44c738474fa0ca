# warp-aggregated atomics in the digit sorts (measured slower and reverted: DESIGN.md section 4): parity tests, the bench, and the windows whose sorts used to double in time
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_msm.py tests/test_gpu_prover.py tests/test_gpu_sharded.py tests/test_gpu_fullsize.py -x -q -m gpu > gpurun_out/pytest_gpu9.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu9.log
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "$name rc=$?"
  python -c "
import json;d=json.load(open('gpurun_out/bench_$name.json'));print(d['value'],d['e2e']['value'],d['stage_ms'])"
}
run aggr_default X=1
run aggr_c19 UGB_MSM_PRE_C=19
run aggr_c22 UGB_MSM_PRE_C=22
