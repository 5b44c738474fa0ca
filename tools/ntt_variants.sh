# NTT tile / thread-count sweep (env knobs read by ntt.cu's pass planner); prints median ms per size.
run() { name=$1; shift; echo "== $name"; env "$@" timeout 120 python sweep.py --msm "" --ntt 20,22,24 2>&1 | grep ntt_fr | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('  2^%d ntt %.3f ms  coset3 %.3f ms' % (j['log2n'], j['ms_median'], j['coset_transform_x3_ms']))"; }
run default X=1
run t128 UGB_NTT_THREADS=128
run t384 UGB_NTT_THREADS=384
run t512 UGB_NTT_THREADS=512
run loge12_t512 UGB_NTT_LOGE=12 UGB_NTT_THREADS=512
