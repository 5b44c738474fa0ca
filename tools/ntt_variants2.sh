# second NTT sweep: pass count from 2^11-element tiles, then narrower strided tiles (UGB_NTT_MAX_LOGC) and fewer threads
run() { name=$1; shift; echo "== $name"; env "$@" timeout 120 python sweep.py --msm "" --ntt 20,22,24,26 2>&1 | grep ntt_fr | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('  2^%d ntt %.3f ms  coset3 %.3f ms' % (j['log2n'], j['ms_median'], j['coset_transform_x3_ms']))"; }
run default X=1
for c in 1 2 3 4; do for t in 128 256; do run c${c}_t$t UGB_NTT_MAX_LOGC=$c UGB_NTT_THREADS=$t; done; done
run loge9_t128 UGB_NTT_LOGE=9 UGB_NTT_THREADS=128
