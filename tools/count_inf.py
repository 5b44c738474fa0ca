"""How many zkey base points of the synthetic circuits are the point at infinity (all-zero encoding)?  Infinity entries cost
a bucket-accumulation slot each without contributing; a large share would justify filtering them at digit-sort time."""
import sys
import numpy as np
sys.path.insert(0, ".")
import ultragroth_b200 as ug
from oracle.formats import ZKey
ctx = ug.Context(0)
for ultra in (True, False):
    zk, wt, vkj = ctx.synth(20, seed=1, ultra=ultra, lookup_bits=14, n_chunks=1 << 18)
    z = ZKey(zk.tobytes())
    print("ultragroth" if ultra else "groth16", "nVars", z.nVars)
    secs = ((5, 64, "A"), (6, 64, "B1"), (7, 128, "B2"), (8, 64, "C/C1"), (9, 64, "H/C2")) + (((12, 64, "H"),) if ultra else ())
    for sec, size, name in secs:
        a = np.frombuffer(z.bin.section(sec), dtype=np.uint8)
        a = a[: (len(a) // size) * size].reshape(-1, size)
        zero = int((~a.any(axis=1)).sum())
        print("  %-5s %9d points, %9d at infinity (%.1f%%)" % (name, a.shape[0], zero, 100.0 * zero / max(1, a.shape[0])))
