"""One fused coset transform of 3 x 2^22 (twice) for an ncu capture; `tma` as argv[1] selects the TMA bulk-copy contiguous pass."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "tma":
    os.environ["UGB_NTT_TMA"] = "1"
    os.environ["UGB_NTT_TMA_THREADS"] = "256"
import numpy as np
import ultragroth_b200 as ug

ctx = ug.Context(0)
rng = np.random.default_rng(1)
v = rng.integers(0, 256, size=(1 << 22, 32), dtype=np.uint8)
v[:, 31] &= 0x1F
ds = [ctx.upload(v) for _ in range(3)]
for _ in range(2):
    ctx.coset_transform(ds, 22)
ctx.sync()
