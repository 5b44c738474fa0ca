"""Runs the modular-product microbenchmarks once each (for an ncu capture of their pipe utilisation): kinds 2 (CIOS fp_mul),
7 (FP64-assisted fp_mul_dfma), 6 (pure DFMA)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ultragroth_b200 as ug

ctx = ug.Context(0)
for k in (2, 7, 6):
    print(k, ctx.bench_int_pipe(k))
