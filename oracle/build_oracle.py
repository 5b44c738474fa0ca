"""Builds the oracle's C++ restatement (oracle/cpu_prover.cpp + cpu_synth.cpp -> oracle/_build/libcpu_prover.so) and, when
/root/reference is present, the reference's own compilable pieces into oracle/_ref/ (see oracle/ref_build.py).
Test infrastructure only."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUT_DIR, "libcpu_prover.so")
SRC = os.path.join(HERE, "cpu_prover.cpp")
SRCS = [SRC, os.path.join(HERE, "cpu_synth.cpp")]
DEPS = SRCS + [os.path.join(HERE, "cpu_field.hpp")]


def build(force=False):
    os.makedirs(OUT_DIR, exist_ok=True)
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < max(os.path.getmtime(f) for f in DEPS):
        tmp = LIB + ".%d.tmp" % os.getpid()  # several ranks may race here: build aside, then rename atomically
        subprocess.check_call(["g++", "-O3", "-std=c++17", "-fopenmp", "-fPIC", "-shared", "-o", tmp] + SRCS)
        os.replace(tmp, LIB)
    try:
        from . import ref_build
        ref_build.build()
    except ImportError:
        pass
    return LIB


if __name__ == "__main__":
    print(build(force=True))
