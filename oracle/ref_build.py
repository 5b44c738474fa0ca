"""Compiles the parts of the reference that DO build here, from the sources where they lie under
/root/reference, into oracle/_ref/libref_field.so:
    build/fr_raw_generic.cpp, build/fq_raw_generic.cpp   (the generated portable field code)
    src/keccak256.cpp                                     (Keccak-256)
The whole prover cannot be built (depends/ffiasm is an empty directory, nasm and gmp.h are absent).  The
field code needs <gmp.h> only for mpn_* prototypes: oracle/ref_shim/gmp.h (ours) declares the handful it
uses against the runtime libgmp.so.10.  No reference source is copied into the repo; outputs go to
oracle/_ref/ only (git-ignored, but shipped to the GPU box).  Test infrastructure."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
OUT = os.path.join(HERE, "_ref")
LIB = os.path.join(OUT, "libref_field.so")
GMP = "/usr/lib/x86_64-linux-gnu/libgmp.so.10"


def available():
    return os.path.exists(LIB)


def build(force=False):
    if not os.path.isdir(os.path.join(REF, "build")):
        return LIB if os.path.exists(LIB) else None  # GPU box: only the prebuilt file travels
    if os.path.exists(LIB) and not force:
        return LIB
    os.makedirs(OUT, exist_ok=True)
    srcs = [os.path.join(REF, "build", "fr_raw_generic.cpp"), os.path.join(REF, "build", "fq_raw_generic.cpp"),
            os.path.join(REF, "src", "keccak256.cpp"), os.path.join(HERE, "ref_wrap.cpp")]
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w", "-I", os.path.join(HERE, "ref_shim"), "-I", os.path.join(REF, "build"),
           "-I", os.path.join(REF, "src"), "-o", LIB] + srcs + [GMP]
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force=True))
