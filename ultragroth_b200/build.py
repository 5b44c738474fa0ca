"""Builds libultragroth_b200.so in-tree with nvcc for sm_100a (no JIT cache, the .so travels with the repo).

    python -m ultragroth_b200.build [--force]
"""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# experiment builds: UGB_BUILD_TAG=x UGB_BUILD_DEFS="-DUG_FOO" -> libultragroth_b200_x.so (load it with UGB_LIB=<path>)
TAG = os.environ.get("UGB_BUILD_TAG", "")
OUT = os.path.join(HERE, "libultragroth_b200%s.so" % ("_" + TAG if TAG else ""))
OBJ = os.path.join(HERE, "build" + ("_" + TAG if TAG else ""))
EXTRA_DEFS = os.environ.get("UGB_BUILD_DEFS", "").split()

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
              "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-Wno-deprecated-gpu-targets"] + EXTRA_DEFS
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-I/usr/local/cuda/include"]


def _sources():
    cu = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    cpp = sorted(f for f in os.listdir(CSRC) if f.endswith(".cpp"))
    return cu, cpp


def _digest():
    h = hashlib.sha256()
    for f in sorted(os.listdir(CSRC)):
        with open(os.path.join(CSRC, f), "rb") as fh:
            h.update(f.encode())
            h.update(fh.read())
    inc = os.path.join(os.path.dirname(HERE), "include")
    for f in sorted(os.listdir(inc)):
        with open(os.path.join(inc, f), "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS + CXX_FLAGS).encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    stamp = os.path.join(OBJ, "stamp")
    dig = _digest()
    if not force and os.path.exists(OUT) and os.path.exists(stamp) and open(stamp).read() == dig:
        return OUT
    cu, cpp = _sources()
    objs = []
    procs = []
    for f in cu:
        o = os.path.join(OBJ, f + ".o")
        cmd = [NVCC] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, f), "-o", o]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
        objs.append(o)
    for f in cpp:
        o = os.path.join(OBJ, f + ".o")
        cmd = ["g++"] + CXX_FLAGS + ["-c", os.path.join(CSRC, f), "-o", o]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
        objs.append(o)
    for cmd, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out.decode(errors="replace"))
        if p.returncode:
            raise RuntimeError("build failed: " + " ".join(cmd))
    cmd = [NVCC, "-shared", "-Wno-deprecated-gpu-targets", "-o", OUT] + objs + ["-lcudart"]
    subprocess.check_call(cmd)
    with open(stamp, "w") as fh:
        fh.write(dig)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
