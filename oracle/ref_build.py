"""Compiles the parts of the reference that DO build here, from the sources where they lie under
/root/reference, into oracle/_ref/libref_field.so:
    build/fr_raw_generic.cpp, build/fq_raw_generic.cpp   (the generated portable field code)
    src/keccak256.cpp                                     (Keccak-256)
and, into oracle/_ref/libref_host.so, the reference's host code that needs no arithmetic engine:
    src/binfile_utils.cpp, fileloader.cpp, zkey_utils.cpp, wtns_utils.cpp   (file loaders)
    build/fr.cpp, fr_generic.cpp, fq.cpp, fq_generic.cpp                    (RawFr / RawFq: set(int), inv, fromMpz, toMpz)
The whole prover cannot be built (depends/ffiasm is an empty directory, nasm and gmp.h are absent).  The
field code needs <gmp.h> only for mpn_* / mpz_* prototypes: oracle/ref_shim/gmp.h (ours) declares the handful it
uses against the runtime libgmp.so.10.  No reference source is copied into the repo; outputs go to
oracle/_ref/ only (git-ignored, but shipped to the GPU box).  Test infrastructure."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
OUT = os.path.join(HERE, "_ref")
LIB = os.path.join(OUT, "libref_field.so")
# the reference's host code: binfile / zkey / wtns loaders, RawFr / RawFq (mpz-based set, inv, fromMpz, toMpz), keccak -- see
# oracle/ref_wrap_host.cpp for what each is used to pin
LIB_HOST = os.path.join(OUT, "libref_host.so")
GMP = "/usr/lib/x86_64-linux-gnu/libgmp.so.10"
HOST_SRCS = ["build/fr.cpp", "build/fr_generic.cpp", "build/fr_raw_generic.cpp", "build/fq.cpp", "build/fq_generic.cpp", "build/fq_raw_generic.cpp",
             "src/binfile_utils.cpp", "src/fileloader.cpp", "src/zkey_utils.cpp", "src/wtns_utils.cpp", "src/keccak256.cpp"]


def available():
    return os.path.exists(LIB)


def host_available():
    return os.path.exists(LIB_HOST)


def build_host(force=False):
    """oracle/_ref/libref_host.so: the reference's loaders + RawFr/RawFq + keccak, compiled in place (needs /root/reference)."""
    if not os.path.isdir(os.path.join(REF, "build")):
        return LIB_HOST if os.path.exists(LIB_HOST) else None
    if os.path.exists(LIB_HOST) and not force and os.path.getmtime(LIB_HOST) >= os.path.getmtime(os.path.join(HERE, "ref_wrap_host.cpp")):
        return LIB_HOST
    os.makedirs(OUT, exist_ok=True)
    srcs = [os.path.join(REF, f) for f in HOST_SRCS] + [os.path.join(HERE, "ref_wrap_host.cpp")]
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w", "-I", os.path.join(HERE, "ref_shim"), "-I", os.path.join(REF, "build"),
           "-I", os.path.join(REF, "src"), "-I", os.path.join(REF, "depends", "json", "single_include"), "-o", LIB_HOST] + srcs + [GMP]
    subprocess.check_call(cmd)
    return LIB_HOST


BIN = os.path.join(OUT, "bin")
REPO = os.path.dirname(HERE)
PRODUCT_LIB_DIR = os.path.join(REPO, "ultragroth_b200")


def dropin_available():
    return all(os.path.exists(os.path.join(BIN, n)) for n in ("prover", "prover_ultra_groth", "verifier", "verifier_ultra_groth", "class_api_prover"))


def build_dropin(force=False):
    """The drop-in demonstration: the reference's CLI mains compiled UNMODIFIED, against the reference's OWN
    prover.h / verifier.h, and linked to libultragroth_b200.so instead of the reference's library:
        src/main_prover.cpp, src/main_prover_ultra_groth.cpp, src/main_verifier.cpp, src/main_verifier_ultra_groth.cpp + src/fileloader.cpp
    plus tests/cpp/class_api_prover.cpp (ours), which drives include/groth16.hpp / include/ultra_groth.hpp through the reference's own
    zkey / wtns loaders.  Outputs: oracle/_ref/bin/ (shipped to the GPU box; rpath points at ../../../ultragroth_b200)."""
    if not os.path.isdir(os.path.join(REF, "src")):
        return BIN if dropin_available() else None
    lib = os.path.join(PRODUCT_LIB_DIR, "libultragroth_b200.so")
    if not os.path.exists(lib):
        return None
    if dropin_available() and not force and os.path.getmtime(os.path.join(BIN, "class_api_prover")) >= max(
            os.path.getmtime(lib), os.path.getmtime(os.path.join(REPO, "tests", "cpp", "class_api_prover.cpp"))):
        return BIN
    os.makedirs(BIN, exist_ok=True)
    link = ["-L", PRODUCT_LIB_DIR, "-lultragroth_b200", "-Wl,-rpath,$ORIGIN/../../../ultragroth_b200"]
    inc_ref = ["-I", os.path.join(REF, "src")]
    src = lambda n: os.path.join(REF, "src", n)
    for out, main in (("prover", "main_prover.cpp"), ("prover_ultra_groth", "main_prover_ultra_groth.cpp"), ("verifier", "main_verifier.cpp"),
                      ("verifier_ultra_groth", "main_verifier_ultra_groth.cpp")):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-w"] + inc_ref + ["-o", os.path.join(BIN, out), src(main), src("fileloader.cpp")] + link)
    subprocess.check_call(
        ["g++", "-O2", "-std=c++17", "-w", "-I", os.path.join(REPO, "include"), "-I", os.path.join(HERE, "ref_shim")] + inc_ref +
        ["-I", os.path.join(REF, "depends", "json", "single_include"), "-o", os.path.join(BIN, "class_api_prover"),
         os.path.join(REPO, "tests", "cpp", "class_api_prover.cpp"), src("binfile_utils.cpp"), src("fileloader.cpp"), src("zkey_utils.cpp"),
         src("wtns_utils.cpp")] + link + [GMP])
    return BIN


def build(force=False):
    if not os.path.isdir(os.path.join(REF, "build")):
        return LIB if os.path.exists(LIB) else None  # GPU box: only the prebuilt file travels
    if os.path.exists(LIB) and not force:
        build_host()
        build_dropin()
        return LIB
    os.makedirs(OUT, exist_ok=True)
    srcs = [os.path.join(REF, "build", "fr_raw_generic.cpp"), os.path.join(REF, "build", "fq_raw_generic.cpp"),
            os.path.join(REF, "src", "keccak256.cpp"), os.path.join(HERE, "ref_wrap.cpp")]
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w", "-I", os.path.join(HERE, "ref_shim"), "-I", os.path.join(REF, "build"),
           "-I", os.path.join(REF, "src"), "-o", LIB] + srcs + [GMP]
    subprocess.check_call(cmd)
    build_host(force)
    build_dropin(force)
    return LIB


if __name__ == "__main__":
    print(build(force=True))
