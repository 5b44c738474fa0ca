"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/*.h declares, mirrors the
reference's validation order and messages for everything that happens before a device is touched, and fails
loudly (no CPU fallback) when there is no GPU."""
import ctypes
import os
import re

import pytest

import ultragroth_b200 as ug
from ultragroth_b200 import prover as up

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "ref_testdata")


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b((?:ugb|groth16|ultra_groth)_\w+)\s*\(", text)))


@pytest.mark.parametrize("header", ["ultragroth_b200.h", "prover.h", "verifier.h"])
def test_library_exports_every_declared_symbol(header):
    L = ug.lib()
    names = _declared(header)
    assert len(names) >= {"prover.h": 16, "verifier.h": 2}.get(header, 20)
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/{header} but not exported"


def test_prover_h_has_the_references_16_entry_points():
    want = []
    for pre in ("groth16", "ultra_groth"):
        want += [pre + s for s in ("_public_size_for_zkey_buf", "_public_size_for_zkey_file", "_proof_size", "_prover_create",
                                   "_prover_create_zkey_file", "_prover_prove", "_prover_destroy", "_prover", "_prover_zkey_file")]
    assert sorted(want) == _declared("prover.h")


def test_sizes_and_header_parsing_without_gpu():
    zk = open(os.path.join(GOLD, "circuit_final.zkey"), "rb").read()
    assert up.proof_size("groth16") == 810 and up.proof_size("ultragroth") == 1400
    assert up.public_size_for_zkey_buf(zk) == 86
    v = ctypes.c_ulonglong()
    e = ctypes.create_string_buffer(256)
    rc = ug.lib().groth16_public_size_for_zkey_file(os.path.join(GOLD, "circuit_final.zkey").encode(), ctypes.byref(v), e, 255)
    assert rc == 0 and v.value == 86
    with pytest.raises(up.ProverError) as ei:
        up.public_size_for_zkey_buf(zk, "ultragroth")
    assert ei.value.code == up.PROVER_ERROR and ei.value.msg == "zkey file is not ultragroth"
    with pytest.raises(up.ProverError) as ei:
        up.public_size_for_zkey_buf(b"wtns" + zk[4:])
    assert ei.value.msg == "Invalid file type. It should be zkey and it is wtns"
    with pytest.raises(up.ProverError) as ei:
        up.public_size_for_zkey_buf(zk[:4] + (2).to_bytes(4, "little") + zk[8:])
    assert ei.value.msg == "Invalid version. It should be <=1 and it is 2"
    with pytest.raises(up.ProverError) as ei:
        up.public_size_for_zkey_buf(zk[:8])
    assert ei.value.msg == "File is too short."
    with pytest.raises(up.ProverError) as ei:
        up.public_size_for_zkey_buf(zk[:1000])
    assert "is invalid" in ei.value.msg


def test_null_arguments_follow_the_reference_order():
    L = ug.lib()
    e = ctypes.create_string_buffer(256)
    assert L.groth16_prover_create(None, b"x", 1, e, 255) == up.PROVER_ERROR and e.value == b"Null prover object"
    h = ctypes.c_void_p()
    assert L.groth16_prover_create(ctypes.byref(h), None, 0, e, 255) == up.PROVER_ERROR and e.value == b"Null zkey buffer"
    n = ctypes.c_ulonglong(0)
    assert L.groth16_prover_prove(None, b"x", 1, e, ctypes.byref(n), e, ctypes.byref(n), e, 255) == up.PROVER_ERROR
    assert e.value == b"Null prover object"
    L.groth16_prover_destroy(None)      # NULL-safe like src/prover.cpp:661-679
    L.ultra_groth_prover_destroy(None)
    # error_msg may be NULL
    assert L.groth16_prover_create(None, b"x", 1, None, 0) == up.PROVER_ERROR


def test_no_gpu_means_loud_failure_not_a_cpu_fallback():
    if ug.lib().ugb_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(ug.UltraGrothError) as ei:
        ug.Context(0)
    assert "no CPU fallback" in str(ei.value)
    zk = open(os.path.join(GOLD, "circuit_final.zkey"), "rb").read()
    with pytest.raises(up.ProverError) as ei:
        up.Prover(zk)
    assert ei.value.code == up.PROVER_ERROR and "no CPU fallback" in ei.value.msg


def _plan(n, resident):
    c, w, mu = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    assert ug.lib().ugb_msm_plan(ctypes.c_uint64(n), 1 if resident else 0, ctypes.byref(c), ctypes.byref(w), ctypes.byref(mu)) == 0
    return c.value, w.value, mu.value


def test_msm_window_planner(monkeypatch):
    """The planner is host arithmetic (csrc/msm.cu: msm_pick_config, msm_reduce_plan), calibrated on measured reduction-tail costs
    (profiles/r02_window_calibration.txt): the windows it must choose at the sizes that were measured, and its invariants."""
    for k in ("UGB_MSM_PRE_C", "UGB_MSM_C", "UGB_RS_MU"):
        monkeypatch.delenv(k, raising=False)
    # resident tables: whole 2^22 UltraGroth prover (witness arrays 5.3 M, H 4.19 M), its 1/2 and 1/8 shards, 2^24 Groth16
    assert _plan(5308415, True)[:2] == (20, 13)
    assert _plan(1 << 22, True)[:2] == (20, 13)
    assert _plan(5308415 // 2, True)[:2] == (20, 13)
    assert _plan(663552, True)[:2] == (17, 15)
    assert _plan(1 << 19, True)[:2] == (17, 15)
    assert _plan((1 << 24) + 3, True)[:2] == (22, 12)
    prev = 0
    for logn in range(8, 27):
        c, w, mu = _plan(1 << logn, True)
        assert w == -(-255 // c) and 4 <= c <= 22
        assert c == 4 or -(-255 // (c - 1)) != w, "only the smallest window of each W is a candidate"
        assert c >= prev, "the window never shrinks as n grows"
        lb = c - 1
        assert mu == min(8, max(2, (lb - 10) // 2))
        prev = c
    # ad-hoc bases (one bucket set per window): unchanged rule, c <= 20, chunks sized for ~64 Ki threads over all windows
    for logn in (10, 16, 20, 24, 26):
        c, w, mu = _plan(1 << logn, False)
        assert 4 <= c <= 20 and w == -(-255 // c) and 2 <= mu <= 8
    # the experiment knobs are honoured
    monkeypatch.setenv("UGB_MSM_PRE_C", "22")
    monkeypatch.setenv("UGB_RS_MU", "6")
    assert _plan(1000, True) == (22, 12, 6)


def test_noalloc_msm_limits_agree_between_header_and_library():
    """ugb_msm_g1_noalloc's stated maximum (include/ultragroth_b200.h) is what the library enforces, its scratch footprint is the
    one the header states, and without a context the call is refused before anything touches a device."""
    text = open(os.path.join(ROOT, "include", "ultragroth_b200.h")).read()
    max_points = int(re.search(r"#define UGB_MSM_NOALLOC_MAX_POINTS (\d+)u", text).group(1))
    L = ug.lib()
    assert int(L.ugb_msm_noalloc_max_points()) == max_points == 16384
    m = re.search(r"#define UGB_MSM_NOALLOC_SCRATCH_BYTES \((\d+)u \* (\d+)u \* (\d+)u \+ (\d+)u\)", text)
    a, b, c, d = (int(x) for x in m.groups())
    # 32 windows x (max_points / 2048) slices x 256 buckets x 128 B, + 33 window / result points (csrc/msm_noalloc.cu)
    assert a * b * c + d == 32 * (max_points // 2048) * 256 * 128 + 33 * 128
    out = ctypes.create_string_buffer(64)
    assert L.ugb_msm_g1_noalloc(None, None, None, 1, out) != 0
    assert b"ctx" in L.ugb_last_error()
