"""GPU (needs >= 2 devices; skipped on a one-GPU box): two contexts / provers on different devices inside ONE process.
ADVICE r1: the > 48 KiB dynamic shared memory opt-in of the NTT kernels is a per-DEVICE function attribute; it used to be set once
per process, so the second device's NTT launches failed."""
import hashlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _two_devices():
    from ultragroth_b200._lib import lib
    if lib().ugb_device_count() < 2:
        pytest.skip("needs two GPUs")


def test_ntt_on_two_devices_in_one_process():
    _two_devices()
    import ultragroth_b200 as ug
    from oracle import cpu_prover as cp
    logn = 16
    rng = np.random.default_rng(3)
    v = rng.integers(0, 256, size=(1 << logn, 32), dtype=np.uint8)
    v[:, 31] &= 0x1F
    want = hashlib.sha256(cp.ntt(v, logn)).hexdigest()
    for dev in (0, 1, 0):
        ctx = ug.Context(dev)
        d = ctx.upload(v)
        ctx.ntt(d, logn)
        assert hashlib.sha256(ctx.download(d)).hexdigest() == want, f"device {dev}"
        ctx.close()


def test_provers_on_two_devices_in_one_process(monkeypatch):
    _two_devices()
    import json
    import os
    from ultragroth_b200 import prover as up
    gold = os.path.join(os.path.dirname(__file__), "golden")
    g = json.load(open(os.path.join(gold, "testdata_proof.json")))
    zk = open(os.path.join(gold, "ref_testdata", "circuit_final.zkey"), "rb").read()
    wt = open(os.path.join(gold, "ref_testdata", "witness.wtns"), "rb").read()
    provers = []
    for dev in (0, 1):
        monkeypatch.setenv("UGB_DEVICE", str(dev))
        p = up.Prover(zk, "groth16")
        p.set_blinding(int(g["r"], 16), int(g["s"], 16))
        provers.append(p)
    for p in provers:
        pj, uj = p.prove(wt)
        assert json.loads(pj) == g["proof"] and json.loads(uj) == g["public"]
    for p in provers:
        p.close()
