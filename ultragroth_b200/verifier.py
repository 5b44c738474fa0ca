"""Host-side mirror of the reference's verifier C API (src/verifier.h) over libultragroth_b200.so: same names, argument
meaning and return codes.  Runs on the CPU (O(1) pairings); needs no GPU."""
import ctypes

from ._lib import lib

VERIFIER_VALID_PROOF, VERIFIER_INVALID_PROOF, VERIFIER_ERROR = 0, 1, 2
ERR_SIZE = 512


class VerifierError(RuntimeError):
    pass


def _call(name, proof, inputs, verification_key):
    fn = getattr(lib(), name)
    fn.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_ulong]
    fn.restype = ctypes.c_int
    enc = lambda s: s if isinstance(s, bytes) else str(s).encode()
    err = ctypes.create_string_buffer(ERR_SIZE)
    rc = fn(enc(proof), enc(inputs), enc(verification_key), err, ERR_SIZE - 1)
    return rc, err.value.decode(errors="replace")


def groth16_verify(proof, inputs, verification_key):
    """-> (code, error message); code 0 valid, 1 invalid, 2 error (src/verifier.h:8-11)."""
    return _call("groth16_verify", proof, inputs, verification_key)


def ultra_groth_verify(proof, inputs, verification_key):
    return _call("ultra_groth_verify", proof, inputs, verification_key)


def verify(protocol, proof, inputs, verification_key) -> bool:
    rc, msg = (ultra_groth_verify if protocol in ("ultragroth", "ultra_groth") else groth16_verify)(proof, inputs, verification_key)
    if rc == VERIFIER_ERROR:
        raise VerifierError(msg)
    return rc == VERIFIER_VALID_PROOF
