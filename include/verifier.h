/* The reference's verifier C API (src/verifier.h:8-38), same names, argument meaning and return codes; implemented on
 * the host in ultragroth_b200/csrc/verifier_api.cpp (O(1) pairings per proof -- no GPU work, no CUDA dependency).
 * 'proof', 'inputs' and 'verification_key' are NUL-terminated JSON strings as snarkjs / the prover emit them. */
#ifndef UGB_VERIFIER_H
#define UGB_VERIFIER_H

#ifdef __cplusplus
extern "C" {
#endif

#define VERIFIER_VALID_PROOF 0x0
#define VERIFIER_INVALID_PROOF 0x1
#define VERIFIER_ERROR 0x2

/* replaces groth16_verify, src/verifier.cpp:149-196 (Groth16::Verifier::verify, src/groth16.cpp:314-364) */
int groth16_verify(const char* proof, const char* inputs, const char* verification_key, char* error_msg, unsigned long error_msg_maxsize);
/* replaces ultra_groth_verify, src/verifier.cpp:198-243 (UltraGroth::Verifier::verify, src/ultra_groth.cpp:581-648) */
int ultra_groth_verify(const char* proof, const char* inputs, const char* verification_key, char* error_msg, unsigned long error_msg_maxsize);

#ifdef __cplusplus
}
#endif
#endif
