/* prover.h -- the reference's prover C API, kept signature-for-signature (rarimo/ultragroth
 * src/prover.h:18-232) and implemented by libultragroth_b200.so on top of sm_100a kernels.
 *
 * Every function returns one of the PROVER_* codes; on failure a message is copied into error_msg
 * (strncpy semantics, may be NULL).  zkey / wtns buffers are borrowed and read-only.  Unlike the
 * reference, the zkey buffer is only needed DURING *_prover_create: bases, coefficients and twiddles
 * are copied to HBM there and stay resident until *_prover_destroy.  Output buffers are caller
 * allocated; *_size arguments are in/out (capacity in, unchanged on success -- outputs are written
 * with strncpy, i.e. zero padded and not NUL terminated when exactly full, as in the reference).
 * A prover object must be used by one caller at a time.
 */
#ifndef UGB_PROVER_H
#define UGB_PROVER_H

#ifdef __cplusplus
extern "C" {
#endif

#define PROVER_OK                     0x0
#define PROVER_ERROR                  0x1
#define PROVER_ERROR_SHORT_BUFFER     0x2
#define PROVER_INVALID_WITNESS_LENGTH 0x3

/* minimum public-signals buffer for a zkey (nPublic * 82 + 4), src/prover.h:18-60 */
int groth16_public_size_for_zkey_buf(const void *zkey_buffer, unsigned long long zkey_size, unsigned long long *public_size,
                                     char *error_msg, unsigned long long error_msg_maxsize);
int ultra_groth_public_size_for_zkey_buf(const void *zkey_buffer, unsigned long long zkey_size, unsigned long long *public_size,
                                         char *error_msg, unsigned long long error_msg_maxsize);
int groth16_public_size_for_zkey_file(const char *zkey_fname, unsigned long long *public_size, char *error_msg,
                                      unsigned long long error_msg_maxsize);
int ultra_groth_public_size_for_zkey_file(const char *zkey_fname, unsigned long long *public_size, char *error_msg,
                                          unsigned long long error_msg_maxsize);

/* minimum proof buffer: 810 (Groth16) / 1400 (UltraGroth), src/prover.h:62-78 */
void groth16_proof_size(unsigned long long *proof_size);
void ultra_groth_proof_size(unsigned long long *proof_size);

/* resident prover: create once (uploads the zkey to HBM), prove many, destroy; src/prover.h:80-164 */
int groth16_prover_create(void **prover_object, const void *zkey_buffer, unsigned long long zkey_size, char *error_msg,
                          unsigned long long error_msg_maxsize);
int ultra_groth_prover_create(void **prover_object, const void *zkey_buffer, unsigned long long zkey_size, char *error_msg,
                              unsigned long long error_msg_maxsize);
int groth16_prover_create_zkey_file(void **prover_object, const char *zkey_file_path, char *error_msg,
                                    unsigned long long error_msg_maxsize);
int ultra_groth_prover_create_zkey_file(void **prover_object, const char *zkey_file_path, char *error_msg,
                                        unsigned long long error_msg_maxsize);
int groth16_prover_prove(void *prover_object, const void *wtns_buffer, unsigned long long wtns_size, char *proof_buffer,
                         unsigned long long *proof_size, char *public_buffer, unsigned long long *public_size, char *error_msg,
                         unsigned long long error_msg_maxsize);
int ultra_groth_prover_prove(void *prover_object, const void *wtns_buffer, unsigned long long wtns_size, char *proof_buffer,
                             unsigned long long *proof_size, char *public_buffer, unsigned long long *public_size, char *error_msg,
                             unsigned long long error_msg_maxsize);
void groth16_prover_destroy(void *prover_object);
void ultra_groth_prover_destroy(void *prover_object);

/* one-shot variants (create + prove + destroy), src/prover.h:166-232 */
int groth16_prover(const void *zkey_buffer, unsigned long long zkey_size, const void *wtns_buffer, unsigned long long wtns_size,
                   char *proof_buffer, unsigned long long *proof_size, char *public_buffer, unsigned long long *public_size,
                   char *error_msg, unsigned long long error_msg_maxsize);
int ultra_groth_prover(const void *zkey_buffer, unsigned long long zkey_size, const void *wtns_buffer, unsigned long long wtns_size,
                       char *proof_buffer, unsigned long long *proof_size, char *public_buffer, unsigned long long *public_size,
                       char *error_msg, unsigned long long error_msg_maxsize);
int groth16_prover_zkey_file(const char *zkey_file_path, const void *wtns_buffer, unsigned long long wtns_size, char *proof_buffer,
                             unsigned long long *proof_size, char *public_buffer, unsigned long long *public_size, char *error_msg,
                             unsigned long long error_msg_maxsize);
int ultra_groth_prover_zkey_file(const char *zkey_file_path, const void *wtns_buffer, unsigned long long wtns_size,
                                 char *proof_buffer, unsigned long long *proof_size, char *public_buffer,
                                 unsigned long long *public_size, char *error_msg, unsigned long long error_msg_maxsize);

#ifdef __cplusplus
}
#endif
#endif
