/* Minimal <gmp.h> stand-in (OURS, test infrastructure): prototypes of the mpn_* entry points that the
 * reference's generated field code (build/f{r,q}_raw_generic.cpp) calls, bound to the runtime
 * libgmp.so.10 of this image (the development header is not installed).  Functions GMP implements as
 * header macros / inlines are spelled out here. */
#ifndef UGB_REF_SHIM_GMP_H
#define UGB_REF_SHIM_GMP_H
#include <stddef.h>
#include <stdint.h>

typedef unsigned long mp_limb_t;
typedef long mp_size_t;
typedef mp_limb_t* mp_ptr;
typedef const mp_limb_t* mp_srcptr;

#ifdef __cplusplus
extern "C" {
#endif
mp_limb_t __gmpn_add_n(mp_ptr, mp_srcptr, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_sub_n(mp_ptr, mp_srcptr, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_add_1(mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
mp_limb_t __gmpn_sub_1(mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
mp_limb_t __gmpn_add(mp_ptr, mp_srcptr, mp_size_t, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_mul_1(mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
mp_limb_t __gmpn_addmul_1(mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
mp_limb_t __gmpn_lshift(mp_ptr, mp_srcptr, mp_size_t, unsigned int);
mp_limb_t __gmpn_rshift(mp_ptr, mp_srcptr, mp_size_t, unsigned int);
int __gmpn_cmp(mp_srcptr, mp_srcptr, mp_size_t);
void __gmpn_copyi(mp_ptr, mp_srcptr, mp_size_t);
void __gmpn_and_n(mp_ptr, mp_srcptr, mp_srcptr, mp_size_t);
#ifdef __cplusplus
}
#endif

#define mpn_add_n __gmpn_add_n
#define mpn_sub_n __gmpn_sub_n
#define mpn_add_1 __gmpn_add_1
#define mpn_sub_1 __gmpn_sub_1
#define mpn_add __gmpn_add
#define mpn_mul_1 __gmpn_mul_1
#define mpn_addmul_1 __gmpn_addmul_1
#define mpn_lshift __gmpn_lshift
#define mpn_rshift __gmpn_rshift
#define mpn_cmp __gmpn_cmp
#define mpn_copyi __gmpn_copyi
#define mpn_and_n __gmpn_and_n

static inline void mpn_ior_n(mp_ptr r, mp_srcptr a, mp_srcptr b, mp_size_t n) { for (mp_size_t i = 0; i < n; i++) r[i] = a[i] | b[i]; }
static inline void mpn_xor_n(mp_ptr r, mp_srcptr a, mp_srcptr b, mp_size_t n) { for (mp_size_t i = 0; i < n; i++) r[i] = a[i] ^ b[i]; }
static inline void mpn_com(mp_ptr r, mp_srcptr a, mp_size_t n) { for (mp_size_t i = 0; i < n; i++) r[i] = ~a[i]; }
static inline int mpn_zero_p(mp_srcptr a, mp_size_t n) { for (mp_size_t i = 0; i < n; i++) if (a[i]) return 0; return 1; }
#endif
