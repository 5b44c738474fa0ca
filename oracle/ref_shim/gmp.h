/* Minimal <gmp.h> stand-in (OURS, test infrastructure): prototypes of the mpn_* / mpz_* entry points that the
 * reference's generated field code (build/f{r,q}_raw_generic.cpp, build/f{r,q}.cpp, build/f{r,q}_generic.cpp) and its file
 * loaders (src/zkey_utils.cpp, src/wtns_utils.cpp) call, bound to the runtime libgmp.so.10 of this image (the development
 * header is not installed).  Functions GMP implements as header macros / inlines are spelled out here; __mpz_struct has
 * GMP's documented layout. */
#ifndef UGB_REF_SHIM_GMP_H
#define UGB_REF_SHIM_GMP_H
#include <stddef.h>
#include <stdint.h>

typedef unsigned long mp_limb_t;
typedef long mp_size_t;
typedef mp_limb_t* mp_ptr;
typedef const mp_limb_t* mp_srcptr;

typedef struct {
    int _mp_alloc;
    int _mp_size;
    mp_limb_t* _mp_d;
} __mpz_struct;
typedef __mpz_struct mpz_t[1];
typedef __mpz_struct* mpz_ptr;
typedef const __mpz_struct* mpz_srcptr;

#ifdef __cplusplus
extern "C" {
#endif
void __gmpz_init(mpz_ptr);
void __gmpz_clear(mpz_ptr);
void __gmpz_import(mpz_ptr, size_t, int, size_t, int, size_t, const void*);
void* __gmpz_export(void*, size_t*, int, size_t, int, size_t, mpz_srcptr);
char* __gmpz_get_str(char*, int, mpz_srcptr);
void __gmpz_fdiv_r(mpz_ptr, mpz_srcptr, mpz_srcptr);
void __gmpz_fdiv_q(mpz_ptr, mpz_srcptr, mpz_srcptr);
void __gmpz_add(mpz_ptr, mpz_srcptr, mpz_srcptr);
void __gmpz_sub(mpz_ptr, mpz_srcptr, mpz_srcptr);
void __gmpz_mul(mpz_ptr, mpz_srcptr, mpz_srcptr);
void __gmpz_set(mpz_ptr, mpz_srcptr);
void __gmpz_set_si(mpz_ptr, long);
void __gmpz_set_ui(mpz_ptr, unsigned long);
int __gmpz_invert(mpz_ptr, mpz_srcptr, mpz_srcptr);
void __gmpz_init_set_ui(mpz_ptr, unsigned long);
void __gmpz_init_set_si(mpz_ptr, long);
int __gmpz_init_set_str(mpz_ptr, const char*, int);
size_t __gmpz_sizeinbase(mpz_srcptr, int);
void __gmpz_powm(mpz_ptr, mpz_srcptr, mpz_srcptr, mpz_srcptr);
void __gmpz_mul_2exp(mpz_ptr, mpz_srcptr, unsigned long);
long __gmpz_get_si(mpz_srcptr);
int __gmpz_fits_sint_p(mpz_srcptr);
int __gmpz_cmp(mpz_srcptr, mpz_srcptr);
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

#define mpz_init __gmpz_init
#define mpz_clear __gmpz_clear
#define mpz_import __gmpz_import
#define mpz_export __gmpz_export
#define mpz_get_str __gmpz_get_str
#define mpz_fdiv_r __gmpz_fdiv_r
#define mpz_fdiv_q __gmpz_fdiv_q
#define mpz_add __gmpz_add
#define mpz_sub __gmpz_sub
#define mpz_mul __gmpz_mul
#define mpz_set __gmpz_set
#define mpz_set_si __gmpz_set_si
#define mpz_set_ui __gmpz_set_ui
#define mpz_invert __gmpz_invert
#define mpz_init_set_ui __gmpz_init_set_ui
#define mpz_init_set_si __gmpz_init_set_si
#define mpz_init_set_str __gmpz_init_set_str
#define mpz_sizeinbase __gmpz_sizeinbase
#define mpz_powm __gmpz_powm
#define mpz_mul_2exp __gmpz_mul_2exp
#define mpz_get_si __gmpz_get_si
#define mpz_fits_sint_p __gmpz_fits_sint_p
#define mpz_cmp __gmpz_cmp
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
