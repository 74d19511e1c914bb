/* Declaration-only stand-in for <gmp.h>.
 *
 * TEST INFRASTRUCTURE ONLY (oracle build). The image ships the GMP runtime
 * (libgmp.so.10, GMP 6.3.0) but not its development header. This file declares
 * exactly the subset of the public GMP ABI (LP64 x86-64) that the reference's
 * sources reach, mapped onto the exported __gmpn_ / __gmpz_ / __gmp_ symbols,
 * so that the unmodified reference can be compiled where it lies and linked
 * against the runtime library by full path.  It contains no GMP code.
 */
#ifndef B200MSM_ORACLE_GMP_SHIM_H
#define B200MSM_ORACLE_GMP_SHIM_H

#include <stddef.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef unsigned long mp_limb_t;
typedef long mp_limb_signed_t;
typedef long mp_size_t;
typedef unsigned long mp_bitcnt_t;
typedef mp_limb_t *mp_ptr;
typedef const mp_limb_t *mp_srcptr;

#define GMP_LIMB_BITS 64
#define GMP_NAIL_BITS 0
#define GMP_NUMB_BITS 64

typedef struct {
    int _mp_alloc;
    int _mp_size;
    mp_limb_t *_mp_d;
} __mpz_struct;
typedef __mpz_struct mpz_t[1];
typedef __mpz_struct *mpz_ptr;
typedef const __mpz_struct *mpz_srcptr;

/* mpn layer */
mp_limb_t __gmpn_add_n(mp_ptr, mp_srcptr, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_sub_n(mp_ptr, mp_srcptr, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_add_1(mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
mp_limb_t __gmpn_sub_1(mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
mp_limb_t __gmpn_add(mp_ptr, mp_srcptr, mp_size_t, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_sub(mp_ptr, mp_srcptr, mp_size_t, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_addmul_1(mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
void __gmpn_mul_n(mp_ptr, mp_srcptr, mp_srcptr, mp_size_t);
int __gmpn_cmp(mp_srcptr, mp_srcptr, mp_size_t);
void __gmpn_copyi(mp_ptr, mp_srcptr, mp_size_t);
void __gmpn_zero(mp_ptr, mp_size_t);
mp_limb_t __gmpn_rshift(mp_ptr, mp_srcptr, mp_size_t, unsigned int);
mp_size_t __gmpn_set_str(mp_ptr, const unsigned char *, size_t, int);
mp_size_t __gmpn_gcdext(mp_ptr, mp_ptr, mp_size_t *, mp_ptr, mp_size_t, mp_ptr, mp_size_t);
void __gmpn_tdiv_qr(mp_ptr, mp_ptr, mp_size_t, mp_srcptr, mp_size_t, mp_srcptr, mp_size_t);

#define mpn_add_n __gmpn_add_n
#define mpn_sub_n __gmpn_sub_n
#define mpn_add_1 __gmpn_add_1
#define mpn_sub_1 __gmpn_sub_1
#define mpn_add __gmpn_add
#define mpn_sub __gmpn_sub
#define mpn_addmul_1 __gmpn_addmul_1
#define mpn_mul_n __gmpn_mul_n
#define mpn_cmp __gmpn_cmp
#define mpn_copyi __gmpn_copyi
#define mpn_zero __gmpn_zero
#define mpn_rshift __gmpn_rshift
#define mpn_set_str __gmpn_set_str
#define mpn_gcdext __gmpn_gcdext
#define mpn_tdiv_qr __gmpn_tdiv_qr

/* mpz layer */
void __gmpz_init(mpz_ptr);
void __gmpz_clear(mpz_ptr);
void __gmpz_init_set(mpz_ptr, mpz_srcptr);
void __gmpz_set_ui(mpz_ptr, unsigned long);
void __gmpz_add_ui(mpz_ptr, mpz_srcptr, unsigned long);
void __gmpz_mul_2exp(mpz_ptr, mpz_srcptr, mp_bitcnt_t);
void __gmpz_fdiv_q_2exp(mpz_ptr, mpz_srcptr, mp_bitcnt_t);
unsigned long __gmpz_get_ui(mpz_srcptr);

#define mpz_init __gmpz_init
#define mpz_clear __gmpz_clear
#define mpz_init_set __gmpz_init_set
#define mpz_set_ui __gmpz_set_ui
#define mpz_add_ui __gmpz_add_ui
#define mpz_mul_2exp __gmpz_mul_2exp
#define mpz_fdiv_q_2exp __gmpz_fdiv_q_2exp
#define mpz_get_ui __gmpz_get_ui
#define mpz_sgn(Z) ((Z)->_mp_size < 0 ? -1 : (Z)->_mp_size > 0)

/* formatted output */
int __gmp_printf(const char *, ...);
int __gmp_snprintf(char *, size_t, const char *, ...);
#define gmp_printf __gmp_printf
#define gmp_snprintf __gmp_snprintf

#ifdef __cplusplus
}
#endif

#endif
