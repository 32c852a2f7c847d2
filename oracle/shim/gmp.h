/* Minimal GMP declarations -- TEST INFRASTRUCTURE ONLY (oracle/_ref build).
 *
 * The GMP development headers are absent from this image but the runtime
 * library /usr/lib/x86_64-linux-gnu/libgmp.so.10 is present.  The reference
 * (src/goldilocks/src/ntt_goldilocks.hpp:61-135) uses a handful of mpz_*
 * calls at constructor time only.  This file declares exactly those entry
 * points with the public, stable GMP ABI (struct layout and __gmpz_* symbol
 * names) so the unmodified reference sources link against libgmp.so.10.
 */
#ifndef GLB200_ORACLE_GMP_SHIM_H
#define GLB200_ORACLE_GMP_SHIM_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef unsigned long mp_limb_t;
typedef unsigned long mp_bitcnt_t;
typedef struct { int _mp_alloc; int _mp_size; mp_limb_t *_mp_d; } __mpz_struct;
typedef __mpz_struct mpz_t[1];
typedef __mpz_struct *mpz_ptr;
typedef const __mpz_struct *mpz_srcptr;

void __gmpz_init(mpz_ptr);
void __gmpz_clear(mpz_ptr);
void __gmpz_set(mpz_ptr, mpz_srcptr);
void __gmpz_set_ui(mpz_ptr, unsigned long);
void __gmpz_init_set(mpz_ptr, mpz_srcptr);
void __gmpz_init_set_ui(mpz_ptr, unsigned long);
int  __gmpz_init_set_str(mpz_ptr, const char *, int);
void __gmpz_import(mpz_ptr, size_t, int, size_t, int, size_t, const void *);
void __gmpz_add(mpz_ptr, mpz_srcptr, mpz_srcptr);
void __gmpz_sub(mpz_ptr, mpz_srcptr, mpz_srcptr);
void __gmpz_mod(mpz_ptr, mpz_srcptr, mpz_srcptr);
void __gmpz_add_ui(mpz_ptr, mpz_srcptr, unsigned long);
void __gmpz_fdiv_q_2exp(mpz_ptr, mpz_srcptr, mp_bitcnt_t);
void __gmpz_powm(mpz_ptr, mpz_srcptr, mpz_srcptr, mpz_srcptr);
int  __gmpz_invert(mpz_ptr, mpz_srcptr, mpz_srcptr);
int  __gmpz_cmp(mpz_srcptr, mpz_srcptr);
int  __gmpz_cmp_ui(mpz_srcptr, unsigned long);
int  __gmpz_tstbit(mpz_srcptr, mp_bitcnt_t);
unsigned long __gmpz_get_ui(mpz_srcptr);
long __gmpz_get_si(mpz_srcptr);
char *__gmpz_get_str(char *, int, mpz_srcptr);

#define mpz_init __gmpz_init
#define mpz_clear __gmpz_clear
#define mpz_set __gmpz_set
#define mpz_set_ui __gmpz_set_ui
#define mpz_init_set __gmpz_init_set
#define mpz_init_set_ui __gmpz_init_set_ui
#define mpz_init_set_str __gmpz_init_set_str
#define mpz_import __gmpz_import
#define mpz_add __gmpz_add
#define mpz_sub __gmpz_sub
#define mpz_mod __gmpz_mod
#define mpz_add_ui __gmpz_add_ui
#define mpz_fdiv_q_2exp __gmpz_fdiv_q_2exp
#define mpz_powm __gmpz_powm
#define mpz_invert __gmpz_invert
#define mpz_cmp __gmpz_cmp
#define mpz_cmp_ui __gmpz_cmp_ui
#define mpz_tstbit __gmpz_tstbit
#define mpz_get_ui __gmpz_get_ui
#define mpz_get_si __gmpz_get_si
#define mpz_get_str __gmpz_get_str
#ifdef __cplusplus
}
#endif
#endif
