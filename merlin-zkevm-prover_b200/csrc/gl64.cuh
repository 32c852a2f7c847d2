// gl64.cuh -- Goldilocks field arithmetic for sm_100a (64-bit integer pipe).
//
// p = 2^64 - 2^32 + 1 (reference: src/goldilocks/src/goldilocks_base_field.hpp:14, non-Montgomery).
// Replaces the reference's scalar x86 asm add/sub/mul (goldilocks_base_field_scalar.hpp:16-160)
// and the AVX2 mult_avx_128/reduce_avx_128_64 (goldilocks_base_field_avx.hpp:79-358).
//
// Representation: one u64 per element.  Like the reference, intermediate values are *lazily
// reduced*: every routine accepts ANY u64 as a representative and returns some representative
// in [0, 2^64); gl::canon() produces the canonical value (Goldilocks::toU64,
// goldilocks_base_field_tools.hpp:152-163).  Everything written to caller-visible memory is
// canonical.
//
// Instruction budget (cuobjdump, sm_100a): mul = 4 IMAD.WIDE + 3 IADD3 (128-bit product) +
// ~8 for the reduction (2^64 = 2^32-1, 2^96 = -1 mod p) ~ 15 SASS instructions.  On this SM the
// integer instructions do not overlap each other (profiles/r1_02_ubench_issue_mix.txt: IMAD.WIDE,
// LOP3, SHF, SEL cost 2 issue clocks per warp, IADD3 1-1.5), so the instruction count is the cost.
#pragma once
#include <stdint.h>

namespace gl {

typedef unsigned long long u64;
typedef unsigned int u32;

constexpr u64 P = 0xFFFFFFFF00000001ULL;
constexpr u32 EPS = 0xFFFFFFFFu; // 2^64 mod p

struct u128 {
    u32 x0, x1, x2, x3;
};

// 2^32 - 1 read from constant memory so that ptxas keeps x*(2^32-1)+y as ONE IMAD.WIDE.U32 with a
// carry-out predicate; with an immediate it strength-reduces to IMAD.IADD + IMAD.HI (half rate).
__constant__ u32 c_eps[2] = {0xFFFFFFFFu, 0u};

__device__ __forceinline__ u64 pack(u32 lo, u32 hi) { return ((u64)hi << 32) | lo; }

// x mod p for x < 2^64: x >= p  <=>  x + (2^32-1) carries out of 64 bits, and then x - p = x + (2^32-1) - 2^64.
// Written so that ptxas emits IADD3, IADD3.X, SEL, IMAD.WIDE (4 instructions; the compare form needs 6).
__device__ __forceinline__ u64 canon_slow(u64 x) { return x >= P ? x - P : x; }

// a*b + c as a 128-bit value (cannot overflow: (2^64-1)^2 + 2^64-1 < 2^128).
__device__ __forceinline__ u128 mad_wide(u64 a, u64 b, u64 c)
{
    u32 a0 = (u32)a, a1 = (u32)(a >> 32), b0 = (u32)b, b1 = (u32)(b >> 32);
    u32 c0 = (u32)c, c1 = (u32)(c >> 32);
    u128 r;
    asm("{\n\t"
        ".reg .u64 y;\n\t"
        ".reg .u32 p1, q0, q1, y0, y1, k;\n\t"
        "mad.lo.cc.u32 %0, %4, %6, %8;\n\t"
        "madc.hi.cc.u32 p1, %4, %6, %9;\n\t"
        "madc.lo.cc.u32 q0, %5, %7, 0;\n\t"
        "madc.hi.u32 q1, %5, %7, 0;\n\t"
        "mul.wide.u32 y, %4, %7;\n\t"
        "mov.b64 {y0, y1}, y;\n\t"
        "mad.lo.cc.u32 y0, %5, %6, y0;\n\t"
        "madc.hi.cc.u32 y1, %5, %6, y1;\n\t"
        "addc.u32 k, 0, 0;\n\t"
        "add.cc.u32 %1, p1, y0;\n\t"
        "addc.cc.u32 %2, q0, y1;\n\t"
        "addc.u32 %3, q1, k;\n\t"
        "}"
        : "=r"(r.x0), "=r"(r.x1), "=r"(r.x2), "=r"(r.x3)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(c0), "r"(c1));
    return r;
}

// x0 + x1 2^32 + x2 2^64 + x3 2^96  ==  {x0,x1} - x3 + x2*(2^32-1)   (mod p)
// One net correction k = carry - borrow in {-1,0,1}; result in [0, 2^64), not nec. canonical.
// Valid for ANY 128-bit input.
__device__ __forceinline__ u64 reduce128(u128 x)
{
    u32 r0, r1;
    asm("{\n\t"
        ".reg .u32 b, k, kl, kh;\n\t"
        "sub.cc.u32 %0, %2, %5;\n\t"
        "subc.cc.u32 %1, %3, 0;\n\t"
        "subc.u32 b, 0, 0;\n\t"                      // b = -borrow
        "mad.lo.cc.u32 %0, %4, %6, %0;\n\t"
        "madc.hi.cc.u32 %1, %4, %6, %1;\n\t"
        "addc.u32 k, b, 0;\n\t"                      // k = carry - borrow
        "neg.s32 kl, k;\n\t"
        "shr.s32 kh, k, 31;\n\t"
        "add.cc.u32 %0, %0, kl;\n\t"                 // r += k*(2^32-1)
        "addc.u32 %1, %1, kh;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(x.x0), "r"(x.x1), "r"(x.x2), "r"(x.x3), "r"(c_eps[0]));
    return pack(r0, r1);
}

// 96-bit input (x3 == 0): {x0,x1} + x2*(2^32-1), then + carry*(2^32-1) (cannot carry again): 2 IMAD.WIDE + SEL.
__device__ __forceinline__ u64 reduce96(u32 x0, u32 x1, u32 x2)
{
    u32 r0, r1;
    asm("{\n\t"
        ".reg .u32 k;\n\t"
        "mad.lo.cc.u32 %0, %4, %5, %2;\n\t"
        "madc.hi.cc.u32 %1, %4, %5, %3;\n\t"
        "addc.u32 k, 0, 0;\n\t"
        "mad.lo.cc.u32 %0, k, %5, %0;\n\t"
        "madc.hi.u32 %1, k, %5, %1;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(x0), "r"(x1), "r"(x2), "r"(c_eps[0]));
    return pack(r0, r1);
}

// reduce96 with ONE IMAD.WIDE: the conditional + (2^32 - 1) of the second step as mask arithmetic instead of a second multiply-add.
// FORM 1: k -> mask with a sign extension of bit 0 (bfe.s32: ALU pipe), add.cc, addc.  FORM 2: {r0 - k, r1 + k - borrow}.
// A/B switch for the MDS folds of the Poseidon layers (GLB_POSEIDON_FOLD); same value as reduce96 for every input.
template <int FORM>
__device__ __forceinline__ u64 reduce96_one(u32 x0, u32 x1, u32 x2)
{
    u32 r0, r1;
    if (FORM == 1) {
        asm("{\n\t"
            ".reg .u32 k, m;\n\t"
            "mad.lo.cc.u32 %0, %4, %5, %2;\n\t"
            "madc.hi.cc.u32 %1, %4, %5, %3;\n\t"
            "addc.u32 k, 0, 0;\n\t"
            "bfe.s32 m, k, 0, 1;\n\t"                 // 0 or 0xFFFFFFFF
            "add.cc.u32 %0, %0, m;\n\t"
            "addc.u32 %1, %1, 0;\n\t"
            "}"
            : "=&r"(r0), "=&r"(r1)
            : "r"(x0), "r"(x1), "r"(x2), "r"(c_eps[0]));
    } else {
        asm("{\n\t"
            ".reg .u32 k, b;\n\t"
            "mad.lo.cc.u32 %0, %4, %5, %2;\n\t"
            "madc.hi.cc.u32 %1, %4, %5, %3;\n\t"
            "addc.u32 k, 0, 0;\n\t"
            "sub.cc.u32 %0, %0, k;\n\t"
            "subc.u32 b, 0, 0;\n\t"                   // -borrow
            "add.u32 %1, %1, k;\n\t"
            "add.u32 %1, %1, b;\n\t"
            "}"
            : "=&r"(r0), "=&r"(r1)
            : "r"(x0), "r"(x1), "r"(x2), "r"(c_eps[0]));
    }
    return pack(r0, r1);
}

// 160-bit input x0 + x1 2^32 + x2 2^64 + x3 2^96 + x4 2^128 with x4 < 2^31 (2^128 = -2^32 mod p):
//   {x0,x1} - {x3,x4} + x2*(2^32-1); same single net correction as reduce128.
__device__ __forceinline__ u64 reduce160(u32 x0, u32 x1, u32 x2, u32 x3, u32 x4)
{
    u32 r0, r1;
    asm("{\n\t"
        ".reg .u32 b, k, kl, kh;\n\t"
        "sub.cc.u32 %0, %2, %5;\n\t"
        "subc.cc.u32 %1, %3, %6;\n\t"
        "subc.u32 b, 0, 0;\n\t"
        "mad.lo.cc.u32 %0, %4, %7, %0;\n\t"
        "madc.hi.cc.u32 %1, %4, %7, %1;\n\t"
        "addc.u32 k, b, 0;\n\t"
        "neg.s32 kl, k;\n\t"
        "shr.s32 kh, k, 31;\n\t"
        "add.cc.u32 %0, %0, kl;\n\t"
        "addc.u32 %1, %1, kh;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(x0), "r"(x1), "r"(x2), "r"(x3), "r"(x4), "r"(c_eps[0]));
    return pack(r0, r1);
}

// ---- ALU-only forms of the two reductions (no IMAD.WIDE for the x2 * (2^32 - 1) term): x2 * (2^32 - 1) = x2 * 2^32 - x2, so
//   {x0,x1} + x2 (2^32 - 1) - x3 = {x0, x1 + x2} - (x2 + x3), again with one net correction k = carry - borrow.
// Three more instructions than reduce128 / four more than reduce96, but all on the ALU pipe.  NOT used by default: they are
// the round-2 experiment for the Poseidon kernels, where ncu shows the FP64 pipe and the IMAD pipe taking turns while the ALU
// pipe is 36 % busy (profiles/r1_27_ubench_fp64_phase.txt); tests/test_reduction_forms.py checks the carry logic exhaustively
// on edge values and on random inputs with an exact simulation.
__device__ __forceinline__ u64 reduce128_alu(u128 x)
{
    u32 r0, r1;
    asm("{\n\t"
        ".reg .u32 h, c1, s0, s1, b, k, kl, kh;\n\t"
        "add.cc.u32 h, %3, %4;\n\t"
        "addc.u32 c1, 0, 0;\n\t"
        "add.cc.u32 s0, %4, %5;\n\t"
        "addc.u32 s1, 0, 0;\n\t"
        "sub.cc.u32 %0, %2, s0;\n\t"
        "subc.cc.u32 %1, h, s1;\n\t"
        "subc.u32 b, 0, 0;\n\t"                      // b = -borrow
        "add.u32 k, c1, b;\n\t"                      // k = carry - borrow
        "neg.s32 kl, k;\n\t"
        "shr.s32 kh, k, 31;\n\t"
        "add.cc.u32 %0, %0, kl;\n\t"                 // r += k*(2^32-1)
        "addc.u32 %1, %1, kh;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(x.x0), "r"(x.x1), "r"(x.x2), "r"(x.x3));
    return pack(r0, r1);
}
__device__ __forceinline__ u64 reduce96_alu(u32 x0, u32 x1, u32 x2)
{
    u32 r0, r1;
    asm("{\n\t"
        ".reg .u32 h, c1, b, k;\n\t"
        "add.cc.u32 h, %3, %4;\n\t"
        "addc.u32 c1, 0, 0;\n\t"
        "sub.cc.u32 %0, %2, %4;\n\t"
        "subc.cc.u32 %1, h, 0;\n\t"
        "subc.u32 b, 0, 0;\n\t"
        "add.u32 k, c1, b;\n\t"                      // 0 or 1
        "neg.s32 k, k;\n\t"                          // 0 or 0xFFFFFFFF
        "add.cc.u32 %0, %0, k;\n\t"
        "addc.u32 %1, %1, 0;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(x0), "r"(x1), "r"(x2));
    return pack(r0, r1);
}

// nvcc's native 128-bit product is one instruction shorter than mul_wide (the cross-term carry rides
// on IMAD.WIDE.U32.X instead of a SEL): 4 IMAD.WIDE + 3 ALU, then the 8-instruction reduction.
__device__ __forceinline__ u128 mul_wide_native(u64 a, u64 b)
{
    const unsigned __int128 x = (unsigned __int128)a * b;
    u128 r;
    r.x0 = (u32)x; r.x1 = (u32)(x >> 32); r.x2 = (u32)(x >> 64); r.x3 = (u32)(x >> 96);
    return r;
}
__device__ __forceinline__ u64 mul(u64 a, u64 b) { return reduce128(mul_wide_native(a, b)); }
__device__ __forceinline__ u64 sqr(u64 a) { return reduce128(mul_wide_native(a, a)); }
// a^2 with THREE 32 x 32 products (a0^2 + 2 a0 a1 2^32 + a1^2 2^64) instead of the four ptxas emits for a * a: one IMAD.WIDE (4 clocks,
// nothing issues beside it) traded for ~4 ALU-pipe instructions that double the cross product.  A/B switch for the Poseidon S-box
// (GLB_POSEIDON_SQR3); tests/test_gpu_parity.py::test_field_ops checks the value as op "sqr3".
__device__ __forceinline__ u64 sqr3(u64 a)
{
    // all in one asm block: with the three products formed in C and their halves passed in as separate operands, ptxas
    // materialised every extracted half with an addition of a zero uniform register (three no-op VIADD / IADD3 per squaring)
    u128 r;
    asm("{\n\t"
        ".reg .u64 p0, p2, c;\n\t"
        ".reg .u32 p0h, p2l, p2h, cl, ch, d0, d1, d2;\n\t"
        "mul.wide.u32 p0, %4, %4;\n\t"
        "mul.wide.u32 c, %4, %5;\n\t"
        "mul.wide.u32 p2, %5, %5;\n\t"
        "mov.b64 {%0, p0h}, p0;\n\t"
        "mov.b64 {cl, ch}, c;\n\t"
        "mov.b64 {p2l, p2h}, p2;\n\t"
        "shl.b32 d0, cl, 1;\n\t"
        "shf.l.wrap.b32 d1, cl, ch, 1;\n\t"
        "shr.b32 d2, ch, 31;\n\t"
        "add.cc.u32 %1, p0h, d0;\n\t"
        "addc.cc.u32 %2, p2l, d1;\n\t"
        "addc.u32 %3, p2h, d2;\n\t"
        "}"
        : "=r"(r.x0), "=r"(r.x1), "=r"(r.x2), "=r"(r.x3)
        : "r"((u32)a), "r"((u32)(a >> 32)));
    return reduce128(r);
}
// a*b + c  (fused: the addend rides on the first partial product)
__device__ __forceinline__ u64 mad(u64 a, u64 b, u64 c) { return reduce128(mad_wide(a, b, c)); }
// the same three with the ALU-only reduction (round-2 experiment, see reduce128_alu)
__device__ __forceinline__ u64 mul_alu(u64 a, u64 b) { return reduce128_alu(mul_wide_native(a, b)); }
__device__ __forceinline__ u64 sqr_alu(u64 a) { return reduce128_alu(mul_wide_native(a, a)); }
__device__ __forceinline__ u64 mad_alu(u64 a, u64 b, u64 c) { return reduce128_alu(mad_wide(a, b, c)); }

// a + b for arbitrary representatives.  Two possible wraps (reference add: scalar.hpp:16-35).
__device__ __forceinline__ u64 add(u64 a, u64 b)
{
    u32 r0, r1;
    asm("{\n\t"
        ".reg .u32 k;\n\t"
        "add.cc.u32 %0, %2, %4;\n\t"
        "addc.cc.u32 %1, %3, %5;\n\t"
        "addc.u32 k, 0, 0;\n\t"
        "neg.s32 k, k;\n\t"
        "add.cc.u32 %0, %0, k;\n\t"
        "addc.cc.u32 %1, %1, 0;\n\t"
        "addc.u32 k, 0, 0;\n\t"
        "neg.s32 k, k;\n\t"
        "add.cc.u32 %0, %0, k;\n\t"
        "addc.u32 %1, %1, 0;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"((u32)a), "r"((u32)(a >> 32)), "r"((u32)b), "r"((u32)(b >> 32)));
    return pack(r0, r1);
}

// a + b where b is CANONICAL (< p): the 65-bit sum folds with one IMAD.WIDE, s + carry*(2^32-1)
// (a + b - 2^64 + 2^32 - 1 < 2^64 because b < p).  IADD3, IADD3.X, SEL, IMAD.WIDE.
__device__ __forceinline__ u64 add_c(u64 a, u64 b)
{
    u32 r0, r1;
    asm("{\n\t"
        ".reg .u32 k;\n\t"
        "add.cc.u32 %0, %2, %4;\n\t"
        "addc.cc.u32 %1, %3, %5;\n\t"
        "addc.u32 k, 0, 0;\n\t"
        "mad.lo.cc.u32 %0, k, %6, %0;\n\t"
        "madc.hi.u32 %1, k, %6, %1;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"((u32)a), "r"((u32)(a >> 32)), "r"((u32)b), "r"((u32)(b >> 32)), "r"(c_eps[0]));
    return pack(r0, r1);
}

__device__ __forceinline__ u64 canon(u64 x)
{
    u32 r0, r1;
    asm("{\n\t"
        ".reg .u32 t0, t1, k;\n\t"
        "add.cc.u32 t0, %2, %4;\n\t"
        "addc.cc.u32 t1, %3, 0;\n\t"
        "addc.u32 k, 0, 0;\n\t"
        "mad.lo.cc.u32 %0, k, %4, %2;\n\t"
        "madc.hi.u32 %1, k, %4, %3;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"((u32)x), "r"((u32)(x >> 32)), "r"(c_eps[0]));
    return pack(r0, r1);
}

// ---- pipe-balanced forms for the NTT passes ------------------------------------------------------------------------------
// Measured issue model of the B200 SM sub-partition (profiles/r2_01_ubench_pipes.txt): an IMAD.WIDE.U32 takes 4 clocks and nothing
// issues beside it; every other integer instruction takes 2 clocks on its pipe -- IADD3 / LOP3 / SHF / SEL / ISETP on the ALU pipe,
// the 32-bit IMAD forms (IMAD, IMAD.X = multiply-add with carry-in) on the IMAD pipe -- and the two pipes run side by side.
// ptxas puts nearly all carry arithmetic on the ALU pipe, which then carries 2.4x the load of the IMAD pipe in the NTT pass and
// bounds it.  The forms below keep on the ALU pipe only what needs a carry-OUT (IADD3 is the one instruction that has it) and write
// every carry-consuming chain end and every mask / negation as a multiply-add (madc.lo / mul.lo -> IMAD.X / IMAD), and they use no
// IMAD.WIDE for the conditional +-(2^32 - 1).  One butterfly (canon_sel + add_cx + sub_cx) is 10 ALU + 9 IMAD + 0 IMAD.WIDE against
// 10 ALU + 1 IMAD + 2 IMAD.WIDE for canon + add_c + sub_c: 21 against 29 clocks in the measured model (tools/sass_cost.py).  The same
// treatment of the multiplication's reduction does NOT pay (ptxas already folds its sign handling into LEA.HI.X: 69 against 62 ALU
// instructions per 8 products), so gl::mul and mul_pow2 stay as they are.  gl_field_op_host ops 7-9 check the values.

// x mod p:  x >= p  <=>  x + (2^32 - 1) carries; then x - p = x + (2^32 - 1) - 2^64.          ALU 3, IMAD 3, no IMAD.WIDE
#ifndef GLB_CANON_SEL2
#define GLB_CANON_SEL2 1 // 1: no second carry chain (x >= p implies hi = 2^32 - 1 and lo >= 1, so x - p = {lo - 1, hi + 1 = 0}): ALU 2, IMAD 3
#endif
__device__ __forceinline__ u64 canon_sel(u64 x)
{
    u32 r0, r1;
#if GLB_CANON_SEL2
    asm("{\n\t"
        ".reg .u32 t, k;\n\t"
        "add.cc.u32 t, %2, 0xFFFFFFFF;\n\t"
        "addc.cc.u32 t, %3, 0;\n\t"
        "madc.lo.u32 k, 0, 0, 0;\n\t"               // k = carry = (x >= p)
        "mad.lo.u32 %0, k, 0xFFFFFFFF, %2;\n\t"     // lo - k   (lo >= 1 when k = 1: no borrow)
        "mad.lo.u32 %1, k, 1, %3;\n\t"              // hi + k   (hi = 2^32 - 1 when k = 1: wraps to 0)
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"((u32)x), "r"((u32)(x >> 32)));
#else
    asm("{\n\t"
        ".reg .u32 t, k, m;\n\t"
        "add.cc.u32 t, %2, 0xFFFFFFFF;\n\t"
        "addc.cc.u32 t, %3, 0;\n\t"
        "madc.lo.u32 k, 0, 0, 0;\n\t"               // k = carry
        "mul.lo.u32 m, k, 0xFFFFFFFF;\n\t"          // m = -k
        "add.cc.u32 %0, %2, m;\n\t"
        "madc.lo.u32 %1, %3, 1, 0;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"((u32)x), "r"((u32)(x >> 32)));
#endif
    return pack(r0, r1);
}
// a + b, b canonical:  s + carry * (2^32 - 1) = {s1, s0 + m} with m = -carry and the carry of the low word into s1.   ALU 3, IMAD 3
__device__ __forceinline__ u64 add_cx(u64 a, u64 b)
{
    u32 r0, r1;
    asm("{\n\t"
        ".reg .u32 k, m;\n\t"
        "add.cc.u32 %0, %2, %4;\n\t"
        "addc.cc.u32 %1, %3, %5;\n\t"
        "madc.lo.u32 k, 0, 0, 0;\n\t"
        "mul.lo.u32 m, k, 0xFFFFFFFF;\n\t"
        "add.cc.u32 %0, %0, m;\n\t"
        "madc.lo.u32 %1, %1, 1, 0;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"((u32)a), "r"((u32)(a >> 32)), "r"((u32)b), "r"((u32)(b >> 32)));
    return pack(r0, r1);
}
// a - b, b canonical:  d - borrow * (2^32 - 1) = {d1 + m, d0 + k} with m = -borrow, k = borrow.                      ALU 3-4, IMAD 2-3
// (The borrow is taken out of the subtract chain with subc and only ADD chains feed madc: ptxas hands madc the hardware carry, which
// after sub.cc is the complement of the borrow PTX documents -- a subtract chain must not end in a madc.)
__device__ __forceinline__ u64 sub_cx(u64 a, u64 b)
{
    u32 r0, r1;
    asm("{\n\t"
        ".reg .u32 m, k;\n\t"
        "sub.cc.u32 %0, %2, %4;\n\t"
        "subc.cc.u32 %1, %3, %5;\n\t"
        "subc.u32 m, 0, 0;\n\t"                     // m = -borrow (0 or 0xFFFFFFFF)
        "mul.lo.u32 k, m, 0xFFFFFFFF;\n\t"          // k = borrow
        "add.cc.u32 %0, %0, k;\n\t"
        "madc.lo.u32 %1, %1, 1, m;\n\t"             // d1 - borrow + carry
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"((u32)a), "r"((u32)(a >> 32)), "r"((u32)b), "r"((u32)(b >> 32)));
    return pack(r0, r1);
}

// x * 2^S mod p for a compile-time 0 <= S < 96 (2^96 = -1): funnel shifts + one short reduction.
template <int S>
__device__ __forceinline__ u64 mul_pow2(u64 x)
{
    static_assert(S >= 0 && S < 96, "shift out of range");
    constexpr int r = S % 32, q = S / 32;
    const u32 x0 = (u32)x, x1 = (u32)(x >> 32);
    u32 y0, y1, y2;
    if (r == 0) { y0 = x0; y1 = x1; y2 = 0; }
    else { y0 = x0 << r; y1 = __funnelshift_l(x0, x1, r); y2 = x1 >> (32 - r); }
    if (q == 0) return reduce96(y0, y1, y2);
    if (q == 1) { u128 v; v.x0 = 0; v.x1 = y0; v.x2 = y1; v.x3 = y2; return reduce128(v); }
    return reduce160(0, 0, y0, y1, y2);
}

// a - b where b is CANONICAL (< p): a single wrap suffices.
__device__ __forceinline__ u64 sub_c(u64 a, u64 b)
{
    u32 r0, r1;
    asm("{\n\t"
        ".reg .u32 k;\n\t"
        "sub.cc.u32 %0, %2, %4;\n\t"
        "subc.cc.u32 %1, %3, %5;\n\t"
        "subc.u32 k, 0, 0;\n\t"                      // 0 or 0xFFFFFFFF
        "sub.cc.u32 %0, %0, k;\n\t"                  // r -= (2^32-1) on borrow
        "subc.u32 %1, %1, 0;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"((u32)a), "r"((u32)(a >> 32)), "r"((u32)b), "r"((u32)(b >> 32)));
    return pack(r0, r1);
}

// a - b for arbitrary representatives.
__device__ __forceinline__ u64 sub(u64 a, u64 b) { return sub_c(a, canon(b)); }

__device__ __forceinline__ u64 neg(u64 a)
{
    a = canon(a);
    return a ? P - a : 0;
}

__device__ __forceinline__ u64 pow(u64 a, u64 e)
{
    u64 r = 1;
    while (e) {
        if (e & 1) r = mul(r, a);
        a = sqr(a);
        e >>= 1;
    }
    return r;
}

} // namespace gl
