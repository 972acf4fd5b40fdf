// arith.cuh — 64-bit modular arithmetic for sm_100a (device side).
//
// Same arithmetic class as the reference (Harvey lazy butterflies with Shoup multipliers, 128-bit
// Barrett: include/butterfly.cuh:10-37, include/uintmodmath.cuh:96-231 of the reference) so every
// canonical result in [0,q) is bit-identical, but written for the Blackwell integer pipes:
//   * q is carried together with -q (two's complement) so  y*w - hi*q  becomes one chain of
//     IMAD / IMAD.WIDE accumulations (6 multiplier ops, no negate/subtract ALU ops);
//   * conditional subtraction is a predicated 64-bit subtract (ISETP, ISETP.EX, 2x @P IADD3);
//   * 128-bit accumulations use mad.lo.cc / madc.hi carry chains.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace b200fhe {

// Op trace: every kernel launcher adds what its launch must move and compute by the algorithm it
// implements — HBM bytes (each operand read once, each result written once; twiddles, keys' pointer tables and constants
// excluded), Harvey butterflies, and 64x64->128 multiply-accumulates (Shoup / Barrett products count as one).  Read through
// b200fhe_trace_read; bench.py turns the sums into the roofline figures of the composite operations.
void trace_add(double hbm_bytes, double butterflies, double macs);


typedef unsigned long long u64;
typedef unsigned int u32;

struct Mod {        // per-prime constants, 32 B, kept in a small device array indexed by prime id
    u64 q;          // modulus (< 2^61)
    u64 nq;         // 2^64 - q
    u64 r0, r1;     // floor(2^128 / q) low / high  (Barrett ratio, reference DModulus::const_ratio)
};

struct u128 {
    u64 lo, hi;
};

__device__ __forceinline__ u64 mulhi(u64 a, u64 b) { return __umul64hi(a, b); }

// x in [0, 2m) -> [0, m)
__device__ __forceinline__ u64 csub(u64 x, u64 m) {
    asm("{\n\t.reg .pred p;\n\tsetp.ge.u64 p, %0, %1;\n\t@p sub.u64 %0, %0, %1;\n\t}" : "+l"(x) : "l"(m));
    return x;
}

// a*w mod q, lazy: result in [0, 2q) for any a < 2^64.  ws = floor(w * 2^64 / q).
__device__ __forceinline__ u64 shoup_lazy(u64 a, u64 w, u64 ws, u64 nq) {
    u64 hi = mulhi(a, ws);
    return a * w + hi * nq;   // == a*w - hi*q  (mod 2^64)
}

__device__ __forceinline__ u64 shoup_mul(u64 a, u64 w, u64 ws, u64 q, u64 nq) { return csub(shoup_lazy(a, w, ws, nq), q); }

// Harvey Cooley-Tukey butterfly: x,y in [0,4q) -> [0,4q)
__device__ __forceinline__ void ct_bfly(u64 &x, u64 &y, u64 w, u64 ws, u64 q2, u64 nq) {
    u64 t = shoup_lazy(y, w, ws, nq);
    x = csub(x, q2);
    y = x + q2 - t;
    x = x + t;
}

// Gentleman-Sande butterfly: x,y in [0,2q) -> [0,2q)
__device__ __forceinline__ void gs_bfly(u64 &x, u64 &y, u64 w, u64 ws, u64 q2, u64 nq) {
    u64 t = x + q2 - y;
    x = csub(x + y, q2);
    y = shoup_lazy(t, w, ws, nq);
}

__device__ __forceinline__ u128 mul128(u64 a, u64 b) {
    u128 r;
    r.lo = a * b;
    r.hi = mulhi(a, b);
    return r;
}

// acc += a*b (128-bit, wraps like the reference's add_uint128_uint128).  Written as 32-bit multiply-add carry
// chains: ptxas turns the four lo/hi pairs into exactly four IMAD.WIDE.U32 with carry predicates (+1 IADD3.X),
// against seven multiplier instructions and a dozen adds for the 64-bit mad.lo.cc / madc.hi form (SASS-counted).
__device__ __forceinline__ void mac128(u128 &acc, u64 a, u64 b) {
    asm("{\n\t.reg .u32 r0, r1, r2, r3, a0, a1, b0, b1;\n\t"
        "mov.b64 {r0, r1}, %0;\n\t"
        "mov.b64 {r2, r3}, %1;\n\t"
        "mov.b64 {a0, a1}, %2;\n\t"
        "mov.b64 {b0, b1}, %3;\n\t"
        "mad.lo.cc.u32  r0, a0, b0, r0;\n\t"
        "madc.hi.cc.u32 r1, a0, b0, r1;\n\t"
        "madc.lo.cc.u32 r2, a1, b1, r2;\n\t"
        "madc.hi.u32    r3, a1, b1, r3;\n\t"
        "mad.lo.cc.u32  r1, a0, b1, r1;\n\t"
        "madc.hi.cc.u32 r2, a0, b1, r2;\n\t"
        "addc.u32       r3, r3, 0;\n\t"
        "mad.lo.cc.u32  r1, a1, b0, r1;\n\t"
        "madc.hi.cc.u32 r2, a1, b0, r2;\n\t"
        "addc.u32       r3, r3, 0;\n\t"
        "mov.b64 %0, {r0, r1};\n\t"
        "mov.b64 %1, {r2, r3};\n\t}"
        : "+l"(acc.lo), "+l"(acc.hi)
        : "l"(a), "l"(b));
}

// Barrett reduction of a 128-bit value to [0,q): floor(x*ratio/2^128) estimate, one correction.
__device__ __forceinline__ u64 barrett128(u128 x, const Mod &m) {
    u64 s;
    asm("{\n\t.reg .u64 tmp;\n\t"
        "mul.hi.u64 tmp, %1, %3;\n\t"
        "mad.lo.cc.u64 tmp, %1, %4, tmp;\n\t"
        "madc.hi.u64 %0, %1, %4, 0;\n\t"
        "mad.lo.cc.u64 tmp, %2, %3, tmp;\n\t"
        "madc.hi.u64 %0, %2, %3, %0;\n\t"
        "mad.lo.u64 %0, %2, %4, %0;\n\t}"
        : "=l"(s)
        : "l"(x.lo), "l"(x.hi), "l"(m.r0), "l"(m.r1));
    return csub(x.lo + s * m.nq, m.q);
}

// Lazy reduction of a 128-bit sum to [0, 4q): the quotient estimate drops the low-word product x.lo * r0 and adds the floors
// of the two middle products separately,
//   qh = x.hi * r1 + floor(x.hi * r0 / 2^64) + floor(x.lo * r1 / 2^64)        (ratio floor(2^128 / q) = r1 * 2^64 + r0)
// so Q - 3 <= qh <= Q for the true quotient Q (the exact Barrett estimate is within 1 of Q, each dropped fraction costs at
// most one more, and the three fractions sum to less than 3) and x - qh * q < 4q.  Needs r1 < 2^32, i.e. q > 2^32.
// For consumers that take lazy inputs (the forward transform's first butterflies): 9 multiplier instructions against the 18 of
// barrett128, no carry chain; the base-conversion loop in registers runs 19 % faster with it (profiles/r02_intpipe.json).
__device__ __forceinline__ u64 barrett128_lazy(u128 x, const Mod &m) {
    const u32 r1 = (u32) m.r1;
    const u64 t = (x.lo & 0xffffffffull) * r1;           // low half of x.lo times r1
    const u64 u = (x.lo >> 32) * r1 + (t >> 32);         // < 2^64: (2^32 - 1) * r1 + r1
    const u64 qh = x.hi * m.r1 + mulhi(x.hi, m.r0) + (u >> 32);
    return x.lo + qh * m.nq;
}

// x mod q for a 64-bit x (reference barrett_reduce_uint64_uint64)
__device__ __forceinline__ u64 barrett64(u64 x, const Mod &m) { return csub(x + mulhi(m.r1, x) * m.nq, m.q); }

__device__ __forceinline__ u64 mulmod(u64 a, u64 b, const Mod &m) { return barrett128(mul128(a, b), m); }

__device__ __forceinline__ u64 addmod(u64 a, u64 b, u64 q) { return csub(a + b, q); }
__device__ __forceinline__ u64 submod(u64 a, u64 b, u64 q) { return csub(a + q - b, q); }

// 128-bit global accesses
__device__ __forceinline__ ulonglong2 ldg2(const u64 *p) { return __ldg(reinterpret_cast<const ulonglong2 *>(p)); }
__device__ __forceinline__ ulonglong2 ld2(const u64 *p) { return *reinterpret_cast<const ulonglong2 *>(p); }
__device__ __forceinline__ void st2(u64 *p, u64 a, u64 b) { *reinterpret_cast<ulonglong2 *>(p) = make_ulonglong2(a, b); }

}  // namespace b200fhe
