// evalk.cu — evaluator / bootstrap kernels for sm_100a.  All of them are HBM-bound streaming kernels:
// one thread owns a PAIR of adjacent coefficients (16-byte accesses, 512 B per warp, whole 128-byte lines),
// grid.y walks limbs, products are accumulated lazily in 128 bits with ONE Barrett reduction per output.
#include <atomic>
#include "evalk.cuh"
#include "launch.cuh"
#include "tma.cuh"

namespace b200fhe {

static constexpr int kT = 256;
static constexpr int kTileBytes = kT * 16;  // one CTA tile: 256 coefficient pairs = 4 KiB per limb stream

// ---------------------------------------------------------------- per-limb scalar ops
template <int OP>
__global__ void __launch_bounds__(kT) limb_scalar_kernel(const Mod *__restrict__ mods, u64 *out, size_t out_stride,
                                                         const u64 *a, size_t a_stride, const u64 *acc,
                                                         size_t acc_stride, const __grid_constant__ LimbScalars s,
                                                         PrimeMap pm, u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2;
    const u32 t = blockIdx.y;
    const Mod md = mods[pm.prime(t)];
    const size_t id = size_t(t) * N + 2 * size_t(pair);
    const ulonglong2 x = ld2(a + blockIdx.z * a_stride + id);
    const u64 w = s.w[t];
    u64 r0, r1;
    if (OP == SC_MUL || OP == SC_MULADD) {
        const u64 ws = s.ws[t];
        r0 = shoup_mul(x.x, w, ws, md.q, md.nq);
        r1 = shoup_mul(x.y, w, ws, md.q, md.nq);
        if (OP == SC_MULADD) {
            const ulonglong2 y = ld2(acc + blockIdx.z * acc_stride + id);
            r0 = addmod(r0, y.x, md.q);
            r1 = addmod(r1, y.y, md.q);
        }
    } else if (OP == SC_ADD) {
        r0 = addmod(x.x, w, md.q);
        r1 = addmod(x.y, w, md.q);
    } else {
        r0 = submod(x.x, w, md.q);
        r1 = submod(x.y, w, md.q);
    }
    st2(out + blockIdx.z * out_stride + id, r0, r1);
}

cudaError_t limb_scalar(const Mod *mods, ScalarOp op, u64 *out, size_t out_stride, const u64 *a, size_t a_stride,
                        const u64 *acc, size_t acc_stride, const LimbScalars &s, u32 limbs, u32 npoly, PrimeMap pm,
                        size_t N, cudaStream_t st) {
    if (limbs == 0 || npoly == 0) return cudaSuccess;
    if (limbs > (u32) kMaxLimbs) return cudaErrorInvalidValue;
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, limbs, npoly);
    trace_add(8.0 * N * limbs * npoly * (op == SC_MULADD ? 3 : 2), 0, (op == SC_MUL || op == SC_MULADD) ? double(N) * limbs * npoly : 0);
    switch (op) {
        case SC_MUL: launch_k(limb_scalar_kernel<SC_MUL>, grid, kT, 0, st, mods, out, out_stride, a, a_stride, acc, acc_stride, s, pm, n_pairs); break;
        case SC_ADD: launch_k(limb_scalar_kernel<SC_ADD>, grid, kT, 0, st, mods, out, out_stride, a, a_stride, acc, acc_stride, s, pm, n_pairs); break;
        case SC_SUB: launch_k(limb_scalar_kernel<SC_SUB>, grid, kT, 0, st, mods, out, out_stride, a, a_stride, acc, acc_stride, s, pm, n_pairs); break;
        default: launch_k(limb_scalar_kernel<SC_MULADD>, grid, kT, 0, st, mods, out, out_stride, a, a_stride, acc, acc_stride, s, pm, n_pairs); break;
    }
    return cudaGetLastError();
}

template <int OP>
__global__ void __launch_bounds__(kT) poly_eltwise_kernel(const Mod *__restrict__ mods, u64 *out, size_t out_stride,
                                                          const u64 *a, size_t a_stride, const u64 *b, size_t b_stride,
                                                          PrimeMap pm, u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2;
    const Mod md = mods[pm.prime(blockIdx.y)];
    const size_t id = size_t(blockIdx.y) * N + 2 * size_t(pair);
    const ulonglong2 x = ld2(a + blockIdx.z * a_stride + id), y = ld2(b + blockIdx.z * b_stride + id);
    u64 r0, r1;
    if (OP == ELT_ADD) {
        r0 = addmod(x.x, y.x, md.q);
        r1 = addmod(x.y, y.y, md.q);
    } else if (OP == ELT_SUB) {
        r0 = submod(x.x, y.x, md.q);
        r1 = submod(x.y, y.y, md.q);
    } else {
        r0 = mulmod(x.x, y.x, md);
        r1 = mulmod(x.y, y.y, md);
    }
    st2(out + blockIdx.z * out_stride + id, r0, r1);
}

cudaError_t poly_eltwise(const Mod *mods, EltOp op, u64 *out, size_t out_stride, const u64 *a, size_t a_stride,
                         const u64 *b, size_t b_stride, u32 limbs, u32 npoly, PrimeMap pm, size_t N, cudaStream_t st) {
    if (limbs == 0 || npoly == 0) return cudaSuccess;
    trace_add(8.0 * N * limbs * (2.0 * npoly + (b_stride ? npoly : 1)), 0, op == ELT_MUL ? double(N) * limbs * npoly : 0);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, limbs, npoly);
    switch (op) {
        case ELT_ADD: launch_k(poly_eltwise_kernel<ELT_ADD>, grid, kT, 0, st, mods, out, out_stride, a, a_stride, b, b_stride, pm, n_pairs); break;
        case ELT_SUB: launch_k(poly_eltwise_kernel<ELT_SUB>, grid, kT, 0, st, mods, out, out_stride, a, a_stride, b, b_stride, pm, n_pairs); break;
        case ELT_MUL: launch_k(poly_eltwise_kernel<ELT_MUL>, grid, kT, 0, st, mods, out, out_stride, a, a_stride, b, b_stride, pm, n_pairs); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------- batched tensor product
template <bool SQUARE>
__global__ void __launch_bounds__(kT) tensor_split_kernel(const Mod *__restrict__ mods, u64 *out01, u64 *out2, const u64 *a,
                                                          const u64 *b, u32 L, u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2, LN = size_t(L) * N;
    const Mod md = mods[blockIdx.y];
    const size_t id = size_t(blockIdx.y) * N + 2 * size_t(pair), z = blockIdx.z;
    const u64 *az = a + z * 2 * LN;
    const ulonglong2 a0 = ld2(az + id), a1 = ld2(az + id + LN);
    const u64 x0[2] = {a0.x, a0.y}, x1[2] = {a1.x, a1.y};
    u64 d0[2], d1[2], d2[2];
    if (SQUARE) {
#pragma unroll
        for (int e = 0; e < 2; e++) {
            d0[e] = mulmod(x0[e], x0[e], md);
            u128 p = mul128(x0[e], x1[e]);  // 2*c0*c1 as one 128-bit value
            p.hi = (p.hi << 1) | (p.lo >> 63);
            p.lo <<= 1;
            d1[e] = barrett128(p, md);
            d2[e] = mulmod(x1[e], x1[e], md);
        }
    } else {
        const u64 *bz = b + z * 2 * LN;
        const ulonglong2 b0 = ld2(bz + id), b1 = ld2(bz + id + LN);
        const u64 y0[2] = {b0.x, b0.y}, y1[2] = {b1.x, b1.y};
#pragma unroll
        for (int e = 0; e < 2; e++) {
            d0[e] = mulmod(x0[e], y0[e], md);
            d2[e] = mulmod(x1[e], y1[e], md);
            u64 m = mulmod(x0[e] + x1[e], y0[e] + y1[e], md);  // Karatsuba middle term
            m = m + 2 * md.q - d0[e] - d2[e];
            d1[e] = csub(csub(m, md.q), md.q);
        }
    }
    st2(out01 + z * 2 * LN + id, d0[0], d0[1]);
    st2(out01 + z * 2 * LN + id + LN, d1[0], d1[1]);
    st2(out2 + z * LN + id, d2[0], d2[1]);
}

cudaError_t tensor_split(const Mod *mods, u64 *out01, u64 *out2, const u64 *a, const u64 *b, u32 L, u32 batch, size_t N,
                         cudaStream_t st) {
    if (L == 0 || batch == 0) return cudaSuccess;
    trace_add(8.0 * N * L * batch * (b ? 7 : 5), 0, 3.0 * N * L * batch);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, L, batch);
    if (b) launch_k(tensor_split_kernel<false>, grid, kT, 0, st, mods, out01, out2, a, b, L, n_pairs);
    else launch_k(tensor_split_kernel<true>, grid, kT, 0, st, mods, out01, out2, a, nullptr, L, n_pairs);
    return cudaGetLastError();
}

// The same product as two launches: d2 alone (the polynomial the key switch starts from) and (d0, d1), which only the last
// step of the key switch adds in, so it can run beside the key switch on another stream.  d1 = x0*y1 + x1*y0 is one 128-bit
// sum and one reduction; every result is the canonical residue, so the bits are tensor_split's.
template <bool SQUARE>
__global__ void __launch_bounds__(kT) tensor_d2_kernel(const Mod *__restrict__ mods, u64 *out2, const u64 *a, const u64 *b, u32 L,
                                                       u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2, LN = size_t(L) * N;
    const Mod md = mods[blockIdx.y];
    const size_t id = size_t(blockIdx.y) * N + 2 * size_t(pair), z = blockIdx.z;
    const ulonglong2 a1 = ld2(a + z * 2 * LN + LN + id);
    const ulonglong2 b1 = SQUARE ? a1 : ld2(b + z * 2 * LN + LN + id);
    st2(out2 + z * LN + id, mulmod(a1.x, b1.x, md), mulmod(a1.y, b1.y, md));
}

template <bool SQUARE>
__global__ void __launch_bounds__(kT) tensor_d01_kernel(const Mod *__restrict__ mods, u64 *out01, const u64 *a, const u64 *b, u32 L,
                                                        u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2, LN = size_t(L) * N;
    const Mod md = mods[blockIdx.y];
    const size_t id = size_t(blockIdx.y) * N + 2 * size_t(pair), z = blockIdx.z;
    const u64 *az = a + z * 2 * LN;
    const ulonglong2 a0 = ld2(az + id), a1 = ld2(az + id + LN);
    const u64 x0[2] = {a0.x, a0.y}, x1[2] = {a1.x, a1.y};
    u64 d0[2], d1[2];
    if (SQUARE) {
#pragma unroll
        for (int e = 0; e < 2; e++) {
            d0[e] = mulmod(x0[e], x0[e], md);
            u128 p = mul128(x0[e], x1[e]);
            p.hi = (p.hi << 1) | (p.lo >> 63);
            p.lo <<= 1;
            d1[e] = barrett128(p, md);
        }
    } else {
        const u64 *bz = b + z * 2 * LN;
        const ulonglong2 b0 = ld2(bz + id), b1 = ld2(bz + id + LN);
        const u64 y0[2] = {b0.x, b0.y}, y1[2] = {b1.x, b1.y};
#pragma unroll
        for (int e = 0; e < 2; e++) {
            d0[e] = mulmod(x0[e], y0[e], md);
            u128 p = mul128(x0[e], y1[e]);   // operands < q < 2^61: the sum of two products stays below 2^123
            mac128(p, x1[e], y0[e]);
            d1[e] = barrett128(p, md);
        }
    }
    st2(out01 + z * 2 * LN + id, d0[0], d0[1]);
    st2(out01 + z * 2 * LN + id + LN, d1[0], d1[1]);
}

cudaError_t tensor_d2(const Mod *mods, u64 *out2, const u64 *a, const u64 *b, u32 L, u32 batch, size_t N, cudaStream_t st) {
    if (L == 0 || batch == 0) return cudaSuccess;
    trace_add(8.0 * N * L * batch * (b ? 3 : 2), 0, 1.0 * N * L * batch);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, L, batch);
    if (b) launch_k(tensor_d2_kernel<false>, grid, kT, 0, st, mods, out2, a, b, L, n_pairs);
    else launch_k(tensor_d2_kernel<true>, grid, kT, 0, st, mods, out2, a, nullptr, L, n_pairs);
    return cudaGetLastError();
}

cudaError_t tensor_d01(const Mod *mods, u64 *out01, const u64 *a, const u64 *b, u32 L, u32 batch, size_t N, cudaStream_t st) {
    if (L == 0 || batch == 0) return cudaSuccess;
    trace_add(8.0 * N * L * batch * (b ? 6 : 4), 0, (b ? 3.0 : 2.0) * N * L * batch);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, L, batch);
    if (b) launch_k(tensor_d01_kernel<false>, grid, kT, 0, st, mods, out01, a, b, L, n_pairs);
    else launch_k(tensor_d01_kernel<true>, grid, kT, 0, st, mods, out01, a, nullptr, L, n_pairs);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- RaiseMod
__device__ __forceinline__ u64 recentre(u64 v, u64 half, u64 om, const Mod &nm) {
    if (nm.q > om) return v > half ? v + (nm.q - om) : v;
    const u64 diff = nm.q - barrett64(om, nm);
    return barrett64(v > half ? v + diff : v, nm);
}

__global__ void __launch_bounds__(kT) switch_modulus_kernel(const Mod *__restrict__ mods, u64 *dst, size_t dst_stride,
                                                            const u64 *src, size_t src_stride, u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2;
    const u32 j = blockIdx.y;
    const ulonglong2 v = ld2(src + blockIdx.z * src_stride + 2 * size_t(pair));
    u64 *o = dst + blockIdx.z * dst_stride + size_t(j) * N + 2 * size_t(pair);
    if (j == 0) {
        st2(o, v.x, v.y);
        return;
    }
    const u64 om = mods[0].q, half = om >> 1;
    const Mod nm = mods[j];
    st2(o, recentre(v.x, half, om, nm), recentre(v.y, half, om, nm));
}

cudaError_t switch_modulus(const Mod *mods, u64 *dst, size_t dst_stride, const u64 *src, size_t src_stride, u32 L,
                           u32 npoly, size_t N, cudaStream_t st) {
    if (L == 0 || npoly == 0) return cudaSuccess;
    trace_add(8.0 * N * npoly * (1 + L), 0, 0);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, L, npoly);
    launch_k(switch_modulus_kernel, grid, kT, 0, st, mods, dst, dst_stride, src, src_stride, n_pairs);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(kT) monomial_kernel(const Mod *__restrict__ mods, u64 *out, u32 index, u32 negative,
                                                      u32 N) {
    pdl_wait();
    pdl_trigger();
    const u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= N) return;
    out[size_t(blockIdx.y) * N + c] = c == index ? (negative ? mods[blockIdx.y].q - 1 : 1) : 0;
}

cudaError_t monomial_init(const Mod *mods, u64 *out, u32 index, bool negative, u32 L, size_t N, cudaStream_t st) {
    if (L == 0) return cudaSuccess;
    trace_add(8.0 * N * L, 0, 0);
    dim3 grid((u32) ((N + kT - 1) / kT), L);
    launch_k(monomial_kernel, grid, kT, 0, st, mods, out, index, negative ? 1u : 0u, (u32) N);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- pending affine combination, one pass
template <int K>
__global__ void __launch_bounds__(kT) affine_kernel(const Mod *__restrict__ mods, u64 *out, size_t out_stride,
                                                    const __grid_constant__ AffineArgs a, u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2;
    const u32 t = blockIdx.y, z = blockIdx.z;
    const Mod md = mods[t];
    const size_t id = size_t(t) * N + 2 * size_t(pair);
    ulonglong2 x[K];
#pragma unroll
    for (int i = 0; i < K; i++) x[i] = ld2(a.src[i] + z * a.stride[i] + id);
    u128 acc0{0, 0}, acc1{0, 0};
#pragma unroll
    for (int i = 0; i < K; i++) {
        const u64 w = a.s[i][t];
        mac128(acc0, x[i].x, w);
        mac128(acc1, x[i].y, w);
    }
    u64 r0 = barrett128(acc0, md), r1 = barrett128(acc1, md);
    if (a.has_kappa && z % a.polys == 0) {
        const u64 k = a.kappa[t];
        r0 = addmod(r0, k, md.q);
        r1 = addmod(r1, k, md.q);
    }
    st2(out + z * out_stride + id, r0, r1);
}

cudaError_t affine_combine(const Mod *mods, u64 *out, size_t out_stride, const AffineArgs &a, u32 limbs, u32 npoly,
                           size_t N, cudaStream_t st) {
    if (limbs == 0 || npoly == 0) return cudaSuccess;
    if (a.nterms < 1 || a.nterms > (u32) kMaxAffine || limbs > (u32) kMaxLimbs) return cudaErrorInvalidValue;
    trace_add(8.0 * N * limbs * npoly * (a.nterms + 1), 0, double(N) * limbs * npoly * a.nterms);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, limbs, npoly);
    switch (a.nterms) {
#define B200FHE_AFFINE_CASE(K) \
    case K: launch_k(affine_kernel<K>, grid, kT, 0, st, mods, out, out_stride, a, n_pairs); break;
        B200FHE_AFFINE_CASE(1) B200FHE_AFFINE_CASE(2) B200FHE_AFFINE_CASE(3) B200FHE_AFFINE_CASE(4)
        B200FHE_AFFINE_CASE(5) B200FHE_AFFINE_CASE(6) B200FHE_AFFINE_CASE(7) B200FHE_AFFINE_CASE(8)
#undef B200FHE_AFFINE_CASE
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------- hoisted rotation, fused
// 16-byte load of the aligned coefficient pair that holds pi(c), pi(c+1) (the pair may come back swapped)
__device__ __forceinline__ ulonglong2 ld_pair_perm(const u64 *limb, u32 base, bool swap) {
    ulonglong2 v = ld2(limb + base);
    if (swap) {
        const u64 tmp = v.x;
        v.x = v.y;
        v.y = tmp;
    }
    return v;
}

template <bool ADD_FIRST, bool ACC>
__global__ void __launch_bounds__(kT) fast_rot_kernel(const Mod *__restrict__ mods, const __grid_constant__ FastRotArgs a,
                                                      const __grid_constant__ LimbScalars PmodQ, PrimeMap pm,
                                                      u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const u32 t = blockIdx.y;
    const size_t N = size_t(n_pairs) * 2;
    const u32 twr = pm.prime(t);
    const Mod md = mods[twr];
    u32 base = 2 * pair;
    bool swap = false;
    if (a.perm) {
        const u32 p0 = __ldg(a.perm + 2 * pair);
        base = p0 & ~1u;
        swap = p0 & 1u;
    }
    const size_t up_stride = size_t(a.size_QlP) * N, k_poly = size_t(a.size_QP) * N;
    const u64 *up = a.up + size_t(t) * N;
    u128 a00 = {0, 0}, a01 = {0, 0}, a10 = {0, 0}, a11 = {0, 0};
    for (u32 d = 0; d < a.beta; d++) {
        const ulonglong2 v = ld_pair_perm(up + d * up_stride, base, swap);
        const u64 *k = a.evk.d[d] + size_t(twr) * N;
        const ulonglong2 k0 = ld_pair_perm(k, base, swap), k1 = ld_pair_perm(k + k_poly, base, swap);
        mac128(a00, v.x, k0.x);
        mac128(a01, v.y, k0.y);
        mac128(a10, v.x, k1.x);
        mac128(a11, v.y, k1.y);
    }
    u64 r0 = barrett128(a00, md), r1 = barrett128(a01, md);
    if (ADD_FIRST && t < a.size_Ql) {
        const ulonglong2 c = ld_pair_perm(a.c0 + size_t(t) * N, base, swap);
        const u64 w = PmodQ.w[t], ws = PmodQ.ws[t];
        r0 = addmod(r0, shoup_mul(c.x, w, ws, md.q, md.nq), md.q);
        r1 = addmod(r1, shoup_mul(c.y, w, ws, md.q, md.nq), md.q);
    }
    u64 *o = a.out + size_t(t) * N + 2 * size_t(pair);
    u64 r2 = barrett128(a10, md), r3 = barrett128(a11, md);
    if (ACC) {
        const ulonglong2 y0 = ld2(o), y1 = ld2(o + up_stride);
        r0 = addmod(r0, y0.x, md.q);
        r1 = addmod(r1, y0.y, md.q);
        r2 = addmod(r2, y1.x, md.q);
        r3 = addmod(r3, y1.y, md.q);
    }
    st2(o, r0, r1);
    st2(o + up_stride, r2, r3);
}

// TMA variant.  The NTT-domain automorphism maps every aligned block of 2^s coefficients onto an aligned block
// (only the order inside changes), so the 512 outputs of a CTA come from ONE contiguous 4 KiB source tile per
// stream: thread 0 issues 3*beta (+1, +2) bulk copies, everybody waits on one mbarrier, and the in-tile
// permutation is applied when reading shared memory (16-byte reads of an aligned permuted pair: conflict-free).
// ~40 KiB per CTA are in flight from the first instruction on, instead of one dependent table load + 9 loads.
__device__ __forceinline__ u32 auto_perm(u32 c, u32 k, int logN) {
    const u32 j = __brev(c) >> (32 - logN);
    const u32 idx = (((2 * j + 1) * k) & ((2u << logN) - 1)) >> 1;
    return __brev(idx) >> (32 - logN);
}

template <bool ADD_FIRST, bool ACC>
__global__ void __launch_bounds__(kT) fast_rot_tma_kernel(const Mod *__restrict__ mods, const __grid_constant__ FastRotArgs a,
                                                          const __grid_constant__ LimbScalars PmodQ, PrimeMap pm) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) u64 bar;
    ulonglong2 *tiles = reinterpret_cast<ulonglong2 *>(smem_raw);  // [3*beta (+1) (+2)][kT]
    const u32 tid = threadIdx.x, t = blockIdx.y;
    const size_t N = size_t(1) << a.logN;
    const u32 twr = pm.prime(t);
    const Mod md = mods[twr];
    const u32 c_tile = blockIdx.x * (2 * kT);
    const u32 src_tile = auto_perm(c_tile, a.auto_index, a.logN) & ~(2u * kT - 1);
    const size_t up_stride = size_t(a.size_QlP) * N, k_poly = size_t(a.size_QP) * N;
    const bool first = ADD_FIRST && t < a.size_Ql;
    const u32 n_in = 3 * a.beta + (first ? 1 : 0);
    if (tid == 0) {
        mbar_init(&bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(&bar, (n_in + (ACC ? 2 : 0)) * kTileBytes);
        const u64 *up = a.up + size_t(t) * N + src_tile;
        for (u32 d = 0; d < a.beta; d++) {
            const u64 *k = a.evk.d[d] + size_t(twr) * N + src_tile;
            tma_load_1d(tiles + (3 * d) * kT, up + d * up_stride, kTileBytes, &bar);
            tma_load_1d(tiles + (3 * d + 1) * kT, k, kTileBytes, &bar);
            tma_load_1d(tiles + (3 * d + 2) * kT, k + k_poly, kTileBytes, &bar);
        }
        if (first) tma_load_1d(tiles + 3 * a.beta * kT, a.c0 + size_t(t) * N + src_tile, kTileBytes, &bar);
        if (ACC) {
            const u64 *o = a.out + size_t(t) * N + c_tile;
            tma_load_1d(tiles + n_in * kT, o, kTileBytes, &bar);
            tma_load_1d(tiles + (n_in + 1) * kT, o + up_stride, kTileBytes, &bar);
        }
    }
    const u32 p0 = auto_perm(c_tile + 2 * tid, a.auto_index, a.logN);
    const u32 slot = (p0 & (2u * kT - 1)) >> 1;
    const bool swap = p0 & 1u;
    __syncthreads();  // barrier initialised
    mbar_wait(&bar, 0);
    auto rd = [&](u32 tile) {
        ulonglong2 v = tiles[tile * kT + slot];
        if (swap) {
            const u64 tmp = v.x;
            v.x = v.y;
            v.y = tmp;
        }
        return v;
    };
    u128 a00 = {0, 0}, a01 = {0, 0}, a10 = {0, 0}, a11 = {0, 0};
    for (u32 d = 0; d < a.beta; d++) {
        const ulonglong2 v = rd(3 * d), k0 = rd(3 * d + 1), k1 = rd(3 * d + 2);
        mac128(a00, v.x, k0.x);
        mac128(a01, v.y, k0.y);
        mac128(a10, v.x, k1.x);
        mac128(a11, v.y, k1.y);
    }
    u64 r0 = barrett128(a00, md), r1 = barrett128(a01, md), r2 = barrett128(a10, md), r3 = barrett128(a11, md);
    if (first) {
        const ulonglong2 c = rd(3 * a.beta);
        const u64 w = PmodQ.w[t], ws = PmodQ.ws[t];
        r0 = addmod(r0, shoup_mul(c.x, w, ws, md.q, md.nq), md.q);
        r1 = addmod(r1, shoup_mul(c.y, w, ws, md.q, md.nq), md.q);
    }
    if (ACC) {
        const ulonglong2 y0 = tiles[n_in * kT + tid], y1 = tiles[(n_in + 1) * kT + tid];
        r0 = addmod(r0, y0.x, md.q);
        r1 = addmod(r1, y0.y, md.q);
        r2 = addmod(r2, y1.x, md.q);
        r3 = addmod(r3, y1.y, md.q);
    }
    u64 *o = a.out + size_t(t) * N + c_tile + 2 * tid;
    st2(o, r0, r1);
    st2(o + up_stride, r2, r3);
}

// ---- all rotations of one hoisting in one launch (see evalk.cuh)
#ifndef B200FHE_ROT_STAGES
#define B200FHE_ROT_STAGES 2
#endif
static constexpr int kRotStages = B200FHE_ROT_STAGES;

template <bool ADD_FIRST>
__global__ void __launch_bounds__(kT) fast_rot_multi_kernel(const Mod *__restrict__ mods,
                                                            const __grid_constant__ FastRotMultiArgs a,
                                                            const __grid_constant__ LimbScalars PmodQ, PrimeMap pm) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) u64 bar_in, full[kRotStages];
    ulonglong2 *tiles = reinterpret_cast<ulonglong2 *>(smem_raw);  // [beta] digits, [1] c0, [stages][2*beta] keys
    const u32 tid = threadIdx.x, t = a.limb0 + blockIdx.y, beta = a.beta;
    const size_t N = size_t(1) << a.logN;
    const u32 twr = pm.prime(t);
    const Mod md = mods[twr];
    const u32 src_tile = blockIdx.x * (2 * kT);
    const size_t up_stride = size_t(a.size_QlP) * N, k_poly = size_t(a.size_QP) * N;
    const bool first = ADD_FIRST && t < a.size_Ql;
    ulonglong2 *keys = tiles + (beta + 1) * kT;
    auto issue_keys = [&](u32 r) {  // thread 0 only
        const u32 s = r % kRotStages;
        mbar_expect_tx(&full[s], 2 * beta * kTileBytes);
        ulonglong2 *dst = keys + size_t(s) * 2 * beta * kT;
        for (u32 d = 0; d < beta; d++) {
            const u64 *k = a.evk[r][d] + size_t(twr) * N + src_tile;
            tma_load_1d(dst + (2 * d) * kT, k, kTileBytes, &full[s]);
            tma_load_1d(dst + (2 * d + 1) * kT, k + k_poly, kTileBytes, &full[s]);
        }
    };
    if (tid == 0) {
        mbar_init(&bar_in, 1);
        for (int s = 0; s < kRotStages; s++) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(&bar_in, (beta + (first ? 1 : 0)) * kTileBytes);
        for (u32 d = 0; d < beta; d++)
            tma_load_1d(tiles + d * kT, a.up + d * up_stride + size_t(t) * N + src_tile, kTileBytes, &bar_in);
        if (first) tma_load_1d(tiles + beta * kT, a.c0 + size_t(t) * N + src_tile, kTileBytes, &bar_in);
        for (u32 r = 0; r < (u32) kRotStages && r < a.nrot; r++) issue_keys(r);
    }
    __syncthreads();  // barriers initialised
    mbar_wait(&bar_in, 0);
    const u64 w = PmodQ.w[t], ws = PmodQ.ws[t];
    for (u32 r = 0; r < a.nrot; r++) {
        const u32 s = r % kRotStages;
        // the tile this source tile lands on, and the source pair my output pair reads
        const u32 c_tile = auto_perm(src_tile, a.k_inv[r], a.logN) & ~(2u * kT - 1);
        const u32 p0 = auto_perm(c_tile + 2 * tid, a.k[r], a.logN);
        const u32 slot = (p0 & (2u * kT - 1)) >> 1;
        const bool swap = p0 & 1u;
        auto rd = [&](const ulonglong2 *tile) {
            ulonglong2 v = tile[slot];
            if (swap) {
                const u64 tmp = v.x;
                v.x = v.y;
                v.y = tmp;
            }
            return v;
        };
        mbar_wait(&full[s], (r / kRotStages) & 1);
        const ulonglong2 *kt = keys + size_t(s) * 2 * beta * kT;
        u128 a00 = {0, 0}, a01 = {0, 0}, a10 = {0, 0}, a11 = {0, 0};
        for (u32 d = 0; d < beta; d++) {
            const ulonglong2 v = rd(tiles + d * kT), k0 = rd(kt + (2 * d) * kT), k1 = rd(kt + (2 * d + 1) * kT);
            mac128(a00, v.x, k0.x);
            mac128(a01, v.y, k0.y);
            mac128(a10, v.x, k1.x);
            mac128(a11, v.y, k1.y);
        }
        __syncthreads();  // every thread has read stage s: refill it while the reductions run
        if (tid == 0 && r + kRotStages < a.nrot) issue_keys(r + kRotStages);
        u64 r0 = barrett128(a00, md), r1 = barrett128(a01, md), r2 = barrett128(a10, md), r3 = barrett128(a11, md);
        if (first) {
            const ulonglong2 c = rd(tiles + beta * kT);
            r0 = addmod(r0, shoup_mul(c.x, w, ws, md.q, md.nq), md.q);
            r1 = addmod(r1, shoup_mul(c.y, w, ws, md.q, md.nq), md.q);
        }
        u64 *o = a.out[r] + size_t(t) * N + c_tile + 2 * tid;
        st2(o, r0, r1);
        st2(o + up_stride, r2, r3);
    }
}

bool fast_rotation_ext_multi_supported(u32 beta, int logN) { return beta >= 1 && beta <= (u32) kRotDigits && logN > 9; }

cudaError_t fast_rotation_ext_multi(const Mod *mods, const FastRotMultiArgs &a, const LimbScalars &PmodQ, PrimeMap pm,
                                    size_t N, cudaStream_t st) {
    if (a.nrot == 0) return cudaSuccess;
    if (a.nrot > (u32) kRotMax || a.size_QlP > (u32) kMaxLimbs || !fast_rotation_ext_multi_supported(a.beta, a.logN) ||
        N % (2 * kT) != 0 || a.limb0 + a.nlimbs > a.size_QlP || (!a.nlimbs && a.limb0))
        return cudaErrorInvalidValue;
    const u32 rows = a.nlimbs ? a.nlimbs : a.size_QlP;
    const double part = double(rows) / a.size_QlP;   // a limb range moves and multiplies its share
    // digits and c0 once per launch; per rotation the key (2 beta QlP limbs) in and the extended ciphertext (2 QlP) out
    trace_add(part * 8.0 * N * (double(a.beta) * a.size_QlP + (a.c0 ? a.size_Ql : 0) + double(a.nrot) * (2.0 * a.beta + 2.0) * a.size_QlP), 0,
              part * double(N) * a.nrot * (2.0 * a.beta * a.size_QlP + (a.c0 ? a.size_Ql : 0)));
    const size_t smem = size_t(a.beta + 1 + kRotStages * 2 * a.beta) * kTileBytes;
    static std::atomic<size_t> attr[2];  // largest dynamic shared memory size enabled so far (host threads of the lanes race here)
    const int v = a.c0 ? 1 : 0;
    if (smem > attr[v]) {
        cudaError_t e = v ? cudaFuncSetAttribute(fast_rot_multi_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem)
                          : cudaFuncSetAttribute(fast_rot_multi_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        if (e != cudaSuccess) return e;
        attr[v] = smem;
    }
    dim3 grid((u32) (N / (2 * kT)), rows);
    if (v) launch_k(fast_rot_multi_kernel<true>, grid, kT, smem, st, mods, a, PmodQ, pm);
    else launch_k(fast_rot_multi_kernel<false>, grid, kT, smem, st, mods, a, PmodQ, pm);
    return cudaGetLastError();
}

template <bool ADD_FIRST, bool ACC>
static cudaError_t fast_rot_tma_launch(const Mod *mods, const FastRotArgs &a, const LimbScalars &PmodQ, PrimeMap pm,
                                       dim3 grid, size_t smem, cudaStream_t st) {
    static std::atomic<size_t> attr{0};  // per instance: largest dynamic shared memory size enabled so far
    if (smem > attr) {
        cudaError_t e = cudaFuncSetAttribute(fast_rot_tma_kernel<ADD_FIRST, ACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        if (e != cudaSuccess) return e;
        attr = smem;
    }
    launch_k(fast_rot_tma_kernel<ADD_FIRST, ACC>, grid, kT, smem, st, mods, a, PmodQ, pm);
    return cudaGetLastError();
}

cudaError_t fast_rotation_ext(const Mod *mods, const FastRotArgs &a, const LimbScalars &PmodQ, PrimeMap pm, size_t N,
                              cudaStream_t st) {
    if (a.size_QlP > (u32) kMaxLimbs || a.beta > (u32) kMaxDigits) return cudaErrorInvalidValue;
    trace_add(8.0 * N * ((3.0 * a.beta + 2.0 + (a.accumulate ? 2 : 0)) * a.size_QlP + (a.c0 ? a.size_Ql : 0)), 0,
              double(N) * (2.0 * a.beta * a.size_QlP + (a.c0 ? a.size_Ql : 0)));
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, a.size_QlP);
    const size_t smem = size_t(3 * a.beta + 3) * kTileBytes;
    if (smem <= 96 * 1024 && N % (2 * kT) == 0 && a.logN > 9 && (a.auto_index & 1)) {
        if (a.c0 && a.accumulate) return fast_rot_tma_launch<true, true>(mods, a, PmodQ, pm, grid, smem, st);
        if (a.c0) return fast_rot_tma_launch<true, false>(mods, a, PmodQ, pm, grid, smem, st);
        if (a.accumulate) return fast_rot_tma_launch<false, true>(mods, a, PmodQ, pm, grid, smem, st);
        return fast_rot_tma_launch<false, false>(mods, a, PmodQ, pm, grid, smem, st);
    }
    if (a.c0 && a.accumulate) launch_k(fast_rot_kernel<true, true>, grid, kT, 0, st, mods, a, PmodQ, pm, n_pairs);
    else if (a.c0) launch_k(fast_rot_kernel<true, false>, grid, kT, 0, st, mods, a, PmodQ, pm, n_pairs);
    else if (a.accumulate) launch_k(fast_rot_kernel<false, true>, grid, kT, 0, st, mods, a, PmodQ, pm, n_pairs);
    else launch_k(fast_rot_kernel<false, false>, grid, kT, 0, st, mods, a, PmodQ, pm, n_pairs);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- BSGS plaintext-matrix product
// HBM-bound: per giant step j a CTA needs one 4 KiB tile of each of 2 + NB streams (the two polynomials of the
// rotated ciphertext j and NB plaintext diagonals).  The tiles are staged through shared memory by 1-D bulk TMA
// copies (cp.async.bulk, one elected thread, mbarrier complete_tx) in a kStages-deep ring, so ~70 KiB per CTA are
// in flight independent of the register file (the accumulators alone take 64 registers per thread).
static constexpr int kStages = 4;

template <int NB>
struct BsgsPack {  // by-value kernel arguments of one launch (NB baby steps)
    const u64 *ct[kMaxBsgs];
    const u64 *pt[NB][kMaxBsgs];
    u64 *out[NB];
    u32 g, size_QlP, accumulate, pt_shift, limb0;
};

template <int NB>
__global__ void __launch_bounds__(kT, 2) bsgs_kernel(const Mod *__restrict__ mods, const __grid_constant__ BsgsPack<NB> a,
                                                     PrimeMap pm) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) u64 full[kStages];
    ulonglong2 *tiles = reinterpret_cast<ulonglong2 *>(smem_raw);  // [kStages][2 + NB][kT]
    const u32 tid = threadIdx.x, t = a.limb0 + blockIdx.y;
    const size_t N = size_t(gridDim.x) * kT * 2;
    const Mod md = mods[pm.prime(t)];
    const size_t tile_off = size_t(t) * N + size_t(blockIdx.x) * kT * 2;  // first coefficient of this CTA's tile
    const size_t off = tile_off + 2 * size_t(tid), poly = size_t(a.size_QlP) * N;
    if (tid == 0) {
        for (int s = 0; s < kStages; s++) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](u32 j) {  // thread 0 only
        const u32 s = j % kStages;
        ulonglong2 *dst = tiles + size_t(s) * (2 + NB) * kT;
        u32 n = 2;
#pragma unroll
        for (int i = 0; i < NB; i++) n += a.pt[i][j] != nullptr;
        mbar_expect_tx(&full[s], 2 * kTileBytes + (n - 2) * (kTileBytes >> a.pt_shift));
        const u64 *c = a.ct[j] + tile_off;
        tma_load_1d(dst, c, kTileBytes, &full[s]);
        tma_load_1d(dst + kT, c + poly, kTileBytes, &full[s]);
#pragma unroll
        for (int i = 0; i < NB; i++)
            if (a.pt[i][j]) tma_load_1d(dst + (2 + i) * kT, a.pt[i][j] + (tile_off >> a.pt_shift), kTileBytes >> a.pt_shift, &full[s]);
    };
    if (tid == 0)
        for (u32 j = 0; j < kStages - 1 && j < a.g; j++) issue(j);

    u128 acc[NB][4];
#pragma unroll
    for (int i = 0; i < NB; i++) {
        if (a.accumulate) {
            const u64 *o = a.out[i] + off;
            const ulonglong2 y0 = ld2(o), y1 = ld2(o + poly);
            acc[i][0] = {y0.x, 0};
            acc[i][1] = {y0.y, 0};
            acc[i][2] = {y1.x, 0};
            acc[i][3] = {y1.y, 0};
        } else {
#pragma unroll
            for (int e = 0; e < 4; e++) acc[i][e] = {0, 0};
        }
    }
    for (u32 j = 0; j < a.g; j++) {
        // every thread has finished reading stage (j-1) % kStages: thread 0 may refill it with step j + kStages - 1
        __syncthreads();
        if (tid == 0 && j + kStages - 1 < a.g) issue(j + kStages - 1);
        const u32 s = j % kStages;
        mbar_wait(&full[s], (j / kStages) & 1);
        const ulonglong2 *src = tiles + size_t(s) * (2 + NB) * kT + tid;
        const ulonglong2 x0 = src[0], x1 = src[kT];
#pragma unroll
        for (int i = 0; i < NB; i++) {
            if (a.pt[i][j]) {
                ulonglong2 m;
                if (a.pt_shift) {  // compact plaintext: both coefficients of the pair share one value
                    const u64 v = reinterpret_cast<const u64 *>(src - tid + (2 + i) * kT)[(2 * tid) >> a.pt_shift];
                    m = make_ulonglong2(v, v);
                } else {
                    m = src[(2 + i) * kT];
                }
                mac128(acc[i][0], x0.x, m.x);
                mac128(acc[i][1], x0.y, m.y);
                mac128(acc[i][2], x1.x, m.x);
                mac128(acc[i][3], x1.y, m.y);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < NB; i++) {
        u64 *o = a.out[i] + off;
        st2(o, barrett128(acc[i][0], md), barrett128(acc[i][1], md));
        st2(o + poly, barrett128(acc[i][2], md), barrett128(acc[i][3], md));
    }
}

template <int NB>
static cudaError_t bsgs_launch(const Mod *mods, const BsgsArgs &a, u32 i0, PrimeMap pm, dim3 grid, cudaStream_t st) {
    BsgsPack<NB> k{};
    for (u32 j = 0; j < a.g; j++) k.ct[j] = a.ct[j];
    for (int i = 0; i < NB; i++) {
        for (u32 j = 0; j < a.g; j++) k.pt[i][j] = a.pt[size_t(i0 + i) * a.g + j];
        k.out[i] = a.out[i0 + i];
    }
    k.g = a.g;
    k.size_QlP = a.size_QlP;
    k.accumulate = a.accumulate;
    k.pt_shift = a.pt_shift;
    k.limb0 = a.nlimbs ? a.limb0 : 0;
    const size_t smem = size_t(kStages) * (2 + NB) * kTileBytes;
    static std::atomic<bool> attr_set{false};  // per template instance
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(bsgs_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    launch_k(bsgs_kernel<NB>, grid, kT, smem, st, mods, k, pm);
    return cudaGetLastError();
}

cudaError_t bsgs_matmul(const Mod *mods, const BsgsArgs &a, PrimeMap pm, size_t N, cudaStream_t st) {
    // 63 products < 2^122 plus a carried-in residue stay below 2^128
    if (a.g == 0 || a.g > (u32) kMaxBsgs - (a.accumulate ? 1u : 0u) || a.nb == 0) return cudaErrorInvalidValue;
    if (N % (2 * kT) || a.pt_shift > 8 || a.limb0 + a.nlimbs > a.size_QlP) return cudaErrorInvalidValue;
    const u32 rows = a.nlimbs ? a.nlimbs : a.size_QlP;
    {   // rotated ciphertexts once for all baby steps, every non-null plaintext once, nb results out (+ in when accumulating)
        double npt = 0;
        for (u32 i = 0; i < a.nb; i++)
            for (u32 j = 0; j < a.g; j++) npt += a.pt[size_t(i) * a.g + j] ? 1 : 0;
        trace_add(8.0 * N * rows * (2.0 * a.g + npt / double(1u << a.pt_shift) + (a.accumulate ? 4.0 : 2.0) * a.nb), 0,
                  2.0 * N * rows * npt);
    }
    dim3 grid((u32) (N / (2 * kT)), rows);
    u32 i0 = 0;
    while (i0 < a.nb) {
        const u32 left = a.nb - i0;
        cudaError_t e;
        if (left >= 4) {
            e = bsgs_launch<4>(mods, a, i0, pm, grid, st);
            i0 += 4;
        } else if (left >= 2) {
            e = bsgs_launch<2>(mods, a, i0, pm, grid, st);
            i0 += 2;
        } else {
            e = bsgs_launch<1>(mods, a, i0, pm, grid, st);
            i0 += 1;
        }
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace b200fhe
