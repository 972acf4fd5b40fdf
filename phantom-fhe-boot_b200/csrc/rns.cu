// rns.cu — RNS kernels for sm_100a: base conversion, evaluation-key inner product, tensor product,
// rescale helper, automorphism gather, element-wise ops.  Each thread owns a PAIR of adjacent
// coefficients, so every global access is a 16-byte vector and consecutive lanes touch consecutive
// 16 B (fully coalesced 512 B per warp) — including the base-conversion stores, which in the
// reference are strided by N because adjacent threads own different output primes
// (src/rns_bconv.cu:143-168).
#include "rns.cuh"

#include "launch.cuh"

#include <atomic>
#include <cmath>

namespace b200fhe {

static constexpr int kT = 256;

// ---------------------------------------------------------------- base conversion
static u32 pick_split(u32 ctas, u32 no);
template <int NI>
__global__ void __launch_bounds__(kT) bconv_kernel(const Mod *__restrict__ mods, BconvArgs a, u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ u64 smat[];  // [no][NI]
    for (u32 i = threadIdx.x; i < a.no * NI; i += blockDim.x) smat[i] = a.mat[size_t(a.row(i / NI)) * NI + i % NI];
    __syncthreads();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2;
    const u64 *in = a.in + blockIdx.y * a.in_stride + 2 * size_t(pair);
    u64 *out = a.out + blockIdx.y * a.out_stride + 2 * size_t(pair);
    u64 x0[NI], x1[NI];
#pragma unroll
    for (int i = 0; i < NI; i++) {
        ulonglong2 v = ld2(in + size_t(i) * N);
        x0[i] = v.x;
        x1[i] = v.y;
    }
    // grid.z splits the OUTPUT limbs: a conversion of one or two polynomials alone is only N/512 CTAs
    const u32 chunk = (a.no + gridDim.z - 1) / gridDim.z;
    const u32 j_end = min(a.no, (blockIdx.z + 1) * chunk);
#pragma unroll 2
    for (u32 j = blockIdx.z * chunk; j < j_end; j++) {  // two output limbs in flight: four independent carry chains
        const u32 ja = a.row(j);
        const u32 pos = ja < a.skip_start ? ja : ja + a.skip_len;
        const Mod md = mods[a.pm.prime(pos)];
        const u64 *row = smat + j * NI;
        u128 acc0 = {0, 0}, acc1 = {0, 0};
#pragma unroll
        for (int i = 0; i < NI; i++) {
            const u64 m = row[i];
            mac128(acc0, x0[i], m);
            mac128(acc1, x1[i], m);
        }
        if (a.lazy) st2(out + size_t(pos) * N, barrett128_lazy(acc0, md), barrett128_lazy(acc1, md));
        else st2(out + size_t(pos) * N, barrett128(acc0, md), barrett128(acc1, md));
    }
}

// generic input count (slow path, inputs re-read through L1)
__global__ void __launch_bounds__(kT) bconv_kernel_generic(const Mod *__restrict__ mods, BconvArgs a, u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ u64 smat[];
    for (u32 i = threadIdx.x; i < a.no * a.ni; i += blockDim.x) smat[i] = a.mat[size_t(a.row(i / a.ni)) * a.ni + i % a.ni];
    __syncthreads();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2;
    const u64 *in = a.in + blockIdx.y * a.in_stride + 2 * size_t(pair);
    u64 *out = a.out + blockIdx.y * a.out_stride + 2 * size_t(pair);
    const u32 chunk = (a.no + gridDim.z - 1) / gridDim.z;
    const u32 j_end = min(a.no, (blockIdx.z + 1) * chunk);
    for (u32 j = blockIdx.z * chunk; j < j_end; j++) {
        const u32 ja = a.row(j);
        const u32 pos = ja < a.skip_start ? ja : ja + a.skip_len;
        const Mod md = mods[a.pm.prime(pos)];
        u128 acc0 = {0, 0}, acc1 = {0, 0};
        for (u32 i = 0; i < a.ni; i++) {
            ulonglong2 v = ld2(in + size_t(i) * N);
            const u64 m = smat[j * a.ni + i];
            mac128(acc0, v.x, m);
            mac128(acc1, v.y, m);
        }
        if (a.lazy) st2(out + size_t(pos) * N, barrett128_lazy(acc0, md), barrett128_lazy(acc1, md));
        else st2(out + size_t(pos) * N, barrett128(acc0, md), barrett128(acc1, md));
    }
}

cudaError_t bconv(const Mod *mods, const BconvArgs &a, size_t N, cudaStream_t st) {
    if (a.no == 0 || a.batch == 0) return cudaSuccess;
    trace_add(8.0 * N * a.batch * (a.ni + a.no), 0, double(N) * a.batch * a.ni * a.no);
    const u32 n_pairs = (u32) (N / 2);
    const u32 ctas = ((n_pairs + kT - 1) / kT) * a.batch;
    const u32 split = pick_split(ctas, a.no);  // each CTA re-reads the ni input limbs (L2 hits)
    dim3 grid((n_pairs + kT - 1) / kT, a.batch, split);
    const size_t sm = size_t(a.no) * a.ni * sizeof(u64);
#define B200_BCONV_CASE(K) \
    case K: launch_k(bconv_kernel<K>, grid, kT, sm, st, mods, a, n_pairs); break;
    switch (a.ni) {
        B200_BCONV_CASE(1) B200_BCONV_CASE(2) B200_BCONV_CASE(3) B200_BCONV_CASE(4) B200_BCONV_CASE(5)
        B200_BCONV_CASE(6) B200_BCONV_CASE(7) B200_BCONV_CASE(8) B200_BCONV_CASE(9) B200_BCONV_CASE(10)
        B200_BCONV_CASE(11) B200_BCONV_CASE(12) B200_BCONV_CASE(13) B200_BCONV_CASE(14) B200_BCONV_CASE(15)
        B200_BCONV_CASE(16)
        default: launch_k(bconv_kernel_generic, grid, kT, sm, st, mods, a, n_pairs); break;
    }
#undef B200_BCONV_CASE
    return cudaGetLastError();
}

// wave-aware split of the output limbs over grid.z: the conversion is integer-bound, so what matters is how full
// the last wave is (3 CTAs of 256 threads per SM at 80 registers)
static u32 pick_split(u32 ctas, u32 no) {
    static std::atomic<int> cached{0};
    int sms = cached.load(std::memory_order_relaxed);
    if (!sms) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
            sms = 148;
        cached.store(sms, std::memory_order_relaxed);
    }
    const double slots = 3.0 * sms;
    u32 best = 1;
    double best_eff = 0;
    for (u32 s = 1; s <= 8 && s <= no; s *= 2) {
        const double waves = ctas * s / slots;
        const double eff = waves / std::ceil(waves) - 0.01 * (s > 1 ? std::log2((double) s) : 0);  // re-reads cost a little
        if (eff > best_eff + 1e-9) {
            best_eff = eff;
            best = s;
        }
    }
    return best;
}

template <int NI>
__global__ void __launch_bounds__(kT) bconv_multi_kernel(const Mod *__restrict__ mods,
                                                         const __grid_constant__ BconvMultiArgs A, u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ u64 smat[];  // [no][ni] of this CTA's digit (+ NI words of padding)
    const u32 batch = A.d[0].batch;
    const u32 dg = blockIdx.y / batch, be = blockIdx.y % batch;
    const BconvArgs &a = A.d[dg];
    const u32 ni = a.ni;
    for (u32 i = threadIdx.x; i < a.no * ni + NI; i += blockDim.x) smat[i] = i < a.no * ni ? a.mat[size_t(a.row(i / ni)) * ni + i % ni] : 0;
    __syncthreads();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2;
    const u64 *in = a.in + be * a.in_stride + 2 * size_t(pair);
    u64 *out = a.out + be * a.out_stride + 2 * size_t(pair);
    u64 x0[NI], x1[NI];
#pragma unroll
    for (int i = 0; i < NI; i++) {
        ulonglong2 v = make_ulonglong2(0, 0);
        if (i < (int) ni) v = ld2(in + size_t(i) * N);
        x0[i] = v.x;
        x1[i] = v.y;
    }
    if (blockIdx.z == 0 && A.own_src[dg]) {  // the digit's own limbs stay as they came
        const u64 *src = A.own_src[dg] + be * A.own_stride + 2 * size_t(pair);
        u64 *dst = out + size_t(a.skip_start) * N;
        const u32 i_lo = A.own_count[dg] ? A.own_first[dg] : 0, i_hi = A.own_count[dg] ? i_lo + A.own_count[dg] : ni;
        for (u32 i = i_lo; i < i_hi; i++) {
            const ulonglong2 v = ld2(src + size_t(i) * N);
            st2(dst + size_t(i) * N, v.x, v.y);
        }
    }
    const u32 chunk = (a.no + gridDim.z - 1) / gridDim.z;
    const u32 j_end = min(a.no, (blockIdx.z + 1) * chunk);
#pragma unroll 2
    for (u32 j = blockIdx.z * chunk; j < j_end; j++) {
        const u32 ja = a.row(j);
        const u32 pos = ja < a.skip_start ? ja : ja + a.skip_len;
        const Mod md = mods[a.pm.prime(pos)];
        const u64 *row = smat + j * ni;
        u128 acc0 = {0, 0}, acc1 = {0, 0};
#pragma unroll
        for (int i = 0; i < NI; i++) {
            if (i >= (int) ni) break;  // uniform over the CTA: a short last digit does not pay for alpha terms
            const u64 m = row[i];
            mac128(acc0, x0[i], m);
            mac128(acc1, x1[i], m);
        }
        if (a.lazy) st2(out + size_t(pos) * N, barrett128_lazy(acc0, md), barrett128_lazy(acc1, md));
        else st2(out + size_t(pos) * N, barrett128(acc0, md), barrett128(acc1, md));
    }
}

cudaError_t bconv_multi(const Mod *mods, const BconvMultiArgs &A, size_t N, cudaStream_t st) {
    if (A.ndig == 0 || A.ndig > (u32) kBconvDigits || A.d[0].batch == 0) return cudaErrorInvalidValue;
    u32 ni = 0, no = 0;
    for (u32 d = 0; d < A.ndig; d++) {
        ni = max(ni, A.d[d].ni);
        no = max(no, A.d[d].no);
        if (A.d[d].batch != A.d[0].batch) return cudaErrorInvalidValue;
    }
    if (ni < 1 || ni > 16) return cudaErrorInvalidValue;
    for (u32 d = 0; d < A.ndig; d++) {
        const double own = A.own_src[d] ? (A.own_count[d] ? A.own_count[d] : A.d[d].skip_len) : 0;
        trace_add(8.0 * N * A.d[d].batch * (A.d[d].ni + A.d[d].no + 2 * own), 0, double(N) * A.d[d].batch * A.d[d].ni * A.d[d].no);
    }
    const u32 n_pairs = (u32) (N / 2);
    const u32 gx = (n_pairs + kT - 1) / kT, gy = A.ndig * A.d[0].batch;
    dim3 grid(gx, gy, pick_split(gx * gy, no));
    const size_t sm = (size_t(no) * ni + ni) * sizeof(u64);
#define B200_BCONVM_CASE(K) \
    case K: launch_k(bconv_multi_kernel<K>, grid, kT, sm, st, mods, A, n_pairs); break;
    switch (ni) {
        B200_BCONVM_CASE(1) B200_BCONVM_CASE(2) B200_BCONVM_CASE(3) B200_BCONVM_CASE(4) B200_BCONVM_CASE(5) B200_BCONVM_CASE(6)
        B200_BCONVM_CASE(7) B200_BCONVM_CASE(8) B200_BCONVM_CASE(9) B200_BCONVM_CASE(10) B200_BCONVM_CASE(11)
        B200_BCONVM_CASE(12) B200_BCONVM_CASE(13) B200_BCONVM_CASE(14) B200_BCONVM_CASE(15) B200_BCONVM_CASE(16)
    }
#undef B200_BCONVM_CASE
    return cudaGetLastError();
}

__global__ void __launch_bounds__(kT) bconv_single_kernel(const Mod *__restrict__ mods, BconvArgs a, u32 in_prime_idx,
                                                          u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2;
    const u64 *in = a.in + blockIdx.y * a.in_stride + 2 * size_t(pair);
    u64 *out = a.out + blockIdx.y * a.out_stride + 2 * size_t(pair);
    const ulonglong2 v = ld2(in);
    const u64 qin = mods[in_prime_idx].q;
    for (u32 j = 0; j < a.no; j++) {
        const u32 pos = j < a.skip_start ? j : j + a.skip_len;
        const Mod md = mods[a.pm.prime(pos)];
        u64 r0 = v.x, r1 = v.y;
        if (qin > md.q) {
            r0 = barrett64(r0, md);
            r1 = barrett64(r1, md);
        }
        st2(out + size_t(pos) * N, r0, r1);
    }
}

cudaError_t bconv_single(const Mod *mods, const BconvArgs &a, u32 in_prime_idx, size_t N, cudaStream_t st) {
    if (a.no == 0 || a.batch == 0) return cudaSuccess;
    trace_add(8.0 * N * a.batch * (1 + a.no), 0, double(N) * a.batch * a.no);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, a.batch);
    launch_k(bconv_single_kernel, grid, kT, 0, st, mods, a, in_prime_idx, n_pairs);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- evaluation-key inner product
template <bool DEVTAB>
__global__ void __launch_bounds__(kT) inner_prod_kernel(const Mod *__restrict__ mods, u64 *__restrict__ cx,
                                                        const u64 *__restrict__ up, const __grid_constant__ EvkPtrs evk,
                                                        u32 beta, u32 size_QlP, u32 size_QP, PrimeMap pm, u32 n_pairs,
                                                        u32 t0, u32 tn, u32 t0b) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const u32 t = blockIdx.y < tn ? t0 + blockIdx.y : t0b + (blockIdx.y - tn);  // second run of rows: limb-sharded key switch
    const size_t N = size_t(n_pairs) * 2;
    const u32 twr = pm.prime(t);
    const Mod md = mods[twr];
    const size_t cidx = 2 * size_t(pair);
    const size_t up_off = size_t(t) * N + cidx, up_stride = size_t(size_QlP) * N;
    up += size_t(blockIdx.z) * beta * up_stride;
    cx += size_t(blockIdx.z) * 2 * up_stride;
    const size_t k_off = size_t(twr) * N + cidx, k_poly = size_t(size_QP) * N;
    u128 a00 = {0, 0}, a01 = {0, 0}, a10 = {0, 0}, a11 = {0, 0};
    for (u32 d = 0; d < beta; d++) {
        const ulonglong2 v = ld2(up + d * up_stride + up_off);
        const u64 *k = (DEVTAB ? reinterpret_cast<const u64 *>(__ldg(reinterpret_cast<const u64 *>(evk.dev_table) + d)) : evk.d[d]) + k_off;
        const ulonglong2 k0 = ldg2(k), k1 = ldg2(k + k_poly);
        mac128(a00, v.x, k0.x);
        mac128(a01, v.y, k0.y);
        mac128(a10, v.x, k1.x);
        mac128(a11, v.y, k1.y);
    }
    st2(cx + up_off, barrett128(a00, md), barrett128(a01, md));
    st2(cx + up_stride + up_off, barrett128(a10, md), barrett128(a11, md));
}

cudaError_t inner_prod(const Mod *mods, u64 *cx, const u64 *up, const EvkPtrs &evk, u32 beta, u32 size_QlP,
                       u32 size_QP, PrimeMap pm, size_t N, cudaStream_t st, u32 batch) {
    if (batch == 0) return cudaSuccess;
    // digits in, both key polynomials in, both accumulators out
    trace_add(8.0 * N * size_QlP * (double(batch) * beta + 2.0 * beta + 2.0 * batch), 0, 2.0 * N * size_QlP * beta * batch);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, size_QlP, batch);
    if (evk.dev_table)
        launch_k(inner_prod_kernel<true>, grid, kT, 0, st, mods, cx, up, evk, beta, size_QlP, size_QP, pm, n_pairs, 0, size_QlP, 0);
    else
        launch_k(inner_prod_kernel<false>, grid, kT, 0, st, mods, cx, up, evk, beta, size_QlP, size_QP, pm, n_pairs, 0, size_QlP, 0);
    return cudaGetLastError();
}

cudaError_t inner_prod_range(const Mod *mods, u64 *cx, const u64 *up, const EvkPtrs &evk, u32 beta, u32 size_QlP,
                             u32 size_QP, PrimeMap pm, u32 t0, u32 tn, size_t N, cudaStream_t st, u32 t0b, u32 tnb, u32 batch) {
    if (tn + tnb == 0 || batch == 0) return cudaSuccess;
    trace_add(8.0 * N * (tn + tnb) * (double(batch) * beta + 2.0 * beta + 2.0 * batch), 0, 2.0 * N * (tn + tnb) * beta * batch);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, tn + tnb, batch);
    if (evk.dev_table)
        launch_k(inner_prod_kernel<true>, grid, kT, 0, st, mods, cx, up, evk, beta, size_QlP, size_QP, pm, n_pairs, t0, tn, t0b);
    else
        launch_k(inner_prod_kernel<false>, grid, kT, 0, st, mods, cx, up, evk, beta, size_QlP, size_QP, pm, n_pairs, t0, tn, t0b);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- tensor product
template <bool SQUARE>
__global__ void __launch_bounds__(kT) tensor_kernel(const Mod *__restrict__ mods, u64 *out, const u64 *a, const u64 *b,
                                                    u32 L, u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2, LN = size_t(L) * N;
    const Mod md = mods[blockIdx.y];
    const size_t id = size_t(blockIdx.y) * N + 2 * size_t(pair);
    const ulonglong2 a0 = ld2(a + id), a1 = ld2(a + id + LN);
    u64 d0[2], d1[2], d2[2];
    if (SQUARE) {
        const u64 x0[2] = {a0.x, a0.y}, x1[2] = {a1.x, a1.y};
#pragma unroll
        for (int e = 0; e < 2; e++) {
            d0[e] = mulmod(x0[e], x0[e], md);
            u128 p = mul128(x0[e], x1[e]);  // 2*c0*c1 as one 128-bit value (polymath.cu:556-559)
            p.hi = (p.hi << 1) | (p.lo >> 63);
            p.lo <<= 1;
            d1[e] = barrett128(p, md);
            d2[e] = mulmod(x1[e], x1[e], md);
        }
    } else {
        const ulonglong2 b0 = ld2(b + id), b1 = ld2(b + id + LN);
        const u64 x0[2] = {a0.x, a0.y}, x1[2] = {a1.x, a1.y}, y0[2] = {b0.x, b0.y}, y1[2] = {b1.x, b1.y};
#pragma unroll
        for (int e = 0; e < 2; e++) {
            d0[e] = mulmod(x0[e], y0[e], md);
            d2[e] = mulmod(x1[e], y1[e], md);
            u64 m = mulmod(x0[e] + x1[e], y0[e] + y1[e], md);  // Karatsuba middle term
            m = m + 2 * md.q - d0[e] - d2[e];
            d1[e] = csub(csub(m, md.q), md.q);
        }
    }
    st2(out + id, d0[0], d0[1]);
    st2(out + id + LN, d1[0], d1[1]);
    st2(out + id + 2 * LN, d2[0], d2[1]);
}

cudaError_t tensor_prod(const Mod *mods, u64 *out, const u64 *a, const u64 *b, u32 L, size_t N, cudaStream_t st) {
    trace_add(8.0 * N * L * 7, 0, 3.0 * N * L);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, L);
    launch_k(tensor_kernel<false>, grid, kT, 0, st, mods, out, a, b, L, n_pairs);
    return cudaGetLastError();
}

cudaError_t tensor_square(const Mod *mods, u64 *out, const u64 *a, u32 L, size_t N, cudaStream_t st) {
    trace_add(8.0 * N * L * 5, 0, 3.0 * N * L);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, L);
    launch_k(tensor_kernel<true>, grid, kT, 0, st, mods, out, a, nullptr, L, n_pairs);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- rescale helper
__global__ void __launch_bounds__(kT) reduce_last_kernel(const Mod *__restrict__ mods, u64 *out, const u64 *in, u32 L,
                                                         size_t in_stride, size_t out_stride, u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2;
    const ulonglong2 v = ld2(in + blockIdx.y * in_stride + 2 * size_t(pair));
    u64 *o = out + blockIdx.y * out_stride + 2 * size_t(pair);
    for (u32 j = 0; j < L; j++) {
        const Mod md = mods[j];
        st2(o + size_t(j) * N, barrett64(v.x, md), barrett64(v.y, md));
    }
}

cudaError_t reduce_last(const Mod *mods, u64 *out, const u64 *in, u32 L, u32 batch, size_t in_stride,
                        size_t out_stride, size_t N, cudaStream_t st) {
    if (L == 0 || batch == 0) return cudaSuccess;
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, batch);
    trace_add(8.0 * N * batch * (1 + L), 0, double(N) * batch * L);
    launch_k(reduce_last_kernel, grid, kT, 0, st, mods, out, in, L, in_stride, out_stride, n_pairs);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- automorphism
__global__ void __launch_bounds__(kT) gather_kernel(u64 *dst, const u64 *src, const u32 *__restrict__ perm, u32 limbs,
                                                    size_t src_stride, size_t dst_stride, u32 N) {
    pdl_wait();
    pdl_trigger();
    const u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= N) return;
    const u32 pc = __ldg(perm + c);
    const u64 *s = src + blockIdx.z * src_stride + size_t(blockIdx.y) * N;
    u64 *d = dst + blockIdx.z * dst_stride + size_t(blockIdx.y) * N;
    d[c] = s[pc];
}

cudaError_t gather_limbs(u64 *dst, const u64 *src, const u32 *perm, u32 limbs, u32 batch, size_t src_stride,
                         size_t dst_stride, size_t N, cudaStream_t st) {
    if (limbs == 0 || batch == 0) return cudaSuccess;
    dim3 grid((u32) ((N + kT - 1) / kT), limbs, batch);
    trace_add(16.0 * N * limbs * batch, 0, 0);
    launch_k(gather_kernel, grid, kT, 0, st, dst, src, perm, limbs, src_stride, dst_stride, (u32) N);
    return cudaGetLastError();
}

__global__ void automap_kernel(u32 *perm, int logN, u32 k) {
    pdl_wait();
    pdl_trigger();
    const u32 n = 1u << logN;
    const u32 j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const u32 m_mask = (n << 1) - 1;
    const u32 idx = (((j << 1) + 1) * k & m_mask) >> 1;
    perm[__brev(j) >> (32 - logN)] = __brev(idx) >> (32 - logN);
}

cudaError_t automap_table(u32 *perm, int logN, u32 k, cudaStream_t st) {
    const u32 n = 1u << logN;
    launch_k(automap_kernel, (n + kT - 1) / kT, kT, 0, st, perm, logN, k);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- element-wise
__global__ void __launch_bounds__(kT) eltwise_kernel(const Mod *__restrict__ mods, int op, u64 *out, const u64 *a,
                                                     const u64 *b, PrimeMap pm, u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2;
    const Mod md = mods[pm.prime(blockIdx.y)];
    const size_t id = size_t(blockIdx.y) * N + 2 * size_t(pair);
    const ulonglong2 x = ld2(a + id);
    ulonglong2 y = make_ulonglong2(0, 0);
    if (op != ELT_NEG) y = ld2(b + id);
    u64 r0, r1;
    switch (op) {
        case ELT_ADD: r0 = addmod(x.x, y.x, md.q); r1 = addmod(x.y, y.y, md.q); break;
        case ELT_SUB: r0 = submod(x.x, y.x, md.q); r1 = submod(x.y, y.y, md.q); break;
        case ELT_MUL: r0 = mulmod(x.x, y.x, md); r1 = mulmod(x.y, y.y, md); break;
        default: r0 = x.x ? md.q - x.x : 0; r1 = x.y ? md.q - x.y : 0; break;
    }
    st2(out + id, r0, r1);
}

cudaError_t eltwise(const Mod *mods, EltOp op, u64 *out, const u64 *a, const u64 *b, u32 limbs, PrimeMap pm, size_t N,
                    cudaStream_t st) {
    if (limbs == 0) return cudaSuccess;
    trace_add(8.0 * N * limbs * (op == ELT_NEG ? 2 : 3), 0, op == ELT_MUL ? double(N) * limbs : 0);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, limbs);
    launch_k(eltwise_kernel, grid, kT, 0, st, mods, (int) op, out, a, b, pm, n_pairs);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(kT) scalar_mul_kernel(const Mod *__restrict__ mods, u64 *out, const u64 *a,
                                                        const u64 *s, const u64 *ss, PrimeMap pm, u32 n_pairs) {
    pdl_wait();
    pdl_trigger();
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const size_t N = size_t(n_pairs) * 2;
    const Mod md = mods[pm.prime(blockIdx.y)];
    const size_t id = size_t(blockIdx.y) * N + 2 * size_t(pair);
    const ulonglong2 x = ld2(a + id);
    const u64 w = s[blockIdx.y], ws = ss[blockIdx.y];
    st2(out + id, shoup_mul(x.x, w, ws, md.q, md.nq), shoup_mul(x.y, w, ws, md.q, md.nq));
}

cudaError_t scalar_mul(const Mod *mods, u64 *out, const u64 *a, const u64 *s, const u64 *s_shoup, u32 limbs,
                       PrimeMap pm, size_t N, cudaStream_t st) {
    if (limbs == 0) return cudaSuccess;
    trace_add(16.0 * N * limbs, 0, double(N) * limbs);
    const u32 n_pairs = (u32) (N / 2);
    dim3 grid((n_pairs + kT - 1) / kT, limbs);
    launch_k(scalar_mul_kernel, grid, kT, 0, st, mods, out, a, s, s_shoup, pm, n_pairs);
    return cudaGetLastError();
}

}  // namespace b200fhe
