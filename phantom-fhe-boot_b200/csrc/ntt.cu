// ntt.cu — forward / inverse negacyclic NTT over RNS limbs for sm_100a.
//
// Replaces the reference's radix-8 2-D kernels (src/ntt/fntt_2d.cu, intt_2d.cu, ntt_modup.cu,
// ntt_moddown.cu) with a two-pass radix-16 design:
//   N = R x 256 (R = N/256 rows).  Forward: pass 1 = the first log2(R) stages on columns (stride 256),
//   pass 2 = the last 8 stages on contiguous 256-element rows.  Inverse mirrors it.
//   Every thread owns 16 coefficients in registers and runs up to 4 stages per "round"; a pass is
//   at most two rounds with ONE shared-memory exchange (the reference: 8 per thread, 2-3 exchanges).
//   All global accesses of a warp are full 128 B segments; pass 2 stores 16 B per lane.
//   One launch covers all limbs of all polynomials of a call (grid.y = limb, grid.z = polynomial).
// Arithmetic: Harvey lazy butterflies in [0,4q) / [0,2q) with Shoup twiddles (arith.cuh), final
// conditional subtractions to the canonical range, so outputs are bit-identical to the reference.
#include "ntt.cuh"

#include <algorithm>
#include <atomic>

#include <cooperative_groups.h>
#include <cstdlib>

#include "launch.cuh"
#include "tma.cuh"

namespace b200fhe {

#ifndef B200FHE_NTT_TPB
#define B200FHE_NTT_TPB 128
#endif
static constexpr int kThreads = B200FHE_NTT_TPB;   // CTA size of every pass: 128 (7 CTAs/SM, finer waves; +7 % over 256 on the B200)
static constexpr int kRows = kThreads / 16;         // rows of 256 coefficients a row-pass CTA owns

__device__ __forceinline__ u32 prime_of(const LimbMap &m, u32 t) {
    return m.size_Ql == 0 ? t : (t < m.size_Ql ? t : m.size_Q + (t - m.size_Ql));
}

__device__ __forceinline__ bool excluded(const LimbMap &m, u32 t, u32 z) {
    if (t >= m.excl_lo && t < m.excl_hi) return true;
    if (m.excl_alpha == 0) return false;
    u32 s = (m.excl_mod ? z % m.excl_mod : z) * m.excl_alpha;
    u32 e = min(s + m.excl_alpha, m.size_Ql);
    return t >= s && t < e;
}

// ---- register rounds over 16 coefficients -------------------------------------------------------
// stage u pairs k and k + (8 >> u); twiddle index (B << u) + (k >> (4 - u)).

template <int NST>
__device__ __forceinline__ void ct_round(u64 (&x)[16], const ulonglong2 *__restrict__ tw, u32 B, u64 q2, u64 nq) {
#pragma unroll
    for (int u = 0; u < NST; u++) {
        const int h = 8 >> u;
#pragma unroll
        for (int g = 0; g < (1 << u); g++) {
            ulonglong2 w = __ldg(tw + ((B << u) + g));
#pragma unroll
            for (int j = 0; j < h; j++) {
                const int k = g * 2 * h + j;
                ct_bfly(x[k], x[k + h], w.x, w.y, q2, nq);
            }
        }
    }
}

template <int NST>
__device__ __forceinline__ void gs_round(u64 (&x)[16], const ulonglong2 *__restrict__ tw, u32 B, u64 q2, u64 nq) {
#pragma unroll
    for (int u = NST - 1; u >= 0; u--) {
        const int h = 8 >> u;
#pragma unroll
        for (int g = 0; g < (1 << u); g++) {
            ulonglong2 w = __ldg(tw + ((B << u) + g));
#pragma unroll
            for (int j = 0; j < h; j++) {
                const int k = g * 2 * h + j;
                gs_bfly(x[k], x[k + h], w.x, w.y, q2, nq);
            }
        }
    }
}

// shared-memory position of column c inside a 256-element row (u64 units): 16-byte pairs are kept
// together and rotated inside each 128 B chunk so both the strided (c = lane + 16k, 8 B) and the
// contiguous (c = 16*lane + k, 16 B) access patterns are bank-conflict free.
__device__ __forceinline__ u32 swz(u32 c) { return (c & ~14u) | ((c ^ ((c >> 4) << 1)) & 14u); }

// ---- forward pass 1: columns ---------------------------------------------------------------------
// RA = log2(R) - 4 in [0,4].  CTA = 256 threads = C columns x 2^RA row-threads, C = 256 >> RA.
// logsub > 0 (N = 2^17): the first logsub stages were done by ntt_fwd_stage0; the CTAs of a limb split into 2^logsub
// groups, group h transforms the contiguous block h of N >> logsub coefficients with the sub-tree rooted at twiddle
// node (1 << logsub) + h.
template <int RA>
__global__ void __launch_bounds__(kThreads) ntt_fwd_p1(NttTables T, u64 *dst, const u64 *src, LimbMap map, u32 logsub) {
    pdl_wait();
    pdl_trigger();
    constexpr int C = kThreads >> RA;
    __shared__ u64 tile[RA > 0 ? 16 * kThreads : 1];
    const u32 t = map.limb(blockIdx.y), z = blockIdx.z;
    if (excluded(map, t, z)) return;
    const u32 p = prime_of(map, t);
    const size_t N = size_t(1) << T.logN;
    const Mod md = T.mods[p];
    const u64 q2 = 2 * md.q, nq = md.nq;
    const ulonglong2 *tw = T.tw + size_t(p) * N;
    const u32 cc = threadIdx.x % C, rt = threadIdx.x / C;
    const u32 tps = gridDim.x >> logsub, h = blockIdx.x / tps, B0 = (1u << logsub) + h;
    const size_t col = size_t(h) * (N >> logsub) + size_t(blockIdx.x % tps) * C + cc;
    const u64 *in = src + z * map.src_poly_stride + (map.src_bcast ? 0 : size_t(t) * N) + col;
    u64 *out = dst + z * map.poly_stride + size_t(t) * N + col;
    u64 x[16];
    if (RA > 0) {
        // round A: rows r = (ka << 4) | (kb << RA) | rt,  k = (ka << (4-RA)) | kb
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const u32 ka = k >> (4 - RA), kb = k & ((1 << (4 - RA)) - 1);
            const u32 r = (ka << 4) | (kb << RA) | rt;
            x[k] = in[size_t(r) * 256];
        }
        ct_round<RA>(x, tw, B0, q2, nq);
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const u32 ka = k >> (4 - RA), kb = k & ((1 << (4 - RA)) - 1);
            const u32 r = (ka << 4) | (kb << RA) | rt;
            tile[r * C + cc] = x[k];
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; k++) x[k] = tile[((rt << 4) | k) * C + cc];
    } else {
#pragma unroll
        for (int k = 0; k < 16; k++) x[k] = in[size_t(k) * 256];
    }
    // round B: rows r = (rt << 4) | k, stages RA..RA+3
    ct_round<4>(x, tw, (B0 << RA) + rt, q2, nq);
#pragma unroll
    for (int k = 0; k < 16; k++) out[size_t((rt << 4) | k) * 256] = x[k];
}


// ---- N = 2^17: one extra radix-2 stage around the two-pass kernels ------------------------------------
// Forward: stage 0 pairs c with c + N/2 (twiddle node 1); values stay lazy in [0,4q).  16 bytes per thread.
__global__ void __launch_bounds__(256) ntt_fwd_stage0(NttTables T, u64 *dst, const u64 *src, LimbMap map) {
    pdl_wait();
    pdl_trigger();
    const u32 t = map.limb(blockIdx.y), z = blockIdx.z;
    if (excluded(map, t, z)) return;
    const u32 p = prime_of(map, t);
    const size_t N = size_t(1) << T.logN, half = N >> 1;
    const Mod md = T.mods[p];
    const u64 q2 = 2 * md.q, nq = md.nq;
    const ulonglong2 w = __ldg(T.tw + size_t(p) * N + 1);
    const size_t c = 2 * (size_t(blockIdx.x) * blockDim.x + threadIdx.x);
    const u64 *in = src + z * map.src_poly_stride + (map.src_bcast ? 0 : size_t(t) * N) + c;
    u64 *out = dst + z * map.poly_stride + size_t(t) * N + c;
    ulonglong2 a = ld2(in), b = ld2(in + half);
    ct_bfly(a.x, b.x, w.x, w.y, q2, nq);
    ct_bfly(a.y, b.y, w.x, w.y, q2, nq);
    st2(out, a.x, a.y);
    st2(out + half, b.x, b.y);
}
// Inverse: last stage (node 1 of the inverse table, which carries N^-1), N^-1 on the sum side, canonical output,
// optional per-limb scale.
template <bool SCALE>
__global__ void __launch_bounds__(256) ntt_inv_stage0(NttTables T, u64 *dst, LimbMap map, const u64 *scale,
                                                       const u64 *scale_shoup) {
    pdl_wait();
    pdl_trigger();
    const u32 t = map.limb(blockIdx.y), z = blockIdx.z;
    const u32 p = prime_of(map, t);
    const size_t N = size_t(1) << T.logN, half = N >> 1;
    const Mod md = T.mods[p];
    const u64 q = md.q, q2 = 2 * md.q, nq = md.nq;
    const ulonglong2 w = __ldg(T.itw + size_t(p) * N + 1);
    const u64 ni = T.ninv[p], nis = T.ninv_shoup[p];
    const size_t c = 2 * (size_t(blockIdx.x) * blockDim.x + threadIdx.x);
    u64 *io = dst + z * map.poly_stride + size_t(t) * N + c;
    ulonglong2 a = ld2(io), b = ld2(io + half);
    gs_bfly(a.x, b.x, w.x, w.y, q2, nq);
    gs_bfly(a.y, b.y, w.x, w.y, q2, nq);
    a.x = csub(shoup_lazy(a.x, ni, nis, nq), q);
    a.y = csub(shoup_lazy(a.y, ni, nis, nq), q);
    b.x = csub(b.x, q);
    b.y = csub(b.y, q);
    if (SCALE) {
        const u64 s = scale[t], ss = scale_shoup[t];
        a.x = shoup_mul(a.x, s, ss, q, nq);
        a.y = shoup_mul(a.y, s, ss, q, nq);
        b.x = shoup_mul(b.x, s, ss, q, nq);
        b.y = shoup_mul(b.y, s, ss, q, nq);
    }
    st2(io, a.x, a.y);
    st2(io + half, b.x, b.y);
}

// ---- forward pass 2: rows --------------------------------------------------------------------------
// CTA = kThreads = kRows rows x 16 lanes = kRows*256 CONTIGUOUS coefficients.
// EPI: 0 plain, 1 fused (other - x) * scale, 2 fused (other - x) * scale + acc.  The epilogue operands of a CTA are
// one contiguous 32 KiB block each: thread 0 starts them towards shared memory with a bulk TMA copy before the
// first butterfly and everybody picks them up after the last one, so their HBM latency hides behind eight stages.
static constexpr int kP2Tile = kRows * 256;  // u64 per CTA

template <int EPI>
__global__ void __launch_bounds__(kThreads) ntt_fwd_p2(NttTables T, u64 *dst, LimbMap map, FwdEpilogue epi) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ __align__(128) u64 dyn[];
    __shared__ __align__(8) u64 bar;
    u64 *tile = dyn;
    const u32 t = map.limb(blockIdx.y), z = blockIdx.z;
    if (excluded(map, t, z)) return;
    const u32 p = prime_of(map, t);
    const int n1 = T.logN - 8;
    const size_t N = size_t(1) << T.logN;
    const Mod md = T.mods[p];
    const u64 q = md.q, q2 = 2 * md.q, nq = md.nq;
    const ulonglong2 *tw = T.tw + size_t(p) * N;
    const u32 lane = threadIdx.x & 15, rowi = threadIdx.x >> 4;
    const u32 r = blockIdx.x * kRows + rowi;
    u64 *row = dst + z * map.poly_stride + size_t(t) * N + size_t(r) * 256;
    u64 *srow = tile + rowi * 256;
    bool with_acc = false;
    if (EPI != 0) {
        const u32 za = epi.acc_period ? z % epi.acc_period : z;
        with_acc = EPI == 2 && (epi.acc_polys == 0 || za < epi.acc_polys);
        if (threadIdx.x == 0) {
            const size_t base = size_t(t) * N + size_t(blockIdx.x) * kP2Tile;
            mbar_init(&bar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            mbar_expect_tx(&bar, (with_acc ? 2 : 1) * kP2Tile * 8);
            tma_load_1d(dyn + kP2Tile, epi.other + z * epi.other_poly_stride + base, kP2Tile * 8, &bar);
            if (with_acc) tma_load_1d(dyn + 2 * kP2Tile, epi.acc + (epi.acc_by_poly ? z : za) * epi.acc_poly_stride + base, kP2Tile * 8, &bar);
        }
        __syncthreads();
    }
    u64 x[16];
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = row[lane + 16 * k];
    const u32 B = (1u << n1) + r;
    ct_round<4>(x, tw, B, q2, nq);  // stages n1 .. n1+3, c = lane + 16k
#pragma unroll
    for (int k = 0; k < 16; k++) srow[swz(lane + 16 * k)] = x[k];
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 16; k += 2) {
        ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(srow + swz(16 * lane + k));
        x[k] = v.x;
        x[k + 1] = v.y;
    }
    ct_round<4>(x, tw, (B << 4) + lane, q2, nq);  // stages n1+4 .. n1+7, c = 16*lane + k
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 16; k += 2) {
        u64 a = x[k], b = x[k + 1];
        if (EPI == 0) {  // the fused epilogue takes the lazy value (< 4q) as it is
            a = csub(csub(a, q2), q);
            b = csub(csub(b, q2), q);
        }
        *reinterpret_cast<ulonglong2 *>(srow + swz(16 * lane + k)) = make_ulonglong2(a, b);
    }
    __syncwarp();
    if (EPI == 0) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const u32 c = 2 * (lane + 16 * j);
            ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(srow + swz(c));
            *reinterpret_cast<ulonglong2 *>(row + c) = v;
        }
    } else {
        const u64 *orow = dyn + kP2Tile + rowi * 256;
        const u64 *arow = dyn + 2 * kP2Tile + rowi * 256;
        u64 *wrow = epi.out + z * epi.out_poly_stride + size_t(t) * N + size_t(r) * 256;
        const u64 s = epi.scale[t], ss = epi.scale_shoup[t];
        mbar_wait(&bar, 0);
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const u32 c = 2 * (lane + 16 * j);
            ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(srow + swz(c));
            ulonglong2 o = *reinterpret_cast<const ulonglong2 *>(orow + c);
            // (other - x) * scale with x < 4q lazy: other + 4q - x < 5q < 2^64, one Shoup product, one csub
            u64 a = shoup_mul(o.x + 2 * q2 - v.x, s, ss, q, nq);
            u64 b = shoup_mul(o.y + 2 * q2 - v.y, s, ss, q, nq);
            if (EPI == 2 && with_acc) {
                ulonglong2 ac = *reinterpret_cast<const ulonglong2 *>(arow + c);
                a = addmod(a, ac.x, q);
                b = addmod(b, ac.y, q);
            }
            st2(wrow + c, a, b);
        }
    }
}

// ---- forward pass 2 of all digits of a modup + evaluation-key inner product ---------------------------------
// CTA = kIpRows rows of limb t, G groups of 16 * kIpRows threads.  Phase 1: group g runs the eight row stages of one digit
// that was converted into t (G = beta for a P limb, beta - 1 for a Q_l limb, whose own digit is read from the source as it
// is); the lazy results (< 4q) go to shared memory, one tile per group, and every thread asks L2 for the key lines it will
// need, so the keys arrive under the butterflies.  Phase 2 (after the only CTA barrier): 128-bit multiply-accumulate over
// the digits against both key polynomials, two coefficients per lane and step, the steps dealt round-robin to the groups,
// coalesced 16-byte key loads, one Barrett reduction per output.
static constexpr int kIpRows = 4;
static constexpr int kIpGroup = 16 * kIpRows;

template <int BETA, int G>
__global__ void __launch_bounds__(G * kIpGroup) ntt_fwd_p2_ip(NttTables T, const __grid_constant__ FwdInnerProd A, u32 t0) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ __align__(128) u64 dyn[];  // [G][kIpRows][256]
    const u32 t = t0 + blockIdx.y, z = blockIdx.z;
    const u32 p = t < A.size_Ql ? t : A.size_Q + (t - A.size_Ql);
    const int n1 = T.logN - 8;
    const size_t N = size_t(1) << T.logN;
    const Mod md = T.mods[p];
    const u64 q2 = 2 * md.q, nq = md.nq;
    const ulonglong2 *tw = T.tw + size_t(p) * N;
    const u32 g = threadIdx.x / kIpGroup, lane = threadIdx.x & 15, rowi = (threadIdx.x % kIpGroup) >> 4;
    const u32 r = blockIdx.x * kIpRows + rowi;
    const u32 od = G == BETA ? (u32) BETA : t / A.alpha;  // G < BETA: the digit t belongs to (not transformed)
    const size_t roff = size_t(t) * N + size_t(r) * 256, dstride = size_t(A.size_QlP) * N;
    const size_t koff = size_t(p) * N + size_t(r) * 256, kpoly = size_t(A.size_QP) * N;
    // phase 2 steps of this thread: j = g, g + G, ...
#pragma unroll
    for (int d = 0; d < BETA; d++)
        for (u32 j = g; j < 8; j += G) {
            const u64 *k = A.key[d] + koff + 2 * (lane + 16 * j);
            asm volatile("prefetch.global.L2 [%0];" ::"l"(k));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(k + kpoly));
        }
    {
        const u32 d = g < od ? g : g + 1;
        const u64 *row = A.up + z * A.up_stride + d * dstride + roff;
        u64 *srow = dyn + (g * kIpRows + rowi) * 256;
        const u32 B = (1u << n1) + r;
        u64 x[16];
#pragma unroll
        for (int k = 0; k < 16; k++) x[k] = row[lane + 16 * k];
        ct_round<4>(x, tw, B, q2, nq);
#pragma unroll
        for (int k = 0; k < 16; k++) srow[swz(lane + 16 * k)] = x[k];
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 16; k += 2) {
            ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(srow + swz(16 * lane + k));
            x[k] = v.x;
            x[k + 1] = v.y;
        }
        ct_round<4>(x, tw, (B << 4) + lane, q2, nq);
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 16; k += 2)
            *reinterpret_cast<ulonglong2 *>(srow + swz(16 * lane + k)) = make_ulonglong2(x[k], x[k + 1]);
    }
    __syncthreads();
    const u64 *own = A.own + z * A.own_stride + roff;
    u64 *c0 = A.cx + z * 2 * dstride + roff, *c1 = c0 + dstride;
#pragma unroll 1
    for (u32 j = g; j < 8; j += G) {
        const u32 c = 2 * (lane + 16 * j);
        u128 a00 = {0, 0}, a01 = {0, 0}, a10 = {0, 0}, a11 = {0, 0};
#pragma unroll
        for (int d = 0; d < BETA; d++) {
            ulonglong2 v;
            if (G < BETA && (u32) d == od) v = ld2(own + c);
            else v = *reinterpret_cast<const ulonglong2 *>(dyn + (((u32) d < od ? d : d - 1) * kIpRows + rowi) * 256 + swz(c));
            const u64 *k = A.key[d] + koff + c;
            const ulonglong2 k0 = ldg2(k), k1 = ldg2(k + kpoly);
            mac128(a00, v.x, k0.x);
            mac128(a01, v.y, k0.y);
            mac128(a10, v.x, k1.x);
            mac128(a11, v.y, k1.y);
        }
        st2(c0 + c, barrett128(a00, md), barrett128(a01, md));
        st2(c1 + c, barrett128(a10, md), barrett128(a11, md));
    }
}

// ---- inverse pass A: rows (stages n1+7 .. n1) -------------------------------------------------------
__global__ void __launch_bounds__(kThreads) ntt_inv_pA(NttTables T, u64 *dst, const u64 *src, LimbMap map) {
    pdl_wait();
    pdl_trigger();
    __shared__ __align__(16) u64 tile[kRows * 256];
    const u32 t = map.limb(blockIdx.y), z = blockIdx.z;
    const u32 p = prime_of(map, t);
    const int n1 = T.logN - 8;
    const size_t N = size_t(1) << T.logN;
    const Mod md = T.mods[p];
    const u64 q2 = 2 * md.q, nq = md.nq;
    const ulonglong2 *tw = T.itw + size_t(p) * N;
    const u32 lane = threadIdx.x & 15, rowi = threadIdx.x >> 4;
    const u32 r = blockIdx.x * kRows + rowi;
    const u64 *irow = src + z * map.src_poly_stride + size_t(t) * N + size_t(r) * 256;
    u64 *orow = dst + z * map.poly_stride + size_t(t) * N + size_t(r) * 256;
    u64 *srow = tile + rowi * 256;
    u64 x[16];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const u32 c = 2 * (lane + 16 * j);
        *reinterpret_cast<ulonglong2 *>(srow + swz(c)) = ld2(irow + c);
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 16; k += 2) {
        ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(srow + swz(16 * lane + k));
        x[k] = v.x;
        x[k + 1] = v.y;
    }
    const u32 B = (1u << n1) + r;
    gs_round<4>(x, tw, (B << 4) + lane, q2, nq);
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 16; k += 2)
        *reinterpret_cast<ulonglong2 *>(srow + swz(16 * lane + k)) = make_ulonglong2(x[k], x[k + 1]);
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = srow[swz(lane + 16 * k)];
    gs_round<4>(x, tw, B, q2, nq);
#pragma unroll
    for (int k = 0; k < 16; k++) orow[lane + 16 * k] = x[k];
}

// ---- inverse pass B: columns (stages n1-1 .. 0), N^-1 and optional scale ------------------------------
// logsub > 0 (N = 2^17): group h of the CTAs of a limb runs the sub-tree rooted at node (1 << logsub) + h on block h and
// leaves lazy values in [0,2q); ntt_inv_stage0 finishes (last stage, N^-1, scale).
template <int RA, bool SCALE, bool MIRROR = false>
__global__ void __launch_bounds__(kThreads) ntt_inv_pB(NttTables T, u64 *dst, LimbMap map, const u64 *scale,
                                                        const u64 *scale_shoup, PeerMirror mirror = PeerMirror{},
                                                        u32 logsub = 0) {
    pdl_wait();
    pdl_trigger();
    constexpr int C = kThreads >> RA;
    __shared__ u64 tile[RA > 0 ? 16 * kThreads : 1];
    const u32 t = map.limb(blockIdx.y), z = blockIdx.z;
    const u32 p = prime_of(map, t);
    const size_t N = size_t(1) << T.logN;
    const Mod md = T.mods[p];
    const u64 q = md.q, q2 = 2 * md.q, nq = md.nq;
    const ulonglong2 *tw = T.itw + size_t(p) * N;
    const u32 cc = threadIdx.x % C, rt = threadIdx.x / C;
    const u32 tps = gridDim.x >> logsub, h = blockIdx.x / tps, B0 = (1u << logsub) + h;
    const size_t col = size_t(h) * (N >> logsub) + size_t(blockIdx.x % tps) * C + cc;
    u64 *io = dst + z * map.poly_stride + size_t(t) * N + col;
    u64 x[16];
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = io[size_t((rt << 4) | k) * 256];
    gs_round<4>(x, tw, (B0 << RA) + rt, q2, nq);
    if (RA > 0) {
#pragma unroll
        for (int k = 0; k < 16; k++) tile[((rt << 4) | k) * C + cc] = x[k];
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const u32 ka = k >> (4 - RA), kb = k & ((1 << (4 - RA)) - 1);
            x[k] = tile[((ka << 4) | (kb << RA) | rt) * C + cc];
        }
        gs_round<RA>(x, tw, B0, q2, nq);
    }
    if (logsub) {
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const u32 ka = RA > 0 ? k >> (4 - RA) : 0, kb = RA > 0 ? k & ((1 << (4 - RA)) - 1) : k;
            io[size_t(RA > 0 ? ((ka << 4) | (kb << RA) | rt) : k) * 256] = x[k];
        }
        return;
    }
    // last stage (s = 0) paired k with k+8: x side still needs N^-1, y side got it through itw[1]
    const u64 ni = T.ninv[p], nis = T.ninv_shoup[p];
    u64 s = 0, ss = 0;
    if (SCALE) {
        s = scale[t];
        ss = scale_shoup[t];
    }
#pragma unroll
    for (int k = 0; k < 16; k++) {
        u64 v = x[k];
        if (k < 8) v = shoup_lazy(v, ni, nis, nq);
        v = csub(v, q);
        if (SCALE) v = shoup_mul(v, s, ss, q, nq);
        u32 r;
        if (RA > 0) {
            const u32 ka = k >> (4 - RA), kb = k & ((1 << (4 - RA)) - 1);
            r = (ka << 4) | (kb << RA) | rt;
        } else {
            r = k;
        }
        io[size_t(r) * 256] = v;
        if (MIRROR) {
            const size_t off = size_t(io - dst) + size_t(r) * 256;
            for (u32 m = 0; m < mirror.n; m++) mirror.base[m][off] = v;
        }
    }
}

// ---- N = 2^16, few limbs: the whole inverse transform in ONE kernel, one HBM round trip ------------------------------
// The two-pass inverse of a 10-limb launch (the digit chains of a key switch, the P limbs of a moddown) is two launches of
// 0.3 waves each: latency, not throughput (12 % of the warps active in ncu).  Here a CLUSTER of 8 thread blocks owns one limb:
// block k runs the eight row stages on rows [32k, 32k + 32) exactly as ntt_inv_pA does, hands every value to the block that
// owns its COLUMN through distributed shared memory (the transpose between the passes: 16 remote 8-byte stores per thread,
// 128-byte runs per half-warp), and after one cluster barrier runs the eight column stages on columns [32k, 32k + 32) out of
// its own shared memory exactly as ntt_inv_pB<4> does.  Same butterflies in the same order on the same values: the result is
// bit-identical; the limb is read once and written once (16 N bytes), and the second launch with its tail is gone.
// MEASURED on the B200 (tools/ntt_small_bench.py, profiles/r02_ntt_cluster_vs_two_pass.json), warm, back to back: one block of
// 512 threads per SM holds 1/8 limb = 65536 butterflies, 20 us at 16 warps per SM whatever the limb count up to ~12 clusters,
// and twice that from 16 limbs on (an 8-block cluster needs 8 free SMs of one GPC).  The two-pass pair takes 11 - 16 us for
// 2 - 8 limbs, 21.8 us for 10, 32 us for 30: its 128-thread blocks spread a few limbs over ALL SMs, which matters more than
// the second launch.  HMult+relin 343 -> 376 us and bootstrap 41.8 -> 42.3 ms with the cluster form in the digit chains.
// Kept as a bit-exact, selectable variant (default off); the two-pass kernels stay the product path.
static constexpr int kClusterCtas = 8;
static constexpr int kClusterThreads = 512;

template <bool SCALE>
__global__ void __cluster_dims__(kClusterCtas, 1, 1) __launch_bounds__(kClusterThreads)
    ntt_inv_cluster(NttTables T, u64 *dst, const u64 *src, LimbMap map, const u64 *scale, const u64 *scale_shoup) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ __align__(16) u64 tile[];  // 8192 words: [32 rows][256] (row pass), then [256 rows][32 columns]
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    const u32 kb = cluster.block_rank();
    const u32 t = map.limb(blockIdx.y), z = blockIdx.z;
    const u32 p = prime_of(map, t);
    const size_t N = size_t(1) << 16;
    const Mod md = T.mods[p];
    const u64 q = md.q, q2 = 2 * md.q, nq = md.nq;
    const ulonglong2 *tw = T.itw + size_t(p) * N;
    u64 x[16];
    {   // rows (stages 15 .. 8)
        const u32 lane = threadIdx.x & 15, rowi = threadIdx.x >> 4, r = kb * 32 + rowi;
        const u64 *irow = src + z * map.src_poly_stride + size_t(t) * N + size_t(r) * 256;
        u64 *srow = tile + rowi * 256;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const u32 c = 2 * (lane + 16 * j);
            *reinterpret_cast<ulonglong2 *>(srow + swz(c)) = ld2(irow + c);
        }
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 16; k += 2) {
            ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(srow + swz(16 * lane + k));
            x[k] = v.x;
            x[k + 1] = v.y;
        }
        const u32 B = (1u << 8) + r;
        gs_round<4>(x, tw, (B << 4) + lane, q2, nq);
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 16; k += 2)
            *reinterpret_cast<ulonglong2 *>(srow + swz(16 * lane + k)) = make_ulonglong2(x[k], x[k + 1]);
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 16; k++) x[k] = srow[swz(lane + 16 * k)];
        gs_round<4>(x, tw, B, q2, nq);
        // x[k] = (row r, column lane + 16 k): to the block that owns the column, into its [256][32] tile
        cluster.sync();  // every block has finished with its row tile (and has started)
#pragma unroll
        for (int k = 0; k < 16; k++) {
            u64 *remote = cluster.map_shared_rank(tile, k >> 1);
            remote[r * 32 + lane + 16 * (k & 1)] = x[k];
        }
        cluster.sync();
    }
    {   // columns (stages 7 .. 0), N^-1, optional scale
        const u32 cc = threadIdx.x & 31, rt = threadIdx.x >> 5;
#pragma unroll
        for (int k = 0; k < 16; k++) x[k] = tile[((rt << 4) | k) * 32 + cc];
        gs_round<4>(x, tw, 16 + rt, q2, nq);
#pragma unroll
        for (int k = 0; k < 16; k++) tile[((rt << 4) | k) * 32 + cc] = x[k];
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; k++) x[k] = tile[((k << 4) | rt) * 32 + cc];
        gs_round<4>(x, tw, 1, q2, nq);
        const u64 ni = T.ninv[p], nis = T.ninv_shoup[p];
        u64 s = 0, ss = 0;
        if (SCALE) {
            s = scale[t];
            ss = scale_shoup[t];
        }
        u64 *out = dst + z * map.poly_stride + size_t(t) * N + kb * 32 + cc;
#pragma unroll
        for (int k = 0; k < 16; k++) {
            u64 v = x[k];
            if (k < 8) v = shoup_lazy(v, ni, nis, nq);
            v = csub(v, q);
            if (SCALE) v = shoup_mul(v, s, ss, q, nq);
            out[size_t((k << 4) | rt) * 256] = v;
        }
    }
}

// ---- N <= 2048: one CTA per limb, the whole transform in shared memory --------------------------------
// Replaces fnwt_1d[_opt] / inwt_1d[_opt] (src/ntt/ntt_1d.cu:146-330): radix-2 stages on an N-word tile, same tables,
// same features as the two-pass kernels (limb map, exclusion, broadcast source, epilogue, inverse scale).
static constexpr int kSmallThreads = 256;

template <bool FWD>
__global__ void __launch_bounds__(kSmallThreads) ntt_small_kernel(NttTables T, u64 *dst, const u64 *src, LimbMap map,
                                                                  FwdEpilogue epi, const u64 *scale,
                                                                  const u64 *scale_shoup) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ u64 sm[];
    const u32 t = map.limb(blockIdx.y), z = blockIdx.z;
    if (FWD && excluded(map, t, z)) return;
    const u32 p = prime_of(map, t);
    const int logN = T.logN;
    const u32 n = 1u << logN, half = n >> 1;
    const Mod md = T.mods[p];
    const u64 q = md.q, q2 = 2 * md.q, nq = md.nq;
    const ulonglong2 *tw = (FWD ? T.tw : T.itw) + size_t(p) * (T.tw_stride ? T.tw_stride : n);
    const u64 *in = src + z * map.src_poly_stride + (FWD && map.src_bcast ? 0 : size_t(t) * n);
    const size_t off = size_t(t) * n;
    for (u32 i = threadIdx.x; i < n; i += blockDim.x) sm[i] = in[i];
    __syncthreads();
    for (int u = 0; u < logN; u++) {
        const int st = FWD ? u : logN - 1 - u;  // stage with 2^st groups of butterflies at distance n >> (st+1)
        const u32 len = n >> (st + 1), m = 1u << st;
        for (u32 j = threadIdx.x; j < half; j += blockDim.x) {
            const u32 g = j >> (logN - 1 - st), k = j & (len - 1), x = 2 * g * len + k;
            const ulonglong2 w = __ldg(tw + m + g);
            u64 a = sm[x], b = sm[x + len];
            if (FWD) ct_bfly(a, b, w.x, w.y, q2, nq);
            else gs_bfly(a, b, w.x, w.y, q2, nq);
            sm[x] = a;
            sm[x + len] = b;
        }
        __syncthreads();
    }
    if (FWD) {
        u64 *out = (epi.other ? epi.out + z * epi.out_poly_stride : dst + z * map.poly_stride) + off;
        const u64 *orow = epi.other ? epi.other + z * epi.other_poly_stride + off : nullptr;
        const u32 za = epi.acc_period ? z % epi.acc_period : z;
        const u64 *arow = (epi.other && epi.acc && (epi.acc_polys == 0 || za < epi.acc_polys)) ? epi.acc + (epi.acc_by_poly ? z : za) * epi.acc_poly_stride + off : nullptr;
        const u64 es = epi.other ? epi.scale[t] : 0, ess = epi.other ? epi.scale_shoup[t] : 0;
        for (u32 i = threadIdx.x; i < n; i += blockDim.x) {
            u64 v = csub(csub(sm[i], q2), q);
            if (orow) {
                v = shoup_mul(submod(orow[i], v, q), es, ess, q, nq);
                if (arow) v = addmod(v, arow[i], q);
            }
            out[i] = v;
        }
    } else {
        // the x side of the last stage (first half) still needs N^-1; the y side got it through itw[1]
        u64 *out = dst + z * map.poly_stride + off;
        const u64 ni = T.ninv[p], nis = T.ninv_shoup[p];
        for (u32 i = threadIdx.x; i < n; i += blockDim.x) {
            u64 v = sm[i];
            if (i < half) v = shoup_lazy(v, ni, nis, nq);
            v = csub(v, q);
            if (scale) v = shoup_mul(v, scale[t], scale_shoup[t], q, nq);
            out[i] = v;
        }
    }
}

// ---- launchers ---------------------------------------------------------------------------------------

// (limb, polynomial) pairs a launch really transforms: the excluded() rule on the host
static double transformed_limbs(const LimbMap &m, bool forward) {
    double n = 0;
    for (u32 z = 0; z < m.npoly; z++)
        for (u32 y = 0; y < m.limbs; y++) {
            const u32 t = m.limb(y);
            bool ex = false;
            if (forward) {
                ex = t >= m.excl_lo && t < m.excl_hi;
                if (!ex && m.excl_alpha) {
                    const u32 s0 = (m.excl_mod ? z % m.excl_mod : z) * m.excl_alpha, e0 = std::min(s0 + m.excl_alpha, m.size_Ql);
                    ex = t >= s0 && t < e0;
                }
            }
            n += ex ? 0 : 1;
        }
    return n;
}
static void trace_ntt(const NttTables &T, const LimbMap &m, bool forward, const FwdEpilogue *epi, bool scaled) {
    const double N = double(size_t(1) << T.logN), n = transformed_limbs(m, forward);
    double bytes = 16.0 * N * n, macs = scaled ? N * n : 0;
    if (forward && m.src_bcast) bytes = 8.0 * N * (n + m.npoly);  // one source row per polynomial
    if (epi && epi->other) {
        bytes += 8.0 * N * n;  // `other` read; the result is the one write already counted
        macs += N * n;
        if (epi->acc) bytes += 8.0 * N * n * (epi->acc_polys ? double(epi->acc_polys) / (epi->acc_period ? epi->acc_period : m.npoly) : 1.0);
    }
    trace_add(bytes, n * (N / 2) * T.logN, macs);
}

// launches of at most this many (limb, polynomial) pairs take the one-kernel cluster transform at N = 2^16
// (B200FHE_CLUSTER_NTT / b200fhe_tune_cluster_ntt; default 0 = never: see the measurements at ntt_inv_cluster)
static std::atomic<int> g_cluster_ntt{-1};
static u32 cluster_ntt_max_limbs() {
    int n = g_cluster_ntt.load(std::memory_order_relaxed);
    if (n < 0) {
        const char *v = getenv("B200FHE_CLUSTER_NTT");
        n = v ? std::max(0, atoi(v)) : 0;  // measured on the B200 (profiles/r02_ntt_cluster_vs_two_pass.json): never faster -> off
        g_cluster_ntt.store(n, std::memory_order_relaxed);
    }
    return (u32) n;
}
void ntt_set_cluster_max_limbs(int n) { g_cluster_ntt.store(n < 0 ? -1 : n, std::memory_order_relaxed); }

int ntt_launch_count_forward(int logN) { return logN < 12 ? 1 : (logN > 16 ? 3 : 2); }
int ntt_launch_count_inverse(int logN) { return logN < 12 ? 1 : (logN > 16 ? 3 : 2); }

// opt in to > 48 KiB dynamic shared memory once per device (callers may be concurrent host threads)
static cudaError_t p2_attributes() {
    static std::atomic<bool> done[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (done[dev].load(std::memory_order_acquire)) return cudaSuccess;
    cudaError_t e1 = cudaFuncSetAttribute(ntt_fwd_p2<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * kP2Tile * 8);
    cudaError_t e2 = cudaFuncSetAttribute(ntt_fwd_p2<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * kP2Tile * 8);
    if (e1 != cudaSuccess) return e1;
    if (e2 != cudaSuccess) return e2;
    done[dev].store(true, std::memory_order_release);
    return cudaSuccess;
}

static cudaError_t launch_fwd_p1(const NttTables &T, u64 *dst, const u64 *src, const LimbMap &map, cudaStream_t st) {
    const u32 logsub = T.logN > 16 ? T.logN - 16 : 0;
    const int ra = T.logN - 12 - (int) logsub;
    dim3 g1(((256u << ra) / kThreads) << logsub, map.limbs, map.npoly);
    LimbMap m1 = map;
    if (logsub) {
        launch_k(ntt_fwd_stage0, dim3((1u << T.logN) / 4 / 256, map.limbs, map.npoly), 256, 0, st, T, dst, src, map);
        src = dst;
        m1.src_bcast = 0;
        m1.src_poly_stride = map.poly_stride;
    }
    switch (ra) {
        case 0: launch_k(ntt_fwd_p1<0>, g1, kThreads, 0, st, T, dst, src, m1, logsub); break;
        case 1: launch_k(ntt_fwd_p1<1>, g1, kThreads, 0, st, T, dst, src, m1, logsub); break;
        case 2: launch_k(ntt_fwd_p1<2>, g1, kThreads, 0, st, T, dst, src, m1, logsub); break;
        case 3: launch_k(ntt_fwd_p1<3>, g1, kThreads, 0, st, T, dst, src, m1, logsub); break;
        default: launch_k(ntt_fwd_p1<4>, g1, kThreads, 0, st, T, dst, src, m1, logsub); break;
    }
    return cudaGetLastError();
}

cudaError_t ntt_forward(const NttTables &T, u64 *dst, const u64 *src, const LimbMap &map, const FwdEpilogue *epi,
                        cudaStream_t st) {
    if (T.logN < 1 || T.logN > 17) return cudaErrorInvalidValue;
    if (map.limbs == 0 || map.npoly == 0) return cudaSuccess;
    trace_ntt(T, map, true, epi, false);
    if (T.logN < 12) {
        FwdEpilogue e{};
        if (epi && epi->other) e = *epi;
        const u32 n = 1u << T.logN;
        launch_k(ntt_small_kernel<true>, dim3(1, map.limbs, map.npoly), std::min<u32>(kSmallThreads, std::max<u32>(n / 2, 1)), n * 8, st, 
            T, dst, src, map, e, nullptr, nullptr);
        return cudaGetLastError();
    }
    // N = 2^17: one radix-2 stage (src -> dst), then the 2^16-point column pass on each half of dst
    const u32 R = 1u << (T.logN - 8);
    dim3 g2(R / kRows, map.limbs, map.npoly);
    if (cudaError_t e = launch_fwd_p1(T, dst, src, map, st)) return e;
    if (cudaError_t e = p2_attributes()) return e;
    if (epi && epi->other && epi->acc) {
        launch_k(ntt_fwd_p2<2>, g2, kThreads, 3 * kP2Tile * 8, st, T, dst, map, *epi);
    } else if (epi && epi->other) {
        launch_k(ntt_fwd_p2<1>, g2, kThreads, 2 * kP2Tile * 8, st, T, dst, map, *epi);
    } else {
        FwdEpilogue e{};
        launch_k(ntt_fwd_p2<0>, g2, kThreads, kP2Tile * 8, st, T, dst, map, e);
    }
    return cudaGetLastError();
}

cudaError_t ntt_forward_p1(const NttTables &T, u64 *dst, const u64 *src, const LimbMap &map, cudaStream_t st) {
    if (T.logN < 12 || T.logN > 17) return cudaErrorInvalidValue;
    if (map.limbs == 0 || map.npoly == 0) return cudaSuccess;
    const double N = double(size_t(1) << T.logN), n = transformed_limbs(map, true);
    trace_add(16.0 * N * n, n * (N / 2) * (T.logN - 8), 0);
    return launch_fwd_p1(T, dst, src, map, st);
}

template <int BETA, int G>
static cudaError_t launch_p2_ip(const NttTables &T, const FwdInnerProd &a, u32 t0, u32 tn, cudaStream_t st) {
    const u32 R = 1u << (T.logN - 8);
    launch_k(ntt_fwd_p2_ip<BETA, G>, dim3(R / kIpRows, tn, a.batch), G * kIpGroup, size_t(G) * kIpRows * 256 * 8, st, T, a, t0);
    return cudaGetLastError();
}

cudaError_t ntt_forward_p2_inner_prod(const NttTables &T, const FwdInnerProd &a, u32 t0, u32 tn, cudaStream_t st) {
    if (T.logN < 12 || T.logN > 17 || a.beta < 2 || a.beta > (u32) kIpDigits || a.alpha == 0) return cudaErrorInvalidValue;
    if (tn == 0 || a.batch == 0) return cudaSuccess;
    if (t0 + tn > a.size_QlP) return cudaErrorInvalidValue;
    // a launch is all P limbs (every digit transformed) or all Q_l limbs (the owning digit read from the source)
    const bool p_rows = t0 >= a.size_Ql;
    if (!p_rows && t0 + tn > a.size_Ql) return cudaErrorInvalidValue;
    const double N = double(size_t(1) << T.logN);
    const double n_tr = double(tn) * (a.beta - (p_rows ? 0 : 1)), n_own = p_rows ? 0 : tn;
    trace_add(8.0 * N * ((n_tr + n_own) * a.batch + 2.0 * a.beta * tn + 2.0 * tn * a.batch), n_tr * a.batch * (N / 2) * 8,
              2.0 * N * tn * a.beta * a.batch);
    switch (a.beta * 2 + (p_rows ? 1 : 0)) {
        case 4: return launch_p2_ip<2, 1>(T, a, t0, tn, st);
        case 5: return launch_p2_ip<2, 2>(T, a, t0, tn, st);
        case 6: return launch_p2_ip<3, 2>(T, a, t0, tn, st);
        case 7: return launch_p2_ip<3, 3>(T, a, t0, tn, st);
        case 8: return launch_p2_ip<4, 3>(T, a, t0, tn, st);
        default: return launch_p2_ip<4, 4>(T, a, t0, tn, st);
    }
}

template <int RA>
static void launch_inv_pB(const NttTables &T, u64 *dst, const LimbMap &map, const u64 *scale, const u64 *scale_shoup,
                          dim3 g, cudaStream_t st, const PeerMirror *mirror, u32 logsub = 0) {
    if (logsub) {
        launch_k(ntt_inv_pB<RA, false>, g, kThreads, 0, st, T, dst, map, nullptr, nullptr, PeerMirror{}, logsub);
    } else if (mirror && mirror->n) {
        if (scale)
            launch_k(ntt_inv_pB<RA, true, true>, g, kThreads, 0, st, T, dst, map, scale, scale_shoup, *mirror, 0u);
        else
            launch_k(ntt_inv_pB<RA, false, true>, g, kThreads, 0, st, T, dst, map, nullptr, nullptr, *mirror, 0u);
    } else if (scale)
        launch_k(ntt_inv_pB<RA, true>, g, kThreads, 0, st, T, dst, map, scale, scale_shoup, PeerMirror{}, 0u);
    else
        launch_k(ntt_inv_pB<RA, false>, g, kThreads, 0, st, T, dst, map, nullptr, nullptr, PeerMirror{}, 0u);
}

cudaError_t ntt_forward_sub(const NttTables &T, int sub_logn, u64 *dst, const u64 *src, const LimbMap &map, cudaStream_t st) {
    if (sub_logn < 1 || sub_logn > 11 || sub_logn > T.logN) return cudaErrorInvalidValue;
    if (map.limbs == 0 || map.npoly == 0) return cudaSuccess;
    trace_add(16.0 * (1u << sub_logn) * map.limbs * map.npoly, double(map.limbs) * map.npoly * (1u << sub_logn) / 2 * sub_logn, 0);
    NttTables S = T;
    S.tw_stride = size_t(1) << T.logN;
    S.logN = sub_logn;
    const u32 n = 1u << sub_logn;
    launch_k(ntt_small_kernel<true>, dim3(1, map.limbs, map.npoly), std::min<u32>(kSmallThreads, std::max<u32>(n / 2, 1)), n * 8, st, 
        S, dst, src, map, FwdEpilogue{}, nullptr, nullptr);
    return cudaGetLastError();
}

cudaError_t ntt_inverse(const NttTables &T, u64 *dst, const u64 *src, const LimbMap &map, const u64 *scale,
                        const u64 *scale_shoup, cudaStream_t st, const PeerMirror *mirror) {
    if (T.logN < 1 || T.logN > 17) return cudaErrorInvalidValue;
    if (map.limbs == 0 || map.npoly == 0) return cudaSuccess;
    if (mirror && mirror->n && (T.logN < 12 || mirror->n > (u32) kMaxMirrors)) return cudaErrorNotSupported;
    trace_ntt(T, map, false, nullptr, scale != nullptr);
    if (T.logN < 12) {
        const u32 n = 1u << T.logN;
        launch_k(ntt_small_kernel<false>, dim3(1, map.limbs, map.npoly), std::min<u32>(kSmallThreads, std::max<u32>(n / 2, 1)), n * 8, st, 
            T, dst, src, map, FwdEpilogue{}, scale, scale_shoup);
        return cudaGetLastError();
    }
    const u32 logsub = T.logN > 16 ? T.logN - 16 : 0;
    if (logsub && mirror && mirror->n) return cudaErrorNotSupported;
    if (T.logN == 16 && !(mirror && mirror->n) && map.limbs * map.npoly <= cluster_ntt_max_limbs()) {
        // few limbs: one kernel, a cluster of 8 blocks per limb (ntt_inv_cluster)
        static std::atomic<bool> done[64];
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
        if (!done[dev].load(std::memory_order_acquire)) {
            cudaError_t e1 = cudaFuncSetAttribute(ntt_inv_cluster<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
            cudaError_t e2 = cudaFuncSetAttribute(ntt_inv_cluster<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
            if (e1 != cudaSuccess) return e1;
            if (e2 != cudaSuccess) return e2;
            done[dev].store(true, std::memory_order_release);
        }
        dim3 g(kClusterCtas, map.limbs, map.npoly);
        if (scale) launch_k(ntt_inv_cluster<true>, g, kClusterThreads, 65536, st, T, dst, src, map, scale, scale_shoup);
        else launch_k(ntt_inv_cluster<false>, g, kClusterThreads, 65536, st, T, dst, src, map, nullptr, nullptr);
        return cudaGetLastError();
    }
    const int ra = T.logN - 12 - (int) logsub;
    const u32 R = 1u << (T.logN - 8);
    dim3 gA(R / kRows, map.limbs, map.npoly), gB(((256u << ra) / kThreads) << logsub, map.limbs, map.npoly);
    launch_k(ntt_inv_pA, gA, kThreads, 0, st, T, dst, src, map);
    switch (ra) {
        case 0: launch_inv_pB<0>(T, dst, map, scale, scale_shoup, gB, st, mirror, logsub); break;
        case 1: launch_inv_pB<1>(T, dst, map, scale, scale_shoup, gB, st, mirror, logsub); break;
        case 2: launch_inv_pB<2>(T, dst, map, scale, scale_shoup, gB, st, mirror, logsub); break;
        case 3: launch_inv_pB<3>(T, dst, map, scale, scale_shoup, gB, st, mirror, logsub); break;
        default: launch_inv_pB<4>(T, dst, map, scale, scale_shoup, gB, st, mirror, logsub); break;
    }
    if (logsub) {
        dim3 g0((1u << T.logN) / 4 / 256, map.limbs, map.npoly);
        if (scale) launch_k(ntt_inv_stage0<true>, g0, 256, 0, st, T, dst, map, scale, scale_shoup);
        else launch_k(ntt_inv_stage0<false>, g0, 256, 0, st, T, dst, map, nullptr, nullptr);
    }
    return cudaGetLastError();
}

}  // namespace b200fhe
