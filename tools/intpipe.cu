// intpipe.cu — integer-pipe microbenchmarks for the sm_100a NTT butterfly (measurement tool, not product).
// Measures (a) raw issue rates of the instruction classes the butterfly is made of and (b) in-register
// butterfly throughput of candidate formulations, i.e. the compute-only ceiling of a limb-NTT.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o tools/intpipe tools/intpipe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../phantom-fhe-boot_b200/csrc/arith.cuh"
using namespace b200fhe;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

// ---------------------------------------------------------------- butterfly variants
// V0: product butterfly (arith.cuh)
struct BflyV0 {
    static __device__ __forceinline__ void round(u64 (&x)[16], const u64 *w, const u64 *ws, u64 q, u64 nq) {
        const u64 q2 = 2 * q;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int h = 8 >> u;
#pragma unroll
            for (int g = 0; g < (1 << u); g++)
#pragma unroll
                for (int j = 0; j < h; j++) {
                    const int k = g * 2 * h + j;
                    ct_bfly(x[k], x[k + h], w[u], ws[u], q2, nq);
                }
        }
    }
};

// sign-test conditional subtract: valid when x - m fits a signed 64-bit value
__device__ __forceinline__ u64 csub_s(u64 x, u64 m) {
    u64 d = x - m;
    return ((long long) d < 0) ? x : d;
}

// V1: no reduction inside a 4-stage round (values grow by 2q per stage, q < 2^60), one csub(8q) at the end
struct BflyV1 {
    static __device__ __forceinline__ void round(u64 (&x)[16], const u64 *w, const u64 *ws, u64 q, u64 nq) {
        const u64 q2 = 2 * q;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int h = 8 >> u;
#pragma unroll
            for (int g = 0; g < (1 << u); g++)
#pragma unroll
                for (int j = 0; j < h; j++) {
                    const int k = g * 2 * h + j;
                    u64 t = shoup_lazy(x[k + h], w[u], ws[u], nq);
                    x[k + h] = x[k] + q2 - t;
                    x[k] = x[k] + t;
                }
        }
        const u64 q8 = 8 * q;
#pragma unroll
        for (int k = 0; k < 16; k++) x[k] = csub_s(x[k], q8);
    }
};

// V2: like V0 but the conditional subtract is the sign-test form
struct BflyV2 {
    static __device__ __forceinline__ void round(u64 (&x)[16], const u64 *w, const u64 *ws, u64 q, u64 nq) {
        const u64 q2 = 2 * q;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int h = 8 >> u;
#pragma unroll
            for (int g = 0; g < (1 << u); g++)
#pragma unroll
                for (int j = 0; j < h; j++) {
                    const int k = g * 2 * h + j;
                    u64 t = shoup_lazy(x[k + h], w[u], ws[u], nq);
                    u64 a = csub_s(x[k], q2);
                    x[k + h] = a + q2 - t;
                    x[k] = a + t;
                }
        }
    }
};

// approximate Shoup quotient: y1*s1 + hi32(y1*s0) + hi32(y0*s1) underestimates floor(y*ws/2^64) by at most 2, so the
// lazy product lands in [0,4q); one wide and two high-half 32-bit multiplies instead of four wide ones
__device__ __forceinline__ u64 shoup_lazy4(u64 y, u64 w, u64 ws, u64 nq) {
    const u32 y0 = (u32) y, y1 = (u32) (y >> 32), s0 = (u32) ws, s1 = (u32) (ws >> 32);
    u64 hi = (u64) y1 * s1 + __umulhi(y1, s0) + __umulhi(y0, s1);
    return y * w + hi * nq;
}
// variant that drops only the lo*lo product (error <= 1, result < 3q)
__device__ __forceinline__ u64 shoup_lazy3(u64 y, u64 w, u64 ws, u64 nq) {
    const u32 y0 = (u32) y, y1 = (u32) (y >> 32), s0 = (u32) ws, s1 = (u32) (ws >> 32);
    u64 mid = (u64) y1 * s0;
    u64 mid2 = (u64) y0 * s1;
    u64 hi = (u64) y1 * s1 + (mid >> 32) + (mid2 >> 32) + ((((mid & 0xffffffffu) + (mid2 & 0xffffffffu))) >> 32);
    return y * w + hi * nq;
}

// V3: values in [0,8q), x reduced against 4q, approximate quotient (t < 4q)
struct BflyV3 {
    static __device__ __forceinline__ void round(u64 (&x)[16], const u64 *w, const u64 *ws, u64 q, u64 nq) {
        const u64 q4 = 4 * q;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int h = 8 >> u;
#pragma unroll
            for (int g = 0; g < (1 << u); g++)
#pragma unroll
                for (int j = 0; j < h; j++) {
                    const int k = g * 2 * h + j;
                    u64 t = shoup_lazy4(x[k + h], w[u], ws[u], nq);
                    u64 a = csub(x[k], q4);
                    x[k + h] = a + q4 - t;
                    x[k] = a + t;
                }
        }
    }
};
struct BflyV4 {
    static __device__ __forceinline__ void round(u64 (&x)[16], const u64 *w, const u64 *ws, u64 q, u64 nq) {
        const u64 q4 = 4 * q;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int h = 8 >> u;
#pragma unroll
            for (int g = 0; g < (1 << u); g++)
#pragma unroll
                for (int j = 0; j < h; j++) {
                    const int k = g * 2 * h + j;
                    u64 t = shoup_lazy3(x[k + h], w[u], ws[u], nq);
                    u64 a = csub(x[k], q4);
                    x[k + h] = a + q4 - t;
                    x[k] = a + t;
                }
        }
    }
};
// V5: V3 arithmetic with the sign-test conditional subtract
struct BflyV5 {
    static __device__ __forceinline__ void round(u64 (&x)[16], const u64 *w, const u64 *ws, u64 q, u64 nq) {
        const u64 q4 = 4 * q;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int h = 8 >> u;
#pragma unroll
            for (int g = 0; g < (1 << u); g++)
#pragma unroll
                for (int j = 0; j < h; j++) {
                    const int k = g * 2 * h + j;
                    u64 t = shoup_lazy4(x[k + h], w[u], ws[u], nq);
                    u64 a = csub_s(x[k], q4);
                    x[k + h] = a + q4 - t;
                    x[k] = a + t;
                }
        }
    }
};

template <class B>
__global__ void __launch_bounds__(256) bfly_kernel(u64 *out, const u64 *tw, u64 q, int rounds) {
    u64 x[16], w[4], ws[4];
    const u64 nq = 0 - q;
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = threadIdx.x * 16 + k + blockIdx.x;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        w[u] = tw[2 * u];
        ws[u] = tw[2 * u + 1];
    }
    for (int r = 0; r < rounds; r++) {
        B::round(x, w, ws, q, nq);
#pragma unroll
        for (int u = 0; u < 4; u++) w[u] ^= (r & 1);  // keep the twiddles loop-variant but cheap
    }
    u64 s = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) s ^= x[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ---------------------------------------------------------------- raw pipes
template <int KIND>
__global__ void __launch_bounds__(256) pipe_kernel(u64 *out, int iters) {
    unsigned a[8], b = threadIdx.x | 1, c = blockIdx.x + 3;
    u64 acc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = threadIdx.x + i; acc[i] = i; }
    for (int r = 0; r < iters; r++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (KIND == 0) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a[i]), "r"(b));
            if (KIND == 1) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
            if (KIND == 2) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
            if (KIND == 3) {  // alternate: one wide multiply + one add per pair
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a[i]), "r"(b));
                asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(c));
            }
            if (KIND == 4) asm volatile("mul.hi.u64 %0, %0, %1;" : "+l"(acc[i]) : "l"((u64) b << 33 | c));
            if (KIND == 5) asm volatile("add.u64 %0, %0, %1;" : "+l"(acc[i]) : "l"((u64) b << 33 | c));
            if (KIND == 6) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
            if (KIND == 7) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
        }
    }
    u64 s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s ^= acc[i] ^ a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ---------------------------------------------------------------- base-conversion inner loop in registers
// One output coefficient = NI multiply-accumulates of 64x64 -> 128 bits + one 128-bit Barrett reduction.  KIND 0: mac128 chain +
// barrett128 (the product's loop), 1: mac128 only, 2: 30-bit limbs (4 plain mad.wide per MAC) + recombination + barrett128,
// 3: 30-bit limbs only.  Two coefficients per thread and two outputs in flight, like bconv_kernel.
__device__ __forceinline__ u64 madw32(unsigned a, unsigned b, u64 c) {
    u64 d;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(d) : "r"(a), "r"(b), "l"(c));
    return d;
}
// lazy reduction of a 128-bit sum x < 2^124 to [0, 4q): quotient estimate from the high word only,
//   qhat = x.hi * r1 + mulhi(x.hi, r0) + ((x.lo >> 32) * r1 >> 32)   (ratio = r1 * 2^64 + r0 = floor(2^128 / q), r1 < 2^7 for q > 2^57)
// which is below the true quotient by at most 3, so x.lo - qhat * q (mod 2^64) < 4q... then one conditional subtraction of 2q is
// NOT taken here: the consumer (a lazy forward butterfly) accepts [0, 4q).
__device__ __forceinline__ u64 barrett_lazy(u128 x, const Mod &m) {
    const u64 qh = x.hi * m.r1 + __umul64hi(x.hi, m.r0) + (((x.lo >> 32) * m.r1) >> 32);
    return x.lo + qh * m.nq;
}
template <int KIND, int NI>
__global__ void __launch_bounds__(256) conv_kernel(u64 *out, const u64 *mat, Mod md, int outputs) {
    __shared__ u64 sm[64 * NI];
    for (int i = threadIdx.x; i < 64 * NI; i += blockDim.x) sm[i] = mat[i % 8] + i;
    __syncthreads();
    u64 x0[NI], x1[NI];
#pragma unroll
    for (int i = 0; i < NI; i++) {
        x0[i] = (mat[i % 8] ^ (threadIdx.x * 0x9E3779B97F4A7C15ull)) >> 4;
        x1[i] = (mat[(i + 3) % 8] + blockIdx.x * 0x2545F4914F6CDD1Dull) >> 4;
    }
    u64 s = 0;
    for (int j = 0; j < outputs; j++) {
        const u64 *row = sm + (j & 63) * NI;
        if (KIND < 2 || KIND == 4) {
            u128 a0 = {0, 0}, a1 = {0, 0};
#pragma unroll
            for (int i = 0; i < NI; i++) {
                const u64 m = row[i];
                mac128(a0, x0[i], m);
                mac128(a1, x1[i], m);
            }
            if (KIND == 0) s ^= barrett128(a0, md) ^ barrett128(a1, md);
            else if (KIND == 4) s ^= barrett_lazy(a0, md) ^ barrett_lazy(a1, md);
            else s ^= a0.lo ^ a0.hi ^ a1.lo ^ a1.hi;
        } else {
            u64 a00 = 0, a01 = 0, a10 = 0, a11 = 0, b00 = 0, b01 = 0, b10 = 0, b11 = 0;
#pragma unroll
            for (int i = 0; i < NI; i++) {
                const u64 m = row[i];
                const unsigned m0 = (unsigned) m & 0x3FFFFFFFu, m1 = (unsigned) (m >> 34);
                const unsigned xa0 = (unsigned) x0[i] & 0x3FFFFFFFu, xa1 = (unsigned) (x0[i] >> 30) & 0x3FFFFFFFu;
                const unsigned xb0 = (unsigned) x1[i] & 0x3FFFFFFFu, xb1 = (unsigned) (x1[i] >> 30) & 0x3FFFFFFFu;
                a00 = madw32(xa0, m0, a00);
                a01 = madw32(xa0, m1, a01);
                a10 = madw32(xa1, m0, a10);
                a11 = madw32(xa1, m1, a11);
                b00 = madw32(xb0, m0, b00);
                b01 = madw32(xb0, m1, b01);
                b10 = madw32(xb1, m0, b10);
                b11 = madw32(xb1, m1, b11);
            }
            if (KIND == 2) {
                u128 ta, tb;
                ta.lo = a00 + (a01 << 30) + (a10 << 30) + (a11 << 60);
                ta.hi = (a01 >> 34) + (a10 >> 34) + (a11 >> 4);
                tb.lo = b00 + (b01 << 30) + (b10 << 30) + (b11 << 60);
                tb.hi = (b01 >> 34) + (b10 >> 34) + (b11 >> 4);
                s ^= barrett128(ta, md) ^ barrett128(tb, md);
            } else {
                s ^= a00 ^ a01 ^ a10 ^ a11 ^ b00 ^ b01 ^ b10 ^ b11;
            }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F>
static float time_ms(F f) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    f();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    f();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

int main() {
    int dev = 0, sms = 0, khz = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
    u64 *out, *tw;
    const int blocks = sms * 8, threads = 256;
    CK(cudaMalloc(&out, size_t(blocks) * threads * 8));
    CK(cudaMalloc(&tw, 64));
    const u64 q = 1152921504606830593ull;  // 60-bit NTT prime
    u64 h[8];
    for (int i = 0; i < 4; i++) {
        h[2 * i] = (0x1234567890abcdefull * (i + 3)) % q;
        h[2 * i + 1] = (u64) (((unsigned __int128) h[2 * i] << 64) / q);
    }
    CK(cudaMemcpy(tw, h, 64, cudaMemcpyHostToDevice));
    printf("{\"sms\": %d, \"clock_khz_attr\": %d", sms, khz);
    const int iters = 4096;
    const char *names[8] = {"mad.wide.u32", "mad.lo.u32", "add.u32", "mad.wide+add pair", "mul.hi.u64", "add.u64", "mul.hi.u32", "mad.hi.u32"};
    const double per[8] = {8, 8, 8, 16, 8, 8, 8, 8};
    float ms;
#define PIPE(K)                                                                                              \
    ms = time_ms([&] { pipe_kernel<K><<<blocks, threads>>>(out, iters); });                                  \
    printf(",\n \"%s_Gops\": %.1f", names[K], double(blocks) * threads * iters * per[K] / ms / 1e6);
    PIPE(0) PIPE(1) PIPE(2) PIPE(3) PIPE(4) PIPE(5) PIPE(6) PIPE(7)
    const int rounds = 512;
#define BF(V, name)                                                                                          \
    ms = time_ms([&] { bfly_kernel<V><<<blocks, threads>>>(out, tw, q, rounds); });                          \
    printf(",\n \"%s_Gbfly\": %.2f", name, double(blocks) * threads * rounds * 32 / ms / 1e6);
    BF(BflyV0, "bfly_v0_product") BF(BflyV1, "bfly_v1_lazy4") BF(BflyV2, "bfly_v2_signcsub") BF(BflyV3, "bfly_v3_approx_hi") BF(BflyV4, "bfly_v4_drop_lolo") BF(BflyV5, "bfly_v5_approx_signcsub")
    {
        Mod md;
        md.q = q;
        md.nq = 0 - q;
        const unsigned __int128 ratio = ~(unsigned __int128) 0 / q;
        md.r0 = (u64) ratio;
        md.r1 = (u64) (ratio >> 64);
        const int outputs = 1024;
#define CONV(K, name)                                                                                        \
    ms = time_ms([&] { conv_kernel<K, 10><<<blocks, threads>>>(out, tw, md, outputs); });                    \
    printf(",\n \"%s_Gmac\": %.1f", name, double(blocks) * threads * outputs * 2 * 10 / ms / 1e6);
        CONV(0, "conv10_mac128_barrett") CONV(1, "conv10_mac128_only") CONV(2, "conv10_limb30_barrett") CONV(3, "conv10_limb30_only") CONV(4, "conv10_mac128_barrett_lazy")
    }
    CK(cudaGetLastError());
    printf("}\n");
    return 0;
}
