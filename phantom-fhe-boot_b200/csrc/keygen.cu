// keygen.cu — key generation, encryption, decryption (see keygen.h).
#include "keygen.h"

#include <cstdio>
#include <cstring>
#include <random>

#include "../../include/b200fhe.h"
#include "evaluator.h"

namespace b200fhe {

static constexpr int kKT = 256;

// ------------------------------------------------------------------------------------------- ChaCha20 block
struct ChaKey {
    u32 k[8];
};

#define B200_QR(a, b, c, d)                \
    a += b; d ^= a; d = __funnelshift_l(d, d, 16); \
    c += d; b ^= c; b = __funnelshift_l(b, b, 12); \
    a += b; d ^= a; d = __funnelshift_l(d, d, 8);  \
    c += d; b ^= c; b = __funnelshift_l(b, b, 7);

// RFC 8439 block function: 32-bit block counter, 96-bit nonce (n0, n1, n2)
__device__ __forceinline__ void chacha20_block(const ChaKey &key, u32 counter, u32 n0, u32 n1, u32 n2, u32 out[16]) {
    const u32 in[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u, key.k[0], key.k[1], key.k[2], key.k[3],
                        key.k[4],    key.k[5],    key.k[6],    key.k[7],    counter,  n0,       n1,       n2};
    u32 x0 = in[0], x1 = in[1], x2 = in[2], x3 = in[3], x4 = in[4], x5 = in[5], x6 = in[6], x7 = in[7];
    u32 x8 = in[8], x9 = in[9], x10 = in[10], x11 = in[11], x12 = in[12], x13 = in[13], x14 = in[14], x15 = in[15];
#pragma unroll
    for (int r = 0; r < 10; r++) {
        B200_QR(x0, x4, x8, x12)
        B200_QR(x1, x5, x9, x13)
        B200_QR(x2, x6, x10, x14)
        B200_QR(x3, x7, x11, x15)
        B200_QR(x0, x5, x10, x15)
        B200_QR(x1, x6, x11, x12)
        B200_QR(x2, x7, x8, x13)
        B200_QR(x3, x4, x9, x14)
    }
    out[0] = x0 + in[0];
    out[1] = x1 + in[1];
    out[2] = x2 + in[2];
    out[3] = x3 + in[3];
    out[4] = x4 + in[4];
    out[5] = x5 + in[5];
    out[6] = x6 + in[6];
    out[7] = x7 + in[7];
    out[8] = x8 + in[8];
    out[9] = x9 + in[9];
    out[10] = x10 + in[10];
    out[11] = x11 + in[11];
    out[12] = x12 + in[12];
    out[13] = x13 + in[13];
    out[14] = x14 + in[14];
    out[15] = x15 + in[15];
}

// known-answer hook for the tests (RFC 8439 section 2.3.2 vector): one block to global memory
__global__ void chacha20_block_kernel(ChaKey key, u32 counter, u32 n0, u32 n1, u32 n2, u32 *out) {
    u32 b[16];
    chacha20_block(key, counter, n0, n1, n2, b);
    for (int i = 0; i < 16; i++) out[i] = b[i];
}

// ------------------------------------------------------------------------------------------- samplers
// Small polynomials (secret, error): the SAME signed value on every limb (src/prng.cu:149-159, 226-241), so one
// thread draws the values of a run of coefficients once and writes them under every prime.
//   TERNARY: 16 coefficients per ChaCha block, one 32-bit word each, Lemire's multiply-shift with rejection
//            (3 * w: the low half is < 2^32 mod 3 = 1 only for w = 0) -> exactly uniform on {-1, 0, 1}
//   CBD:     8 coefficients per block, one 64-bit word each: popcount of 21 bits minus popcount of 21 bits
//            (the reference's 8 + 8 + 5 coins a side, sigma = sqrt(21 / 2) = 3.24)
template <bool TERNARY>
__global__ void sample_small_kernel(u64 *out, ChaKey key, u32 n0, u32 n1, const Mod *__restrict__ mods, PrimeMap pm,
                                    u32 limbs, u32 N) {
    constexpr int PER = TERNARY ? 16 : 8;
    const u32 idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N / PER) return;
    u32 b[16];
    chacha20_block(key, idx, n0, n1, TERNARY ? 1u : 2u, b);
    int v[PER];
    if (TERNARY) {
        u32 spare[16];
        int sp = 16;
        u32 tries = 0;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            u64 m = (u64) b[i] * 3u;
            while ((u32) m == 0u) {  // probability 2^-32 per draw
                if (sp == 16) {
                    tries++;
                    chacha20_block(key, idx + tries * (N / PER), n0, n1, 1u, spare);
                    sp = 0;
                }
                u32 w = 0;
                for (int j = 0; j < 16; j++) w = j == sp ? spare[j] : w;
                sp++;
                m = (u64) w * 3u;
            }
            v[i] = (int) (m >> 32) - 1;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const u32 lo = b[2 * i], hi = b[2 * i + 1];
            v[i] = __popc(lo & 0x1FFFFFu) - __popc(hi & 0x1FFFFFu);
        }
    }
    const size_t base = (size_t) idx * PER;
    for (u32 l = 0; l < limbs; l++) {
        const u64 q = mods[pm.prime(l)].q;
        u64 *o = out + (size_t) l * N + base;
#pragma unroll
        for (int i = 0; i < PER; i += 2)
            st2(o + i, v[i] < 0 ? q + (u64) (long long) v[i] : (u64) v[i],
                v[i + 1] < 0 ? q + (u64) (long long) v[i + 1] : (u64) v[i + 1]);
    }
}

// Uniform residues (src/prng.cu:171-205): 8 coefficients of one limb per ChaCha block; a 64-bit word is accepted
// when it lies below the largest multiple of q that fits 64 bits, rejected words are replaced from further blocks
// of the same stream (counter + tries * total).
__global__ void sample_uniform_kernel(u64 *out, ChaKey key, u32 n0, u32 n1, const Mod *__restrict__ mods, PrimeMap pm,
                                      u32 limbs, u32 N) {
    const u32 idx = blockIdx.x * blockDim.x + threadIdx.x, per_limb = N / 8, total = limbs * per_limb;
    if (idx >= total) return;
    const Mod m = mods[pm.prime(idx / per_limb)];
    const u64 limit = 0ull - (m.nq % m.q);  // 2^64 - (2^64 mod q): accept w < limit
    u32 b[16], spare[16];
    chacha20_block(key, idx, n0, n1, 3u, b);
    int sp = 16;
    u32 tries = 0;
    u64 r[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        u64 w = (u64) b[2 * i] | ((u64) b[2 * i + 1] << 32);
        while (w >= limit) {  // probability q / 2^64 < 1/8
            if (sp == 16) {
                tries++;
                chacha20_block(key, idx + tries * total, n0, n1, 3u, spare);
                sp = 0;
            }
            u32 lo = 0, hi = 0;
            for (int j = 0; j < 16; j += 2)
                if (j == sp) {
                    lo = spare[j];
                    hi = spare[j + 1];
                }
            sp += 2;
            w = (u64) lo | ((u64) hi << 32);
        }
        r[i] = barrett64(w, m);
    }
    u64 *o = out + (size_t) idx * 8;
#pragma unroll
    for (int i = 0; i < 8; i += 2) st2(o + i, r[i], r[i + 1]);
}

// ------------------------------------------------------------------------------------------- reference-compatible samplers
// The reference's own generator and samplers, restated so that the same 64-byte seeds give the same polynomials bit for
// bit (src/prng.cu:17-133 salsa20_gpu, :142-160 sample_ternary_poly, :171-205 sample_uniform_poly, :222-243
// sample_error_poly).  Its block function is the Salsa20 double round run on a state WITHOUT the sigma constants: words
// 0..7 and 10..15 are the first 56 key bytes, words 8, 9 the 64-bit nonce; there is no block counter (every sampler asks for
// one 64-byte block per nonce).  The output block is state + input, little endian.
struct RefSeed {
    u32 w[14];  // key bytes 0..55 as little-endian words
};
__device__ __forceinline__ u32 rotl32(u32 v, int c) { return (v << c) | (v >> (32 - c)); }
__device__ __forceinline__ void salsa_qr(u32 &a, u32 &b, u32 &c, u32 &d) {  // b ^= (a+d)<<<7; c ^= (b+a)<<<9; d ^= (c+b)<<<13; a ^= (d+c)<<<18
    b ^= rotl32(a + d, 7);
    c ^= rotl32(b + a, 9);
    d ^= rotl32(c + b, 13);
    a ^= rotl32(d + c, 18);
}
__device__ __forceinline__ void ref_block(const RefSeed &k, u64 nonce, u32 (&out)[16]) {
    u32 j[16], x[16];
#pragma unroll
    for (int i = 0; i < 8; i++) j[i] = k.w[i];
    j[8] = (u32) nonce;
    j[9] = (u32) (nonce >> 32);
#pragma unroll
    for (int i = 0; i < 6; i++) j[10 + i] = k.w[8 + i];
#pragma unroll
    for (int i = 0; i < 16; i++) x[i] = j[i];
#pragma unroll 1
    for (int r = 0; r < 10; r++) {
        salsa_qr(x[0], x[4], x[8], x[12]);    // columns
        salsa_qr(x[5], x[9], x[13], x[1]);
        salsa_qr(x[10], x[14], x[2], x[6]);
        salsa_qr(x[15], x[3], x[7], x[11]);
        salsa_qr(x[0], x[1], x[2], x[3]);     // rows
        salsa_qr(x[5], x[6], x[7], x[4]);
        salsa_qr(x[10], x[11], x[8], x[9]);
        salsa_qr(x[15], x[12], x[13], x[14]);
    }
#pragma unroll
    for (int i = 0; i < 16; i++) out[i] = x[i] + j[i];
}

// kind 0: ternary, byte 0 of the block mod 3 (0 -> -1, 1 -> 0, 2 -> 1); kind 1: centred binomial from bytes 0..5 (8 + 8 + 5
// coins a side).  The nonce is the coefficient index, the same value goes under every prime.
template <bool TERNARY>
__global__ void ref_sample_small_kernel(u64 *out, RefSeed seed, const Mod *__restrict__ mods, PrimeMap pm, u32 limbs, u32 N) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    u32 b[16];
    ref_block(seed, i, b);
    int v;
    if (TERNARY) {
        v = (int) ((b[0] & 0xFFu) % 3u) - 1;
    } else {
        v = __popc(b[0] & 0xFFu) + __popc((b[0] >> 8) & 0xFFu) + __popc((b[0] >> 16) & 0x1Fu) - __popc((b[0] >> 24) & 0xFFu) -
            __popc(b[1] & 0xFFu) - __popc((b[1] >> 8) & 0x1Fu);
    }
    for (u32 l = 0; l < limbs; l++) {
        const u64 q = mods[pm.prime(l)].q;
        out[(size_t) l * N + i] = v < 0 ? q + (u64) (long long) v : (u64) v;
    }
}

// 8 coefficients of one limb per block; a word above max_multiple is replaced by regenerating the WHOLE block under the
// nonce tid + tries * N * limbs and looking at the same word again, as the reference's loop does
__global__ void ref_sample_uniform_kernel(u64 *out, RefSeed seed, const Mod *__restrict__ mods, PrimeMap pm, u32 limbs, u32 N) {
    const u32 tid = blockIdx.x * blockDim.x + threadIdx.x, per_limb = N / 8;
    if (tid >= limbs * per_limb) return;
    const Mod m = mods[pm.prime(tid / per_limb)];
    const u64 max_random = ~0ull;
    const u64 max_multiple = max_random - barrett64(max_random, m) - 1;
    u32 b[16];
    ref_block(seed, tid, b);
    u64 tries = 1;
    for (int index = 0; index < 8; index++) {
        u64 r;
        for (;;) {
            r = 0;
#pragma unroll
            for (int k = 0; k < 8; k++)   // select without dynamic register indexing
                if (k == index) r = (u64) b[2 * k] | ((u64) b[2 * k + 1] << 32);
            if (r <= max_multiple) break;
            ref_block(seed, (u64) tid + tries * (u64) N * limbs, b);
            tries++;
        }
        out[(size_t) tid * 8 + index] = barrett64(r, m);
    }
}

// ------------------------------------------------------------------------------------------- fused element-wise
// c0[l][c] = -(a[l][c] * s_new[p][c] + e[l][c])  (+ pmodq[l] * s_old[p][c] on limbs [inj_lo, inj_hi))  (+ plain[l][c])
// with p = prime(l): multiply_and_add_negate_rns_poly (+ multiply_temp_mod_and_add_rns_poly / add_rns_poly) of
// src/secretkey.cu:276-277, 403-405, 613-615 in one pass.  e is read from c0 (NTT form), the secrets are laid out by
// prime index.
struct InjectScalars {
    u64 w[kMaxLimbs], ws[kMaxLimbs];
};
__global__ void enc_zero_finish_kernel(u64 *c0, const u64 *__restrict__ a, const u64 *__restrict__ s_new,
                                       const u64 *__restrict__ s_old, const u64 *__restrict__ plain,
                                       const Mod *__restrict__ mods, PrimeMap pm, u32 N, u32 inj_lo, u32 inj_hi,
                                       const __grid_constant__ InjectScalars inj) {
    const u32 l = blockIdx.y, p = pm.prime(l);
    const u32 c = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    if (c >= N) return;
    const Mod m = mods[p];
    const size_t o = (size_t) l * N + c, k = (size_t) p * N + c;
    const ulonglong2 av = ldg2(a + o), sv = ldg2(s_new + k), ev = ld2(c0 + o);
    u64 r0 = m.q - addmod(mulmod(av.x, sv.x, m), ev.x, m.q), r1 = m.q - addmod(mulmod(av.y, sv.y, m), ev.y, m.q);
    r0 = csub(r0, m.q);  // -(0) = q -> 0
    r1 = csub(r1, m.q);
    if (s_old != nullptr && l >= inj_lo && l < inj_hi) {
        const ulonglong2 ov = ldg2(s_old + k);
        r0 = addmod(r0, shoup_mul(ov.x, inj.w[l], inj.ws[l], m.q, m.nq), m.q);
        r1 = addmod(r1, shoup_mul(ov.y, inj.w[l], inj.ws[l], m.q, m.nq), m.q);
    }
    if (plain != nullptr) {
        const ulonglong2 pv = ldg2(plain + o);
        r0 = addmod(r0, pv.x, m.q);
        r1 = addmod(r1, pv.y, m.q);
    }
    st2(c0 + o, r0, r1);
}

// out[z][l][c] = pk[z][p][c] * u[l][c] + out[z][l][c]   (z < 2; out holds NTT(e_z); multiply_and_add_rns_poly,
// src/secretkey.cu:61-62)
__global__ void pk_mul_add_kernel(u64 *out, const u64 *__restrict__ pk, const u64 *__restrict__ u,
                                  const Mod *__restrict__ mods, PrimeMap pm, u32 N, u32 limbs, u32 size_QP) {
    const u32 l = blockIdx.y, z = blockIdx.z, p = pm.prime(l);
    const u32 c = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    if (c >= N) return;
    const Mod m = mods[p];
    const size_t o = ((size_t) z * limbs + l) * N + c;
    const ulonglong2 kv = ldg2(pk + ((size_t) z * size_QP + p) * N + c), uv = ldg2(u + (size_t) l * N + c), ev = ld2(out + o);
    st2(out + o, addmod(mulmod(kv.x, uv.x, m), ev.x, m.q), addmod(mulmod(kv.y, uv.y, m), ev.y, m.q));
}

// out[l][c] = c0 + c1 * s (+ c2 * s2)   (ckks_decrypt, src/secretkey.cu:668-678); secrets by prime index = limb
__global__ void decrypt_kernel(u64 *out, const u64 *__restrict__ ct, const u64 *__restrict__ s,
                               const u64 *__restrict__ s2, const Mod *__restrict__ mods, u32 N, u32 limbs, u32 polys) {
    const u32 l = blockIdx.y;
    const u32 c = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    if (c >= N) return;
    const Mod m = mods[l];
    const size_t o = (size_t) l * N + c, P = (size_t) limbs * N;
    ulonglong2 r = ldg2(ct + o);
    if (polys > 1) {
        const ulonglong2 a = ldg2(ct + P + o), sv = ldg2(s + o);
        r.x = addmod(r.x, mulmod(a.x, sv.x, m), m.q);
        r.y = addmod(r.y, mulmod(a.y, sv.y, m), m.q);
    }
    if (polys > 2) {
        const ulonglong2 a = ldg2(ct + 2 * P + o), sv = ldg2(s2 + o);
        r.x = addmod(r.x, mulmod(a.x, sv.x, m), m.q);
        r.y = addmod(r.y, mulmod(a.y, sv.y, m), m.q);
    }
    st2(out + o, r.x, r.y);
}

// ------------------------------------------------------------------------------------------- host side
void KeyGen::cu(cudaError_t e, const char *what) const {
    if (e != cudaSuccess) throw EvalError(B200FHE_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

static uint32_t inverse_mod_pow2_u32(uint32_t a, uint32_t mask) {  // a odd; Newton: x <- x (2 - a x)
    uint32_t x = a;
    for (int i = 0; i < 5; i++) x *= 2u - a * x;
    return x & mask;
}

KeyGen::KeyGen(Context *ctx, uint64_t reference_master_seed, cudaStream_t stream, RefPrng) : c(ctx), st(stream) {
    init_common();
    ref_mode_ = true;
    ref_master_ = reference_master_seed;
    if (c->size_Q % c->size_P) throw EvalError(B200FHE_ERR_UNSUPPORTED, "reference-PRNG key generator needs |P| to divide |Q| (the reference's dnum = |Q| / |P|)");
    // gen_secretkey (src/secretkey.cu:454-491): two seeds are drawn, the second one is used
    uint8_t seed[64];
    next_seed(seed);
    next_seed(seed);
    sample_ref(0, sk_.as<u64>(), c->size_QP, 0, seed);
    const LimbMap m = c->limb_map(0, c->size_QP, 0, 1, 0);
    cu(ntt_forward(c->T, sk_.as<u64>(), sk_.as<u64>(), m, nullptr, st), "ntt(secret)");
    count_launches(ntt_launch_count_forward(c->logN));
}

void KeyGen::init_common() {
    if (!c) throw EvalError(B200FHE_ERR_ARG, "null context");
    if (c->host_only) throw EvalError(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    if (c->size_P == 0) throw EvalError(B200FHE_ERR_ARG, "context has no special primes (size_P == 0)");
    if (c->N < 16) throw EvalError(B200FHE_ERR_UNSUPPORTED, "key generation needs N >= 16");
    if (c->size_QP > (size_t) kMaxLimbs) throw EvalError(B200FHE_ERR_UNSUPPORTED, "more limbs than kMaxLimbs");
    dnum = (c->size_Q + c->size_P - 1) / c->size_P;
    const size_t words = c->size_QP * c->N;
    cu(sk_.alloc(words * 8), "secret key");
    cu(sk2_.alloc(words * 8), "secret key square");
    cu(tmp_.alloc(3 * words * 8), "key generation workspace");
    cu(tmp2_.alloc(words * 8), "key generation workspace");
}

KeyGen::KeyGen(Context *ctx, const uint8_t *seed, cudaStream_t stream) : c(ctx), st(stream) {
    if (!c) throw EvalError(B200FHE_ERR_ARG, "null context");
    if (c->host_only) throw EvalError(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    if (c->size_P == 0) throw EvalError(B200FHE_ERR_ARG, "context has no special primes (size_P == 0)");
    if (c->N < 16) throw EvalError(B200FHE_ERR_UNSUPPORTED, "key generation needs N >= 16");
    if (c->size_QP > (size_t) kMaxLimbs) throw EvalError(B200FHE_ERR_UNSUPPORTED, "more limbs than kMaxLimbs");
    if (seed) {
        std::memcpy(key_, seed, 32);
    } else {
        std::random_device rd;
        for (int i = 0; i < 8; i++) key_[i] = rd();
    }
    dnum = (c->size_Q + c->size_P - 1) / c->size_P;
    const size_t words = c->size_QP * c->N;
    cu(sk_.alloc(words * 8), "secret key");
    cu(sk2_.alloc(words * 8), "secret key square");
    cu(tmp_.alloc(3 * words * 8), "key generation workspace");
    cu(tmp2_.alloc(words * 8), "key generation workspace");
    // gen_secretkey (src/secretkey.cu:454-491): ternary, the same coefficients under every prime, kept in NTT form
    sample(0, sk_.as<u64>(), c->size_QP, 0);
    const LimbMap m = c->limb_map(0, c->size_QP, 0, 1, 0);
    cu(ntt_forward(c->T, sk_.as<u64>(), sk_.as<u64>(), m, nullptr, st), "ntt(secret)");
    count_launches(ntt_launch_count_forward(c->logN));
}

void KeyGen::set_secret(const u64 *dev) {
    if (!dev) throw EvalError(B200FHE_ERR_ARG, "null secret key");
    cu(cudaMemcpyAsync(sk_.p, dev, c->size_QP * c->N * 8, cudaMemcpyDeviceToDevice, st), "copy secret");
    have_sk2_ = false;
}

void KeyGen::sample(int kind, u64 *out, size_t limbs, size_t size_Ql) {
    if (!out || limbs == 0 || limbs > c->size_QP || kind < 0 || kind > 2) throw EvalError(B200FHE_ERR_ARG, "bad sampler argument");
    if (ref_mode_) throw EvalError(B200FHE_ERR_UNSUPPORTED, "reference-PRNG key generator: samplers take explicit seeds");
    ChaKey k;
    std::memcpy(k.k, key_, 32);
    const uint64_t stream = next_stream_++;
    const u32 n0 = (u32) stream, n1 = (u32) (stream >> 32), N = (u32) c->N;
    const PrimeMap pm = c->prime_map(size_Ql);
    const Mod *mods = c->d_mods.as<Mod>();
    if (kind == 0)
        sample_small_kernel<true><<<(N / 16 + kKT - 1) / kKT, kKT, 0, st>>>(out, k, n0, n1, mods, pm, (u32) limbs, N);
    else if (kind == 1)
        sample_small_kernel<false><<<(N / 8 + kKT - 1) / kKT, kKT, 0, st>>>(out, k, n0, n1, mods, pm, (u32) limbs, N);
    else
        sample_uniform_kernel<<<((u32) limbs * (N / 8) + kKT - 1) / kKT, kKT, 0, st>>>(out, k, n0, n1, mods, pm, (u32) limbs, N);
    cu(cudaGetLastError(), "sampler");
    count_launches(1);
}

// std::random_device stand-in of the reference-compatible mode: byte i of the stream is the top byte of
// splitmix64(master + i); the reference's random_bytes() (include/prng.cuh:12-21) takes one byte per 32-bit draw
static uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
void KeyGen::next_seed(uint8_t out[64]) {
    for (int i = 0; i < 64; i++) out[i] = (uint8_t) (splitmix64(ref_master_ + ref_drawn_++) >> 56);
}
void KeyGen::sample_ref(int kind, u64 *out, size_t limbs, size_t size_Ql, const uint8_t seed[64]) {
    RefSeed k;
    std::memcpy(k.w, seed, 56);
    const u32 N = (u32) c->N;
    const PrimeMap pm = c->prime_map(size_Ql);
    const Mod *mods = c->d_mods.as<Mod>();
    if (kind == 0) ref_sample_small_kernel<true><<<(N + kKT - 1) / kKT, kKT, 0, st>>>(out, k, mods, pm, (u32) limbs, N);
    else if (kind == 1) ref_sample_small_kernel<false><<<(N + kKT - 1) / kKT, kKT, 0, st>>>(out, k, mods, pm, (u32) limbs, N);
    else ref_sample_uniform_kernel<<<((u32) limbs * (N / 8) + kKT - 1) / kKT, kKT, 0, st>>>(out, k, mods, pm, (u32) limbs, N);
    cu(cudaGetLastError(), "reference sampler");
    count_launches(1);
}

void KeyGen::enc_zero(u64 *c0, u64 *c1, const u64 *s_new, size_t limbs, PrimeMap pm, const u64 *s_old, size_t inj_lo,
                      size_t inj_hi, const u64 *plain) {
    const u32 N = (u32) c->N;
    const LimbMap m = c->limb_map(0, limbs, pm.size_Ql, 1, 0);
    if (ref_mode_) {
        // the reference draws the seed of `a` before it calls encrypt_zero_symmetric, which draws the error seed
        // (src/secretkey.cu:388-390, 254-265)
        uint8_t seed_a[64], seed_e[64];
        next_seed(seed_a);
        next_seed(seed_e);
        sample_ref(1, c0, limbs, pm.size_Ql, seed_e);
        cu(ntt_forward(c->T, c0, c0, m, nullptr, st), "ntt(error)");
        sample_ref(2, c1, limbs, pm.size_Ql, seed_a);
    } else {
        sample(1, c0, limbs, pm.size_Ql);
        cu(ntt_forward(c->T, c0, c0, m, nullptr, st), "ntt(error)");
        sample(2, c1, limbs, pm.size_Ql);
    }
    InjectScalars inj{};
    if (s_old) {
        // P mod q_j (DRNSTool bigP_mod_q, src/rns.cu): the factor of multiply_temp_mod_and_add_rns_poly
        DeviceLevel *L = nullptr;
        if (int rc = get_level(c, c->size_Q, &L)) throw EvalError(rc, b200fhe_last_error());
        for (size_t l = inj_lo; l < inj_hi; l++) {
            inj.w[l] = L->h.bigP_mod_q[l];
            inj.ws[l] = L->h.bigP_mod_q_shoup[l];
        }
    }
    dim3 grid((N / 2 + kKT - 1) / kKT, (u32) limbs);
    enc_zero_finish_kernel<<<grid, kKT, 0, st>>>(c0, c1, s_new, s_old, plain, c->d_mods.as<Mod>(), pm, N, (u32) inj_lo,
                                                 (u32) inj_hi, inj);
    cu(cudaGetLastError(), "enc_zero_finish");
    count_launches(1 + ntt_launch_count_forward(c->logN));
}

// generate_one_kswitch_key[_fused] (src/secretkey.cu:362-452): digit d is an encryption of zero under s_new over all
// of QP whose first polynomial carries P * s_old on the limbs of digit d
std::unique_ptr<SwitchKey> KeyGen::switch_key(const u64 *s_old, const u64 *s_new) {
    auto key = std::make_unique<SwitchKey>();
    const size_t words = c->size_QP * c->N, alpha = c->size_P;
    for (size_t d = 0; d < dnum; d++) {
        auto buf = std::make_unique<DevBuf>();
        if (buf->alloc(2 * words * 8) != cudaSuccess) throw EvalError(B200FHE_ERR_ALLOC, "device memory for an evaluation key");
        u64 *p = buf->as<u64>();
        enc_zero(p, p + words, s_new, c->size_QP, PrimeMap{0, (u32) c->size_Q}, s_old, d * alpha,
                 std::min((d + 1) * alpha, c->size_Q), nullptr);
        key->digits.push_back(std::move(buf));
    }
    return key;
}

std::unique_ptr<SwitchKey> KeyGen::relin_key() {
    if (!have_sk2_) {  // compute_secret_key_array (src/secretkey.cu:196-230), power 2
        cu(eltwise(c->d_mods.as<Mod>(), ELT_MUL, sk2_.as<u64>(), sk_.as<u64>(), sk_.as<u64>(), (u32) c->size_QP,
                   PrimeMap{0, (u32) c->size_Q}, c->N, st),
           "s^2");
        count_launches(1);
        have_sk2_ = true;
    }
    return switch_key(sk2_.as<u64>(), sk_.as<u64>());
}

std::unique_ptr<SwitchKey> KeyGen::galois_key(uint32_t auto_index) {
    const uint32_t M = 2 * (uint32_t) c->N;
    if (!(auto_index & 1) || auto_index >= M) throw EvalError(B200FHE_ERR_ARG, "Invalid in EvalAtIndexKeyGen");
    const uint32_t inv = inverse_mod_pow2_u32(auto_index, M - 1);
    cu(perm_.alloc(c->N * 4), "perm");
    u32 *perm = perm_.as<u32>();
    cu(automap_table(perm, c->logN, inv, st), "automap_table");
    cu(gather_limbs(tmp2_.as<u64>(), sk_.as<u64>(), perm, (u32) c->size_QP, 1, 0, 0, c->N, st), "sigma(secret)");
    count_launches(2);
    return switch_key(sk_.as<u64>(), tmp2_.as<u64>());
}

// PhantomSecretKey::create_galois_keys (src/secretkey.cu:532-574): the key of Galois element g switches FROM sigma_g(s)
// back to s, for the apply-the-automorphism-then-key-switch rotation of apply_galois_inplace / hoisting_inplace
std::unique_ptr<SwitchKey> KeyGen::galois_key_seal(uint32_t galois_elt) {
    const uint32_t M = 2 * (uint32_t) c->N;
    if (!(galois_elt & 1) || galois_elt >= M) throw EvalError(B200FHE_ERR_ARG, "Galois element is not valid");
    cu(perm_.alloc(c->N * 4), "perm");
    u32 *perm = perm_.as<u32>();
    cu(automap_table(perm, c->logN, galois_elt, st), "automap_table");
    cu(gather_limbs(tmp2_.as<u64>(), sk_.as<u64>(), perm, (u32) c->size_QP, 1, 0, 0, c->N, st), "sigma(secret)");
    count_launches(2);
    return switch_key(tmp2_.as<u64>(), sk_.as<u64>());
}

void KeyGen::public_key(u64 *out) {  // gen_publickey (src/secretkey.cu:493-505)
    if (!out) throw EvalError(B200FHE_ERR_ARG, "null output");
    enc_zero(out, out + c->size_QP * c->N, sk_.as<u64>(), c->size_QP, PrimeMap{0, (u32) c->size_Q}, nullptr, 0, 0, nullptr);
}

void KeyGen::encrypt_symmetric(const u64 *plain, size_t size_Ql, u64 *ct) {
    if (!plain || !ct || size_Ql == 0 || size_Ql > c->size_Q) throw EvalError(B200FHE_ERR_ARG, "bad encryption argument");
    enc_zero(ct, ct + size_Ql * c->N, sk_.as<u64>(), size_Ql, PrimeMap{0, (u32) c->size_Q}, nullptr, 0, 0, plain);
}

void KeyGen::encrypt_asymmetric(const u64 *pk, const u64 *plain, size_t size_Ql, u64 *ct) {
    if (!pk || !plain || !ct || size_Ql == 0 || size_Ql > c->size_Q) throw EvalError(B200FHE_ERR_ARG, "bad encryption argument");
    const u32 N = (u32) c->N;
    const size_t QlP = size_Ql + c->size_P, LN = QlP * c->N;
    const PrimeMap pm{(u32) size_Ql, (u32) c->size_Q};
    u64 *u = tmp_.as<u64>(), *e = u + c->size_QP * c->N;  // e: [2][QlP][N]
    sample(0, u, QlP, size_Ql);
    sample(1, e, QlP, size_Ql);
    sample(1, e + LN, QlP, size_Ql);
    cu(ntt_forward(c->T, u, u, c->limb_map(0, QlP, size_Ql, 1, 0), nullptr, st), "ntt(u)");
    cu(ntt_forward(c->T, e, e, c->limb_map(0, QlP, size_Ql, 2, LN), nullptr, st), "ntt(e)");
    dim3 grid((N / 2 + kKT - 1) / kKT, (u32) QlP, 2);
    pk_mul_add_kernel<<<grid, kKT, 0, st>>>(e, pk, u, c->d_mods.as<Mod>(), pm, N, (u32) QlP, (u32) c->size_QP);
    cu(cudaGetLastError(), "pk_mul_add");
    count_launches(1 + 2 * ntt_launch_count_forward(c->logN));
    // divide by P (the reference's moddown of the level above, src/secretkey.cu:112-121) and add the plaintext
    DeviceLevel *L = nullptr;
    if (int rc = get_level(c, size_Ql, &L)) throw EvalError(rc, b200fhe_last_error());
    if (int rc = do_moddown(c, L, ct, size_Ql * c->N, e, LN, 2, plain, 0, 1, st)) throw EvalError(rc, b200fhe_last_error());
}

void KeyGen::decrypt(const u64 *ct, size_t size_Ql, u32 polys, u64 *plain_out) {
    if (!ct || !plain_out || size_Ql == 0 || size_Ql > c->size_Q || polys == 0 || polys > 3)
        throw EvalError(B200FHE_ERR_ARG, "bad decryption argument");
    if (polys == 3 && !have_sk2_) {
        cu(eltwise(c->d_mods.as<Mod>(), ELT_MUL, sk2_.as<u64>(), sk_.as<u64>(), sk_.as<u64>(), (u32) c->size_QP,
                   PrimeMap{0, (u32) c->size_Q}, c->N, st),
           "s^2");
        count_launches(1);
        have_sk2_ = true;
    }
    const u32 N = (u32) c->N;
    dim3 grid((N / 2 + kKT - 1) / kKT, (u32) size_Ql);
    decrypt_kernel<<<grid, kKT, 0, st>>>(plain_out, ct, sk_.as<u64>(), sk2_.as<u64>(), c->d_mods.as<Mod>(), N, (u32) size_Ql,
                                         polys);
    cu(cudaGetLastError(), "decrypt");
    count_launches(1);
}

int chacha20_block_device(const uint32_t *key8, uint32_t counter, const uint32_t *nonce3, uint32_t *out16_host) {
    ChaKey k;
    std::memcpy(k.k, key8, 32);
    DevBuf d;
    if (d.alloc(64) != cudaSuccess) return fail(B200FHE_ERR_ALLOC, "device memory");
    chacha20_block_kernel<<<1, 1>>>(k, counter, nonce3[0], nonce3[1], nonce3[2], d.as<u32>());
    const cudaError_t e = cudaMemcpy(out16_host, d.p, 64, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return fail(B200FHE_ERR_CUDA, "chacha20 block: %s", cudaGetErrorString(e));
    count_launches(1);
    return B200FHE_OK;
}

}  // namespace b200fhe
