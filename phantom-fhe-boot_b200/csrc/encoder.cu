// encoder.cu — CKKS encoder / decoder (see encoder.h).
#include "encoder.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <stdexcept>

#include "../../include/b200fhe.h"

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(B200FHE_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

namespace b200fhe {

static constexpr int kET = 256;
static constexpr int kMaxWords = 48;  // multi-precision words of a composed coefficient (|Q_l P| <= 48 limbs)

// ------------------------------------------------------------------------------------------- small host bigints
namespace {
typedef std::vector<uint64_t> Big;  // little-endian 64-bit words
Big big_mul_word(const Big &a, uint64_t w) {
    Big r(a.size() + 1, 0);
    unsigned __int128 carry = 0;
    for (size_t i = 0; i < a.size(); i++) {
        carry += (unsigned __int128) a[i] * w;
        r[i] = (uint64_t) carry;
        carry >>= 64;
    }
    r[a.size()] = (uint64_t) carry;
    return r;
}
uint64_t big_mod_word(const Big &a, uint64_t q) {
    unsigned __int128 r = 0;
    for (size_t i = a.size(); i--;) r = ((r << 64) | a[i]) % q;
    return (uint64_t) r;
}
int big_bits(const Big &a) {
    for (size_t i = a.size(); i--;)
        if (a[i]) return (int) (64 * i + 64 - __builtin_clzll(a[i]));
    return 0;
}
Big big_product(const std::vector<uint64_t> &primes, size_t skip) {
    Big r{1};
    for (size_t i = 0; i < primes.size(); i++)
        if (i != skip) r = big_mul_word(r, primes[i]);
    r.resize(primes.size(), 0);  // the product of n word-sized primes fits n words
    return r;
}
}  // namespace

// ------------------------------------------------------------------------------------------- root table (host)
// util::ComplexRoots (src/fft.cu:15-58): one eighth of the m-th roots from cos/sin, the rest by symmetry — the same
// expressions, so the table holds the same doubles.
namespace {
struct ComplexRoots {
    size_t m;
    std::vector<cuDoubleComplex> r;
    explicit ComplexRoots(size_t degree) : m(degree), r(degree / 8 + 1) {
        const double PI = 3.1415926535897932384626433832795028842;
        for (size_t i = 0; i <= m / 8; i++) {
            const double theta = 2 * PI * static_cast<double>(i) / static_cast<double>(m);
            r[i] = make_cuDoubleComplex(1.0 * std::cos(theta), 1.0 * std::sin(theta));
        }
    }
    cuDoubleComplex get(size_t index) const {
        index &= m - 1;
        if (index <= m / 8) return r[index];
        if (index <= m / 4) {
            const cuDoubleComplex a = r[m / 4 - index];
            return make_cuDoubleComplex(a.y, a.x);
        }
        if (index <= m / 2) return cuCsub(make_cuDoubleComplex(0, 0), cuConj(get(m / 2 - index)));
        if (index <= 3 * m / 4) return cuCsub(make_cuDoubleComplex(0, 0), get(index - m / 2));
        return cuConj(get(m - index));
    }
};
}  // namespace

// carried_group: the reference's encode_sparse_ext tables (set_sparse_encode, include/ckks.h:400-447) start the
// rotation group not at 1 but at the value its running `pos` was left with by the loop over the half-size ring
// (5^(n/4) mod 2n): the slots come out cyclically rotated.  Reproduced so that sparse extended plaintexts match.
const FftTables &Encoder::tables(uint32_t n, bool carried_group) {
    const uint32_t key = n | (carried_group ? 0x80000000u : 0u);
    auto it = tabs_.find(key);
    if (it != tabs_.end()) return *it->second;
    auto t = std::make_unique<FftTables>();
    t->n = n;
    t->logn = 0;
    while ((1u << t->logn) < n) t->logn++;
    t->m = 4 * n;
    std::vector<cuDoubleComplex> roots(t->m);
    if (t->m >= 8) {
        ComplexRoots cr(t->m);
        for (size_t i = 0; i < t->m; i++) roots[i] = cr.get(i);
    } else {
        roots[0] = make_cuDoubleComplex(1, 0);
        roots[1] = make_cuDoubleComplex(0, 1);
        roots[2] = make_cuDoubleComplex(-1, 0);
        roots[3] = make_cuDoubleComplex(0, -1);
    }
    std::vector<uint32_t> group(std::max<uint32_t>(n / 2, 1));
    uint32_t pos = 1;
    if (carried_group)
        for (uint32_t i = 0; i < n / 4; i++) pos = (pos * 5u) & (2 * n - 1);
    for (uint32_t i = 0; i < n / 2; i++) {
        group[i] = pos;
        pos = (pos * 5u) & (t->m - 1);
    }
    if (t->roots.upload(roots.data(), roots.size() * sizeof(cuDoubleComplex)) != cudaSuccess ||
        t->group.upload(group.data(), group.size() * sizeof(uint32_t)) != cudaSuccess)
        throw std::runtime_error("encoder tables: device allocation failed");
    return *(tabs_[key] = std::move(t));
}

Encoder::Encoder(Context *ctx) : c(ctx), slots(ctx->N / 2) {}

// ------------------------------------------------------------------------------------------- kernels
__global__ void bit_reverse_scatter(cuDoubleComplex *dst, const cuDoubleComplex *src, uint32_t count, uint32_t logn) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) dst[logn ? __brev(i) >> (32 - logn) : 0] = src[i];
}

// Complex product with the rounding sequence pinned.  nvcc contracts cuCmul's  a.x*w.y + a.y*w.x  differently in the
// reference's two kernel families (SASS of src/fft.cu built for sm_100a): the shared-memory "base" kernels, which run
// the stages with butterfly distance < 2048, round w.y*a.x first and fuse the other product; the global-memory "iter"
// kernels (distance >= 2048) round w.x*a.y first.  The real part is fma(w.x, a.x, -(w.y*a.y)) in both.  One double
// rounded the other way moves an encoded coefficient by one, so the choice is made per stage here.
template <bool ITER_FORM>
__device__ __forceinline__ cuDoubleComplex cmul_pinned(cuDoubleComplex a, cuDoubleComplex w) {
    const double re = __fma_rn(w.x, a.x, -__dmul_rn(w.y, a.y));
    const double im = ITER_FORM ? __fma_rn(w.y, a.x, __dmul_rn(w.x, a.y)) : __fma_rn(w.x, a.y, __dmul_rn(w.y, a.x));
    return make_cuDoubleComplex(re, im);
}

// one stage of the special FFT (src/fft.cu:275-316 forward, :547-594 inverse): butterflies at distance 2^logPairs,
// twiddle = root^(5^brev(k) * 2^logPairs) (forward) or its inverse
template <bool INVERSE, bool ITER_FORM>
__global__ void special_fft_stage(cuDoubleComplex *io, const cuDoubleComplex *__restrict__ roots,
                                  const uint32_t *__restrict__ group, uint32_t n, uint32_t logn, uint32_t logPairs,
                                  uint32_t M, double scalar) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= (n >> 1)) return;
    const uint32_t pairs = 1u << logPairs;
    const uint32_t k = tid >> logPairs, j = tid & (pairs - 1), g = 2 * k * pairs + j;
    uint32_t psi = logn > 1 ? group[__brev(k << logPairs) >> (33 - logn)] : 1u;
    psi = (psi << logPairs) & (M - 1);
    cuDoubleComplex x0 = io[g], x1 = io[g + pairs];
    if (INVERSE) {
        const cuDoubleComplex w = roots[M - psi];
        const cuDoubleComplex s0 = make_cuDoubleComplex(__dadd_rn(x0.x, x1.x), __dadd_rn(x0.y, x1.y));
        const cuDoubleComplex s1 = make_cuDoubleComplex(__dsub_rn(x0.x, x1.x), __dsub_rn(x0.y, x1.y));
        x0 = s0;
        x1 = cmul_pinned<ITER_FORM>(s1, w);
        if (scalar != 0.0) {  // last stage only (the caller passes 0 otherwise)
            x0 = make_cuDoubleComplex(__dmul_rn(x0.x, scalar), __dmul_rn(x0.y, scalar));
            x1 = make_cuDoubleComplex(__dmul_rn(x1.x, scalar), __dmul_rn(x1.y, scalar));
        }
    } else {
        const cuDoubleComplex w = roots[psi];
        const cuDoubleComplex t = cmul_pinned<ITER_FORM>(x1, w);
        x1 = make_cuDoubleComplex(__dsub_rn(x0.x, t.x), __dsub_rn(x0.y, t.y));
        x0 = make_cuDoubleComplex(__dadd_rn(x0.x, t.x), __dadd_rn(x0.y, t.y));
    }
    io[g] = x0;
    io[g + pairs] = x1;
}

template <bool INVERSE>
static void launch_fft_stage(unsigned blocks, cudaStream_t st, cuDoubleComplex *io, const FftTables &T, uint32_t logPairs,
                             double scalar) {
    // SWITCH_POINT = 2048 (include/common.h:13): distances below it are the reference's shared-memory kernels
    if (logPairs >= 11)
        special_fft_stage<INVERSE, true><<<blocks, kET, 0, st>>>(io, T.roots.as<cuDoubleComplex>(), T.group.as<uint32_t>(), T.n,
                                                                 T.logn, logPairs, T.m, scalar);
    else
        special_fft_stage<INVERSE, false><<<blocks, kET, 0, st>>>(io, T.roots.as<cuDoubleComplex>(), T.group.as<uint32_t>(), T.n,
                                                                  T.logn, logPairs, T.m, scalar);
}

__global__ void max_abs_kernel(const cuDoubleComplex *v, uint32_t n, unsigned long long *out) {
    double m = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        m = fmax(m, fmax(fabs(v[i].x), fabs(v[i].y)));
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(out, (unsigned long long) __double_as_longlong(m));  // m >= 0: order-preserving
}

// DRNSBase::decompose_array (src/rns_base.cu:49-171): round, split the magnitude into 64-bit words, reduce modulo
// every prime, negate by q - r (a negative zero gives q, as in the reference; the NTT that follows canonicalises it)
__global__ void decompose_kernel(const Mod *__restrict__ mods, PrimeMap pm, u64 *dst, const cuDoubleComplex *src,
                                 uint32_t degree, uint32_t half, size_t dst_stride, uint32_t coeff_step, int words) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= degree) return;
    const Mod md = mods[pm.prime(blockIdx.y)];
    double v = i < half ? src[i].x : src[i - half].y;
    v = round(v);
    const bool neg = signbit(v);
    v = fabs(v);
    const double two64 = 18446744073709551616.0;
    u64 r;
    if (words <= 1) {
        r = barrett64((u64) v, md);
    } else if (words == 2) {
        u128 x;
        x.lo = (u64) fmod(v, two64);
        x.hi = (u64) (v / two64);
        r = barrett128(x, md);
    } else {
        u64 w[24];  // up to 1536-bit magnitudes
        for (int k = 0; k < words && k < 24; k++) {
            if (v >= 1) {
                w[k] = (u64) fmod(v, two64);
                v /= two64;
            } else {
                w[k] = 0;
            }
        }
        const int nw = words < 24 ? words : 24;
        u64 hi = w[nw - 1] % md.q;
        for (int k = nw - 2; k >= 0; k--) {
            u128 x;
            x.lo = w[k];
            x.hi = hi;
            hi = barrett128(x, md);
        }
        r = hi;
    }
    if (neg) r = md.q - r;
    dst[size_t(blockIdx.y) * dst_stride + size_t(i) * coeff_step] = r;
}

// Whole inverse special FFT of one SMALL message (n <= 2048 values) per CTA, in shared memory: bit-reversal, the logn
// stages of special_fft_stage<true, false> (every butterfly distance is below the reference's SWITCH_POINT, so all
// stages use the shared-memory rounding form), the final scaling, and the running maximum of the batch item.
// Same operations on the same operands as the stage-by-stage path, hence the same doubles.
__global__ void __launch_bounds__(kET) special_ifft_small_batch(cuDoubleComplex *out, const cuDoubleComplex *values,
                                                                const cuDoubleComplex *__restrict__ roots,
                                                                const uint32_t *__restrict__ group, uint32_t n, uint32_t logn,
                                                                uint32_t M, double scalar, unsigned long long *maxes) {
    extern __shared__ __align__(16) unsigned char fft_smem[];
    cuDoubleComplex *io = reinterpret_cast<cuDoubleComplex *>(fft_smem);
    const size_t b = blockIdx.x;
    const cuDoubleComplex *src = values + b * n;
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) io[__brev(i) >> (32 - logn)] = src[i];
    __syncthreads();
    for (uint32_t logPairs = 0; logPairs < logn; logPairs++) {
        const uint32_t pairs = 1u << logPairs;
        const bool last = logPairs + 1 == logn;
        for (uint32_t tid = threadIdx.x; tid < (n >> 1); tid += blockDim.x) {
            const uint32_t k = tid >> logPairs, j = tid & (pairs - 1), g = 2 * k * pairs + j;
            uint32_t psi = logn > 1 ? group[__brev(k << logPairs) >> (33 - logn)] : 1u;
            psi = (psi << logPairs) & (M - 1);
            cuDoubleComplex x0 = io[g], x1 = io[g + pairs];
            const cuDoubleComplex w = roots[M - psi];
            const cuDoubleComplex s0 = make_cuDoubleComplex(__dadd_rn(x0.x, x1.x), __dadd_rn(x0.y, x1.y));
            const cuDoubleComplex s1 = make_cuDoubleComplex(__dsub_rn(x0.x, x1.x), __dsub_rn(x0.y, x1.y));
            x0 = s0;
            x1 = cmul_pinned<false>(s1, w);
            if (last) {
                x0 = make_cuDoubleComplex(__dmul_rn(x0.x, scalar), __dmul_rn(x0.y, scalar));
                x1 = make_cuDoubleComplex(__dmul_rn(x1.x, scalar), __dmul_rn(x1.y, scalar));
            }
            io[g] = x0;
            io[g + pairs] = x1;
        }
        __syncthreads();
    }
    double m = 0;
    cuDoubleComplex *dst = out + b * n;
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        const cuDoubleComplex v = io[i];
        dst[i] = v;
        m = fmax(m, fmax(fabs(v.x), fabs(v.y)));
    }
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(maxes + b, (unsigned long long) __double_as_longlong(m));
}

// decompose_kernel over a batch: item z reads src + z * half complex values, writes dst + z * limbs * dst_stride;
// the word count of each item comes from its own maximum (words8[z])
__global__ void decompose_batch_kernel(const Mod *__restrict__ mods, u64 *dst, const cuDoubleComplex *src, uint32_t degree,
                                       uint32_t half, size_t dst_stride, uint32_t coeff_step, const uint8_t *__restrict__ words8) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= degree) return;
    const size_t z = blockIdx.z;
    const Mod md = mods[blockIdx.y];
    const cuDoubleComplex *s = src + z * half;
    double v = i < half ? s[i].x : s[i - half].y;
    v = round(v);
    const bool neg = signbit(v);
    v = fabs(v);
    const double two64 = 18446744073709551616.0;
    const int words = words8[z];
    u64 r;
    if (words <= 1) {
        r = barrett64((u64) v, md);
    } else {
        u128 x;
        x.lo = (u64) fmod(v, two64);
        x.hi = (u64) (v / two64);
        r = barrett128(x, md);
    }
    if (neg) r = md.q - r;
    dst[(z * gridDim.y + blockIdx.y) * dst_stride + size_t(i) * coeff_step] = r;
}

// decompose_batch_kernel fused with what DNN::Conv does to every weight plaintext that meets a NoiseScaleDeg-2 input
// (EvalMultAutoInplace, src/evaluate.cu:2813-2825): EvalMultConstInplaceCore(pt, 1.0) -- a per-limb scalar `lift` -- and
// EvalModReduceInPlace -- (a_j - [a_last]_{q_j}) * q_last^-1 mod q_j.  Both are exact arithmetic mod q_j on each
// COEFFICIENT and the NTT is linear mod q_j, so doing them here, before the one forward NTT, gives the residues the
// reference gets from NTT -> scalar multiply -> INTT(last limb) -> NTT -> epilogue: 2.2x fewer limb transforms.
// Output: Ql - 1 limbs per item.  grid.y = output limb j, the item's last limb is recomputed by every j.
__global__ void decompose_lift_rescale_kernel(const Mod *__restrict__ mods, u64 *dst, const cuDoubleComplex *src, uint32_t degree,
                                              uint32_t half, size_t dst_stride, uint32_t coeff_step,
                                              const uint8_t *__restrict__ words8, LimbScalars lift, uint32_t Ql,
                                              const u64 *__restrict__ inv_q_last) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= degree) return;
    const size_t z = blockIdx.z;
    const uint32_t j = blockIdx.y;
    const Mod mj = mods[j], ml = mods[Ql - 1];
    const cuDoubleComplex *s = src + z * half;
    double v = i < half ? s[i].x : s[i - half].y;
    v = round(v);
    const bool neg = signbit(v);
    v = fabs(v);
    const double two64 = 18446744073709551616.0;
    u64 rj, rl;
    if (words8[z] <= 1) {
        rj = barrett64((u64) v, mj);
        rl = barrett64((u64) v, ml);
    } else {
        u128 x;
        x.lo = (u64) fmod(v, two64);
        x.hi = (u64) (v / two64);
        rj = barrett128(x, mj);
        rl = barrett128(x, ml);
    }
    if (neg) {  // q - r, canonical (the stand-alone kernel leaves q for a negative zero and lets the NTT fold it)
        rj = rj ? mj.q - rj : 0;
        rl = rl ? ml.q - rl : 0;
    }
    const u64 vj = mulmod(rj, lift.w[j], mj), vl = mulmod(rl, lift.w[Ql - 1], ml);
    const u64 t = barrett64(vl, mj);
    dst[(z * gridDim.y + j) * dst_stride + size_t(i) * coeff_step] = mulmod(submod(vj, t, mj.q), inv_q_last[j], mj);
}

// ------------------------------------------------------------------------------------------- encode
// encode_sparse of `batch` messages of n values each in ONE pass (the convolution layer encodes thousands of weight
// plaintexts per call): one CTA per message for the whole FFT, one host synchronisation for the magnitude check of all
// of them, one decomposition launch, one NTT launch pair over every limb of every plaintext.  Falls back to the
// per-message path for n > 2048 or coefficients wider than 128 bits.
int Encoder::encode_sparse_batch(const cuDoubleComplex *values, size_t n, size_t batch, size_t size_Ql, double scale, u64 *out,
                                 cudaStream_t st, const LimbScalars *lift, const u64 *inv_q_last, bool compact) {
    if (c->host_only) return fail(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    if (!values || n == 0 || batch == 0) return fail(B200FHE_ERR_ARG, "Input vector is empty");
    if (n > slots) return fail(B200FHE_ERR_ARG, "Input vector exceeds max slots");
    if (!out || size_Ql == 0 || size_Ql > c->size_Q) return fail(B200FHE_ERR_ARG, "bad argument");
    if (n & (n - 1)) return fail(B200FHE_ERR_ARG, "sparse encoding needs a power-of-two number of values");
    const size_t N = c->N;
    auto one_by_one = [&]() {
        for (size_t z = 0; z < batch; z++)
            if (int rc = encode(values + z * n, n, size_Ql, scale, false, true, out + z * size_Ql * N, st)) return rc;
        return (int) B200FHE_OK;
    };
    const bool fused = lift != nullptr;  // out: [batch][size_Ql - 1][N], lifted and rescaled
    if (fused && (!inv_q_last || size_Ql < 2)) return fail(B200FHE_ERR_ARG, "bad rescale argument");
    if (compact && (n > 1024 || n < 2 || batch > 65535)) return fail(B200FHE_ERR_UNSUPPORTED, "compact encode needs 2 <= n <= 1024");
    if (n > 2048 || n < 2 || batch > 65535) return fused ? fail(B200FHE_ERR_UNSUPPORTED, "fused encode needs n <= 2048") : one_by_one();
    const int total_bits = big_bits(big_product(std::vector<uint64_t>(c->primes.begin(), c->primes.begin() + size_Ql), size_Ql));
    if (scale <= 0 || static_cast<int>(std::log2(scale)) + 1 >= total_bits) return fail(B200FHE_ERR_ARG, "scale out of bounds");
    try {
        const FftTables &T = tables((uint32_t) n, false);
        const size_t vbytes = batch * n * sizeof(cuDoubleComplex);
        CU(buf_batch_in_.alloc(vbytes));
        CU(buf_batch_res_.alloc(vbytes));
        CU(buf_batch_aux_.alloc(batch * 16));
        cuDoubleComplex *vin = buf_batch_in_.as<cuDoubleComplex>(), *res = buf_batch_res_.as<cuDoubleComplex>();
        unsigned long long *d_max = buf_batch_aux_.as<unsigned long long>();
        uint8_t *d_words = reinterpret_cast<uint8_t *>(d_max + batch);
        CU(cudaMemcpyAsync(vin, values, vbytes, cudaMemcpyHostToDevice, st));
        CU(cudaMemsetAsync(d_max, 0, batch * 8, st));
        const double fix = scale / static_cast<double>(n);
        special_ifft_small_batch<<<(unsigned) batch, kET, n * sizeof(cuDoubleComplex), st>>>(
            res, vin, T.roots.as<cuDoubleComplex>(), T.group.as<uint32_t>(), (uint32_t) n, T.logn, T.m, fix, d_max);
        std::vector<unsigned long long> hmax(batch);
        CU(cudaMemcpyAsync(hmax.data(), d_max, batch * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        std::vector<uint8_t> words(batch);
        for (size_t z = 0; z < batch; z++) {
            double max_coeff;
            std::memcpy(&max_coeff, &hmax[z], 8);
            const int bits = static_cast<int>(std::ceil(std::log2(std::max(max_coeff, 1.0)))) + 1;
            if (bits >= total_bits) return fail(B200FHE_ERR_ARG, "encoded values are too large");
            if (bits > 128 && compact) return fail(B200FHE_ERR_UNSUPPORTED, "compact encode needs coefficients below 2^128");
            if (bits > 128) return fused ? fail(B200FHE_ERR_UNSUPPORTED, "fused encode needs coefficients below 2^128") : one_by_one();
            words[z] = bits <= 64 ? 1 : 2;
        }
        CU(cudaMemcpyAsync(d_words, words.data(), batch, cudaMemcpyHostToDevice, st));
        const uint32_t degree = 2 * (uint32_t) n, step = (uint32_t) (slots / n);
        const size_t out_limbs = fused ? size_Ql - 1 : size_Ql;
        // compact: only the `degree` coefficients that can be non-zero are stored ([batch][limbs][degree]) and transformed
        // (ntt_forward_sub): value k of the result stands for the N / degree entries [k * step, (k + 1) * step)
        const size_t row = compact ? degree : N;
        const uint32_t cstep = compact ? 1 : step;
        if (cstep > 1) CU(cudaMemsetAsync(out, 0, batch * out_limbs * N * 8, st));
        const dim3 dgrid((degree + kET - 1) / kET, (unsigned) out_limbs, (unsigned) batch);
        if (fused)
            decompose_lift_rescale_kernel<<<dgrid, kET, 0, st>>>(c->T.mods, out, res, degree, (uint32_t) n, row, cstep, d_words, *lift,
                                                                 (uint32_t) size_Ql, inv_q_last);
        else
            decompose_batch_kernel<<<dgrid, kET, 0, st>>>(c->T.mods, out, res, degree, (uint32_t) n, row, cstep, d_words);
        CU(cudaStreamSynchronize(st));  // `words` (pageable host memory) must outlive its asynchronous copy
        LimbMap m = c->limb_map(0, out_limbs, 0, batch, out_limbs * row);
        if (compact) {
            int lg = 0;
            while ((1u << lg) < degree) lg++;
            CU(ntt_forward_sub(c->T, lg, out, out, m, st));
            count_launches(3);
        } else {
            CU(ntt_forward(c->T, out, out, m, nullptr, st));
            count_launches(2 + ntt_launch_count_forward(c->logN));
        }
    } catch (const std::exception &e) {
        return fail(B200FHE_ERR_ALLOC, "%s", e.what());
    }
    return B200FHE_OK;
}

int Encoder::encode(const cuDoubleComplex *values, size_t n, size_t size_Ql, double scale, bool ext, bool sparse, u64 *out,
                    cudaStream_t st) {
    if (c->host_only) return fail(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    if (!values || n == 0) return fail(B200FHE_ERR_ARG, "Input vector is empty");
    if (n > slots) return fail(B200FHE_ERR_ARG, "Input vector exceeds max slots");
    if (!out || size_Ql == 0 || size_Ql > c->size_Q) return fail(B200FHE_ERR_ARG, "bad argument");
    if (c->logN < 2 || c->logN > 16) return fail(B200FHE_ERR_UNSUPPORTED, "encoder supports N = 2^2 .. 2^16");
    const size_t limbs = ext ? size_Ql + c->size_P : size_Ql;
    const PrimeMap pm = ext ? c->prime_map(size_Ql) : PrimeMap{0, 0};
    // total_coeff_modulus_bit_count of the level: significant bits of q_0 ... q_{l-1}
    const int total_bits = big_bits(big_product(std::vector<uint64_t>(c->primes.begin(), c->primes.begin() + size_Ql), size_Ql));
    if (scale <= 0 || static_cast<int>(std::log2(scale)) + 1 >= total_bits) return fail(B200FHE_ERR_ARG, "scale out of bounds");
    if (sparse && (n & (n - 1))) return fail(B200FHE_ERR_ARG, "sparse encoding needs a power-of-two number of values");
    const uint32_t fft_n = sparse ? (uint32_t) n : (uint32_t) slots;
    try {
        const FftTables &T = tables(fft_n, sparse && ext);
        CU(buf_in_.alloc(slots * sizeof(cuDoubleComplex)));
        CU(buf_tmp_.alloc(slots * sizeof(cuDoubleComplex) + 8));
        cuDoubleComplex *in = buf_in_.as<cuDoubleComplex>(), *tmp = buf_tmp_.as<cuDoubleComplex>();
        unsigned long long *d_max = reinterpret_cast<unsigned long long *>(tmp + slots);
        CU(cudaMemcpyAsync(tmp, values, n * sizeof(cuDoubleComplex), cudaMemcpyHostToDevice, st));
        CU(cudaMemsetAsync(in, 0, fft_n * sizeof(cuDoubleComplex), st));
        CU(cudaMemsetAsync(d_max, 0, 8, st));
        bit_reverse_scatter<<<(unsigned) ((n + kET - 1) / kET), kET, 0, st>>>(in, tmp, (uint32_t) n, T.logn);
        const double fix = scale / static_cast<double>(fft_n);
        const unsigned gb = std::max<unsigned>(1, (fft_n / 2 + kET - 1) / kET);
        if (T.logn == 0) return fail(B200FHE_ERR_UNSUPPORTED, "a single slot cannot be encoded");
        for (uint32_t lp = 0; lp < T.logn; lp++)  // inverse transform: adjacent pairs first
            launch_fft_stage<true>(gb, st, in, T, lp, lp + 1 == T.logn ? fix : 0.0);
        max_abs_kernel<<<std::min<unsigned>(gb, 148), kET, 0, st>>>(in, fft_n, d_max);
        unsigned long long hmax = 0;
        CU(cudaMemcpyAsync(&hmax, d_max, 8, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));  // the reference synchronises here too (ckks.cu:149)
        double max_coeff;
        std::memcpy(&max_coeff, &hmax, 8);
        const int bits = static_cast<int>(std::ceil(std::log2(std::max(max_coeff, 1.0)))) + 1;
        if (bits >= total_bits) return fail(B200FHE_ERR_ARG, "encoded values are too large");
        const int words = bits <= 64 ? 1 : (bits <= 128 ? 2 : (int) limbs);
        if (words > 24) return fail(B200FHE_ERR_UNSUPPORTED, "coefficients wider than 1536 bits");
        const uint32_t degree = 2 * fft_n, step = (uint32_t) (slots / fft_n);
        if (step > 1) CU(cudaMemsetAsync(out, 0, limbs * c->N * 8, st));
        decompose_kernel<<<dim3((degree + kET - 1) / kET, (unsigned) limbs), kET, 0, st>>>(c->T.mods, pm, out, in, degree, fft_n, c->N,
                                                                                    step, words);
        LimbMap m = c->limb_map(0, limbs, ext ? size_Ql : 0, 1, 0);
        CU(ntt_forward(c->T, out, out, m, nullptr, st));
        count_launches(T.logn + 3 + ntt_launch_count_forward(c->logN));
    } catch (const std::exception &e) {
        return fail(B200FHE_ERR_ALLOC, "%s", e.what());
    }
    return B200FHE_OK;
}

// ------------------------------------------------------------------------------------------- decode
struct ComposeTables {
    const u64 *inv_punct, *inv_punct_shoup;  // [L]   (Q/q_i)^-1 mod q_i
    const u64 *punct;                         // [L][L] Q/q_i as L words
    const u64 *big_q, *half;                  // [L]   Q and (Q+1)/2
};

// DRNSBase::compose_array (src/rns_base.cu:173-238): CRT-compose every coefficient into a multi-precision integer
// modulo Q, centre it, and turn it into a double word by word with the scale folded in
__global__ void compose_kernel(const Mod *__restrict__ mods, PrimeMap pm, ComposeTables T, cuDoubleComplex *dst,
                               const u64 *src, uint32_t L, uint32_t N, double inv_scale) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= N) return;
    u64 acc[kMaxWords];
    for (uint32_t w = 0; w < L; w++) acc[w] = 0;
    if (L > 1) {
        for (uint32_t i = 0; i < L; i++) {
            const Mod md = mods[pm.prime(i)];
            const u64 prod = shoup_mul(src[size_t(i) * N + tid], T.inv_punct[i], T.inv_punct_shoup[i], md.q, md.nq);
            // acc += punct_i * prod  (fits L words: both terms are below Q), then one conditional subtraction of Q
            const u64 *p = T.punct + size_t(i) * L;
            u64 carry = 0, c2 = 0;
            for (uint32_t w = 0; w < L; w++) {
                const u128 m = mul128(p[w], prod);
                u64 lo = m.lo + carry;
                u64 hi = m.hi + (lo < carry);
                u64 s = acc[w] + lo;
                u64 cs = s < lo;
                s += c2;
                cs += s < c2;
                acc[w] = s;
                carry = hi;
                c2 = cs;
            }
            bool ge = (carry | c2) != 0;  // overflow past L words cannot happen (sum < 2Q < 2^(64L))
            if (!ge) {
                ge = true;
                for (uint32_t w = L; w--;) {
                    if (acc[w] != T.big_q[w]) {
                        ge = acc[w] > T.big_q[w];
                        break;
                    }
                }
            }
            if (ge) {
                u64 borrow = 0;
                for (uint32_t w = 0; w < L; w++) {
                    const u64 b = T.big_q[w];
                    const u64 d = acc[w] - b - borrow;
                    borrow = (acc[w] < b) || (acc[w] == b && borrow);
                    acc[w] = d;
                }
            }
        }
    } else {
        acc[0] = src[tid];
    }
    bool upper = true;
    for (uint32_t w = L; w--;) {
        if (acc[w] != T.half[w]) {
            upper = acc[w] > T.half[w];
            break;
        }
    }
    const double two64 = 18446744073709551616.0;
    double res = 0.0, sc = inv_scale;
    if (upper) {
        for (uint32_t w = 0; w < L; w++, sc *= two64) {
            if (acc[w] > T.big_q[w]) {
                const u64 diff = acc[w] - T.big_q[w];
                res += diff ? static_cast<double>(diff) * sc : 0.0;
            } else {
                const u64 diff = T.big_q[w] - acc[w];
                res -= diff ? static_cast<double>(diff) * sc : 0.0;
            }
        }
    } else {
        for (uint32_t w = 0; w < L; w++, sc *= two64) {
            const u64 diff = acc[w];
            res += diff ? static_cast<double>(diff) * sc : 0.0;
        }
    }
    if (tid < N / 2) dst[tid].x = res;
    else dst[tid - N / 2].y = res;
}

__global__ void bit_reverse_gather(cuDoubleComplex *dst, const cuDoubleComplex *src, uint32_t count, uint32_t logn) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) dst[__brev(i) >> (32 - logn)] = src[i];
}

int Encoder::decode(const u64 *plain, size_t size_Ql, bool ext, double scale, cuDoubleComplex *out, cudaStream_t st) {
    if (c->host_only) return fail(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    if (!plain || !out || size_Ql == 0 || size_Ql > c->size_Q) return fail(B200FHE_ERR_ARG, "bad argument");
    if (c->logN < 2 || c->logN > 16) return fail(B200FHE_ERR_UNSUPPORTED, "encoder supports N = 2^2 .. 2^16");
    const size_t L = ext ? size_Ql + c->size_P : size_Ql, N = c->N;
    if (L > (size_t) kMaxWords) return fail(B200FHE_ERR_UNSUPPORTED, "more than %d limbs", kMaxWords);
    const PrimeMap pm = ext ? c->prime_map(size_Ql) : PrimeMap{0, 0};
    std::vector<uint64_t> pr(L);
    for (size_t i = 0; i < L; i++) pr[i] = c->primes[ext ? pm.prime((u32) i) : i];
    const Big Q = big_product(pr, L);
    if (scale <= 0 || static_cast<int>(std::log2(scale)) >= big_bits(Q)) return fail(B200FHE_ERR_ARG, "scale out of bounds");
    try {
        // constants of the composition (host, once per call: L^2 words)
        std::vector<uint64_t> h(2 * L + L * L + 2 * L);
        uint64_t *inv = h.data(), *invs = inv + L, *punct = invs + L, *bq = punct + L * L, *half = bq + L;
        for (size_t i = 0; i < L; i++) {
            const Big P = big_product(pr, i);
            std::copy(P.begin(), P.end(), punct + i * L);
            inv[i] = L > 1 ? host::invmod(big_mod_word(P, pr[i]), pr[i]) : 1;
            invs[i] = host::shoup(inv[i], pr[i]);
            bq[i] = Q[i];
        }
        {  // (Q + 1) >> 1
            Big t = Q;
            for (size_t i = 0; i < L; i++)
                if (++t[i]) break;
            for (size_t i = 0; i < L; i++) half[i] = (t[i] >> 1) | (i + 1 < L ? t[i + 1] << 63 : 0);
        }
        CU(buf_res_.alloc(h.size() * 8 + L * N * 8));
        u64 *d_tab = buf_res_.as<u64>(), *coef = d_tab + h.size();
        CU(cudaMemcpyAsync(d_tab, h.data(), h.size() * 8, cudaMemcpyHostToDevice, st));
        CU(buf_in_.alloc(slots * sizeof(cuDoubleComplex)));
        CU(buf_tmp_.alloc(slots * sizeof(cuDoubleComplex) + 8));
        cuDoubleComplex *in = buf_in_.as<cuDoubleComplex>(), *tmp = buf_tmp_.as<cuDoubleComplex>();
        LimbMap m = c->limb_map(0, L, ext ? size_Ql : 0, 1, 0);
        CU(ntt_inverse(c->T, coef, plain, m, nullptr, nullptr, st));
        ComposeTables T{d_tab, d_tab + L, d_tab + 2 * L, d_tab + 2 * L + L * L, d_tab + 3 * L + L * L};
        compose_kernel<<<(unsigned) ((N + kET - 1) / kET), kET, 0, st>>>(c->T.mods, pm, T, in, coef, (uint32_t) L, (uint32_t) N,
                                                                      1.0 / scale);
        const FftTables &F = tables((uint32_t) slots, false);
        const unsigned gb = std::max<unsigned>(1, (unsigned) ((slots / 2 + kET - 1) / kET));
        for (uint32_t it = 0; it < F.logn; it++)  // forward transform: widest butterflies first
            launch_fft_stage<false>(gb, st, in, F, F.logn - it - 1, 0.0);
        bit_reverse_gather<<<(unsigned) ((slots + kET - 1) / kET), kET, 0, st>>>(tmp, in, (uint32_t) slots, F.logn);
        CU(cudaMemcpyAsync(out, tmp, slots * sizeof(cuDoubleComplex), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        count_launches(F.logn + 2 + ntt_launch_count_inverse(c->logN));
    } catch (const std::exception &e) {
        return fail(B200FHE_ERR_ALLOC, "%s", e.what());
    }
    return B200FHE_OK;
}

}  // namespace b200fhe
