// ref_shim.cu — OUR flat C interface over the UNMODIFIED reference GPU library
// (/root/reference/src/**/*.cu compiled for sm_100a by oracle/Makefile.ref into
// oracle/_ref/libphantom_ref.so).  TEST / BASELINE INFRASTRUCTURE ONLY:
//   * tests/test_ref_gpu_parity.py feeds identical buffers, keys and ciphertexts to the reference's
//     own kernels / evaluator and to libb200fhe.so on the B200 and memcmp's the outputs;
//   * bench.py times the reference's GPU build next to ours ("reference_gpu" in the JSON line).
// Nothing under phantom-fhe-boot_b200/ links or loads this.
#include <cstdint>
#include <cstring>
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include <complex>
#include <iostream>
#include <map>
#include <sstream>

// the precomputed plaintext diagonals live in a private member of FHECKKSRNS (m_bootPrecomMap): the checker reads
// them through the class definition with access control switched off.  The reference sources stay unmodified.
#define private public
#include "phantom.h"
#include "bootstrap.cuh"
#include "dnn.cuh"
#undef private

using namespace phantom;
using namespace phantom::arith;
using namespace phantom::util;

namespace {
struct Ref {
    EncryptionParameters parms{scheme_type::ckks};
    std::unique_ptr<PhantomContext> ctx;
    std::unique_ptr<PhantomSecretKey> sk;
    std::unique_ptr<PhantomCKKSEncoder> enc;
    size_t N = 0, size_QP = 0, size_P = 0;
};
thread_local std::string g_err;
template <class F>
int guard(F &&f) {
    try {
        f();
        return 0;
    } catch (const std::exception &e) {
        g_err = e.what();
        return -1;
    }
}
// cudaStreamPerThread is a macro for a constant handle: keep one object to hand out references to
const cudaStream_t g_stream = cudaStreamPerThread;
const cudaStream_t &S() { return g_stream; }
}  // namespace

// ---- std::random_device made seedable (TEST INFRASTRUCTURE): the reference draws every 64-byte PRNG seed from
// std::random_device (include/prng.cuh:12-32), one byte per 32-bit draw (uniform_int_distribution<unsigned short>(0, 255)
// over a 2^32-range source keeps the top byte).  libphantom_ref.so is linked with -Bsymbolic-functions, so the reference's
// calls of random_device::_M_getval() bind to this definition: after rref_set_random_seed(master) draw i returns
// top8(splitmix64(master + i)) << 24, the byte stream b200fhe_keygen_create_reference_prng consumes; before it (or after
// rref_set_random_seed_off) the bytes come from /dev/urandom as usual.
#include <atomic>
#include <cstdio>
#include <random>
static std::atomic<uint64_t> g_rd_master{0}, g_rd_drawn{0};
static std::atomic<bool> g_rd_seeded{false}, g_rd_pending{false};
static uint64_t shim_splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
std::random_device::result_type std::random_device::_M_getval() {
    if (g_rd_seeded.load()) return (result_type) ((shim_splitmix64(g_rd_master.load() + g_rd_drawn.fetch_add(1)) >> 56) << 24);
    static FILE *f = fopen("/dev/urandom", "rb");
    result_type v = 0;
    if (!f || fread(&v, sizeof v, 1, f) != 1) v = (result_type) rand();
    return v;
}

extern "C" {
void rref_set_random_seed(uint64_t master) {
    g_rd_master = master;
    g_rd_drawn = 0;
    g_rd_seeded = true;
}
// seed the source from the NEXT rref_create's secret key on (context construction before it stays unseeded)
void rref_seed_next_secret_key(uint64_t master) {
    g_rd_master = master;
    g_rd_pending = true;
}
void rref_set_random_seed_off(void) {
    g_rd_seeded = false;
    g_rd_pending = false;
}
uint64_t rref_random_bytes_drawn(void) { return g_rd_drawn.load(); }

const char *rref_last_error() { return g_err.c_str(); }

// EncryptionParameters(ckks) + CoeffModulus::Create + PhantomContext + PhantomSecretKey (+ encoder)
void *rref_create(int logN, const int *bits, size_t nbits, size_t special_size) {
    Ref *r = new Ref();
    int rc = guard([&] {
        r->N = size_t(1) << logN;
        r->parms.set_poly_modulus_degree(r->N);
        r->parms.set_special_modulus_size(special_size);
        r->parms.set_coeff_modulus(CoeffModulus::Create(r->N, std::vector<int>(bits, bits + nbits)));
        r->ctx = std::make_unique<PhantomContext>(r->parms);
        // the host number theory of the context also draws from std::random_device (primitive-root search): a pending
        // seed (rref_seed_next_secret_key) starts counting at the secret key, where key generation begins
        if (g_rd_pending.exchange(false)) {
            g_rd_drawn = 0;
            g_rd_seeded = true;
        }
        r->sk = std::make_unique<PhantomSecretKey>(*r->ctx);
        r->enc = std::make_unique<PhantomCKKSEncoder>(*r->ctx);
        r->size_QP = nbits;
        r->size_P = special_size;
    });
    if (rc) {
        delete r;
        return nullptr;
    }
    return r;
}
void rref_destroy(void *h) { delete static_cast<Ref *>(h); }
void rref_sync() { cudaStreamSynchronize(S()); cudaDeviceSynchronize(); }

int rref_primes(void *h, uint64_t *out) {
    auto *r = static_cast<Ref *>(h);
    auto &m = r->ctx->get_context_data(0).parms().coeff_modulus();
    for (size_t i = 0; i < m.size(); i++) out[i] = m[i].value();
    return (int) m.size();
}

// ---- L1: NTT launchers (include/ntt.cuh:157-226) on raw device pointers
int rref_ntt_forward(void *h, uint64_t *d, size_t limbs, size_t start) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] { nwt_2d_radix8_forward_inplace(d, r->ctx->gpu_rns_tables(), limbs, start, S()); });
}
int rref_ntt_inverse(void *h, uint64_t *d, size_t limbs, size_t start) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] { nwt_2d_radix8_backward_inplace(d, r->ctx->gpu_rns_tables(), limbs, start, S()); });
}
int rref_ntt_forward_special(void *h, uint64_t *d, size_t limbs, size_t start) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        nwt_2d_radix8_forward_inplace_include_special_mod(d, r->ctx->gpu_rns_tables(), limbs, start, r->size_QP,
                                                          r->size_P, S());
    });
}
int rref_ntt_inverse_special(void *h, uint64_t *d, size_t limbs, size_t start) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        nwt_2d_radix8_backward_inplace_include_special_mod(d, r->ctx->gpu_rns_tables(), limbs, start, r->size_QP,
                                                           r->size_P, S());
    });
}

// ---- L2: DRNSTool of a level (levels_dropped = |Q| - |Ql|)
static const DRNSTool &tool(Ref *r, size_t levels_dropped) {
    return r->ctx->get_context_data(1 + levels_dropped).gpu_rns_tool();
}
int rref_modup(void *h, size_t levels_dropped, uint64_t *dst, const uint64_t *c2) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] { tool(r, levels_dropped).modup(dst, c2, r->ctx->gpu_rns_tables(), scheme_type::ckks, S()); });
}
int rref_moddown_from_ntt(void *h, size_t levels_dropped, uint64_t *ct, uint64_t *cx) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        tool(r, levels_dropped).moddown_from_NTT(ct, cx, r->ctx->gpu_rns_tables(), scheme_type::ckks, S());
    });
}
int rref_inner_prod(void *h, size_t levels_dropped, uint64_t *cx, const uint64_t *up, void *key) {
    auto *r = static_cast<Ref *>(h);
    auto *k = static_cast<const PhantomRelinKey *>(key);
    return guard([&] {
        key_switch_inner_prod(cx, up, k->public_keys_ptr(), tool(r, levels_dropped),
                              r->ctx->gpu_rns_tables().modulus(), 7, S());
    });
}

// ---- keys
void *rref_gen_relinkey(void *h) {
    auto *r = static_cast<Ref *>(h);
    PhantomRelinKey *k = nullptr;
    if (guard([&] { k = new PhantomRelinKey(r->sk->gen_relinkey(*r->ctx)); })) return nullptr;
    return k;
}
void rref_free_relinkey(void *k) { delete static_cast<PhantomRelinKey *>(k); }
// device pointers of the digits (each [2][size_QP][N]); returns dnum
int rref_relinkey_digits(void *h, void *key, uint64_t **ptrs_out, size_t max) {
    auto *r = static_cast<Ref *>(h);
    auto *k = static_cast<const PhantomRelinKey *>(key);
    const size_t size_Q = r->size_QP - r->size_P;
    const size_t dnum = size_Q / r->size_P;  // src/secretkey.cu:420
    if (dnum > max) return -1;
    cudaStreamSynchronize(S());
    cudaMemcpy(ptrs_out, k->public_keys_ptr(), dnum * sizeof(uint64_t *), cudaMemcpyDeviceToHost);
    return (int) dnum;
}
void *rref_gen_rotkeys(void *h, const int32_t *idx, size_t n) {
    auto *r = static_cast<Ref *>(h);
    PhantomGaloisKeyFused *k = nullptr;
    if (guard([&] { k = new PhantomGaloisKeyFused(r->sk->EvalAtIndexKeyGen(*r->ctx, std::vector<int32_t>(idx, idx + n))); }))
        return nullptr;
    return k;
}
void rref_free_rotkeys(void *k) { delete static_cast<PhantomGaloisKeyFused *>(k); }
uint32_t rref_auto_index(void *h, int32_t index) {
    auto *r = static_cast<Ref *>(h);
    return FindAutomorphismIndex2nComplex(index, 2 * r->N);
}
void *rref_rotkey_get(void *keys, uint32_t auto_index) {
    auto *k = static_cast<PhantomGaloisKeyFused *>(keys);
    const PhantomRelinKey *p = nullptr;
    if (guard([&] { p = &k->getKey(auto_index); })) return nullptr;
    return const_cast<PhantomRelinKey *>(p);
}

// ---- ciphertexts: encode + encrypt a real vector, or wrap raw limbs
void *rref_encrypt(void *h, const double *vals, size_t n, double scale, size_t chain_index) {
    auto *r = static_cast<Ref *>(h);
    PhantomCiphertext *ct = nullptr;
    if (guard([&] {
            std::vector<double> v(vals, vals + n);
            PhantomPlaintext pt;
            r->enc->encode(*r->ctx, v, scale, pt, chain_index);
            ct = new PhantomCiphertext();
            r->sk->encrypt_symmetric(*r->ctx, pt, *ct);
        }))
        return nullptr;
    return ct;
}
void *rref_ct_from_raw(void *h, size_t chain_index, size_t size, const uint64_t *dev_data, double scale) {
    auto *r = static_cast<Ref *>(h);
    PhantomCiphertext *ct = nullptr;
    if (guard([&] {
            ct = new PhantomCiphertext();
            ct->resize(*r->ctx, chain_index, size, S());
            ct->set_scale(scale);
            ct->set_ntt_form(true);
            ct->SetNoiseScaleDeg(1);
            cudaMemcpyAsync(ct->data(), dev_data, size * ct->coeff_modulus_size() * r->N * 8, cudaMemcpyDeviceToDevice, S());
            cudaStreamSynchronize(S());
        }))
        return nullptr;
    return ct;
}
void *rref_ct_clone(void *c) { return new PhantomCiphertext(*static_cast<PhantomCiphertext *>(c)); }
void rref_ct_free(void *c) { delete static_cast<PhantomCiphertext *>(c); }
uint64_t *rref_ct_data(void *c) { return static_cast<PhantomCiphertext *>(c)->data(); }
size_t rref_ct_size(void *c) { return static_cast<PhantomCiphertext *>(c)->size(); }
size_t rref_ct_limbs(void *c) { return static_cast<PhantomCiphertext *>(c)->coeff_modulus_size(); }
size_t rref_ct_chain_index(void *c) { return static_cast<PhantomCiphertext *>(c)->chain_index(); }
double rref_ct_scale(void *c) { return static_cast<PhantomCiphertext *>(c)->scale(); }
int rref_encode_decode(void *h, const double *vals, size_t n, double scale, size_t chain_index, double *out) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        std::vector<double> v(vals, vals + n), w;
        PhantomPlaintext pt;
        r->enc->encode(*r->ctx, v, scale, pt, chain_index);
        r->enc->decode(*r->ctx, pt, w);
        for (size_t i = 0; i < n && i < w.size(); i++) out[i] = w[i];
    });
}
// PhantomCKKSEncoder::encode / encode_ext / encode_sparse / encode_sparse_ext of complex values: the plaintext limbs
// ([limbs][N], NTT form) are copied to out_dev; returns the limb count.  mode: 0 plain, 1 ext, 2 sparse, 3 sparse ext
int rref_encode_raw(void *h, const double *re_im, size_t n, double scale, size_t chain_index, int mode, uint64_t *out_dev) {
    auto *r = static_cast<Ref *>(h);
    int limbs = -1;
    if (guard([&] {
            std::vector<cuDoubleComplex> v(n);
            for (size_t i = 0; i < n; i++) v[i] = make_cuDoubleComplex(re_im[2 * i], re_im[2 * i + 1]);
            PhantomPlaintext pt;
            // encode_sparse works on the ring of degree 2n; encode_sparse_ext on the "bootstrap" tables, which
            // set_sparse_encode builds for TWICE the degree it is given
            if (mode == 2) r->enc->set_sparse_encode(r->parms, 2 * n);
            if (mode == 3) r->enc->set_sparse_encode(r->parms, n);
            if (mode == 0) r->enc->encode(*r->ctx, v, scale, pt, chain_index);
            else if (mode == 1) r->enc->encode_ext(*r->ctx, v, scale, pt, chain_index);
            else if (mode == 2) r->enc->encode_sparse(*r->ctx, v, scale, pt, chain_index);
            else r->enc->encode_sparse_ext(*r->ctx, v, scale, pt, chain_index);
            cudaStreamSynchronize(S());
            const size_t Ql = r->size_QP - r->size_P - chain_index + 1;
            limbs = (int) ((mode & 1) ? Ql + r->size_P : Ql);
            cudaMemcpyAsync(out_dev, pt.data(), size_t(limbs) * r->N * 8, cudaMemcpyDeviceToDevice, S());
            cudaStreamSynchronize(S());
        }))
        return -1;
    return limbs;
}
// decode of raw plaintext limbs at chain_index with the given scale: `slots` complex numbers to out_re_im
int rref_decode_raw(void *h, const uint64_t *dev, size_t chain_index, double scale, double *out_re_im) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        const size_t Ql = r->size_QP - r->size_P - chain_index + 1;
        PhantomPlaintext pt;
        pt.resize(Ql, r->N, S());
        pt.chain_index_ = chain_index;
        pt.scale_ = scale;
        cudaMemcpyAsync(pt.data(), dev, Ql * r->N * 8, cudaMemcpyDeviceToDevice, S());
        std::vector<cuDoubleComplex> w;
        r->enc->decode(*r->ctx, pt, w);
        for (size_t i = 0; i < w.size(); i++) {
            out_re_im[2 * i] = w[i].x;
            out_re_im[2 * i + 1] = w[i].y;
        }
    });
}
// decrypt only: the plaintext limbs [limbs][N] (NTT form) are copied to out_dev; returns the limb count
int rref_decrypt_raw(void *h, void *c, uint64_t *out_dev) {
    auto *r = static_cast<Ref *>(h);
    int limbs = -1;
    if (guard([&] {
            auto *ct = static_cast<PhantomCiphertext *>(c);
            PhantomPlaintext pt = r->sk->decrypt(*r->ctx, *ct);
            limbs = (int) ct->coeff_modulus_size();
            cudaMemcpyAsync(out_dev, pt.data(), size_t(limbs) * r->N * 8, cudaMemcpyDeviceToDevice, S());
            cudaStreamSynchronize(S());
        }))
        return -1;
    return limbs;
}
int rref_decrypt_decode(void *h, void *c, double *out, size_t n) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        PhantomPlaintext pt = r->sk->decrypt(*r->ctx, *static_cast<PhantomCiphertext *>(c));
        std::vector<double> v;
        r->enc->decode(*r->ctx, pt, v);
        for (size_t i = 0; i < n && i < v.size(); i++) out[i] = v[i];
    });
}
// ---- L3 evaluator
int rref_multiply_and_relin(void *h, void *a, void *b, void *rlk) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        multiply_and_relin_inplace(*r->ctx, *static_cast<PhantomCiphertext *>(a), *static_cast<PhantomCiphertext *>(b),
                                   *static_cast<PhantomRelinKey *>(rlk));
    });
}
int rref_multiply(void *h, void *a, void *b) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] { multiply_inplace(*r->ctx, *static_cast<PhantomCiphertext *>(a), *static_cast<PhantomCiphertext *>(b)); });
}
int rref_relinearize(void *h, void *a, void *rlk) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] { relinearize_inplace(*r->ctx, *static_cast<PhantomCiphertext *>(a), *static_cast<PhantomRelinKey *>(rlk)); });
}
int rref_rescale(void *h, void *a) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] { rescale_to_next_inplace(*r->ctx, *static_cast<PhantomCiphertext *>(a)); });
}
int rref_rotate_fused(void *h, void *keys, void *in, void *out, int32_t index) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        EvalRotateFused(*r->ctx, *static_cast<PhantomGaloisKeyFused *>(keys), *static_cast<PhantomCiphertext *>(in),
                        *static_cast<PhantomCiphertext *>(out), index);
    });
}
int rref_keyswitch(void *h, void *ct, uint64_t *c2, void *key) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        keyswitch_inplace(*r->ctx, *static_cast<PhantomCiphertext *>(ct), c2, *static_cast<PhantomRelinKey *>(key), false, S());
    });
}

// ---- timing of the reference's own code path (CUDA events on its stream)
static float timed(int iters, const std::function<void()> &f) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaStreamSynchronize(S());
    cudaEventRecord(e0, S());
    for (int i = 0; i < iters; i++) f();
    cudaEventRecord(e1, S());
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return ms / iters;
}
// forward then inverse NTT over `limbs` limbs of each of `npoly` polys at d (+ p*poly_stride); returns ms per pass
int rref_time_ntt(void *h, uint64_t *d, size_t limbs, size_t npoly, size_t poly_stride, int iters, float *ms_fwd,
                  float *ms_inv) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        auto &T = r->ctx->gpu_rns_tables();
        *ms_fwd = timed(iters, [&] { for (size_t p = 0; p < npoly; p++) nwt_2d_radix8_forward_inplace(d + p * poly_stride, T, limbs, 0, S()); });
        *ms_inv = timed(iters, [&] { for (size_t p = 0; p < npoly; p++) nwt_2d_radix8_backward_inplace(d + p * poly_stride, T, limbs, 0, S()); });
    });
}
// The way the reference's own harness batches (benchmark/ntt_bench.cu:84-119): ONE launch over a DNTTTable with one row per
// limb of the whole batch (row r carries the tables of prime r % limbs), data [npoly][limbs][N] contiguous
int rref_time_ntt_wide(void *h, uint64_t *d, size_t limbs, size_t npoly, int iters, float *ms_fwd, float *ms_inv) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        auto &T = r->ctx->gpu_rns_tables();
        const size_t n = T.n(), rows = limbs * npoly;
        DNTTTable W;
        W.init(n, rows, S());
        for (size_t i = 0; i < rows; i++) {
            const size_t p = i % limbs;
            cudaMemcpyAsync(W.modulus() + i, T.modulus() + p, sizeof(DModulus), cudaMemcpyDeviceToDevice, S());
            cudaMemcpyAsync(W.twiddle() + i * n, T.twiddle() + p * n, n * 8, cudaMemcpyDeviceToDevice, S());
            cudaMemcpyAsync(W.twiddle_shoup() + i * n, T.twiddle_shoup() + p * n, n * 8, cudaMemcpyDeviceToDevice, S());
            cudaMemcpyAsync(W.itwiddle() + i * n, T.itwiddle() + p * n, n * 8, cudaMemcpyDeviceToDevice, S());
            cudaMemcpyAsync(W.itwiddle_shoup() + i * n, T.itwiddle_shoup() + p * n, n * 8, cudaMemcpyDeviceToDevice, S());
            cudaMemcpyAsync(W.n_inv_mod_q() + i, T.n_inv_mod_q() + p, 8, cudaMemcpyDeviceToDevice, S());
            cudaMemcpyAsync(W.n_inv_mod_q_shoup() + i, T.n_inv_mod_q_shoup() + p, 8, cudaMemcpyDeviceToDevice, S());
        }
        cudaStreamSynchronize(S());
        *ms_fwd = timed(iters, [&] { nwt_2d_radix8_forward_inplace(d, W, rows, 0, S()); });
        *ms_inv = timed(iters, [&] { nwt_2d_radix8_backward_inplace(d, W, rows, 0, S()); });
        cudaStreamSynchronize(S());
    });
}
// multiply_and_relin_inplace on fresh copies of (a, b): the copy cost is excluded by pre-cloning `iters` inputs
int rref_time_hmult(void *h, void *a, void *b, void *rlk, int iters, float *ms) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        std::vector<PhantomCiphertext> as(iters, *static_cast<PhantomCiphertext *>(a));
        int i = 0;
        *ms = timed(iters, [&] {
            multiply_and_relin_inplace(*r->ctx, as[i++], *static_cast<PhantomCiphertext *>(b), *static_cast<PhantomRelinKey *>(rlk));
        });
    });
}
int rref_time_rotate(void *h, void *keys, void *in, int32_t index, int iters, float *ms) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        PhantomCiphertext out;
        *ms = timed(iters, [&] {
            EvalRotateFused(*r->ctx, *static_cast<PhantomGaloisKeyFused *>(keys), *static_cast<PhantomCiphertext *>(in), out, index);
        });
    });
}
int rref_time_rescale(void *h, void *a, int iters, float *ms) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        *ms = timed(iters, [&] { PhantomCiphertext o = rescale_to_next(*r->ctx, *static_cast<PhantomCiphertext *>(a)); });
    });
}

// ---- ciphertext metadata of the FLEXIBLEAUTO layer
uint32_t rref_ct_noise(void *c) { return (uint32_t) static_cast<PhantomCiphertext *>(c)->GetNoiseScaleDeg(); }
void rref_ct_set_noise(void *c, uint32_t n) { static_cast<PhantomCiphertext *>(c)->SetNoiseScaleDeg(n); }
void rref_ct_set_scale(void *c, double s) { static_cast<PhantomCiphertext *>(c)->set_scale(s); }

// ---- bootstrap: FHECKKSRNS with its own keys and precomputed diagonals (bootstrapping_example.cu:121-160)
struct RefBoot {
    Ref *r;
    std::unique_ptr<FHECKKSRNS> fhe;
    std::vector<double> sf, sf_big;
    uint32_t slots;
};
void *rref_boot_create(void *h, uint32_t budget_enc, uint32_t budget_dec, double scale, uint32_t num_slots, int keygen) {
    auto *r = static_cast<Ref *>(h);
    RefBoot *b = new RefBoot();
    b->r = r;
    if (guard([&] {
            PhantomCiphertext tmp;
            tmp.PreComputeScale(*r->ctx, scale);
            b->sf = tmp.getScalingFactorsReal();
            b->sf_big = tmp.getScalingFactorsRealBig();
            b->fhe = std::make_unique<FHECKKSRNS>(*r->enc);
            std::vector<uint32_t> budget = {budget_enc, budget_dec};
            b->slots = num_slots ? num_slots : (uint32_t) (r->N / 2);
            // sparse packing encodes through the sparse tables (bootstrapping_example.cu:270)
            if (b->slots != r->N / 2) r->enc->set_sparse_encode(r->parms, size_t(b->slots) * 2);
            b->fhe->EvalBootstrapSetup(*r->ctx, budget, scale, b->sf, b->sf_big, {0, 0}, num_slots);
            if (keygen) {
                b->fhe->EvalMultKeyGen(*r->sk, *r->ctx);
                b->fhe->EvalBootstrapKeyGen(*r->sk, *r->ctx, b->slots);
            }
            cudaStreamSynchronize(S());
        })) {
        delete b;
        return nullptr;
    }
    return b;
}
void rref_boot_destroy(void *b) { delete static_cast<RefBoot *>(b); }
int rref_boot_scaling_factors(void *bh, int which, double *out, size_t cap) {
    auto *b = static_cast<RefBoot *>(bh);
    const auto &v = which == 0 ? b->sf : b->sf_big;
    for (size_t i = 0; i < v.size() && i < cap; i++) out[i] = v[i];
    return (int) v.size();
}
uint32_t rref_boot_correction_factor(void *bh) { return static_cast<RefBoot *>(bh)->fhe->m_correctionFactor; }
int rref_boot_params(void *bh, int which, int32_t *out9) {
    auto *b = static_cast<RefBoot *>(bh);
    auto &pre = *b->fhe->m_bootPrecomMap.at(b->slots);
    const auto &v = which == 0 ? pre.m_paramsEnc : pre.m_paramsDec;
    for (size_t i = 0; i < 9 && i < v.size(); i++) out9[i] = v[i];
    return (int) v.size();
}
int rref_boot_rotation_indices(void *bh, int32_t *out, size_t cap) {
    auto *b = static_cast<RefBoot *>(bh);
    std::vector<int32_t> v;
    if (guard([&] { v = b->fhe->FindBootstrapRotationIndices(b->slots, 2 * (uint32_t) b->r->N); })) return -1;
    for (size_t i = 0; i < v.size() && i < cap; i++) out[i] = v[i];
    return (int) v.size();
}
// diagonal idx of stage `stage` of m_U0hatTPreFFT (which = 0) / m_U0PreFFT (1): device pointer, chain index, scale.
// returns the number of diagonals of the stage, or -1
int rref_boot_diag(void *bh, int which, size_t stage, size_t idx, uint64_t **ptr, size_t *chain, double *scale,
                   size_t *limbs) {
    auto *b = static_cast<RefBoot *>(bh);
    auto &pre = *b->fhe->m_bootPrecomMap.at(b->slots);
    auto &A = which == 0 ? pre.m_U0hatTPreFFT : pre.m_U0PreFFT;
    if (stage >= A.size()) return -1;
    if (idx < A[stage].size() && A[stage][idx]) {
        auto &pt = *A[stage][idx];
        if (ptr) *ptr = pt.data();
        if (chain) *chain = pt.chain_index();
        if (scale) *scale = pt.scale();
        if (limbs) *limbs = pt.coeff_modulus_size();
    }
    return (int) A[stage].size();
}
int rref_boot_nstages(void *bh, int which) {
    auto *b = static_cast<RefBoot *>(bh);
    auto &pre = *b->fhe->m_bootPrecomMap.at(b->slots);
    return (int) (which == 0 ? pre.m_U0hatTPreFFT.size() : pre.m_U0PreFFT.size());
}
void *rref_boot_mult_key(void *bh) { return &static_cast<RefBoot *>(bh)->fhe->mul_key; }
void *rref_boot_galois_keys(void *bh) { return &static_cast<RefBoot *>(bh)->fhe->galois_keys; }
int rref_rotkey_has(void *keys, uint32_t auto_index) {
    return static_cast<PhantomGaloisKeyFused *>(keys)->key_map.count((int32_t) auto_index) ? 1 : 0;
}
static PhantomCiphertext *newct(PhantomCiphertext &&c) { return new PhantomCiphertext(std::move(c)); }
#define RB(bh) auto *b = static_cast<RefBoot *>(bh); auto &ctx = *b->r->ctx; (void) ctx
void *rref_boot_eval(void *bh, void *ct) {
    RB(bh);
    PhantomCiphertext *o = nullptr;
    if (guard([&] { o = newct(b->fhe->EvalBootstrap(*static_cast<PhantomCiphertext *>(ct), ctx, b->slots == b->r->N / 2 ? 0 : b->slots)); })) return nullptr;
    return o;
}
void *rref_boot_eval_iter(void *bh, void *ct, uint32_t iters, uint32_t precision) {
    RB(bh);
    PhantomCiphertext *o = nullptr;
    if (guard([&] { o = newct(b->fhe->EvalBootstrap(*static_cast<PhantomCiphertext *>(ct), ctx, b->slots == b->r->N / 2 ? 0 : b->slots, iters, precision)); })) return nullptr;
    return o;
}
void *rref_boot_c2s(void *bh, void *ct) {
    RB(bh);
    PhantomCiphertext *o = nullptr;
    auto &pre = *b->fhe->m_bootPrecomMap.at(b->slots);
    if (guard([&] { o = newct(b->fhe->EvalCoeffsToSlots(pre.m_U0hatTPreFFT, *static_cast<PhantomCiphertext *>(ct), ctx, b->slots)); }))
        return nullptr;
    return o;
}
void *rref_boot_s2c(void *bh, void *ct) {
    RB(bh);
    PhantomCiphertext *o = nullptr;
    auto &pre = *b->fhe->m_bootPrecomMap.at(b->slots);
    if (guard([&] { o = newct(b->fhe->EvalSlotsToCoeffs(pre.m_U0PreFFT, *static_cast<PhantomCiphertext *>(ct), ctx, b->slots)); }))
        return nullptr;
    return o;
}
int rref_boot_mult_const(void *bh, void *ct, double v) {
    RB(bh);
    return guard([&] { EvalMultConstInplace(ctx, *static_cast<PhantomCiphertext *>(ct), v, b->sf); });
}
int rref_boot_add_const(void *bh, void *ct, double v) {
    RB(bh);
    return guard([&] { EvalAddConstInPlaceWrap(ctx, *static_cast<PhantomCiphertext *>(ct), v, b->sf, b->sf_big); });
}
int rref_boot_mod_reduce(void *bh, void *ct, size_t levels) {
    RB(bh);
    return guard([&] { EvalModReduceInPlace(ctx, *static_cast<PhantomCiphertext *>(ct), levels); });
}
int rref_boot_mod_switch_level(void *bh, void *ct, size_t levels) {
    RB(bh);
    return guard([&] { ModSwitchLevelInPlace(ctx, *static_cast<PhantomCiphertext *>(ct), levels); });
}
int rref_boot_add_auto(void *bh, void *a, void *c2, int sub) {
    RB(bh);
    return guard([&] {
        if (sub) EvalSubAutoInplace(ctx, *static_cast<PhantomCiphertext *>(a), *static_cast<PhantomCiphertext *>(c2), b->sf, b->sf_big);
        else EvalAddAutoInplace(ctx, *static_cast<PhantomCiphertext *>(a), *static_cast<PhantomCiphertext *>(c2), b->sf, b->sf_big);
    });
}
void *rref_boot_mult_auto(void *bh, void *a, void *c2) {
    RB(bh);
    PhantomCiphertext *o = nullptr;
    if (guard([&] {
            o = newct(EvalMultAuto(ctx, *static_cast<PhantomCiphertext *>(a), *static_cast<PhantomCiphertext *>(c2),
                                   b->fhe->mul_key, b->sf, b->sf_big));
        }))
        return nullptr;
    return o;
}
void *rref_boot_square(void *bh, void *a) {
    RB(bh);
    PhantomCiphertext *o = nullptr;
    if (guard([&] { o = newct(EvalSquare(ctx, *static_cast<PhantomCiphertext *>(a), b->fhe->mul_key, b->sf, b->sf_big)); })) return nullptr;
    return o;
}
void *rref_boot_chebyshev(void *bh, void *ct, const double *coeffs, size_t n, double lo, double hi) {
    RB(bh);
    PhantomCiphertext *o = nullptr;
    if (guard([&] {
            o = newct(EvalChebyshevSeries(ctx, b->fhe->mul_key, *static_cast<PhantomCiphertext *>(ct),
                                          std::vector<double>(coeffs, coeffs + n), lo, hi, b->sf, b->sf_big));
        }))
        return nullptr;
    return o;
}
void *rref_boot_raise_mod(void *bh, void *ct) {
    RB(bh);
    PhantomCiphertext *o = nullptr;
    if (guard([&] {
            PhantomCiphertext tmp = *static_cast<PhantomCiphertext *>(ct);  // RaiseMod leaves its input in coefficient form
            o = newct(RaiseMod(ctx, tmp));
        }))
        return nullptr;
    return o;
}
int rref_boot_mult_monomial(void *bh, void *ct, uint32_t power) {
    RB(bh);
    return guard([&] { MultByMonomialInPlace(ctx, *static_cast<PhantomCiphertext *>(ct), power); });
}
int rref_boot_mult_integer(void *bh, void *ct, uint64_t v) {
    RB(bh);
    return guard([&] { MultByIntegerInPlace(ctx, *static_cast<PhantomCiphertext *>(ct), v); });
}
void *rref_boot_conj(void *bh, void *ct) {
    RB(bh);
    PhantomCiphertext *o = new PhantomCiphertext();
    if (guard([&] { EvalConjFused(ctx, b->fhe->galois_keys, *static_cast<PhantomCiphertext *>(ct), *o); })) {
        delete o;
        return nullptr;
    }
    return o;
}
// EvalFastRotationPrecompute + EvalFastRotationExt (index != 0) or KeySwitchExt (index == 0)
void *rref_boot_fast_rotation_ext(void *bh, void *ct, int32_t index, int add_first) {
    RB(bh);
    PhantomCiphertext *o = nullptr;
    if (guard([&] {
            auto &x = *static_cast<PhantomCiphertext *>(ct);
            if (index == 0) {
                o = newct(KeySwitchExt(ctx, x));
            } else {
                auto digits = EvalFastRotationPrecompute(ctx, x);
                o = newct(EvalFastRotationExt(ctx, x, b->fhe->galois_keys, index, std::move(digits), add_first != 0));
            }
        }))
        return nullptr;
    return o;
}
void *rref_boot_keyswitch_down(void *bh, void *ct, int first_only) {
    RB(bh);
    PhantomCiphertext *o = nullptr;
    if (guard([&] {
            auto &x = *static_cast<PhantomCiphertext *>(ct);
            o = newct(first_only ? KeySwitchDownFirstElement(ctx, x) : KeySwitchDown(ctx, x));
        }))
        return nullptr;
    return o;
}
int rref_time_bootstrap(void *bh, void *ct, int iters, float *ms) {
    RB(bh);
    return guard([&] {
        *ms = timed(iters, [&] { PhantomCiphertext o = b->fhe->EvalBootstrap(*static_cast<PhantomCiphertext *>(ct), ctx); });
    });
}

// ---- key material exchange (tests/test_keygen_gpu.py): the reference's secret key out, OUR evaluation keys in.
// The reference's key classes only hold device pointers to the digits; a key built here points at caller-owned memory.
int rref_secret_key(void *h, uint64_t *out_dev) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        cudaStreamSynchronize(S());
        cudaMemcpy(out_dev, r->sk->secret_key_array(), r->size_QP * r->N * 8, cudaMemcpyDeviceToDevice);
    });
}
// gen_publickey (src/secretkey.cu:493-505): [2][size_QP][N] copied out
int rref_gen_publickey(void *h, uint64_t *out_dev) {
    auto *r = static_cast<Ref *>(h);
    return guard([&] {
        PhantomPublicKey pk = r->sk->gen_publickey(*r->ctx);
        cudaStreamSynchronize(S());
        cudaMemcpy(out_dev, pk.pk_.data(), 2 * r->size_QP * r->N * 8, cudaMemcpyDeviceToDevice);
    });
}
static void fill_relinkey(PhantomRelinKey &k, const uint64_t *const *digits, size_t n) {
    k.public_keys_ptr_ = make_cuda_auto_ptr<uint64_t *>(n, S());
    cudaMemcpyAsync(k.public_keys_ptr_.get(), digits, n * sizeof(uint64_t *), cudaMemcpyHostToDevice, S());
    cudaStreamSynchronize(S());
    k.gen_flag_ = true;
}
void *rref_relinkey_from_raw(const uint64_t *const *digits, size_t n) {
    PhantomRelinKey *k = nullptr;
    if (guard([&] {
            k = new PhantomRelinKey();
            fill_relinkey(*k, digits, n);
        }))
        return nullptr;
    return k;
}
void *rref_rotkeys_new() { return new PhantomGaloisKeyFused(); }
int rref_rotkeys_set_raw(void *keys, uint32_t auto_index, const uint64_t *const *digits, size_t n) {
    return guard([&] {
        PhantomRelinKey k;
        fill_relinkey(k, digits, n);
        static_cast<PhantomGaloisKeyFused *>(keys)->setKey((int32_t) auto_index, std::move(k));
    });
}

// ---- the reference's own save() / load() through memory buffers (tests/test_serial_gpu.py)
// kind 0: PhantomCiphertext*, 1: PhantomRelinKey*, 2: the context's PhantomSecretKey (obj = Ref handle)
long rref_save(int kind, void *obj, uint8_t *buf, size_t cap) {
    std::string bytes;
    if (guard([&] {
            std::ostringstream os(std::ios::binary);
            if (kind == 0) static_cast<PhantomCiphertext *>(obj)->save(os);
            else if (kind == 1) static_cast<PhantomRelinKey *>(obj)->save(os);
            else static_cast<Ref *>(obj)->sk->save(os);
            bytes = os.str();
        }))
        return -1;
    if (buf && bytes.size() <= cap) std::memcpy(buf, bytes.data(), bytes.size());
    return (long) bytes.size();
}
void *rref_load(int kind, void *h, const uint8_t *buf, size_t len) {
    void *out = nullptr;
    if (guard([&] {
            std::istringstream is(std::string(reinterpret_cast<const char *>(buf), len), std::ios::binary);
            if (kind == 0) {
                auto *c = new PhantomCiphertext();
                c->load(is);
                out = c;
            } else if (kind == 1) {
                auto *k = new PhantomRelinKey();
                k->load(is);
                out = k;
            } else {
                static_cast<Ref *>(h)->sk->load(is);
                out = h;
            }
            cudaStreamSynchronize(S());
        }))
        return nullptr;
    return out;
}

// ---- encrypted CNN layers: phantom::DNN (include/dnn.cuh, src/dnn.cu) on the bootstrapper's context
// (tests/test_dnn_gpu.py).  Tensors are TensorCT objects; ciphertext handles handed out are borrowed.
struct RefDnn {
    RefBoot *b;
    std::unique_ptr<DNN> dnn;
    std::unique_ptr<PhantomPublicKey> pk;
};
void *rref_dnn_create(void *bh, double scale, const int32_t *rots, size_t nrots) {
    auto *b = static_cast<RefBoot *>(bh);
    auto *d = new RefDnn();
    d->b = b;
    if (guard([&] {
            d->dnn = std::make_unique<DNN>((size_t) scale, *b->r->enc, *b->r->ctx);
            d->dnn->setScalingFactors(b->sf, b->sf_big);
            std::vector<int32_t> v(rots, rots + nrots);
            d->dnn->BuildGaloisKey(*b->r->sk, v);
            d->dnn->RelinKeyGen(*b->r->sk);
            d->pk = std::make_unique<PhantomPublicKey>(b->r->sk->gen_publickey(*b->r->ctx));
            cudaStreamSynchronize(S());
        })) {
        delete d;
        return nullptr;
    }
    return d;
}
void rref_dnn_destroy(void *d) { delete static_cast<RefDnn *>(d); }
void *rref_dnn_galois_keys(void *d) { return &static_cast<RefDnn *>(d)->dnn->gk_; }
void *rref_dnn_mult_key(void *d) { return &static_cast<RefDnn *>(d)->dnn->mul_key_; }
int rref_dnn_rotation_indices(void *dh, int32_t *inout, size_t count, size_t cap, size_t width, size_t kh, size_t slotstr, int avgpool) {
    auto *d = static_cast<RefDnn *>(dh);
    std::vector<int32_t> v(inout, inout + count);
    if (avgpool) d->dnn->AddAvgPoolRotationsTo(v, width, slotstr);
    else d->dnn->AddRotationIndicesTo(v, width, kh, slotstr);
    for (size_t i = 0; i < v.size() && i < cap; i++) inout[i] = v[i];
    return (int) v.size();
}
static TensorCT *newt(TensorCT &&t) { return new TensorCT(std::move(t)); }
void *rref_tensor_new(void *const *cts, size_t n, size_t width, size_t slotstr, size_t num_ch) {
    auto *t = new TensorCT();
    for (size_t i = 0; i < n; i++) t->ctks_.push_back(*static_cast<PhantomCiphertext *>(cts[i]));
    t->width_ = width;
    t->slotstr_ = slotstr;
    t->num_ch_ = num_ch;
    cudaStreamSynchronize(S());
    return t;
}
void rref_tensor_free(void *t) { delete static_cast<TensorCT *>(t); }
size_t rref_tensor_size(void *t) { return static_cast<TensorCT *>(t)->ctks_.size(); }
size_t rref_tensor_width(void *t) { return static_cast<TensorCT *>(t)->width_; }
size_t rref_tensor_slotstr(void *t) { return static_cast<TensorCT *>(t)->slotstr_; }
size_t rref_tensor_num_ch(void *t) { return static_cast<TensorCT *>(t)->num_ch_; }
void *rref_tensor_ct(void *t, size_t i) { return &static_cast<TensorCT *>(t)->ctks_[i]; }
#define RD(dh) auto *d = static_cast<RefDnn *>(dh); auto &T = *static_cast<TensorCT *>(th); TensorCT *o = nullptr
void *rref_dnn_enc_tensor(void *dh, const double *image, size_t width, size_t num_ch) {
    auto *d = static_cast<RefDnn *>(dh);
    TensorCT *o = nullptr;
    if (guard([&] {
            std::vector<std::vector<std::vector<double>>> img(width, std::vector<std::vector<double>>(width, std::vector<double>(num_ch)));
            for (size_t i = 0; i < width; i++)
                for (size_t j = 0; j < width; j++)
                    for (size_t k = 0; k < num_ch; k++) img[i][j][k] = image[(i * width + j) * num_ch + k];
            o = newt(d->dnn->EncTensor(img, *d->pk));
        }))
        return nullptr;
    return o;
}
int rref_dnn_dec_tensor(void *dh, void *th, double *out) {
    auto *d = static_cast<RefDnn *>(dh);
    auto &T = *static_cast<TensorCT *>(th);
    return guard([&] {
        auto v = d->dnn->DecTensor(T, *d->b->r->sk);
        for (size_t i = 0; i < T.width_; i++)
            for (size_t j = 0; j < T.width_; j++)
                for (size_t k = 0; k < T.num_ch_; k++) out[(i * T.width_ + j) * T.num_ch_ + k] = v[i][j][k];
    });
}
void *rref_dnn_conv(void *dh, void *th, const double *w, size_t kh, size_t in_ch, size_t out_ch, size_t stride) {
    RD(dh);
    if (guard([&] {
            std::vector<std::vector<std::vector<std::vector<double>>>> W(
                kh, std::vector<std::vector<std::vector<double>>>(kh, std::vector<std::vector<double>>(in_ch, std::vector<double>(out_ch))));
            for (size_t i = 0; i < kh; i++)
                for (size_t j = 0; j < kh; j++)
                    for (size_t k = 0; k < in_ch; k++)
                        for (size_t h = 0; h < out_ch; h++) W[i][j][k][h] = w[((i * kh + j) * in_ch + k) * out_ch + h];
            o = newt(d->dnn->Conv(T, W, stride));
        }))
        return nullptr;
    return o;
}
void *rref_dnn_batch_norm(void *dh, void *th, const double *w, const double *bias, const double *mean, const double *var) {
    RD(dh);
    const size_t n = T.num_ch_;
    if (guard([&] {
            o = newt(d->dnn->BatchNorm(T, std::vector<double>(w, w + n), std::vector<double>(bias, bias + n),
                                       std::vector<double>(mean, mean + n), std::vector<double>(var, var + n)));
        }))
        return nullptr;
    return o;
}
void *rref_dnn_add(void *dh, void *th, void *th2) {
    RD(dh);
    if (guard([&] { o = newt(d->dnn->Add(T, *static_cast<TensorCT *>(th2))); })) return nullptr;
    return o;
}
void *rref_dnn_relu(void *dh, void *th, int a, int b, int deg) {
    RD(dh);
    if (guard([&] { o = newt(d->dnn->Relu(T, d->dnn->mul_key_, a, b, deg)); })) return nullptr;
    return o;
}
void *rref_dnn_relu_deg_two(void *dh, void *th) {
    RD(dh);
    if (guard([&] { o = newt(d->dnn->ReluDegTwo(T, d->dnn->mul_key_)); })) return nullptr;
    return o;
}
void *rref_dnn_sign(void *dh, void *th, size_t k) {
    RD(dh);
    if (guard([&] { o = newt(d->dnn->Sign(T, d->dnn->mul_key_, k)); })) return nullptr;
    return o;
}
void *rref_dnn_relu_composite(void *dh, void *th) {
    RD(dh);
    if (guard([&] { o = newt(d->dnn->ReluComposite(T, *d->b->fhe)); })) return nullptr;
    return o;
}
void *rref_dnn_bootstrap(void *dh, void *th) {
    RD(dh);
    if (guard([&] { o = newt(d->dnn->BootStrap(T, *d->b->fhe)); })) return nullptr;
    return o;
}
void *rref_dnn_avgpool_fullcon(void *dh, void *th, const double *w, size_t Tn, size_t t, const double *bias) {
    RD(dh);
    if (guard([&] {
            std::vector<std::vector<double>> W(Tn, std::vector<double>(t));
            for (size_t u = 0; u < Tn; u++)
                for (size_t k = 0; k < t; k++) W[u][k] = w[u * t + k];
            o = newt(d->dnn->AvgPoolFullCon(T, W, std::vector<double>(bias, bias + Tn)));
        }))
        return nullptr;
    return o;
}
void *rref_dnn_softmax(void *dh, void *th, size_t BoundB, size_t BoundR, size_t GumbelLambda, size_t GoldSchmidtD) {
    RD(dh);
    if (guard([&] { o = newt(d->dnn->SoftMax(T, d->dnn->mul_key_, *d->b->fhe, BoundB, BoundR, GumbelLambda, GoldSchmidtD)); })) return nullptr;
    return o;
}
// wall time of fn(layer) with a device synchronise on both sides, for bench legs: 0 conv, 1 relu_composite
int rref_dnn_time_conv(void *dh, void *th, const double *w, size_t kh, size_t in_ch, size_t out_ch, size_t stride, int iters, float *ms) {
    float total = 0;
    for (int it = 0; it < iters; it++) {
        cudaDeviceSynchronize();
        cudaEvent_t a, b;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        cudaEventRecord(a, S());
        void *o = rref_dnn_conv(dh, th, w, kh, in_ch, out_ch, stride);
        cudaEventRecord(b, S());
        cudaEventSynchronize(b);
        float t = 0;
        cudaEventElapsedTime(&t, a, b);
        total += t;
        cudaEventDestroy(a);
        cudaEventDestroy(b);
        if (!o) return -1;
        rref_tensor_free(o);
    }
    *ms = total / iters;
    return 0;
}

// ---- SEAL-style evaluator surface (tests/test_seal_api_gpu.py): PhantomGaloisKey, rotate / hoisting, plain ops
void *rref_gen_galois_keys(void *h) {
    auto *r = static_cast<Ref *>(h);
    PhantomGaloisKey *k = nullptr;
    if (guard([&] { k = new PhantomGaloisKey(r->sk->create_galois_keys(*r->ctx)); })) return nullptr;
    return k;
}
void rref_free_galois_keys(void *k) { delete static_cast<PhantomGaloisKey *>(k); }
// Galois elements the context holds keys for, in key order; returns the count
int rref_galois_elts(void *h, uint32_t *out, size_t cap) {
    auto *r = static_cast<Ref *>(h);
    auto &e = r->ctx->key_galois_tool_->galois_elts();
    for (size_t i = 0; i < e.size() && i < cap; i++) out[i] = e[i];
    return (int) e.size();
}
void *rref_galois_key_get(void *keys, size_t idx) {
    return const_cast<PhantomRelinKey *>(&static_cast<PhantomGaloisKey *>(keys)->get_relin_keys(idx));
}
void *rref_encode_plain(void *h, const double *vals, size_t n, double scale, size_t chain_index) {
    auto *r = static_cast<Ref *>(h);
    PhantomPlaintext *pt = nullptr;
    if (guard([&] {
            pt = new PhantomPlaintext();
            r->enc->encode(*r->ctx, std::vector<double>(vals, vals + n), scale, *pt, chain_index);
            cudaStreamSynchronize(S());
        }))
        return nullptr;
    return pt;
}
void rref_plain_free(void *p) { delete static_cast<PhantomPlaintext *>(p); }
uint64_t *rref_plain_data(void *p) { return static_cast<PhantomPlaintext *>(p)->data(); }
#define CT(x) (*static_cast<PhantomCiphertext *>(x))
int rref_seal_op(void *h, int op, void *a, void *b, void *keys, int arg, const int *steps, size_t nsteps) {
    auto *r = static_cast<Ref *>(h);
    auto &ctx = *r->ctx;
    return guard([&] {
        switch (op) {
            case 0: negate_inplace(ctx, CT(a)); break;
            case 1: add_plain_inplace(ctx, CT(a), *static_cast<PhantomPlaintext *>(b)); break;
            case 2: sub_plain_inplace(ctx, CT(a), *static_cast<PhantomPlaintext *>(b)); break;
            case 3: multiply_plain_inplace(ctx, CT(a), *static_cast<PhantomPlaintext *>(b)); break;
            case 4: add_inplace(ctx, CT(a), CT(b)); break;
            case 5: sub_inplace(ctx, CT(a), CT(b)); break;
            case 6: multiply_inplace(ctx, CT(a), CT(b)); break;
            case 7: relinearize_inplace(ctx, CT(a), *static_cast<PhantomRelinKey *>(keys)); break;
            case 8: CT(a) = rescale_to_next(ctx, CT(a)); break;
            case 9: CT(a) = mod_switch_to_next(ctx, CT(a)); break;
            case 10: apply_galois_inplace(ctx, CT(a), (size_t) (uint32_t) arg, *static_cast<PhantomGaloisKey *>(keys)); break;
            case 11: rotate_inplace(ctx, CT(a), arg, *static_cast<PhantomGaloisKey *>(keys)); break;
            case 12: hoisting_inplace(ctx, CT(a), *static_cast<PhantomGaloisKey *>(keys), std::vector<int>(steps, steps + nsteps)); break;
            default: throw std::invalid_argument("unknown op");
        }
        cudaStreamSynchronize(S());
    });
}
void *rref_add_many(void *h, void *const *cts, size_t n) {
    auto *r = static_cast<Ref *>(h);
    PhantomCiphertext *o = nullptr;
    if (guard([&] {
            std::vector<PhantomCiphertext> v;
            for (size_t i = 0; i < n; i++) v.push_back(CT(cts[i]));
            o = new PhantomCiphertext();
            add_many(*r->ctx, v, *o);
            cudaStreamSynchronize(S());
        }))
        return nullptr;
    return o;
}
}
