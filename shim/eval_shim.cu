// eval_shim.cu — the evaluator entries that are one composite call in libb200fhe.so: multiply_and_relin_inplace
// (src/evaluate.cu:1220), EvalRotateFused in both forms (:2154, :2198) and EvalConjFused (:2242).  CKKS, two-polynomial
// ciphertexts in NTT form at N = 2^12 .. 2^17; everything else goes to the reference's own bodies (b200ref_* aliases).
// The bookkeeping the reference does around the kernels (checks, scale, which fields the result inherits) is kept.
#include "shim_common.h"

using namespace phantom;
using b200shim::ck;

extern "C" {
void b200ref_multiply_and_relin_inplace(const PhantomContext *context, PhantomCiphertext *encrypted1, const PhantomCiphertext *encrypted2,
                                        const PhantomRelinKey *relin_keys);
void b200ref_EvalRotateFused_out(const PhantomContext *context, const PhantomGaloisKeyFused *galois_keys, PhantomCiphertext *in,
                                 PhantomCiphertext *out, int32_t index);
void b200ref_EvalRotateFused_ret(PhantomCiphertext *ret, const PhantomContext *context, const PhantomGaloisKeyFused *galois_keys,
                                 PhantomCiphertext *in, int32_t index);
void b200ref_EvalConjFused_out(const PhantomContext *context, const PhantomGaloisKeyFused *galois_keys, PhantomCiphertext *in,
                               PhantomCiphertext *out);
}

namespace {

std::vector<const uint64_t *> digits_of(const PhantomRelinKey &k) {
    std::vector<const uint64_t *> d;
    for (auto &pk : k.public_keys_) d.push_back(pk.pk_.data());
    return d;
}

// the composite path applies: CKKS, level 3 of the shim, the ring sizes the composite kernels cover, special primes present
bool fast(const PhantomContext &context, const PhantomCiphertext &ct) {
    if (b200shim::level() < 3 || ct.size() != 2 || !ct.is_ntt_form() || ct.chain_index() < 1) return false;
    auto &parms = context.get_context_data(ct.chain_index()).parms();
    const size_t n = parms.poly_modulus_degree();
    return parms.scheme() == scheme_type::ckks && n >= 4096 && n <= 131072 && context.get_context_data(0).parms().special_modulus_size() >= 1;
}

b200fhe_ctx *ctx_at(const PhantomContext &context, const PhantomCiphertext &ct, cudaStream_t s) {
    return b200shim::ctx_for_tool(context.gpu_rns_tables(), context.get_context_data(ct.chain_index()).gpu_rns_tool(), s);
}

// key switch c1 under the key of automorphism k, then the automorphism of both polynomials (src/evaluate.cu:2154-2196)
void rotate_into(const PhantomContext &context, const PhantomRelinKey &key, PhantomCiphertext &in, PhantomCiphertext &out, uint32_t k) {
    const cudaStream_t s = cudaStreamPerThread;
    if (in.size() > 2) throw std::invalid_argument("encrypted size must be 2");
    out = in;  // metadata and a buffer of its own, as the reference's `out_ciphertext = in_ciphertext`
    auto d = digits_of(key);
    ck(b200fhe_rotate_fused(ctx_at(context, in, s), in.coeff_modulus_size(), out.data(), in.data(), k, d.data(), (void *) s));
}

}  // namespace

namespace phantom {

void multiply_and_relin_inplace(const PhantomContext &context, PhantomCiphertext &encrypted1, const PhantomCiphertext &encrypted2,
                                const PhantomRelinKey &relin_keys) {
    if (!fast(context, encrypted1) || encrypted2.size() != 2 || !encrypted2.is_ntt_form())
        return b200ref_multiply_and_relin_inplace(&context, &encrypted1, &encrypted2, &relin_keys);
    // the reference's checks (src/evaluate.cu:1225-1243; its scale check is commented out)
    if (encrypted1.chain_index() != encrypted2.chain_index()) throw std::invalid_argument("encrypted1 and encrypted2 parameter mismatch");
    if (encrypted1.GetNoiseScaleDeg() != encrypted2.GetNoiseScaleDeg()) throw std::invalid_argument("noise mismatch");
    const cudaStream_t s = cudaStreamPerThread;
    auto d = digits_of(relin_keys);
    // tensor product, key switch of c2, both moddowns and the additions in one call, in place in encrypted1 (the reference
    // grows the ciphertext to three polynomials and shrinks it again, two reallocations)
    ck(b200fhe_hmult_relin(ctx_at(context, encrypted1, s), encrypted1.coeff_modulus_size(), encrypted1.data(), encrypted1.data(),
                           encrypted2.data(), d.data(), (void *) s));
    encrypted1.set_scale(encrypted1.scale() * encrypted2.scale());
}

void EvalRotateFused(const PhantomContext &context, const PhantomGaloisKeyFused &galois_keys, PhantomCiphertext &in_ciphertext,
                     PhantomCiphertext &out_ciphertext, int32_t index) {
    if (!fast(context, in_ciphertext)) return b200ref_EvalRotateFused_out(&context, &galois_keys, &in_ciphertext, &out_ciphertext, index);
    const size_t N = context.get_context_data(in_ciphertext.chain_index()).parms().poly_modulus_degree();
    const uint32_t k = FindAutomorphismIndex2nComplex(index, 2 * N);
    if (&in_ciphertext == &out_ciphertext) {
        PhantomCiphertext tmp;
        rotate_into(context, galois_keys.getKey(k), in_ciphertext, tmp, k);
        out_ciphertext = std::move(tmp);
        return;
    }
    rotate_into(context, galois_keys.getKey(k), in_ciphertext, out_ciphertext, k);
}

PhantomCiphertext EvalRotateFused(const PhantomContext &context, const PhantomGaloisKeyFused &galois_keys, PhantomCiphertext &in_ciphertext,
                                  int32_t index) {
    if (!fast(context, in_ciphertext)) {
        PhantomCiphertext ret;
        b200ref_EvalRotateFused_ret(&ret, &context, &galois_keys, &in_ciphertext, index);
        return ret;
    }
    const size_t N = context.get_context_data(in_ciphertext.chain_index()).parms().poly_modulus_degree();
    const uint32_t k = FindAutomorphismIndex2nComplex(index, 2 * N);
    PhantomCiphertext out;
    rotate_into(context, galois_keys.getKey(k), in_ciphertext, out, k);
    return out;
}

void EvalConjFused(const PhantomContext &context, const PhantomGaloisKeyFused &galois_keys, PhantomCiphertext &in_ciphertext,
                   PhantomCiphertext &out_ciphertext) {
    if (!fast(context, in_ciphertext) || &in_ciphertext == &out_ciphertext)
        return b200ref_EvalConjFused_out(&context, &galois_keys, &in_ciphertext, &out_ciphertext);
    const size_t N = context.get_context_data(in_ciphertext.chain_index()).parms().poly_modulus_degree();
    const uint32_t k = (uint32_t) (2 * N - 1);
    rotate_into(context, galois_keys.getKey(k), in_ciphertext, out_ciphertext, k);
}

}  // namespace phantom
