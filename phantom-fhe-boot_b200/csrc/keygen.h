// keygen.h — secret / public / evaluation key generation, encryption and decryption on the device
// (SURVEY.md 8f rank 3).  Replaces PhantomSecretKey::gen_secretkey / gen_publickey / gen_relinkey /
// EvalAtIndexKeyGen / encrypt_symmetric / ckks_decrypt and PhantomPublicKey::encrypt_asymmetric
// (src/secretkey.cu:12-190, 196-530, 576-683, 956-1024) and the samplers of src/prng.cu:139-244.
//
// What is kept from the reference: the distributions (ternary secret, centred binomial error with 21 + 21 coins,
// uniform mask by rejection), the key shapes ([2][|QP|][N] per digit, digit d carries P * s_old on the limbs of
// digit d), which secret a Galois key encrypts under (sigma_{k^-1}(s), for the key-switch-then-permute rotation)
// and the ciphertext layout.  What is not: the reference draws a fresh Salsa20 seed from the host for every
// polynomial; here ONE 32-byte master seed keys a ChaCha20 stream per sampled polynomial (stream number in the
// nonce), so a key set is reproducible from its seed, and the ternary sampler has no modulo bias.
// Keys are random: parity with the reference is functional (tests/test_keygen_gpu.py: each side decrypts and
// evaluates what the other side produced), bit-exact only for the deterministic decryption.
#pragma once
#include <memory>
#include <vector>

#include "context.h"

namespace b200fhe {

struct SwitchKey {  // one evaluation key: dnum digits, each [2][size_QP][N] in NTT form
    std::vector<std::unique_ptr<DevBuf>> digits;
    std::vector<const uint64_t *> ptrs() const {
        std::vector<const uint64_t *> p;
        for (auto &d : digits) p.push_back(d->as<uint64_t>());
        return p;
    }
};

class KeyGen {
  public:
    // seed: 32 bytes, or nullptr for one drawn from the operating system
    KeyGen(Context *ctx, const uint8_t *seed, cudaStream_t stream);
    // reference-compatible mode (same-seed key parity with the reference): every 64-byte seed the reference would draw from
    // std::random_device comes from the byte stream top8(splitmix64(master + i)), in the reference's order of draws, and the
    // samplers are the reference's own (Salsa20-core blocks keyed by the seed, nonce = coefficient index).  With the
    // reference's random_device answering from the same stream (oracle/ref_shim.cu: rref_seed_next_secret_key) secret, public,
    // relinearisation and rotation keys are bit-identical (tests/test_keygen_gpu.py).
    struct RefPrng {};
    KeyGen(Context *ctx, uint64_t reference_master_seed, cudaStream_t stream, RefPrng);
    Context *c;
    cudaStream_t st;
    size_t dnum;  // ceil(|Q| / |P|) digits per evaluation key (the reference's |Q| / |P| when |P| divides |Q|)

    const u64 *secret() const { return sk_.as<u64>(); }  // [size_QP][N], NTT form
    void set_secret(const u64 *dev);                     // adopt a secret key generated elsewhere (same layout)

    // enc_0 under s with P * s^2 (relinearisation), or under sigma_{k^-1}(s) with P * s (rotation by the
    // automorphism x -> x^k: EvalAtIndexKeyGen); conjugation is auto_index 2N - 1
    std::unique_ptr<SwitchKey> relin_key();
    std::unique_ptr<SwitchKey> galois_key(uint32_t auto_index);
    // SEAL-style Galois key of element g (create_galois_keys): from sigma_g(s) to s
    std::unique_ptr<SwitchKey> galois_key_seal(uint32_t galois_elt);
    // public key: an encryption of zero over all of QP, [2][size_QP][N]
    void public_key(u64 *out);

    // plain: [Ql][N] NTT form over q_0..q_{Ql-1}; ct: [2][Ql][N]
    void encrypt_symmetric(const u64 *plain, size_t size_Ql, u64 *ct);
    // c = (pk0 u + e0, pk1 u + e1) over Q_l P, divided by P (moddown), + plain on polynomial 0
    void encrypt_asymmetric(const u64 *pk, const u64 *plain, size_t size_Ql, u64 *ct);
    // plain_out [Ql][N] = c0 + c1 s (+ c2 s^2)
    void decrypt(const u64 *ct, size_t size_Ql, u32 polys, u64 *plain_out);

    // samplers (exposed for the distribution tests): out [limbs][N] over the primes of Q_l P (size_Ql == 0: the
    // first `limbs` primes); kind 0 ternary, 1 centred binomial, 2 uniform.  Coefficient form for 0 and 1.
    void sample(int kind, u64 *out, size_t limbs, size_t size_Ql);

  private:
    std::unique_ptr<SwitchKey> switch_key(const u64 *s_old, const u64 *s_new);
    // c0 = -(a s_new + e) (+ pmodq * s_old on limbs [inj_lo, inj_hi)) (+ plain), c1 = a; limbs over pm
    void enc_zero(u64 *c0, u64 *c1, const u64 *s_new, size_t limbs, PrimeMap pm, const u64 *s_old, size_t inj_lo,
                  size_t inj_hi, const u64 *plain);
    void cu(cudaError_t e, const char *what) const;
    void init_common();
    void next_seed(uint8_t out[64]);
    void sample_ref(int kind, u64 *out, size_t limbs, size_t size_Ql, const uint8_t seed[64]);
    bool ref_mode_ = false;
    uint64_t ref_master_ = 0, ref_drawn_ = 0;
    uint32_t key_[8];
    uint64_t next_stream_ = 0;
    DevBuf sk_, sk2_, tmp_, tmp2_, perm_;
    bool have_sk2_ = false;
};

int chacha20_block_device(const uint32_t *key8, uint32_t counter, const uint32_t *nonce3, uint32_t *out16_host);

}  // namespace b200fhe
