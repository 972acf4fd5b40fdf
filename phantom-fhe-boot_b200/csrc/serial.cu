// serial.cu — byte formats of the reference's save / load for the objects on the path, so that files written by
// either library load in the other:
//   PhantomCiphertext::save / load      include/ciphertext.h:184-225   (header of 58 bytes, then the limbs)
//   PhantomPlaintext::save / load       include/plaintext.h:91-122
//   PhantomPublicKey::save / load       include/secretkey.h:84-97      (= the ciphertext format)
//   PhantomRelinKey::save / load        include/secretkey.h:129-163    (digit count, then one public key per digit)
//   PhantomSecretKey::save / load       include/secretkey.h (sk_max_power, N, |QP|, then the powers of s)
// PhantomGaloisKeyFused has no save / load in the reference; the container format here is ours: magic "B2GK",
// key count, then (automorphism index u32, relinearisation-key record) per key.
// Everything goes through HOST byte buffers supplied by the caller; the device copies are synchronous.
#include <cstring>

#include "../../include/b200fhe.h"
#include "context.h"

using namespace b200fhe;

namespace {
struct Writer {
    uint8_t *p;
    size_t cap, n = 0;
    bool ok = true;
    void put(const void *src, size_t len) {
        if (p && n + len <= cap) std::memcpy(p + n, src, len);
        else if (p) ok = false;
        n += len;
    }
    template <class T>
    void val(T v) { put(&v, sizeof v); }
    // device words -> buffer
    int dev(const uint64_t *d, size_t words) {
        if (p && n + words * 8 <= cap) {
            if (cudaMemcpy(p + n, d, words * 8, cudaMemcpyDeviceToHost) != cudaSuccess)
                return fail(B200FHE_ERR_CUDA, "device to host copy failed");
        } else if (p) ok = false;
        n += words * 8;
        return B200FHE_OK;
    }
};
struct Reader {
    const uint8_t *p;
    size_t len, n = 0;
    bool ok = true;
    void get(void *dst, size_t l) {
        if (n + l <= len) std::memcpy(dst, p + n, l);
        else ok = false;
        n += l;
    }
    template <class T>
    T val() {
        T v{};
        get(&v, sizeof v);
        return v;
    }
};

void put_ct_header(Writer &w, const b200fhe_ct_header &h) {
    w.val<uint64_t>(h.chain_index);
    w.val<uint64_t>(h.polys);
    w.val<uint64_t>(h.N);
    w.val<uint64_t>(h.limbs);
    w.val<double>(h.scale);
    w.val<uint64_t>(h.correction_factor);
    w.val<uint64_t>(h.noise_scale_deg);
    w.val<uint8_t>(h.is_ntt_form ? 1 : 0);
    w.val<uint8_t>(h.is_asymmetric ? 1 : 0);
}
bool get_ct_header(Reader &r, b200fhe_ct_header &h) {
    h.chain_index = r.val<uint64_t>();
    h.polys = r.val<uint64_t>();
    h.N = r.val<uint64_t>();
    h.limbs = r.val<uint64_t>();
    h.scale = r.val<double>();
    h.correction_factor = r.val<uint64_t>();
    h.noise_scale_deg = r.val<uint64_t>();
    h.is_ntt_form = r.val<uint8_t>();
    h.is_asymmetric = r.val<uint8_t>();
    // sanity: a corrupt header must not turn into a huge allocation or copy
    return r.ok && h.N >= 2 && h.N <= (1u << 20) && (h.N & (h.N - 1)) == 0 && h.polys >= 1 && h.polys <= 8 && h.limbs >= 1 &&
           h.limbs <= 4096;
}
int finish(const Writer &w, size_t *written) {
    if (written) *written = w.n;
    if (!w.ok) return fail(B200FHE_ERR_ARG, "capacity too small: %zu bytes needed", w.n);
    return B200FHE_OK;
}
}  // namespace

extern "C" {

int b200fhe_ct_save(const b200fhe_ct_header *hdr, const uint64_t *dev, uint8_t *out_host, size_t capacity, size_t *written) {
    if (!hdr || (!dev && out_host)) return fail(B200FHE_ERR_ARG, "null argument");
    Writer w{out_host, capacity};
    put_ct_header(w, *hdr);
    if (int rc = w.dev(dev, hdr->polys * hdr->limbs * hdr->N)) return rc;
    return finish(w, written);
}

int b200fhe_ct_load(const uint8_t *in_host, size_t len, b200fhe_ct_header *hdr, uint64_t *dev_out, size_t capacity_words,
                    size_t *consumed) {
    if (!in_host || !hdr) return fail(B200FHE_ERR_ARG, "null argument");
    Reader r{in_host, len};
    if (!get_ct_header(r, *hdr)) return fail(B200FHE_ERR_ARG, "not a ciphertext record (bad or truncated header)");
    const size_t words = hdr->polys * hdr->limbs * hdr->N;
    if (r.n + words * 8 > len) return fail(B200FHE_ERR_ARG, "truncated ciphertext record");
    if (dev_out) {
        if (capacity_words < words) return fail(B200FHE_ERR_ARG, "capacity too small: %zu words needed", words);
        if (cudaMemcpy(dev_out, in_host + r.n, words * 8, cudaMemcpyHostToDevice) != cudaSuccess)
            return fail(B200FHE_ERR_CUDA, "host to device copy failed");
    }
    if (consumed) *consumed = r.n + words * 8;
    return B200FHE_OK;
}

int b200fhe_plain_save(uint64_t chain_index, uint64_t N, uint64_t limbs, double scale, const uint64_t *dev, uint8_t *out_host,
                       size_t capacity, size_t *written) {
    if (!dev && out_host) return fail(B200FHE_ERR_ARG, "null argument");
    Writer w{out_host, capacity};
    w.val<uint64_t>(chain_index);
    w.val<uint64_t>(N);
    w.val<uint64_t>(limbs);
    w.val<double>(scale);
    if (int rc = w.dev(dev, limbs * N)) return rc;
    return finish(w, written);
}

int b200fhe_plain_load(const uint8_t *in_host, size_t len, uint64_t *chain_index, uint64_t *N, uint64_t *limbs, double *scale,
                       uint64_t *dev_out, size_t capacity_words, size_t *consumed) {
    if (!in_host) return fail(B200FHE_ERR_ARG, "null argument");
    Reader r{in_host, len};
    const uint64_t ci = r.val<uint64_t>(), n = r.val<uint64_t>(), l = r.val<uint64_t>();
    const double sc = r.val<double>();
    if (!r.ok || n < 2 || n > (1u << 20) || (n & (n - 1)) || l < 1 || l > 4096 || r.n + l * n * 8 > len)
        return fail(B200FHE_ERR_ARG, "not a plaintext record (bad or truncated)");
    if (dev_out) {
        if (capacity_words < l * n) return fail(B200FHE_ERR_ARG, "capacity too small: %zu words needed", (size_t) (l * n));
        if (cudaMemcpy(dev_out, in_host + r.n, l * n * 8, cudaMemcpyHostToDevice) != cudaSuccess)
            return fail(B200FHE_ERR_CUDA, "host to device copy failed");
    }
    if (chain_index) *chain_index = ci;
    if (N) *N = n;
    if (limbs) *limbs = l;
    if (scale) *scale = sc;
    if (consumed) *consumed = r.n + l * n * 8;
    return B200FHE_OK;
}

// relinearisation / Galois key record: dnum, then per digit a ciphertext record at chain index 0
int b200fhe_evk_save(b200fhe_ctx *ctx, const uint64_t *const *digits, size_t dnum, uint8_t *out_host, size_t capacity,
                     size_t *written) {
    auto *c = reinterpret_cast<Context *>(ctx);
    if (!c || !digits || dnum == 0) return fail(B200FHE_ERR_ARG, "null argument");
    Writer w{out_host, capacity};
    w.val<uint64_t>(dnum);
    const b200fhe_ct_header h{0, 2, c->N, c->size_QP, 1.0, 1, 1, 1, 0};
    for (size_t d = 0; d < dnum; d++) {
        put_ct_header(w, h);
        if (int rc = w.dev(digits[d], 2 * c->size_QP * c->N)) return rc;
    }
    return finish(w, written);
}

// digits_out: `capacity` device buffers of 2 * size_QP * N words each (may be NULL to query dnum and the length)
int b200fhe_evk_load(b200fhe_ctx *ctx, const uint8_t *in_host, size_t len, uint64_t *const *digits_out, size_t capacity,
                     size_t *dnum_out, size_t *consumed) {
    auto *c = reinterpret_cast<Context *>(ctx);
    if (!c || !in_host) return fail(B200FHE_ERR_ARG, "null argument");
    Reader r{in_host, len};
    const uint64_t dnum = r.val<uint64_t>();
    if (!r.ok || dnum == 0 || dnum > (uint64_t) kMaxDigits) return fail(B200FHE_ERR_ARG, "not an evaluation-key record");
    if (digits_out && capacity < dnum) return fail(B200FHE_ERR_ARG, "capacity too small: %zu digits", (size_t) dnum);
    const size_t words = 2 * c->size_QP * c->N;
    for (uint64_t d = 0; d < dnum; d++) {
        b200fhe_ct_header h;
        if (!get_ct_header(r, h)) return fail(B200FHE_ERR_ARG, "truncated evaluation-key record");
        if (h.polys != 2 || h.N != c->N || h.limbs != c->size_QP)
            return fail(B200FHE_ERR_ARG, "evaluation key of another parameter set (N %zu, limbs %zu)", (size_t) h.N,
                        (size_t) h.limbs);
        if (r.n + words * 8 > len) return fail(B200FHE_ERR_ARG, "truncated evaluation-key record");
        if (digits_out && cudaMemcpy(digits_out[d], in_host + r.n, words * 8, cudaMemcpyHostToDevice) != cudaSuccess)
            return fail(B200FHE_ERR_CUDA, "host to device copy failed");
        r.n += words * 8;
    }
    if (dnum_out) *dnum_out = dnum;
    if (consumed) *consumed = r.n;
    return B200FHE_OK;
}

// PhantomSecretKey record with sk_max_power = 1
int b200fhe_secret_save(b200fhe_ctx *ctx, const uint64_t *secret_dev, uint8_t *out_host, size_t capacity, size_t *written) {
    auto *c = reinterpret_cast<Context *>(ctx);
    if (!c || (!secret_dev && out_host)) return fail(B200FHE_ERR_ARG, "null argument");
    Writer w{out_host, capacity};
    w.val<uint64_t>(1);
    w.val<uint64_t>(c->N);
    w.val<uint64_t>(c->size_QP);
    if (int rc = w.dev(secret_dev, c->size_QP * c->N)) return rc;
    return finish(w, written);
}

// reads the first power only (the reference stores sk_max_power of them)
int b200fhe_secret_load(b200fhe_ctx *ctx, const uint8_t *in_host, size_t len, uint64_t *secret_dev_out, size_t *consumed) {
    auto *c = reinterpret_cast<Context *>(ctx);
    if (!c || !in_host) return fail(B200FHE_ERR_ARG, "null argument");
    Reader r{in_host, len};
    const uint64_t powers = r.val<uint64_t>(), n = r.val<uint64_t>(), limbs = r.val<uint64_t>();
    if (!r.ok || powers == 0 || powers > 16) return fail(B200FHE_ERR_ARG, "not a secret-key record");
    if (n != c->N || limbs != c->size_QP) return fail(B200FHE_ERR_ARG, "secret key of another parameter set");
    const size_t words = c->size_QP * c->N;
    if (r.n + powers * words * 8 > len) return fail(B200FHE_ERR_ARG, "truncated secret-key record");
    if (secret_dev_out && cudaMemcpy(secret_dev_out, in_host + r.n, words * 8, cudaMemcpyHostToDevice) != cudaSuccess)
        return fail(B200FHE_ERR_CUDA, "host to device copy failed");
    if (consumed) *consumed = r.n + powers * words * 8;
    return B200FHE_OK;
}

}  // extern "C"
