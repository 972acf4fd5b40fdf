// capi_tables.cu — the entries a C++ shim over the reference's own types needs (shim/): a context built on twiddle tables
// the caller already holds on the device (DNTTTable, include/ntt.cuh:34-129 of the reference), transforms over an explicit
// limb -> prime map (the reference's launcher variants, include/ntt.cuh:174-226), and the key switch with the digit
// pointers in DEVICE memory (PhantomRelinKey::public_keys_ptr(), include/secretkey.h:124).
#include <cstring>

#include "../../include/b200fhe.h"
#include "context.h"

using namespace b200fhe;

#define CU(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) return fail(B200FHE_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)
#define NEED_DEV(ctx)                                                                      \
    do {                                                                                   \
        if (!(ctx)) return fail(B200FHE_ERR_ARG, "null context");                          \
        if ((ctx)->host_only) return fail(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback"); \
    } while (0)

static inline cudaStream_t S(void *s) { return reinterpret_cast<cudaStream_t>(s); }

// (w, w_shoup) pairs: one 16-byte load per butterfly twiddle
__global__ void interleave_kernel(ulonglong2 *out, const u64 *a, const u64 *b, size_t n) {
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        out[i] = make_ulonglong2(a[i], b[i]);
}

int Context::init_from_tables(int logN_, const uint64_t *primes_QP, size_t size_QP_, size_t size_P_, const uint64_t *tw,
                              const uint64_t *tw_shoup, const uint64_t *itw, const uint64_t *itw_shoup,
                              const uint64_t *ninv, const uint64_t *ninv_shoup, cudaStream_t st) {
    if (logN_ < 1 || logN_ > 17) return fail(B200FHE_ERR_ARG, "logN out of range");
    if (!primes_QP || size_QP_ == 0 || size_P_ >= size_QP_) return fail(B200FHE_ERR_ARG, "bad prime list");
    if (!tw || !tw_shoup || !itw || !itw_shoup || !ninv || !ninv_shoup) return fail(B200FHE_ERR_ARG, "null table");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(B200FHE_ERR_CUDA, "no CUDA device available (no CPU fallback)");
    logN = logN_;
    N = size_t(1) << logN;
    size_QP = size_QP_;
    size_P = size_P_;
    size_Q = size_QP - size_P;
    host_only = false;
    external_tables = true;
    primes.assign(primes_QP, primes_QP + size_QP);
    hmods.resize(size_QP);
    for (size_t i = 0; i < size_QP; i++) {
        const uint64_t q = primes[i];
        if (q >> 61 || q < 2) return fail(B200FHE_ERR_ARG, "modulus must be at most 61 bits");
            if (!(q >> 33)) lazy_conv = false;
        hmods[i].q = q;
        hmods[i].nq = 0 - q;
        uint64_t r0, r1;
        host::barrett_ratio(q, r0, r1);
        hmods[i].r0 = r0;
        hmods[i].r1 = r1;
    }
    CU(cudaGetDevice(&device));
    CU(d_mods.upload(hmods.data(), size_QP * sizeof(Mod)));
    const size_t n = size_QP * N;
    CU(d_tw.alloc(n * sizeof(ulonglong2)));
    CU(d_itw.alloc(n * sizeof(ulonglong2)));
    CU(d_ninv.alloc(size_QP * 8));
    CU(d_ninv_shoup.alloc(size_QP * 8));
    const unsigned blocks = (unsigned) std::min<size_t>((n + 255) / 256, 148 * 8);
    interleave_kernel<<<blocks, 256, 0, st>>>(d_tw.as<ulonglong2>(), (const u64 *) tw, (const u64 *) tw_shoup, n);
    interleave_kernel<<<blocks, 256, 0, st>>>(d_itw.as<ulonglong2>(), (const u64 *) itw, (const u64 *) itw_shoup, n);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(d_ninv.p, ninv, size_QP * 8, cudaMemcpyDeviceToDevice, st));
    CU(cudaMemcpyAsync(d_ninv_shoup.p, ninv_shoup, size_QP * 8, cudaMemcpyDeviceToDevice, st));
    CU(cudaStreamSynchronize(st));
    count_launches(2);
    T.logN = logN;
    T.size_QP = (u32) size_QP;
    T.mods = d_mods.as<Mod>();
    T.tw = d_tw.as<ulonglong2>();
    T.itw = d_itw.as<ulonglong2>();
    T.ninv = d_ninv.as<u64>();
    T.ninv_shoup = d_ninv_shoup.as<u64>();
    CU(ws_perm.alloc(N * sizeof(u32)));
    return B200FHE_OK;
}

extern "C" int b200fhe_ctx_create_from_tables(int logN, const uint64_t *primes_QP_host, size_t size_QP, size_t size_P,
                                              const uint64_t *tw, const uint64_t *tw_shoup, const uint64_t *itw,
                                              const uint64_t *itw_shoup, const uint64_t *ninv, const uint64_t *ninv_shoup,
                                              void *stream, b200fhe_ctx **out) {
    if (!out) return fail(B200FHE_ERR_ARG, "null out");
    auto *c = new Context();
    int rc = c->init_from_tables(logN, primes_QP_host, size_QP, size_P, tw, tw_shoup, itw, itw_shoup, ninv, ninv_shoup,
                                 S(stream));
    if (rc != B200FHE_OK) {
        delete c;
        return rc;
    }
    *out = reinterpret_cast<b200fhe_ctx *>(c);
    return B200FHE_OK;
}

extern "C" int b200fhe_ctx_set_special(b200fhe_ctx *ctx_, size_t size_P) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    if (!c) return fail(B200FHE_ERR_ARG, "null context");
    std::lock_guard<std::mutex> g(c->mu);
    if (size_P == c->size_P) return B200FHE_OK;
    if (size_P >= c->size_QP) return fail(B200FHE_ERR_ARG, "size_P must be below the number of primes");
    if (!c->levels.empty() || !c->host_levels.empty())
        return fail(B200FHE_ERR_ARG, "the special-prime count is fixed once a level has been used (%zu)", c->size_P);
    c->size_P = size_P;
    c->size_Q = c->size_QP - size_P;
    return B200FHE_OK;
}

extern "C" int b200fhe_ctx_info(const b200fhe_ctx *ctx_, int *logN, size_t *size_QP, size_t *size_P, int *device) {
    auto *c = reinterpret_cast<const Context *>(ctx_);
    if (!c) return fail(B200FHE_ERR_ARG, "null context");
    if (logN) *logN = c->logN;
    if (size_QP) *size_QP = c->size_QP;
    if (size_P) *size_P = c->size_P;
    if (device) *device = c->device;
    return B200FHE_OK;
}

// limb range -> LimbMap.  special_from == 0 means the identity map.
static int make_map(const Context *c, const b200fhe_limb_range *r, LimbMap &m) {
    if (!r) return fail(B200FHE_ERR_ARG, "null limb range");
    m = LimbMap{};
    m.start = (u32) r->start;
    m.limbs = (u32) r->limbs;
    m.npoly = 1;
    m.size_Ql = (u32) r->special_from;
    m.size_Q = (u32) r->special_base;
    m.excl_lo = (u32) r->excl_start;
    m.excl_hi = (u32) r->excl_end;
    if (r->limbs) {
        auto prime = [&](size_t t) { return r->special_from == 0 || t < r->special_from ? t : r->special_base + (t - r->special_from); };
        const size_t last = r->start + r->limbs - 1;
        bool ok = prime(r->start) < c->size_QP && prime(last) < c->size_QP;
        if (r->special_from > r->start && r->special_from <= last) ok = ok && r->special_from - 1 < c->size_QP;
        if (!ok) return fail(B200FHE_ERR_ARG, "limb range maps outside the %zu primes of the tables", c->size_QP);
    }
    return B200FHE_OK;
}

extern "C" int b200fhe_ntt_forward_ex(b200fhe_ctx *ctx_, uint64_t *out, const uint64_t *in, const b200fhe_limb_range *r,
                                      void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    if (!out || !in) return fail(B200FHE_ERR_ARG, "null argument");
    LimbMap m;
    if (int rc = make_map(c, r, m)) return rc;
    if (m.limbs == 0) return B200FHE_OK;
    CU(ntt_forward(c->T, (u64 *) out, (const u64 *) in, m, nullptr, S(stream)));
    count_launches(ntt_launch_count_forward(c->logN));
    return B200FHE_OK;
}

extern "C" int b200fhe_ntt_inverse_ex(b200fhe_ctx *ctx_, uint64_t *out, const uint64_t *in, const b200fhe_limb_range *r,
                                      const uint64_t *scale, const uint64_t *scale_shoup, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    if (!out || !in) return fail(B200FHE_ERR_ARG, "null argument");
    if ((scale == nullptr) != (scale_shoup == nullptr)) return fail(B200FHE_ERR_ARG, "scale and scale_shoup go together");
    LimbMap m;
    if (int rc = make_map(c, r, m)) return rc;
    if (m.excl_hi > m.excl_lo) return fail(B200FHE_ERR_ARG, "the inverse transform has no excluded range");
    if (m.limbs == 0) return B200FHE_OK;
    CU(ntt_inverse(c->T, (u64 *) out, (const u64 *) in, m, (const u64 *) scale, (const u64 *) scale_shoup, S(stream)));
    count_launches(ntt_launch_count_inverse(c->logN));
    return B200FHE_OK;
}

// ---- key switch with the digit pointers in device memory
extern "C" int b200fhe_inner_prod_devkeys(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *cx, const uint64_t *t_mod_up,
                                          const uint64_t *const *evk_digits_dev, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    DeviceLevel *L;
    if (int rc = get_level(c, size_Ql, &L)) return rc;
    if (!cx || !t_mod_up || !evk_digits_dev) return fail(B200FHE_ERR_ARG, "null argument");
    EvkPtrs e{};
    e.dev_table = (const u64 *const *) evk_digits_dev;
    return do_inner_prod(c, L, (u64 *) cx, (const u64 *) t_mod_up, e, S(stream));
}

extern "C" int b200fhe_keyswitch_devkeys(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *ct, const uint64_t *c2,
                                         const uint64_t *const *evk_digits_dev, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    DeviceLevel *L;
    if (int rc = get_level(c, size_Ql, &L)) return rc;
    if (!ct || !c2 || !evk_digits_dev) return fail(B200FHE_ERR_ARG, "null argument");
    EvkPtrs e{};
    e.dev_table = (const u64 *const *) evk_digits_dev;
    return do_keyswitch(c, L, (u64 *) ct, (const u64 *) ct, 0, (const u64 *) c2, e, S(stream));
}
