// capi.cu — the C ABI (include/b200fhe.h) over the sm_100a kernels.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/b200fhe.h"
#include "context.h"
#include "evalk.cuh"
#include "launch.cuh"

using namespace b200fhe;

static thread_local std::string g_err;
static std::atomic<uint64_t> g_launches{0};
void b200fhe::count_launches(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int b200fhe::fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(B200FHE_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)
#define LAUNCHED(n) g_launches.fetch_add(n, std::memory_order_relaxed)
#define NEED_DEV(ctx)                                                                      \
    do {                                                                                   \
        if (!(ctx)) return fail(B200FHE_ERR_ARG, "null context");                          \
        if ((ctx)->host_only) return fail(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback"); \
        int cur_ = -1;                                                                     \
        if (cudaGetDevice(&cur_) == cudaSuccess && cur_ != (ctx)->device)                  \
            return fail(B200FHE_ERR_ARG, "current CUDA device %d is not the context's device %d", cur_, (ctx)->device); \
    } while (0)

// ------------------------------------------------------------------------------------------- DevBuf
cudaError_t DevBuf::alloc(size_t n) {
    if (n <= bytes) return cudaSuccess;
    if (pinned) return cudaErrorNotSupported;  // see b200fhe_boot_graph_capture
    release();
    cudaError_t e = cudaMalloc(&p, n);
    if (e == cudaSuccess) bytes = n;
    return e;
}
cudaError_t DevBuf::upload(const void *h, size_t n) {
    if (n == 0) return cudaSuccess;
    cudaError_t e = alloc(n);
    if (e != cudaSuccess) return e;
    return cudaMemcpy(p, h, n, cudaMemcpyHostToDevice);
}
void DevBuf::release() {
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
}

// ------------------------------------------------------------------------------------------- Context
int Context::init(int logN_, const uint64_t *primes_QP, size_t size_QP_, size_t size_P_, int device_, bool host_only_) {
    if (logN_ < 1 || logN_ > 17) return fail(B200FHE_ERR_ARG, "logN out of range");
    if (!primes_QP || size_QP_ == 0 || size_P_ >= size_QP_) return fail(B200FHE_ERR_ARG, "bad prime list");
    logN = logN_;
    N = size_t(1) << logN;
    size_QP = size_QP_;
    size_P = size_P_;
    size_Q = size_QP - size_P;
    host_only = host_only_;
    primes.assign(primes_QP, primes_QP + size_QP);
    tabs.resize(size_QP);
    hmods.resize(size_QP);
    try {
        for (size_t i = 0; i < size_QP; i++) {
            const uint64_t q = primes[i];
            if (q >> 61) return fail(B200FHE_ERR_ARG, "modulus must be at most 61 bits");
            if (!(q >> 33)) lazy_conv = false;
            if (!host::is_prime(q) || (q - 1) % (2 * N)) return fail(B200FHE_ERR_ARG, "modulus %llu is not an NTT prime", (unsigned long long) q);
            tabs[i] = host::ntt_tables(logN, q);
            hmods[i].q = q;
            hmods[i].nq = 0 - q;
            uint64_t r0, r1;
            host::barrett_ratio(q, r0, r1);
            hmods[i].r0 = r0;
            hmods[i].r1 = r1;
        }
    } catch (const std::exception &e) {
        return fail(B200FHE_ERR_ARG, "table generation failed: %s", e.what());
    }
    if (host_only) return B200FHE_OK;

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(B200FHE_ERR_CUDA, "no CUDA device available (no CPU fallback)");
    if (device_ >= 0) CU(cudaSetDevice(device_));
    CU(cudaGetDevice(&device));
    CU(d_mods.upload(hmods.data(), size_QP * sizeof(Mod)));
    {
        std::vector<ulonglong2> buf(size_QP * N);
        for (size_t i = 0; i < size_QP; i++)
            for (size_t k = 0; k < N; k++) buf[i * N + k] = make_ulonglong2(tabs[i].tw[k], tabs[i].tw_shoup[k]);
        CU(d_tw.upload(buf.data(), buf.size() * sizeof(ulonglong2)));
        for (size_t i = 0; i < size_QP; i++)
            for (size_t k = 0; k < N; k++) buf[i * N + k] = make_ulonglong2(tabs[i].itw[k], tabs[i].itw_shoup[k]);
        CU(d_itw.upload(buf.data(), buf.size() * sizeof(ulonglong2)));
        std::vector<uint64_t> a(size_QP), b(size_QP);
        for (size_t i = 0; i < size_QP; i++) {
            a[i] = tabs[i].ninv;
            b[i] = tabs[i].ninv_shoup;
        }
        CU(d_ninv.upload(a.data(), size_QP * 8));
        CU(d_ninv_shoup.upload(b.data(), size_QP * 8));
    }
    T.logN = logN;
    T.size_QP = (u32) size_QP;
    T.mods = d_mods.as<Mod>();
    T.tw = d_tw.as<ulonglong2>();
    T.itw = d_itw.as<ulonglong2>();
    T.ninv = d_ninv.as<u64>();
    T.ninv_shoup = d_ninv_shoup.as<u64>();
    // eager per-level constants + workspace for the top level, so the hot path never allocates
    if (size_P > 0) {
        for (size_t l = 1; l <= size_Q; l++)
            if (!level(l)) return B200FHE_ERR_CUDA;
        const size_t Ql = size_Q, QlP = size_QP, beta = (Ql + size_P - 1) / size_P;
        CU(ws_tcks.alloc(Ql * N * 8));
        CU(ws_up.alloc(beta * QlP * N * 8));
        CU(ws_cx.alloc(2 * QlP * N * 8));
        CU(ws_delta.alloc(2 * Ql * N * 8));
    }
    CU(ws_tmp.alloc(3 * size_Q * N * 8));
    CU(ws_perm.alloc(N * sizeof(u32)));
    return B200FHE_OK;
}

const host::LevelConsts *Context::host_level(size_t size_Ql) {
    std::lock_guard<std::mutex> g(mu);
    auto it = host_levels.find(size_Ql);
    if (it != host_levels.end()) return it->second.get();
    try {
        auto p = std::make_unique<host::LevelConsts>(host::level_consts(primes, size_P, size_Ql));
        auto *raw = p.get();
        host_levels[size_Ql] = std::move(p);
        return raw;
    } catch (const std::exception &e) {
        fail(B200FHE_ERR_ARG, "level constants: %s", e.what());
        return nullptr;
    }
}

DeviceLevel *Context::level(size_t size_Ql) {
    {
        std::lock_guard<std::mutex> g(mu);
        auto it = levels.find(size_Ql);
        if (it != levels.end()) return it->second.get();
    }
    const host::LevelConsts *hc = host_level(size_Ql);
    if (!hc) return nullptr;
    auto L = std::make_unique<DeviceLevel>();
    L->h = *hc;
    const size_t Ql = size_Ql, P = size_P;
    bool ok = true;
    auto up = [&](DevBuf &b, const std::vector<uint64_t> &v) { ok = ok && b.upload(v.data(), v.size() * 8) == cudaSuccess; };
    up(L->partQlHatInv, hc->partQlHatInv);
    up(L->partQlHatInv_shoup, hc->partQlHatInv_shoup);
    for (auto &m : hc->modup_mat) {
        L->modup_mat.emplace_back(new DevBuf());
        up(*L->modup_mat.back(), m);
    }
    up(L->bigPInv, hc->bigPInv_mod_q);
    up(L->bigPInv_shoup, hc->bigPInv_mod_q_shoup);
    up(L->bigP, hc->bigP_mod_q);
    up(L->bigP_shoup, hc->bigP_mod_q_shoup);
    std::vector<uint64_t> s(Ql + P, 0), ss(Ql + P, 0);
    for (size_t i = 0; i < P; i++) {
        s[Ql + i] = hc->PHatInv_mod_p[i];
        ss[Ql + i] = hc->PHatInv_mod_p_shoup[i];
    }
    up(L->moddown_scale, s);
    up(L->moddown_scale_shoup, ss);
    up(L->PHat_mod_q, hc->PHat_mod_q);
    up(L->inv_q_last, hc->inv_q_last_mod_q);
    up(L->inv_q_last_shoup, hc->inv_q_last_mod_q_shoup);
    if (!ok) {
        fail(B200FHE_ERR_CUDA, "uploading level constants failed: %s", cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    std::lock_guard<std::mutex> g(mu);
    auto ins = levels.emplace(size_Ql, std::move(L));  // a concurrent first touch keeps the entry that got there first
    return ins.first->second.get();
}

LimbMap Context::limb_map(size_t start, size_t limbs, size_t size_Ql, size_t npoly, size_t stride) const {
    LimbMap m{};
    m.start = (u32) start;
    m.limbs = (u32) limbs;
    m.size_Ql = (u32) size_Ql;
    m.size_Q = (u32) size_Q;
    m.npoly = (u32) npoly;
    m.excl_alpha = 0;
    m.poly_stride = stride;
    m.src_poly_stride = stride;
    return m;
}

// ------------------------------------------------------------------------------------------- misc
extern "C" const char *b200fhe_last_error(void) { return g_err.c_str(); }
extern "C" const char *b200fhe_version(void) { return "b200fhe 0.1 (sm_100a)"; }
extern "C" uint64_t b200fhe_launch_count(void) { return g_launches.load(); }

extern "C" int b200fhe_coeff_modulus_create(size_t N, const int *bit_sizes, size_t count, uint64_t *out) {
    if (!bit_sizes || !out) return fail(B200FHE_ERR_ARG, "null argument");
    try {
        auto v = host::coeff_modulus_create(N, std::vector<int>(bit_sizes, bit_sizes + count));
        std::memcpy(out, v.data(), count * 8);
        return B200FHE_OK;
    } catch (const std::exception &e) {
        return fail(B200FHE_ERR_ARG, "%s", e.what());
    }
}

static int make_ctx(int logN, const uint64_t *primes, size_t size_QP, size_t size_P, int device, bool host_only,
                    b200fhe_ctx **out) {
    if (!out) return fail(B200FHE_ERR_ARG, "null out");
    auto *c = new Context();
    int rc = c->init(logN, primes, size_QP, size_P, device, host_only);
    if (rc != B200FHE_OK) {
        delete c;
        return rc;
    }
    *out = reinterpret_cast<b200fhe_ctx *>(c);
    return B200FHE_OK;
}

extern "C" int b200fhe_ctx_create(int logN, const uint64_t *primes, size_t size_QP, size_t size_P, int device,
                                  b200fhe_ctx **out) {
    return make_ctx(logN, primes, size_QP, size_P, device, false, out);
}
extern "C" int b200fhe_ctx_create_host(int logN, const uint64_t *primes, size_t size_QP, size_t size_P,
                                       b200fhe_ctx **out) {
    return make_ctx(logN, primes, size_QP, size_P, -1, true, out);
}
extern "C" int b200fhe_ctx_destroy(b200fhe_ctx *ctx) {
    delete reinterpret_cast<Context *>(ctx);
    return B200FHE_OK;
}

extern "C" int b200fhe_ctx_export_table(const b200fhe_ctx *ctx_, int which, uint64_t *out) {
    auto *c = reinterpret_cast<const Context *>(ctx_);
    if (!c || !out) return fail(B200FHE_ERR_ARG, "null argument");
    if (c->external_tables && which < 6) return fail(B200FHE_ERR_UNSUPPORTED, "context was built on caller-owned tables");
    for (size_t i = 0; i < c->size_QP; i++) {
        static const host::PrimeTables none{};
        const auto &t = c->external_tables ? none : c->tabs[i];
        switch (which) {
            case 0: std::memcpy(out + i * c->N, t.tw.data(), c->N * 8); break;
            case 1: std::memcpy(out + i * c->N, t.tw_shoup.data(), c->N * 8); break;
            case 2: std::memcpy(out + i * c->N, t.itw.data(), c->N * 8); break;
            case 3: std::memcpy(out + i * c->N, t.itw_shoup.data(), c->N * 8); break;
            case 4: out[i] = t.ninv; break;
            case 5: out[i] = t.ninv_shoup; break;
            case 6: out[i] = c->hmods[i].q; break;
            case 7: out[i] = c->hmods[i].r0; break;
            case 8: out[i] = c->hmods[i].r1; break;
            default: return fail(B200FHE_ERR_ARG, "unknown table id %d", which);
        }
    }
    return B200FHE_OK;
}

extern "C" int b200fhe_level_beta(b200fhe_ctx *ctx_, size_t size_Ql, size_t *beta) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    if (!c || !beta) return fail(B200FHE_ERR_ARG, "null argument");
    auto *h = c->host_level(size_Ql);
    if (!h) return B200FHE_ERR_ARG;
    *beta = h->beta;
    return B200FHE_OK;
}

extern "C" int b200fhe_level_export(b200fhe_ctx *ctx_, size_t size_Ql, int which, uint64_t *out, size_t cap,
                                    size_t *count) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    if (!c) return fail(B200FHE_ERR_ARG, "null context");
    auto *h = c->host_level(size_Ql);
    if (!h) return B200FHE_ERR_ARG;
    const std::vector<uint64_t> *v = nullptr;
    switch (which) {
        case 0: v = &h->partQlHatInv; break;
        case 1: v = &h->partQlHatInv_shoup; break;
        case 2: v = &h->bigP_mod_q; break;
        case 3: v = &h->bigPInv_mod_q; break;
        case 4: v = &h->bigPInv_mod_q_shoup; break;
        case 5: v = &h->PHatInv_mod_p; break;
        case 6: v = &h->PHatInv_mod_p_shoup; break;
        case 7: v = &h->PHat_mod_q; break;
        case 8: v = &h->inv_q_last_mod_q; break;
        case 9: v = &h->inv_q_last_mod_q_shoup; break;
        default:
            if (which >= 100 && size_t(which - 100) < h->beta) v = &h->modup_mat[which - 100];
            else return fail(B200FHE_ERR_ARG, "unknown level constant id %d", which);
    }
    if (count) *count = v->size();
    if (out) {
        if (cap < v->size()) return fail(B200FHE_ERR_ARG, "capacity too small");
        std::memcpy(out, v->data(), v->size() * 8);
    }
    return B200FHE_OK;
}

extern "C" int b200fhe_galois_table_host(int logN, uint32_t elt, uint32_t *perm) {
    if (!perm || logN < 1 || logN > 17) return fail(B200FHE_ERR_ARG, "bad argument");
    const uint32_t n = 1u << logN;
    for (uint32_t i = n; i < 2 * n; i++) {
        const uint32_t rev = host::reverse_bits(i, logN + 1);
        const uint64_t raw = (((uint64_t) elt * rev) >> 1) & (n - 1);
        perm[i - n] = host::reverse_bits((uint32_t) raw, logN);
    }
    return B200FHE_OK;
}

// ------------------------------------------------------------------------------------------- NTT
static inline cudaStream_t S(void *s) { return reinterpret_cast<cudaStream_t>(s); }

static int check_limbs(const Context *c, size_t limbs, size_t start, size_t size_Ql) {
    if (size_Ql > c->size_Q) return fail(B200FHE_ERR_ARG, "size_Ql exceeds |Q|");
    const size_t maxl = size_Ql ? size_Ql + c->size_P : c->size_QP;
    if (start + limbs > maxl) return fail(B200FHE_ERR_ARG, "limb range [%zu,%zu) exceeds %zu", start, start + limbs, maxl);
    return B200FHE_OK;
}

extern "C" int b200fhe_ntt_forward(b200fhe_ctx *ctx_, uint64_t *inout, size_t limbs, size_t start, size_t size_Ql,
                                   size_t npoly, size_t poly_stride, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    if (int rc = check_limbs(c, limbs, start, size_Ql)) return rc;
    if (limbs == 0 || npoly == 0) return B200FHE_OK;
    LimbMap m = c->limb_map(start, limbs, size_Ql, npoly, poly_stride);
    CU(ntt_forward(c->T, (u64 *) inout, (const u64 *) inout, m, nullptr, S(stream)));
    LAUNCHED(ntt_launch_count_forward(c->logN));
    return B200FHE_OK;
}

extern "C" int b200fhe_ntt_inverse(b200fhe_ctx *ctx_, uint64_t *inout, size_t limbs, size_t start, size_t size_Ql,
                                   size_t npoly, size_t poly_stride, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    if (int rc = check_limbs(c, limbs, start, size_Ql)) return rc;
    if (limbs == 0 || npoly == 0) return B200FHE_OK;
    LimbMap m = c->limb_map(start, limbs, size_Ql, npoly, poly_stride);
    CU(ntt_inverse(c->T, (u64 *) inout, (const u64 *) inout, m, nullptr, nullptr, S(stream)));
    LAUNCHED(ntt_launch_count_forward(c->logN));
    return B200FHE_OK;
}

extern "C" int b200fhe_ntt_inverse_scale(b200fhe_ctx *ctx_, uint64_t *out, const uint64_t *in, size_t limbs,
                                         size_t start, size_t size_Ql, const uint64_t *scale,
                                         const uint64_t *scale_shoup, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    if (int rc = check_limbs(c, limbs, start, size_Ql)) return rc;
    if ((scale == nullptr) != (scale_shoup == nullptr)) return fail(B200FHE_ERR_ARG, "scale and scale_shoup go together");
    if (limbs == 0) return B200FHE_OK;
    LimbMap m = c->limb_map(start, limbs, size_Ql, 1, 0);
    CU(ntt_inverse(c->T, (u64 *) out, (const u64 *) in, m, (const u64 *) scale, (const u64 *) scale_shoup, S(stream)));
    LAUNCHED(ntt_launch_count_forward(c->logN));
    return B200FHE_OK;
}

extern "C" int b200fhe_ntt_forward_fuse_moddown(b200fhe_ctx *ctx_, uint64_t *ct, const uint64_t *cx,
                                                const uint64_t *PInv, const uint64_t *PInv_shoup, uint64_t *delta,
                                                size_t limbs, size_t start, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    if (int rc = check_limbs(c, limbs, start, 0)) return rc;
    if (!ct || !cx || !PInv || !PInv_shoup || !delta) return fail(B200FHE_ERR_ARG, "null argument");
    if (limbs == 0) return B200FHE_OK;
    LimbMap m = c->limb_map(start, limbs, 0, 1, 0);
    FwdEpilogue e{};
    e.other = (const u64 *) cx;
    e.scale = (const u64 *) PInv;
    e.scale_shoup = (const u64 *) PInv_shoup;
    e.out = (u64 *) ct;
    CU(ntt_forward(c->T, (u64 *) delta, (const u64 *) delta, m, &e, S(stream)));
    LAUNCHED(ntt_launch_count_forward(c->logN));
    return B200FHE_OK;
}

static int host_roundtrip(Context *c, uint64_t *h, size_t limbs, size_t start, size_t size_Ql, bool fwd) {
    NEED_DEV(c);
    if (int rc = check_limbs(c, limbs, start, size_Ql)) return rc;
    if (limbs == 0) return B200FHE_OK;
    const size_t bytes = limbs * c->N * 8;
    CU(c->ws_tmp.alloc((start + limbs) * c->N * 8));
    u64 *d = c->ws_tmp.as<u64>();
    CU(cudaMemcpyAsync(d + start * c->N, h + start * c->N, bytes, cudaMemcpyHostToDevice, 0));
    LimbMap m = c->limb_map(start, limbs, size_Ql, 1, 0);
    if (fwd) CU(ntt_forward(c->T, d, d, m, nullptr, 0));
    else CU(ntt_inverse(c->T, d, d, m, nullptr, nullptr, 0));
    LAUNCHED(ntt_launch_count_forward(c->logN));
    CU(cudaMemcpyAsync(h + start * c->N, d + start * c->N, bytes, cudaMemcpyDeviceToHost, 0));
    CU(cudaStreamSynchronize(0));
    return B200FHE_OK;
}

Context::~Context() {
    for (int i = 0; i < kLanes; i++) {
        if (lane[i]) cudaStreamDestroy(lane[i]);
        if (lane_join[i]) cudaEventDestroy(lane_join[i]);
    }
    if (lane_fork) cudaEventDestroy(lane_fork);
    if (bg) cudaStreamDestroy(bg);
    for (cudaEvent_t e : {bg_start, bg_done, ipq_done})
        if (e) cudaEventDestroy(e);
    for (int i = 0; i < 3; i++)
        if (pipe_stream[i]) cudaStreamDestroy(pipe_stream[i]);
    for (int i = 0; i < kStageSlots; i++)
        for (int j = 0; j < 3; j++)
            if (pipe_ev[i][j]) cudaEventDestroy(pipe_ev[i][j]);
}

// Host-buffer batch: runs of limbs go H2D -> NTT -> D2H through four staging slots on three streams, so the copy-in of the
// next run and the copy-out of the previous one overlap the transform of the current one (full-duplex PCIe + compute).
extern "C" int b200fhe_ntt_host_batch(b200fhe_ctx *ctx_, uint64_t *inout_host, size_t limbs, size_t size_Ql,
                                      size_t npoly, int ops) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    if (int rc = check_limbs(c, limbs, 0, size_Ql)) return rc;
    if (ops < 1 || ops > 3 || !inout_host) return fail(B200FHE_ERR_ARG, "bad argument");
    if (limbs == 0 || npoly == 0) return B200FHE_OK;
    // Units of the pipeline: a polynomial may be cut into runs of limbs (limbs are independent transforms) so that the first
    // copy-in and the last copy-out, which nothing overlaps, get shorter (B200FHE_HOST_CHUNKS: runs per polynomial).  Measured
    // on the B200 at the headline shape (16 x 36 limbs, 302 MB each way): 7.17 / 7.72 / 8.47 / 9.70 ms per step with 1 / 2 / 4 /
    // 8 runs -- 19 MB copies already leave no gap on the duplex PCIe link and shorter ones only add per-copy cost -- so the
    // default stays one run = the whole polynomial.
    static const size_t kChunks = [] {
        const char *v = getenv("B200FHE_HOST_CHUNKS");
        return (size_t) std::max(1, v ? atoi(v) : 1);
    }();
    const size_t words = limbs * c->N;
    const size_t cl = (limbs + kChunks - 1) / kChunks, nchunk = (limbs + cl - 1) / cl;   // limbs per run, runs per polynomial
    constexpr int kSlots = Context::kStageSlots;
    for (int i = 0; i < kSlots; i++) {
        CU(c->stage[i].alloc(cl * c->N * 8));
        for (int j = 0; j < 3; j++)
            if (!c->pipe_ev[i][j]) CU(cudaEventCreateWithFlags(&c->pipe_ev[i][j], cudaEventDisableTiming));
    }
    for (int i = 0; i < 3; i++)
        if (!c->pipe_stream[i]) CU(cudaStreamCreateWithFlags(&c->pipe_stream[i], cudaStreamNonBlocking));
    cudaStream_t s_in = c->pipe_stream[0], s_cmp = c->pipe_stream[1], s_out = c->pipe_stream[2];
    // a failure in the middle must not return while copies into the caller's buffer are still in flight
#undef CU
#define CU(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) {                                                                       \
            for (int i_ = 0; i_ < 3; i_++) cudaStreamSynchronize(c->pipe_stream[i_]);                  \
            return fail(B200FHE_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_));                    \
        }                                                                                              \
    } while (0)
    for (size_t u = 0; u < npoly * nchunk; u++) {
        const size_t p = u / nchunk, l0 = (u % nchunk) * cl, ln = std::min(cl, limbs - l0);
        const int slot = (int) (u % kSlots);
        // the staging slot holds limbs [l0, l0 + ln) at the offsets the limb map expects: its base is moved back by l0 rows
        u64 *d = c->stage[slot].as<u64>();
        uint64_t *h = inout_host + p * words + l0 * c->N;
        const size_t bytes = ln * c->N * 8;
        LimbMap m = c->limb_map(l0, ln, size_Ql, 1, 0);
        if (u >= (size_t) kSlots) CU(cudaStreamWaitEvent(s_in, c->pipe_ev[slot][2], 0));  // slot drained by its last copy-out
        CU(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, s_in));
        CU(cudaEventRecord(c->pipe_ev[slot][0], s_in));
        CU(cudaStreamWaitEvent(s_cmp, c->pipe_ev[slot][0], 0));
        if (ops & 1) {
            CU(ntt_forward(c->T, d - l0 * c->N, d - l0 * c->N, m, nullptr, s_cmp));
            LAUNCHED(ntt_launch_count_forward(c->logN));
        }
        if (ops & 2) {
            CU(ntt_inverse(c->T, d - l0 * c->N, d - l0 * c->N, m, nullptr, nullptr, s_cmp));
            LAUNCHED(ntt_launch_count_forward(c->logN));
        }
        CU(cudaEventRecord(c->pipe_ev[slot][1], s_cmp));
        CU(cudaStreamWaitEvent(s_out, c->pipe_ev[slot][1], 0));
        CU(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, s_out));
        CU(cudaEventRecord(c->pipe_ev[slot][2], s_out));
    }
    CU(cudaStreamSynchronize(s_out));
#undef CU
#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(B200FHE_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)
    return B200FHE_OK;
}

extern "C" int b200fhe_ntt_forward_host(b200fhe_ctx *ctx, uint64_t *h, size_t limbs, size_t start, size_t size_Ql) {
    return host_roundtrip(reinterpret_cast<Context *>(ctx), h, limbs, start, size_Ql, true);
}
extern "C" int b200fhe_ntt_inverse_host(b200fhe_ctx *ctx, uint64_t *h, size_t limbs, size_t start, size_t size_Ql) {
    return host_roundtrip(reinterpret_cast<Context *>(ctx), h, limbs, start, size_Ql, false);
}

// ------------------------------------------------------------------------------------------- key switching
int b200fhe::get_level(Context *c, size_t size_Ql, DeviceLevel **out) {
    if (c->size_P == 0) return fail(B200FHE_ERR_ARG, "context has no special primes (size_P == 0)");
    if (size_Ql == 0 || size_Ql > c->size_Q) return fail(B200FHE_ERR_ARG, "size_Ql out of range");
    if (c->logN < 12 || c->logN > 17) return fail(B200FHE_ERR_UNSUPPORTED, "kernels support N = 2^12 .. 2^17");
    *out = c->level(size_Ql);
    return *out ? B200FHE_OK : B200FHE_ERR_CUDA;
}

// P -> Ql conversion of `batch` polys; src = the P limbs (coefficient form, already x PHatInv)
static int bconv_down(Context *c, DeviceLevel *L, u64 *dst, size_t dst_stride, const u64 *srcP, size_t src_stride,
                      u32 batch, cudaStream_t st, bool lazy = false) {
    BconvArgs a{};
    a.lazy = lazy && c->lazy_conv ? 1u : 0u;
    a.in = srcP;
    a.out = dst;
    a.mat = L->PHat_mod_q.as<u64>();
    a.ni = (u32) c->size_P;
    a.no = (u32) L->h.size_Ql;
    a.skip_start = a.no;
    a.skip_len = 0;
    a.pm = PrimeMap{0, 0};
    a.in_stride = src_stride;
    a.out_stride = dst_stride;
    a.batch = batch;
    if (c->size_P == 1) CU(bconv_single(c->T.mods, a, (u32) c->size_Q, c->N, st));
    else CU(bconv(c->T.mods, a, c->N, st));
    LAUNCHED(1);
    return B200FHE_OK;
}

extern "C" int b200fhe_bconv_P_to_Ql(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *dst, const uint64_t *src,
                                     void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    DeviceLevel *L;
    if (int rc = get_level(c, size_Ql, &L)) return rc;
    if (c->size_P == 1) return bconv_down(c, L, (u64 *) dst, 0, (const u64 *) src, 0, 1, S(stream));
    // reference bConv_BEHZ = bconv_mult (x QHatInvModq) then matmul: do the scaling into workspace
    const size_t P = c->size_P;
    CU(c->ws_tmp.alloc(P * c->N * 8));
    CU(scalar_mul(c->T.mods + c->size_Q, c->ws_tmp.as<u64>(), (const u64 *) src, L->moddown_scale.as<u64>() + size_Ql,
                  L->moddown_scale_shoup.as<u64>() + size_Ql, (u32) P, PrimeMap{0, 0}, c->N, S(stream)));
    LAUNCHED(1);
    return bconv_down(c, L, (u64 *) dst, 0, c->ws_tmp.as<u64>(), 0, 1, S(stream));
}

int b200fhe::do_modup(Context *c, DeviceLevel *L, u64 *dst, const u64 *c2, cudaStream_t st) {
    return do_modup_batch(c, L, dst, c2, 0, 1, st);
}

// `batch` polynomials c2 + z*c2_stride ([Ql][N]) -> dst + z*beta*QlP*N ([beta][QlP][N]); one launch set for all
int b200fhe::do_modup_batch(Context *c, DeviceLevel *L, u64 *dst, const u64 *c2, size_t c2_stride, u32 batch,
                            cudaStream_t st) {
    const size_t N = c->N, Ql = L->h.size_Ql, P = c->size_P, QlP = Ql + P, alpha = L->h.alpha, beta = L->h.beta;
    if (batch == 0) return B200FHE_OK;
    CU(c->ws_tcks.alloc(size_t(batch) * Ql * N * 8));
    u64 *t = c->ws_tcks.as<u64>();
    const size_t dst_stride = beta * QlP * N;
    LimbMap mi = c->limb_map(0, Ql, 0, batch, Ql * N);
    mi.src_poly_stride = c2_stride;
    if (alpha == 1) CU(ntt_inverse(c->T, t, c2, mi, nullptr, nullptr, st));
    else CU(ntt_inverse(c->T, t, c2, mi, L->partQlHatInv.as<u64>(), L->partQlHatInv_shoup.as<u64>(), st));
    LAUNCHED(2);
    // one launch for the beta conversions (and the copies of each digit's own limbs) when the digits are wide enough
    const bool merged = alpha >= 2 && alpha <= 16 && beta <= (size_t) kBconvDigits;
    BconvMultiArgs M{};
    for (size_t d = 0; d < beta; d++) {
        const size_t s = alpha * d, ni = L->h.digit_size(d);
        BconvArgs a{};
        a.in = t + s * N;
        a.out = dst + d * QlP * N;
        a.mat = L->modup_mat[d]->as<u64>();
        a.ni = (u32) ni;
        a.no = (u32) (QlP - ni);
        a.skip_start = (u32) s;
        a.skip_len = (u32) ni;
        a.pm = c->prime_map(Ql);
        a.in_stride = Ql * N;
        a.out_stride = dst_stride;
        a.batch = batch;
        a.lazy = c->lazy_conv ? 1u : 0u;   // the converted limbs go straight into the forward transform below
        if (merged && ni >= 1) {
            M.d[d] = a;
            M.own_src[d] = c2 + s * N;
            continue;
        }
        if (alpha == 1) CU(bconv_single(c->T.mods, a, (u32) s, N, st));
        else CU(bconv(c->T.mods, a, N, st));
        LAUNCHED(1);
        // the digit's own limbs stay as they came, in NTT form (modup_copy_partQl_kernel)
        CU(cudaMemcpy2DAsync(dst + d * QlP * N + s * N, dst_stride * 8, c2 + s * N, (batch > 1 ? c2_stride : Ql * N) * 8,
                             ni * N * 8, batch, cudaMemcpyDeviceToDevice, st));
    }
    if (merged) {
        M.ndig = (u32) beta;
        M.own_stride = batch > 1 ? c2_stride : Ql * N;
        CU(bconv_multi(c->T.mods, M, N, st));
        LAUNCHED(1);
    }
    LimbMap mf = c->limb_map(0, QlP, Ql, batch * beta, QlP * N);
    mf.excl_alpha = (u32) alpha;
    mf.excl_mod = (u32) beta;
    CU(ntt_forward(c->T, dst, dst, mf, nullptr, st));
    LAUNCHED(2);
    return B200FHE_OK;
}

int b200fhe::make_evk(const Context *c, const DeviceLevel *L, const uint64_t *const *digits, EvkPtrs &e) {
    if (!digits) return fail(B200FHE_ERR_ARG, "null evaluation key");
    if (L->h.beta > (size_t) kMaxDigits) return fail(B200FHE_ERR_UNSUPPORTED, "more than %d digits", kMaxDigits);
    for (size_t d = 0; d < L->h.beta; d++) {
        if (!digits[d]) return fail(B200FHE_ERR_ARG, "evaluation key digit %zu is null", d);
        e.d[d] = (const u64 *) digits[d];
    }
    (void) c;
    return B200FHE_OK;
}

int b200fhe::do_inner_prod(Context *c, DeviceLevel *L, u64 *cx, const u64 *up, const EvkPtrs &e, cudaStream_t st,
                           u32 batch) {
    const size_t Ql = L->h.size_Ql;
    CU(inner_prod(c->T.mods, cx, up, e, (u32) L->h.beta, (u32) (Ql + c->size_P), (u32) c->size_QP, c->prime_map(Ql),
                  c->N, st, batch));
    LAUNCHED(1);
    return B200FHE_OK;
}

// batch polys: cx + z*cx_stride ([QlP][N], P part clobbered) -> out + z*out_stride ([Ql][N]) (+ acc)
int b200fhe::do_moddown(Context *c, DeviceLevel *L, u64 *out, size_t out_stride, u64 *cx, size_t cx_stride, u32 batch,
                      const u64 *acc, size_t acc_stride, u32 acc_polys, cudaStream_t st, u32 acc_period, bool acc_by_poly) {
    const size_t N = c->N, Ql = L->h.size_Ql, P = c->size_P;
    CU(c->ws_delta.alloc(size_t(batch) * Ql * N * 8));
    u64 *delta = c->ws_delta.as<u64>();
    LimbMap mi = c->limb_map(Ql, P, Ql, batch, cx_stride);
    if (P == 1) CU(ntt_inverse(c->T, cx, cx, mi, nullptr, nullptr, st));
    else CU(ntt_inverse(c->T, cx, cx, mi, L->moddown_scale.as<u64>(), L->moddown_scale_shoup.as<u64>(), st));
    LAUNCHED(2);
    if (int rc = bconv_down(c, L, delta, Ql * N, cx + Ql * N, cx_stride, batch, st, /*lazy: delta goes into the forward transform*/ true)) return rc;
    LimbMap mf = c->limb_map(0, Ql, 0, batch, Ql * N);
    FwdEpilogue e{};
    e.other = cx;
    e.other_poly_stride = cx_stride;
    e.acc = acc;
    e.acc_poly_stride = acc_stride;
    e.acc_polys = acc_polys;
    e.acc_period = acc_period;
    e.acc_by_poly = acc_by_poly ? 1u : 0u;
    e.scale = L->bigPInv.as<u64>();
    e.scale_shoup = L->bigPInv_shoup.as<u64>();
    e.out = out;
    e.out_poly_stride = out_stride;
    CU(ntt_forward(c->T, delta, delta, mf, &e, st));
    LAUNCHED(2);
    return B200FHE_OK;
}

void Context::read_overlap_env() {
    if (overlap_env_read) return;
    if (const char *v = getenv("B200FHE_OVERLAP")) overlap = atoi(v);
    overlap_env_read = true;
}

int Context::ensure_lanes() {
    if (lane_fork) return B200FHE_OK;
    read_overlap_env();
    int least = 0, greatest = 0;
    CU(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    for (int i = 0; i < kLanes; i++) {
        CU(cudaStreamCreateWithPriority(&lane[i], cudaStreamNonBlocking, overlap ? greatest : least));
        CU(cudaEventCreateWithFlags(&lane_join[i], cudaEventDisableTiming));
    }
    CU(cudaStreamCreateWithPriority(&bg, cudaStreamNonBlocking, least));
    CU(cudaEventCreateWithFlags(&bg_start, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&bg_done, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ipq_done, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&lane_fork, cudaEventDisableTiming));
    return B200FHE_OK;
}

// One key switch with its independent chains on side streams: digit d of the modup = inverse transform of its own limbs,
// conversion into the other limbs (+ copy of the own ones), forward transform of those; polynomial p of the moddown =
// inverse transform of its P limbs, conversion, forward transform with the (cx - .) P^-1 (+ acc) epilogue.  Same kernels,
// same numbers; only the order in which independent launches reach the GPU changes.
static int keyswitch_lanes_impl(Context *c, DeviceLevel *L, u64 *out, const u64 *acc, u32 acc_polys, const u64 *c2,
                                size_t c2_stride, u32 batch, const EvkPtrs &e, cudaStream_t st);
static int keyswitch_lanes(Context *c, DeviceLevel *L, u64 *out, const u64 *acc, u32 acc_polys, const u64 *c2,
                           size_t c2_stride, u32 batch, const EvkPtrs &e, cudaStream_t st) {
    pdl_suppress(true);   // launch.cuh: no early placement between kernels that run beside each other
    const int rc = keyswitch_lanes_impl(c, L, out, acc, acc_polys, c2, c2_stride, batch, e, st);
    pdl_suppress(false);
    return rc;
}
static int keyswitch_lanes_impl(Context *c, DeviceLevel *L, u64 *out, const u64 *acc, u32 acc_polys, const u64 *c2,
                                size_t c2_stride, u32 batch, const EvkPtrs &e, cudaStream_t st) {
    const size_t N = c->N, Ql = L->h.size_Ql, P = c->size_P, QlP = Ql + P, alpha = L->h.alpha, beta = L->h.beta;
    if (int rc = c->ensure_lanes()) return rc;
    u64 *t = c->ws_tcks.as<u64>(), *up = c->ws_up.as<u64>(), *cx = c->ws_cx.as<u64>(), *delta = c->ws_delta.as<u64>();
    const size_t up_stride = beta * QlP * N;   // ciphertext z: up + z * up_stride = [beta][QlP][N]
    // overlap bit 2: the row pass of the digits' forward transforms and the inner product are one kernel (the transformed
    // digits never reach HBM; ntt.cuh FwdInnerProd); the lanes then stop after the column pass
    const bool fused = (c->overlap & 4) && beta <= (size_t) kIpDigits && c->logN >= 12 && !e.dev_table;
    FwdInnerProd fa{};
    if (fused) {
        fa.up = up;
        fa.own = c2;
        fa.cx = cx;
        for (size_t d = 0; d < beta; d++) fa.key[d] = e.d[d];
        fa.up_stride = up_stride;
        fa.own_stride = batch > 1 ? c2_stride : Ql * N;
        fa.beta = (u32) beta;
        fa.alpha = (u32) alpha;
        fa.size_Ql = (u32) Ql;
        fa.size_QlP = (u32) QlP;
        fa.size_Q = (u32) c->size_Q;
        fa.size_QP = (u32) c->size_QP;
        fa.batch = batch;
    }
    CU(cudaEventRecord(c->lane_fork, st));
    for (size_t d = 0; d < beta; d++) {
        cudaStream_t ls = c->lane[d];
        const size_t s = alpha * d, ni = L->h.digit_size(d);
        CU(cudaStreamWaitEvent(ls, c->lane_fork, 0));
        LimbMap mi = c->limb_map(s, ni, 0, batch, Ql * N);
        mi.src_poly_stride = batch > 1 ? c2_stride : Ql * N;
        CU(ntt_inverse(c->T, t, c2, mi, L->partQlHatInv.as<u64>(), L->partQlHatInv_shoup.as<u64>(), ls));
        BconvMultiArgs M{};
        BconvArgs &a = M.d[0];
        a.in = t + s * N;
        a.out = up + d * QlP * N;
        a.mat = L->modup_mat[d]->as<u64>();
        a.ni = (u32) ni;
        a.no = (u32) (QlP - ni);
        a.skip_start = (u32) s;
        a.skip_len = (u32) ni;
        a.pm = c->prime_map(Ql);
        a.in_stride = Ql * N;
        a.out_stride = up_stride;
        a.batch = batch;
        a.lazy = c->lazy_conv ? 1u : 0u;
        M.own_src[0] = fused ? nullptr : c2 + s * N;   // the fused kernel reads the digit's own limbs from c2 itself
        M.own_stride = batch > 1 ? c2_stride : Ql * N;
        M.ndig = 1;
        CU(bconv_multi(c->T.mods, M, N, ls));
        LimbMap mf = c->limb_map(0, QlP, Ql, batch, up_stride);
        mf.excl_lo = (u32) s;
        mf.excl_hi = (u32) (s + ni);
        if (fused) CU(ntt_forward_p1(c->T, up + d * QlP * N, up + d * QlP * N, mf, ls));
        else CU(ntt_forward(c->T, up + d * QlP * N, up + d * QlP * N, mf, nullptr, ls));
        CU(cudaEventRecord(c->lane_join[d], ls));
        CU(cudaStreamWaitEvent(st, c->lane_join[d], 0));
        LAUNCHED(5);
    }
    // The moddown chains need only the P rows of the inner product until their last step; with overlap on, those rows are
    // their own launch and the Q_l rows (HBM-bound: the keys) stream in under the chains' transforms and conversions.
    const bool split = c->overlap & 2;
    if (fused) {
        CU(ntt_forward_p2_inner_prod(c->T, fa, (u32) Ql, (u32) P, st));
        CU(cudaEventRecord(c->lane_fork, st));
        CU(ntt_forward_p2_inner_prod(c->T, fa, 0, (u32) Ql, st));
        if (split) CU(cudaEventRecord(c->ipq_done, st));
        else CU(cudaEventRecord(c->lane_fork, st));
        LAUNCHED(2);
    } else if (split) {
        CU(inner_prod_range(c->T.mods, cx, up, e, (u32) beta, (u32) QlP, (u32) c->size_QP, c->prime_map(Ql), (u32) Ql, (u32) P, N, st, 0, 0,
                            batch));
        CU(cudaEventRecord(c->lane_fork, st));
        CU(inner_prod_range(c->T.mods, cx, up, e, (u32) beta, (u32) QlP, (u32) c->size_QP, c->prime_map(Ql), 0, (u32) Ql, N, st, 0, 0,
                            batch));
        CU(cudaEventRecord(c->ipq_done, st));
        LAUNCHED(2);
    } else {
        if (int rc = do_inner_prod(c, L, cx, up, e, st, batch)) return rc;
        CU(cudaEventRecord(c->lane_fork, st));
    }
    for (u32 j = 0; j < 2; j++) {   // polynomial j of every ciphertext of the batch: stride two polynomials
        cudaStream_t ls = c->lane[j];
        CU(cudaStreamWaitEvent(ls, c->lane_fork, 0));
        u64 *cz = cx + j * QlP * N, *dz = delta + j * Ql * N;
        LimbMap mi = c->limb_map(Ql, P, Ql, batch, 2 * QlP * N);
        CU(ntt_inverse(c->T, cz, cz, mi, L->moddown_scale.as<u64>(), L->moddown_scale_shoup.as<u64>(), ls));
        BconvArgs a{};
        a.in = cz + Ql * N;
        a.out = dz;
        a.mat = L->PHat_mod_q.as<u64>();
        a.ni = (u32) P;
        a.no = (u32) Ql;
        a.skip_start = a.no;
        a.skip_len = 0;
        a.pm = PrimeMap{0, 0};
        a.in_stride = 2 * QlP * N;
        a.out_stride = 2 * Ql * N;
        a.batch = batch;
        a.lazy = c->lazy_conv ? 1u : 0u;
        CU(bconv(c->T.mods, a, N, ls));
        LimbMap mf = c->limb_map(0, Ql, 0, batch, 2 * Ql * N);
        FwdEpilogue ep{};
        ep.other = cz;
        ep.other_poly_stride = 2 * QlP * N;
        const bool with_acc = acc && (acc_polys == 0 || j < acc_polys);
        ep.acc = with_acc ? acc + j * Ql * N : nullptr;
        ep.acc_poly_stride = 2 * Ql * N;
        ep.scale = L->bigPInv.as<u64>();
        ep.scale_shoup = L->bigPInv_shoup.as<u64>();
        ep.out = out + j * Ql * N;
        ep.out_poly_stride = 2 * Ql * N;
        if (split) CU(cudaStreamWaitEvent(ls, c->ipq_done, 0));
        if (with_acc && c->acc_wait) CU(cudaStreamWaitEvent(ls, c->acc_wait, 0));
        CU(ntt_forward(c->T, dz, dz, mf, &ep, ls));
        CU(cudaEventRecord(c->lane_join[j], ls));
        CU(cudaStreamWaitEvent(st, c->lane_join[j], 0));
        LAUNCHED(5);
    }
    return B200FHE_OK;
}

// modup of `batch` polynomials and the inner product with the key, on one stream: cx [batch][2][QlP][N].  With overlap bit 2
// and digits of 2 .. 16 limbs: inverse transform, ONE conversion launch for all digits (no copies of the own limbs), column
// pass, then the fused row pass + inner product (ntt.cuh FwdInnerProd) — the transformed digits never reach HBM, which is
// where a large batch, HBM-bound in the inner product, gains.  Otherwise do_modup_batch + do_inner_prod.
static int modup_inner_prod(Context *c, DeviceLevel *L, u64 *cx, u64 *up, const u64 *c2, size_t c2_stride, u32 batch,
                            const EvkPtrs &e, cudaStream_t st) {
    const size_t N = c->N, Ql = L->h.size_Ql, P = c->size_P, QlP = Ql + P, alpha = L->h.alpha, beta = L->h.beta;
    c->read_overlap_env();
    const bool fused = (c->overlap & 4) && alpha >= 2 && alpha <= 16 && beta >= 2 && beta <= (size_t) kIpDigits &&
                       beta <= (size_t) kBconvDigits && c->logN >= 12 && !e.dev_table;
    if (!fused) {
        if (int rc = do_modup_batch(c, L, up, c2, c2_stride, batch, st)) return rc;
        return do_inner_prod(c, L, cx, up, e, st, batch);
    }
    CU(c->ws_tcks.alloc(size_t(batch) * Ql * N * 8));
    u64 *t = c->ws_tcks.as<u64>();
    const size_t up_stride = beta * QlP * N;
    LimbMap mi = c->limb_map(0, Ql, 0, batch, Ql * N);
    mi.src_poly_stride = c2_stride;
    CU(ntt_inverse(c->T, t, c2, mi, L->partQlHatInv.as<u64>(), L->partQlHatInv_shoup.as<u64>(), st));
    BconvMultiArgs M{};
    for (size_t d = 0; d < beta; d++) {
        const size_t s = alpha * d, ni = L->h.digit_size(d);
        BconvArgs &a = M.d[d];
        a.in = t + s * N;
        a.out = up + d * QlP * N;
        a.mat = L->modup_mat[d]->as<u64>();
        a.ni = (u32) ni;
        a.no = (u32) (QlP - ni);
        a.skip_start = (u32) s;
        a.skip_len = (u32) ni;
        a.pm = c->prime_map(Ql);
        a.in_stride = Ql * N;
        a.out_stride = up_stride;
        a.batch = batch;
        a.lazy = c->lazy_conv ? 1u : 0u;
        M.own_src[d] = nullptr;
    }
    M.ndig = (u32) beta;
    M.own_stride = c2_stride;
    CU(bconv_multi(c->T.mods, M, N, st));
    LimbMap mf = c->limb_map(0, QlP, Ql, batch * beta, QlP * N);
    mf.excl_alpha = (u32) alpha;
    mf.excl_mod = (u32) beta;
    CU(ntt_forward_p1(c->T, up, up, mf, st));
    FwdInnerProd fa{};
    fa.up = up;
    fa.own = c2;
    fa.cx = cx;
    for (size_t d = 0; d < beta; d++) fa.key[d] = e.d[d];
    fa.up_stride = up_stride;
    fa.own_stride = c2_stride;
    fa.beta = (u32) beta;
    fa.alpha = (u32) alpha;
    fa.size_Ql = (u32) Ql;
    fa.size_QlP = (u32) QlP;
    fa.size_Q = (u32) c->size_Q;
    fa.size_QP = (u32) c->size_QP;
    fa.batch = batch;
    CU(ntt_forward_p2_inner_prod(c->T, fa, (u32) Ql, (u32) P, st));
    CU(ntt_forward_p2_inner_prod(c->T, fa, 0, (u32) Ql, st));
    LAUNCHED(6 + (c->logN > 16 ? 2 : 0));
    return B200FHE_OK;
}

static bool lanes_apply(const Context *c, const DeviceLevel *L) {
    return c->ks_lanes && L->h.alpha >= 2 && L->h.alpha <= 16 && c->size_P >= 2 && L->h.beta >= 2 && L->h.beta <= (size_t) Context::kLanes;
}

// out[2][Ql][N] = acc (first acc_polys polys) + KeySwitch(c2)
int b200fhe::do_keyswitch(Context *c, DeviceLevel *L, u64 *out, const u64 *acc, u32 acc_polys, const u64 *c2,
                        const EvkPtrs &e, cudaStream_t st) {
    const size_t N = c->N, Ql = L->h.size_Ql, QlP = Ql + c->size_P;
    CU(c->ws_up.alloc(L->h.beta * QlP * N * 8));
    CU(c->ws_cx.alloc(2 * QlP * N * 8));
    if (lanes_apply(c, L)) {
        CU(c->ws_tcks.alloc(Ql * N * 8));
        CU(c->ws_delta.alloc(2 * Ql * N * 8));
        return keyswitch_lanes(c, L, out, acc, acc_polys, c2, Ql * N, 1, e, st);
    }
    if (int rc = modup_inner_prod(c, L, c->ws_cx.as<u64>(), c->ws_up.as<u64>(), c2, Ql * N, 1, e, st)) return rc;
    return do_moddown(c, L, out, Ql * N, c->ws_cx.as<u64>(), QlP * N, 2, acc, Ql * N, acc_polys, st);
}

int b200fhe::do_hmult(Context *c, DeviceLevel *L, u64 *out, const u64 *a, const u64 *b, const EvkPtrs &e, u32 batch, cudaStream_t st) {
    const size_t N = c->N, Ql = L->h.size_Ql, LN = Ql * N;
    if (b == a) b = nullptr;
    CU(c->ws_tmp.alloc(3 * size_t(batch) * LN * 8));
    u64 *d01 = c->ws_tmp.as<u64>(), *d2 = d01 + 2 * size_t(batch) * LN;
    const bool lanes = lanes_apply(c, L) && batch <= (u32) c->ks_lanes_max_batch;
    if (lanes)
        if (int rc = c->ensure_lanes()) return rc;
    if (lanes && (c->overlap & 1)) {
        // d2 feeds the key switch; (d0, d1) are only added in by its last transforms and go to the background stream
        CU(cudaEventRecord(c->bg_start, st));
        CU(tensor_d2(c->T.mods, d2, a, b, (u32) Ql, batch, N, st));
        CU(cudaStreamWaitEvent(c->bg, c->bg_start, 0));
        CU(tensor_d01(c->T.mods, d01, a, b, (u32) Ql, batch, N, c->bg));
        CU(cudaEventRecord(c->bg_done, c->bg));
        c->acc_wait = c->bg_done;
        LAUNCHED(2);
    } else {
        CU(tensor_split(c->T.mods, d01, d2, a, b, (u32) Ql, batch, N, st));
        LAUNCHED(1);
    }
    const int rc = batch == 1 ? do_keyswitch(c, L, out, d01, 0, d2, e, st) : do_keyswitch_batch(c, L, out, d01, d2, LN, e, batch, st);
    c->acc_wait = nullptr;
    return rc;
}

// batch ciphertexts: out / acc [batch][2][Ql][N] (acc may be null), c2 + z*c2_stride
int b200fhe::do_keyswitch_batch(Context *c, DeviceLevel *L, u64 *out, const u64 *acc, const u64 *c2, size_t c2_stride,
                                const EvkPtrs &e, u32 batch, cudaStream_t st, u32 acc_polys) {
    const size_t N = c->N, Ql = L->h.size_Ql, QlP = Ql + c->size_P;
    CU(c->ws_up.alloc(size_t(batch) * L->h.beta * QlP * N * 8));
    CU(c->ws_cx.alloc(size_t(batch) * 2 * QlP * N * 8));
    if (lanes_apply(c, L) && batch <= (u32) c->ks_lanes_max_batch) {
        CU(c->ws_tcks.alloc(size_t(batch) * Ql * N * 8));
        CU(c->ws_delta.alloc(size_t(batch) * 2 * Ql * N * 8));
        return keyswitch_lanes(c, L, out, acc, acc_polys, c2, c2_stride, batch, e, st);
    }
    if (int rc = modup_inner_prod(c, L, c->ws_cx.as<u64>(), c->ws_up.as<u64>(), c2, c2_stride, batch, e, st)) return rc;
    return do_moddown(c, L, out, Ql * N, c->ws_cx.as<u64>(), QlP * N, 2 * batch, acc, Ql * N, acc_polys, st, acc_polys ? 2 : 0,
                      acc_polys != 0);
}

extern "C" int b200fhe_ctx_set_lanes(b200fhe_ctx *ctx_, int on) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    if (!c) return fail(B200FHE_ERR_ARG, "null context");
    c->ks_lanes = on ? 1 : 0;
    return B200FHE_OK;
}

extern "C" int b200fhe_modup(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *dst, const uint64_t *c2, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    DeviceLevel *L;
    if (int rc = get_level(c, size_Ql, &L)) return rc;
    if (!dst || !c2) return fail(B200FHE_ERR_ARG, "null argument");
    return do_modup(c, L, (u64 *) dst, (const u64 *) c2, S(stream));
}

extern "C" int b200fhe_inner_prod(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *cx, const uint64_t *t_mod_up,
                                  const uint64_t *const *evk, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    DeviceLevel *L;
    if (int rc = get_level(c, size_Ql, &L)) return rc;
    if (!cx || !t_mod_up) return fail(B200FHE_ERR_ARG, "null argument");
    EvkPtrs e{};
    if (int rc = make_evk(c, L, evk, e)) return rc;
    return do_inner_prod(c, L, (u64 *) cx, (const u64 *) t_mod_up, e, S(stream));
}

extern "C" int b200fhe_moddown_from_ntt(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *ct, uint64_t *cx, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    DeviceLevel *L;
    if (int rc = get_level(c, size_Ql, &L)) return rc;
    if (!ct || !cx) return fail(B200FHE_ERR_ARG, "null argument");
    return do_moddown(c, L, (u64 *) ct, 0, (u64 *) cx, 0, 1, nullptr, 0, 0, S(stream));
}

extern "C" int b200fhe_keyswitch(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *ct, const uint64_t *c2,
                                 const uint64_t *const *evk, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    DeviceLevel *L;
    if (int rc = get_level(c, size_Ql, &L)) return rc;
    if (!ct || !c2) return fail(B200FHE_ERR_ARG, "null argument");
    EvkPtrs e{};
    if (int rc = make_evk(c, L, evk, e)) return rc;
    return do_keyswitch(c, L, (u64 *) ct, (const u64 *) ct, 0, (const u64 *) c2, e, S(stream));
}

// ------------------------------------------------------------------------------------------- evaluator
extern "C" int b200fhe_tensor(b200fhe_ctx *ctx_, size_t L, uint64_t *out, const uint64_t *a, const uint64_t *b,
                              void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    if (L == 0 || L > c->size_Q || !out || !a || !b) return fail(B200FHE_ERR_ARG, "bad argument");
    if (a == b) CU(tensor_square(c->T.mods, (u64 *) out, (const u64 *) a, (u32) L, c->N, S(stream)));
    else CU(tensor_prod(c->T.mods, (u64 *) out, (const u64 *) a, (const u64 *) b, (u32) L, c->N, S(stream)));
    LAUNCHED(1);
    return B200FHE_OK;
}

extern "C" int b200fhe_hmult_relin(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *out, const uint64_t *a,
                                   const uint64_t *b, const uint64_t *const *rlk, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    DeviceLevel *L;
    if (int rc = get_level(c, size_Ql, &L)) return rc;
    if (!out || !a || !b) return fail(B200FHE_ERR_ARG, "null argument");
    EvkPtrs e{};
    if (int rc = make_evk(c, L, rlk, e)) return rc;
    return do_hmult(c, L, (u64 *) out, (const u64 *) a, (const u64 *) b, e, 1, S(stream));
}

int b200fhe::do_rescale(Context *c, size_t size_Ql, u64 *dst, const u64 *src, size_t polys, cudaStream_t st) {
    if (size_Ql < 2 || size_Ql > c->size_Q) return fail(B200FHE_ERR_ARG, "rescale needs 2 <= size_Ql <= |Q|");
    if (c->logN < 12 || c->logN > 17) return fail(B200FHE_ERR_UNSUPPORTED, "kernels support N = 2^12 .. 2^17");
    if (!dst || !src || polys == 0) return fail(B200FHE_ERR_ARG, "bad argument");
    if (!c->size_P) return fail(B200FHE_ERR_UNSUPPORTED, "rescale without special primes is not wired yet");
    const size_t N = c->N, Lq = size_Ql, Ln = Lq - 1;
    DeviceLevel *L;
    if (int rc = get_level(c, size_Ql, &L)) return rc;
    CU(c->ws_tmp.alloc((polys * Ln + polys) * N * 8));
    u64 *last = c->ws_tmp.as<u64>();          // [polys][N]
    u64 *red = last + polys * N;              // [polys][Ln][N]
    LimbMap mi = c->limb_map(Ln, 1, 0, polys, N);
    mi.src_poly_stride = Lq * N;
    CU(ntt_inverse(c->T, last - Ln * N, src, mi, nullptr, nullptr, st));
    LAUNCHED(2);
    // The reference reduces the last limb modulo every remaining prime before transforming it
    // (divide_and_round_reduce_q_last_kernel).  The lazy butterflies take any input below 4q, so when
    // q_last < 4 q_i for every i the forward transform reads the unreduced row directly (same canonical result)
    // and the reduce pass with its Ln*N write and re-read disappears.
    bool direct = true;
    for (size_t i = 0; i < Ln; i++) direct = direct && (c->primes[Ln] >> 2) < c->primes[i];
    LimbMap mf = c->limb_map(0, Ln, 0, polys, Ln * N);
    const u64 *fsrc = red;
    if (direct) {
        mf.src_bcast = 1;
        mf.src_poly_stride = N;
        fsrc = last;
    } else {
        CU(reduce_last(c->T.mods, red, last, (u32) Ln, (u32) polys, N, Ln * N, N, st));
        LAUNCHED(1);
    }
    FwdEpilogue e{};
    e.other = src;
    e.other_poly_stride = Lq * N;
    e.scale = L->inv_q_last.as<u64>();
    e.scale_shoup = L->inv_q_last_shoup.as<u64>();
    e.out = dst;
    e.out_poly_stride = Ln * N;
    CU(ntt_forward(c->T, red, fsrc, mf, &e, st));
    LAUNCHED(2);
    return B200FHE_OK;
}

extern "C" int b200fhe_rescale(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *dst, const uint64_t *src, size_t polys,
                               void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    return do_rescale(c, size_Ql, (u64 *) dst, (const u64 *) src, polys, S(stream));
}

extern "C" int b200fhe_automap_table(b200fhe_ctx *ctx_, uint32_t k, uint32_t *perm, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    if (!perm || !(k & 1)) return fail(B200FHE_ERR_ARG, "automorphism index must be odd");
    CU(automap_table(perm, c->logN, k, S(stream)));
    LAUNCHED(1);
    return B200FHE_OK;
}

extern "C" int b200fhe_apply_galois_ntt(b200fhe_ctx *ctx_, uint64_t *dst, const uint64_t *src, const uint32_t *perm,
                                        size_t limbs, size_t npoly, size_t poly_stride, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    if (!dst || !src || !perm || dst == src) return fail(B200FHE_ERR_ARG, "bad argument (dst must not alias src)");
    CU(gather_limbs((u64 *) dst, (const u64 *) src, perm, (u32) limbs, (u32) npoly, poly_stride, poly_stride, c->N,
                    S(stream)));
    LAUNCHED(1);
    return B200FHE_OK;
}

extern "C" int b200fhe_rotate_fused(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *out, const uint64_t *ct, uint32_t k,
                                    const uint64_t *const *glk, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    DeviceLevel *L;
    if (int rc = get_level(c, size_Ql, &L)) return rc;
    if (!out || !ct || out == ct || !(k & 1)) return fail(B200FHE_ERR_ARG, "bad argument");
    EvkPtrs e{};
    if (int rc = make_evk(c, L, glk, e)) return rc;
    const size_t LN = size_Ql * c->N;
    cudaStream_t st = S(stream);
    CU(c->ws_tmp.alloc(2 * LN * 8));
    u64 *tmp = c->ws_tmp.as<u64>();
    // (c0 + ks0, ks1): the addend exists only for polynomial 0 (the reference zeroes c1 first)
    if (int rc = do_keyswitch(c, L, tmp, (const u64 *) ct, 1, (const u64 *) ct + LN, e, st)) return rc;
    // the permutation of automorphism k: built once per index and kept (outside a stream capture; inside one it is rebuilt
    // in the workspace so that the graph owns every launch it replays)
    const u32 *perm = nullptr;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    CU(cudaStreamIsCapturing(st, &cap));
    if (cap == cudaStreamCaptureStatusNone) {
        std::lock_guard<std::mutex> g(c->mu);
        auto &slot = c->perm_cache[k];
        if (!slot) {
            auto buf = std::make_unique<DevBuf>();
            CU(buf->alloc(c->N * sizeof(u32)));
            CU(automap_table(buf->as<u32>(), c->logN, k, st));
            CU(cudaStreamSynchronize(st));
            LAUNCHED(1);
            slot = std::move(buf);
        }
        perm = slot->as<u32>();
    } else {
        CU(automap_table(c->ws_perm.as<u32>(), c->logN, k, st));
        LAUNCHED(1);
        perm = c->ws_perm.as<u32>();
    }
    CU(gather_limbs((u64 *) out, tmp, perm, (u32) size_Ql, 2, LN, LN, c->N, st));
    LAUNCHED(1);
    return B200FHE_OK;
}

extern "C" int b200fhe_eltwise(b200fhe_ctx *ctx_, int op, uint64_t *out, const uint64_t *a, const uint64_t *b,
                               size_t limbs, size_t size_Ql, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    NEED_DEV(c);
    if (op < 0 || op > 3 || !out || !a || (op != 3 && !b)) return fail(B200FHE_ERR_ARG, "bad argument");
    if (int rc = check_limbs(c, limbs, 0, size_Ql)) return rc;
    CU(eltwise(c->T.mods, (EltOp) op, (u64 *) out, (const u64 *) a, (const u64 *) b, (u32) limbs,
               PrimeMap{(u32) size_Ql, (u32) c->size_Q}, c->N, S(stream)));
    LAUNCHED(1);
    return B200FHE_OK;
}
