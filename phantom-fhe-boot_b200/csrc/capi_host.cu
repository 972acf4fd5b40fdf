// capi_host.cu — the op trace (b200fhe_trace_*) and the host-buffer forms of the composite operations: ciphertext limbs
// in HOST memory -> copy in, compute, copy out, all inside the call (the reference-facing end-to-end path that bench.py
// times as `e2e`).  Keys stay device-resident, as in the reference (PhantomRelinKey / PhantomGaloisKeyFused live in HBM).
#include <algorithm>
#include <atomic>
#include <string>

#include "../../include/b200fhe.h"
#include "bootstrapper.h"
#include "context.h"
#include "evaluator.h"
#include "launch.cuh"

using namespace b200fhe;

#define CU(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) return fail(B200FHE_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

// ------------------------------------------------------------------------------------------- op trace
// doubles accumulated as fixed point (bytes and operation counts are integers far below 2^63 per bench run)
static std::atomic<uint64_t> g_bytes{0}, g_bfly{0}, g_macs{0}, g_calls{0};
void b200fhe::trace_add(double hbm_bytes, double butterflies, double macs) {
    g_bytes.fetch_add((uint64_t) hbm_bytes, std::memory_order_relaxed);
    g_bfly.fetch_add((uint64_t) butterflies, std::memory_order_relaxed);
    g_macs.fetch_add((uint64_t) macs, std::memory_order_relaxed);
    g_calls.fetch_add(1, std::memory_order_relaxed);
}
extern "C" void b200fhe_trace_reset(void) {
    g_bytes = 0;
    g_bfly = 0;
    g_macs = 0;
    g_calls = 0;
}
extern "C" void b200fhe_trace_read(uint64_t out4[4]) {
    out4[0] = g_bytes.load();
    out4[1] = g_bfly.load();
    out4[2] = g_macs.load();
    out4[3] = g_calls.load();
}

// programmatic dependent launch (launch.cuh)
unsigned b200fhe::pdl_min_blocks() {
    static const unsigned n = [] {
        const char *off = getenv("B200FHE_PDL");
        if (off && off[0] == '0') return 0u;
        const char *v = getenv("B200FHE_PDL_MIN");
        return v ? (unsigned) std::max(1, atoi(v)) : 1u;
    }();
    return n;
}
extern "C" int b200fhe_tune_cluster_ntt(int max_limb_polys) {
    b200fhe::ntt_set_cluster_max_limbs(max_limb_polys);
    return B200FHE_OK;
}
static thread_local int g_pdl_suppress = 0;
void b200fhe::pdl_suppress(bool on) { g_pdl_suppress += on ? 1 : -1; }
bool b200fhe::pdl_suppressed() { return g_pdl_suppress > 0; }

// ------------------------------------------------------------------------------------------- host-buffer composites
static inline cudaStream_t S(void *s) { return reinterpret_cast<cudaStream_t>(s); }

extern "C" int b200fhe_hmult_relin_host(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *out_host, const uint64_t *a_host,
                                        const uint64_t *b_host, const uint64_t *const *rlk_digits, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    if (!c || c->host_only) return fail(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    if (!out_host || !a_host || !b_host) return fail(B200FHE_ERR_ARG, "null argument");
    if (size_Ql == 0 || size_Ql > c->size_Q) return fail(B200FHE_ERR_ARG, "size_Ql out of range");
    const size_t bytes = 2 * size_Ql * c->N * 8;
    for (int i = 0; i < 3; i++) CU(c->stage[i].alloc(bytes));
    uint64_t *da = c->stage[0].as<uint64_t>(), *db = c->stage[1].as<uint64_t>(), *dout = c->stage[2].as<uint64_t>();
    cudaStream_t st = S(stream);
    CU(cudaMemcpyAsync(da, a_host, bytes, cudaMemcpyHostToDevice, st));
    if (b_host != a_host) CU(cudaMemcpyAsync(db, b_host, bytes, cudaMemcpyHostToDevice, st));
    if (int rc = b200fhe_hmult_relin(ctx_, size_Ql, dout, da, b_host == a_host ? da : db, rlk_digits, stream)) return rc;
    CU(cudaMemcpyAsync(out_host, dout, bytes, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return B200FHE_OK;
}

extern "C" int b200fhe_rotate_fused_host(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *out_host, const uint64_t *ct_host, uint32_t k,
                                         const uint64_t *const *glk_digits, void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    if (!c || c->host_only) return fail(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    if (!out_host || !ct_host) return fail(B200FHE_ERR_ARG, "null argument");
    if (size_Ql == 0 || size_Ql > c->size_Q) return fail(B200FHE_ERR_ARG, "size_Ql out of range");
    const size_t bytes = 2 * size_Ql * c->N * 8;
    for (int i = 0; i < 2; i++) CU(c->stage[i].alloc(bytes));
    uint64_t *din = c->stage[0].as<uint64_t>(), *dout = c->stage[1].as<uint64_t>();
    cudaStream_t st = S(stream);
    CU(cudaMemcpyAsync(din, ct_host, bytes, cudaMemcpyHostToDevice, st));
    if (int rc = b200fhe_rotate_fused(ctx_, size_Ql, dout, din, k, glk_digits, stream)) return rc;
    CU(cudaMemcpyAsync(out_host, dout, bytes, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return B200FHE_OK;
}

// EvalBootstrap on a ciphertext held in host memory: [2][size_Ql][N] in, [2][*out_size_Ql][N] out (capacity in words)
extern "C" int b200fhe_boot_eval_host(b200fhe_boot *bt, const uint64_t *ct_host, size_t size_Ql, double scale,
                                      uint32_t noise_scale_deg, uint64_t *out_host, size_t capacity_words, b200fhe_ct_desc *out_desc) {
    if (!bt || !ct_host || !out_host) return fail(B200FHE_ERR_ARG, "null argument");
    b200fhe_eval *ev = b200fhe_boot_evaluator(bt);
    Evaluator *E = eval_of(ev);
    if (!E) return fail(B200FHE_ERR_ARG, "bootstrapper has no evaluator");
    Context *c = E->c;
    cudaStream_t st = E->st;
    const size_t in_bytes = 2 * size_Ql * c->N * 8;
    CU(c->stage[0].alloc(in_bytes));
    CU(cudaMemcpyAsync(c->stage[0].p, ct_host, in_bytes, cudaMemcpyHostToDevice, st));
    b200fhe_ct *in = nullptr, *out = nullptr;
    if (int rc = b200fhe_ct_import(ev, c->stage[0].as<uint64_t>(), size_Ql, 2, 0, scale, noise_scale_deg, &in)) return rc;
    int rc = b200fhe_boot_eval(bt, in, &out);
    b200fhe_ct_free(in);
    if (rc) return rc;
    b200fhe_ct_desc d{};
    b200fhe_ct_describe(out, &d);
    const size_t words = size_t(d.polys) * d.limbs * c->N;
    if (words > capacity_words) {
        b200fhe_ct_free(out);
        return fail(B200FHE_ERR_ARG, "output needs %zu words, capacity %zu", words, capacity_words);
    }
    cudaError_t e = c->stage[1].alloc(words * 8);
    if (e == cudaSuccess) {
        rc = b200fhe_ct_export(ev, out, c->stage[1].as<uint64_t>(), words);
        if (rc == B200FHE_OK) e = cudaMemcpyAsync(out_host, c->stage[1].p, words * 8, cudaMemcpyDeviceToHost, st);
    }
    b200fhe_ct_free(out);
    if (rc) return rc;
    CU(e);
    CU(cudaStreamSynchronize(st));
    if (out_desc) *out_desc = d;
    return B200FHE_OK;
}
