// capi_noalloc.cu — b200fhe_rescale_noalloc: the rescale (divide_and_round_q_last_ntt) for callers that cannot
// call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels and the
// same numbers as b200fhe_rescale; the only scratch the operation needs — the coefficient form of the last limb
// of every polynomial — lives in a file-scope __device__ array instead of the context's grown-on-demand workspace,
// and the forward transform runs in place in the caller's dst instead of through a workspace copy.
//
// No counterpart in the reference.
#include <mutex>

#include "../../include/b200fhe.h"
#include "context.h"

using namespace b200fhe;

// 2 polynomials of N = 2^16 coefficients (or 1 of 2^17): 1 MiB on every GPU that loads the library, held until exit.
static constexpr size_t kScratchWords = B200FHE_NOALLOC_MAX_WORDS;
static __device__ __align__(256) u64 g_rescale_scratch[kScratchWords];

// In host code `g_rescale_scratch` is a host-side stand-in, not a device address: the address the kernels get is
// looked up once per device, with that device current.
static constexpr int kMaxDevices = 64;
static std::mutex g_mu;
static u64 *g_scratch_addr[kMaxDevices] = {};

static int scratch_for_current_device(int dev, u64 **out) {
    std::lock_guard<std::mutex> g(g_mu);
    if (!g_scratch_addr[dev]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_rescale_scratch);
        if (e != cudaSuccess)
            return fail(B200FHE_ERR_CUDA, "cudaGetSymbolAddress(g_rescale_scratch) on device %d: %s", dev, cudaGetErrorString(e));
        g_scratch_addr[dev] = static_cast<u64 *>(p);
    }
    *out = g_scratch_addr[dev];
    return B200FHE_OK;
}

#define CU(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) return fail(B200FHE_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

extern "C" int b200fhe_rescale_noalloc(b200fhe_ctx *ctx_, size_t size_Ql, uint64_t *dst, const uint64_t *src, size_t polys,
                                       void *stream) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    if (!c) return fail(B200FHE_ERR_ARG, "null context");
    if (c->host_only) return fail(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    if (size_Ql < 2 || size_Ql > c->size_Q) return fail(B200FHE_ERR_ARG, "rescale needs 2 <= size_Ql <= |Q|");
    if (c->logN < 12 || c->logN > 17) return fail(B200FHE_ERR_UNSUPPORTED, "kernels support N = 2^12 .. 2^17");
    if (!dst || !src || polys == 0 || dst == src) return fail(B200FHE_ERR_ARG, "bad argument (dst must not alias src)");
    const size_t N = c->N, Lq = size_Ql, Ln = Lq - 1;
    if (polys * N > kScratchWords)
        return fail(B200FHE_ERR_ARG,
                    "b200fhe_rescale_noalloc: polys * N = %zu words exceeds the static scratch of %zu words "
                    "(2 polynomials at N = 2^16); use b200fhe_rescale",
                    polys * N, kScratchWords);
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev != c->device || dev < 0 || dev >= kMaxDevices)
        return fail(B200FHE_ERR_ARG, "current device %d is not the context's device %d", dev, c->device);
    // per-level constants must already be resident (b200fhe_ctx_create builds every level when |P| > 0): never built here
    DeviceLevel *L = nullptr;
    {
        std::lock_guard<std::mutex> g(c->mu);
        auto it = c->levels.find(size_Ql);
        if (it != c->levels.end()) L = it->second.get();
    }
    if (!L || !L->inv_q_last.p)
        return fail(B200FHE_ERR_UNSUPPORTED, "level %zu constants are not resident; this entry point never allocates", size_Ql);
    // the forward transform reads the unreduced last-limb row for every remaining prime: needs q_last < 4 q_i
    for (size_t i = 0; i < Ln; i++)
        if ((c->primes[Ln] >> 2) >= c->primes[i])
            return fail(B200FHE_ERR_UNSUPPORTED, "q_last >= 4 q_%zu: needs the reduce pass and its workspace; use b200fhe_rescale", i);
    u64 *last = nullptr;  // [polys][N], every word written by the inverse transform before it is read
    if (int rc = scratch_for_current_device(dev, &last)) return rc;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);

    LimbMap mi = c->limb_map(Ln, 1, 0, polys, N);
    mi.src_poly_stride = Lq * N;
    CU(ntt_inverse(c->T, last - Ln * N, (const u64 *) src, mi, nullptr, nullptr, st));
    LimbMap mf = c->limb_map(0, Ln, 0, polys, Ln * N);
    mf.src_bcast = 1;
    mf.src_poly_stride = N;
    FwdEpilogue e{};
    e.other = (const u64 *) src;
    e.other_poly_stride = Lq * N;
    e.scale = L->inv_q_last.as<u64>();
    e.scale_shoup = L->inv_q_last_shoup.as<u64>();
    e.out = (u64 *) dst;  // pass 1 writes dst, pass 2 reads each 256-word row into registers and writes the same row
    e.out_poly_stride = Ln * N;
    CU(ntt_forward(c->T, (u64 *) dst, last, mf, &e, st));
    count_launches(ntt_launch_count_inverse(c->logN) + ntt_launch_count_forward(c->logN));
    return B200FHE_OK;
}
