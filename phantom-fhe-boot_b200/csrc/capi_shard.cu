// capi_shard.cu — limb-sharded key switch (second mode): ONE ciphertext's key switch spread over G GPUs.
// NTT, inverse NTT, the evaluation-key inner product and the moddown epilogue are per-limb; only the two base
// conversions mix limbs.  Rank g owns the ciphertext limbs [q_lo, q_hi) of Q_l and the special limbs [p_lo, p_hi) of P
// and computes exactly those rows of every intermediate; the rows a base conversion needs from the other ranks are
// exchanged between the three phases below (the caller's all-gather over NVLink: the alpha-limb digits in coefficient
// form before the modup conversions, the P limbs of the two accumulators before the moddown conversion).
//
// Every buffer keeps the FULL single-GPU layout ([Ql][N], [2][QlP][N] ...) so that a row has the same address
// arithmetic on every rank; a rank touches only its rows (+ the rows the exchange filled in).  Each row is computed by
// the same kernels with the same operands as in do_keyswitch, hence the same bits.  The reference has no multi-GPU path.
#include "../../include/b200fhe.h"
#include "context.h"

using namespace b200fhe;

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(B200FHE_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

namespace {
struct Shard {
    Context *c;
    DeviceLevel *L;
    size_t N, Ql, P, QlP, alpha, beta, q_lo, q_hi, p_lo, p_hi;
    cudaStream_t st;
};
int open_shard(b200fhe_ctx *ctx_, size_t size_Ql, size_t q_lo, size_t q_hi, size_t p_lo, size_t p_hi, void *stream, Shard &s) {
    auto *c = reinterpret_cast<Context *>(ctx_);
    if (!c) return fail(B200FHE_ERR_ARG, "null context");
    if (c->host_only) return fail(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    DeviceLevel *L = nullptr;
    if (int rc = get_level(c, size_Ql, &L)) return rc;
    if (q_lo > q_hi || q_hi > size_Ql || p_lo > p_hi || p_hi > c->size_P) return fail(B200FHE_ERR_ARG, "limb range out of bounds");
    if (L->h.alpha < 2 || c->size_P < 2) return fail(B200FHE_ERR_UNSUPPORTED, "limb sharding needs digits of at least 2 limbs");
    s = Shard{c, L, c->N, size_Ql, c->size_P, size_Ql + c->size_P, L->h.alpha, L->h.beta, q_lo, q_hi, p_lo, p_hi,
              reinterpret_cast<cudaStream_t>(stream)};
    return B200FHE_OK;
}
}  // namespace

// Phase 1: own rows of tcks = INTT(c2) x (Q_hat_d^-1 mod q_t)   (DRNSTool::modup prologue, src/rns_bconv.cu:544-560)
extern "C" int b200fhe_shard_ks_begin(b200fhe_ctx *ctx, size_t size_Ql, size_t q_lo, size_t q_hi, const uint64_t *c2,
                                      uint64_t *tcks, void *stream) {
    Shard s{};
    if (int rc = open_shard(ctx, size_Ql, q_lo, q_hi, 0, 0, stream, s)) return rc;
    if (!c2 || !tcks) return fail(B200FHE_ERR_ARG, "null buffer");
    if (q_hi == q_lo) return B200FHE_OK;
    LimbMap mi = s.c->limb_map(q_lo, q_hi - q_lo, 0, 1, s.Ql * s.N);
    CU(ntt_inverse(s.c->T, (u64 *) tcks, (const u64 *) c2, mi, s.L->partQlHatInv.as<u64>(), s.L->partQlHatInv_shoup.as<u64>(), s.st));
    count_launches(2);
    return B200FHE_OK;
}

// Phase 2 (tcks complete): per digit the base conversion into the OWN rows of Q_l P, their NTT, the copy of the own rows
// of the digit itself; inner product with the key on the own rows; inverse NTT x P_hat^-1 of the own P rows of cx.
extern "C" int b200fhe_shard_ks_middle(b200fhe_ctx *ctx, size_t size_Ql, size_t q_lo, size_t q_hi, size_t p_lo, size_t p_hi,
                                       const uint64_t *tcks, const uint64_t *c2, const uint64_t *const *evk, uint64_t *cx,
                                       void *stream) {
    Shard s{};
    if (int rc = open_shard(ctx, size_Ql, q_lo, q_hi, p_lo, p_hi, stream, s)) return rc;
    if (!tcks || !c2 || !cx) return fail(B200FHE_ERR_ARG, "null buffer");
    EvkPtrs e{};
    if (int rc = make_evk(s.c, s.L, evk, e)) return rc;
    Context *c = s.c;
    const size_t N = s.N, Ql = s.Ql, QlP = s.QlP, nq = q_hi - q_lo, np = p_hi - p_lo;
    CU(c->ws_up.alloc(s.beta * QlP * N * 8));
    u64 *up = c->ws_up.as<u64>();
    for (size_t d = 0; d < s.beta; d++) {
        const size_t ds = s.alpha * d, ni = s.L->h.digit_size(d), de = ds + ni;
        auto below = [&](size_t p) { return p < ds ? p : std::max(p, de) - ni; };  // non-digit positions below p
        BconvArgs a{};
        a.in = (const u64 *) tcks + ds * N;
        a.out = up + d * QlP * N;
        a.mat = s.L->modup_mat[d]->as<u64>();
        a.ni = (u32) ni;
        a.skip_start = (u32) ds;
        a.skip_len = (u32) ni;
        a.pm = c->prime_map(Ql);
        a.in_stride = a.out_stride = 0;
        a.batch = 1;
        const size_t ranges[2][2] = {{below(q_lo), below(q_hi)}, {below(Ql + p_lo), below(Ql + p_hi)}};
        for (const auto &r : ranges) {
            if (r[1] <= r[0]) continue;
            a.j0 = (u32) r[0];
            a.no = (u32) (r[1] - r[0]);
            CU(bconv(c->T.mods, a, N, s.st));
            count_launches(1);
        }
        const size_t o0 = std::max(q_lo, ds), o1 = std::min(q_hi, de);  // own rows of the digit itself stay in NTT form
        if (o1 > o0)
            CU(cudaMemcpyAsync(up + d * QlP * N + o0 * N, (const u64 *) c2 + o0 * N, (o1 - o0) * N * 8, cudaMemcpyDeviceToDevice, s.st));
    }
    const size_t starts[2] = {q_lo, Ql + p_lo}, counts[2] = {nq, np};
    for (int k = 0; k < 2; k++) {
        if (!counts[k]) continue;
        LimbMap mf = c->limb_map(starts[k], counts[k], Ql, s.beta, QlP * N);
        mf.excl_alpha = (u32) s.alpha;
        mf.excl_mod = (u32) s.beta;
        CU(ntt_forward(c->T, up, up, mf, nullptr, s.st));
        CU(inner_prod_range(c->T.mods, (u64 *) cx, up, e, (u32) s.beta, (u32) QlP, (u32) c->size_QP, c->prime_map(Ql),
                            (u32) starts[k], (u32) counts[k], N, s.st));
        count_launches(3);
    }
    if (np) {
        LimbMap mi = c->limb_map(Ql + p_lo, np, Ql, 2, QlP * N);
        CU(ntt_inverse(c->T, (u64 *) cx, (const u64 *) cx, mi, s.L->moddown_scale.as<u64>(), s.L->moddown_scale_shoup.as<u64>(), s.st));
        count_launches(2);
    }
    return B200FHE_OK;
}

// Phase 3 (P rows of cx complete, coefficient form x P_hat^-1): conversion P -> own rows of Q_l, NTT with the
// (cx - .) x P^-1 (+ acc) epilogue   (DRNSTool::moddown_from_NTT, src/rns_bconv.cu:791-843 + add_to_ct_kernel)
extern "C" int b200fhe_shard_ks_end(b200fhe_ctx *ctx, size_t size_Ql, size_t q_lo, size_t q_hi, const uint64_t *cx,
                                    const uint64_t *acc, uint32_t acc_polys, uint64_t *out, void *stream) {
    Shard s{};
    if (int rc = open_shard(ctx, size_Ql, q_lo, q_hi, 0, 0, stream, s)) return rc;
    if (!cx || !out) return fail(B200FHE_ERR_ARG, "null buffer");
    if (q_hi == q_lo) return B200FHE_OK;
    Context *c = s.c;
    const size_t N = s.N, Ql = s.Ql, QlP = s.QlP, nq = q_hi - q_lo;
    CU(c->ws_delta.alloc(2 * Ql * N * 8));
    u64 *delta = c->ws_delta.as<u64>();
    BconvArgs a{};
    a.in = (const u64 *) cx + Ql * N;
    a.out = delta;
    a.mat = s.L->PHat_mod_q.as<u64>();
    a.ni = (u32) s.P;
    a.no = (u32) nq;
    a.j0 = (u32) q_lo;
    a.skip_start = (u32) Ql;
    a.skip_len = 0;
    a.pm = PrimeMap{0, 0};
    a.in_stride = QlP * N;
    a.out_stride = Ql * N;
    a.batch = 2;
    CU(bconv(c->T.mods, a, N, s.st));
    LimbMap mf = c->limb_map(q_lo, nq, 0, 2, Ql * N);
    FwdEpilogue e{};
    e.other = (const u64 *) cx;
    e.other_poly_stride = QlP * N;
    e.acc = (const u64 *) acc;
    e.acc_poly_stride = Ql * N;
    e.acc_polys = acc_polys;
    e.scale = s.L->bigPInv.as<u64>();
    e.scale_shoup = s.L->bigPInv_shoup.as<u64>();
    e.out = (u64 *) out;
    e.out_poly_stride = Ql * N;
    CU(ntt_forward(c->T, delta, delta, mf, &e, s.st));
    count_launches(3);
    return B200FHE_OK;
}
