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

namespace {
int make_mirror(uint64_t *const *peers, uint32_t n, PeerMirror &m) {
    m = PeerMirror{};
    if (n > (uint32_t) kMaxMirrors) return fail(B200FHE_ERR_ARG, "at most 8 peer copies");
    if (n && !peers) return fail(B200FHE_ERR_ARG, "null peer list");
    for (uint32_t i = 0; i < n; i++) {
        if (!peers[i]) return fail(B200FHE_ERR_ARG, "null peer pointer");
        m.base[i] = reinterpret_cast<u64 *>(peers[i]);
    }
    m.n = n;
    return B200FHE_OK;
}
}  // namespace

// The rows [q_lo, q_hi) of Q_l and [p_lo, p_hi) of P of every digit of a modup (DRNSTool::modup after its inverse transform):
// tcks = INTT(c2) x (Q_hat_d^-1 mod q_t) complete, c2 the source in NTT form; up [beta][QlP][N].  ONE conversion launch for all
// digits -- digit d converts into the wanted rows (two runs of rows of the Q_l P basis minus the digit) and copies the wanted
// rows of the digit itself, which stay in NTT form -- and one forward transform of the converted rows.
int b200fhe::modup_rows(Context *c, DeviceLevel *L, u64 *up, const u64 *tcks, const u64 *c2, size_t q_lo, size_t q_hi, size_t p_lo,
                        size_t p_hi, cudaStream_t st) {
    const size_t N = c->N, Ql = L->h.size_Ql, QlP = Ql + c->size_P, alpha = L->h.alpha, beta = L->h.beta;
    const size_t nq = q_hi - q_lo, np = p_hi - p_lo;
    if (beta > (size_t) kBconvDigits) return fail(B200FHE_ERR_UNSUPPORTED, "more than %d digits", kBconvDigits);
    BconvMultiArgs A{};
    A.ndig = (u32) beta;
    A.own_stride = 0;
    for (size_t d = 0; d < beta; d++) {
        const size_t ds = alpha * d, ni = L->h.digit_size(d), de = ds + ni;
        auto below = [&](size_t p) { return p < ds ? p : std::max(p, de) - ni; };  // non-digit positions below p
        BconvArgs &a = A.d[d];
        a.in = tcks + ds * N;
        a.out = up + d * QlP * N;
        a.mat = L->modup_mat[d]->as<u64>();
        a.ni = (u32) ni;
        a.skip_start = (u32) ds;
        a.skip_len = (u32) ni;
        a.pm = c->prime_map(Ql);
        a.in_stride = a.out_stride = 0;
        a.batch = 1;
        a.lazy = c->lazy_conv ? 1u : 0u;   // into the forward transform
        const size_t r0 = below(q_lo), n0 = below(q_hi) - r0, r1 = below(Ql + p_lo), n1 = below(Ql + p_hi) - r1;
        a.j0 = (u32) (n0 ? r0 : r1);
        a.no = (u32) (n0 + n1);
        a.split_at = (u32) (n0 && n1 ? n0 : 0);
        a.j0b = (u32) r1;
        const size_t o0 = std::max(q_lo, ds), o1 = std::min(q_hi, de);
        A.own_src[d] = o1 > o0 ? c2 + ds * N : nullptr;
        A.own_first[d] = o1 > o0 ? (u32) (o0 - ds) : 0;
        A.own_count[d] = o1 > o0 ? (u32) (o1 - o0) : 0;
    }
    CU(bconv_multi(c->T.mods, A, N, st));
    if (nq + np) {
        LimbMap mf = c->limb_map(nq ? q_lo : Ql + p_lo, nq + np, Ql, beta, QlP * N);
        mf.split_at = (u32) (nq && np ? nq : 0);
        mf.start_b = (u32) (Ql + p_lo);
        mf.excl_alpha = (u32) alpha;
        mf.excl_mod = (u32) beta;
        CU(ntt_forward(c->T, up, up, mf, nullptr, st));
    }
    return B200FHE_OK;
}

// Phase 1: own rows of tcks = INTT(c2) x (Q_hat_d^-1 mod q_t)   (DRNSTool::modup prologue, src/rns_bconv.cu:544-560)
extern "C" int b200fhe_shard_ks_begin(b200fhe_ctx *ctx, size_t size_Ql, size_t q_lo, size_t q_hi, const uint64_t *c2,
                                      uint64_t *tcks, void *stream) {
    return b200fhe_shard_ks_begin_mirrored(ctx, size_Ql, q_lo, q_hi, c2, tcks, nullptr, 0, stream);
}

extern "C" int b200fhe_shard_ks_begin_mirrored(b200fhe_ctx *ctx, size_t size_Ql, size_t q_lo, size_t q_hi, const uint64_t *c2,
                                               uint64_t *tcks, uint64_t *const *peer_tcks, uint32_t n_peers, void *stream) {
    PeerMirror pm;
    if (int rc = make_mirror(peer_tcks, n_peers, pm)) return rc;
    Shard s{};
    if (int rc = open_shard(ctx, size_Ql, q_lo, q_hi, 0, 0, stream, s)) return rc;
    if (!c2 || !tcks) return fail(B200FHE_ERR_ARG, "null buffer");
    if (q_hi == q_lo) return B200FHE_OK;
    LimbMap mi = s.c->limb_map(q_lo, q_hi - q_lo, 0, 1, s.Ql * s.N);
    CU(ntt_inverse(s.c->T, (u64 *) tcks, (const u64 *) c2, mi, s.L->partQlHatInv.as<u64>(), s.L->partQlHatInv_shoup.as<u64>(), s.st, &pm));
    count_launches(2);
    return B200FHE_OK;
}

// Phase 2 (tcks complete): per digit the base conversion into the OWN rows of Q_l P, their NTT, the copy of the own rows
// of the digit itself; inner product with the key on the own rows; inverse NTT x P_hat^-1 of the own P rows of cx.
extern "C" int b200fhe_shard_ks_middle(b200fhe_ctx *ctx, size_t size_Ql, size_t q_lo, size_t q_hi, size_t p_lo, size_t p_hi,
                                       const uint64_t *tcks, const uint64_t *c2, const uint64_t *const *evk, uint64_t *cx,
                                       void *stream) {
    return b200fhe_shard_ks_middle_mirrored(ctx, size_Ql, q_lo, q_hi, p_lo, p_hi, tcks, c2, evk, cx, nullptr, 0, stream);
}

extern "C" int b200fhe_shard_ks_middle_mirrored(b200fhe_ctx *ctx, size_t size_Ql, size_t q_lo, size_t q_hi, size_t p_lo,
                                                size_t p_hi, const uint64_t *tcks, const uint64_t *c2,
                                                const uint64_t *const *evk, uint64_t *cx, uint64_t *const *peer_cx,
                                                uint32_t n_peers, void *stream) {
    PeerMirror pm;
    if (int rc = make_mirror(peer_cx, n_peers, pm)) return rc;
    Shard s{};
    if (int rc = open_shard(ctx, size_Ql, q_lo, q_hi, p_lo, p_hi, stream, s)) return rc;
    if (!tcks || !c2 || !cx) return fail(B200FHE_ERR_ARG, "null buffer");
    EvkPtrs e{};
    if (int rc = make_evk(s.c, s.L, evk, e)) return rc;
    Context *c = s.c;
    const size_t N = s.N, Ql = s.Ql, QlP = s.QlP, nq = q_hi - q_lo, np = p_hi - p_lo;
    CU(c->ws_up.alloc(s.beta * QlP * N * 8));
    u64 *up = c->ws_up.as<u64>();
    if (int rc = modup_rows(c, s.L, up, (const u64 *) tcks, (const u64 *) c2, q_lo, q_hi, p_lo, p_hi, s.st)) return rc;
    if (nq + np)   // own Q rows and own P rows of all digits: one inner product
        CU(inner_prod_range(c->T.mods, (u64 *) cx, up, e, (u32) s.beta, (u32) QlP, (u32) c->size_QP, c->prime_map(Ql),
                            (u32) q_lo, (u32) nq, N, s.st, (u32) (Ql + p_lo), (u32) np));
    count_launches(4);
    if (np) {
        LimbMap mi = c->limb_map(Ql + p_lo, np, Ql, 2, QlP * N);
        CU(ntt_inverse(c->T, (u64 *) cx, (const u64 *) cx, mi, s.L->moddown_scale.as<u64>(), s.L->moddown_scale_shoup.as<u64>(), s.st, &pm));
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
    a.lazy = c->lazy_conv ? 1u : 0u;
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

// ---------------------------------------------------------------------------------------------------------------------
// Row exchange over NVLink peer memory (no NCCL on the data path).  Every rank keeps the exchanged buffers (tcks, cx,
// optionally the result) at the SAME offsets of a symmetric arena whose base on every peer is mapped into this process
// (CUDA IPC / fabric handles; torch.distributed._symmetric_memory does the handle exchange in the Python host).
// One launch per boundary: the blocks store the rank's own rows straight into every peer's arena (16-byte stores over
// NVLink), the last block to finish publishes `epoch` in each peer's flag row and waits for every peer's epoch in its
// own, so the kernel ends only when this GPU's copy of the rows is complete -- the next kernel on the stream (the base
// conversion) needs no other synchronisation.  Reuse is safe without double buffering: a rank overwrites tcks rows of
// key switch k+1 only after the cx boundary of key switch k, which every rank reaches after it has read tcks (and the
// other way round for cx).
namespace {
constexpr u32 kMaxPeers = 8, kMaxSegs = 4;
struct PeerPutArgs {
    const char *src;
    char *peer[kMaxPeers];        // arena bases of the OTHER ranks
    char *mc;                     // multicast address of the arena (or nullptr)
    u64 *peer_flag[kMaxPeers];    // their flag rows
    u32 peer_rank[kMaxPeers];
    size_t off[kMaxSegs], len[kMaxSegs];
    u64 *my_flags;
    unsigned *counter;
    u64 epoch;
    u32 npeers, nseg, rank;
    u32 rep;              // every segment is stored `rep` times, copy k at off + k * rep_stride (the own limb rows of every
    size_t rep_stride;    // polynomial of a block of extended ciphertexts: one segment, rep = polynomials)
};

// MC: a.mc is the arena's NVSwitch multicast address -- ONE store is replicated by the switch into every rank's arena
// (this rank's included, same bytes), so the egress is the rows once instead of once per peer
template <bool MC>
__device__ __forceinline__ void put16(uint4 *d, const uint4 &v) {
    if (MC)
        asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(d), "f"(__uint_as_float(v.x)),
                     "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
                     : "memory");
    else
        *d = v;
}

static constexpr unsigned long long kPeerWaitNs = 20ull * 1000 * 1000 * 1000;  // 20 s: far above any skew between healthy ranks

// End of an exchange kernel: the last block to finish (ticket counter after a system fence) publishes the exchange number in
// every peer's flag row and waits for every peer's number in its own.  epoch 0: the exchanges are numbered on the device
// (counter[2..3] as one u64), so a captured CUDA graph can be replayed -- every rank runs the same sequence of exchanges.
__device__ __forceinline__ void peer_barrier_tail(unsigned *counter, u64 epoch_arg, u32 npeers, u32 rank, u64 *const *peer_flag,
                                                  const u32 *peer_rank, u64 *my_flags) {
    __threadfence_system();
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) last = atomicAdd(counter, 1u) == gridDim.x * gridDim.y - 1;
    __syncthreads();
    if (!last) return;
    u64 *auto_epoch = reinterpret_cast<u64 *>(counter + 2);
    const u64 epoch = epoch_arg ? epoch_arg : *auto_epoch + 1;
    __syncthreads();
    if (threadIdx.x == 0) {
        *counter = 0;  // for the next launch (stream ordered)
        if (!epoch_arg) *auto_epoch = epoch;
    }
    __threadfence_system();
    if (threadIdx.x < npeers) {
        u64 *f = peer_flag[threadIdx.x] + rank;
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(f), "l"(epoch) : "memory");
        const u64 *w = my_flags + peer_rank[threadIdx.x];
        u64 v, t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        do {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(w) : "memory");
            if (v >= epoch) break;
            // a peer that never arrives (its process died, or it skipped an exchange after an error) must not hang this GPU:
            // after kPeerWaitNs the wait is given up and counter[1] counts it -- the data of this exchange is then incomplete
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > kPeerWaitNs) {
                atomicAdd(counter + 1, 1u);
                break;
            }
        } while (true);
    }
    __threadfence_system();
}

template <bool MC>
__global__ void __launch_bounds__(256) peer_put_kernel(const PeerPutArgs a) {
    const u32 p = blockIdx.y;
    for (u32 sk = 0; sk < a.nseg * a.rep; sk++) {
        const u32 s = sk / a.rep;
        const size_t o = a.off[s] + size_t(sk % a.rep) * a.rep_stride;
        const uint4 *s4 = reinterpret_cast<const uint4 *>(a.src + o);
        uint4 *d4 = reinterpret_cast<uint4 *>((MC ? a.mc : a.peer[p]) + o);
        const size_t n = a.len[s] / 16;
        const size_t step = size_t(gridDim.x) * blockDim.x;
        size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
        for (; i + 3 * step < n; i += 4 * step) {  // four loads in flight per thread before the first NVLink store
            const uint4 v0 = s4[i], v1 = s4[i + step], v2 = s4[i + 2 * step], v3 = s4[i + 3 * step];
            put16<MC>(d4 + i, v0);
            put16<MC>(d4 + i + step, v1);
            put16<MC>(d4 + i + 2 * step, v2);
            put16<MC>(d4 + i + 3 * step, v3);
        }
        for (; i < n; i += step) put16<MC>(d4 + i, s4[i]);
    }
    peer_barrier_tail(a.counter, a.epoch, a.npeers, a.rank, a.peer_flag, a.peer_rank, a.my_flags);
}
}  // namespace

extern "C" int b200fhe_peer_exchange(const void *local_base, void *const *peer_bases, uint64_t *const *peer_flags,
                                     uint32_t *counter, uint32_t world, uint32_t rank, const size_t *seg_offset,
                                     const size_t *seg_bytes, uint32_t nseg, uint64_t epoch, void *stream) {
    return b200fhe_peer_exchange_mc(local_base, peer_bases, nullptr, peer_flags, counter, world, rank, seg_offset, seg_bytes,
                                    nseg, epoch, stream);
}

extern "C" int b200fhe_peer_exchange_mc(const void *local_base, void *const *peer_bases, void *multicast_base,
                                        uint64_t *const *peer_flags, uint32_t *counter, uint32_t world, uint32_t rank,
                                        const size_t *seg_offset, const size_t *seg_bytes, uint32_t nseg, uint64_t epoch,
                                        void *stream) {
    return b200fhe::peer_exchange_strided(local_base, peer_bases, multicast_base, peer_flags, counter, world, rank, seg_offset,
                                          seg_bytes, nseg, 1, 0, epoch, reinterpret_cast<cudaStream_t>(stream));
}

int b200fhe::peer_exchange_strided(const void *local_base, void *const *peer_bases, void *multicast_base,
                                   uint64_t *const *peer_flags, uint32_t *counter, uint32_t world, uint32_t rank,
                                   const size_t *seg_offset, const size_t *seg_bytes, uint32_t nseg, uint32_t rep,
                                   size_t rep_stride, uint64_t epoch, cudaStream_t stream) {
    if (world <= 1) return B200FHE_OK;
    if (rep == 0 || (rep > 1 && rep_stride % 16)) return fail(B200FHE_ERR_ARG, "peer exchange: bad repetition");
    if (world > kMaxPeers + 1 || rank >= world || nseg > kMaxSegs) return fail(B200FHE_ERR_ARG, "peer exchange: at most 9 ranks, 4 segments");
    if (!local_base || !peer_bases || !peer_flags || !counter || (nseg && (!seg_offset || !seg_bytes)))
        return fail(B200FHE_ERR_ARG, "null buffer");
    PeerPutArgs a{};
    a.src = static_cast<const char *>(local_base);
    size_t most = 0;
    for (u32 s = 0; s < nseg; s++) {
        if (seg_offset[s] % 16 || seg_bytes[s] % 16) return fail(B200FHE_ERR_ARG, "peer exchange: segments must be 16-byte aligned");
        a.off[a.nseg] = seg_offset[s];
        a.len[a.nseg] = seg_bytes[s];
        if (seg_bytes[s]) a.nseg++;
        most = std::max(most, seg_bytes[s]);
    }
    for (u32 g = 0; g < world; g++) {
        if (g == rank) continue;
        if (!peer_bases[g] || !peer_flags[g]) return fail(B200FHE_ERR_ARG, "peer exchange: null peer pointer");
        a.peer[a.npeers] = static_cast<char *>(peer_bases[g]);
        a.peer_flag[a.npeers] = reinterpret_cast<u64 *>(peer_flags[g]);
        a.peer_rank[a.npeers] = g;
        a.npeers++;
    }
    a.my_flags = reinterpret_cast<u64 *>(peer_flags[rank]);
    a.counter = counter;
    a.epoch = epoch;
    a.rank = rank;
    a.rep = rep;
    a.rep_stride = rep_stride;
    most *= rep;
    // enough blocks to keep the NVLink ports busy, few enough that the tail (ticket + flag round trip) stays short
    a.mc = static_cast<char *>(multicast_base);
    if (a.mc && a.nseg) {
        const u32 blocks = (u32) std::min<size_t>(std::max<size_t>(1, most / (16 * 256 * 8)), 296);
        peer_put_kernel<true><<<dim3(blocks, 1), 256, 0, stream>>>(a);
    } else {
        const u32 per_peer = (u32) std::min<size_t>(std::max<size_t>(1, most / (16 * 256 * 8)), std::max<u32>(1, 296 / a.npeers));
        peer_put_kernel<false><<<dim3(per_peer, a.npeers), 256, 0, stream>>>(a);
    }
    CU(cudaGetLastError());
    count_launches(1);
    return B200FHE_OK;
}

// ---- all-to-all form: a different set of rows for every peer (the baby steps of a limb-sharded linear-transform stage, the
// reduce-scatter of its partial sums, the halves of EvalMod between partner ranks), same barrier at the end
namespace {
struct PeerA2AArgs {
    char *peer[kMaxPeers];
    u64 *peer_flag[kMaxPeers];
    u32 peer_rank[kMaxPeers];
    b200fhe::PeerSeg seg[kMaxPeers][b200fhe::kPeerSegs];
    u64 *my_flags;
    unsigned *counter;
    u32 npeers, rank;
};

__global__ void __launch_bounds__(256) peer_a2a_kernel(const __grid_constant__ PeerA2AArgs a) {
    const u32 p = blockIdx.y;
    const size_t step = size_t(gridDim.x) * blockDim.x;
    for (int s = 0; s < b200fhe::kPeerSegs; s++) {
        const b200fhe::PeerSeg &g = a.seg[p][s];
        const size_t n = g.len / 16;
        for (u32 k = 0; k < g.rep; k++) {
            const uint4 *s4 = reinterpret_cast<const uint4 *>(g.src + size_t(k) * g.src_stride);
            uint4 *d4 = reinterpret_cast<uint4 *>(a.peer[p] + g.dst_off + size_t(k) * g.dst_stride);
            size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
            for (; i + 3 * step < n; i += 4 * step) {
                const uint4 v0 = s4[i], v1 = s4[i + step], v2 = s4[i + 2 * step], v3 = s4[i + 3 * step];
                d4[i] = v0;
                d4[i + step] = v1;
                d4[i + 2 * step] = v2;
                d4[i + 3 * step] = v3;
            }
            for (; i < n; i += step) d4[i] = s4[i];
        }
    }
    peer_barrier_tail(a.counter, 0, a.npeers, a.rank, a.peer_flag, a.peer_rank, a.my_flags);
}
}  // namespace

int b200fhe::peer_all_to_all(void *const *peer_bases, uint64_t *const *peer_flags, uint32_t *counter, uint32_t world, uint32_t rank,
                             const PeerSeg (*segs)[kPeerSegs], cudaStream_t stream) {
    if (world <= 1) return B200FHE_OK;
    if (world > kMaxPeers + 1 || rank >= world) return fail(B200FHE_ERR_ARG, "peer all-to-all: at most 9 ranks");
    if (!peer_bases || !peer_flags || !counter || !segs) return fail(B200FHE_ERR_ARG, "null buffer");
    PeerA2AArgs a{};
    size_t most = 0;
    for (u32 g = 0; g < world; g++) {
        if (g == rank) continue;
        if (!peer_bases[g] || !peer_flags[g]) return fail(B200FHE_ERR_ARG, "peer all-to-all: null peer pointer");
        a.peer[a.npeers] = static_cast<char *>(peer_bases[g]);
        a.peer_flag[a.npeers] = reinterpret_cast<u64 *>(peer_flags[g]);
        a.peer_rank[a.npeers] = g;
        size_t bytes = 0;
        for (int s = 0; s < kPeerSegs; s++) {
            const PeerSeg &sg = segs[g][s];
            if (sg.len % 16 || sg.dst_off % 16 || sg.src_stride % 16 || sg.dst_stride % 16 || (reinterpret_cast<uintptr_t>(sg.src) % 16))
                return fail(B200FHE_ERR_ARG, "peer all-to-all: segments must be 16-byte aligned");
            a.seg[a.npeers][s] = sg;
            if (!sg.len || !sg.src) a.seg[a.npeers][s].rep = 0;
            bytes += sg.len * a.seg[a.npeers][s].rep;
        }
        most = std::max(most, bytes);
        a.npeers++;
    }
    a.my_flags = reinterpret_cast<u64 *>(peer_flags[rank]);
    a.counter = counter;
    a.rank = rank;
    const u32 per_peer = (u32) std::min<size_t>(std::max<size_t>(1, most / (16 * 256 * 8)), std::max<u32>(1, 296 / a.npeers));
    peer_a2a_kernel<<<dim3(per_peer, a.npeers), 256, 0, stream>>>(a);
    CU(cudaGetLastError());
    count_launches(1);
    return B200FHE_OK;
}

// ---- sum of the ranks' partial rows (the reduce step of a limb-sharded linear-transform stage) ------------------------------
// slots: [nsrc][rows][N] (slot g = rank g's partial sums of this rank's rows); out[r] = sum_g slots[g][r] mod the prime of row
// row0 + r, where row R of the virtual block [2][QlP][N] ++ [Ql][N] is limb R % QlP of Q_l P for R < 2 QlP, else limb R - 2 QlP.
namespace {
__global__ void __launch_bounds__(256) row_sum_kernel(const Mod *__restrict__ mods, u64 *out, const u64 *slots, size_t slot_stride,
                                                      u32 nsrc, u32 row0, u32 size_QlP, PrimeMap pm, u32 n_pairs, u32 periodic) {
    const u32 pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n_pairs) return;
    const u32 R = row0 + blockIdx.y;
    // periodic: the rows are polynomials of size_QlP limbs one after the other (a block of ciphertexts)
    const bool ext = periodic || R < 2 * size_QlP;
    const u32 t = ext ? R % size_QlP : R - 2 * size_QlP;
    const u64 q = mods[ext ? pm.prime(t) : t].q;
    const size_t id = size_t(blockIdx.y) * n_pairs * 2 + 2 * size_t(pair);
    ulonglong2 acc = ld2(slots + id);
    for (u32 g = 1; g < nsrc; g++) {
        const ulonglong2 v = ld2(slots + g * slot_stride + id);
        acc.x = addmod(acc.x, v.x, q);
        acc.y = addmod(acc.y, v.y, q);
    }
    st2(out + id, acc.x, acc.y);
}
}  // namespace

int b200fhe::peer_row_sum(const Mod *mods, u64 *out, const u64 *slots, size_t slot_stride, u32 nsrc, u32 row0, u32 rows,
                          u32 size_QlP, PrimeMap pm, size_t N, cudaStream_t st, bool periodic) {
    if (rows == 0 || nsrc == 0) return B200FHE_OK;
    const u32 n_pairs = (u32) (N / 2);
    row_sum_kernel<<<dim3((n_pairs + 255) / 256, rows), 256, 0, st>>>(mods, out, slots, slot_stride, nsrc, row0, size_QlP, pm, n_pairs,
                                                                     periodic ? 1u : 0u);
    CU(cudaGetLastError());
    count_launches(1);
    return B200FHE_OK;
}
