// context.h — device-resident tables, per-level constants and workspace of one b200fhe context.
#pragma once
#include <map>
#include <memory>
#include <mutex>
#include <vector>

#include "hostmath.h"
#include "ntt.cuh"
#include "rns.cuh"

namespace b200fhe {

struct DevBuf {  // cudaMalloc RAII
    void *p = nullptr;
    size_t bytes = 0;
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    ~DevBuf() { release(); }
    bool pinned = false;  // a captured graph holds this address: growing (= moving) the buffer is refused
    cudaError_t alloc(size_t n);
    cudaError_t upload(const void *h, size_t n);  // alloc + sync copy
    void release();
    template <class T>
    T *as() const { return reinterpret_cast<T *>(p); }
};

struct DeviceLevel {
    host::LevelConsts h;
    DevBuf partQlHatInv, partQlHatInv_shoup;  // [Ql]
    std::vector<std::unique_ptr<DevBuf>> modup_mat;
    DevBuf bigPInv, bigPInv_shoup;            // [Ql]
    DevBuf bigP, bigP_shoup;                  // [Ql]
    DevBuf moddown_scale, moddown_scale_shoup;  // [QlP]: entries Ql+i = PHatInv_mod_p[i]
    DevBuf PHat_mod_q;                        // [Ql][P]
    DevBuf inv_q_last, inv_q_last_shoup;      // [Ql-1]
};

struct Context {
    int logN = 0;
    size_t N = 0, size_QP = 0, size_P = 0, size_Q = 0;
    int device = -1;
    bool host_only = false;
    bool lazy_conv = true;   // every prime > 2^32: internal base conversions may leave results in [0, 4q) (barrett128_lazy)
    bool external_tables = false;  // built by init_from_tables: no host copies of the twiddle tables
    std::vector<uint64_t> primes;
    std::vector<host::PrimeTables> tabs;  // host copies, reference layout
    std::vector<Mod> hmods;
    DevBuf d_mods, d_tw, d_itw, d_ninv, d_ninv_shoup;
    NttTables T{};
    std::mutex mu;
    std::map<size_t, std::unique_ptr<DeviceLevel>> levels;
    std::map<size_t, std::unique_ptr<host::LevelConsts>> host_levels;
    // workspace (grown on demand, never shrunk): composite ops on one context are single-stream
    DevBuf ws_tcks, ws_up, ws_cx, ws_delta, ws_tmp, ws_perm;
    int ws_pins = 0;  // captured graphs alive that bake the workspace addresses (b200fhe_boot_graph_capture)
    void pin_workspaces(bool on) {
        ws_pins += on ? 1 : -1;
        for (DevBuf *w : {&ws_tcks, &ws_up, &ws_cx, &ws_delta, &ws_tmp, &ws_perm}) w->pinned = ws_pins > 0;
    }
    // host-buffer pipeline (b200fhe_ntt_host_batch): 3 staging slots, copy-in / compute / copy-out streams
    static constexpr int kStageSlots = 4;
    DevBuf stage[kStageSlots];
    cudaStream_t pipe_stream[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t pipe_ev[kStageSlots][3] = {};
    // digit lanes of a single key switch (capi.cu do_keyswitch): the beta digit chains of the modup and the two polynomials
    // of the moddown are independent until the inner product / the result, and each of their launches is below one wave at
    // batch 1, so they run on side streams forked from and joined to the caller's stream with events (graph-capturable)
    static constexpr int kLanes = 4;
    cudaStream_t lane[kLanes] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t lane_fork = nullptr, lane_join[kLanes] = {nullptr, nullptr, nullptr, nullptr};
    // Work that is off the critical path of a composite operation — the (d0, d1) half of the tensor product, the Q_l rows of
    // the inner product — is HBM-bound while the chain it runs beside (transforms, base conversions) is integer-bound.  The
    // lanes are created with the greatest stream priority, so their thread blocks are placed first and the background
    // launch fills the memory system under them.
    cudaStream_t bg = nullptr;
    cudaEvent_t bg_start = nullptr, bg_done = nullptr, ipq_done = nullptr;
    cudaEvent_t acc_wait = nullptr;   // set by do_hmult for the key switch it calls: the addend is ready after this event
    int overlap = 7;   // bit 0: (d0, d1) of the tensor product in the background; bit 1: Q_l rows of the inner product beside the moddown
                       // chains; bit 2: row pass of the modup transforms fused with the inner product (ntt.cuh FwdInnerProd); 0: none of them,
                       // lanes at normal priority (env B200FHE_OVERLAP, read when the lanes are created)
    bool overlap_env_read = false;
    void read_overlap_env();
    int ks_lanes = 1;  // 0: everything on the caller's stream (b200fhe_ctx_set_lanes)
    int ks_lanes_max_batch = 4;  // larger batches already fill the GPU launch by launch
    int ensure_lanes();
    // permutation tables of the automorphisms b200fhe_rotate_fused has been called with (built once per index)
    std::map<uint32_t, std::unique_ptr<DevBuf>> perm_cache;
    ~Context();

    int init(int logN, const uint64_t *primes_QP, size_t size_QP, size_t size_P, int device, bool host_only);
    // tables the caller already holds on the device (the reference's DNTTTable): interleaved copies are made on `st`
    int init_from_tables(int logN, const uint64_t *primes_QP, size_t size_QP, size_t size_P, const uint64_t *tw,
                         const uint64_t *tw_shoup, const uint64_t *itw, const uint64_t *itw_shoup, const uint64_t *ninv,
                         const uint64_t *ninv_shoup, cudaStream_t st);
    DeviceLevel *level(size_t size_Ql);
    const host::LevelConsts *host_level(size_t size_Ql);
    LimbMap limb_map(size_t start, size_t limbs, size_t size_Ql, size_t npoly, size_t stride) const;
    PrimeMap prime_map(size_t size_Ql) const { return PrimeMap{(u32) size_Ql, (u32) size_Q}; }
};

// ---- composite operations (capi.cu), shared by the C ABI and the evaluator / bootstrap layer.
// All return a B200FHE_* status and set the thread-local error text through fail().
int fail(int code, const char *fmt, ...);
void count_launches(uint64_t n);
int get_level(Context *c, size_t size_Ql, DeviceLevel **out);
int make_evk(const Context *c, const DeviceLevel *L, const uint64_t *const *digits, EvkPtrs &e);
// DRNSTool::modup: c2 [Ql][N] -> dst [beta][QlP][N]
int do_modup(Context *c, DeviceLevel *L, u64 *dst, const u64 *c2, cudaStream_t st);
int do_modup_batch(Context *c, DeviceLevel *L, u64 *dst, const u64 *c2, size_t c2_stride, u32 batch, cudaStream_t st);
int do_inner_prod(Context *c, DeviceLevel *L, u64 *cx, const u64 *up, const EvkPtrs &e, cudaStream_t st, u32 batch = 1);
// acc_polys == 1: the addend exists only for polynomial 0 of every ciphertext (EvalRotateFused on a batch)
int do_keyswitch_batch(Context *c, DeviceLevel *L, u64 *out, const u64 *acc, const u64 *c2, size_t c2_stride,
                       const EvkPtrs &e, u32 batch, cudaStream_t st, u32 acc_polys = 0);
// DRNSTool::moddown_from_NTT on `batch` polynomials: cx + z*cx_stride ([QlP][N], P part clobbered) ->
// out + z*out_stride ([Ql][N]) (+ acc for polynomials z < acc_polys; acc_polys == 0 with acc != null: all)
int do_moddown(Context *c, DeviceLevel *L, u64 *out, size_t out_stride, u64 *cx, size_t cx_stride, u32 batch,
               const u64 *acc, size_t acc_stride, u32 acc_polys, cudaStream_t st, u32 acc_period = 0, bool acc_by_poly = false);
// multiply_and_relin on `batch` ciphertext pairs a[z], b[z] ([2][Ql][N]; b == nullptr: square) -> out [batch][2][Ql][N];
// out may alias a or b
int do_hmult(Context *c, DeviceLevel *L, u64 *out, const u64 *a, const u64 *b, const EvkPtrs &e, u32 batch, cudaStream_t st);
// out [2][Ql][N] = acc (first acc_polys polynomials) + KeySwitch(c2)
int do_keyswitch(Context *c, DeviceLevel *L, u64 *out, const u64 *acc, u32 acc_polys, const u64 *c2, const EvkPtrs &e,
                 cudaStream_t st);
// DRNSTool::divide_and_round_q_last_ntt: src [polys][Ql][N] -> dst [polys][Ql-1][N]
int do_rescale(Context *c, size_t size_Ql, u64 *dst, const u64 *src, size_t polys, cudaStream_t st);

// rows [q_lo, q_hi) of Q_l and [p_lo, p_hi) of P of every digit of a modup, given tcks = the scaled inverse transform of c2
// (capi_shard.cu): the conversion launch for all digits + the forward transform of the converted rows
int modup_rows(Context *c, DeviceLevel *L, u64 *up, const u64 *tcks, const u64 *c2, size_t q_lo, size_t q_hi, size_t p_lo,
               size_t p_hi, cudaStream_t st);

// b200fhe_peer_exchange_mc (capi_shard.cu) with every segment stored `rep` times, copy k at offset + k * rep_stride bytes:
// the own limb rows of a block of polynomials as ONE launch and ONE flag round trip
int peer_exchange_strided(const void *local_base, void *const *peer_bases, void *multicast_base, uint64_t *const *peer_flags,
                          uint32_t *counter, uint32_t world, uint32_t rank, const size_t *seg_offset, const size_t *seg_bytes,
                          uint32_t nseg, uint32_t rep, size_t rep_stride, uint64_t epoch, cudaStream_t stream);


// All-to-all over the same arenas and flag rows: to rank g go the (up to kPeerSegs) row blocks segs[g][.]; entry [rank] is
// ignored.  A block is `rep` runs of `len` bytes, run k read at src + k * src_stride and stored at arena offset
// dst_off + k * dst_stride of the destination.  Ends with the same flag barrier: when the kernel is done, every peer's rows
// for this rank have arrived.  Every rank must call it the same number of times (ranks with nothing to send pass rep = 0).
static constexpr int kMaxRanks = 9;   // ranks of one peer-memory group (the exchange kernels carry 8 peers by value)
static constexpr int kPeerSegs = 2;
struct PeerSeg {
    const char *src;
    size_t dst_off, len;
    uint32_t rep;
    size_t src_stride, dst_stride;
};
// out[r][.] = sum over nsrc slots of slots[g][r][.] for `rows` rows; row r is row row0 + r of the block [2][QlP][N] ++ [Ql][N]
// periodic: the rows are polynomials of size_QlP limbs one after the other instead (row R is limb R % size_QlP)
int peer_row_sum(const Mod *mods, u64 *out, const u64 *slots, size_t slot_stride, u32 nsrc, u32 row0, u32 rows, u32 size_QlP,
                 PrimeMap pm, size_t N, cudaStream_t st, bool periodic = false);
int peer_all_to_all(void *const *peer_bases, uint64_t *const *peer_flags, uint32_t *counter, uint32_t world, uint32_t rank,
                    const PeerSeg (*segs)[kPeerSegs], cudaStream_t stream);

}  // namespace b200fhe
