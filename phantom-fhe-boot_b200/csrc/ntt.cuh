// ntt.cuh — launch interface of the negacyclic NTT kernels (ntt.cu).
#pragma once
#include "arith.cuh"

namespace b200fhe {

// Which prime a limb uses and which limbs a launch touches.  Generalises the reference's launcher
// variants (plain / include_special_mod / exclude_range: src/ntt/fntt_2d.cu:620-736,
// src/ntt/ntt_modup.cu:610-657) into one descriptor:
//   limb t = start + blockIdx.y;   prime(t) = size_Ql == 0 ? t : (t < size_Ql ? t : size_Q + t - size_Ql)
//   poly z = blockIdx.z at data + z * poly_stride; if excl_alpha > 0 poly z skips limbs
//   [z*excl_alpha, min((z+1)*excl_alpha, size_Ql))  (the modup digit that is already in NTT form).
struct LimbMap {
    u32 start, limbs;
    u32 size_Ql, size_Q;
    u32 npoly;
    u32 excl_alpha;
    u32 excl_mod;            // > 0: the digit of polynomial z is z % excl_mod (batched modup: [batch][beta] polynomials)
    u32 src_bcast;           // forward only: every limb of polynomial z reads the SAME source row src + z*src_poly_stride
                             // (rescale: the last limb's coefficients, values < 4q of every target prime)
    size_t poly_stride;      // in u64
    size_t src_poly_stride;  // in u64 (out-of-place source)
    u32 excl_lo, excl_hi;    // forward only: limbs [excl_lo, excl_hi) of every polynomial are skipped (the reference's
                             // exclude_range launcher, src/ntt/ntt_modup.cu:610)
    u32 split_at, start_b;   // split_at != 0: grid rows >= split_at are the limbs start_b + (y - split_at): two runs of limbs
                             // in one launch (limb-sharded key switch: a rank's Q rows and P rows); `limbs` counts both
    __host__ __device__ u32 limb(u32 y) const { return split_at && y >= split_at ? start_b + (y - split_at) : start + y; }
};

// Optional fused epilogue of the forward transform (pass 2):
//   out = (other - NTT(x)) * scale[limb]  (+ acc)      mod q
// = nwt_2d_radix8_forward_inplace_fuse_moddown (src/ntt/ntt_moddown.cu:209-214) followed by
// add_to_ct_kernel (src/rns_bconv.cu:763), and the rescale epilogue (src/rns.cu:1141-1158).
struct FwdEpilogue {
    const u64 *other;  // [poly][limb][N] indexed like dst with other_poly_stride; nullptr = plain NTT
    const u64 *acc;    // optional addend, same indexing as other (acc_poly_stride)
    const u64 *scale, *scale_shoup;  // per limb (index = limb t)
    size_t other_poly_stride, acc_poly_stride;
    u64 *out;          // destination (may differ from the transform buffer); indexed with out_poly_stride
    size_t out_poly_stride;
    u32 acc_polys;     // acc is added only for polynomials z < acc_polys (0 = all)
    u32 acc_period;    // != 0: z is taken modulo acc_period for the test and for the acc index (R rotations of one
                       // ciphertext moddown'ed as one batch all add the same c0 to their polynomial 0)
    u32 acc_by_poly;   // with acc_period: the test uses z modulo acc_period, the acc index is z itself (a batch of DIFFERENT
                       // ciphertexts, each adding its own c0 to its polynomial 0)
};

// Inverse transform only: every word of the result is ALSO stored at the same offset behind each of these bases (the
// other ranks' copies of dst in NVLink peer memory; limb-sharded key switch): the transfer rides on the last pass.
static constexpr int kMaxMirrors = 8;
struct PeerMirror {
    u64 *base[kMaxMirrors];
    u32 n;
};

struct NttTables {
    int logN;
    u32 size_QP;
    const Mod *mods;            // [size_QP]
    const ulonglong2 *tw;       // [size_QP][N]  (w, w_shoup) bit-reversed psi powers (reference order)
    const ulonglong2 *itw;      // [size_QP][N]  inverse, entry 1 carries N^-1 (reference host/ntt.cu:53-55)
    const u64 *ninv, *ninv_shoup;  // [size_QP]
    size_t tw_stride;           // per-prime stride of tw / itw in entries; 0 = 2^logN (set by ntt_forward_sub only)
};

// Forward: natural order in -> bit-reversed order out, canonical [0,q).  dst may equal src.
cudaError_t ntt_forward(const NttTables &T, u64 *dst, const u64 *src, const LimbMap &map, const FwdEpilogue *epi,
                        cudaStream_t stream);
// Inverse incl. N^-1; optional per-limb Shoup scale (index = limb t), as nwt_2d_radix8_backward_scale.
cudaError_t ntt_inverse(const NttTables &T, u64 *dst, const u64 *src, const LimbMap &map, const u64 *scale,
                        const u64 *scale_shoup, cudaStream_t stream, const PeerMirror *mirror = nullptr);

// Forward transform of polynomials in X^(N/n'), n' = 2^sub_logn <= 2048, given COMPACT (n' coefficients per limb, the
// others are zero): the first sub_logn stages of the N-point transform touch only those coefficients and use the first
// n' entries of the same twiddle table, the remaining stages pair every value with a zero, i.e. copy it.  So
// NTT_N(p)[ (N/n') k + r ] = out[k] for all r < N/n': the N/n' times smaller transform is the whole result, bit for bit.
// dst / src: [npoly][..limbs..][n'] (limb stride n', polynomial strides from the map).
cudaError_t ntt_forward_sub(const NttTables &T, int sub_logn, u64 *dst, const u64 *src, const LimbMap &map, cudaStream_t stream);

// The forward transform of a modup with the evaluation-key inner product as its epilogue (N >= 2^12).
// ntt_forward_p1 runs the column pass only (for N = 2^17 with the radix-2 stage before it); ntt_forward_p2_inner_prod then
// takes, per CTA, the same tile of 256-coefficient rows of limb t from EVERY digit — finishing the transform of the digits that
// were converted into t, reading the digit that owns t from the source polynomial as it is — and writes
//   cx[p][t] = sum_d up_d[t] * key_d[p][prime(t)]          (eval_key_switch.cu:26-85 after ntt_modup.cu:610)
// so the transformed digits never go to HBM and the keys stream in under the butterflies of the other CTAs of the SM.
static constexpr int kIpDigits = 4;
struct FwdInnerProd {
    const u64 *up;           // [batch][beta][size_QlP][N]: pass-1 output in the converted limbs (a digit's own limbs unused)
    const u64 *own;          // source polynomials in NTT form, [size_Ql][N] each, own_stride apart
    u64 *cx;                 // [batch][2][size_QlP][N]
    const u64 *key[kIpDigits];  // digit d -> [2][size_QP][N]
    size_t up_stride, own_stride;   // per batch element (u64)
    u32 beta, alpha, size_Ql, size_QlP, size_Q, size_QP, batch;
};
cudaError_t ntt_forward_p1(const NttTables &T, u64 *dst, const u64 *src, const LimbMap &map, cudaStream_t stream);
// limbs t0 .. t0 + tn - 1 of the extended basis (0 .. size_QlP - 1)
cudaError_t ntt_forward_p2_inner_prod(const NttTables &T, const FwdInnerProd &a, u32 t0, u32 tn, cudaStream_t stream);

// inverse transforms at N = 2^16 of at most n (limb, polynomial) pairs take the one-kernel cluster form (ntt_inv_cluster);
// 0 = never, < 0 = back to the default / B200FHE_CLUSTER_NTT
void ntt_set_cluster_max_limbs(int n);
int ntt_launch_count_forward(int logN);
int ntt_launch_count_inverse(int logN);

}  // namespace b200fhe
