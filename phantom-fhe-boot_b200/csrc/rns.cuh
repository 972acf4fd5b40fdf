// rns.cuh — launch interface of the RNS kernels (rns.cu): base conversion, evaluation-key inner
// product, tensor product, rescale helper, automorphism gather, element-wise ops.
#pragma once
#include "arith.cuh"

namespace b200fhe {

// limb position -> prime index map shared by all kernels (reference twr mapping,
// src/eval_key_switch.cu:35, src/ntt/fntt_2d.cu:434-436)
struct PrimeMap {
    u32 size_Ql, size_Q;  // size_Ql == 0: identity
    __host__ __device__ u32 prime(u32 t) const { return size_Ql == 0 ? t : (t < size_Ql ? t : size_Q + (t - size_Ql)); }
};

// out[pos(j)][c] = Barrett( sum_i in[i][c] * mat[j][i] )  mod prime(pos(j)),  j < no
// pos(j) = j < skip_start ? j : j + skip_len   (modup writes around its own digit: rns_bconv.cu:479)
// grid.y batches polynomials: in + y*in_stride, out + y*out_stride (same matrix).
struct BconvArgs {
    const u64 *in;
    u64 *out;
    const u64 *mat;  // [no][ni] device
    u32 ni, no;
    u32 skip_start, skip_len;
    PrimeMap pm;
    size_t in_stride, out_stride;  // per batch element (u64)
    u32 batch;
    u32 j0;  // first output row of this launch (limb-sharded key switch: a rank converts into its own limbs only);
             // `no` rows from j0 on, `mat` still points at row 0
    u32 lazy;  // != 0: outputs in [0, 4q) instead of [0, q) (barrett128_lazy): for results that go straight into a forward
               // transform, whose first butterflies take lazy inputs; every prime must exceed 2^32
    u32 split_at, j0b;  // split_at != 0: launch rows >= split_at are the rows j0b + (j - split_at) (a rank's own Q rows
                        // and own P rows of the Q_l P basis are two separate runs)
    __host__ __device__ u32 row(u32 j) const { return split_at && j >= split_at ? j0b + (j - split_at) : j0 + j; }
};
cudaError_t bconv(const Mod *mods, const BconvArgs &a, size_t N, cudaStream_t st);

// The beta conversions of one modup (one per digit, same batch) as ONE launch: grid.y = digit * batch + element.
// Each digit's CTAs also copy the digit's own limbs (already in NTT form) from `own_src` to their place in the
// output (modup_copy_partQl_kernel, src/rns_bconv.cu:455), so a modup has no separate copies.
static constexpr int kBconvDigits = 4;
struct BconvMultiArgs {
    BconvArgs d[kBconvDigits];
    const u64 *own_src[kBconvDigits];  // digit's own limbs in the source polynomial (c2 + skip_start * N); nullptr = no copy
    size_t own_stride;                 // per batch element (u64)
    u32 own_first[kBconvDigits], own_count[kBconvDigits];  // own_count != 0: copy only limbs [own_first, +own_count) of the digit
    u32 ndig;
};
cudaError_t bconv_multi(const Mod *mods, const BconvMultiArgs &a, size_t N, cudaStream_t st);

// alpha == 1 fast paths (modup_bconv_single_p_kernel rns_bconv.cu:432, moddown_bconv_single_p_kernel :691):
// out[pos(j)][c] = in_prime > out_prime ? in[c] mod out_prime : in[c]
cudaError_t bconv_single(const Mod *mods, const BconvArgs &a, u32 in_prime_idx, size_t N, cudaStream_t st);

static constexpr int kMaxDigits = 64;
struct EvkPtrs {
    const u64 *d[kMaxDigits];  // digit d -> [2][size_QP][N]
    const u64 *const *dev_table;  // != nullptr: the digit pointers are read from this DEVICE array instead (the reference's
                                  // PhantomRelinKey::public_keys_ptr(), include/secretkey.h:124); inner_prod only
};
// cx[p][t][c] = sum_d up[d][t][c] * evk[d][p][prime(t)][c]   (eval_key_switch.cu:26-85)
// perm (optional, N u32): output coefficient c reads input coefficient perm[c] of both operands'
// ... not used here; automorphism is applied by gather_limbs.
// batch > 1: up + z*beta*QlP*N, cx + z*2*QlP*N for z < batch (same key)
cudaError_t inner_prod(const Mod *mods, u64 *cx, const u64 *up, const EvkPtrs &evk, u32 beta, u32 size_QlP,
                       u32 size_QP, PrimeMap pm, size_t N, cudaStream_t st, u32 batch = 1);
// the same for the limbs t0 <= t < t0 + tn only (limb-sharded key switch)
cudaError_t inner_prod_range(const Mod *mods, u64 *cx, const u64 *up, const EvkPtrs &evk, u32 beta, u32 size_QlP,
                             u32 size_QP, PrimeMap pm, u32 t0, u32 tn, size_t N, cudaStream_t st, u32 t0b = 0, u32 tnb = 0,
                             u32 batch = 1);

// (d0,d1,d2) = (a0*b0, a0*b1 + a1*b0, a1*b1) per limb, Karatsuba like polymath.cu:501-536;
// out may alias a (in-place multiply_inplace semantics: out[0..2L) over a's storage needs 3L limbs).
cudaError_t tensor_prod(const Mod *mods, u64 *out, const u64 *a, const u64 *b, u32 L, size_t N, cudaStream_t st);
cudaError_t tensor_square(const Mod *mods, u64 *out, const u64 *a, u32 L, size_t N, cudaStream_t st);

// out[y][j][c] = in[y][c] mod q_j, j < L   (divide_and_round_reduce_q_last_kernel, rns.cu:1128)
cudaError_t reduce_last(const Mod *mods, u64 *out, const u64 *in, u32 L, u32 batch, size_t in_stride,
                        size_t out_stride, size_t N, cudaStream_t st);

// dst[y][l][c] = src[y][l][perm[c]]   (apply_galois_ntt_permutation, galois.cu:11-18)
cudaError_t gather_limbs(u64 *dst, const u64 *src, const u32 *perm, u32 limbs, u32 batch, size_t src_stride,
                         size_t dst_stride, size_t N, cudaStream_t st);
// PrecomputeAutoMapKernel (util.cu:941-961)
cudaError_t automap_table(u32 *perm, int logN, u32 k, cudaStream_t st);

enum EltOp { ELT_ADD = 0, ELT_SUB = 1, ELT_MUL = 2, ELT_NEG = 3 };
// out[l][c] = a[l][c] (op) b[l][c] mod prime(l)
cudaError_t eltwise(const Mod *mods, EltOp op, u64 *out, const u64 *a, const u64 *b, u32 limbs, PrimeMap pm, size_t N,
                    cudaStream_t st);
// out[l][c] = a[l][c] * s[l] mod prime(l)  (Shoup scalar per limb)
cudaError_t scalar_mul(const Mod *mods, u64 *out, const u64 *a, const u64 *s, const u64 *s_shoup, u32 limbs,
                       PrimeMap pm, size_t N, cudaStream_t st);

}  // namespace b200fhe
