// evalk.cuh — launch interface of the evaluator / bootstrap kernels (evalk.cu): per-limb scalar ops,
// modulus raising, hoisted ("fast") rotation in the extended basis Q_l P, and the baby-step giant-step
// plaintext-matrix product of the homomorphic DFT stages.
#pragma once
#include "rns.cuh"

namespace b200fhe {

static constexpr int kMaxLimbs = 64;   // limbs of one polynomial in Q_l P that the by-value tables below can carry
static constexpr int kMaxBsgs = 64;    // giant-step width g handled by one matrix-product launch

// per-limb scalars passed BY VALUE in the kernel-argument block (no device-side constant upload per call)
struct LimbScalars {
    u64 w[kMaxLimbs];   // scalar mod prime(limb)
    u64 ws[kMaxLimbs];  // floor(w * 2^64 / q)  (Shoup companion)
};

enum ScalarOp {
    SC_MUL = 0,     // out = a * w                 multiply_scalar_rns_poly   (src/polymath.cu)
    SC_ADD = 1,     // out = a + w                 add_scalar_rns_poly
    SC_SUB = 2,     // out = a - w                 sub_scalar_rns_poly
    SC_MULADD = 3,  // out = acc + a * w           multiply_scalar_rns_poly + add_to_ct_kernel (evaluate.cu:3726-3736)
};
// limbs x npoly polynomials; a/out/acc are indexed [poly*stride + limb*N + c]
cudaError_t limb_scalar(const Mod *mods, ScalarOp op, u64 *out, size_t out_stride, const u64 *a, size_t a_stride,
                        const u64 *acc, size_t acc_stride, const LimbScalars &s, u32 limbs, u32 npoly, PrimeMap pm,
                        size_t N, cudaStream_t st);

// out[z][t][c] = a[z][t][c] (op) b[z][t][c] mod prime(t), z < npoly; b_stride == 0 broadcasts one polynomial
// (multiply_rns_poly / add_rns_poly / sub_rns_poly over the polynomials of a ciphertext, src/polymath.cu)
// out[z][t][c] = sum_i s_i[t] * src_i[z][t][c]  (+ kappa[t] on polynomial 0 of every ciphertext)   mod q_t
// The evaluator keeps chains of scalar multiplications, additions and level drops (EvalMultConst, EvalAdd/Sub,
// EvalAddConst, MultByInteger, ModSwitchLevel: multiply_scalar/add/sub_rns_poly and add/sub_scalar_rns_poly of
// src/polymath.cu, one pass each in the reference) as a pending affine combination and evaluates it in ONE pass.
static constexpr int kMaxAffine = 8;
struct AffineArgs {
    const u64 *src[kMaxAffine];
    size_t stride[kMaxAffine];        // polynomial stride of each source, in u64 (its own limb count * N)
    u64 s[kMaxAffine][kMaxLimbs];     // per-limb multipliers
    u64 kappa[kMaxLimbs];
    u32 nterms, polys, has_kappa;
};
cudaError_t affine_combine(const Mod *mods, u64 *out, size_t out_stride, const AffineArgs &a, u32 limbs, u32 npoly,
                           size_t N, cudaStream_t st);

cudaError_t poly_eltwise(const Mod *mods, EltOp op, u64 *out, size_t out_stride, const u64 *a, size_t a_stride,
                         const u64 *b, size_t b_stride, u32 limbs, u32 npoly, PrimeMap pm, size_t N, cudaStream_t st);

// Batched tensor product of `batch` ciphertext pairs a[z], b[z] ([2][L][N] each, stride 2*L*N; b == nullptr: square):
// (d0, d1) -> out01[z] ([2][L][N]), d2 -> out2[z] ([L][N])   (tensor_prod_2x2 / tensor_square_2x2, polymath.cu:501-570)
cudaError_t tensor_split(const Mod *mods, u64 *out01, u64 *out2, const u64 *a, const u64 *b, u32 L, u32 batch, size_t N,
                         cudaStream_t st);

// tensor_split as two launches (d2 first; (d0, d1) may run beside the key switch of d2 on another stream)
cudaError_t tensor_d2(const Mod *mods, u64 *out2, const u64 *a, const u64 *b, u32 L, u32 batch, size_t N, cudaStream_t st);
cudaError_t tensor_d01(const Mod *mods, u64 *out01, const u64 *a, const u64 *b, u32 L, u32 batch, size_t N, cudaStream_t st);

// RaiseMod (src/evaluate.cu:2414-2457 switchModulusKernel): limb 0 of `src` (coefficient form, mod q_0) is
// re-centred into every prime j < L of dst ([npoly][L][N]); limb 0 is copied.
cudaError_t switch_modulus(const Mod *mods, u64 *dst, size_t dst_stride, const u64 *src, size_t src_stride, u32 L,
                           u32 npoly, size_t N, cudaStream_t st);

// init_monomial_kernel (src/evaluate.cu:2506-2519): +-x^index in coefficient form over L limbs
cudaError_t monomial_init(const Mod *mods, u64 *out, u32 index, bool negative, u32 L, size_t N, cudaStream_t st);

// One hoisted rotation in the extended basis (EvalFastRotationExt, src/evaluate.cu:3660-3755), fused:
//   out[p][t][c] = sum_d up[d][t][pi(c)] * evk[d][p][prime(t)][pi(c)]  (+ c0[t][pi(c)] * PmodQ[t]  for p == 0, t < Ql)
// pi = perm of the automorphism (NTT-domain gather, src/galois.cu:11-18).  The automorphism permutes aligned
// blocks of coefficients among themselves, so the permuted reads stay whole 32-byte sectors.
struct FastRotArgs {
    const u64 *up;        // [beta][QlP][N] digits (EvalFastRotationPrecompute)
    EvkPtrs evk;          // key of this automorphism
    const u32 *perm;      // [N] or nullptr (identity)
    const u64 *c0;        // [Ql][N] or nullptr (add_first == false)
    u64 *out;             // [2][QlP][N]
    u32 beta, size_Ql, size_QlP, size_QP;
    u32 accumulate;       // != 0: out += result (EvalAddExtInPlace of the baby-step loop folded in)
    u32 auto_index;       // the automorphism x -> x^k the table `perm` belongs to (1 = identity); the TMA kernel
    int logN;             // derives the permutation arithmetically from (k, logN) instead of reading the table
};
cudaError_t fast_rotation_ext(const Mod *mods, const FastRotArgs &a, const LimbScalars &PmodQ, PrimeMap pm, size_t N,
                              cudaStream_t st);

// All hoisted rotations of one ciphertext in ONE launch (the g - 1 giant-step rotations of a linear-transform stage share
// the digits `up` and c0): a CTA owns one 4 KiB SOURCE tile of one limb, keeps its beta digit tiles and the c0 tile in
// shared memory, and streams the key tiles of rotation after rotation through a TMA ring, writing each result to the
// tile the inverse automorphism sends it to.  The digits are read once per stage instead of once per rotation
// (-32 % HBM bytes per rotation at the top level) and the launch count of a stage drops from g - 1 to 1.
static constexpr int kRotMax = 64;     // rotations per launch
static constexpr int kRotDigits = 4;   // beta supported by the one-launch form
struct FastRotMultiArgs {
    const u64 *up;                          // [beta][QlP][N]
    const u64 *c0;                          // [Ql][N] or nullptr
    const u64 *evk[kRotMax][kRotDigits];    // key digits of rotation r
    u64 *out[kRotMax];                      // [2][QlP][N] each
    u32 k[kRotMax], k_inv[kRotMax];         // automorphism index and its inverse mod 2N
    u32 nrot, beta, size_Ql, size_QlP, size_QP;
    int logN;
    u32 limb0, nlimbs;                      // nlimbs != 0: only the limbs [limb0, limb0 + nlimbs) of Q_l P (a rank's rows of a
                                            // limb-sharded linear transform); the other rows of every output stay untouched
};
cudaError_t fast_rotation_ext_multi(const Mod *mods, const FastRotMultiArgs &a, const LimbScalars &PmodQ, PrimeMap pm,
                                    size_t N, cudaStream_t st);
bool fast_rotation_ext_multi_supported(u32 beta, int logN);

// Baby-step giant-step plaintext-matrix product of one linear-transform stage
// (the EvalMultExt / EvalAddExtInPlace double loop of FHECKKSRNS::EvalCoeffsToSlots, src/bootstrap.cu:1290-1305):
//   out[i][p][t][c] = sum_{j < g, pt[i][j] != null} ct[j][p][t][c] * pt[i][j][t][c]      mod prime(t)
// for nb baby steps i in one launch: the g rotated ciphertexts are read once for all i.
struct BsgsArgs {
    const u64 *ct[kMaxBsgs];  // g extended ciphertexts [2][QlP][N]
    const u64 *const *pt;     // HOST table [nb][g] of device plaintext pointers ([QlP][N] each), null = skip
    u64 *const *out;          // HOST table [nb] of device outputs [2][QlP][N]
                              // (both are packed into the kernel-argument block, <= 4 baby steps per launch)
    u32 g, nb, size_QlP;
    u32 accumulate;           // != 0: out[i] += ... (giant steps wider than kMaxBsgs are chunked by the caller)
    u32 limb0, nlimbs;        // nlimbs != 0: only the limbs [limb0, limb0 + nlimbs) of every output (limb-sharded stage)
    u32 pt_shift;             // != 0: the plaintexts are COMPACT, [QlP][N >> pt_shift], value k standing for the coefficients
                              // [k << pt_shift, (k + 1) << pt_shift) (ntt_forward_sub); 1 <= pt_shift <= 8
};
cudaError_t bsgs_matmul(const Mod *mods, const BsgsArgs &a, PrimeMap pm, size_t N, cudaStream_t st);

}  // namespace b200fhe
