/*
 * b200fhe.h — C ABI of the B200-native RNS polynomial hot path (libb200fhe.so).
 *
 * Drop-in boundary for the hot path of alexlee838/phantom-fhe-boot: the reference has no FFI; the
 * path sits behind C++ link-time symbols of libPhantom.so whose L1/L2 layer already takes raw
 * device pointers.  Each entry point below names the reference interface it replaces
 * (paths relative to the reference tree).  INTEGRATION.md shows the C++ shim that re-exports the
 * reference names on top of this ABI.
 *
 * Conventions
 *   - all data pointers are DEVICE pointers to uint64_t residues, layout [poly][limb][N] contiguous
 *     (reference PhantomCiphertext::data_, include/ciphertext.h), unless a name ends in _host;
 *   - `stream` is a cudaStream_t passed as void*; calls are asynchronous on it; a context's internal
 *     workspace is used by the composite calls (keyswitch/hmult/rotate/rescale), so composite calls
 *     on ONE context must be issued on one stream at a time;
 *   - size_Ql selects the level (number of ciphertext primes, reference DRNSTool of
 *     context_data[1 + levels_dropped]); prime index of limb t in a QlP buffer is
 *     t < size_Ql ? t : size_Q + (t - size_Ql);
 *   - return value: 0 = ok, negative = error; b200fhe_last_error() returns a thread-local message.
 *     (The reference throws std::invalid_argument / exits; the C++ shim maps codes to exceptions.)
 *   - no CPU fallback: every compute entry point fails with B200FHE_ERR_CUDA when no GPU is present.
 *   - a context belongs to the device that was current when it was created (b200fhe_ctx_create(device >= 0) makes that
 *     device current); the entry points of the key-switch / evaluator path check that the same device is current and fail
 *     with B200FHE_ERR_ARG otherwise; a context is used from one host thread at a time.
 */
#ifndef B200FHE_H
#define B200FHE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200FHE_OK 0
#define B200FHE_ERR_ARG (-1)
#define B200FHE_ERR_CUDA (-2)
#define B200FHE_ERR_ALLOC (-3)
#define B200FHE_ERR_UNSUPPORTED (-4)

typedef struct b200fhe_ctx b200fhe_ctx;

const char *b200fhe_last_error(void);
const char *b200fhe_version(void);
/* number of kernels launched by this library since process start (for bench.py's gpu_launches) */
uint64_t b200fhe_launch_count(void);

/* ---- parameter generation (host only; usable without a GPU) -------------------------------------
 * CoeffModulus::Create                       src/host/modulus.cu:80-111 */
int b200fhe_coeff_modulus_create(size_t N, const int *bit_sizes, size_t count, uint64_t *primes_out);

/* ---- context --------------------------------------------------------------------------------------
 * PhantomContext (src/context.cu:121-232): global DNTTTable over all QP primes + one DRNSTool per
 * level (src/rns.cu:11-195).  primes_QP = {q_0..q_{Q-1}, p_0..p_{P-1}}.  device < 0: current device. */
int b200fhe_ctx_create(int logN, const uint64_t *primes_QP, size_t size_QP, size_t size_P, int device,
                       b200fhe_ctx **out);
int b200fhe_ctx_destroy(b200fhe_ctx *ctx);
/* host-only twin (no device memory) so tables/constants can be checked on a CPU-only machine */
int b200fhe_ctx_create_host(int logN, const uint64_t *primes_QP, size_t size_QP, size_t size_P, b200fhe_ctx **out);
/* reference-layout tables, host copies: which = 0 twiddle, 1 twiddle_shoup, 2 itwiddle, 3 itwiddle_shoup
 * (each size_QP*N), 4 n_inv_mod_q, 5 n_inv_mod_q_shoup, 6 modulus value, 7 const_ratio[0], 8 const_ratio[1]
 * (each size_QP).  DNTTTable / DModulus, include/ntt.cuh:6-129 */
int b200fhe_ctx_export_table(const b200fhe_ctx *ctx, int which, uint64_t *host_out);
/* DRNSTool constants of one level, host copies: which = 0 partQlHatInv_mod_Ql_concat [Ql], 1 its shoup,
 * 2 bigP_mod_q [Ql], 3 bigPInv_mod_q [Ql], 4 bigPInv_mod_q_shoup [Ql], 5 P->Ql QHatInvModq [P], 6 its shoup,
 * 7 P->Ql QHatModp [Ql][P], 8 inv_q_last_mod_q [Ql-1], 9 its shoup, 100+d modup QHatModp of digit d
 * [QlP-|digit d|][|digit d|].  src/rns.cu:60-195 */
int b200fhe_level_export(b200fhe_ctx *ctx, size_t size_Ql, int which, uint64_t *host_out, size_t capacity,
                         size_t *count_out);
int b200fhe_level_beta(b200fhe_ctx *ctx, size_t size_Ql, size_t *beta_out);

/* ---- NTT ------------------------------------------------------------------------------------------
 * nwt_2d_radix8_forward_inplace                                   src/ntt/fntt_2d.cu:620
 * nwt_2d_radix8_forward_inplace_include_special_mod               src/ntt/fntt_2d.cu:695   (size_Ql > 0)
 * `limbs` limbs starting at limb index `start` of `npoly` polynomials spaced poly_stride (u64). */
int b200fhe_ntt_forward(b200fhe_ctx *ctx, uint64_t *inout, size_t limbs, size_t start, size_t size_Ql,
                        size_t npoly, size_t poly_stride, void *stream);
/* nwt_2d_radix8_backward_inplace                                  src/ntt/intt_2d.cu:724
 * nwt_2d_radix8_backward_inplace_include_special_mod              src/ntt/intt_2d.cu:796   (size_Ql > 0) */
int b200fhe_ntt_inverse(b200fhe_ctx *ctx, uint64_t *inout, size_t limbs, size_t start, size_t size_Ql,
                        size_t npoly, size_t poly_stride, void *stream);
/* nwt_2d_radix8_backward / _backward_scale / _backward_inplace_scale   src/ntt/ntt_modup.cu:320,356
 * out-of-place (out may equal in); scale/scale_shoup may be NULL; indexed by limb (start + i). */
int b200fhe_ntt_inverse_scale(b200fhe_ctx *ctx, uint64_t *out, const uint64_t *in, size_t limbs, size_t start,
                              size_t size_Ql, const uint64_t *scale, const uint64_t *scale_shoup, void *stream);
/* nwt_2d_radix8_forward_inplace_fuse_moddown                      src/ntt/ntt_moddown.cu:222
 * ct[i] = (cx[i] - NTT(delta)[i]) * PInv[i] mod q_i over `limbs` limbs from `start`; delta is clobbered. */
int b200fhe_ntt_forward_fuse_moddown(b200fhe_ctx *ctx, uint64_t *ct, const uint64_t *cx, const uint64_t *PInv,
                                     const uint64_t *PInv_shoup, uint64_t *delta, size_t limbs, size_t start,
                                     void *stream);

/* ---- key switching --------------------------------------------------------------------------------
 * DBaseConverter::bConv_BEHZ (P -> Ql of this level)              src/rns_bconv.cu:212-229 */
int b200fhe_bconv_P_to_Ql(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *dst, const uint64_t *src, void *stream);
/* DRNSTool::modup (CKKS)                                          src/rns_bconv.cu:530-628
 * c2 [Ql][N] (NTT) -> dst [beta][QlP][N] (NTT) */
int b200fhe_modup(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *dst, const uint64_t *c2, void *stream);
/* A single key switch (b200fhe_keyswitch / _hmult_relin / _rotate_fused and the evaluator calls built on them) runs the
 * independent chains of its modup (one per digit) and of its moddown (one per polynomial) on internal side streams, forked
 * from and joined to the caller's stream with events (capturable in a CUDA graph): same kernels, same numbers, fuller GPU.
 * on = 0 keeps every launch on the caller's stream.  Default on. */
int b200fhe_ctx_set_lanes(b200fhe_ctx *ctx, int on);
/* key_switch_inner_prod                                           src/eval_key_switch.cu:87-109
 * evk_digits: HOST array of `beta` DEVICE pointers, digit d -> [2][size_QP][N] (PhantomRelinKey layout,
 * include/secretkey.h:103-164) */
int b200fhe_inner_prod(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *cx, const uint64_t *t_mod_up,
                       const uint64_t *const *evk_digits, void *stream);
/* DRNSTool::moddown_from_NTT (CKKS)                               src/rns_bconv.cu:791-843
 * cx [QlP][N] (NTT; P part clobbered) -> ct [Ql][N]; ct may alias cx (eval_key_switch.cu:190) */
int b200fhe_moddown_from_ntt(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *ct, uint64_t *cx, void *stream);
/* keyswitch_inplace                                               src/eval_key_switch.cu:112-212
 * ct [2][Ql][N] += KeySwitch(c2 [Ql][N]) */
int b200fhe_keyswitch(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *ct, const uint64_t *c2,
                      const uint64_t *const *evk_digits, void *stream);

/* ---- evaluator ------------------------------------------------------------------------------------
 * bgv_ckks_multiply -> tensor_prod_2x2 / tensor_square_2x2        src/evaluate.cu:415, src/polymath.cu:501,538
 * out [3][L][N]; b == a selects the square kernel like the reference (evaluate.cu:437). out may alias a. */
int b200fhe_tensor(b200fhe_ctx *ctx, size_t L, uint64_t *out, const uint64_t *a, const uint64_t *b, void *stream);
/* multiply_and_relin_inplace                                      src/evaluate.cu:1220
 * a, b [2][L][N] -> out [2][L][N] (out may alias a) */
int b200fhe_hmult_relin(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *out, const uint64_t *a, const uint64_t *b,
                        const uint64_t *const *rlk_digits, void *stream);
/* DRNSTool::divide_and_round_q_last_ntt (rescale_to_next)         src/rns.cu:1160-1184
 * src [polys][L][N] -> dst [polys][L-1][N]; unlike the reference src is left untouched */
int b200fhe_rescale(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *dst, const uint64_t *src, size_t polys, void *stream);
/* The same rescale for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  NO COUNTERPART IN THE REFERENCE.
 * Allocates nothing (no cudaMalloc / cudaFree / allocating library call): the one scratch the operation needs, the
 * coefficient form of the last limb of each polynomial, is a static __device__ array of B200FHE_NOALLOC_MAX_WORDS
 * uint64_t (1 MiB, held by every GPU that loads the library until exit), and the forward transform runs in place in dst.
 * Limit: polys * N <= B200FHE_NOALLOC_MAX_WORDS (2 polynomials at N = 2^16), otherwise B200FHE_ERR_ARG; dst must not
 * alias src; the level's constants must already be resident (b200fhe_ctx_create builds all of them); the context's device
 * must be current.  Every call on a device shares the array (it keeps the previous call's contents; nothing in it has to
 * start at zero, the inverse transform overwrites every word it later reads): calls that overlap on the GPU — different
 * streams, or graphs replayed concurrently — race.  Output is bit-identical to b200fhe_rescale. */
#define B200FHE_NOALLOC_MAX_WORDS ((size_t) 131072)
int b200fhe_rescale_noalloc(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *dst, const uint64_t *src, size_t polys,
                            void *stream);
/* PrecomputeAutoMapKernel                                         src/util.cu:941-961 */
int b200fhe_automap_table(b200fhe_ctx *ctx, uint32_t k, uint32_t *perm_dev, void *stream);
/* PhantomGaloisTool permutation table for galois element (host)   include/galois.cuh:99-114 */
int b200fhe_galois_table_host(int logN, uint32_t galois_elt, uint32_t *perm_host);
/* apply_galois_ntt_permutation[_direct]                           src/galois.cu:11-18,104
 * dst[y][l][c] = src[y][l][perm[c]] ; dst must not alias src */
int b200fhe_apply_galois_ntt(b200fhe_ctx *ctx, uint64_t *dst, const uint64_t *src, const uint32_t *perm_dev,
                             size_t limbs, size_t npoly, size_t poly_stride, void *stream);
/* EvalRotateFused / EvalConjFused                                 src/evaluate.cu:2154,2242
 * key switch c1 with the key of automorphism index k, then automorphism of both polys.
 * ct [2][L][N] -> out [2][L][N] (no aliasing) */
int b200fhe_rotate_fused(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *out, const uint64_t *ct, uint32_t k,
                         const uint64_t *const *glk_digits, void *stream);
/* add_rns_poly / sub_rns_poly / multiply_rns_poly / negate        src/polymath.cu
 * op: 0 add, 1 sub, 2 mul, 3 negate (b ignored).  size_Ql selects the limb->prime map (0 = identity). */
int b200fhe_eltwise(b200fhe_ctx *ctx, int op, uint64_t *out, const uint64_t *a, const uint64_t *b, size_t limbs,
                    size_t size_Ql, void *stream);

/* ---- evaluator with FLEXIBLEAUTO bookkeeping (the OpenFHE-style layer the bootstrap is built from) ------------
 * include/evaluate.cuh:262-470, src/evaluate.cu:2154-4036.  A b200fhe_eval owns a stream-ordered block pool; every
 * call on one evaluator is issued on the stream given at creation.  Ciphertexts are opaque handles (the
 * reference's PhantomCiphertext: data + chain index + scale + NoiseScaleDeg); reference chain_index =
 * |Q| - size_Ql + 1.  Handles are value-like: operations never change another handle's content. */
typedef struct b200fhe_eval b200fhe_eval;
typedef struct b200fhe_ct b200fhe_ct;
typedef struct {
    size_t size_Ql;           /* ciphertext primes q_0..q_{size_Ql-1} */
    size_t limbs;             /* size_Ql, or size_Ql + |P| for an extended (Q_l P) ciphertext */
    uint32_t polys;
    uint32_t noise_scale_deg; /* PhantomCiphertext::GetNoiseScaleDeg */
    double scale;
} b200fhe_ct_desc;

int b200fhe_eval_create(b200fhe_ctx *ctx, void *stream, b200fhe_eval **out);
int b200fhe_eval_destroy(b200fhe_eval *ev);
/* PhantomCiphertext::PreComputeScale (include/ciphertext.h:320-366): which = 0 m_scalingFactorsReal [|Q|],
 * 1 m_scalingFactorsRealBig [|Q|-1] */
int b200fhe_eval_scaling_factors(b200fhe_eval *ev, int which, double *out, size_t capacity, size_t *count);
/* PhantomRelinKey / PhantomGaloisKeyFused::getKey(auto_index): HOST arrays of DEVICE digit pointers
 * ([2][size_QP][N] each, include/secretkey.h:103-164); the device memory stays owned by the caller */
int b200fhe_eval_set_relin_key(b200fhe_eval *ev, const uint64_t *const *digits, size_t ndigits);
int b200fhe_eval_set_galois_key(b200fhe_eval *ev, uint32_t auto_index, const uint64_t *const *digits, size_t ndigits);
/* FindAutomorphismIndex2nComplex (src/util.cu:908) */
uint32_t b200fhe_auto_index(int logN, int32_t rotation);

/* import copies [polys][limbs][N] device words into a pool block; export copies them out (capacity in u64 words) */
int b200fhe_ct_import(b200fhe_eval *ev, const uint64_t *dev, size_t size_Ql, uint32_t polys, int extended, double scale,
                      uint32_t noise_scale_deg, b200fhe_ct **out);
int b200fhe_ct_export(b200fhe_eval *ev, const b200fhe_ct *ct, uint64_t *dev_out, size_t capacity_words);
int b200fhe_ct_describe(const b200fhe_ct *ct, b200fhe_ct_desc *out);
int b200fhe_ct_clone(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out);
int b200fhe_ct_free(b200fhe_ct *ct);

/* EvalModReduceInPlace (rescale) / ModSwitchLevelInPlace (drop primes)          include/evaluate.cuh:272,304 */
int b200fhe_eval_mod_reduce(b200fhe_eval *ev, b200fhe_ct *ct, size_t levels);
int b200fhe_eval_mod_switch_level(b200fhe_eval *ev, b200fhe_ct *ct, size_t levels);
/* EvalMultConstInplace / EvalAddConstInPlaceWrap / MultByIntegerInPlace / MultByMonomialInPlace
 *                                                 include/evaluate.cuh:317,425; src/evaluate.cu:3942,2521 */
int b200fhe_eval_mult_const(b200fhe_eval *ev, b200fhe_ct *ct, double scalar);
int b200fhe_eval_add_const(b200fhe_eval *ev, b200fhe_ct *ct, double operand);
int b200fhe_eval_mult_by_integer(b200fhe_eval *ev, b200fhe_ct *ct, uint64_t v);
int b200fhe_eval_mult_by_monomial(b200fhe_eval *ev, b200fhe_ct *ct, uint32_t power);
/* EvalAddAutoInplace / EvalSubAutoInplace (a op= b)                              src/evaluate.cu:2856,2865 */
int b200fhe_eval_add_auto(b200fhe_eval *ev, b200fhe_ct *a, const b200fhe_ct *b);
int b200fhe_eval_sub_auto(b200fhe_eval *ev, b200fhe_ct *a, const b200fhe_ct *b);
/* EvalMultAuto / EvalSquare (multiply_and_relin_inplace inside)                  src/evaluate.cu:2794,3611 */
int b200fhe_eval_mult_auto(b200fhe_eval *ev, const b200fhe_ct *a, const b200fhe_ct *b, b200fhe_ct **out);
int b200fhe_eval_square(b200fhe_eval *ev, const b200fhe_ct *a, b200fhe_ct **out);
/* RaiseMod                                                                       src/evaluate.cu:2459 */
int b200fhe_eval_raise_mod(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out);
/* EvalRotateFused / EvalConjFused                                                src/evaluate.cu:2154,2242 */
int b200fhe_eval_rotate_fused(b200fhe_eval *ev, const b200fhe_ct *ct, int32_t index, b200fhe_ct **out);
int b200fhe_eval_conj_fused(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out);
/* EvalChebyshevSeries (Paterson-Stockmeyer for degree >= 5)                      src/evaluate.cu:3176 */
int b200fhe_eval_chebyshev(b200fhe_eval *ev, const b200fhe_ct *ct, const double *coeffs, size_t ncoeffs, double a,
                           double b, b200fhe_ct **out);
/* hoisting in the extended basis: EvalFastRotationPrecompute + EvalFastRotationExt, KeySwitchExt, KeySwitchDown,
 * KeySwitchDownFirstElement, EvalMultExtInPlace (plaintext [QlP][N] at the ciphertext's level, with its scale),
 * EvalAddExtInPlace                                                              src/evaluate.cu:3631-3940 */
int b200fhe_eval_fast_rotations_ext(b200fhe_eval *ev, const b200fhe_ct *ct, const int32_t *indices, size_t n,
                                    int add_first, b200fhe_ct **outs);
int b200fhe_eval_keyswitch_ext(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out);
int b200fhe_eval_keyswitch_down(b200fhe_eval *ev, const b200fhe_ct *ct, int first_only, b200fhe_ct **out);
int b200fhe_eval_mult_ext(b200fhe_eval *ev, b200fhe_ct *ct, const uint64_t *plain_dev, double plain_scale);
int b200fhe_eval_add_ext(b200fhe_eval *ev, b200fhe_ct *a, const b200fhe_ct *b);

/* ---- CKKS bootstrapping ------------------------------------------------------------------------------------
 * FHECKKSRNS (include/bootstrap.cuh:98-160, src/bootstrap.cu).  num_slots = 0: N/2 (fully packed);
 * correction_factor = 0: the reference's default rule; dim1 = 0: automatic giant step. */
typedef struct b200fhe_boot b200fhe_boot;
typedef struct b200fhe_encoder b200fhe_encoder;
int b200fhe_boot_create(b200fhe_ctx *ctx, void *stream, uint32_t budget_enc, uint32_t budget_dec, uint32_t dim1_enc,
                        uint32_t dim1_dec, uint32_t num_slots, uint32_t correction_factor, b200fhe_boot **out);
int b200fhe_boot_destroy(b200fhe_boot *bt);
/* the evaluator the bootstrap runs on (keys are set on it; ciphertext handles belong to it) */
b200fhe_eval *b200fhe_boot_evaluator(b200fhe_boot *bt);
/* GetCollapsedFFTParams (src/util.cu:768): which = 0 CoeffsToSlots, 1 SlotsToCoeffs; out[9] in CKKS_BOOT_PARAMS order */
int b200fhe_boot_params(b200fhe_boot *bt, int which, int32_t *out9);
/* FindBootstrapRotationIndices (src/bootstrap.cu:608): rotation keys EvalBootstrapKeyGen must provide (+ conjugation) */
int b200fhe_boot_rotation_indices(b200fhe_boot *bt, int32_t *out, size_t capacity, size_t *count);
/* shape of m_U0hatTPreFFT (which = 0) / m_U0PreFFT (which = 1): stages, diagonals per stage, |Q_l| of a stage */
int b200fhe_boot_stage_info(b200fhe_boot *bt, int which, size_t stage, size_t *n_stages, size_t *n_diagonals,
                            size_t *size_Ql);
/* one pre-encoded diagonal (EvalCoeffsToSlotsPrecompute / EvalSlotsToCoeffsPrecompute output, encode_ext):
 * DEVICE pointer [size_Ql + |P|][N], NTT form, caller-owned */
int b200fhe_boot_set_diagonal(b200fhe_boot *bt, int which, size_t stage, size_t index, const uint64_t *dev,
                              size_t size_Ql, double scale);
/* EvalBootstrapSetup with precompute = true (src/bootstrap.cu:88-177, EvalCoeffsToSlotsPrecompute :183,
 * EvalSlotsToCoeffsPrecompute :396): compute every diagonal on the host in double precision, encode it in the extended
 * basis with `enc` and keep it in device memory owned by the bootstrapper (29 GiB at the reference configuration). */
int b200fhe_boot_precompute(b200fhe_boot *bt, b200fhe_encoder *enc);
/* the diagonal currently installed at (which, stage, index): device pointer, |Q_l| and scale */
int b200fhe_boot_get_diagonal(b200fhe_boot *bt, int which, size_t stage, size_t index, const uint64_t **dev,
                              size_t *size_Ql, double *scale);
/* EvalBootstrap (src/bootstrap.cu:843), numIterations = 1 */
int b200fhe_boot_eval(b200fhe_boot *bt, const b200fhe_ct *ct, b200fhe_ct **out);
/* EvalBootstrap with numIterations > 1 (Meta-BTS, src/bootstrap.cu:858-910): bootstrap, bootstrap the error measured at
 * 2^precision times the message, subtract.  num_iterations <= 1 is b200fhe_boot_eval.  Like the reference, the input comes
 * back unchanged when the first bootstrap does not end above the input's level. */
int b200fhe_boot_eval_iter(b200fhe_boot *bt, const b200fhe_ct *ct, uint32_t num_iterations, uint32_t precision, b200fhe_ct **out);
/* EvalCoeffsToSlots / EvalSlotsToCoeffs alone (src/bootstrap.cu:1157,1407) */
int b200fhe_boot_coeffs_to_slots(b200fhe_boot *bt, const b200fhe_ct *ct, b200fhe_ct **out);
int b200fhe_boot_slots_to_coeffs(b200fhe_boot *bt, const b200fhe_ct *ct, b200fhe_ct **out);
/* ONE bootstrap on several GPUs: the linear-transform stages limb-sharded (no counterpart in the reference, which is
 * single-GPU).  Every rank calls b200fhe_boot_eval on the same input with the same keys and diagonals and gets
 * the same bits as the single-GPU call; inside a CoeffsToSlots / SlotsToCoeffs stage the hoisted giant-step rotations and the
 * plaintext-matrix product (HBM-bound, limb-wise) compute only the rank's rows of Q_l P, and the rows of the inner sums are
 * exchanged once per stage over NVLink peer memory by the store kernel of b200fhe_peer_exchange.  arena_bases[g]: rank g's
 * symmetric arena of b200fhe_boot_peer_arena_bytes() bytes, ZERO-INITIALISED, mapped into this process (entry [rank] is
 * the own one; two halves of inner sums used alternately, then the flag rows); counter: uint32_t[4] of own device memory,
 * zero-initialised.  virtual_ranks != 0: no arena and no exchange, this GPU computes the rows of all `world` ranks in turn
 * (single-GPU check of the row ranges).  world <= 1 switches back. */
size_t b200fhe_boot_peer_arena_bytes(b200fhe_boot *bt);
int b200fhe_boot_set_peer_shard(b200fhe_boot *bt, uint32_t world, uint32_t rank, void *const *arena_bases, size_t arena_bytes,
                                uint32_t *counter, int virtual_ranks);
/* EvalBootstrap of ONE input buffer captured as a CUDA graph (no counterpart in the reference; its EvalBootstrap issues every
 * kernel from the host each time).  capture: runs the bootstrap once eagerly (tables and workspaces reach their final size),
 * then once under stream capture on a block pool of its own; launch: one cudaGraphLaunch on the bootstrapper's stream.
 * The input is the buffer of `ct` as it is at capture ([2][size_Ql][N], NTT form): to bootstrap another ciphertext of the same
 * level, scale and NoiseScaleDeg, overwrite *input_limbs on the bootstrapper's stream before launching.  *output is a handle on
 * the graph's result buffer (rewritten by every launch; free it with b200fhe_ct_free).  While a graph exists the context's
 * key-switch workspaces are pinned: an operation that needs larger ones (a larger batch) fails instead of moving them.
 * With b200fhe_boot_set_peer_shard active the exchanges are part of the graph; every rank captures, then every rank launches. */
/* Tuning knob, process-wide: inverse transforms at N = 2^16 over at most max_limb_polys (limb, polynomial) pairs run as ONE
 * kernel, a cluster of 8 thread blocks per limb with the transpose between the row and the column stages through distributed
 * shared memory (one HBM round trip; replaces the pair nwt_2d_radix8_backward phase 1 / phase 2 for the few-limb launches
 * inside a key switch).  0 = always the two-pass kernels; < 0 = default (B200FHE_CLUSTER_NTT, else 0: measured slower on the B200, DESIGN.md
 * section 7).  Same bits either way. */
int b200fhe_tune_cluster_ntt(int max_limb_polys);
typedef struct b200fhe_boot_graph b200fhe_boot_graph;
int b200fhe_boot_graph_capture(b200fhe_boot *bt, const b200fhe_ct *ct, b200fhe_boot_graph **out);
int b200fhe_boot_graph_launch(b200fhe_boot_graph *g);
int b200fhe_boot_graph_io(b200fhe_boot_graph *g, uint64_t **input_limbs, b200fhe_ct **output);
int b200fhe_boot_graph_destroy(b200fhe_boot_graph *g);
/* bytes the evaluator's pool holds (for sizing) */
size_t b200fhe_eval_pool_bytes(b200fhe_eval *ev);

/* ---- CKKS encoder / decoder ----------------------------------------------------
 * Replaces PhantomCKKSEncoder (include/ckks.h:175-345, src/ckks.cu:45-669): special FFT (src/fft.cu),
 * DRNSBase::decompose_array / compose_array (src/rns_base.cu:49-260), then the NTT.  Values are complex numbers as
 * interleaved (re, im) doubles in HOST memory; plaintexts are DEVICE buffers [limbs][N] in NTT form. */
int b200fhe_encoder_create(b200fhe_ctx *ctx, b200fhe_encoder **out);
void b200fhe_encoder_destroy(b200fhe_encoder *enc);
size_t b200fhe_encoder_slots(const b200fhe_encoder *enc);
/* encode (ext = 0, sparse = 0), encode_ext (ext = 1: limbs of Q_l then the special primes), encode_sparse[_ext]
 * (sparse = 1: n a power of two, the message lives on every (slots/n)-th coefficient).  n <= slots; with sparse = 0 the
 * vector is zero-padded.  size_Ql = |Q| - chain_index + 1.  out_dev: [size_Ql (+|P|)][N]. */
int b200fhe_encode(b200fhe_encoder *enc, const double *values_re_im, size_t n, size_t size_Ql, double scale, int ext,
                   int sparse, uint64_t *out_dev, void *stream);
/* encode_sparse of `batch` messages of n values each (HOST [batch][n] complex) in one pass -> out_dev [batch][size_Ql][N];
 * the same residues as `batch` calls of b200fhe_encode(ext = 0, sparse = 1) (DNN::Conv encodes thousands of weight
 * plaintexts per layer, src/dnn.cu:122-136) */
int b200fhe_encode_sparse_batch(b200fhe_encoder *enc, const double *values_re_im, size_t n, size_t batch, size_t size_Ql,
                                double scale, uint64_t *out_dev, void *stream);
/* decode / decode_ext: out_re_im receives `slots` complex numbers (host); synchronises the stream */
int b200fhe_decode(b200fhe_encoder *enc, const uint64_t *plain_dev, size_t size_Ql, int ext, double scale,
                   double *out_re_im, void *stream);

/* ---- key generation, encryption, decryption ---------------------------------------
 * Replaces PhantomSecretKey (include/secretkey.h:232-420, src/secretkey.cu:196-530, 576-683, 956-1024),
 * PhantomPublicKey::encrypt_asymmetric (src/secretkey.cu:12-190) and the samplers of src/prng.cu:139-244.
 * Same distributions and key shapes; the randomness is ChaCha20 (RFC 8439) keyed by ONE 32-byte seed, a stream per
 * sampled polynomial.  All buffers are DEVICE memory in NTT form; work goes to the stream given at creation. */
typedef struct b200fhe_keygen b200fhe_keygen;
typedef struct b200fhe_evk b200fhe_evk; /* one evaluation key: dnum digits of [2][size_QP][N] */
/* PhantomSecretKey(context): draws the ternary secret.  seed32 == NULL: seeded from the operating system */
int b200fhe_keygen_create(b200fhe_ctx *ctx, const uint8_t *seed32, void *stream, b200fhe_keygen **out);
/* Same-seed key parity with the reference: a generator whose samplers ARE the reference's (src/prng.cu:17-244: Salsa20-core
 * blocks keyed by a 64-byte seed, nonce = coefficient index; ternary = byte mod 3, centred binomial from 42 coins, uniform by
 * rejection with whole-block regeneration) and whose 64-byte seeds are drawn in the reference's order (gen_secretkey,
 * gen_publickey, gen_relinkey, EvalAtIndexKeyGen: src/secretkey.cu:454-530, 956-1024) from the byte stream
 * top8(splitmix64(master_seed + i)) standing in for std::random_device.  With the reference's random_device answering from
 * the same stream, secret / public / relinearisation / rotation keys are bit-identical (tests/test_keygen_gpu.py).
 * Needs |P| to divide |Q| (the reference's dnum = |Q| / |P|).  Encryption in this mode is not provided. */
int b200fhe_keygen_create_reference_prng(b200fhe_ctx *ctx, uint64_t master_seed, void *stream, b200fhe_keygen **out);
void b200fhe_keygen_destroy(b200fhe_keygen *kg);
/* the secret key [size_QP][N] (device pointer owned by the generator) / adopt one generated elsewhere */
const uint64_t *b200fhe_keygen_secret(const b200fhe_keygen *kg);
int b200fhe_keygen_set_secret(b200fhe_keygen *kg, const uint64_t *secret_dev);
int b200fhe_keygen_get_secret(b200fhe_keygen *kg, uint64_t *secret_dev_out); /* copy out, on the generator's stream */
/* gen_relinkey (src/secretkey.cu:507-530) */
int b200fhe_keygen_relin_key(b200fhe_keygen *kg, b200fhe_evk **out);
/* one key of EvalAtIndexKeyGen (src/secretkey.cu:956-1024) for the automorphism x -> x^auto_index
 * (b200fhe_auto_index(logN, rotation); 2N - 1 for conjugation) */
int b200fhe_keygen_galois_key(b200fhe_keygen *kg, uint32_t auto_index, b200fhe_evk **out);
/* digit pointers of a key, for b200fhe_eval_set_relin_key / _set_galois_key / b200fhe_keyswitch; returns the count */
size_t b200fhe_evk_digits(const b200fhe_evk *key, const uint64_t **ptrs, size_t capacity);
void b200fhe_evk_free(b200fhe_evk *key);
/* gen_publickey (src/secretkey.cu:493-505): pk_dev_out [2][size_QP][N] */
int b200fhe_keygen_public_key(b200fhe_keygen *kg, uint64_t *pk_dev_out);
/* encrypt_symmetric / encrypt_asymmetric (CKKS): plain [size_Ql][N] -> ct [2][size_Ql][N]; scale and level are the
 * plaintext's, NoiseScaleDeg 1 */
int b200fhe_encrypt_symmetric(b200fhe_keygen *kg, const uint64_t *plain_dev, size_t size_Ql, uint64_t *ct_dev_out);
int b200fhe_encrypt_asymmetric(b200fhe_keygen *kg, const uint64_t *pk_dev, const uint64_t *plain_dev, size_t size_Ql,
                               uint64_t *ct_dev_out);
/* ckks_decrypt (src/secretkey.cu:646-682): plain = c0 + c1 s (+ c2 s^2); polys = 1..3 */
int b200fhe_decrypt(b200fhe_keygen *kg, const uint64_t *ct_dev, size_t size_Ql, uint32_t polys, uint64_t *plain_dev_out);
/* the samplers alone (tests): kind 0 ternary, 1 centred binomial (coefficient form, same value on every limb),
 * 2 uniform; out [limbs][N] over the first `limbs` primes (size_Ql == 0) or over Q_l then P */
int b200fhe_keygen_sample(b200fhe_keygen *kg, int kind, uint64_t *out_dev, size_t limbs, size_t size_Ql);
/* known-answer hook: one ChaCha20 block (key 8 words, counter, nonce 3 words) computed ON THE DEVICE -> 16 host words */
int b200fhe_chacha20_block(b200fhe_ctx *ctx, const uint32_t *key8, uint32_t counter, const uint32_t *nonce3,
                           uint32_t *out16_host);

/* ---- serialization: the reference's byte formats, through caller-owned HOST buffers ----
 * A record written here loads with the reference's load() and the other way round (tests/test_serial_gpu.py).
 * Saving with out_host == NULL only reports the size in *written.  Device copies are synchronous. */
typedef struct {
    uint64_t chain_index;       /* |Q| - size_Ql + 1 for a data-level ciphertext, 0 for key material */
    uint64_t polys;             /* size_ */
    uint64_t N;                 /* poly_modulus_degree_ */
    uint64_t limbs;             /* coeff_modulus_size_ */
    double scale;
    uint64_t correction_factor;
    uint64_t noise_scale_deg;
    uint8_t is_ntt_form, is_asymmetric;
} b200fhe_ct_header;
/* PhantomCiphertext::save / load (include/ciphertext.h:184-225); also PhantomPublicKey (include/secretkey.h:84-97) */
int b200fhe_ct_save(const b200fhe_ct_header *hdr, const uint64_t *dev, uint8_t *out_host, size_t capacity, size_t *written);
int b200fhe_ct_load(const uint8_t *in_host, size_t len, b200fhe_ct_header *hdr, uint64_t *dev_out, size_t capacity_words,
                    size_t *consumed);
/* PhantomPlaintext::save / load (include/plaintext.h:91-122) */
int b200fhe_plain_save(uint64_t chain_index, uint64_t N, uint64_t limbs, double scale, const uint64_t *dev, uint8_t *out_host,
                       size_t capacity, size_t *written);
int b200fhe_plain_load(const uint8_t *in_host, size_t len, uint64_t *chain_index, uint64_t *N, uint64_t *limbs, double *scale,
                       uint64_t *dev_out, size_t capacity_words, size_t *consumed);
/* PhantomRelinKey::save / load (include/secretkey.h:129-163): digits are HOST arrays of DEVICE pointers, each digit
 * [2][size_QP][N].  Loading with digits_out == NULL reports dnum and the record length only. */
int b200fhe_evk_save(b200fhe_ctx *ctx, const uint64_t *const *digits, size_t dnum, uint8_t *out_host, size_t capacity,
                     size_t *written);
int b200fhe_evk_load(b200fhe_ctx *ctx, const uint8_t *in_host, size_t len, uint64_t *const *digits_out, size_t capacity,
                     size_t *dnum_out, size_t *consumed);
/* PhantomSecretKey::save / load (include/secretkey.h): the first power of s, [size_QP][N] */
int b200fhe_secret_save(b200fhe_ctx *ctx, const uint64_t *secret_dev, uint8_t *out_host, size_t capacity, size_t *written);
int b200fhe_secret_load(b200fhe_ctx *ctx, const uint8_t *in_host, size_t len, uint64_t *secret_dev_out, size_t *consumed);

/* ---- limb-sharded key switch (no counterpart in the reference) -------------------------------
 * ONE key switch spread over G GPUs by RNS limb: rank g owns the ciphertext limbs [q_lo, q_hi) of Q_l and the special
 * limbs [p_lo, p_hi) of P, computes those rows of every intermediate, and exchanges rows with the other ranks only where
 * a base conversion mixes limbs:  begin -> all-gather the rows of tcks -> middle -> all-gather the P rows (limbs
 * size_Ql + [0, |P|)) of both polynomials of cx -> end.  All buffers have the FULL single-GPU layout (tcks, c2:
 * [size_Ql][N]; cx: [2][size_Ql + |P|][N]; acc, out: [2][size_Ql][N]) and a rank touches only its rows (plus what the
 * exchange filled in), so the union of the ranks' `out` rows is b200fhe_keyswitch's result bit for bit.  acc (optional) is
 * added to the first acc_polys polynomials (0 = both).  phantom-fhe-boot_b200/sharding.py drives it over torch.distributed. */
int b200fhe_shard_ks_begin(b200fhe_ctx *ctx, size_t size_Ql, size_t q_lo, size_t q_hi, const uint64_t *c2, uint64_t *tcks,
                           void *stream);
int b200fhe_shard_ks_middle(b200fhe_ctx *ctx, size_t size_Ql, size_t q_lo, size_t q_hi, size_t p_lo, size_t p_hi,
                            const uint64_t *tcks, const uint64_t *c2, const uint64_t *const *evk, uint64_t *cx, void *stream);
int b200fhe_shard_ks_end(b200fhe_ctx *ctx, size_t size_Ql, size_t q_lo, size_t q_hi, const uint64_t *cx, const uint64_t *acc,
                         uint32_t acc_polys, uint64_t *out, void *stream);

/* begin / middle with the transfer fused into the producing kernel: the last pass of the inverse transform stores every
 * word of the rank's own rows of tcks (of the P rows of cx) into its own buffer AND at the same offset behind each of the
 * n_peers pointers (the other ranks' tcks / cx in NVLink peer memory).  Only the flag exchange is left for
 * b200fhe_peer_exchange (nseg = 0).  n_peers <= 8; N >= 2^12. */
int b200fhe_shard_ks_begin_mirrored(b200fhe_ctx *ctx, size_t size_Ql, size_t q_lo, size_t q_hi, const uint64_t *c2,
                                    uint64_t *tcks, uint64_t *const *peer_tcks, uint32_t n_peers, void *stream);
int b200fhe_shard_ks_middle_mirrored(b200fhe_ctx *ctx, size_t size_Ql, size_t q_lo, size_t q_hi, size_t p_lo, size_t p_hi,
                                     const uint64_t *tcks, const uint64_t *c2, const uint64_t *const *evk, uint64_t *cx,
                                     uint64_t *const *peer_cx, uint32_t n_peers, void *stream);

/* Row exchange of the limb-sharded key switch over NVLink PEER MEMORY instead of NCCL.  The exchanged buffers live at the
 * same offsets of a symmetric arena on every rank; peer_bases[g] is rank g's arena base mapped into this process (CUDA IPC
 * or fabric handle; entry [rank] is ignored), peer_flags[g] rank g's flag row (uint64_t[world], zero-initialised, inside
 * or outside the arena; entry [rank] is this rank's own row), counter a zeroed uint32_t in local device memory.  ONE
 * launch stores the nseg byte ranges [seg_offset, +seg_bytes) of the local arena (16-byte aligned; the rows this rank
 * owns) into every peer's arena, publishes `epoch` to the peers and waits until every peer has published it: when the
 * kernel ends this rank holds everybody's rows.  epoch must grow by one per call, identically on all ranks; epoch == 0
 * numbers the calls on the device instead (counter is then uint32_t[4], 16-byte aligned, zeroed once), which is what a
 * captured CUDA graph of the key switch needs to be replayable.  world <= 9.  world == 1 is a no-op.  A peer that never
 * arrives does not hang the GPU: after 20 s the wait is given up and counter[1] is incremented (the rows of that exchange
 * are then incomplete; a caller that cannot rule out a dead peer checks counter[1] == 0 after synchronising). */
int b200fhe_peer_exchange(const void *local_base, void *const *peer_bases, uint64_t *const *peer_flags, uint32_t *counter,
                          uint32_t world, uint32_t rank, const size_t *seg_offset, const size_t *seg_bytes, uint32_t nseg,
                          uint64_t epoch, void *stream);

/* The same with the data stores going ONCE to the arena's NVSwitch multicast address (cuMulticast* binding of the same
 * arena on all ranks; torch's symmetric memory exposes it as multicast_ptr): the switch replicates every store into all
 * arenas, so a rank's NVLink egress is its rows once, not once per peer.  multicast_base == NULL: b200fhe_peer_exchange. */
int b200fhe_peer_exchange_mc(const void *local_base, void *const *peer_bases, void *multicast_base,
                             uint64_t *const *peer_flags, uint32_t *counter, uint32_t world, uint32_t rank,
                             const size_t *seg_offset, const size_t *seg_bytes, uint32_t nseg, uint64_t epoch, void *stream);

/* ---- batched ciphertexts (no counterpart in the reference) -----------------------------------------------
 * n independent ciphertexts with identical level, scale and NoiseScaleDeg become ONE handle whose operations run as one
 * launch set ([batch][polys][limbs][N]): b200fhe_eval_mult_auto / _square / _mod_reduce / _mult_const / _add_const /
 * _add_auto / _sub_auto / _mult_by_integer / _mult_by_monomial / _chebyshev accept it (rotations and the bootstrap take
 * single ciphertexts).  Each unpacked result equals the one the operation gives on the single ciphertext. */
int b200fhe_ct_pack(b200fhe_eval *ev, const b200fhe_ct *const *cts, size_t n, b200fhe_ct **out);
int b200fhe_ct_unpack(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **outs, size_t capacity, size_t *count);

/* ---- SEAL-style evaluator surface --------------------------------------------------
 * The reference's non-FLEXIBLEAUTO entry points (src/evaluate.cu:95-410, 1183-1590, 1700-2149; examples/3_ckks.cu):
 * the caller manages scales and levels.  Plaintexts are DEVICE buffers [size_Ql][N] in NTT form at the ciphertext's level.
 * Rotations here apply the automorphism FIRST and then key-switch with a PhantomGaloisKey-style key of the Galois element
 * g (sigma_g(s) -> s, PhantomSecretKey::create_galois_keys, src/secretkey.cu:532-574) -- not the fused keys above. */
int b200fhe_eval_set_seal_galois_key(b200fhe_eval *ev, uint32_t galois_elt, const uint64_t *const *digits, size_t ndigits);
int b200fhe_keygen_galois_key_seal(b200fhe_keygen *kg, uint32_t galois_elt, b200fhe_evk **out);
int b200fhe_eval_negate(b200fhe_eval *ev, b200fhe_ct *ct);                                   /* negate_inplace */
/* add_plain_inplace (sub = 0) / sub_plain_inplace (sub = 1): scales must agree (are_close, 1e-6 relative) */
int b200fhe_eval_add_plain(b200fhe_eval *ev, b200fhe_ct *ct, const uint64_t *plain_dev, double plain_scale, int sub);
int b200fhe_eval_multiply_plain(b200fhe_eval *ev, b200fhe_ct *ct, const uint64_t *plain_dev, double plain_scale);
int b200fhe_eval_add_many(b200fhe_eval *ev, const b200fhe_ct *const *cts, size_t n, b200fhe_ct **out);
int b200fhe_eval_multiply(b200fhe_eval *ev, const b200fhe_ct *a, const b200fhe_ct *b, b200fhe_ct **out); /* 3 polys */
int b200fhe_eval_relinearize(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out);
int b200fhe_eval_rescale_to_next(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out);
int b200fhe_eval_mod_switch_to_next(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out);
int b200fhe_eval_apply_galois(b200fhe_eval *ev, const b200fhe_ct *ct, uint32_t galois_elt, b200fhe_ct **out);
/* rotate_inplace: the step's own key if present, else the NAF decomposition over the keys present */
int b200fhe_eval_rotate(b200fhe_eval *ev, const b200fhe_ct *ct, int step, b200fhe_ct **out);
/* hoisting_inplace: sum over `steps` of rotate(ct, step) with one modup and one moddown */
int b200fhe_eval_hoisting(b200fhe_eval *ev, const b200fhe_ct *ct, const int *steps, size_t n, b200fhe_ct **out);

/* ---- encrypted CNN layers ---------------------------------------------------------
 * Replaces phantom::DNN / TensorCT (include/dnn.cuh:33-80, src/dnn.cu).  A tensor is an array of ciphertext handles,
 * one per channel, plus its descriptor: channel k holds a width x width image on the (width << slotstr)^2 sparse
 * slots, pixel (i, j) in slot ((i * L + j) << slotstr), L = width << slotstr.  All layers run on ONE evaluator (the
 * bootstrapper's when bootstraps are involved: b200fhe_boot_evaluator); the rotation keys of
 * b200fhe_dnn_add_rotation_indices / _add_avgpool_rotations (index 0 included: the reference key-switches the centre
 * tap with the identity-automorphism key) and the relinearisation key must be set on it.  Every output limb equals
 * the reference's.  Outputs are NEW handles (b200fhe_ct_free); capacity counts handles in `outs`. */
typedef struct b200fhe_dnn b200fhe_dnn;
typedef struct {
    size_t width, slotstr, num_ch;
    size_t num_cts; /* handles in the array; 0 = num_ch (AvgPoolFullCon returns more handles than channels, see below) */
} b200fhe_tensor_desc;
/* DNN::DNN(scale, encoder, context) (include/dnn.cuh:44) */
int b200fhe_dnn_create(b200fhe_eval *ev, b200fhe_encoder *enc, double scale, b200fhe_dnn **out);
void b200fhe_dnn_destroy(b200fhe_dnn *d);
/* DNN::RelinKeyGen / BuildGaloisKey (src/dnn.cu:332-337, 501-504): in the reference the layers own key material that is
 * separate from the bootstrapper's.  Keys set here serve the layers' own multiplications and rotations; bootstraps keep
 * the evaluator's keys.  Without these calls the evaluator's keys serve both.  Digits: DEVICE pointers, caller-owned. */
int b200fhe_dnn_set_relin_key(b200fhe_dnn *d, const uint64_t *const *digits, size_t ndigits);
int b200fhe_dnn_set_galois_key(b200fhe_dnn *d, uint32_t auto_index, const uint64_t *const *digits, size_t ndigits);
/* Lanes (no counterpart in the reference): the independent channels of ReluComposite / BootStrap run concurrently on n
 * extra streams, channel i on lane i % n.  One sparse bootstrap fills a fraction of a B200; a lane is a private workspace,
 * pool and stream sharing the key and diagonal memory of `bt`.  Call after all keys and diagonals are installed; the
 * results are the same limbs.  n = 0 switches lanes off. */
int b200fhe_dnn_set_lanes(b200fhe_dnn *d, b200fhe_boot *bt, size_t n);
/* ONE image on several GPUs (no counterpart in the reference): every rank runs the same network on the same encrypted input
 * with the same keys; in ReluComposite / BootStrap rank r computes the channels [n r / world, n (r + 1) / world) of the layer
 * and the results are all-gathered over NVLink peer memory (the store kernel of b200fhe_peer_exchange); in Conv rank r rotates and
 * multiplies its run of INPUT channels and the partial sums of every output channel are reduced across the ranks (all-to-all,
 * row sum, all-gather; exact), so every rank goes on with the complete layer and the single-GPU limbs.  arena_bases[g]: rank g's
 * arena mapped into this process, zero-initialised, arena_bytes = 2 halves (alternate layers; each >= 2 x channels x 2 x |Q_l| x N x 8
 * of the widest layer: partial-sum slots + sums) + 256 bytes of flag rows;
 * counter: uint32_t[4], own device memory, zero-initialised.  world <= 1 switches it off. */
int b200fhe_dnn_set_peer_ranks(b200fhe_dnn *d, uint32_t world, uint32_t rank, void *const *arena_bases, size_t arena_bytes,
                               uint32_t *counter);
/* Channel batches (no counterpart in the reference, which walks the channels one ciphertext at a time, src/dnn.cu:198-279): up to
 * n channels of a layer go through ReluComposite / BootStrap as one batched ciphertext, every key switch, rescale and polynomial
 * step one launch set over all of them; with lanes, each lane takes batches of its channels.  Same limbs per channel.
 * Default 4 (environment B200FHE_DNN_BATCH); n <= 1: one channel at a time. */
int b200fhe_dnn_set_batch(b200fhe_dnn *d, size_t n);
/* DNN::AddRotationIndicesTo / AddAvgPoolRotationsTo (src/dnn.cu:283-330): inout holds *count indices on entry, the new
 * ones are appended; inout == NULL only counts */
int b200fhe_dnn_add_rotation_indices(int32_t *inout, size_t capacity, size_t *count, size_t width, size_t kernel_h,
                                     size_t slotstr);
int b200fhe_dnn_add_avgpool_rotations(int32_t *inout, size_t capacity, size_t *count, size_t width, size_t slotstr);
/* DNN::EncTensor / DecTensor (src/dnn.cu:10-76): image and out are HOST arrays [width][width][num_ch] */
int b200fhe_dnn_enc_tensor(b200fhe_dnn *d, b200fhe_keygen *kg, const uint64_t *pk_dev, const double *image, size_t width,
                           size_t num_ch, b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *out_desc);
int b200fhe_dnn_dec_tensor(b200fhe_dnn *d, b200fhe_keygen *kg, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc,
                           double *out);
/* DNN::Conv (src/dnn.cu:82-150): weight HOST [kernel_h][kernel_h][in_ch][out_ch], stride 1 or 2 */
int b200fhe_dnn_conv(b200fhe_dnn *d, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc, const double *weight,
                     size_t kernel_h, size_t out_ch, size_t stride, b200fhe_ct **outs, size_t capacity,
                     b200fhe_tensor_desc *out_desc);
/* DNN::BatchNorm (src/dnn.cu:449-476): per-channel arrays of num_ch doubles, eps = 1e-5 */
int b200fhe_dnn_batch_norm(b200fhe_dnn *d, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc, const double *weight,
                           const double *bias, const double *mean, const double *var, b200fhe_ct **outs, size_t capacity,
                           b200fhe_tensor_desc *out_desc);
/* DNN::Add (src/dnn.cu:478-499) */
int b200fhe_dnn_add(b200fhe_dnn *d, const b200fhe_ct *const *a, const b200fhe_tensor_desc *a_desc, const b200fhe_ct *const *b,
                    const b200fhe_tensor_desc *b_desc, b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *out_desc);
/* DNN::Relu (Chebyshev interpolation of max(0,x) on [a,b]), ReluDegTwo (x + x^2), Sign (stage k = 0..2 of the composite
 * sign polynomial) (src/dnn.cu:152-196, 250-264) */
int b200fhe_dnn_relu(b200fhe_dnn *d, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc, int a, int b, int deg,
                     b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *out_desc);
int b200fhe_dnn_relu_deg_two(b200fhe_dnn *d, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc, b200fhe_ct **outs,
                             size_t capacity, b200fhe_tensor_desc *out_desc);
int b200fhe_dnn_sign(b200fhe_dnn *d, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc, size_t k, b200fhe_ct **outs,
                     size_t capacity, b200fhe_tensor_desc *out_desc);
/* DNN::ReluComposite / BootStrap (src/dnn.cu:198-246, 266-279): the bootstrapper must have been created for
 * (width << slotstr)^2 slots and `d` must run on its evaluator */
int b200fhe_dnn_relu_composite(b200fhe_dnn *d, b200fhe_boot *bt, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc,
                               b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *out_desc);
int b200fhe_dnn_bootstrap(b200fhe_dnn *d, b200fhe_boot *bt, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc,
                          b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *out_desc);
/* DNN::AvgPoolFullCon (src/dnn.cu:397-447): weight HOST [T][t], bias [T].  As in the reference the result array holds t
 * handles (out_desc->num_cts = t, num_ch = T): its u == 0 pass appends one ciphertext per input channel. */
int b200fhe_dnn_avgpool_fullcon(b200fhe_dnn *d, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc, const double *weight,
                                size_t T, size_t t, const double *bias, b200fhe_ct **outs, size_t capacity,
                                b200fhe_tensor_desc *out_desc);
/* DNN::SoftMax (src/dnn.cu:339-395) */
int b200fhe_dnn_softmax(b200fhe_dnn *d, b200fhe_boot *bt, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc,
                        size_t BoundB, size_t BoundR, size_t GumbelLambda, size_t GoldSchmidtD, b200fhe_ct **outs, size_t capacity,
                        b200fhe_tensor_desc *out_desc);

/* ---- ResNet-20 driver (Resnet/models/model_resnet20.cu:3-153: PrePareResNet20, ResNet20_infer) ------------------
 * conv1-bn1-relu, three stages of `blocks` residual blocks (stages 2, 3 open with a stride-2 block whose skip path is a
 * 1x1 stride-2 convolution + batch norm), AvgPoolFullCon.  The reference reads .npy files that are not in its tree; here
 * the parameters are ONE flat HOST array in the order the reference loads them: conv1 [3][3][in][base], bn1; per block
 * conv1, bn1, conv2, bn2 (, downsample conv [1][1][in][out], downsample bn); fc weight [classes][ch], fc bias.
 * Convolutions in the layers' [kh][kw][in][out] layout, a batch norm = weight, bias, running_mean, running_var.
 * The reference network is (in_ch 3, base_ch 16, blocks 3, classes 10) on a 32 x 32 image at 1024 sparse slots.
 * layer_ms4 (optional): host milliseconds spent in conv / batch norm + add / ReluComposite / pool + fc. */
size_t b200fhe_resnet20_param_count(size_t in_ch, size_t base_ch, size_t blocks, size_t classes);
int b200fhe_resnet20_rotation_indices(size_t width, int32_t *out, size_t capacity, size_t *count);
int b200fhe_resnet20_infer(b200fhe_dnn *d, b200fhe_boot *bt, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc,
                           const double *params, size_t nparams, size_t base_ch, size_t blocks, size_t classes, b200fhe_ct **outs,
                           size_t capacity, b200fhe_tensor_desc *out_desc, double *layer_ms4);

/* ---- entries for a C++ shim over the reference's own types (shim/) ----------------------------
 * DNTTTable (include/ntt.cuh:34-129): a context on twiddle tables the caller already holds on the DEVICE (each
 * [size_QP][N], ninv [size_QP]); the moduli come as a HOST array.  Interleaved (w, w_shoup) copies are made on
 * `stream`, which is synchronised before returning.  size_P may be 0 and set later (once, before any level is used)
 * with b200fhe_ctx_set_special: a DNTTTable does not know how many of its primes are special. */
int b200fhe_ctx_create_from_tables(int logN, const uint64_t *primes_QP_host, size_t size_QP, size_t size_P,
                                   const uint64_t *tw, const uint64_t *tw_shoup, const uint64_t *itw,
                                   const uint64_t *itw_shoup, const uint64_t *ninv, const uint64_t *ninv_shoup,
                                   void *stream, b200fhe_ctx **out);
int b200fhe_ctx_set_special(b200fhe_ctx *ctx, size_t size_P);
int b200fhe_ctx_info(const b200fhe_ctx *ctx, int *logN, size_t *size_QP, size_t *size_P, int *device);
/* The reference's launcher variants (include/ntt.cuh:174-226) as ONE limb -> prime description:
 *   limbs [start, start + limbs);  limb t uses prime t when special_from == 0 or t < special_from,
 *   else prime special_base + (t - special_from)   (include_special_mod: special_from = start + size - size_P,
 *   special_base = size_QP - size_P; include_temp_mod: special_from = size - 1, special_base = total - 1);
 *   limbs [excl_start, excl_end) are skipped by a forward transform (exclude_range). */
typedef struct {
    size_t start, limbs;
    size_t special_from, special_base;
    size_t excl_start, excl_end;
} b200fhe_limb_range;
/* out-of-place forms (out may equal in); limb t of `in` lands in limb t of `out`.  scale / scale_shoup (inverse, optional)
 * are indexed by limb t like nwt_2d_radix8_backward_scale (src/ntt/intt_2d.cu:308) */
int b200fhe_ntt_forward_ex(b200fhe_ctx *ctx, uint64_t *out, const uint64_t *in, const b200fhe_limb_range *r, void *stream);
int b200fhe_ntt_inverse_ex(b200fhe_ctx *ctx, uint64_t *out, const uint64_t *in, const b200fhe_limb_range *r,
                           const uint64_t *scale, const uint64_t *scale_shoup, void *stream);
/* key_switch_inner_prod / keyswitch_inplace with the digit pointers in DEVICE memory, as the reference passes them
 * (PhantomRelinKey::public_keys_ptr(), include/secretkey.h:124; src/eval_key_switch.cu:87,112) */
int b200fhe_inner_prod_devkeys(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *cx, const uint64_t *t_mod_up,
                               const uint64_t *const *evk_digits_dev, void *stream);
int b200fhe_keyswitch_devkeys(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *ct, const uint64_t *c2,
                              const uint64_t *const *evk_digits_dev, void *stream);

/* ---- host-buffer entry points (the reference-facing e2e path: H2D, compute, D2H inside) ------------ */
int b200fhe_ntt_forward_host(b200fhe_ctx *ctx, uint64_t *inout_host, size_t limbs, size_t start, size_t size_Ql);
int b200fhe_ntt_inverse_host(b200fhe_ctx *ctx, uint64_t *inout_host, size_t limbs, size_t start, size_t size_Ql);
/* benchmark/ntt_bench.cu:84-152 shape: `npoly` polynomials of `limbs` limbs in (preferably pinned) host memory,
 * [npoly][limbs][N]; ops: 1 forward, 2 inverse, 3 forward then inverse.  Copies and transforms are pipelined
 * (3 staging slots, separate copy-in / compute / copy-out streams); returns after the last copy-out. */
int b200fhe_ntt_host_batch(b200fhe_ctx *ctx, uint64_t *inout_host, size_t limbs, size_t size_Ql, size_t npoly, int ops);

/* the composite operations on ciphertexts held in HOST memory (preferably pinned): copy in, compute, copy out and
 * synchronise inside the call; the keys stay device-resident as in the reference.  [2][size_Ql][N] in and out. */
int b200fhe_hmult_relin_host(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *out_host, const uint64_t *a_host,
                             const uint64_t *b_host, const uint64_t *const *rlk_digits, void *stream);
int b200fhe_rotate_fused_host(b200fhe_ctx *ctx, size_t size_Ql, uint64_t *out_host, const uint64_t *ct_host, uint32_t k,
                              const uint64_t *const *glk_digits, void *stream);
/* EvalBootstrap (src/bootstrap.cu:843) on a host ciphertext; the result [2][out_desc->size_Ql][N] goes to out_host */
int b200fhe_boot_eval_host(b200fhe_boot *bt, const uint64_t *ct_host, size_t size_Ql, double scale, uint32_t noise_scale_deg,
                           uint64_t *out_host, size_t capacity_words, b200fhe_ct_desc *out_desc);

/* ---- op trace ("each C-ABI call logs its algorithmic bytes/ops") ----------------------------------
 * Every kernel launcher adds the HBM bytes its launch must move by the algorithm it implements (operands once, results
 * once), its Harvey butterflies and its 64x64->128 multiply-accumulates to process-wide counters.
 * out4 = {hbm_bytes, butterflies, macs, kernel-launcher calls} since the last reset. */
void b200fhe_trace_reset(void);
void b200fhe_trace_read(uint64_t out4[4]);

#ifdef __cplusplus
}
#endif
#endif
