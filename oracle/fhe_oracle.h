/*
 * fhe_oracle.h — CPU restatement (plain C) of the RNS polynomial hot path of
 * alexlee838/phantom-fhe-boot (PhantomFHE + CKKS bootstrapping).
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product
 * (phantom-fhe-boot_b200/) never links, imports or calls anything in oracle/.
 *
 * Parity pinning: the reference ships NO golden vectors for this path (SURVEY.md §8c:
 * test/ntt_test.cu only checks INTT(NTT(x)) == x).  The oracle is pinned two ways:
 *   (1) tables/constants: against the reference's own SEAL-derived host code compiled from
 *       /root/reference/src/host/*.cu (oracle/_ref/libphantom_hostref.so; tests/test_oracle_pin.py),
 *       with the resulting vectors committed under tests/golden/;
 *   (2) transforms / key switch: against the reference's GPU kernels compiled unmodified for
 *       sm_100a (oracle/_ref/libphantom_ref.so, oracle/Makefile.ref) and run on the B200 box
 *       (tests/test_ref_gpu_parity.py, -m gpu), plus independent mathematical identities
 *       (O(N^2) negacyclic evaluation, schoolbook convolution, exact CRT checks).
 *
 * Every function cites the reference file:line whose arithmetic it restates
 * (paths relative to /root/reference).
 */
#ifndef FHE_ORACLE_H
#define FHE_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------- scalar arithmetic */
uint64_t orc_mulmod(uint64_t a, uint64_t b, uint64_t q);
uint64_t orc_powmod(uint64_t a, uint64_t e, uint64_t q);
uint64_t orc_invmod(uint64_t a, uint64_t q);           /* q prime */
uint64_t orc_shoup(uint64_t w, uint64_t q);            /* floor(w*2^64/q), include/host/uintarithsmallmod.h:119 */
void     orc_barrett_ratio(uint64_t q, uint64_t ratio[2]); /* floor(2^128/q), src/host/modulus.cu:29-38 */
int      orc_is_prime(uint64_t q);

/* ---------------------------------------------------------------- parameter generation */
/* src/host/numth.cu:207-233  get_primes */
int orc_get_primes(size_t ntt_size, int bit_size, size_t count, uint64_t *out);
/* src/host/modulus.cu:80-111 CoeffModulus::Create */
int orc_coeff_modulus_create(size_t N, const int *bit_sizes, size_t count, uint64_t *out);
/* src/host/numth.cu:309-332 try_minimal_primitive_root (deterministic: the minimum is unique) */
uint64_t orc_minimal_primitive_root(uint64_t degree, uint64_t q);
/* src/host/ntt.cu:11-56 NTT::NTT — reference DNTTTable layout for one prime
 * (bit-reversed psi powers, Shoup companions, itw[1] pre-multiplied by N^-1). */
void orc_ntt_tables(int logN, uint64_t q, uint64_t *tw, uint64_t *tw_shoup, uint64_t *itw,
                    uint64_t *itw_shoup, uint64_t *ninv, uint64_t *ninv_shoup);

/* ---------------------------------------------------------------- transforms (one limb) */
/* src/ntt/fntt_2d.cu:9-198 + include/butterfly.cuh:10-24 : natural in -> bit-reversed out, [0,q) */
void orc_ntt_fwd(uint64_t *a, int logN, uint64_t q, const uint64_t *tw, const uint64_t *tw_shoup);
/* src/ntt/intt_2d.cu:9-207 + include/butterfly.cuh:30-37 : inverse incl. N^-1, [0,q) */
void orc_ntt_inv(uint64_t *a, int logN, uint64_t q, const uint64_t *itw, const uint64_t *itw_shoup,
                 uint64_t ninv, uint64_t ninv_shoup);
/* O(N^2) definition: out[brv(k)] = sum_j a[j] * psi^{(2k+1) j}  (independent cross-check) */
void orc_ntt_fwd_naive(const uint64_t *a, uint64_t *out, int logN, uint64_t q, uint64_t psi);
/* schoolbook negacyclic product c = a*b mod (X^N+1, q) */
void orc_negacyclic_mul_naive(const uint64_t *a, const uint64_t *b, uint64_t *c, size_t N, uint64_t q);

/* ---------------------------------------------------------------- context (all QP primes) */
typedef struct orc_ctx orc_ctx;
/* primes_QP = {q_0..q_{Q-1}, p_0..p_{P-1}} as PhantomContext orders them (src/context.cu:121-232) */
orc_ctx *orc_ctx_create(int logN, const uint64_t *primes_QP, size_t size_QP, size_t size_P);
void     orc_ctx_destroy(orc_ctx *c);
size_t   orc_ctx_N(const orc_ctx *c);
const uint64_t *orc_ctx_tables(const orc_ctx *c, int which); /* 0 tw 1 tw_shoup 2 itw 3 itw_shoup 4 ninv 5 ninv_shoup 6 primes 7 ratio0 8 ratio1 */

/* batched transforms with the reference's limb -> prime maps:
 * prime index of limb t (t = start + i):  t < size_Ql ? t : size_Q + (t - size_Ql)
 * (twr_idx2 in src/ntt/fntt_2d.cu:434-436).  size_Ql == 0 means "plain": prime index = t. */
void orc_ntt_fwd_limbs(const orc_ctx *c, uint64_t *data, size_t limbs, size_t start, size_t size_Ql);
void orc_ntt_inv_limbs(const orc_ctx *c, uint64_t *data, size_t limbs, size_t start, size_t size_Ql);

/* ---------------------------------------------------------------- level constants (DRNSTool, src/rns.cu:97-195) */
typedef struct {
    size_t size_Ql, size_P, size_Q, alpha, beta;
    uint64_t *partQlHatInv_mod_Ql;        /* [size_Ql]            src/rns.cu:176-178 */
    uint64_t *partQlHatInv_mod_Ql_shoup;  /* [size_Ql] */
    uint64_t *modup_QHatModp;             /* beta matrices concatenated; digit d: [obase_d][ibase_d] */
    size_t   *modup_off;                  /* [beta+1] offsets into modup_QHatModp */
    uint64_t *bigP_mod_q, *bigP_mod_q_shoup;        /* [size_Ql]  src/rns.cu:113-125 */
    uint64_t *bigPInv_mod_q, *bigPInv_mod_q_shoup;  /* [size_Ql] */
    uint64_t *PHatInv_mod_p, *PHatInv_mod_p_shoup;  /* [size_P]   RNSBase(P).QHatInvModq */
    uint64_t *PHat_mod_q;                 /* [size_Ql][size_P]    BaseConverter(P->Ql).QHatModp, src/host/rns.cu:449-458 */
    uint64_t *inv_q_last_mod_q, *inv_q_last_mod_q_shoup; /* [size_Ql-1] src/rns.cu:66-81 */
} orc_level;
orc_level *orc_level_create(const orc_ctx *c, size_t size_Ql);
void       orc_level_destroy(orc_level *l);

/* ---------------------------------------------------------------- key-switch pipeline */
/* DBaseConverter::bConv_BEHZ, src/rns_bconv.cu:212-229 (generic ibase -> obase, approximate) */
void orc_bconv(uint64_t *dst, const uint64_t *src, size_t N, const uint64_t *ibase, size_t ni,
               const uint64_t *obase, size_t no);
/* DRNSTool::modup (CKKS branch), src/rns_bconv.cu:530-628: c2[size_Ql][N] (NTT) -> dst[beta][QlP][N] (NTT) */
void orc_modup(const orc_ctx *c, const orc_level *l, uint64_t *dst, const uint64_t *c2);
/* key_switch_inner_prod_c2_and_evk, src/eval_key_switch.cu:26-85.  evk[d] -> [2][size_QP][N] */
void orc_inner_prod(const orc_ctx *c, const orc_level *l, uint64_t *cx, const uint64_t *t_mod_up,
                    const uint64_t *const *evk);
/* DRNSTool::moddown_from_NTT (CKKS branch), src/rns_bconv.cu:791-843; cx[QlP][N] is clobbered (P part INTT'd) */
void orc_moddown_from_ntt(const orc_ctx *c, const orc_level *l, uint64_t *ct, uint64_t *cx);
/* keyswitch_inplace, src/eval_key_switch.cu:112-212: ct[2][size_Ql][N] += KS(c2) */
void orc_keyswitch(const orc_ctx *c, const orc_level *l, uint64_t *ct, const uint64_t *c2,
                   const uint64_t *const *evk);

/* ---------------------------------------------------------------- evaluator pieces */
/* tensor_prod_2x2_rns_poly, src/polymath.cu:501-536: out[3][L][N] */
void orc_tensor_prod(const orc_ctx *c, size_t L, const uint64_t *a, const uint64_t *b, uint64_t *out);
/* tensor_square_2x2_rns_poly, src/polymath.cu:538-570 */
void orc_tensor_square(const orc_ctx *c, size_t L, const uint64_t *a, uint64_t *out);
/* DRNSTool::divide_and_round_q_last_ntt, src/rns.cu:1160-1184: src[polys][L][N] -> dst[polys][L-1][N]; src clobbered like the reference */
void orc_rescale(const orc_ctx *c, const orc_level *l, uint64_t *src, size_t polys, uint64_t *dst);
/* permutation table of apply_galois_ntt: include/galois.cuh:99-114 */
void orc_galois_table(int logN, uint32_t galois_elt, uint32_t *table);
/* PrecomputeAutoMapKernel, src/util.cu:941-961 */
void orc_automap_table(int logN, uint32_t k, uint32_t *table);
/* apply_galois_ntt_permutation, src/galois.cu:11-18: dst[l][i] = src[l][table[i]] */
void orc_apply_galois_ntt(const uint64_t *src, uint64_t *dst, size_t limbs, size_t N, const uint32_t *table);
/* element-wise helpers over L limbs with primes 0..L-1 (src/polymath.cu add/sub/multiply_rns_poly) */
void orc_add(const orc_ctx *c, size_t L, const uint64_t *a, const uint64_t *b, uint64_t *out);
void orc_sub(const orc_ctx *c, size_t L, const uint64_t *a, const uint64_t *b, uint64_t *out);
void orc_mul(const orc_ctx *c, size_t L, const uint64_t *a, const uint64_t *b, uint64_t *out);

/* whole ops on the CPU (used as the timed cpu_baseline): multiply_and_relin_inplace, src/evaluate.cu:1220 */
void orc_hmult_relin(const orc_ctx *c, const orc_level *l, const uint64_t *a, const uint64_t *b,
                     const uint64_t *const *rlk, uint64_t *out /* [2][size_Ql][N] */);
/* EvalRotateFused order (src/evaluate.cu:2154): key switch c1 first, then automorphism of both polys */
void orc_rotate_fused(const orc_ctx *c, const orc_level *l, const uint64_t *ct, const uint64_t *const *glk,
                      const uint32_t *table, uint64_t *out);

int orc_set_threads(int n); /* OpenMP threads over limbs for the batched functions (0 = all cores) */

#ifdef __cplusplus
}
#endif
#endif
