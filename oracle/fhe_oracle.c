/*
 * fhe_oracle.c — CPU restatement of the reference's RNS polynomial hot path.  See fhe_oracle.h.
 * TEST INFRASTRUCTURE ONLY (parity oracle + timed CPU baseline).  Plain C, unsigned __int128.
 */
#include "fhe_oracle.h"

#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;

static int g_threads = 0;
int orc_set_threads(int n) {
#ifdef _OPENMP
    g_threads = n > 0 ? n : omp_get_num_procs();
    omp_set_num_threads(g_threads);
    return g_threads;
#else
    (void) n;
    g_threads = 1;
    return 1;
#endif
}

/* ------------------------------------------------------------------ scalar arithmetic */

uint64_t orc_mulmod(uint64_t a, uint64_t b, uint64_t q) { return (uint64_t) (((u128) a * b) % q); }

uint64_t orc_powmod(uint64_t a, uint64_t e, uint64_t q) {
    uint64_t r = 1;
    a %= q;
    while (e) {
        if (e & 1) r = orc_mulmod(r, a, q);
        a = orc_mulmod(a, a, q);
        e >>= 1;
    }
    return r;
}

uint64_t orc_invmod(uint64_t a, uint64_t q) { return orc_powmod(a, q - 2, q); }

/* include/host/uintarithsmallmod.h:119-124 */
uint64_t orc_shoup(uint64_t w, uint64_t q) { return (uint64_t) (((u128) w << 64) / q); }

/* src/host/modulus.cu:29-38: quotient of 2^128 / q, low and high words */
void orc_barrett_ratio(uint64_t q, uint64_t ratio[2]) {
    /* floor(2^128/q) = floor((2^128-1)/q) unless q | 2^128 (q odd > 1: never) */
    u128 r = (~(u128) 0) / q;
    ratio[0] = (uint64_t) r;
    ratio[1] = (uint64_t) (r >> 64);
}

/* deterministic Miller-Rabin for 64-bit (the reference's is_prime is probabilistic MR, numth.cu:150-206;
 * primality itself is a mathematical fact so any correct test selects the same primes) */
int orc_is_prime(uint64_t n) {
    static const uint64_t bases[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    if (n < 2) return 0;
    for (size_t i = 0; i < 12; i++) {
        if (n == bases[i]) return 1;
        if (n % bases[i] == 0) return 0;
    }
    uint64_t d = n - 1;
    int r = 0;
    while (!(d & 1)) { d >>= 1; r++; }
    for (size_t i = 0; i < 12; i++) {
        uint64_t x = orc_powmod(bases[i], d, n);
        if (x == 1 || x == n - 1) continue;
        int comp = 1;
        for (int j = 1; j < r; j++) {
            x = orc_mulmod(x, x, n);
            if (x == n - 1) { comp = 0; break; }
        }
        if (comp) return 0;
    }
    return 1;
}

/* ------------------------------------------------------------------ parameter generation */

/* src/host/numth.cu:207-233: descend from 2^bits - 2N + 1 in steps of 2N */
int orc_get_primes(size_t ntt_size, int bit_size, size_t count, uint64_t *out) {
    uint64_t factor = 2 * (uint64_t) ntt_size;
    uint64_t value = ((uint64_t) 1 << bit_size) - factor + 1;
    uint64_t lower = (uint64_t) 1 << (bit_size - 1);
    size_t k = 0;
    while (k < count && value > lower) {
        if (orc_is_prime(value)) out[k++] = value;
        value -= factor;
    }
    return k == count ? 0 : -1;
}

/* src/host/modulus.cu:96-110: per bit size a descending prime list; each request pops from the BACK */
int orc_coeff_modulus_create(size_t N, const int *bit_sizes, size_t count, uint64_t *out) {
    size_t cnt[64] = {0};
    uint64_t *tab[64] = {0};
    int rc = 0;
    for (size_t i = 0; i < count; i++) {
        if (bit_sizes[i] < 2 || bit_sizes[i] > 61) return -2;
        cnt[bit_sizes[i]]++;
    }
    for (int b = 0; b < 64; b++) {
        if (!cnt[b]) continue;
        tab[b] = (uint64_t *) malloc(cnt[b] * sizeof(uint64_t));
        if (orc_get_primes(N, b, cnt[b], tab[b])) rc = -1;
    }
    if (!rc)
        for (size_t i = 0; i < count; i++) {
            int b = bit_sizes[i];
            out[i] = tab[b][--cnt[b]];
        }
    for (int b = 0; b < 64; b++) free(tab[b]);
    return rc;
}

/* src/host/numth.cu:309-332.  The reference starts from a random primitive root and walks all
 * odd powers keeping the minimum; the minimum does not depend on the start, so we start from a
 * deterministic one. */
uint64_t orc_minimal_primitive_root(uint64_t degree, uint64_t q) {
    uint64_t group = (q - 1) / degree;
    if ((q - 1) % degree) return 0;
    uint64_t root = 0;
    for (uint64_t a = 2; a < 1000; a++) {
        uint64_t r = orc_powmod(a, group, q);
        if (orc_powmod(r, degree >> 1, q) == q - 1) { root = r; break; }
    }
    if (!root) return 0;
    uint64_t gsq = orc_mulmod(root, root, q);
    uint64_t cur = root, best = root;
    for (uint64_t i = 0; i < degree / 2; i++) {
        if (cur < best) best = cur;
        cur = orc_mulmod(cur, gsq, q);
    }
    return best;
}

static uint32_t brv(uint32_t x, int bits) {
    uint32_t r = 0;
    for (int i = 0; i < bits; i++) {
        r = (r << 1) | (x & 1);
        x >>= 1;
    }
    return r;
}

/* src/host/ntt.cu:11-56 */
void orc_ntt_tables(int logN, uint64_t q, uint64_t *tw, uint64_t *tw_shoup, uint64_t *itw,
                    uint64_t *itw_shoup, uint64_t *ninv, uint64_t *ninv_shoup) {
    size_t N = (size_t) 1 << logN;
    uint64_t psi = orc_minimal_primitive_root(2 * N, q);
    uint64_t ipsi = orc_invmod(psi, q);
    uint64_t p = psi, ip = ipsi;
    for (size_t i = 1; i < N; i++) {
        uint32_t r = brv((uint32_t) i, logN);
        tw[r] = p;
        tw_shoup[r] = orc_shoup(p, q);
        itw[r] = ip;
        itw_shoup[r] = orc_shoup(ip, q);
        p = orc_mulmod(p, psi, q);
        ip = orc_mulmod(ip, ipsi, q);
    }
    tw[0] = 1;
    tw_shoup[0] = orc_shoup(1, q);
    itw[0] = 1;
    itw_shoup[0] = orc_shoup(1, q);
    uint64_t ni = orc_invmod((uint64_t) N % q, q);
    *ninv = ni;
    *ninv_shoup = orc_shoup(ni, q);
    /* host/ntt.cu:53-55: the last inverse stage's twiddle carries N^-1 */
    itw[1] = orc_mulmod(itw[1], ni, q);
    itw_shoup[1] = orc_shoup(itw[1], q);
}

/* ------------------------------------------------------------------ butterflies (include/butterfly.cuh) */

static inline uint64_t mulhi64(uint64_t a, uint64_t b) { return (uint64_t) (((u128) a * b) >> 64); }

/* include/uintmodmath.cuh:18-21 */
static inline uint64_t csub(uint64_t x, uint64_t q) {
    uint64_t t = x - q;
    return t + (t >> 63) * q;
}

/* include/uintmodmath.cuh:222-231: result in [0,2q) */
static inline uint64_t shoup_lazy(uint64_t a, uint64_t w, uint64_t ws, uint64_t q) {
    return a * w - mulhi64(a, ws) * q;
}

/* include/uintmodmath.cuh:207-216 */
static inline uint64_t shoup_mul(uint64_t a, uint64_t w, uint64_t ws, uint64_t q) {
    return csub(shoup_lazy(a, w, ws, q), q);
}

/* include/butterfly.cuh:10-24 (Harvey CT, values in [0,4q)) */
static inline void ct_bfly(uint64_t *x, uint64_t *y, uint64_t w, uint64_t ws, uint64_t q) {
    uint64_t t = shoup_lazy(*y, w, ws, q);
    uint64_t q2 = 2 * q;
    uint64_t xx = csub(*x, q2);
    *y = xx + q2 - t;
    *x = xx + t;
}

/* include/butterfly.cuh:30-37 (GS) */
static inline void gs_bfly(uint64_t *x, uint64_t *y, uint64_t w, uint64_t ws, uint64_t q) {
    uint64_t q2 = 2 * q;
    uint64_t t = *x + q2 - *y;
    uint64_t s = csub(*x + *y, q2);
    *x = s;
    *y = shoup_lazy(t, w, ws, q);
}

/* Forward negacyclic NTT.  Stage order, twiddle indexing (tw[m+i]) and lazy ranges follow
 * src/ntt/fntt_2d.cu:9-198 (phase1 = first stages, phase2 = rest, final csub 2q then q at :185-190);
 * the 2-D radix-8 blocking of the reference only regroups these same butterflies. */
void orc_ntt_fwd(uint64_t *a, int logN, uint64_t q, const uint64_t *tw, const uint64_t *tw_shoup) {
    size_t N = (size_t) 1 << logN;
    for (size_t m = 1, gap = N >> 1; m < N; m <<= 1, gap >>= 1) {
        for (size_t i = 0; i < m; i++) {
            uint64_t w = tw[m + i], ws = tw_shoup[m + i];
            uint64_t *x = a + 2 * i * gap, *y = x + gap;
            for (size_t j = 0; j < gap; j++) ct_bfly(x + j, y + j, w, ws, q);
        }
    }
    for (size_t j = 0; j < N; j++) a[j] = csub(csub(a[j], 2 * q), q);
}

/* Inverse negacyclic NTT.  src/ntt/intt_2d.cu:9-207: GS stages with itw[m+i]; the last stage (m=1)
 * uses itw[1] which already carries N^-1 (host/ntt.cu:53-55) and the x outputs are multiplied by
 * N^-1 (:196-199); final csub q (:203). */
void orc_ntt_inv(uint64_t *a, int logN, uint64_t q, const uint64_t *itw, const uint64_t *itw_shoup,
                 uint64_t ninv, uint64_t ninv_shoup) {
    size_t N = (size_t) 1 << logN;
    for (size_t m = N >> 1, gap = 1; m >= 1; m >>= 1, gap <<= 1) {
        for (size_t i = 0; i < m; i++) {
            uint64_t w = itw[m + i], ws = itw_shoup[m + i];
            uint64_t *x = a + 2 * i * gap, *y = x + gap;
            for (size_t j = 0; j < gap; j++) gs_bfly(x + j, y + j, w, ws, q);
        }
    }
    size_t h = N >> 1;
    for (size_t j = 0; j < h; j++) {
        a[j] = csub(shoup_lazy(a[j], ninv, ninv_shoup, q), q);
        a[j + h] = csub(a[j + h], q);
    }
}

void orc_ntt_fwd_naive(const uint64_t *a, uint64_t *out, int logN, uint64_t q, uint64_t psi) {
    size_t N = (size_t) 1 << logN;
    for (size_t k = 0; k < N; k++) {
        uint64_t wk = orc_powmod(psi, 2 * k + 1, q);
        uint64_t acc = 0, p = 1;
        for (size_t j = 0; j < N; j++) {
            acc = (uint64_t) (((u128) a[j] * p + acc) % q);
            p = orc_mulmod(p, wk, q);
        }
        out[brv((uint32_t) k, logN)] = acc;
    }
}

void orc_negacyclic_mul_naive(const uint64_t *a, const uint64_t *b, uint64_t *c, size_t N, uint64_t q) {
    for (size_t k = 0; k < N; k++) {
        uint64_t pos = 0, neg = 0;
        for (size_t i = 0; i < N; i++) {
            size_t j = (k + N - i) % N;
            uint64_t p = orc_mulmod(a[i], b[j], q);
            if (i <= k) pos = (pos + p) % q; else neg = (neg + p) % q;
        }
        c[k] = (pos + q - neg) % q;
    }
}

/* ------------------------------------------------------------------ context */

struct orc_ctx {
    int logN;
    size_t N, size_QP, size_P, size_Q;
    uint64_t *primes, *ratio0, *ratio1;
    uint64_t *tw, *tw_shoup, *itw, *itw_shoup, *ninv, *ninv_shoup;
};

orc_ctx *orc_ctx_create(int logN, const uint64_t *primes_QP, size_t size_QP, size_t size_P) {
    orc_ctx *c = (orc_ctx *) calloc(1, sizeof(orc_ctx));
    size_t N = (size_t) 1 << logN;
    c->logN = logN; c->N = N; c->size_QP = size_QP; c->size_P = size_P; c->size_Q = size_QP - size_P;
    c->primes = (uint64_t *) malloc(size_QP * 8);
    c->ratio0 = (uint64_t *) malloc(size_QP * 8);
    c->ratio1 = (uint64_t *) malloc(size_QP * 8);
    c->tw = (uint64_t *) malloc(size_QP * N * 8);
    c->tw_shoup = (uint64_t *) malloc(size_QP * N * 8);
    c->itw = (uint64_t *) malloc(size_QP * N * 8);
    c->itw_shoup = (uint64_t *) malloc(size_QP * N * 8);
    c->ninv = (uint64_t *) malloc(size_QP * 8);
    c->ninv_shoup = (uint64_t *) malloc(size_QP * 8);
    memcpy(c->primes, primes_QP, size_QP * 8);
#pragma omp parallel for schedule(dynamic)
    for (size_t i = 0; i < size_QP; i++) {
        uint64_t r[2];
        orc_barrett_ratio(primes_QP[i], r);
        c->ratio0[i] = r[0]; c->ratio1[i] = r[1];
        orc_ntt_tables(logN, primes_QP[i], c->tw + i * N, c->tw_shoup + i * N, c->itw + i * N,
                       c->itw_shoup + i * N, c->ninv + i, c->ninv_shoup + i);
    }
    return c;
}

void orc_ctx_destroy(orc_ctx *c) {
    if (!c) return;
    free(c->primes); free(c->ratio0); free(c->ratio1); free(c->tw); free(c->tw_shoup);
    free(c->itw); free(c->itw_shoup); free(c->ninv); free(c->ninv_shoup); free(c);
}

size_t orc_ctx_N(const orc_ctx *c) { return c->N; }

const uint64_t *orc_ctx_tables(const orc_ctx *c, int which) {
    switch (which) {
        case 0: return c->tw; case 1: return c->tw_shoup; case 2: return c->itw; case 3: return c->itw_shoup;
        case 4: return c->ninv; case 5: return c->ninv_shoup; case 6: return c->primes;
        case 7: return c->ratio0; case 8: return c->ratio1; default: return NULL;
    }
}

/* twr_idx2 of src/ntt/fntt_2d.cu:434-436 */
static inline size_t prime_of_limb(const orc_ctx *c, size_t t, size_t size_Ql) {
    if (size_Ql == 0) return t;
    return t < size_Ql ? t : c->size_Q + (t - size_Ql);
}

void orc_ntt_fwd_limbs(const orc_ctx *c, uint64_t *data, size_t limbs, size_t start, size_t size_Ql) {
#pragma omp parallel for schedule(dynamic)
    for (size_t i = 0; i < limbs; i++) {
        size_t t = start + i, p = prime_of_limb(c, t, size_Ql);
        orc_ntt_fwd(data + t * c->N, c->logN, c->primes[p], c->tw + p * c->N, c->tw_shoup + p * c->N);
    }
}

void orc_ntt_inv_limbs(const orc_ctx *c, uint64_t *data, size_t limbs, size_t start, size_t size_Ql) {
#pragma omp parallel for schedule(dynamic)
    for (size_t i = 0; i < limbs; i++) {
        size_t t = start + i, p = prime_of_limb(c, t, size_Ql);
        orc_ntt_inv(data + t * c->N, c->logN, c->primes[p], c->itw + p * c->N, c->itw_shoup + p * c->N,
                    c->ninv[p], c->ninv_shoup[p]);
    }
}

/* ------------------------------------------------------------------ level constants (src/rns.cu:11-195) */

static uint64_t prod_mod_except(const uint64_t *base, size_t n, size_t skip, uint64_t q) {
    uint64_t r = 1 % q;
    for (size_t k = 0; k < n; k++)
        if (k != skip) r = orc_mulmod(r, base[k] % q, q);
    return r;
}

static inline size_t digit_size(const orc_level *l, size_t d) {
    return d == l->beta - 1 ? l->size_Ql - l->alpha * (l->beta - 1) : l->alpha;
}

orc_level *orc_level_create(const orc_ctx *c, size_t size_Ql) {
    orc_level *l = (orc_level *) calloc(1, sizeof(orc_level));
    size_t P = c->size_P, Q = c->size_Q, alpha = P;
    l->size_Ql = size_Ql; l->size_P = P; l->size_Q = Q; l->alpha = alpha;
    l->beta = (size_Ql + alpha - 1) / alpha; /* ceil, src/rns.cu:147 */
    const uint64_t *q = c->primes, *p = c->primes + Q;
    size_t QlP = size_Ql + P;
    uint64_t *base_QlP = (uint64_t *) malloc(QlP * 8);
    memcpy(base_QlP, q, size_Ql * 8);
    memcpy(base_QlP + size_Ql, p, P * 8);

    l->partQlHatInv_mod_Ql = (uint64_t *) malloc(size_Ql * 8);
    l->partQlHatInv_mod_Ql_shoup = (uint64_t *) malloc(size_Ql * 8);
    l->modup_off = (size_t *) malloc((l->beta + 1) * sizeof(size_t));
    size_t tot = 0;
    for (size_t d = 0; d < l->beta; d++) {
        l->modup_off[d] = tot;
        size_t ni = digit_size(l, d);
        tot += ni * (QlP - ni);
    }
    l->modup_off[l->beta] = tot;
    l->modup_QHatModp = (uint64_t *) malloc((tot ? tot : 1) * 8);
    for (size_t d = 0; d < l->beta; d++) {
        size_t s = alpha * d, ni = digit_size(l, d), no = QlP - ni;
        const uint64_t *ib = q + s;
        for (size_t i = 0; i < ni; i++) {
            /* RNSBase(part).QHatInvModq, src/rns.cu:164-168 */
            uint64_t hat = prod_mod_except(ib, ni, i, ib[i]);
            uint64_t inv = orc_invmod(hat, ib[i]);
            l->partQlHatInv_mod_Ql[s + i] = inv;
            l->partQlHatInv_mod_Ql_shoup[s + i] = orc_shoup(inv, ib[i]);
        }
        /* complement base = QlP with [s, s+ni) erased (src/rns.cu:158-162); matrix [obase][ibase] (host/rns.cu:449-458) */
        uint64_t *M = l->modup_QHatModp + l->modup_off[d];
        for (size_t j = 0; j < no; j++) {
            uint64_t pj = base_QlP[j < s ? j : j + ni];
            for (size_t i = 0; i < ni; i++) M[j * ni + i] = prod_mod_except(ib, ni, i, pj);
        }
    }
    l->bigP_mod_q = (uint64_t *) malloc(size_Ql * 8);
    l->bigP_mod_q_shoup = (uint64_t *) malloc(size_Ql * 8);
    l->bigPInv_mod_q = (uint64_t *) malloc(size_Ql * 8);
    l->bigPInv_mod_q_shoup = (uint64_t *) malloc(size_Ql * 8);
    for (size_t i = 0; i < size_Ql; i++) {
        uint64_t pm = prod_mod_except(p, P, (size_t) -1, q[i]);
        l->bigP_mod_q[i] = pm;
        l->bigP_mod_q_shoup[i] = orc_shoup(pm, q[i]);
        uint64_t pi = orc_invmod(pm, q[i]);
        l->bigPInv_mod_q[i] = pi;
        l->bigPInv_mod_q_shoup[i] = orc_shoup(pi, q[i]);
    }
    l->PHatInv_mod_p = (uint64_t *) malloc(P * 8);
    l->PHatInv_mod_p_shoup = (uint64_t *) malloc(P * 8);
    for (size_t i = 0; i < P; i++) {
        uint64_t inv = orc_invmod(prod_mod_except(p, P, i, p[i]), p[i]);
        l->PHatInv_mod_p[i] = inv;
        l->PHatInv_mod_p_shoup[i] = orc_shoup(inv, p[i]);
    }
    l->PHat_mod_q = (uint64_t *) malloc(size_Ql * P * 8);
    for (size_t j = 0; j < size_Ql; j++)
        for (size_t i = 0; i < P; i++) l->PHat_mod_q[j * P + i] = prod_mod_except(p, P, i, q[j]);
    size_t nl = size_Ql > 1 ? size_Ql - 1 : 1;
    l->inv_q_last_mod_q = (uint64_t *) malloc(nl * 8);
    l->inv_q_last_mod_q_shoup = (uint64_t *) malloc(nl * 8);
    for (size_t i = 0; i + 1 < size_Ql; i++) {
        uint64_t inv = orc_invmod(q[size_Ql - 1] % q[i], q[i]);
        l->inv_q_last_mod_q[i] = inv;
        l->inv_q_last_mod_q_shoup[i] = orc_shoup(inv, q[i]);
    }
    free(base_QlP);
    return l;
}

void orc_level_destroy(orc_level *l) {
    if (!l) return;
    free(l->partQlHatInv_mod_Ql); free(l->partQlHatInv_mod_Ql_shoup); free(l->modup_QHatModp);
    free(l->modup_off); free(l->bigP_mod_q); free(l->bigP_mod_q_shoup); free(l->bigPInv_mod_q);
    free(l->bigPInv_mod_q_shoup); free(l->PHatInv_mod_p); free(l->PHatInv_mod_p_shoup);
    free(l->PHat_mod_q); free(l->inv_q_last_mod_q); free(l->inv_q_last_mod_q_shoup); free(l);
}

/* ------------------------------------------------------------------ Barrett (include/uintmodmath.cuh:96-136) */

static inline uint64_t barrett128(u128 x, uint64_t q, uint64_t r0, uint64_t r1) {
    uint64_t lo = (uint64_t) x, hi = (uint64_t) (x >> 64);
    /* floor(x * ratio / 2^128), same partial-product sum as the PTX sequence at :108-124 */
    uint64_t carry = mulhi64(lo, r0);
    u128 t1 = (u128) lo * r1 + carry;
    u128 t2 = (u128) hi * r0 + (uint64_t) t1;
    uint64_t s = hi * r1 + (uint64_t) (t1 >> 64) + (uint64_t) (t2 >> 64);
    return csub(lo - s * q, q);
}

/* include/uintmodmath.cuh:144-151 */
static inline uint64_t barrett64(uint64_t x, uint64_t q, uint64_t r1) {
    return csub(x - mulhi64(r1, x) * q, q);
}

/* ------------------------------------------------------------------ base conversion */

/* generic BEHZ conversion used for cross-checks: bconv_mult (src/rns_bconv.cu:22-30) then
 * bconv_matmul (:109-131) with constants derived from the bases. */
void orc_bconv(uint64_t *dst, const uint64_t *src, size_t N, const uint64_t *ibase, size_t ni,
               const uint64_t *obase, size_t no) {
    uint64_t *tmp = (uint64_t *) malloc(ni * N * 8);
    for (size_t i = 0; i < ni; i++) {
        uint64_t inv = orc_invmod(prod_mod_except(ibase, ni, i, ibase[i]), ibase[i]);
        uint64_t invs = orc_shoup(inv, ibase[i]);
        for (size_t k = 0; k < N; k++) tmp[i * N + k] = shoup_mul(src[i * N + k], inv, invs, ibase[i]);
    }
    for (size_t j = 0; j < no; j++) {
        uint64_t r[2];
        orc_barrett_ratio(obase[j], r);
        uint64_t *row = (uint64_t *) malloc(ni * 8);
        for (size_t i = 0; i < ni; i++) row[i] = prod_mod_except(ibase, ni, i, obase[j]);
        for (size_t k = 0; k < N; k++) {
            u128 acc = 0;
            for (size_t i = 0; i < ni; i++) acc += (u128) tmp[i * N + k] * row[i];
            dst[j * N + k] = barrett128(acc, obase[j], r[0], r[1]);
        }
        free(row);
    }
    free(tmp);
}

/* DRNSTool::modup, CKKS branch, alpha > 1 and alpha == 1 (src/rns_bconv.cu:530-628) */
void orc_modup(const orc_ctx *c, const orc_level *l, uint64_t *dst, const uint64_t *c2) {
    size_t N = c->N, Ql = l->size_Ql, P = l->size_P, QlP = Ql + P, alpha = l->alpha;
    uint64_t *t = (uint64_t *) malloc(Ql * N * 8);
    /* nwt_2d_radix8_backward_scale (src/ntt/ntt_modup.cu:356): INTT then x partQlHatInv (Shoup) */
#pragma omp parallel for schedule(dynamic)
    for (size_t i = 0; i < Ql; i++) {
        uint64_t q = c->primes[i];
        memcpy(t + i * N, c2 + i * N, N * 8);
        orc_ntt_inv(t + i * N, c->logN, q, c->itw + i * N, c->itw_shoup + i * N, c->ninv[i], c->ninv_shoup[i]);
        if (alpha != 1) /* alpha == 1: "no need to multiply qiHatInv_mod_qi" (:551-553) */
            for (size_t k = 0; k < N; k++)
                t[i * N + k] = shoup_mul(t[i * N + k], l->partQlHatInv_mod_Ql[i], l->partQlHatInv_mod_Ql_shoup[i], q);
    }
    for (size_t d = 0; d < l->beta; d++) {
        size_t s = alpha * d, ni = digit_size(l, d), no = QlP - ni;
        uint64_t *out = dst + d * QlP * N;
        const uint64_t *M = l->modup_QHatModp + l->modup_off[d];
#pragma omp parallel for schedule(dynamic)
        for (size_t j = 0; j < no; j++) {
            size_t pos = j < s ? j : j + ni; /* padded_out_prime_idx, :479 */
            size_t pidx = prime_of_limb(c, pos, Ql);
            uint64_t pj = c->primes[pidx];
            uint64_t *o = out + pos * N;
            if (alpha == 1) {
                /* modup_bconv_single_p_kernel :432-453 */
                uint64_t qin = c->primes[s];
                for (size_t k = 0; k < N; k++) {
                    uint64_t v = t[s * N + k];
                    o[k] = qin > pj ? barrett64(v, pj, c->ratio1[pidx]) : v;
                }
            } else {
                for (size_t k = 0; k < N; k++) {
                    u128 acc = 0;
                    for (size_t i = 0; i < ni; i++) acc += (u128) t[(s + i) * N + k] * M[j * ni + i];
                    o[k] = barrett128(acc, pj, c->ratio0[pidx], c->ratio1[pidx]);
                }
            }
            /* forward NTT with the special-prime map, own digit excluded (:618-621) */
            orc_ntt_fwd(o, c->logN, pj, c->tw + pidx * N, c->tw_shoup + pidx * N);
        }
        /* own digit copied in NTT form (modup_copy_partQl_kernel :522-528) */
        memcpy(out + s * N, c2 + s * N, ni * N * 8);
    }
    free(t);
}

/* key_switch_inner_prod_c2_and_evk, src/eval_key_switch.cu:26-85 */
void orc_inner_prod(const orc_ctx *c, const orc_level *l, uint64_t *cx, const uint64_t *t_mod_up,
                    const uint64_t *const *evk) {
    size_t N = c->N, Ql = l->size_Ql, QlP = Ql + l->size_P, QP = c->size_QP;
#pragma omp parallel for schedule(dynamic)
    for (size_t t = 0; t < QlP; t++) {
        size_t twr = prime_of_limb(c, t, Ql);
        uint64_t q = c->primes[twr], r0 = c->ratio0[twr], r1 = c->ratio1[twr];
        for (size_t k = 0; k < N; k++) {
            u128 a0 = 0, a1 = 0;
            for (size_t d = 0; d < l->beta; d++) {
                uint64_t v = t_mod_up[(d * QlP + t) * N + k];
                a0 += (u128) v * evk[d][twr * N + k];
                a1 += (u128) v * evk[d][(QP + twr) * N + k];
            }
            cx[t * N + k] = barrett128(a0, q, r0, r1);
            cx[(QlP + t) * N + k] = barrett128(a1, q, r0, r1);
        }
    }
}

/* DRNSTool::moddown_from_NTT, CKKS branch (src/rns_bconv.cu:791-843) */
void orc_moddown_from_ntt(const orc_ctx *c, const orc_level *l, uint64_t *ct, uint64_t *cx) {
    size_t N = c->N, Ql = l->size_Ql, P = l->size_P, Q = l->size_Q;
    /* INTT of the P limbs (:803) */
    orc_ntt_inv_limbs(c, cx, P, Ql, Ql);
    uint64_t *tmp = (uint64_t *) malloc(P * N * 8);
    const uint64_t *p = c->primes + Q;
    if (P != 1)
        for (size_t i = 0; i < P; i++) /* bconv_mult_unroll2_kernel :40-60 */
            for (size_t k = 0; k < N; k++)
                tmp[i * N + k] = shoup_mul(cx[(Ql + i) * N + k], l->PHatInv_mod_p[i], l->PHatInv_mod_p_shoup[i], p[i]);
#pragma omp parallel for schedule(dynamic)
    for (size_t j = 0; j < Ql; j++) {
        uint64_t q = c->primes[j];
        uint64_t *delta = (uint64_t *) malloc(N * 8);
        if (P == 1) { /* moddown_bconv_single_p_kernel :691-707 */
            for (size_t k = 0; k < N; k++) {
                uint64_t v = cx[Ql * N + k];
                delta[k] = p[0] > q ? barrett64(v, q, c->ratio1[j]) : v;
            }
        } else { /* bconv_matmul_unroll2_kernel :143-168 */
            for (size_t k = 0; k < N; k++) {
                u128 acc = 0;
                for (size_t i = 0; i < P; i++) acc += (u128) tmp[i * N + k] * l->PHat_mod_q[j * P + i];
                delta[k] = barrett128(acc, q, c->ratio0[j], c->ratio1[j]);
            }
        }
        /* nwt_2d_radix8_forward_inplace_fuse_moddown, src/ntt/ntt_moddown.cu:209-214 */
        orc_ntt_fwd(delta, c->logN, q, c->tw + j * N, c->tw_shoup + j * N);
        for (size_t k = 0; k < N; k++) {
            uint64_t v = csub(cx[j * N + k] + q - delta[k], q);
            ct[j * N + k] = shoup_mul(v, l->bigPInv_mod_q[j], l->bigPInv_mod_q_shoup[j], q);
        }
        free(delta);
    }
    free(tmp);
}

/* keyswitch_inplace, src/eval_key_switch.cu:112-212 */
void orc_keyswitch(const orc_ctx *c, const orc_level *l, uint64_t *ct, const uint64_t *c2,
                   const uint64_t *const *evk) {
    size_t N = c->N, Ql = l->size_Ql, QlP = Ql + l->size_P;
    uint64_t *up = (uint64_t *) malloc(l->beta * QlP * N * 8);
    uint64_t *cx = (uint64_t *) malloc(2 * QlP * N * 8);
    orc_modup(c, l, up, c2);
    orc_inner_prod(c, l, cx, up, evk);
    for (size_t i = 0; i < 2; i++) {
        uint64_t *cxi = cx + i * QlP * N;
        orc_moddown_from_ntt(c, l, cxi, cxi); /* in-place aliasing as :190 */
        uint64_t *cti = ct + i * Ql * N;
        for (size_t j = 0; j < Ql; j++) /* add_to_ct_kernel, src/rns_bconv.cu:763-785 */
            for (size_t k = 0; k < N; k++) cti[j * N + k] = csub(cti[j * N + k] + cxi[j * N + k], c->primes[j]);
    }
    free(up);
    free(cx);
}

/* ------------------------------------------------------------------ evaluator pieces */

/* include/uintmodmath.cuh:161-199 */
static inline uint64_t mul_barrett(uint64_t a, uint64_t b, uint64_t q, uint64_t r0, uint64_t r1) {
    return barrett128((u128) a * b, q, r0, r1);
}

/* tensor_prod_2x2_rns_poly, src/polymath.cu:501-536 */
void orc_tensor_prod(const orc_ctx *c, size_t L, const uint64_t *a, const uint64_t *b, uint64_t *out) {
    size_t N = c->N, LN = L * N;
#pragma omp parallel for schedule(dynamic)
    for (size_t t = 0; t < L; t++) {
        uint64_t q = c->primes[t], r0 = c->ratio0[t], r1 = c->ratio1[t];
        for (size_t k = 0; k < N; k++) {
            size_t id = t * N + k;
            uint64_t c00 = a[id], c01 = a[id + LN], c10 = b[id], c11 = b[id + LN];
            uint64_t d0 = mul_barrett(c00, c10, q, r0, r1);
            uint64_t d2 = mul_barrett(c01, c11, q, r0, r1);
            uint64_t d1 = mul_barrett(c00 + c01, c10 + c11, q, r0, r1);
            d1 = d1 + 2 * q - d0 - d2;
            if (d1 >= q) d1 -= q;
            if (d1 >= q) d1 -= q;
            out[id] = d0; out[id + LN] = d1; out[id + 2 * LN] = d2;
        }
    }
}

/* tensor_square_2x2_rns_poly, src/polymath.cu:538-570 */
void orc_tensor_square(const orc_ctx *c, size_t L, const uint64_t *a, uint64_t *out) {
    size_t N = c->N, LN = L * N;
#pragma omp parallel for schedule(dynamic)
    for (size_t t = 0; t < L; t++) {
        uint64_t q = c->primes[t], r0 = c->ratio0[t], r1 = c->ratio1[t];
        for (size_t k = 0; k < N; k++) {
            size_t id = t * N + k;
            uint64_t c0 = a[id], c1 = a[id + LN];
            out[id] = mul_barrett(c0, c0, q, r0, r1);
            out[id + LN] = barrett128(((u128) c0 * c1) << 1, q, r0, r1);
            out[id + 2 * LN] = mul_barrett(c1, c1, q, r0, r1);
        }
    }
}

/* DRNSTool::divide_and_round_q_last_ntt, src/rns.cu:1160-1184 */
void orc_rescale(const orc_ctx *c, const orc_level *l, uint64_t *src, size_t polys, uint64_t *dst) {
    size_t N = c->N, L = l->size_Ql, Ln = L - 1;
    for (size_t i = 0; i < polys; i++) {
        uint64_t *in = src + i * N * L, *out = dst + i * N * Ln;
        orc_ntt_inv_limbs(c, in, 1, L - 1, 0); /* last limb to coefficient form, in place (:1170) */
#pragma omp parallel for schedule(dynamic)
        for (size_t j = 0; j < Ln; j++) {
            uint64_t q = c->primes[j];
            for (size_t k = 0; k < N; k++) out[j * N + k] = barrett64(in[Ln * N + k], q, c->ratio1[j]); /* :1128-1139 */
            orc_ntt_fwd(out + j * N, c->logN, q, c->tw + j * N, c->tw_shoup + j * N);
            for (size_t k = 0; k < N; k++) { /* :1141-1158 */
                uint64_t v = csub(in[j * N + k] + q - out[j * N + k], q);
                out[j * N + k] = shoup_mul(v, l->inv_q_last_mod_q[j], l->inv_q_last_mod_q_shoup[j], q);
            }
        }
    }
}

/* include/galois.cuh:99-114 (PhantomGaloisTool ctor, permutation_tables_) */
void orc_galois_table(int logN, uint32_t galois_elt, uint32_t *table) {
    uint32_t N = 1u << logN;
    for (uint32_t i = N; i < 2 * N; i++) {
        uint32_t rev = brv(i, logN + 1);
        uint64_t raw = ((uint64_t) galois_elt * rev) >> 1;
        raw &= (uint64_t) (N - 1);
        table[i - N] = brv((uint32_t) raw, logN);
    }
}

/* PrecomputeAutoMapKernel, src/util.cu:941-961 */
void orc_automap_table(int logN, uint32_t k, uint32_t *table) {
    uint32_t n = 1u << logN, logm = logN + 1;
    for (uint32_t j = 0; j < n; j++) {
        uint32_t jt = (j << 1) + 1;
        uint32_t idx = ((jt * k) - (((jt * k) >> logm) << logm)) >> 1;
        table[brv(j, logN)] = brv(idx, logN);
    }
}

/* apply_galois_ntt_permutation, src/galois.cu:11-18 */
void orc_apply_galois_ntt(const uint64_t *src, uint64_t *dst, size_t limbs, size_t N, const uint32_t *table) {
    for (size_t t = 0; t < limbs; t++)
        for (size_t k = 0; k < N; k++) dst[t * N + k] = src[t * N + table[k]];
}

/* add_rns_poly / sub_rns_poly / multiply_rns_poly, src/polymath.cu */
void orc_add(const orc_ctx *c, size_t L, const uint64_t *a, const uint64_t *b, uint64_t *out) {
    for (size_t t = 0; t < L; t++)
        for (size_t k = 0; k < c->N; k++) out[t * c->N + k] = csub(a[t * c->N + k] + b[t * c->N + k], c->primes[t]);
}
void orc_sub(const orc_ctx *c, size_t L, const uint64_t *a, const uint64_t *b, uint64_t *out) {
    for (size_t t = 0; t < L; t++)
        for (size_t k = 0; k < c->N; k++)
            out[t * c->N + k] = csub(a[t * c->N + k] + c->primes[t] - b[t * c->N + k], c->primes[t]);
}
void orc_mul(const orc_ctx *c, size_t L, const uint64_t *a, const uint64_t *b, uint64_t *out) {
    for (size_t t = 0; t < L; t++)
        for (size_t k = 0; k < c->N; k++)
            out[t * c->N + k] = mul_barrett(a[t * c->N + k], b[t * c->N + k], c->primes[t], c->ratio0[t], c->ratio1[t]);
}

/* multiply_and_relin_inplace, src/evaluate.cu:1220: tensor -> relinearize(c2 = d2) -> 2 polys */
void orc_hmult_relin(const orc_ctx *c, const orc_level *l, const uint64_t *a, const uint64_t *b,
                     const uint64_t *const *rlk, uint64_t *out) {
    size_t LN = l->size_Ql * c->N;
    uint64_t *d = (uint64_t *) malloc(3 * LN * 8);
    if (a == b) orc_tensor_square(c, l->size_Ql, a, d); else orc_tensor_prod(c, l->size_Ql, a, b, d);
    orc_keyswitch(c, l, d, d + 2 * LN, rlk);
    memcpy(out, d, 2 * LN * 8);
    free(d);
}

/* EvalRotateFused, src/evaluate.cu:2154-2240: ks(c1) with key[autoIndex] on (c0, 0), then permute both */
void orc_rotate_fused(const orc_ctx *c, const orc_level *l, const uint64_t *ct, const uint64_t *const *glk,
                      const uint32_t *table, uint64_t *out) {
    size_t N = c->N, LN = l->size_Ql * N;
    uint64_t *tmp = (uint64_t *) calloc(2 * LN, 8);
    memcpy(tmp, ct, LN * 8); /* c0 kept, c1 slot zeroed */
    orc_keyswitch(c, l, tmp, ct + LN, glk);
    orc_apply_galois_ntt(tmp, out, l->size_Ql, N, table);
    orc_apply_galois_ntt(tmp + LN, out + LN, l->size_Ql, N, table);
    free(tmp);
}
