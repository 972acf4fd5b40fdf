// hostmath.cpp — see hostmath.h.
#include "hostmath.h"

#include <map>
#include <stdexcept>

namespace b200fhe {
namespace host {

uint64_t powmod(uint64_t a, uint64_t e, uint64_t q) {
    uint64_t r = 1 % q;
    a %= q;
    for (; e; e >>= 1) {
        if (e & 1) r = mulmod(r, a, q);
        a = mulmod(a, a, q);
    }
    return r;
}

bool is_prime(uint64_t n) {
    if (n < 2) return false;
    static const uint64_t small[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (uint64_t p : small) {
        if (n == p) return true;
        if (n % p == 0) return false;
    }
    uint64_t d = n - 1;
    int s = 0;
    while ((d & 1) == 0) {
        d >>= 1;
        s++;
    }
    for (uint64_t a : small) {  // deterministic for 64-bit inputs
        uint64_t x = powmod(a, d, n);
        if (x == 1 || x == n - 1) continue;
        bool witness = true;
        for (int i = 1; i < s && witness; i++) {
            x = mulmod(x, x, n);
            if (x == n - 1) witness = false;
        }
        if (witness) return false;
    }
    return true;
}

void barrett_ratio(uint64_t q, uint64_t &r0, uint64_t &r1) {
    u128_t r = (~(u128_t) 0) / q;  // q odd: floor((2^128-1)/q) == floor(2^128/q)
    r0 = (uint64_t) r;
    r1 = (uint64_t) (r >> 64);
}

std::vector<uint64_t> get_primes(size_t ntt_size, int bit_size, size_t count) {
    std::vector<uint64_t> out;
    if (bit_size < 2 || bit_size > 61) throw std::invalid_argument("bit_size is invalid");
    const uint64_t step = 2 * (uint64_t) ntt_size;
    const uint64_t lower = uint64_t(1) << (bit_size - 1);
    uint64_t v = (uint64_t(1) << bit_size) - step + 1;
    for (; out.size() < count && v > lower; v -= step)
        if (is_prime(v)) out.push_back(v);
    if (out.size() < count) throw std::logic_error("failed to find enough qualifying primes");
    return out;
}

std::vector<uint64_t> coeff_modulus_create(size_t N, const std::vector<int> &bit_sizes) {
    std::map<int, size_t> need;
    for (int b : bit_sizes) need[b]++;
    std::map<int, std::vector<uint64_t>> pool;
    for (auto &kv : need) pool[kv.first] = get_primes(N, kv.first, kv.second);
    std::vector<uint64_t> out;
    for (int b : bit_sizes) {  // the reference hands out each size's primes smallest-first
        out.push_back(pool[b].back());
        pool[b].pop_back();
    }
    return out;
}

uint64_t minimal_primitive_root(uint64_t degree, uint64_t q) {
    if ((q - 1) % degree) throw std::invalid_argument("modulus does not support the NTT size");
    const uint64_t cof = (q - 1) / degree;
    uint64_t g = 0;
    for (uint64_t a = 2; a < 4096 && !g; a++) {
        uint64_t c = powmod(a, cof, q);
        if (powmod(c, degree / 2, q) == q - 1) g = c;
    }
    if (!g) throw std::logic_error("no primitive root found");
    const uint64_t g2 = mulmod(g, g, q);
    uint64_t best = g, cur = g;
    for (uint64_t i = 1; i < degree / 2; i++) {  // all odd powers of g
        cur = mulmod(cur, g2, q);
        if (cur < best) best = cur;
    }
    return best;
}

uint32_t reverse_bits(uint32_t x, int bits) {
    uint32_t r = 0;
    for (int i = 0; i < bits; i++, x >>= 1) r = (r << 1) | (x & 1);
    return r;
}

PrimeTables ntt_tables(int logN, uint64_t q) {
    const size_t N = size_t(1) << logN;
    PrimeTables t;
    t.tw.assign(N, 0);
    t.tw_shoup.assign(N, 0);
    t.itw.assign(N, 0);
    t.itw_shoup.assign(N, 0);
    const uint64_t psi = minimal_primitive_root(2 * N, q), ipsi = invmod(psi, q);
    uint64_t a = 1, b = 1;
    for (size_t i = 0; i < N; i++) {
        const uint32_t r = reverse_bits((uint32_t) i, logN);
        t.tw[r] = a;
        t.itw[r] = b;
        a = mulmod(a, psi, q);
        b = mulmod(b, ipsi, q);
    }
    t.ninv = invmod((uint64_t) N, q);
    t.ninv_shoup = shoup(t.ninv, q);
    t.itw[1] = mulmod(t.itw[1], t.ninv, q);
    for (size_t i = 0; i < N; i++) {
        t.tw_shoup[i] = shoup(t.tw[i], q);
        t.itw_shoup[i] = shoup(t.itw[i], q);
    }
    return t;
}

// product of base[] except index skip, mod m
static uint64_t punctured(const uint64_t *base, size_t n, size_t skip, uint64_t m) {
    uint64_t r = 1 % m;
    for (size_t i = 0; i < n; i++)
        if (i != skip) r = mulmod(r, base[i] % m, m);
    return r;
}

LevelConsts level_consts(const std::vector<uint64_t> &QP, size_t size_P, size_t size_Ql) {
    LevelConsts L;
    const size_t size_Q = QP.size() - size_P;
    if (size_Ql == 0 || size_Ql > size_Q) throw std::invalid_argument("size_Ql out of range");
    L.size_Ql = size_Ql;
    L.size_P = size_P;
    L.size_Q = size_Q;
    L.alpha = size_P;
    L.beta = size_P ? (size_Ql + size_P - 1) / size_P : 0;
    const uint64_t *q = QP.data(), *p = QP.data() + size_Q;
    std::vector<uint64_t> QlP(q, q + size_Ql);
    QlP.insert(QlP.end(), p, p + size_P);

    L.partQlHatInv.resize(size_Ql);
    L.partQlHatInv_shoup.resize(size_Ql);
    L.modup_mat.resize(L.beta);
    for (size_t d = 0; d < L.beta; d++) {
        const size_t s = L.alpha * d, ni = L.digit_size(d), no = QlP.size() - ni;
        for (size_t i = 0; i < ni; i++) {
            const uint64_t inv = invmod(punctured(q + s, ni, i, q[s + i]), q[s + i]);
            L.partQlHatInv[s + i] = inv;
            L.partQlHatInv_shoup[s + i] = shoup(inv, q[s + i]);
        }
        auto &M = L.modup_mat[d];
        M.resize(no * ni);
        for (size_t j = 0; j < no; j++) {
            const uint64_t pj = QlP[j < s ? j : j + ni];
            for (size_t i = 0; i < ni; i++) M[j * ni + i] = punctured(q + s, ni, i, pj);
        }
    }
    L.bigP_mod_q.resize(size_Ql);
    L.bigP_mod_q_shoup.resize(size_Ql);
    L.bigPInv_mod_q.resize(size_Ql);
    L.bigPInv_mod_q_shoup.resize(size_Ql);
    for (size_t i = 0; i < size_Ql && size_P; i++) {
        const uint64_t pm = punctured(p, size_P, (size_t) -1, q[i]);
        L.bigP_mod_q[i] = pm;
        L.bigP_mod_q_shoup[i] = shoup(pm, q[i]);
        L.bigPInv_mod_q[i] = invmod(pm, q[i]);
        L.bigPInv_mod_q_shoup[i] = shoup(L.bigPInv_mod_q[i], q[i]);
    }
    L.PHatInv_mod_p.resize(size_P);
    L.PHatInv_mod_p_shoup.resize(size_P);
    for (size_t i = 0; i < size_P; i++) {
        L.PHatInv_mod_p[i] = invmod(punctured(p, size_P, i, p[i]), p[i]);
        L.PHatInv_mod_p_shoup[i] = shoup(L.PHatInv_mod_p[i], p[i]);
    }
    L.PHat_mod_q.resize(size_Ql * size_P);
    for (size_t j = 0; j < size_Ql; j++)
        for (size_t i = 0; i < size_P; i++) L.PHat_mod_q[j * size_P + i] = punctured(p, size_P, i, q[j]);
    L.inv_q_last_mod_q.resize(size_Ql - 1);
    L.inv_q_last_mod_q_shoup.resize(size_Ql - 1);
    for (size_t i = 0; i + 1 < size_Ql; i++) {
        L.inv_q_last_mod_q[i] = invmod(q[size_Ql - 1], q[i]);
        L.inv_q_last_mod_q_shoup[i] = shoup(L.inv_q_last_mod_q[i], q[i]);
    }
    return L;
}

}  // namespace host
}  // namespace b200fhe
