// evalseal.cu — the SEAL-style evaluator surface of the reference: negate, plaintext add / sub /
// multiply, add_many, multiply / relinearize as separate steps, rescale_to_next, mod_switch_to_next, apply_galois,
// rotate (NAF decomposition over the Galois keys present) and hoisting (src/evaluate.cu:95-410, 1183-1590, 1700-2149).
// These rotate by applying the automorphism FIRST and key-switching with the key of sigma_g(s) -> s (PhantomGaloisKey),
// the other order than EvalRotateFused.
#include <algorithm>
#include <cmath>

#include "../../include/b200fhe.h"
#include "evaluator.h"

namespace b200fhe {

// are_same_scale -> are_close (src/evaluate.cu:89-92, include/host/common.h:342-345)
static bool same_scale(double a, double b) { return std::fabs(a - b) < 1e-6 * std::max({std::fabs(a), std::fabs(b), 1.0}); }

// negate_inplace (src/evaluate.cu:95-121)
void Evaluator::negate(Ct &x) {
    std::vector<uint64_t> w(x.Ql);
    for (size_t i = 0; i < x.Ql; i++) w[i] = c->primes[i] - 1;
    lazy_scale(x, w);
}

// add_plain_inplace / sub_plain_inplace (src/evaluate.cu:1281-1420): (c0 +- pt, c1 ...)
void Evaluator::add_plain(Ct &x, const u64 *pt, double pt_scale, bool sub) {
    if (x.ext() || x.batch != 1 || !pt) throw EvalError(B200FHE_ERR_ARG, "add_plain: bad argument");
    if (!same_scale(x.scale, pt_scale)) throw EvalError(B200FHE_ERR_ARG, "scale mismatch");
    const size_t LN = x.Ql * c->N;
    Ct y = alloc(x.Ql, x.polys, false);
    y.scale = x.scale;
    y.noise = x.noise;
    const u64 *src = x.data();
    cu(poly_eltwise(c->T.mods, sub ? ELT_SUB : ELT_ADD, y.data(), LN, src, LN, pt, 0, (u32) x.Ql, 1, PrimeMap{0, 0}, c->N, st), "add_plain");
    cu(cudaMemcpyAsync(y.data() + LN, src + LN, (x.polys - 1) * LN * 8, cudaMemcpyDeviceToDevice, st), "add_plain copy");
    count_launches(1);
    x = y;
}

// multiply_plain_inplace -> multiply_plain_ntt (src/evaluate.cu:1420-1448): every polynomial times the plaintext
void Evaluator::multiply_plain(Ct &x, const u64 *pt, double pt_scale) {
    if (x.ext() || x.batch != 1 || !pt) throw EvalError(B200FHE_ERR_ARG, "multiply_plain: bad argument");
    const size_t LN = x.Ql * c->N;
    Ct y = alloc(x.Ql, x.polys, false);
    y.scale = x.scale * pt_scale;
    y.noise = x.noise;
    cu(poly_eltwise(c->T.mods, ELT_MUL, y.data(), LN, x.data(), LN, pt, 0, (u32) x.Ql, x.polys, PrimeMap{0, 0}, c->N, st), "multiply_plain");
    count_launches(1);
    x = y;
}

// add_many (src/evaluate.cu:232-306)
Ct Evaluator::add_many(const std::vector<Ct> &xs) {
    if (xs.empty()) throw EvalError(B200FHE_ERR_ARG, "encrypteds cannot be empty");
    Ct acc = xs[0];
    for (size_t i = 1; i < xs.size(); i++) {
        if (xs[i].Ql != acc.Ql) throw EvalError(B200FHE_ERR_ARG, "encrypteds parameter mismatch");
        if (!same_scale(xs[i].scale, acc.scale)) throw EvalError(B200FHE_ERR_ARG, "encrypteds scale mismatch");
        if (xs[i].polys != acc.polys) throw EvalError(B200FHE_ERR_ARG, "encrypteds size mismatch");
        Ct b = xs[i];
        b.noise = acc.noise;
        add_inplace(acc, b, false);
    }
    return acc;
}

// multiply_inplace -> bgv_ckks_multiply (src/evaluate.cu:415-480, 1183-1218): (d0, d1, d2), scale = product
Ct Evaluator::multiply(const Ct &a, const Ct &b) {
    if (a.Ql != b.Ql) throw EvalError(B200FHE_ERR_ARG, "encrypted1 and encrypted2 parameter mismatch");
    if (a.ext() || b.ext() || a.polys != 2 || b.polys != 2 || a.batch != 1 || b.batch != 1)
        throw EvalError(B200FHE_ERR_ARG, "multiply takes two 2-polynomial ciphertexts");
    const size_t LN = a.Ql * c->N;
    Ct out = alloc(a.Ql, 3, false);
    const bool sq = a.buf && a.buf == b.buf && !a.lazy && !b.lazy;
    cu(tensor_split(c->T.mods, out.data(), out.data() + 2 * LN, a.data(), sq ? nullptr : b.data(), (u32) a.Ql, 1, c->N, st), "tensor");
    count_launches(1);
    out.scale = a.scale * b.scale;
    out.noise = a.noise + b.noise;
    return out;
}

// relinearize_inplace (src/evaluate.cu:1552-1589)
Ct Evaluator::relinearize(const Ct &x) {
    if (x.ext() || x.polys != 3 || x.batch != 1) throw EvalError(B200FHE_ERR_ARG, "destination_size must be 3");
    if (!have_relin) throw EvalError(B200FHE_ERR_ARG, "relinearization key not set");
    DeviceLevel *L = level(x.Ql);
    const size_t LN = x.Ql * c->N;
    Ct out = alloc(x.Ql, 2, false);
    out.scale = x.scale;
    out.noise = x.noise;
    ok(do_keyswitch(c, L, out.data(), x.data(), 2, x.data() + 2 * LN, evk_of(relin, L), st));
    return out;
}

// rescale_to_next -> mod_switch_scale_to_next (src/evaluate.cu:1591-1646, 1779-1801)
Ct Evaluator::rescale_to_next(const Ct &x) {
    if (x.ext() || x.Ql < 2) throw EvalError(B200FHE_ERR_ARG, "end of modulus switching chain reached");
    Ct y = alloc(x.Ql - 1, x.polys, false, x.batch);
    ok(do_rescale(c, x.Ql, y.data(), x.data(), x.npoly(), st));
    y.scale = x.scale / static_cast<double>(c->primes[x.Ql - 1]);
    y.noise = x.noise;
    return y;
}

// mod_switch_to_next for CKKS -> mod_switch_drop_to_next (src/evaluate.cu:1649-1698, 1733-1777)
Ct Evaluator::mod_switch_to_next(const Ct &x) {
    Ct y = x;
    mod_switch_level(y, 1);
    return y;
}

static const std::vector<const uint64_t *> &seal_key(const Evaluator &ev, uint32_t elt) {
    auto it = ev.seal_galois.find(elt);
    if (it == ev.seal_galois.end()) throw EvalError(B200FHE_ERR_ARG, "Galois key not present");
    return it->second;
}

// apply_galois_inplace (src/evaluate.cu:1803-1874): (sigma(c0), 0) + KeySwitch(sigma(c1)) with the key of sigma_g(s) -> s
Ct Evaluator::apply_galois(const Ct &x, uint32_t elt) {
    if (x.ext() || x.polys != 2 || x.batch != 1) throw EvalError(B200FHE_ERR_ARG, "encrypted size must be 2");
    DeviceLevel *L = level(x.Ql);
    const EvkPtrs e = evk_of(seal_key(*this, elt), L);
    const size_t LN = x.Ql * c->N;
    auto tmp = pool.get(2 * LN * 8);
    u64 *t = reinterpret_cast<u64 *>(tmp->p);
    cu(gather_limbs(t, x.data(), perm_table(elt), (u32) x.Ql, 2, LN, LN, c->N, st), "automorphism");
    count_launches(1);
    Ct out = alloc(x.Ql, 2, false);
    out.scale = x.scale;
    out.noise = x.noise;
    ok(do_keyswitch(c, L, out.data(), t, 1, t + LN, e, st));
    return out;
}

// get_elt_from_step (include/galois.cuh:17-50)
static uint32_t elt_from_step(int step, size_t N) {
    const uint32_t n = (uint32_t) N, m32 = 2 * n;
    if (step == 0) return m32 - 1;
    const bool sign = step < 0;
    uint32_t pos = (uint32_t) std::abs(step);
    if (pos >= (n >> 1)) throw EvalError(B200FHE_ERR_ARG, "step count too large");
    pos &= m32 - 1;
    int s = sign ? (int) (n >> 1) - (int) pos : (int) pos;
    uint64_t g = 1;
    while (s--) g = (g * 5) & (uint64_t(m32) - 1);
    return (uint32_t) g;
}

// arith::naf (include/host/numth.h:17-34)
static std::vector<int> naf(int value) {
    std::vector<int> res;
    const bool sign = value < 0;
    value = std::abs(value);
    for (int i = 0; value; i++) {
        const int zi = (value & 1) ? 2 - (value & 3) : 0;
        value = (value - zi) >> 1;
        if (zi) res.push_back((sign ? -zi : zi) * (1 << i));
    }
    return res;
}

// rotate_inplace -> rotate_internal (src/evaluate.cu:1876-1920)
Ct Evaluator::rotate(const Ct &x, int step) {
    const uint32_t elt = elt_from_step(step, c->N);
    if (seal_galois.count(elt)) return apply_galois(x, elt);
    const std::vector<int> parts = naf(step);
    if (parts.size() == 1) throw EvalError(B200FHE_ERR_ARG, "Galois key not present");
    Ct y = x;
    for (int p : parts)
        if ((size_t) std::abs(p) != (c->N >> 1)) y = rotate(y, p);
    return y;
}

// hoisting_inplace (src/evaluate.cu:1922-2149): sum over the steps of rotate(ct, step) with ONE modup of c1 and ONE
// moddown: the digits are permuted per step, multiplied with that step's key and accumulated in Q_l P
Ct Evaluator::hoisting(const Ct &x, const std::vector<int> &steps) {
    if (x.ext() || x.polys != 2 || x.batch != 1) throw EvalError(B200FHE_ERR_ARG, "ciphertext size must be 2");
    if (steps.empty()) throw EvalError(B200FHE_ERR_ARG, "no steps");
    DeviceLevel *L = level(x.Ql);
    const size_t N = c->N, LN = x.Ql * N, QlP = x.Ql + c->size_P, beta = L->h.beta;
    auto digits = fast_rotation_precompute(x);
    auto pd = pool.get(beta * QlP * N * 8), cxa = pool.get(2 * QlP * N * 8), cxt = pool.get(2 * QlP * N * 8);
    auto c0a = pool.get(LN * 8), c0t = pool.get(LN * 8);
    u64 *perm_digits = reinterpret_cast<u64 *>(pd->p), *acc_cx = reinterpret_cast<u64 *>(cxa->p), *tmp_cx = reinterpret_cast<u64 *>(cxt->p);
    u64 *acc_c0 = reinterpret_cast<u64 *>(c0a->p), *tmp_c0 = reinterpret_cast<u64 *>(c0t->p);
    const PrimeMap pm = c->prime_map(x.Ql);
    for (size_t i = 0; i < steps.size(); i++) {
        const uint32_t elt = elt_from_step(steps[i], N);
        const EvkPtrs e = evk_of(seal_key(*this, elt), L);
        const u32 *perm = perm_table(elt);
        cu(gather_limbs(i ? tmp_c0 : acc_c0, x.data(), perm, (u32) x.Ql, 1, LN, LN, N, st), "hoisting c0");
        cu(gather_limbs(perm_digits, reinterpret_cast<const u64 *>(digits->p), perm, (u32) QlP, (u32) beta, QlP * N, QlP * N, N, st),
           "hoisting digits");
        ok(do_inner_prod(c, L, i ? tmp_cx : acc_cx, perm_digits, e, st));
        count_launches(2);
        if (i) {
            cu(poly_eltwise(c->T.mods, ELT_ADD, acc_c0, LN, acc_c0, LN, tmp_c0, LN, (u32) x.Ql, 1, PrimeMap{0, 0}, N, st), "hoisting add c0");
            cu(poly_eltwise(c->T.mods, ELT_ADD, acc_cx, QlP * N, acc_cx, QlP * N, tmp_cx, QlP * N, (u32) QlP, 2, pm, N, st), "hoisting add cx");
            count_launches(2);
        }
    }
    Ct out = alloc(x.Ql, 2, false);
    out.scale = x.scale;
    out.noise = x.noise;
    ok(do_moddown(c, L, out.data(), LN, acc_cx, QlP * N, 2, acc_c0, LN, 1, st));
    return out;
}

}  // namespace b200fhe
