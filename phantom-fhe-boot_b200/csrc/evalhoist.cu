// evalhoist.cu — hoisted rotations and the extended-basis (Q_l P) helpers of the evaluator
// (EvalFastRotationPrecompute / EvalFastRotationExt / KeySwitchExt / KeySwitchDown / EvalMultExt / EvalAddExt,
// src/evaluate.cu:3631-3940).
#include <cmath>

#include "../../include/b200fhe.h"
#include "evaluator.h"

#include <algorithm>

namespace b200fhe {

// src/util.cu:355-367
uint32_t reduce_rotation(int32_t index, uint32_t slots) {
    const int32_t islots = int32_t(slots);
    if ((slots & (slots - 1)) == 0) {
        const int32_t n = (int32_t) std::log2(slots);
        if (index >= 0) return index - ((index >> n) << n);
        return index + islots + ((int32_t(std::fabs(index)) >> n) << n);
    }
    return (islots + index % islots) % islots;
}

// src/util.cu:908-937: 5^i (or 5^-|i|) mod m; m-1 is the conjugation
uint32_t find_automorphism_index_2n_complex(int32_t i, uint32_t m) {
    if (i == 0) return 1;
    if (i == int32_t(m - 1)) return uint32_t(i);
    uint64_t g0 = 5;
    if (i < 0) {
        // inverse of 5 modulo the power of two m (Newton iteration doubles the valid bits)
        uint64_t inv = 1;
        for (int it = 0; it < 6; it++) inv = inv * (2 - 5 * inv);
        g0 = inv & (uint64_t(m) - 1);
    }
    const uint32_t steps = (uint32_t) std::abs(i);
    uint64_t g = g0;
    for (uint32_t j = 1; j < steps; j++) g = (g * g0) % m;
    return uint32_t(g);
}

// EvalFastRotationPrecompute: digits = modup(c1)
std::shared_ptr<Pool::Block> Evaluator::fast_rotation_precompute(const Ct &x, const std::vector<std::pair<u32, u32>> *ranges) {
    if (x.ext() || x.polys != 2 || x.batch != 1) throw EvalError(B200FHE_ERR_ARG, "fast_rotation_precompute needs a 2-polynomial ciphertext");
    DeviceLevel *L = level(x.Ql);
    const size_t Ql = x.Ql, QlP = Ql + c->size_P, N = c->N;
    auto digits = pool.get(L->h.beta * QlP * N * 8);
    u64 *up = reinterpret_cast<u64 *>(digits->p);
    const u64 *c1 = x.data() + Ql * N;
    if (!ranges || L->h.alpha < 2 || c->size_P < 2 || L->h.beta > (size_t) kBconvDigits) {
        ok(do_modup(c, L, up, c1, st));
        return digits;
    }
    // the inverse transform of all limbs (every conversion needs the whole digit), then conversion + forward transform of the
    // wanted rows only: the rows outside the runs are never read
    cu(c->ws_tcks.alloc(Ql * N * 8), "workspace");
    u64 *t = c->ws_tcks.as<u64>();
    LimbMap mi = c->limb_map(0, Ql, 0, 1, Ql * N);
    cu(ntt_inverse(c->T, t, c1, mi, L->partQlHatInv.as<u64>(), L->partQlHatInv_shoup.as<u64>(), st), "modup inverse transform");
    count_launches(2);
    for (const auto &r : *ranges) {
        if (r.second == 0) continue;
        const size_t lo = r.first, hi = lo + r.second;
        const size_t q_lo = std::min(lo, Ql), q_hi = std::min(hi, Ql), p_lo = std::max(lo, Ql) - Ql, p_hi = std::max(hi, Ql) - Ql;
        ok(modup_rows(c, L, up, t, c1, q_lo, q_hi, p_lo, p_hi, st));
        count_launches(3);
    }
    return digits;
}

// EvalFastRotationExt: inner product with the key of the automorphism (+ P*c0), then the automorphism, one kernel
Ct Evaluator::fast_rotation_ext(const Ct &x, int32_t index, const std::shared_ptr<Pool::Block> &digits, bool add_first) {
    if (x.ext() || x.polys != 2 || x.batch != 1) throw EvalError(B200FHE_ERR_ARG, "encrypted size must be 2");
    const uint32_t k = find_automorphism_index_2n_complex(index, 2 * (uint32_t) c->N);
    DeviceLevel *L = level(x.Ql);
    Ct out = alloc(x.Ql, 2, true);
    out.scale = x.scale;
    out.noise = x.noise;
    FastRotArgs a{};
    a.up = reinterpret_cast<const u64 *>(digits->p);
    a.evk = evk_of(galois_key(k), L);
    a.perm = perm_table(k);
    a.c0 = add_first ? x.data() : nullptr;
    a.out = out.data();
    a.beta = (u32) L->h.beta;
    a.size_Ql = (u32) x.Ql;
    a.size_QlP = (u32) out.limbs;
    a.size_QP = (u32) c->size_QP;
    a.auto_index = k;
    a.logN = c->logN;
    cu(b200fhe::fast_rotation_ext(c->T.mods, a, p_mod_q(x.Ql), c->prime_map(x.Ql), c->N, st), "fast_rotation_ext");
    count_launches(1);
    return out;
}

static uint32_t inverse_mod_pow2(uint32_t k, uint32_t mask) {  // k odd; Newton doubles the valid bits: 5 steps >= 32
    uint32_t x = k;
    for (int i = 0; i < 5; i++) x *= 2u - k * x;
    return x & mask;
}

std::vector<Ct> Evaluator::fast_rotations_ext(const Ct &x, const std::vector<int32_t> &indices,
                                              const std::shared_ptr<Pool::Block> &digits, bool add_first,
                                              const std::vector<std::pair<u32, u32>> *ranges) {
    if (x.ext() || x.polys != 2 || x.batch != 1) throw EvalError(B200FHE_ERR_ARG, "encrypted size must be 2");
    DeviceLevel *L = level(x.Ql);
    std::vector<Ct> outs(indices.size());
    if (ranges && !fast_rotation_ext_multi_supported((u32) L->h.beta, c->logN))
        throw EvalError(B200FHE_ERR_UNSUPPORTED, "limb-sharded hoisted rotations need the one-launch kernel (beta <= 4, N > 512)");
    if (!fast_rotation_ext_multi_supported((u32) L->h.beta, c->logN)) {
        for (size_t i = 0; i < indices.size(); i++) outs[i] = fast_rotation_ext(x, indices[i], digits, add_first);
        return outs;
    }
    const uint32_t M = 2 * (uint32_t) c->N;
    const u64 *c0 = add_first ? x.data() : nullptr;
    for (size_t i0 = 0; i0 < indices.size(); i0 += kRotMax) {
        const size_t n = std::min<size_t>(kRotMax, indices.size() - i0);
        FastRotMultiArgs a{};
        a.up = reinterpret_cast<const u64 *>(digits->p);
        a.c0 = c0;
        a.nrot = (u32) n;
        a.beta = (u32) L->h.beta;
        a.size_Ql = (u32) x.Ql;
        a.size_QlP = (u32) (x.Ql + c->size_P);
        a.size_QP = (u32) c->size_QP;
        a.logN = c->logN;
        for (size_t i = 0; i < n; i++) {
            const uint32_t k = find_automorphism_index_2n_complex(indices[i0 + i], M);
            const EvkPtrs e = evk_of(galois_key(k), L);
            for (size_t d = 0; d < L->h.beta; d++) a.evk[i][d] = e.d[d];
            Ct o = alloc(x.Ql, 2, true);
            o.scale = x.scale;
            o.noise = x.noise;
            a.out[i] = o.data();
            a.k[i] = k;
            a.k_inv[i] = inverse_mod_pow2(k, M - 1);
            outs[i0 + i] = o;
        }
        if (ranges) {
            for (const auto &r : *ranges) {
                if (r.second == 0) continue;
                a.limb0 = r.first;
                a.nlimbs = r.second;
                cu(fast_rotation_ext_multi(c->T.mods, a, p_mod_q(x.Ql), c->prime_map(x.Ql), c->N, st), "fast_rotation_ext_multi");
                count_launches(1);
            }
        } else {
            cu(fast_rotation_ext_multi(c->T.mods, a, p_mod_q(x.Ql), c->prime_map(x.Ql), c->N, st), "fast_rotation_ext_multi");
            count_launches(1);
        }
    }
    return outs;
}

// KeySwitchExt: lift to Q_l P by multiplying with P mod q_i; the P limbs are zero
Ct Evaluator::keyswitch_ext(const Ct &x) {
    if (x.ext() || x.batch != 1) throw EvalError(B200FHE_ERR_ARG, "keyswitch_ext on an extended ciphertext");
    Ct out = alloc(x.Ql, x.polys, true);
    out.scale = x.scale;
    out.noise = x.noise;
    const size_t N = c->N, P = c->size_P;
    cu(limb_scalar(c->T.mods, SC_MUL, out.data(), out.limbs * N, x.data(), x.limbs * N, nullptr, 0, p_mod_q(x.Ql),
                   (u32) x.Ql, x.polys, PrimeMap{0, 0}, N, st),
       "keyswitch_ext");
    cu(cudaMemset2DAsync(out.data() + x.Ql * N, out.limbs * N * 8, 0, P * N * 8, x.polys, st), "keyswitch_ext zero");
    count_launches(1);
    return out;
}

// KeySwitchDown: moddown of both polynomials in one batched call
Ct Evaluator::keyswitch_down(Ct &x) {
    if (!x.ext() || x.polys != 2 || x.batch != 1) throw EvalError(B200FHE_ERR_ARG, "keyswitch_down needs an extended 2-polynomial ciphertext");
    DeviceLevel *L = level(x.Ql);
    own(x);  // the P limbs are transformed in place
    Ct out = alloc(x.Ql, 2, false);
    out.scale = x.scale;
    out.noise = x.noise;
    ok(do_moddown(c, L, out.data(), x.Ql * c->N, x.data(), x.limbs * c->N, 2, nullptr, 0, 0, st));
    return out;
}

// KeySwitchDownFirstElement: moddown of polynomial 0 only
Ct Evaluator::keyswitch_down_first(Ct &x) {
    if (!x.ext() || x.batch != 1) throw EvalError(B200FHE_ERR_ARG, "keyswitch_down_first needs an extended ciphertext");
    DeviceLevel *L = level(x.Ql);
    own(x);
    Ct out = alloc(x.Ql, 1, false);
    out.scale = x.scale;
    out.noise = x.noise;
    ok(do_moddown(c, L, out.data(), 0, x.data(), 0, 1, nullptr, 0, 0, st));
    return out;
}

// EvalMultExtInPlace (scale / NoiseScaleDeg are updated by the caller, who knows the plaintext's)
void Evaluator::mult_ext(Ct &x, const u64 *pt) {
    if (!x.ext()) throw EvalError(B200FHE_ERR_ARG, "mult_ext needs an extended ciphertext");
    own(x);
    const size_t stride = x.limbs * c->N;
    cu(poly_eltwise(c->T.mods, ELT_MUL, x.data(), stride, x.data(), stride, pt, 0, (u32) x.limbs, x.npoly(),
                    c->prime_map(x.Ql), c->N, st),
       "mult_ext");
    count_launches(1);
}

// EvalAddExtInPlace
void Evaluator::add_ext(Ct &a, const Ct &b) {
    if (!a.ext() || !b.ext() || a.noise != b.noise || a.Ql != b.Ql) throw EvalError(B200FHE_ERR_ARG, "Eval Add Ext Failed.");
    add_inplace(a, b, false);
}

}  // namespace b200fhe
