// evaluator.cu — FLEXIBLEAUTO CKKS evaluator over the sm_100a kernels (see evaluator.h).
#include "evaluator.h"

#include <algorithm>
#include <cmath>
#include <cstring>

#include "../../include/b200fhe.h"

namespace b200fhe {

// ------------------------------------------------------------------------------------------- pool
Pool::~Pool() {
    sh_->alive = false;
    trim();
}

void Pool::trim() {
    for (auto &kv : sh_->idle) cudaFree(kv.second);
    sh_->idle.clear();
}

std::shared_ptr<Pool::Block> Pool::get(size_t bytes) {
    if (bytes == 0) bytes = 256;
    void *p = nullptr;
    size_t got = 0;
    auto it = sh_->idle.lower_bound(bytes);
    if (it != sh_->idle.end() && it->first <= bytes + bytes / 2) {
        got = it->first;
        p = it->second;
        sh_->idle.erase(it);
    } else {
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) {
            trim();  // give idle blocks of other size classes back and retry once
            e = cudaMalloc(&p, bytes);
        }
        if (e != cudaSuccess) throw EvalError(B200FHE_ERR_ALLOC, std::string("cudaMalloc: ") + cudaGetErrorString(e));
        got = bytes;
        reserved_ += bytes;
    }
    auto *b = new Block{p, got};
    std::weak_ptr<Shared> w = sh_;
    return std::shared_ptr<Block>(b, [w](Block *blk) {
        auto s = w.lock();
        if (s && s->alive) s->idle.emplace(blk->bytes, blk->p);
        else cudaFree(blk->p);
        delete blk;
    });
}

// ------------------------------------------------------------------------------------------- helpers
Evaluator::Evaluator(Context *ctx, cudaStream_t stream) : c(ctx), st(stream) {
    // PhantomCiphertext::PreComputeScale (include/ciphertext.h:320-366)
    const size_t sizeQ = c->size_Q;
    sf.resize(sizeQ);
    if (sizeQ == 1) {
        sf[0] = 1.0;  // caller-provided scale is irrelevant for depth 0; never used by the bootstrap
    } else {
        sf[0] = static_cast<double>(c->primes[sizeQ - 1]);
        const double last = sf[0];
        for (size_t k = 1; k < sizeQ; k++) {
            const double prev = sf[k - 1];
            sf[k] = prev * prev / static_cast<double>(c->primes[sizeQ - k]);
            const double ratio = sf[k] / last;
            if (ratio <= 0.5 || ratio >= 2.0)
                throw EvalError(B200FHE_ERR_ARG, "FLEXIBLEAUTO cannot support this number of levels in this parameter setting");
        }
    }
    sf_big.resize(sizeQ - 1);
    for (size_t k = 0; k + 1 < sizeQ; k++) sf_big[k] = sf[k] * sf[k];
}

void Evaluator::cu(cudaError_t e, const char *what) const {
    if (e != cudaSuccess) throw EvalError(B200FHE_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
void Evaluator::ok(int rc) const {
    if (rc != B200FHE_OK) throw EvalError(rc, b200fhe_last_error());
}

DeviceLevel *Evaluator::level(size_t Ql) {
    DeviceLevel *L = nullptr;
    ok(get_level(c, Ql, &L));
    return L;
}

EvkPtrs Evaluator::evk_of(const std::vector<const uint64_t *> &digits, DeviceLevel *L) const {
    if (digits.size() < L->h.beta) throw EvalError(B200FHE_ERR_ARG, "evaluation key has fewer digits than the level needs");
    EvkPtrs e{};
    ok(make_evk(c, L, digits.data(), e));
    return e;
}

const std::vector<const uint64_t *> &Evaluator::galois_key(uint32_t auto_index) const {
    auto it = galois.find(auto_index);
    if (it == galois.end())
        throw EvalError(B200FHE_ERR_ARG, "no Galois key for automorphism index " + std::to_string(auto_index));
    return it->second;
}

const u32 *Evaluator::perm_table(uint32_t auto_index) {
    auto it = perms_.find(auto_index);
    if (it != perms_.end()) return it->second->as<u32>();
    auto b = std::make_unique<DevBuf>();
    cu(b->alloc(c->N * sizeof(u32)), "perm alloc");
    cu(automap_table(b->as<u32>(), c->logN, auto_index, st), "automap_table");
    count_launches(1);
    const u32 *p = b->as<u32>();
    perms_[auto_index] = std::move(b);
    return p;
}

LimbScalars Evaluator::scalars(const std::vector<uint64_t> &w, size_t Ql) const {
    if (w.size() > (size_t) kMaxLimbs) throw EvalError(B200FHE_ERR_UNSUPPORTED, "more limbs than kMaxLimbs");
    LimbScalars s{};
    const PrimeMap pm = c->prime_map(Ql);
    for (size_t t = 0; t < w.size(); t++) {
        const uint64_t q = c->primes[w.size() > Ql ? pm.prime((u32) t) : t];
        s.w[t] = w[t] % q;
        s.ws[t] = host::shoup(s.w[t], q);
    }
    return s;
}

const LimbScalars &Evaluator::p_mod_q(size_t Ql) {
    auto it = pmodq_.find(Ql);
    if (it != pmodq_.end()) return it->second;
    // BaseConverter(Ql -> P).PModq() (src/evaluate.cu:3707-3721) = prod(p_i) mod q_j = DRNSTool bigP_mod_q
    DeviceLevel *L = level(Ql);
    pmodq_[Ql] = scalars(L->h.bigP_mod_q, Ql);
    return pmodq_[Ql];
}

Ct Evaluator::alloc(size_t Ql, u32 polys, bool ext, u32 batch) {
    if (c->host_only) throw EvalError(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    if (Ql == 0 || Ql > c->size_Q) throw EvalError(B200FHE_ERR_ARG, "level out of range");
    Ct x;
    x.ev = this;
    x.Ql = Ql;
    x.limbs = ext ? Ql + c->size_P : Ql;
    x.polys = polys;
    x.batch = batch;
    x.buf = pool.get(size_t(x.npoly()) * x.limbs * c->N * 8);
    return x;
}

// Independent ciphertexts whose bookkeeping is identical run through the evaluator as ONE batched ciphertext: every
// kernel launch covers all of them (twice the grid for the real / imaginary halves of EvalMod).
Ct Evaluator::pack(const std::vector<Ct> &xs) {
    if (xs.empty()) throw EvalError(B200FHE_ERR_ARG, "nothing to pack");
    const Ct &f = xs[0];
    Ct y = alloc(f.Ql, f.polys, f.ext(), (u32) xs.size());
    y.scale = f.scale;
    y.noise = f.noise;
    const size_t bytes = size_t(f.polys) * f.limbs * c->N * 8;
    for (size_t i = 0; i < xs.size(); i++) {
        const Ct &x = xs[i];
        if (x.batch != 1 || x.Ql != f.Ql || x.limbs != f.limbs || x.polys != f.polys || x.scale != f.scale || x.noise != f.noise)
            throw EvalError(B200FHE_ERR_ARG, "pack: ciphertexts differ in level, scale or NoiseScaleDeg");
        cu(cudaMemcpyAsync(reinterpret_cast<char *>(y.data()) + i * bytes, x.data(), bytes, cudaMemcpyDeviceToDevice, st), "pack");
    }
    return y;
}

std::vector<Ct> Evaluator::unpack(const Ct &x) {
    std::vector<Ct> out(x.batch);
    const size_t bytes = size_t(x.polys) * x.limbs * c->N * 8;
    for (u32 i = 0; i < x.batch; i++) {
        Ct y = alloc(x.Ql, x.polys, x.ext());
        y.scale = x.scale;
        y.noise = x.noise;
        cu(cudaMemcpyAsync(y.data(), reinterpret_cast<const char *>(x.data()) + i * bytes, bytes, cudaMemcpyDeviceToDevice, st), "unpack");
        out[i] = y;
    }
    return out;
}

Ct Evaluator::wrap_copy(const u64 *dev, size_t Ql, u32 polys, bool ext, double scale, u32 noise) {
    Ct x = alloc(Ql, polys, ext);
    x.scale = scale;
    x.noise = noise;
    cu(cudaMemcpyAsync(x.data(), dev, size_t(polys) * x.limbs * c->N * 8, cudaMemcpyDeviceToDevice, st), "copy in");
    return x;
}

void Evaluator::own(Ct &x) {
    x.data();  // a pending combination is evaluated first; its block may still be shared with other copies
    if (x.buf.use_count() <= 1) return;
    auto nb = pool.get(size_t(x.npoly()) * x.limbs * c->N * 8);
    cu(cudaMemcpyAsync(nb->p, x.buf->p, size_t(x.npoly()) * x.limbs * c->N * 8, cudaMemcpyDeviceToDevice, st), "cow copy");
    x.buf = nb;
}

const u64 *Evaluator::begin_update(Ct &x) {
    const u64 *src = x.data();
    if (x.buf.use_count() > 1) x.buf = pool.get(size_t(x.npoly()) * x.limbs * c->N * 8);  // old block lives on in its other owners
    return src;
}

// ------------------------------------------------------------------------------------------- pending combinations
u64 *Ct::data() const {
    if (lazy) ev->materialize(*this);
    return buf ? reinterpret_cast<u64 *>(buf->p) : nullptr;
}

void Evaluator::materialize(const Ct &x) {
    const Lazy &L = *x.lazy;
    if (!L.done) {
        const Lazy::Term &t0 = L.terms[0];
        bool identity = L.terms.size() == 1 && L.kappa.empty() && t0.limbs == x.Ql;
        for (size_t t = 0; identity && t < x.Ql; t++) identity = t0.s[t] == 1;
        if (identity) {
            L.done = t0.buf;
        } else {
            AffineArgs a{};
            a.nterms = (u32) L.terms.size();
            a.polys = x.polys;
            a.has_kappa = L.kappa.empty() ? 0 : 1;
            for (size_t i = 0; i < L.terms.size(); i++) {
                a.src[i] = reinterpret_cast<const u64 *>(L.terms[i].buf->p);
                a.stride[i] = L.terms[i].limbs * c->N;
                for (size_t t = 0; t < x.Ql; t++) a.s[i][t] = L.terms[i].s[t];
            }
            for (size_t t = 0; t < x.Ql && a.has_kappa; t++) a.kappa[t] = L.kappa[t];
            auto nb = pool.get(size_t(x.npoly()) * x.Ql * c->N * 8);
            cu(affine_combine(c->T.mods, reinterpret_cast<u64 *>(nb->p), x.Ql * c->N, a, (u32) x.Ql, x.npoly(), c->N, st),
               "affine_combine");
            count_launches(1);
            L.done = nb;
        }
    }
    x.buf = L.done;
    x.lazy.reset();
}

std::shared_ptr<Lazy> Evaluator::lazy_of(const Ct &x) const {
    if (x.ext()) throw EvalError(B200FHE_ERR_ARG, "pending combinations are not kept for extended ciphertexts");
    if (x.Ql > (size_t) kMaxLimbs) throw EvalError(B200FHE_ERR_UNSUPPORTED, "more limbs than kMaxLimbs");
    auto L = std::make_shared<Lazy>();
    if (x.lazy && !x.lazy->done) {
        L->terms = x.lazy->terms;
        L->kappa = x.lazy->kappa;
    } else {
        L->terms.push_back(Lazy::Term{x.lazy ? x.lazy->done : x.buf, x.Ql, std::vector<uint64_t>(x.Ql, 1)});
    }
    return L;
}

void Evaluator::lazy_scale(Ct &x, const std::vector<uint64_t> &w) {
    auto L = lazy_of(x);
    for (auto &t : L->terms)
        for (size_t i = 0; i < x.Ql; i++) t.s[i] = host::mulmod(t.s[i], w[i] % c->primes[i], c->primes[i]);
    for (size_t i = 0; i < x.Ql && !L->kappa.empty(); i++) L->kappa[i] = host::mulmod(L->kappa[i], w[i] % c->primes[i], c->primes[i]);
    x.lazy = L;
    x.buf.reset();
}

void Evaluator::lazy_add(Ct &a, const Ct &b, bool sub) {
    auto La = lazy_of(a);
    auto Lb = lazy_of(b);
    size_t extra = 0;
    for (const auto &tb : Lb->terms) {
        bool found = false;
        for (const auto &ta : La->terms) found = found || (ta.buf == tb.buf && ta.limbs == tb.limbs);
        extra += found ? 0 : 1;
    }
    if (La->terms.size() + extra > (size_t) kMaxAffine) {  // too wide for one pass: evaluate both sides first
        a.data();
        b.data();
        La = lazy_of(a);
        Lb = lazy_of(b);
    }
    const size_t n = a.Ql;
    auto signed_of = [&](uint64_t v, size_t i) { return sub && v ? c->primes[i] - v : v; };
    for (const auto &tb : Lb->terms) {
        Lazy::Term *hit = nullptr;
        for (auto &ta : La->terms)
            if (ta.buf == tb.buf && ta.limbs == tb.limbs) hit = &ta;
        if (hit) {
            for (size_t i = 0; i < n; i++) hit->s[i] = (hit->s[i] + signed_of(tb.s[i], i)) % c->primes[i];
        } else {
            Lazy::Term t = tb;
            for (size_t i = 0; i < n; i++) t.s[i] = signed_of(tb.s[i], i);
            La->terms.push_back(std::move(t));
        }
    }
    if (!Lb->kappa.empty()) {
        if (La->kappa.empty()) La->kappa.assign(n, 0);
        for (size_t i = 0; i < n; i++) La->kappa[i] = (La->kappa[i] + signed_of(Lb->kappa[i], i)) % c->primes[i];
    }
    a.lazy = La;
    a.buf.reset();
}

void Evaluator::lazy_add_const(Ct &x, const std::vector<uint64_t> &k, bool sub) {
    auto L = lazy_of(x);
    if (L->kappa.empty()) L->kappa.assign(x.Ql, 0);
    for (size_t i = 0; i < x.Ql; i++) {
        const uint64_t q = c->primes[i], v = k[i] % q;
        L->kappa[i] = (L->kappa[i] + (sub && v ? q - v : v)) % q;
    }
    x.lazy = L;
    x.buf.reset();
}

// ------------------------------------------------------------------------------------------- level / scale
// ModReduce (src/evaluate.cu:2284-2297): every loop iteration rescales the ORIGINAL ciphertext, so any
// levels >= 1 drops exactly one prime; NoiseScaleDeg -= levels.  levels == 0 leaves an empty ciphertext in the
// reference (a later crash); here it is a no-op.
void Evaluator::mod_reduce(Ct &x, size_t levels) {
    if (levels == 0) return;
    if (x.ext() || x.Ql < 2) throw EvalError(B200FHE_ERR_ARG, "end of modulus switching chain reached");
    Ct y = alloc(x.Ql - 1, x.polys, false, x.batch);
    ok(do_rescale(c, x.Ql, y.data(), x.data(), x.npoly(), st));
    y.scale = x.scale / static_cast<double>(c->primes[x.Ql - 1]);  // evaluate.cu:1636
    y.noise = x.noise - (u32) levels;
    x = y;
}

// ModSwitchLevelInPlace -> mod_switch_drop_to_next (src/evaluate.cu:1649-1698): drop the last primes, keep the scale
void Evaluator::mod_switch_level(Ct &x, size_t levels) {
    if (levels == 0) return;
    if (x.ext() || levels >= x.Ql) throw EvalError(B200FHE_ERR_ARG, "end of modulus switching chain reached");
    // dropping limbs is a view: the surviving limbs are picked up with the source's own stride when the
    // combination is evaluated (the reference copies the ciphertext)
    auto L = lazy_of(x);
    x.Ql -= levels;
    x.limbs = x.Ql;
    x.lazy = L;
    x.buf.reset();
}

static void crt_mult(std::vector<uint64_t> &a, const std::vector<uint64_t> &b, const std::vector<uint64_t> &primes) {
    for (size_t i = 0; i < a.size(); i++) a[i] = host::mulmod(a[i], b[i], primes[i]);
}

// scale-up by 2^logApprox inside the CRT representation (tail of GetElementForEvalMult / ...AddOrSub)
static void crt_scale_pow2(std::vector<uint64_t> &v, int32_t logApprox, const std::vector<uint64_t> &primes) {
    if (logApprox <= 0) return;
    const size_t n = v.size();
    int32_t step = logApprox <= 60 ? logApprox : 60;
    std::vector<uint64_t> approx(n, uint64_t(1) << step);
    logApprox -= step;
    while (logApprox > 0) {
        step = logApprox <= 60 ? logApprox : 60;
        std::vector<uint64_t> sfv(n, uint64_t(1) << step);
        crt_mult(approx, sfv, primes);
        logApprox -= step;
    }
    crt_mult(v, approx, primes);
}

// GetElementForEvalMult (src/evaluate.cu:2332-2412)
std::vector<uint64_t> Evaluator::element_for_mult(const Ct &x, double operand) const {
    const size_t n = x.Ql;
    const double scFactor = level_sf(x);
    typedef __int128_t DoubleInteger;
    const int32_t MAX_BITS_IN_WORD_LOCAL = 125;
    int32_t logApprox = 0;
    const double res = std::fabs(operand * scFactor);
    if (res > 0) {
        const int32_t logSF = static_cast<int32_t>(std::ceil(std::log2(res)));
        const int32_t logValid = (logSF <= MAX_BITS_IN_WORD_LOCAL) ? logSF : MAX_BITS_IN_WORD_LOCAL;
        logApprox = logSF - logValid;
    }
    const double approxFactor = pow(2, logApprox);
    const DoubleInteger large = static_cast<DoubleInteger>(operand / approxFactor * scFactor + 0.5);
    const DoubleInteger large_abs = (large < 0 ? -large : large);
    const DoubleInteger bound = (uint64_t) 1 << 63;
    std::vector<uint64_t> factors(n);
    if (large_abs >= bound) {
        for (size_t i = 0; i < n; i++) {
            const DoubleInteger reduced = large % c->primes[i];
            factors[i] = (reduced < 0) ? static_cast<uint64_t>(reduced + c->primes[i]) : static_cast<uint64_t>(reduced);
        }
    } else {
        const int64_t scConstant = static_cast<int64_t>(large);
        for (size_t i = 0; i < n; i++) {
            const int64_t reduced = scConstant % static_cast<int64_t>(c->primes[i]);
            factors[i] = (reduced < 0) ? reduced + c->primes[i] : reduced;
        }
    }
    crt_scale_pow2(factors, logApprox, c->primes);
    return factors;
}

// GetElementForEvalAddOrSub (src/evaluate.cu:2894-2949)
std::vector<uint64_t> Evaluator::element_for_add(const Ct &x, double operand) const {
    const size_t n = x.Ql;
    const double scFactor = level_sf(x);
    int32_t logApprox = 0;
    const double res = std::fabs(operand * scFactor);
    if (res > 0) {
        const int32_t logSF = static_cast<int32_t>(std::ceil(std::log2(res)));
        const int32_t logValid = (logSF <= 61) ? logSF : 61;
        logApprox = logSF - logValid;
    }
    const double approxFactor = pow(2, logApprox);
    const uint64_t scConstant = static_cast<uint64_t>(operand * scFactor / approxFactor + 0.5);
    std::vector<uint64_t> crt(n, scConstant);
    crt_scale_pow2(crt, logApprox, c->primes);
    const uint64_t intScFactor = static_cast<uint64_t>(scFactor + 0.5);
    std::vector<uint64_t> crtSc(n, intScFactor);
    for (uint32_t i = 1; i < x.noise; i++) crt_mult(crt, crtSc, c->primes);
    return crt;
}

// EvalMultConstInplaceCore (src/evaluate.cu:2299-2330)
void Evaluator::mult_const_core(Ct &x, double scalar) {
    if (x.ext()) throw EvalError(B200FHE_ERR_ARG, "mult_const on an extended ciphertext");
    lazy_scale(x, element_for_mult(x, scalar));
    const double scFactor = level_sf(x);
    x.noise += 1;
    x.scale = x.scale * scFactor;
}

// EvalMultConstInplace (include/evaluate.cuh:317-326)
void Evaluator::mult_const(Ct &x, double scalar) {
    if (x.noise == 2) mod_reduce(x, 1);
    mult_const_core(x, scalar);
}

// EvalAddConstInPlaceWrap -> EvalAddConstInPlace / EvalSubConstInPlace (evaluate.cu:2951-2996): polynomial 0 only
void Evaluator::add_const(Ct &x, double operand) {
    if (operand == 0) return;
    const bool sub = operand < 0;
    lazy_add_const(x, element_for_add(x, sub ? -operand : operand), sub);
}

// AdjustLevelsAndDepthInPlace (src/evaluate.cu:2611-2779)
void Evaluator::adjust_levels_and_depth(Ct &a, Ct &b) {
    const uint32_t c1lvl = (uint32_t) (chain_index(a) - 1), c2lvl = (uint32_t) (chain_index(b) - 1);
    const uint32_t c1depth = a.noise, c2depth = b.noise;
    auto one_side = [&](Ct &lo, Ct &hi, uint32_t llvl, uint32_t hlvl, uint32_t ldepth, uint32_t hdepth) {
        // lo sits at the smaller chain index (more primes) and is brought down to hi
        if (ldepth == 2) {
            if (hdepth == 2) {
                const double scf1 = lo.scale, scf2 = hi.scale, scf = sf[llvl];
                const double q1 = static_cast<double>(c->primes[lo.Ql - 1]);
                mult_const_core(lo, scf2 / scf1 * q1 / scf);
                mod_reduce(lo, 1);
                if (llvl + 1 < hlvl) mod_switch_level(lo, hlvl - llvl - 1);
                lo.scale = hi.scale;
            } else {
                if (llvl + 1 == hlvl) {
                    mod_reduce(lo, 1);
                } else {
                    const double scf1 = lo.scale, scf2 = sf_big[hlvl - 1], scf = sf[llvl];
                    const double q1 = static_cast<double>(c->primes[lo.Ql - 1]);
                    mult_const_core(lo, scf2 / scf1 * q1 / scf);
                    mod_reduce(lo, 1);
                    if (llvl + 2 < hlvl) mod_switch_level(lo, hlvl - llvl - 2);
                    mod_reduce(lo, 1);
                    lo.scale = hi.scale;
                }
            }
        } else {
            if (hdepth == 2) {
                const double scf1 = lo.scale, scf2 = hi.scale, scf = sf[llvl];
                mult_const_core(lo, scf2 / scf1 / scf);
                mod_switch_level(lo, hlvl - llvl);
                lo.scale = scf2;
            } else {
                const double scf1 = lo.scale, scf2 = sf_big[hlvl - 1], scf = sf[llvl];
                mult_const_core(lo, scf2 / scf1 / scf);
                if (llvl + 1 < hlvl) mod_switch_level(lo, hlvl - llvl - 1);
                mod_reduce(lo, 1);
                lo.scale = hi.scale;
            }
        }
    };
    if (c1lvl < c2lvl) one_side(a, b, c1lvl, c2lvl, c1depth, c2depth);
    else if (c1lvl > c2lvl) one_side(b, a, c2lvl, c1lvl, c2depth, c1depth);
    else if (c1depth < c2depth) mult_const_core(a, 1.0);
    else if (c2depth < c1depth) mult_const_core(b, 1.0);
}

// add_inplace / sub_inplace (src/evaluate.cu:120-410): same level and NoiseScaleDeg required, scale of `a` kept
void Evaluator::add_inplace(Ct &a, const Ct &b, bool sub) {
    if (a.Ql != b.Ql || a.limbs != b.limbs) throw EvalError(B200FHE_ERR_ARG, "encrypted1 and encrypted2 parameter mismatch");
    if (a.polys != b.polys || a.batch != b.batch) throw EvalError(B200FHE_ERR_ARG, "poly number mismatch");
    if (a.noise != b.noise) throw EvalError(B200FHE_ERR_ARG, "noise mismatch");
    if (!a.ext()) {
        lazy_add(a, b, sub);
        return;
    }
    const u64 *bsrc = b.data();  // read before a's block may be replaced (a and b can share one)
    const u64 *src = begin_update(a);
    const size_t stride = a.limbs * c->N;
    cu(poly_eltwise(c->T.mods, sub ? ELT_SUB : ELT_ADD, a.data(), stride, src, stride, bsrc, stride,
                    (u32) a.limbs, a.npoly(), a.ext() ? c->prime_map(a.Ql) : PrimeMap{0, 0}, c->N, st),
       "add");
    count_launches(1);
}

void Evaluator::add_auto(Ct &a, const Ct &b) {
    Ct b2 = b;
    adjust_levels_and_depth(a, b2);
    add_inplace(a, b2, false);
}

void Evaluator::sub_auto(Ct &a, const Ct &b) {
    Ct b2 = b;
    adjust_levels_and_depth(a, b2);
    add_inplace(a, b2, true);
}

// multiply_and_relin_inplace on (a, b) at the same level -> 2-polynomial product (evaluate.cu:1220, :415, :1552);
// batched ciphertexts go through the tensor product and the key switch as one launch set
static Ct hmult(Evaluator &ev, const Ct &a, const Ct &b, bool is_square) {
    Context *c = ev.c;
    if (a.Ql != b.Ql) throw EvalError(B200FHE_ERR_ARG, "encrypted1 and encrypted2 parameter mismatch");
    if (a.polys != 2 || b.polys != 2 || a.batch != b.batch) throw EvalError(B200FHE_ERR_ARG, "poly number mismatch");
    if (a.noise != b.noise) throw EvalError(B200FHE_ERR_ARG, "noise mismatch");
    if (!ev.have_relin) throw EvalError(B200FHE_ERR_ARG, "relinearization key not set");
    DeviceLevel *L = ev.level(a.Ql);
    const EvkPtrs e = ev.evk_of(ev.relin, L);
    const u64 *pa = a.data(), *pb = is_square ? nullptr : b.data();
    Ct out = ev.alloc(a.Ql, 2, false, a.batch);
    ev.ok(do_hmult(c, L, out.data(), pa, pb, e, a.batch, ev.st));
    out.scale = a.scale * b.scale;
    return out;
}

// EvalMultAuto (src/evaluate.cu:2794-2811)
Ct Evaluator::mult_auto(const Ct &a, const Ct &b) {
    Ct x = a, y = b;
    adjust_levels_and_depth(x, y);
    if (x.noise == 2) {
        mod_reduce(x, 1);
        mod_reduce(y, 1);
    }
    const u32 n1 = x.noise, n2 = y.noise;
    Ct out = hmult(*this, x, y, false);
    out.noise = n1 + n2;
    return out;
}

// EvalSquare (src/evaluate.cu:3611-3629)
Ct Evaluator::square(const Ct &a) {
    Ct x = a;
    if (x.noise == 1) {
        Ct out = hmult(*this, x, x, true);
        out.noise = 2;
        return out;
    }
    mod_reduce(x, 1);
    const u32 n = x.noise;
    Ct out = hmult(*this, x, x, true);
    out.noise = 2 * n;
    return out;
}

// MultByIntegerInPlace (src/evaluate.cu:3942-3961)
void Evaluator::mult_by_integer(Ct &x, uint64_t v) {
    lazy_scale(x, std::vector<uint64_t>(x.Ql, v));
}

// MultByMonomialInPlace (src/evaluate.cu:2521-2554).  NTT(+-x^index) of limb j depends only on prime j, so it is
// computed once per power over all |Q| limbs and reused at every level (the reference rebuilds it per call).
void Evaluator::mult_by_monomial(Ct &x, uint32_t power) {
    const uint32_t N = (uint32_t) c->N, M = 2 * N;
    const uint32_t powerReduced = power % M, index = power % N;
    auto it = monomials_.find(powerReduced);
    if (it == monomials_.end()) {
        auto b = std::make_unique<DevBuf>();
        cu(b->alloc(c->size_Q * c->N * 8), "monomial alloc");
        cu(monomial_init(c->T.mods, b->as<u64>(), index, powerReduced >= N, (u32) c->size_Q, c->N, st), "monomial_init");
        LimbMap m = c->limb_map(0, c->size_Q, 0, 1, 0);
        cu(ntt_forward(c->T, b->as<u64>(), b->as<u64>(), m, nullptr, st), "monomial ntt");
        count_launches(3);
        it = monomials_.emplace(powerReduced, std::move(b)).first;
    }
    const u64 *src = begin_update(x);
    const size_t stride = x.limbs * c->N;
    cu(poly_eltwise(c->T.mods, ELT_MUL, x.data(), stride, src, stride, it->second->as<u64>(), 0, (u32) x.Ql,
                    x.npoly(), PrimeMap{0, 0}, c->N, st),
       "mult_by_monomial");
    count_launches(1);
}

// RaiseMod (src/evaluate.cu:2459-2504): only limb 0 of the input feeds switchModulusKernel, so only limb 0 is
// brought to coefficient form (the reference inverse-transforms every limb in place and leaves them so).
Ct Evaluator::raise_mod(const Ct &x) {
    if (x.ext() || x.polys != 2) throw EvalError(B200FHE_ERR_ARG, "raise_mod needs a 2-polynomial ciphertext");
    const size_t N = c->N, Q = c->size_Q;
    const u32 np = x.npoly();   // a batch is one launch set over all its polynomials
    auto tmp = pool.get(size_t(np) * N * 8);
    u64 *t = reinterpret_cast<u64 *>(tmp->p);
    LimbMap mi = c->limb_map(0, 1, 0, np, N);
    mi.src_poly_stride = x.limbs * N;
    cu(ntt_inverse(c->T, t, x.data(), mi, nullptr, nullptr, st), "raise_mod intt");
    Ct y = alloc(Q, 2, false, x.batch);
    y.scale = x.scale;
    y.noise = x.noise;
    cu(switch_modulus(c->T.mods, y.data(), Q * N, t, N, (u32) Q, np, N, st), "switch_modulus");
    LimbMap mf = c->limb_map(0, Q, 0, np, Q * N);
    cu(ntt_forward(c->T, y.data(), y.data(), mf, nullptr, st), "raise_mod ntt");
    count_launches(5);
    return y;
}

// EvalRotateFused / EvalConjFused (src/evaluate.cu:2154-2282): key switch c1 first, then the automorphism
static Ct rotate_by_auto_index(Evaluator &ev, const Ct &x, uint32_t k) {
    Context *c = ev.c;
    if (x.ext() || x.polys != 2) throw EvalError(B200FHE_ERR_ARG, "encrypted size must be 2");
    DeviceLevel *L = ev.level(x.Ql);
    const EvkPtrs e = ev.evk_of(ev.galois_key(k), L);
    const size_t LN = x.Ql * c->N, B = x.batch;
    auto tmp = ev.pool.get(B * 2 * LN * 8);
    u64 *t = reinterpret_cast<u64 *>(tmp->p);
    // (c0 + ks0, ks1) of every ciphertext of the batch: the addend exists for polynomial 0 only
    if (B == 1) ev.ok(do_keyswitch(c, L, t, x.data(), 1, x.data() + LN, e, ev.st));
    else ev.ok(do_keyswitch_batch(c, L, t, x.data(), x.data() + LN, 2 * LN, e, (u32) B, ev.st, 1));
    Ct out = ev.alloc(x.Ql, 2, false, x.batch);
    out.scale = x.scale;
    out.noise = x.noise;
    ev.cu(gather_limbs(out.data(), t, ev.perm_table(k), (u32) x.Ql, (u32) (2 * B), LN, LN, c->N, ev.st), "automorphism");
    count_launches(1);
    return out;
}

// EvalRotateFused of ONE ciphertext by several indices (the kernel taps of DNN::Conv, src/dnn.cu:82-150, rotate the
// same channel kernel_h^2 times): modup(c1) once, one inner product per key, ONE batched moddown of all 2R polynomials
// (+ c0 on polynomial 0 of each), then the automorphisms.  Each result is the reference's EvalRotateFused bit for bit
// (the digits of c1 do not depend on the key).  dst: [R][2][Ql][N]; index 0 is the reference's key switch with the
// identity-automorphism key (its rotation-key set holds one for rot = 0).
void Evaluator::rotate_fused_many(const Ct &x, const std::vector<int32_t> &indices, u64 *dst) {
    if (x.ext() || x.polys != 2 || x.batch != 1) throw EvalError(B200FHE_ERR_ARG, "encrypted size must be 2");
    const size_t R = indices.size();
    if (R == 0) return;
    DeviceLevel *L = level(x.Ql);
    const size_t N = c->N, LN = x.Ql * N, QlP = x.Ql + c->size_P;
    auto digits = fast_rotation_precompute(x);
    auto cxb = pool.get(R * 2 * QlP * N * 8);
    auto tb = pool.get(R * 2 * LN * 8);
    u64 *cx = reinterpret_cast<u64 *>(cxb->p), *t = reinterpret_cast<u64 *>(tb->p);
    std::vector<uint32_t> ks(R);
    for (size_t r = 0; r < R; r++) {
        ks[r] = find_automorphism_index_2n_complex(indices[r], 2 * (uint32_t) N);
        ok(do_inner_prod(c, L, cx + r * 2 * QlP * N, reinterpret_cast<const u64 *>(digits->p), evk_of(galois_key(ks[r]), L), st));
    }
    ok(do_moddown(c, L, t, LN, cx, QlP * N, (u32) (2 * R), x.data(), LN, 1, st, 2));
    for (size_t r = 0; r < R; r++) {
        cu(gather_limbs(dst + r * 2 * LN, t + r * 2 * LN, perm_table(ks[r]), (u32) x.Ql, 2, LN, LN, N, st), "automorphism");
        count_launches(1);
    }
}

Ct Evaluator::rotate_fused(const Ct &x, int32_t index) {
    return rotate_by_auto_index(*this, x, find_automorphism_index_2n_complex(index, 2 * (uint32_t) c->N));
}
Ct Evaluator::conj_fused(const Ct &x) { return rotate_by_auto_index(*this, x, 2 * (uint32_t) c->N - 1); }

}  // namespace b200fhe
