// evalpoly.cu — Chebyshev series evaluation (Paterson-Stockmeyer) of the evaluator: EvalChebyshevSeries[PS|Linear],
// InnerEvalChebyshevPS, EvalLinearWSumMutable (src/evaluate.cu:2998-3583) and the double-precision polynomial
// helpers they depend on (src/util.cu:76-300).  The floating-point operation ORDER of the reference is kept:
// every coefficient feeds an integer scalar, so a different rounding would change ciphertext limbs.
#include <algorithm>
#include <cmath>

#include "../../include/b200fhe.h"
#include "evaluator.h"

namespace b200fhe {

static bool not_one(double v) {  // IsNotEqualOne, PREC = 2^-20
    static const double prec = std::pow(2, -20);
    return (1 - prec >= v) || (1 + prec <= v);
}

uint32_t poly_degree(const std::vector<double> &cf) {
    if (cf.empty()) throw EvalError(B200FHE_ERR_ARG, "The coefficients vector can not be empty");
    int32_t i = (int32_t) cf.size();
    while (--i >= 0)
        if (cf[i]) break;
    return (uint32_t) (i < 0 ? 0 : i);
}

static void scale_all(std::vector<double> &d, double by, bool divide) {
    for (double &v : d) v = divide ? v / by : v * by;
}

static void subtract_and_trim(std::vector<double> &r, const std::vector<double> &d, uint32_t &n) {
    for (size_t i = 0; i < r.size(); i++) r[i] = r[i] - d[i];
    if (r.size() > 1) {
        n = poly_degree(r);
        r.resize(n + 1);
    }
}

// Division of Chebyshev-basis series: T_a * T_b = (T_{a+b} + T_{|a-b|}) / 2 drives the shape of the subtrahend.
LongDiv long_division_chebyshev(const std::vector<double> &f, const std::vector<double> &g) {
    uint32_t n = poly_degree(f);
    const uint32_t k = poly_degree(g);
    if (n != f.size() - 1) throw EvalError(B200FHE_ERR_ARG, "LongDivisionChebyshev: The dominant coefficient of the divident is zero.");
    if (k != g.size() - 1) throw EvalError(B200FHE_ERR_ARG, "LongDivisionChebyshev: The dominant coefficient of the divisor is zero.");
    LongDiv out;
    if (int32_t(n - k) < 0) {
        out.q.assign(1, 0.0);
        out.r = f;
        return out;
    }
    std::vector<double> q(n - k + 1), r(f), d;
    const double lead_g = g.back();
    const bool g_scaled = not_one(g[k]);
    while (int32_t(n - k) > 0) {
        d.assign(n + 1, 0.0);
        const uint32_t s = n - k;  // degree of the quotient term
        q[s] = 2 * r.back();
        if (g_scaled) q[s] /= lead_g;
        if (k == s) {
            d[0] = 2 * g[s];
            for (uint32_t i = 1; i < 2 * k + 1; i++) d[i] = g[std::abs(int32_t(s - i))];
        } else if (int32_t(k) > int32_t(s)) {
            d[0] = 2 * g[s];
            for (uint32_t i = 1; i < k - s + 1; i++) d[i] = g[std::abs(int32_t(s - i))] + g[int32_t(s + i)];
            for (uint32_t i = k - s + 1; i < n + 1; i++) d[i] = g[std::abs(int32_t(i - s))];
        } else {
            d[s] = g.front();
            for (uint32_t i = n - 2 * k; i < n + 1; i++)
                if (i != s) d[i] = g[std::abs(int32_t(i - s))];
        }
        const double lead_r = r.back();
        if (not_one(lead_r)) scale_all(d, lead_r, false);
        if (not_one(lead_g)) scale_all(d, lead_g, true);
        subtract_and_trim(r, d, n);
    }
    if (n == k) {
        d = g;
        q.front() = r.back();
        if (not_one(lead_g)) q.front() /= lead_g;
        const double lead_r = r.back();
        if (not_one(lead_r)) scale_all(d, lead_r, false);
        if (not_one(lead_g)) scale_all(d, lead_g, true);
        subtract_and_trim(r, d, n);
    }
    q.front() *= 2;  // the constant term is kept as c0, not c0/2
    out.q = std::move(q);
    out.r = std::move(r);
    return out;
}

// ComputeDegreesPS (src/util.cu:267-300, table :15-38): degree n -> (k, m) with n < k(2^m - 1)
void compute_degrees_ps(uint32_t n, uint32_t &k, uint32_t &m) {
    if (n == 0) throw EvalError(B200FHE_ERR_ARG, "ComputeDegreesPS: The degree is zero.");
    static const struct {
        uint32_t upto, m;
    } table[] = {{2, 1},    {11, 2},   {13, 3},   {17, 2},   {55, 3},   {59, 4},   {76, 3},   {239, 4},  {247, 5},
                 {284, 4},  {991, 5},  {1007, 6}, {1083, 5}, {2015, 6}, {2031, 7}, {2204, 6}};
    if (n <= 2204) {
        for (const auto &e : table)
            if (n <= e.upto) {
                m = e.m;
                break;
            }
        k = (uint32_t) std::floor(n / ((1 << m) - 1)) + 1;
        return;
    }
    // heuristic for larger degrees: fewest homomorphic multiplications k + 2m + 2^(m-1) - 4 with k near sqrt(n/2)
    uint32_t best = ~0u;
    for (uint32_t kk = 1; kk <= n; kk++)
        for (uint32_t mm = 1; mm <= std::ceil(log2(n / kk) + 1) + 1; mm++)
            if (int32_t(n) - int32_t(kk * ((1 << mm) - 1)) < 0 &&
                std::abs(std::floor(log2(kk)) - std::floor(log2(sqrt(n / 2)))) <= 1) {
                const uint32_t mults = kk + 2 * mm + (1 << (mm - 1)) - 4;
                if (mults < best) {
                    best = mults;
                    k = kk;
                    m = mm;
                }
            }
}

// EvalLinearWSumMutable (src/evaluate.cu:3537-3583): the inputs are adjusted IN PLACE to the deepest one
Ct Evaluator::linear_wsum(CtVec &cts, const std::vector<double> &w) {
    size_t maxLevel = chain_index(*cts[0]);
    uint32_t maxIdx = 0;
    for (uint32_t i = 1; i < cts.size(); i++) {
        const size_t ci = chain_index(*cts[i]);
        if (ci > maxLevel || (ci == maxLevel && cts[i]->noise == 2)) {
            maxLevel = ci;
            maxIdx = i;
        }
    }
    for (uint32_t i = 0; i < cts.size(); i++)
        if (i != maxIdx) adjust_levels_and_depth(*cts[i], *cts[maxIdx]);
    if (cts[maxIdx]->noise == 2)
        for (auto &p : cts) mod_reduce(*p, 1);
    Ct sum = *cts[0];
    mult_const(sum, w[0]);
    for (uint32_t i = 1; i < cts.size(); i++) {
        Ct tmp = *cts[i];
        mult_const(tmp, w[i]);
        add_auto(sum, tmp);
    }
    return sum;
}

Ct Evaluator::chebyshev_series(const Ct &x, const std::vector<double> &cf, double a, double b) {
    return poly_degree(cf) < 5 ? chebyshev_linear(x, cf, a, b) : chebyshev_ps(x, cf, a, b);
}

// T_1 .. T_k by the binary-tree recurrences (shared by the Linear and PS variants, evaluate.cu:3195-3240, :3283-3330)
static Evaluator::CtVec chebyshev_basis(Evaluator &ev, const Ct &x, uint32_t k, double a, double b) {
    Evaluator::CtVec T(k);
    if ((a - std::round(a) < 1e-10) && (b - std::round(b) < 1e-10) && (std::round(a) == -1) && (std::round(b) == 1)) {
        T[0] = std::make_shared<Ct>(x);
    } else {
        const double alpha = 2 / (b - a), beta = 2 * a / (b - a);
        T[0] = std::make_shared<Ct>(x);
        ev.mult_const(*T[0], alpha);
        ev.add_const(*T[0], -1.0 - beta);
    }
    const Ct y = *T[0];
    for (uint32_t i = 2; i <= k; i++) {
        if (!(i & (i - 1)) || i % 2 == 0) {
            // T_i = 2 T_{i/2}^2 - 1
            Ct sq = ev.square(*T[i / 2 - 1]);
            Ct dbl = sq;
            ev.add_auto(dbl, sq);
            ev.add_const(dbl, -1.0);
            T[i - 1] = std::make_shared<Ct>(dbl);
        } else {
            // T_i = 2 T_{(i-1)/2} T_{(i+1)/2} - T_1
            Ct prod = ev.mult_auto(*T[i / 2 - 1], *T[i / 2]);
            Ct dbl = prod;
            ev.add_auto(dbl, prod);
            ev.sub_auto(dbl, y);
            T[i - 1] = std::make_shared<Ct>(dbl);
        }
    }
    return T;
}

Ct Evaluator::chebyshev_linear(const Ct &x, const std::vector<double> &cf, double a, double b) {
    const uint32_t k = (uint32_t) cf.size() - 1;
    CtVec T = chebyshev_basis(*this, x, k, a, b);
    Ct result = *T[k - 1];
    mult_const(result, cf[k]);
    for (size_t i = 0; i + 1 < k; i++) {
        if (cf[i + 1] != 0) {
            mult_const(*T[i], cf[i + 1]);
            add_auto(result, *T[i]);
        }
    }
    add_const(result, cf[0] / 2);
    return result;
}

// sum_{i < n} w[i+1] T_{i+1}, through EvalLinearWSumMutable on the first n basis ciphertexts
static Ct wsum_prefix(Evaluator &ev, Evaluator::CtVec &T, const std::vector<double> &w, uint32_t n) {
    Evaluator::CtVec ctxs(T.begin(), T.begin() + n);
    std::vector<double> weights(w.begin() + 1, w.begin() + 1 + n);
    return ev.linear_wsum(ctxs, weights);
}

// Shared body of EvalChebyshevSeriesPS (top == true, evaluate.cu:3378-3532) and InnerEvalChebyshevPS
// (top == false, :2998-3174): split f = q * T_{k 2^(m-1)} + r, r2 = r - T_{k(2^(m-1)-1)} = c * q + s, and evaluate
// (T_{k 2^(m-1)} + c) * q + s  with  s2 = s + T_{k(2^(m-1)-1)}.
Ct Evaluator::ps_combine(const std::vector<double> &f, uint32_t k, uint32_t m, CtVec &T, CtVec &T2, bool top) {
    const uint32_t k2m2k = k * (1 << (m - 1)) - k;
    std::vector<double> Tkm(int32_t(k2m2k + k) + 1, 0.0);
    Tkm.back() = 1;
    const LongDiv divqr = long_division_chebyshev(f, Tkm);
    std::vector<double> r2 = divqr.r;
    if (int32_t(k2m2k - poly_degree(divqr.r)) <= 0) {
        r2[int32_t(k2m2k)] -= 1;
        r2.resize(poly_degree(r2) + 1);
    } else {
        r2.resize(int32_t(k2m2k + 1), 0.0);
        r2.back() = -1;
    }
    const LongDiv divcs = long_division_chebyshev(r2, divqr.q);
    std::vector<double> s2 = divcs.r;
    s2.resize(int32_t(k2m2k + 1), 0.0);
    s2.back() = 1;

    // c(u)
    Ct cu_;
    const uint32_t dc = poly_degree(divcs.q);
    bool flag_c = false;
    if (dc >= 1) {
        if (dc == 1) {
            cu_ = *T.front();
            if (divcs.q[1] != 1) mult_const(cu_, divcs.q[1]);
        } else {
            cu_ = wsum_prefix(*this, T, divcs.q, dc);
        }
        add_const(cu_, divcs.q.front() / 2);
        flag_c = true;
    }

    // q(u)
    Ct qu;
    if (poly_degree(divqr.q) > k) {
        qu = inner_chebyshev_ps(divqr.q, k, m - 1, T, T2);
    } else {
        std::vector<double> qcopy = divqr.q;
        qcopy.resize(k);
        const uint32_t dq = poly_degree(qcopy);
        if (dq > 0) {
            qu = wsum_prefix(*this, T, divqr.q, dq);
            Ct sum = *T[k - 1];
            if (top) {
                // leading coefficient is 2 after one Chebyshev division
                add_auto(sum, *T[k - 1]);
            } else {
                // leading coefficient is a power of two up to 2^(m-1)
                for (uint32_t i = 0; i < log2(divqr.q.back()); i++) {
                    Ct s = sum;
                    add_auto(sum, s);
                }
            }
            add_auto(qu, sum);
        } else if (top) {
            qu = *T[k - 1];
            for (uint32_t i = 1; i < divqr.q.back(); i++) add_auto(qu, *T[k - 1]);
        } else {
            Ct sum = *T[k - 1];
            for (uint32_t i = 0; i < log2(divqr.q.back()); i++) {
                Ct s = sum;
                add_auto(sum, s);
            }
            qu = sum;
        }
        add_const(qu, divqr.q.front() / 2);
    }

    // s(u)
    Ct su;
    if (poly_degree(s2) > k) {
        su = inner_chebyshev_ps(s2, k, m - 1, T, T2);
    } else {
        std::vector<double> scopy = s2;
        scopy.resize(k);
        const uint32_t ds = poly_degree(scopy);
        if (ds > 0) {
            su = wsum_prefix(*this, T, s2, ds);
            add_auto(su, *T[k - 1]);  // s2 is monic
        } else {
            su = *T[k - 1];
        }
        add_const(su, s2.front() / 2);
    }

    Ct result = *T2[m - 1];
    if (flag_c) add_auto(result, cu_);
    else add_const(result, divcs.q.front() / 2);
    result = mult_auto(result, qu);
    add_auto(result, su);
    return result;
}

Ct Evaluator::inner_chebyshev_ps(const std::vector<double> &cf, uint32_t k, uint32_t m, CtVec &T, CtVec &T2) {
    return ps_combine(cf, k, m, T, T2, false);
}

// EvalChebyshevSeriesPS (src/evaluate.cu:3264-3535)
Ct Evaluator::chebyshev_ps(const Ct &x, const std::vector<double> &cf, double a, double b) {
    const uint32_t n = poly_degree(cf);
    std::vector<double> f2 = cf;
    if (cf[cf.size() - 1] == 0) f2.resize(n + 1);
    uint32_t k = 0, m = 0;
    compute_degrees_ps(n, k, m);
    CtVec T = chebyshev_basis(*this, x, k, a, b);
    for (size_t i = 1; i < k; i++) adjust_levels_and_depth(*T[i - 1], *T[k - 1]);

    // T_k, T_2k, T_4k, ... ; T2[0] IS T[k-1] (same object, as in the reference)
    CtVec T2(m);
    T2.front() = T.back();
    for (uint32_t i = 1; i < m; i++) {
        Ct sq = square(*T2[i - 1]);
        Ct dbl = sq;
        add_auto(dbl, sq);
        add_const(dbl, -1.0);
        T2[i] = std::make_shared<Ct>(dbl);
    }
    // T_{k(2^m - 1)}
    std::shared_ptr<Ct> T2km1 = T2.front();
    for (uint32_t i = 1; i < m; i++) {
        Ct prod = mult_auto(*T2km1, *T2[i]);
        Ct dbl = prod;
        add_auto(dbl, prod);
        sub_auto(dbl, *T2.front());
        T2km1 = std::make_shared<Ct>(dbl);
    }
    const uint32_t k2m2k = k * (1 << (m - 1)) - k;
    f2.resize(2 * k2m2k + k + 1, 0.0);
    f2.back() = 1;
    Ct result = ps_combine(f2, k, m, T, T2, true);
    sub_auto(result, *T2km1);
    return result;
}

}  // namespace b200fhe
