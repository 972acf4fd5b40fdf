// bootsetup.cu — EvalBootstrapSetup with precompute == true: the plaintext diagonals of the collapsed-FFT linear
// transforms, computed on the host in double precision and encoded in the extended basis with the B200 encoder.
// Replaces src/bootstrap.cu:88-177 (setup), :183-394 (EvalCoeffsToSlotsPrecompute), :396-593
// (EvalSlotsToCoeffsPrecompute) and src/util.cu:328-366 (Rotate), :381-476 (CoeffEncoding/DecodingOneLevel),
// :480-726 (CoeffEncoding/DecodingCollapse).  The floating-point expressions and their order are the reference's
// (OpenFHE's): every coefficient is later rounded to an integer, so the diagonals come out bit-identical
// (tests/test_bootstrap_gpu.py::test_setup_diagonals_match_reference).
#include <cmath>
#include <complex>

#include "../../include/b200fhe.h"
#include "bootstrapper.h"
#include "encoder.h"

namespace b200fhe {

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

typedef std::complex<double> cd;
typedef std::vector<cd> Row;
typedef std::vector<Row> Rows;

static const cd kI(0.0, 1.0);

// the three weight vectors of each FFT layer: [s] input shifted right, [s + log] unshifted, [s + 2 log] shifted left
static Rows one_level(const Row &pows, const std::vector<uint32_t> &rotGroup, bool flag_i, bool decoding) {
    const uint32_t dim = (uint32_t) pows.size() - 1, slots = (uint32_t) rotGroup.size();
    const uint32_t lg = (uint32_t) std::log2(slots);
    Rows coeff(3 * lg, Row(slots));
    for (uint32_t m = 2; m <= slots; m <<= 1) {  // the layers are independent of each other: any order
        const uint32_t s = (uint32_t) std::log2(m) - 1, lenh = m >> 1, lenq = m << 2;
        for (uint32_t k = 0; k < slots; k += m) {
            for (uint32_t j = 0; j < lenh; j++) {
                const uint32_t r = rotGroup[j] % lenq;
                const uint32_t jTwiddle = (decoding ? r : lenq - r) * (dim / lenq);
                if (!decoding) {
                    if (flag_i && m == 2) {
                        const cd e = std::exp(-M_PI / 2 * kI), w = e * pows[jTwiddle];
                        coeff[s + lg][j + k] = e;
                        coeff[s + 2 * lg][j + k] = e;
                        coeff[s + lg][j + k + lenh] = -w;
                        coeff[s][j + k + lenh] = w;
                    } else {
                        const cd w = pows[jTwiddle];
                        coeff[s + lg][j + k] = 1;
                        coeff[s + 2 * lg][j + k] = 1;
                        coeff[s + lg][j + k + lenh] = -w;
                        coeff[s][j + k + lenh] = w;
                    }
                } else {
                    if (flag_i && m == 2) {
                        const cd e = std::exp(M_PI / 2 * kI), w = e * pows[jTwiddle];
                        coeff[s + lg][j + k] = e;
                        coeff[s + 2 * lg][j + k] = w;
                        coeff[s + lg][j + k + lenh] = -w;
                        coeff[s][j + k + lenh] = e;
                    } else {
                        const cd w = pows[jTwiddle];
                        coeff[s + lg][j + k] = 1;
                        coeff[s + 2 * lg][j + k] = w;
                        coeff[s + lg][j + k + lenh] = -w;
                        coeff[s][j + k + lenh] = 1;
                    }
                }
            }
        }
    }
    return coeff;
}

// merge `layers` consecutive FFT layers of one stage into 2^(layers+1) - 1 diagonals (encoding direction)
static Rows collapse_encoding(const Rows &c1, uint32_t slots, int32_t top, int32_t layers, uint32_t n_out) {
    const uint32_t lg = (uint32_t) std::log2(slots);
    Rows cur(n_out, Row(slots));
    for (int32_t l = 0; l < layers; l++) {
        if (l == 0) {
            cur[0] = c1[top];
            cur[1] = c1[top + lg];
            cur[2] = c1[top + 2 * lg];
            continue;
        }
        const Rows temp = cur;
        cur.assign(n_out, Row(slots, 0.0));
        const int32_t shift = 1 << (top - l);
        uint32_t t = 0;
        for (int32_t u = 0; u < (1 << (l + 1)) - 1; u++) {
            for (uint32_t k = 0; k < slots; k++) {
                cur[u + t][k] += c1[top - l][k] * temp[u][reduce_rotation(int32_t(k) - shift, slots)];
                cur[u + t + 1][k] += c1[top - l + lg][k] * temp[u][k];
                cur[u + t + 2][k] += c1[top - l + 2 * lg][k] * temp[u][reduce_rotation(int32_t(k) + shift, slots)];
            }
            t += 1;
        }
    }
    return cur;
}

// decoding direction: the layers of a stage start at `first`
static Rows collapse_decoding(const Rows &c1, uint32_t slots, int32_t first, int32_t layers, uint32_t n_out) {
    const uint32_t lg = (uint32_t) std::log2(slots);
    Rows cur(n_out, Row(slots));
    for (int32_t l = 0; l < layers; l++) {
        if (l == 0) {
            cur[0] = c1[first];
            cur[1] = c1[lg + first];
            cur[2] = c1[2 * lg + first];
            continue;
        }
        const Rows temp = cur;
        cur.assign(n_out, Row(slots, 0.0));
        for (uint32_t t = 0; t < 3; t++) {
            const Row &wt = c1[first + l + t * lg];
            const uint32_t off = t == 0 ? 0 : (t == 1 ? (1u << l) : (1u << (l + 1)));
            for (int32_t u = 0; u < (1 << (l + 1)) - 1; u++)
                for (uint32_t k = 0; k < slots; k++) cur[u + off][k] += wt[k] * temp[u][k];
        }
    }
    return cur;
}

// all stages of CoeffEncodingCollapse / CoeffDecodingCollapse
static std::vector<Rows> collapse(const Row &pows, const std::vector<uint32_t> &rotGroup, const FFTParams &p, bool flag_i,
                                  bool decoding) {
    const uint32_t slots = (uint32_t) rotGroup.size();
    const int32_t lb = p.levelBudget, flagRem = p.remCollapse != 0 ? 1 : 0;
    const Rows c1 = one_level(pows, rotGroup, flag_i, decoding);
    std::vector<Rows> coeff(lb);
    if (!decoding) {
        const int32_t stop = flagRem ? 0 : -1;
        for (int32_t s = lb - 1; s > stop; s--) {
            const int32_t top = int32_t(std::log2(slots)) - (lb - 1 - s) * p.layersCollapse - 1;
            coeff[s] = collapse_encoding(c1, slots, top, p.layersCollapse, (uint32_t) p.numRotations);
        }
        if (flagRem) {
            const int32_t top = int32_t(std::log2(slots)) - (lb - 1) * p.layersCollapse - 1;
            coeff[0] = collapse_encoding(c1, slots, top, p.remCollapse, (uint32_t) p.numRotationsRem);
        }
    } else {
        const int32_t rows = lb - flagRem;  // SelectLayers: rowsCollapse
        for (int32_t s = 0; s < rows; s++)
            coeff[s] = collapse_decoding(c1, slots, s * p.layersCollapse, p.layersCollapse, (uint32_t) p.numRotations);
        if (flagRem)
            coeff[rows] = collapse_decoding(c1, slots, rows * p.layersCollapse, p.remCollapse, (uint32_t) p.numRotationsRem);
    }
    return coeff;
}

// Rotate (src/util.cu:328-353): cyclic left rotation of the slot vector
static std::vector<cuDoubleComplex> rotate(const Row &a, int32_t index) {
    const int32_t n = (int32_t) a.size();
    if (index < 0 || index > n) index = (int32_t) reduce_rotation(index, (uint32_t) n);
    std::vector<cuDoubleComplex> r(n);
    for (int32_t i = 0; i < n; i++) {
        const cd &v = a[(i + index) % n];
        r[i] = make_cuDoubleComplex(v.real(), v.imag());
    }
    return r;
}

void Bootstrapper::precompute(Encoder &encoder) {
    Context *c = ev.c;
    if (c->host_only) throw EvalError(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    const uint32_t N = (uint32_t) c->N, M = 2 * N, m = 4 * slots;
    const bool sparse = M != m;
    std::vector<uint32_t> rotGroup(slots);
    uint32_t fivePows = 1;
    for (uint32_t i = 0; i < slots; i++) {
        rotGroup[i] = fivePows;
        fivePows *= 5;
        fivePows %= m;
    }
    Row ksiPows(m + 1);
    for (uint32_t j = 0; j < m; j++) {
        const double angle = 2.0 * M_PI * j / m;
        ksiPows[j].real(cos(angle));
        ksiPows[j].imag(sin(angle));
    }
    ksiPows[m] = ksiPows[0];
    const double qDouble = static_cast<double>(c->primes[0]);
    const unsigned __int128 factor = (unsigned __int128) 1 << ((uint32_t) std::round(std::log2(qDouble)));
    const double pre = qDouble / factor;
    const double scaleEnc = pre / 1.0, scaleDec = 1 / pre;

    owned_.clear();
    for (int which = 0; which < 2; which++) {
        const FFTParams &p = which == 0 ? enc : dec;
        const double scale = which == 0 ? scaleEnc : scaleDec;
        const int32_t lb = p.levelBudget, flagRem = p.remCollapse != 0 ? 1 : 0;
        std::vector<Rows> coeff = collapse(ksiPows, rotGroup, p, false, which == 1), coeffi;
        if (sparse) coeffi = collapse(ksiPows, rotGroup, p, true, which == 1);
        for (int32_t s = 0; s < lb; s++) {
            // the remainder stage is the first one when encoding and the last one when decoding
            const bool rem = flagRem && (which == 0 ? s == 0 : s == lb - 1);
            const int32_t g = rem ? p.gRem : p.g, b = rem ? p.bRem : p.b, nrot = rem ? p.numRotationsRem : p.numRotations;
            // the constant factor goes into the stage that is applied last
            const bool scaled = rem || (!flagRem && (which == 0 ? s == 0 : s == lb - 1));
            const int32_t sh = which == 0 ? (rem ? 1 : 1 << ((s - flagRem) * p.layersCollapse + p.remCollapse))
                                          : 1 << (s * p.layersCollapse);
            const size_t Ql = stage_level(which, (size_t) s), QlP = Ql + c->size_P;
            const double pt_scale = ev.sf[c->size_Q - Ql];
            auto block = std::make_unique<DevBuf>();
            if (block->alloc(size_t(nrot) * QlP * N * 8) != cudaSuccess)
                throw EvalError(B200FHE_ERR_ALLOC, "device memory for the plaintext diagonals");
            for (int32_t i = 0; i < b; i++) {
                for (int32_t j = 0; j < g; j++) {
                    const int32_t idx = g * i + j;
                    if (idx == nrot) continue;
                    const int32_t rot = (int32_t) reduce_rotation(-g * i * sh, sparse ? M / 4 : slots);
                    Row v = coeff[s][idx];
                    if (sparse) v.insert(v.end(), coeffi[s][idx].begin(), coeffi[s][idx].end());
                    if (scaled)
                        for (cd &x : v) x *= scale;
                    const std::vector<cuDoubleComplex> r = rotate(v, rot);
                    u64 *dst = block->as<u64>() + size_t(idx) * QlP * N;
                    const int rc = encoder.encode(r.data(), r.size(), Ql, pt_scale, true, sparse, dst, ev.st);
                    if (rc != B200FHE_OK) throw EvalError(rc, b200fhe_last_error());
                    set_diagonal(which, (size_t) s, (size_t) idx, dst, Ql, pt_scale);
                }
            }
            owned_.push_back(std::move(block));
        }
    }
}

}  // namespace b200fhe
