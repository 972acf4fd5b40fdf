// dnn.cu — encrypted CNN layers (see dnn.h).  Follows phantom::DNN (src/dnn.cu) operation by operation for the
// bookkeeping (levels, scales, NoiseScaleDeg) and the plaintext contents, so every output limb equals the reference's.
#include <cstdlib>
#include "dnn.h"

#include <algorithm>
#include <cmath>
#include <exception>
#include <thread>

#include "../../include/b200fhe.h"

namespace b200fhe {

// EvalChebyshevCoefficients (src/evaluate.cu:3585-3609)
std::vector<double> chebyshev_coefficients(const std::function<double(double)> &func, double a, double b, uint32_t degree) {
    if (!degree) throw EvalError(B200FHE_ERR_ARG, "The degree of approximation can not be zero");
    const size_t coeffTotal = degree + 1;
    const double bMinusA = 0.5 * (b - a), bPlusA = 0.5 * (b + a);
    const double PiByDeg = 3.14159265358979323846 / static_cast<double>(coeffTotal);
    std::vector<double> functionPoints(coeffTotal);
    for (size_t i = 0; i < coeffTotal; ++i) functionPoints[i] = func(std::cos(PiByDeg * (i + 0.5)) * bMinusA + bPlusA);
    const double multFactor = 2.0 / static_cast<double>(coeffTotal);
    std::vector<double> coefficients(coeffTotal);
    for (size_t i = 0; i < coeffTotal; ++i) {
        for (size_t j = 0; j < coeffTotal; ++j) coefficients[i] += functionPoints[j] * std::cos(PiByDeg * i * (j + 0.5));
        coefficients[i] *= multFactor;
    }
    return coefficients;
}

// src/dnn.cu:283-306
void DNN::add_rotation_indices(std::vector<int32_t> &v, size_t width, size_t kernel_h, size_t slotstr) {
    const size_t large_L = width * (size_t(1) << slotstr);
    for (int j = 0; j < (int) kernel_h; ++j)
        for (int i = 0; i < (int) kernel_h; ++i) {
            const int rot = (i - (int) (kernel_h / 2)) * (int) large_L + (j - (int) (kernel_h / 2));
            const int32_t idx = rot << (int) slotstr;
            if (std::find(v.begin(), v.end(), idx) == v.end()) v.push_back(idx);
        }
}

// src/dnn.cu:308-330
void DNN::add_avgpool_rotations(std::vector<int32_t> &v, size_t width, size_t slotstr) {
    const int log_l = (int) std::log2(width), pow_slotstr = 1 << slotstr;
    auto put = [&](int32_t r) {
        if (std::find(v.begin(), v.end(), r) == v.end()) v.push_back(r);
    };
    for (int i = 0; i < log_l; ++i) put(pow_slotstr << i);
    for (int j = 0; j < log_l; ++j) put((pow_slotstr << j) * (int) width);
}

// EncTensor (src/dnn.cu:10-41): channel k -> encode_sparse at the first chain index, encrypt_asymmetric
TensorCT DNN::enc_tensor(const double *image, size_t width, size_t num_ch, KeyGen &kg, const u64 *pk) {
    TensorCT r;
    r.width = width;
    r.num_ch = num_ch;
    r.slotstr = 0;
    const size_t Ql = ev->c->size_Q, N = ev->c->N, n = width * width;
    auto ptb = ev->pool.get(Ql * N * 8);
    std::vector<cuDoubleComplex> vk(n);
    for (size_t k = 0; k < num_ch; k++) {
        for (size_t x = 0; x < n; x++) vk[x] = make_cuDoubleComplex(image[x * num_ch + k], 0.0);
        ev->ok(enc->encode(vk.data(), n, Ql, scale, false, true, reinterpret_cast<u64 *>(ptb->p), ev->st));
        Ct ct = ev->alloc(Ql, 2, false);
        ct.scale = scale;
        ct.noise = 1;
        if (kg.st != ev->st) ev->cu(cudaStreamSynchronize(ev->st), "enc_tensor");
        kg.encrypt_asymmetric(pk, reinterpret_cast<const u64 *>(ptb->p), Ql, ct.data());
        if (kg.st != ev->st) ev->cu(cudaStreamSynchronize(kg.st), "enc_tensor");
        r.ctks.push_back(ct);
    }
    return r;
}

// DecTensor (src/dnn.cu:43-76).  A sparsely encoded message repeats with the period of its slot count, so the first
// (width << slotstr)^2 values of the full decode are the sparse decode.
void DNN::dec_tensor(const TensorCT &t, KeyGen &kg, double *out) {
    const size_t N = ev->c->N;
    std::vector<cuDoubleComplex> vals(N / 2);
    for (size_t k = 0; k < t.num_ch && k < t.ctks.size(); k++) {
        const Ct &ct = t.ctks[k];
        auto pb = ev->pool.get(ct.Ql * N * 8);
        const u64 *src = ct.data();
        if (kg.st != ev->st) ev->cu(cudaStreamSynchronize(ev->st), "dec_tensor");
        kg.decrypt(src, ct.Ql, ct.polys, reinterpret_cast<u64 *>(pb->p));
        if (kg.st != ev->st) ev->cu(cudaStreamSynchronize(kg.st), "dec_tensor");
        ev->ok(enc->decode(reinterpret_cast<const u64 *>(pb->p), ct.Ql, false, ct.scale, vals.data(), ev->st));
        for (size_t x = 0; x < t.width * t.width; x++) out[x * t.num_ch + k] = vals[x].x;
    }
}

// Conv (src/dnn.cu:82-150)
TensorCT DNN::conv(const TensorCT &in, const double *weight, size_t kh, size_t out_ch, size_t stride) {
    KeyScope keys(*this);
    if (stride != 1 && stride != 2) throw EvalError(B200FHE_ERR_ARG, "stride must be 1 or 2");
    const size_t in_ch = in.num_ch;
    if (in_ch == 0 || in.ctks.size() < in_ch || out_ch == 0 || kh == 0) throw EvalError(B200FHE_ERR_ARG, "empty tensor");
    Context *c = ev->c;
    const size_t N = c->N;
    const int pow_slotstr = 1 << in.slotstr, width = (int) in.width, large_L = width * pow_slotstr, half = (int) (kh / 2);
    const size_t n_slots = size_t(large_L) * large_L, R = kh * kh, J = in_ch * R;
    const Ct &f = in.ctks[0];
    for (size_t k = 0; k < in_ch; k++) {
        const Ct &x = in.ctks[k];
        if (x.ext() || x.polys != 2 || x.batch != 1) throw EvalError(B200FHE_ERR_ARG, "encrypted size must be 2");
        if (x.Ql != f.Ql || x.scale != f.scale || x.noise != f.noise)
            throw EvalError(B200FHE_ERR_UNSUPPORTED, "conv: input channels differ in level, scale or NoiseScaleDeg");
    }
    // taps in the reference's loop order (j outer, i inner)
    std::vector<int32_t> rots;
    for (int j = 0; j < (int) kh; j++)
        for (int i = 0; i < (int) kh; i++) rots.push_back(((i - half) * large_L + j - half) * pow_slotstr);

    // ONE image on several GPUs (dnn.h `Ranks`): a convolution is a sum over the input channels, so rank r rotates and
    // multiplies only ITS run of input channels [k0, k1) -- rotations, plaintext encodes and products all shrink by the number
    // of ranks -- and the partial sums of every output channel are added up across the ranks below (exact: modular addition
    // of canonical residues).  Fewer input channels than ranks (the first layer): replicated.
    const bool by_rank = ranks.world > 1 && in_ch >= ranks.world;
    const size_t k0 = by_rank ? in_ch * ranks.rank / ranks.world : 0, k1 = by_rank ? in_ch * (ranks.rank + 1) / ranks.world : in_ch;
    const size_t J_all = J;
    (void) J_all;
    const size_t Jm = (k1 - k0) * R;   // products this process computes per output channel

    // 1. every input channel (of this rank) rotated by every tap, once: X [Jm][2][Ql][N]
    Ct X = ev->alloc(f.Ql, 2, false, (u32) Jm);
    X.scale = f.scale;
    X.noise = f.noise;
    for (size_t k = k0; k < k1; k++) ev->rotate_fused_many(in.ctks[k], rots, X.data() + (k - k0) * R * 2 * f.Ql * N);

    // 2. the ciphertext side of EvalMultAutoInplace (src/evaluate.cu:2813-2825; plaintext at the same level with
    //    NoiseScaleDeg 1): nothing for a NoiseScaleDeg-1 input, one rescale for a NoiseScaleDeg-2 input
    const u32 pt_noise0 = 1;
    const bool reduce = X.noise == 2;
    const bool lift_pt = pt_noise0 < f.noise;
    if (reduce) ev->mod_reduce(X, 1);
    const size_t Qm = X.Ql;  // level of the products
    const u64 *xd = X.data();

    // 3. output channels, HB at a time: encode the Jm plaintexts of each, bring them to the product level as one
    //    batched single-polynomial ciphertext, and take the plaintext-matrix product
    // Compact plaintexts: a weight plaintext is a polynomial in X^(N / 2 n_slots), whose NTT repeats every value
    // N / 2 n_slots times (ntt.cuh, ntt_forward_sub) -- only the 2 n_slots distinct values are computed and kept, and the
    // product kernel reads value c >> pt_shift for coefficient c.  Same limbs as the reference's full-size plaintexts.
    const bool fused_lift = lift_pt && reduce && f.Ql >= 2;
    u32 pt_shift = 0;
    while (((2 * n_slots) << pt_shift) < N) pt_shift++;
    const bool compact = 2 * n_slots <= 2048 && pt_shift >= 1 && pt_shift <= 8 && (fused_lift || (!lift_pt && !reduce));
    if (!compact) pt_shift = 0;
    const size_t row = compact ? 2 * n_slots : N;  // words per plaintext limb
    const size_t pt_bytes = Jm * f.Ql * row * 8;
    const size_t HB = std::max<size_t>(1, std::min<size_t>({size_t(4), out_ch, (size_t(4) << 30) / pt_bytes}));
    TensorCT result;
    result.width = in.width / stride;
    result.num_ch = out_ch;
    result.slotstr = stride == 2 ? in.slotstr + 1 : in.slotstr;
    result.ctks.resize(out_ch);
    std::vector<cuDoubleComplex> wsp;
    for (size_t h0 = 0; h0 < out_ch; h0 += HB) {
        const size_t hb = std::min(HB, out_ch - h0);
        // A NoiseScaleDeg-2 input: the plaintexts are lifted by EvalMultConstInplaceCore(pt, 1.0) and rescaled; both
        // happen on the coefficients inside the encoder, ahead of its one forward transform (encoder.cu)
        const bool fused_pt = fused_lift && n_slots <= 2048;
        const size_t pt_limbs = fused_pt ? f.Ql - 1 : f.Ql;
        Ct P = ev->alloc(pt_limbs, 1, false, (u32) ((hb * Jm * row + N - 1) / N));
        P.scale = scale;
        P.noise = pt_noise0;
        wsp.assign(hb * Jm * n_slots, make_cuDoubleComplex(0.0, 0.0));
        for (size_t hh = 0; hh < hb; hh++) {
            const size_t h = h0 + hh;
            size_t jj = 0;
            for (size_t k = k0; k < k1; k++)
                for (int j = 0; j < (int) kh; j++)
                    for (int i = 0; i < (int) kh; i++, jj++) {
                        cuDoubleComplex *v = wsp.data() + (hh * Jm + jj) * n_slots;
                        const double w = weight[((size_t(i) * kh + j) * in_ch + k) * out_ch + h];
                        for (int jp = 0; jp < width; jp++)
                            for (int ip = 0; ip < width; ip++)
                                if (ip + i - half <= width - 1 && jp + j - half <= width - 1 && ip + i - half >= 0 && jp + j - half >= 0)
                                    v[size_t(ip * large_L + jp) * pow_slotstr] = make_cuDoubleComplex(w, 0.0);
                    }
        }
        if (fused_pt) {
            Ct top;  // metadata of the plaintext at the input's level, for the scalar GetElementForEvalMult builds
            top.Ql = top.limbs = f.Ql;
            top.scale = scale;
            top.noise = pt_noise0;
            const std::vector<uint64_t> w = ev->element_for_mult(top, 1.0);
            LimbScalars lift{};
            for (size_t i = 0; i < f.Ql; i++) lift.w[i] = w[i] % c->primes[i];
            DeviceLevel *L;
            ev->ok(get_level(c, f.Ql, &L));
            int rc = enc->encode_sparse_batch(wsp.data(), n_slots, hb * Jm, f.Ql, scale, P.data(), ev->st, &lift, L->inv_q_last.as<u64>(),
                                              compact);
            if (rc == B200FHE_OK) {
                P.scale = scale * ev->level_sf(top) / static_cast<double>(c->primes[f.Ql - 1]);
                P.noise = pt_noise0 + 1 - 1;
            } else if (rc == B200FHE_ERR_UNSUPPORTED && !compact) {  // coefficients beyond 2^128: the three-step way
                P = ev->alloc(f.Ql, 1, false, (u32) (hb * Jm));
                P.scale = scale;
                P.noise = pt_noise0;
                ev->ok(enc->encode_sparse_batch(wsp.data(), n_slots, hb * Jm, f.Ql, scale, P.data(), ev->st));
                ev->mult_const_core(P, 1.0);
                ev->mod_reduce(P, 1);
            } else
                ev->ok(rc);
        } else {
            ev->ok(enc->encode_sparse_batch(wsp.data(), n_slots, hb * Jm, f.Ql, scale, P.data(), ev->st, nullptr, nullptr, compact));
            if (lift_pt) ev->mult_const_core(P, 1.0);
            if (reduce) ev->mod_reduce(P, 1);
        }
        if (P.Ql != Qm) throw EvalError(B200FHE_ERR_ARG, "conv: plaintext level mismatch");
        const u64 *pd = P.data();
        std::vector<u64 *> outs(hb);
        for (size_t hh = 0; hh < hb; hh++) {
            Ct o = ev->alloc(Qm, 2, false);
            o.scale = X.scale * P.scale;
            o.noise = X.noise + P.noise;
            outs[hh] = o.data();
            result.ctks[h0 + hh] = o;
        }
        std::vector<const u64 *> ptab;
        for (size_t j0 = 0; j0 < Jm;) {
            const size_t g = std::min<size_t>(Jm - j0, j0 == 0 ? kMaxBsgs : kMaxBsgs - 1);
            BsgsArgs a{};
            ptab.assign(hb * g, nullptr);
            for (size_t j = 0; j < g; j++) {
                a.ct[j] = xd + (j0 + j) * 2 * Qm * N;
                for (size_t hh = 0; hh < hb; hh++) ptab[hh * g + j] = pd + (hh * Jm + j0 + j) * Qm * row;
            }
            a.pt = ptab.data();
            a.out = outs.data();
            a.g = (u32) g;
            a.nb = (u32) hb;
            a.size_QlP = (u32) Qm;
            a.accumulate = j0 == 0 ? 0 : 1;
            a.pt_shift = pt_shift;
            ev->cu(bsgs_matmul(c->T.mods, a, PrimeMap{0, 0}, N, ev->st), "conv product");
            count_launches((hb + 3) / 4);
            j0 += g;
        }
    }
    if (by_rank) {
        // partial sums -> sums.  Arena half: [slots: one block of partial ciphertexts per sending rank | sums: every output
        // channel].  Rank p owns the output channels [out_ch p / G, out_ch (p + 1) / G): every rank stores its partial sums of
        // those channels into p's slot for the sender (all-to-all), p adds the G slots up (row-sum kernel) and the sums are
        // all-gathered.
        const u32 G = ranks.world, me = ranks.rank;
        char *base = static_cast<char *>(ranks.bases[me]);
        const size_t ctb = size_t(2) * Qm * N * 8;                       // one result ciphertext
        auto h_lo = [&](u32 r) { return out_ch * r / G; };
        const size_t own_max = (out_ch + G - 1) / G;                     // channels a rank owns at most
        const size_t slots_bytes = size_t(G) * own_max * ctb, sums_bytes = out_ch * ctb;
        if (slots_bytes + sums_bytes > ranks.half_bytes) throw EvalError(B200FHE_ERR_ARG, "peer arena too small for this convolution");
        const size_t off0 = size_t(ranks.parity) * ranks.half_bytes, sums_off = off0 + slots_bytes;
        ranks.parity ^= 1;
        // the partial ciphertexts are separate blocks: stage them contiguously, in channel order, behind the sums region's
        // place for them is not free (peers write there), so use a pool block
        auto stage = ev->pool.get(out_ch * ctb);
        char *stg = static_cast<char *>(stage->p);
        for (size_t h = 0; h < out_ch; h++)
            ev->cu(cudaMemcpyAsync(stg + h * ctb, result.ctks[h].data(), ctb, cudaMemcpyDeviceToDevice, ev->st), "stage partial sums");
        PeerSeg segs[kMaxRanks][kPeerSegs] = {};
        for (u32 r = 0; r < G; r++) {
            const size_t a0 = h_lo(r), a1 = h_lo(r + 1);
            if (a1 == a0) continue;
            if (r == me) {
                ev->cu(cudaMemcpyAsync(base + off0 + me * own_max * ctb, stg + a0 * ctb, (a1 - a0) * ctb, cudaMemcpyDeviceToDevice, ev->st),
                       "own partial sums");
            } else {
                segs[r][0] = PeerSeg{stg + a0 * ctb, off0 + me * own_max * ctb, (a1 - a0) * ctb, 1, 0, 0};
            }
        }
        ev->ok(peer_all_to_all(ranks.bases, ranks.flags, ranks.counter, G, me, segs, ev->st));
        const size_t m0 = h_lo(me), m1 = h_lo(me + 1);
        if (m1 > m0)
            ev->ok(peer_row_sum(c->T.mods, reinterpret_cast<u64 *>(base + sums_off + m0 * ctb), reinterpret_cast<const u64 *>(base + off0),
                                own_max * ctb / 8, G, 0, (u32) ((m1 - m0) * 2 * Qm), (u32) Qm, PrimeMap{0, 0}, N, ev->st, true));
        const size_t off = sums_off + m0 * ctb, len = (m1 - m0) * ctb;
        ev->ok(peer_exchange_strided(base, ranks.bases, nullptr, ranks.flags, ranks.counter, G, me, &off, &len, 1, 1, 0, 0, ev->st));
        for (size_t h = 0; h < out_ch; h++) {
            const Ct &pc = result.ctks[h];
            result.ctks[h] = ev->wrap_copy(reinterpret_cast<const u64 *>(base + sums_off + h * ctb), Qm, 2, false, pc.scale, pc.noise);
        }
    }
    return result;
}

// BatchNorm (src/dnn.cu:449-476)
TensorCT DNN::batch_norm(const TensorCT &in, const double *weight, const double *bias, const double *mean, const double *var) {
    const double eps = 1e-5;
    TensorCT out = like(in);
    out.ctks.resize(in.num_ch);
    for (size_t ch = 0; ch < in.num_ch; ++ch) {
        const double a = weight[ch] / std::sqrt(var[ch] + eps);
        const double b = bias[ch] - a * mean[ch];
        Ct x = in.ctks[ch];
        ev->mult_const(x, a);
        ev->add_const(x, b);
        out.ctks[ch] = x;
    }
    return out;
}

// Add (src/dnn.cu:478-499)
TensorCT DNN::add(const TensorCT &a, const TensorCT &b) {
    if (a.num_ch != b.num_ch || a.width != b.width || a.slotstr != b.slotstr)
        throw EvalError(B200FHE_ERR_ARG, "TensorCT dimension mismatch in Add()");
    TensorCT r = like(a);
    r.ctks.resize(a.num_ch);
    for (size_t u = 0; u < a.num_ch; ++u) {
        Ct x = a.ctks[u];
        ev->add_auto(x, b.ctks[u]);
        r.ctks[u] = x;
    }
    return r;
}

// Relu (src/dnn.cu:152-170): Chebyshev interpolation of max(0, x) on [a, b]
TensorCT DNN::relu(const TensorCT &in, int a, int b, int deg) {
    KeyScope keys(*this);
    const std::vector<double> coeffs = chebyshev_coefficients([](double x) { return std::max(0.0, x); }, a, b, (uint32_t) deg);
    TensorCT r = like(in);
    for (size_t i = 0; i < in.num_ch; i++) r.ctks.push_back(ev->chebyshev_series(in.ctks[i], coeffs, a, b));
    return r;
}

// ReluDegTwo (src/dnn.cu:250-264): x + x^2
TensorCT DNN::relu_deg_two(const TensorCT &in) {
    KeyScope keys(*this);
    TensorCT r = like(in);
    for (size_t i = 0; i < in.num_ch; i++) {
        Ct x2 = ev->square(in.ctks[i]);
        Ct x = in.ctks[i];
        ev->add_auto(x, x2);
        r.ctks.push_back(x);
    }
    return r;
}

// Sign (src/dnn.cu:172-196): stage k of the composite minimax sign polynomial
static const std::vector<std::vector<double>> kSignCoefficients = {
    {0, 0.667972070856, 0, -0.223989523020, 0, 0.136121229346, 0, -0.099160550898, 0, 0.079224867308, 0, -0.067250088206, 0,
     0.059852569462, 0, -0.503955481350},
    {0, 0.955669291788, 0, -0.317870998995, 0, 0.189953989728, 0, -0.134924463410, 0, 0.104260767625, 0, -0.084798113265, 0,
     0.071534728674, 0, -0.282024623439},
    {0, 1.254717353059, 0, -0.371638622338, 0, 0.175181567419, 0, -0.085946606966, 0, 0.039326533561, 0, -0.015616729371, 0,
     0.004903749402, 0, -0.000987938705}};
static double sign_bound(size_t k) { return (k == 0) ? 1 : (k == 1) ? 1.908 : 1.332; }

TensorCT DNN::sign(const TensorCT &in, size_t k) {
    KeyScope keys(*this);
    if (k > 2) throw EvalError(B200FHE_ERR_ARG, "sign stage must be 0, 1 or 2");
    TensorCT r = like(in);
    for (size_t i = 0; i < in.num_ch; i++)
        r.ctks.push_back(ev->chebyshev_series(in.ctks[i], kSignCoefficients[k], -sign_bound(k), sign_bound(k)));
    return r;
}

static void check_boot(const DNN &d, const TensorCT &in, const Bootstrapper &bt) {
    if (&bt.ev != d.ev) throw EvalError(B200FHE_ERR_ARG, "the layers must run on the bootstrapper's evaluator");
    if (in.width == 0) return;  // plain list of ciphertexts (no image geometry): the caller vouches for the packing
    const size_t s = (in.width << in.slotstr) * (in.width << in.slotstr);
    if (s != bt.slots) throw EvalError(B200FHE_ERR_ARG, "bootstrapper was set up for a different slot count");
}

// ---- lanes
struct DNN::Lane {
    Context ctx;
    cudaStream_t st = nullptr;
    cudaEvent_t done = nullptr;
    std::unique_ptr<Bootstrapper> bt;
    std::unique_ptr<DNN> dnn;
    ~Lane() {
        dnn.reset();
        bt.reset();
        if (done) cudaEventDestroy(done);
        if (st) cudaStreamDestroy(st);
    }
};

DNN::DNN(Evaluator *ev_, Encoder *enc_, double scale_) : ev(ev_), enc(enc_), scale(scale_) {
    if (const char *v = getenv("B200FHE_DNN_BATCH")) set_batch((size_t) std::max(1, atoi(v)));
}
DNN::~DNN() = default;

void DNN::set_lanes(Bootstrapper &bt, size_t n) {
    if (&bt.ev != ev) throw EvalError(B200FHE_ERR_ARG, "the layers must run on the bootstrapper's evaluator");
    ev->cu(cudaStreamSynchronize(ev->st), "set_lanes");
    lanes_.clear();
    Context *c = ev->c;
    for (size_t l = 0; l < n; l++) {
        auto L = std::make_unique<Lane>();
        ev->ok(L->ctx.init(c->logN, c->primes.data(), c->size_QP, c->size_P, c->device, false));
        // the channel lanes already keep several key switches in flight: their digit lanes would only add launches and events
        L->ctx.ks_lanes = getenv("B200FHE_DNN_KS_LANES") ? atoi(getenv("B200FHE_DNN_KS_LANES")) : 0;
        ev->cu(cudaStreamCreateWithFlags(&L->st, cudaStreamNonBlocking), "lane stream");
        ev->cu(cudaEventCreateWithFlags(&L->done, cudaEventDisableTiming), "lane event");
        const uint32_t *a = bt.ctor_args;
        L->bt = std::make_unique<Bootstrapper>(&L->ctx, L->st, a[0], a[1], a[2], a[3], a[4], a[5]);
        for (int which = 0; which < 2; which++) {
            const auto &A = which == 0 ? bt.A_enc : bt.A_dec;
            for (size_t s = 0; s < A.size(); s++)
                for (size_t i = 0; i < A[s].size(); i++)
                    if (A[s][i].p) L->bt->set_diagonal(which, s, i, A[s][i].p, A[s][i].Ql, A[s][i].scale);
        }
        L->bt->ev.relin = ev->relin;
        L->bt->ev.have_relin = ev->have_relin;
        L->bt->ev.galois = ev->galois;
        L->dnn = std::make_unique<DNN>(&L->bt->ev, enc, scale);
        L->dnn->relin = relin;
        L->dnn->have_relin = have_relin;
        L->dnn->galois = galois;
        L->dnn->have_galois_ = have_galois_;
        lanes_.push_back(std::move(L));
    }
}

void DNN::run_batched(DNN &d, Bootstrapper &b, const TensorCT &in, const std::vector<size_t> &idx, const u64 *const *src, size_t batch,
                      const std::function<Ct(DNN &, Bootstrapper &, const Ct &)> &f, std::vector<Ct> &ys) {
    Evaluator &E = b.ev;
    const size_t N = E.c->N;
    for (size_t g0 = 0; g0 < idx.size();) {
        const Ct &x0 = in.ctks[idx[g0]];
        // the longest run of channels (up to `batch`) whose bookkeeping equals the first one's
        size_t gn = 1;
        while (g0 + gn < idx.size() && gn < batch) {
            const Ct &x = in.ctks[idx[g0 + gn]];
            if (x.Ql != x0.Ql || x.limbs != x0.limbs || x.polys != x0.polys || x.scale != x0.scale || x.noise != x0.noise) break;
            gn++;
        }
        const size_t bytes = size_t(x0.polys) * x0.limbs * N * 8;
        Ct X = E.alloc(x0.Ql, x0.polys, false, (u32) gn);
        X.scale = x0.scale;
        X.noise = x0.noise;
        for (size_t k = 0; k < gn; k++) {
            const u64 *from = src ? src[idx[g0 + k]] : in.ctks[idx[g0 + k]].data();
            E.cu(cudaMemcpyAsync(reinterpret_cast<char *>(X.data()) + k * bytes, from, bytes, cudaMemcpyDeviceToDevice, E.st), "channel batch");
        }
        Ct Y = f(d, b, X);
        if (gn == 1) {
            Y.data();  // a pending combination is evaluated here, on this evaluator's stream
            ys[idx[g0]] = Y;
        } else {
            std::vector<Ct> parts = E.unpack(Y);
            for (size_t k = 0; k < gn; k++) ys[idx[g0 + k]] = parts[k];
        }
        g0 += gn;
    }
}

TensorCT DNN::run_on_lanes(const TensorCT &in, const std::function<Ct(DNN &, Bootstrapper &, const Ct &)> &f) {
    const size_t n = in.num_ch, G = lanes_.size(), N = ev->c->N;
    std::vector<const u64 *> src(n);
    for (size_t i = 0; i < n; i++) {
        const Ct &x = in.ctks[i];
        if (x.ext() || x.batch != 1) throw EvalError(B200FHE_ERR_ARG, "lanes take plain ciphertexts");
        src[i] = x.data();  // pending combinations are evaluated on the main stream, before the fork
    }
    cudaEvent_t fork;
    ev->cu(cudaEventCreateWithFlags(&fork, cudaEventDisableTiming), "fork event");
    ev->cu(cudaEventRecord(fork, ev->st), "fork");
    for (auto &L : lanes_) ev->cu(cudaStreamWaitEvent(L->st, fork, 0), "fork wait");
    TensorCT r = like(in);
    r.ctks.resize(n);
    // One host thread per lane.  The driver's launch queue is about a thousand launches deep and a channel is ~3000
    // launches: issued from one thread, the window the GPU sees is almost always a single lane's work and nothing
    // overlaps (measured: 1270 ms for 16 channels with 1, 2, 4 or 8 lanes).  Threads interleave the submissions.
    // Each thread touches only its lane's context, evaluator and pool; this evaluator's pool is not used meanwhile.
    std::vector<Ct> ys(n);
    std::vector<std::exception_ptr> errs(G);
    std::vector<std::thread> threads;
    const int device = ev->c->device;
    for (size_t l = 0; l < G && l < n; l++)
        threads.emplace_back([&, l] {
            try {
                if (device >= 0) cudaSetDevice(device);
                Lane &L = *lanes_[l];
                std::vector<size_t> mine;
                for (size_t i = l; i < n; i += G) mine.push_back(i);
                run_batched(*L.dnn, *L.bt, in, mine, src.data(), batch_, f, ys);
            } catch (...) {
                errs[l] = std::current_exception();
            }
        });
    for (auto &t : threads) t.join();
    for (size_t l = 0; l < G; l++)
        if (errs[l]) {
            for (auto &L : lanes_) cudaStreamSynchronize(L->st);
            cudaEventDestroy(fork);
            std::rethrow_exception(errs[l]);
        }
    for (size_t i = 0; i < n; i++) {
        Lane &L = *lanes_[i % G];
        Ct &y = ys[i];
        Ct o = ev->alloc(y.Ql, y.polys, false);
        o.scale = y.scale;
        o.noise = y.noise;
        // written on the LANE stream into a block of the main pool: every earlier use of that block was queued on the
        // main stream before `fork`, which the lane has waited for
        ev->cu(cudaMemcpyAsync(o.data(), y.data(), size_t(y.polys) * y.Ql * N * 8, cudaMemcpyDeviceToDevice, L.st), "lane result");
        r.ctks[i] = o;
        y = Ct();
    }
    for (auto &L : lanes_) {
        ev->cu(cudaEventRecord(L->done, L->st), "join");
        ev->cu(cudaStreamWaitEvent(ev->st, L->done, 0), "join wait");
    }
    cudaEventDestroy(fork);
    return r;
}

// no lanes: the channels in batches on this object's evaluator
TensorCT DNN::run_here(const TensorCT &in, Bootstrapper &bt, const std::function<Ct(DNN &, Bootstrapper &, const Ct &)> &f) {
    std::vector<size_t> all(in.num_ch);
    for (size_t i = 0; i < in.num_ch; i++) {
        all[i] = i;
        if (in.ctks[i].ext() || in.ctks[i].batch != 1) throw EvalError(B200FHE_ERR_ARG, "the layers take plain ciphertexts");
    }
    TensorCT r = like(in);
    r.ctks.resize(in.num_ch);
    run_batched(*this, bt, in, all, nullptr, batch_, f, r.ctks);
    return r;
}

// the channels of `in` on this GPU: on the lanes when there are any, else on this evaluator
TensorCT DNN::run_local(const TensorCT &in, Bootstrapper &bt, const std::function<Ct(DNN &, Bootstrapper &, const Ct &)> &f) {
    if (!lanes_.empty() && in.num_ch > 1) return run_on_lanes(in, f);
    return run_here(in, bt, f);
}

// ONE image on several GPUs (dnn.h `Ranks`): this rank's run of channels locally, results all-gathered through the arena
TensorCT DNN::run_on_ranks(const TensorCT &in, Bootstrapper &bt, const std::function<Ct(DNN &, Bootstrapper &, const Ct &)> &f) {
    const size_t n = in.num_ch, G = ranks.world, N = ev->c->N;
    bool uniform = n >= G;
    for (size_t i = 1; i < n && uniform; i++) {
        const Ct &a = in.ctks[0], &x = in.ctks[i];
        uniform = x.Ql == a.Ql && x.limbs == a.limbs && x.polys == a.polys && x.scale == a.scale && x.noise == a.noise && !x.ext() && x.batch == 1;
    }
    if (G <= 1 || !uniform) return run_local(in, bt, f);   // the same decision on every rank: the inputs are the same
    auto lo = [&](size_t r) { return n * r / G; };
    const size_t c0 = lo(ranks.rank), c1 = lo(ranks.rank + 1);
    TensorCT sub = like(in);
    sub.num_ch = c1 - c0;
    sub.ctks.assign(in.ctks.begin() + c0, in.ctks.begin() + c1);
    TensorCT mine = run_local(sub, bt, f);
    const Ct &y0 = mine.ctks[0];
    if (y0.ext() || y0.batch != 1) throw EvalError(B200FHE_ERR_ARG, "rank-sharded layer: unexpected result form");
    const size_t slot = size_t(y0.polys) * y0.Ql * N * 8;
    if (n * slot > ranks.half_bytes) throw EvalError(B200FHE_ERR_ARG, "peer arena too small for this layer");
    char *base = static_cast<char *>(ranks.bases[ranks.rank]);
    const size_t off0 = size_t(ranks.parity) * ranks.half_bytes;
    ranks.parity ^= 1;
    for (size_t i = c0; i < c1; i++) {
        const Ct &y = mine.ctks[i - c0];
        if (y.Ql != y0.Ql || y.polys != y0.polys || y.scale != y0.scale || y.noise != y0.noise)
            throw EvalError(B200FHE_ERR_ARG, "rank-sharded layer: the channels' results differ in level or scale");
        ev->cu(cudaMemcpyAsync(base + off0 + i * slot, y.data(), slot, cudaMemcpyDeviceToDevice, ev->st), "own channels into the arena");
    }
    const size_t off = off0 + c0 * slot, len = (c1 - c0) * slot;
    ev->ok(peer_exchange_strided(base, ranks.bases, nullptr, ranks.flags, ranks.counter, ranks.world, ranks.rank, &off, &len, 1, 1, 0, 0,
                                 ev->st));
    TensorCT r = like(in);
    r.ctks.resize(n);
    for (size_t i = 0; i < n; i++) {
        if (i >= c0 && i < c1) r.ctks[i] = mine.ctks[i - c0];
        else r.ctks[i] = ev->wrap_copy(reinterpret_cast<const u64 *>(base + off0 + i * slot), y0.Ql, y0.polys, false, y0.scale, y0.noise);
    }
    return r;
}

// BootStrap (src/dnn.cu:266-279)
TensorCT DNN::bootstrap(const TensorCT &in, Bootstrapper &bt) {
    check_boot(*this, in, bt);
    const auto f = [](DNN &, Bootstrapper &b, const Ct &x) { return b.eval(x); };
    if (ranks.world > 1) return run_on_ranks(in, bt, f);
    return run_local(in, bt, f);
}

// ReluComposite (src/dnn.cu:198-246): relu(x) = x/2 * (1 + sign(x/10)), sign = three composed degree-15 polynomials with
// a bootstrap after each and one after the product.  The reference walks the channels stage by stage; the channels are
// independent, so one channel at a time gives the same limbs.  The sign polynomials and the product use the layers' own
// relinearisation key, the bootstraps the bootstrapper's keys.
Ct DNN::relu_composite_one(const Ct &in, Bootstrapper &bt) {
    Ct s = in;
    ev->mult_const(s, 0.1);  // into [-1, 1]
    for (size_t k = 0; k < 3; k++) {
        {
            KeyScope keys(*this);
            s = ev->chebyshev_series(s, kSignCoefficients[k], -sign_bound(k), sign_bound(k));
        }
        if (k < 2) s = bt.eval(s);
    }
    Ct r;
    {
        KeyScope keys(*this);
        ev->add_const(s, 1);
        Ct half = in;
        ev->mult_const(half, 0.5);
        r = ev->mult_auto(s, half);
    }
    return bt.eval(r);
}

TensorCT DNN::relu_composite(const TensorCT &in, Bootstrapper &bt) {
    check_boot(*this, in, bt);
    const auto f = [](DNN &d, Bootstrapper &b, const Ct &x) { return d.relu_composite_one(x, b); };
    if (ranks.world > 1) return run_on_ranks(in, bt, f);
    return run_local(in, bt, f);
}

// AvgPoolFullCon (src/dnn.cu:397-447).  Reproduced as written, including its indexing: the u == 0 pass appends one
// ciphertext PER INPUT CHANNEL (weight[0][k] * pooled[k]) and the passes u >= 1 accumulate row u into entry u, so the
// result holds t ciphertexts, entry u >= 1 = weight[0][u] * pooled[u] + sum_k weight[u][k] * pooled[k] (+ bias[u]).
TensorCT DNN::avgpool_fullcon(const TensorCT &in, const double *weight, size_t T, size_t t, const double *bias) {
    KeyScope keys(*this);
    if (t > in.num_ch || T > t) throw EvalError(B200FHE_ERR_ARG, "AvgPoolFullCon: weight shape does not fit the tensor");
    const int pow_slotstr = 1 << in.slotstr;
    const size_t log_l = (size_t) std::log2(in.width);
    std::vector<Ct> pooled;
    for (size_t k = 0; k < in.num_ch; k++) {
        pooled.push_back(in.ctks[k]);
        for (size_t i = 0; i < log_l; i++) {
            Ct tmp = ev->rotate_fused(pooled[k], pow_slotstr << i);
            ev->add_auto(tmp, pooled[k]);
            pooled[k] = tmp;
        }
        for (size_t j = 0; j < log_l; j++) {
            Ct tmp = ev->rotate_fused(pooled[k], (pow_slotstr << j) * (int) in.width);
            ev->add_auto(tmp, pooled[k]);
            pooled[k] = tmp;
        }
    }
    TensorCT r;
    for (size_t u = 0; u < T; u++)
        for (size_t k = 0; k < t; k++) {
            Ct tmp = pooled[k];
            ev->mult_const(tmp, weight[u * t + k]);
            if (u == 0) {
                r.ctks.push_back(tmp);
            } else {
                ev->add_auto(tmp, r.ctks[u]);
                r.ctks[u] = tmp;
            }
        }
    for (size_t u = 0; u < T; u++) ev->add_const(r.ctks[u], bias[u]);
    r.width = in.width;
    r.num_ch = T;
    r.slotstr = in.slotstr;
    return r;
}

// SoftMax (src/dnn.cu:339-395), as written: `1 / BoundB` is an integer division there
TensorCT DNN::softmax(const TensorCT &in, Bootstrapper &bt, size_t BoundB, size_t BoundR, size_t GumbelLambda, size_t GoldSchmidtD) {
    KeyScope keys(*this);
    (void) bt;
    TensorCT r = like(in);
    const size_t log_B = (size_t) std::log2(BoundB), log_Lambda = (size_t) std::log2(GumbelLambda);
    const std::vector<double> coeffs = chebyshev_coefficients([](double x) { return std::exp(x); }, -1, 1, 12);
    Ct sum;
    for (size_t k = 0; k < in.num_ch; k++) {
        Ct x = in.ctks[k];
        ev->mult_const(x, (double) (1 / BoundB));
        x = ev->chebyshev_series(x, coeffs, -1, 1);
        for (int i = 0; i < (int) (log_B - log_Lambda); i++) x = ev->square(x);
        r.ctks.push_back(x);
        if (k == 0) {
            sum = x;
        } else {
            ev->add_auto(sum, x);
        }
    }
    ev->mult_const(sum, (-1.0) / BoundR);
    ev->add_const(sum, 2);
    Ct temp = sum;
    ev->add_const(temp, -1);
    for (size_t j = 0; j < GoldSchmidtD; j++) {
        temp = ev->square(temp);
        Ct temp2 = temp;
        ev->add_const(temp2, 1);
        sum = ev->mult_auto(sum, temp2);
    }
    for (size_t k = 0; k < in.num_ch; k++) r.ctks[k] = ev->mult_auto(r.ctks[k], sum);
    return r;
}

}  // namespace b200fhe
