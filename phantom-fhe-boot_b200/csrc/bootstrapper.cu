// bootstrapper.cu — CKKS bootstrapping over the B200 evaluator (see bootstrapper.h).
#include "bootstrapper.h"

#include <algorithm>
#include <cmath>

#include "../../include/b200fhe.h"

namespace b200fhe {

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

static const uint32_t K_UNIFORM = 512;  // include/bootstrap.cuh:202
static const uint32_t R_UNIFORM = 6;    // include/bootstrap.cuh:203

// Chebyshev coefficients of (1/2pi) sin(2 pi K x) for the uniform-ternary secret, degree 88: numerical DATA of the
// algorithm (include/bootstrap.cuh:232-255, the published OpenFHE table) — any other digits give another ciphertext.
static const std::vector<double> &coefficients_uniform() {
    static const std::vector<double> v{
        0.15421426400235561,    -0.0037671538417132409,  0.16032011744533031,      -0.0034539657223742453,
        0.17711481926851286,    -0.0027619720033372291,  0.19949802549604084,      -0.0015928034845171929,
        0.21756948616367638,    0.00010729951647566607,  0.21600427371240055,      0.0022171399198851363,
        0.17647500259573556,    0.0042856217194480991,   0.086174491919472254,     0.0054640252312780444,
        -0.046667988130649173,  0.0047346914623733714,   -0.17712686172280406,     0.0016205080004247200,
        -0.22703114241338604,   -0.0028145845916205865,  -0.13123089730288540,     -0.0056345646688793190,
        0.078818395388692147,   -0.0037868875028868542,  0.23226434602675575,      0.0021116338645426574,
        0.13985510526186795,    0.0059365649669377071,   -0.13918475289368595,     0.0018580676740836374,
        -0.23254376365752788,   -0.0054103844866927788,  0.056840618403875359,     -0.0035227192748552472,
        0.25667909012207590,    0.0055029673963982112,   -0.073334392714092062,    0.0027810273357488265,
        -0.24912792167850559,   -0.0069524866497120566,  0.21288810409948347,      0.0017810057298691725,
        0.088760951809475269,   0.0055957188940032095,   -0.31937177676259115,     -0.0087539416335935556,
        0.34748800245527145,    0.0075378299617709235,   -0.25116537379803394,     -0.0047285674679876204,
        0.13970502851683486,    0.0023672533925155220,   -0.063649401080083698,    -0.00098993213448982727,
        0.024597838934816905,   0.00035553235917057483,  -0.0082485030307578155,   -0.00011176184313622549,
        0.0024390574829093264,  0.000031180384864488629, -0.00064373524734389861,  -7.8036008952377965e-6,
        0.00015310015145922058, 1.7670804180220134e-6,   -0.000033066844379476900, -3.6460909134279425e-7,
        6.5276969021754105e-6,  6.8957843666189918e-8,   -1.1842811187642386e-6,   -1.2015133285307312e-8,
        1.9839339947648331e-7,  1.9372045971100854e-9,   -3.0815418032523593e-8,   -2.9013806338735810e-10,
        4.4540904298173700e-9,  4.0505136697916078e-11,  -6.0104912807134771e-10,  -5.2873323696828491e-12,
        7.5943206779351725e-11, 6.4679566322060472e-13,  -9.0081200925539902e-12,  -7.4396949275292252e-14,
        1.0057423059167244e-12, 8.1701187638005194e-15,  -1.0611736208855373e-13,  -8.9597492970451533e-16,
        1.1421575296031385e-14};
    return v;
}

// SelectLayers (src/util.cu:733-766): split logSlots FFT layers over `budget` levels -> {layers per level, rows, rem}
static void select_layers(uint32_t logSlots, uint32_t budget, uint32_t &layers, uint32_t &rows, uint32_t &rem) {
    layers = (uint32_t) std::ceil(static_cast<double>(logSlots) / budget);
    rows = logSlots / layers;
    rem = logSlots % layers;
    uint32_t dim = rem != 0 ? rows + 1 : rows;
    if (dim < budget) {
        layers -= 1;
        rows = logSlots / layers;
        rem = logSlots - rows * layers;
        dim = rem != 0 ? rows + 1 : rows;
        while (dim != budget) {
            rows -= 1;
            rem = logSlots - rows * layers;
            dim = rem != 0 ? rows + 1 : rows;
        }
    }
}

FFTParams collapsed_fft_params(uint32_t slots, uint32_t levelBudget, uint32_t dim1) {
    uint32_t logSlots = (uint32_t) std::log2(slots);
    if (logSlots == 0) logSlots = 1;
    uint32_t layers, rows, rem;
    select_layers(logSlots, levelBudget, layers, rows, rem);
    FFTParams p{};
    p.levelBudget = (int32_t) levelBudget;
    p.layersCollapse = (int32_t) layers;
    p.remCollapse = (int32_t) rem;
    const uint32_t numRot = (1u << (layers + 1)) - 1, numRotRem = (1u << (rem + 1)) - 1;
    p.numRotations = (int32_t) numRot;
    p.numRotationsRem = (int32_t) numRotRem;
    if (dim1 == 0 || dim1 > numRot) p.g = 1 << (p.layersCollapse / 2 + (numRot > 7 ? 2 : 1));
    else p.g = (int32_t) dim1;
    p.b = (int32_t) ((numRot + 1) / p.g);
    if (rem != 0) {
        p.gRem = 1 << (p.remCollapse / 2 + (numRotRem > 7 ? 2 : 1));
        p.bRem = (int32_t) ((numRotRem + 1) / p.gRem);
    }
    return p;
}

Bootstrapper::Bootstrapper(Context *ctx, cudaStream_t stream, uint32_t budgetEnc, uint32_t budgetDec, uint32_t dim1Enc,
                           uint32_t dim1Dec, uint32_t numSlots, uint32_t correctionFactor)
    : ev(ctx, stream) {
    const uint32_t args[6] = {budgetEnc, budgetDec, dim1Enc, dim1Dec, numSlots, correctionFactor};
    for (int i = 0; i < 6; i++) ctor_args[i] = args[i];
    // EvalBootstrapSetup (src/bootstrap.cu:15-89)
    const uint32_t N = (uint32_t) ctx->N, M = 2 * N;
    slots = numSlots == 0 ? N / 2 : numSlots;
    if (slots == 0 || (slots & (slots - 1)) || slots > N / 2) throw EvalError(B200FHE_ERR_ARG, "slots must be a power of two <= N/2");
    if (correctionFactor == 0) {
        const double tmp = std::round(-0.265 * (2 * std::log2(M / 2) + std::log2(slots)) + 19.1);
        correction = tmp < 7 ? 7 : (tmp > 13 ? 13 : (uint32_t) static_cast<int>(tmp));
    } else {
        correction = correctionFactor;
    }
    uint32_t logSlots = (uint32_t) std::log2(slots);
    if (logSlots == 0) logSlots = 1;
    const uint32_t be = std::min(std::max(budgetEnc, 1u), logSlots), bd = std::min(std::max(budgetDec, 1u), logSlots);
    enc = collapsed_fft_params(slots, be, dim1Enc);
    dec = collapsed_fft_params(slots, bd, dim1Dec);
    if (be == 1 && bd == 1)
        throw EvalError(B200FHE_ERR_UNSUPPORTED, "level budget {1,1} (EvalLinearTransform) is disabled in the reference too");
    A_enc.resize(enc.levelBudget);
    for (int32_t i = 0; i < enc.levelBudget; i++)
        A_enc[i].resize((enc.remCollapse != 0 && i == 0) ? enc.numRotationsRem : enc.numRotations);
    A_dec.resize(dec.levelBudget);
    for (int32_t i = 0; i < dec.levelBudget; i++)
        A_dec[i].resize((dec.remCollapse != 0 && i == dec.levelBudget - 1) ? dec.numRotationsRem : dec.numRotations);
}

// one half of the arena: the inner sums of a stage (b x [2][QP][N]) | a slot of partial rows per rank | the summed rows
size_t Bootstrapper::peer_sums_bytes() const {
    const int32_t bmax = std::max(std::max(enc.b, enc.bRem), std::max(dec.b, dec.bRem));
    return size_t(bmax) * 2 * ev.c->size_QP * ev.c->N * 8;
}
size_t Bootstrapper::peer_total_bytes() const { return (2 * ev.c->size_QP + ev.c->size_Q) * ev.c->N * 8; }
size_t Bootstrapper::peer_slots_bytes() const { return peer_total_bytes() + 9 * ev.c->N * 8; }  // G ceil(R / G) <= R + G - 1 rows
size_t Bootstrapper::peer_arena_half_bytes() const { return peer_sums_bytes() + peer_slots_bytes() + peer_total_bytes(); }

size_t Bootstrapper::peer_arena_bytes() const { return 2 * peer_arena_half_bytes() + 2 * ev.c->size_Q * ev.c->N * 8 + 256; }

size_t Bootstrapper::stage_size(int which, size_t stage) const {
    const auto &A = which == 0 ? A_enc : A_dec;
    if (stage >= A.size()) throw EvalError(B200FHE_ERR_ARG, "stage out of range");
    return A[stage].size();
}

// GetDepthByDegree (src/util.cu:44-73)
static uint32_t depth_by_degree(size_t degree) {
    static const struct {
        size_t upto;
        uint32_t depth;
    } t[] = {{4, 3}, {5, 4}, {13, 5}, {27, 6}, {59, 7}, {119, 8}, {247, 9}, {495, 10}, {1007, 11}, {2031, 12}};
    for (const auto &e : t)
        if (degree <= e.upto) return e.depth;
    throw EvalError(B200FHE_ERR_ARG, "Polynomial degree is supported from 5 to 2031 inclusive");
}

// level (|Q_l|) at which EvalCoeffsToSlotsPrecompute / EvalSlotsToCoeffsPrecompute encode stage `stage`
// (src/bootstrap.cu:126-133, :239-262, :452-474)
size_t Bootstrapper::stage_level(int which, size_t stage) const {
    const size_t Q = ev.c->size_Q;
    if (which == 0) return Q - 1 - (size_t) (enc.levelBudget - 1 - (int32_t) stage);
    const size_t modDepth = depth_by_degree(coefficients_uniform().size() - 1) + R_UNIFORM;
    return Q - (modDepth + enc.levelBudget) - stage;
}

void Bootstrapper::set_diagonal(int which, size_t stage, size_t idx, const u64 *dev, size_t Ql, double scale) {
    auto &A = which == 0 ? A_enc : A_dec;
    if (stage >= A.size() || idx >= A[stage].size()) throw EvalError(B200FHE_ERR_ARG, "diagonal index out of range");
    if (Ql == 0 || Ql > ev.c->size_Q) throw EvalError(B200FHE_ERR_ARG, "diagonal level out of range");
    A[stage][idx] = Diagonal{dev, Ql, scale};
}

// FindBootstrapRotationIndices (src/bootstrap.cu:608-826)
std::vector<int32_t> Bootstrapper::rotation_indices() const {
    const uint32_t M = 2 * (uint32_t) ev.c->N;
    std::vector<int32_t> list;
    {  // CoeffsToSlots
        const FFTParams &p = enc;
        const int32_t flagRem = p.remCollapse != 0 ? 1 : 0, stop = flagRem ? 0 : -1;
        for (int32_t s = p.levelBudget - 1; s > stop; s--) {
            const int32_t sh = 1 << ((s - flagRem) * p.layersCollapse + p.remCollapse);
            for (int32_t j = 0; j < p.g; j++) list.push_back(reduce_rotation((j - int32_t((p.numRotations + 1) / 2) + 1) * sh, slots));
            for (int32_t i = 0; i < p.b; i++) list.push_back(reduce_rotation((p.g * i) * sh, M / 4));
        }
        if (flagRem) {
            for (int32_t j = 0; j < p.gRem; j++) list.push_back(reduce_rotation(j - int32_t((p.numRotationsRem + 1) / 2) + 1, slots));
            for (int32_t i = 0; i < p.bRem; i++) list.push_back(reduce_rotation(p.gRem * i, M / 4));
        }
    }
    {  // SlotsToCoeffs
        const FFTParams &p = dec;
        const int32_t flagRem = p.remCollapse != 0 ? 1 : 0;
        for (int32_t s = 0; s < p.levelBudget - flagRem; s++) {
            const int32_t sh = 1 << (s * p.layersCollapse);
            for (int32_t j = 0; j < p.g; j++) list.push_back(reduce_rotation((j - (p.numRotations + 1) / 2 + 1) * sh, M / 4));
            for (int32_t i = 0; i < p.b; i++) list.push_back(reduce_rotation((p.g * i) * sh, M / 4));
        }
        if (flagRem) {
            const int32_t s = p.levelBudget - flagRem, sh = 1 << (s * p.layersCollapse);
            for (int32_t j = 0; j < p.gRem; j++) list.push_back(reduce_rotation((j - (p.numRotationsRem + 1) / 2 + 1) * sh, M / 4));
            for (int32_t i = 0; i < p.bRem; i++) list.push_back(reduce_rotation((p.gRem * i) * sh, M / 4));
        }
    }
    if (4 * slots != M)  // sparse packing: partial sums
        for (uint32_t j = 1; j < M / (4 * slots); j <<= 1) list.push_back((int32_t) (j * slots));
    std::sort(list.begin(), list.end());
    list.erase(std::unique(list.begin(), list.end()), list.end());
    list.erase(std::remove(list.begin(), list.end(), 0), list.end());
    list.erase(std::remove(list.begin(), list.end(), int32_t(M / 4)), list.end());
    return list;
}

// One collapsed-FFT level (loop bodies of EvalCoeffsToSlots / EvalSlotsToCoeffs, src/bootstrap.cu:1262-1333).
// Same arithmetic as the reference, different schedule: the g hoisted rotations are single fused kernels, and the
// g x b plaintext products + sums run as ONE matrix-product pass that reads every rotated ciphertext once per
// four baby steps instead of once per product.
Ct Bootstrapper::bsgs_stage(const Ct &in, const std::vector<Diagonal> &A, const std::vector<int32_t> &rot_in,
                            const std::vector<int32_t> &rot_out, int32_t g, int32_t b, int32_t numRot) {
    Context *c = ev.c;
    const size_t N = c->N;
    if (in.ext() || in.polys != 2) throw EvalError(B200FHE_ERR_ARG, "linear transform needs a 2-polynomial ciphertext");
    if (in.noise != 1) throw EvalError(B200FHE_ERR_ARG, "Eval Mult in Place Failed Due to Difference in Noise Scale");
    double pt_scale = 0;
    for (int32_t idx = 0; idx < numRot; idx++) {
        const Diagonal &d = A[idx];
        if (!d.p) throw EvalError(B200FHE_ERR_ARG, "plaintext diagonal " + std::to_string(idx) + " was not set");
        if (d.Ql != in.Ql) throw EvalError(B200FHE_ERR_ARG, "Eval Mult in Place Failed Due to Difference in Chain Index");
        pt_scale = d.scale;
    }
    // limb-sharded stage (PeerShard): the limb runs of Q_l P this process computes in the two limb-wise kernels
    const size_t QlP_in = in.Ql + c->size_P;
    const bool sharded = shard.world > 1 && in.batch == 1;
    const bool peers = sharded && !shard.virtual_ranks;
    std::vector<std::pair<u32, u32>> ranges;
    if (sharded) {
        if (!fast_rotation_ext_multi_supported((u32) ev.level(in.Ql)->h.beta, c->logN))
            throw EvalError(B200FHE_ERR_UNSUPPORTED, "limb-sharded stage needs the one-launch rotation kernel (beta <= 4)");
        for (u32 r = 0; r < shard.world; r++) {
            if (peers && r != shard.rank) continue;
            const u32 lo = (u32) (QlP_in * r / shard.world), hi = (u32) (QlP_in * (r + 1) / shard.world);
            ranges.push_back({lo, hi - lo});
        }
    }
    // giant steps: hoisted rotations in the extended basis
    auto digits = ev.fast_rotation_precompute(in, sharded ? &ranges : nullptr);  // sharded: the digits' rows of the runs only
    std::vector<Ct> rot(g);
    {
        std::vector<int32_t> idx;
        for (int32_t j = 0; j < g; j++)
            if (rot_in[j] != 0) idx.push_back(rot_in[j]);
        std::vector<Ct> r = ev.fast_rotations_ext(in, idx, digits, true, sharded ? &ranges : nullptr);  // one launch for all of them
        size_t n = 0;
        for (int32_t j = 0; j < g; j++) rot[j] = rot_in[j] != 0 ? r[n++] : ev.keyswitch_ext(in);
    }
    digits.reset();

    // inner[i] = sum_j rot[j] * A[g i + j]   (index numRot = g b - 1 does not exist)
    // all b results live in ONE block [b][2][QlP][N] so the baby steps can be batched
    Ct innerAll;
    size_t arena_off = 0;
    if (peers) {  // the block is the current half of the symmetric arena: the peers' rows land in it
        const size_t need = size_t(b) * 2 * QlP_in * N * 8;
        if (need > shard.sums_bytes) throw EvalError(B200FHE_ERR_ARG, "peer arena too small for this stage");
        arena_off = size_t(shard.parity) * shard.half_bytes;
        shard.parity ^= 1;
        auto blk = std::make_shared<Pool::Block>();
        blk->p = static_cast<char *>(shard.bases[shard.rank]) + arena_off;
        blk->bytes = need;
        innerAll.buf = blk;
        innerAll.ev = &ev;
        innerAll.Ql = in.Ql;
        innerAll.limbs = QlP_in;
        innerAll.polys = 2;
        innerAll.batch = (u32) b;
    } else {
        innerAll = ev.alloc(in.Ql, 2, true, (u32) b);
    }
    innerAll.scale = rot[0].scale * pt_scale;
    innerAll.noise = rot[0].noise + 1;
    const size_t QlP = innerAll.limbs, ext_words = 2 * QlP * N;
    std::vector<const u64 *> pts(size_t(b) * g, nullptr);
    std::vector<u64 *> outs(b);
    for (int32_t i = 0; i < b; i++) {
        outs[i] = innerAll.data() + size_t(i) * ext_words;
        for (int32_t j = 0; j < g; j++)
            if (g * i + j != numRot) pts[size_t(i) * g + j] = A[g * i + j].p;
    }
    // one pass for g <= 64; wider giant steps are chunked with a carried-in partial sum (63 products + carry < 2^128)
    const int32_t W = g <= kMaxBsgs ? g : kMaxBsgs - 1;
    for (int32_t j0 = 0; j0 < g; j0 += W) {
        const int32_t gw = std::min<int32_t>(W, g - j0);
        BsgsArgs a{};
        std::vector<const u64 *> sub(size_t(b) * gw);
        for (int32_t i = 0; i < b; i++)
            for (int32_t j = 0; j < gw; j++) sub[size_t(i) * gw + j] = pts[size_t(i) * g + j0 + j];
        for (int32_t j = 0; j < gw; j++) a.ct[j] = rot[j0 + j].data();
        a.pt = sub.data();
        a.out = outs.data();
        a.g = (u32) gw;
        a.nb = (u32) b;
        a.size_QlP = (u32) QlP;
        a.accumulate = j0 != 0;
        if (sharded) {
            for (const auto &r : ranges) {
                if (r.second == 0) continue;
                a.limb0 = r.first;
                a.nlimbs = r.second;
                ev.cu(bsgs_matmul(c->T.mods, a, c->prime_map(in.Ql), N, ev.st), "bsgs_matmul");
                count_launches((b + 3) / 4);
            }
        } else {
            ev.cu(bsgs_matmul(c->T.mods, a, c->prime_map(in.Ql), N, ev.st), "bsgs_matmul");
            count_launches((b + 3) / 4);
        }
    }
    rot.clear();
    // limb-sharded stage: the baby steps are dealt to the ranks in contiguous runs [i_lo(r), i_lo(r + 1)); each rank sends its
    // rows of the inner sums to the rank that owns the step (all-to-all), runs its steps into PARTIAL sums of `first` and
    // `outer`, and the partial sums are added up below
    auto step_lo = [&](u32 r) { return (int32_t) (size_t(b) * r / shard.world); };
    const int32_t my_lo = peers ? step_lo(shard.rank) : 0, my_hi = peers ? step_lo(shard.rank + 1) : b;
    auto mine = [&](int32_t i) { return i >= my_lo && i < my_hi; };
    if (peers) {
        char *base = static_cast<char *>(shard.bases[shard.rank]);
        PeerSeg segs[kMaxRanks][kPeerSegs] = {};
        for (u32 r = 0; r < shard.world; r++) {
            const int32_t lo = step_lo(r), hi = step_lo(r + 1);
            if (r == shard.rank || hi == lo || ranges[0].second == 0) continue;
            const size_t off = arena_off + (size_t(lo) * 2 * QlP_in + ranges[0].first) * N * 8;
            segs[r][0] = PeerSeg{base + off, off, size_t(ranges[0].second) * N * 8, 2 * (u32) (hi - lo), QlP_in * N * 8, QlP_in * N * 8};
        }
        ev.ok(peer_all_to_all(shard.bases, shard.flags, shard.counter, shard.world, shard.rank, segs, ev.st));
    }

    // baby steps.  inner[0]: its c0 goes down to Q_l alone ("first"); the block then serves as the extended
    // accumulator "outer" with c0 restarted from zero (KeySwitchDownFirstElement + reset_poly_ext).
    DeviceLevel *L = ev.level(in.Ql);
    const size_t Ql = in.Ql, LN = Ql * N;
    Ct first = ev.alloc(Ql, 1, false);
    Ct outer;
    if (mine(0)) {
        ev.ok(do_moddown(c, L, first.data(), 0, innerAll.data(), 0, 1, nullptr, 0, 0, ev.st));
        outer = std::move(innerAll);  // view on element 0 of the block (the block stays alive through `outer`)
        outer.batch = 1;
        ev.cu(cudaMemsetAsync(outer.data(), 0, QlP * N * 8, ev.st), "reset_poly_ext");
    } else {  // a rank without step 0 starts its partial sums from zero
        const double sc = innerAll.scale;
        const u32 nz = innerAll.noise;
        ev.cu(cudaMemsetAsync(first.data(), 0, LN * 8, ev.st), "zero partial first");
        outer = ev.alloc(Ql, 2, true);
        outer.scale = sc;
        outer.noise = nz;
        ev.cu(cudaMemsetAsync(outer.data(), 0, 2 * QlP * N * 8, ev.st), "zero partial outer");
    }
    auto add_to_first = [&](const u64 *poly) {
        ev.cu(poly_eltwise(c->T.mods, ELT_ADD, first.data(), 0, first.data(), 0, poly, 0, (u32) Ql, 1, PrimeMap{0, 0}, N, ev.st),
              "add_two_poly_inplace");
        count_launches(1);
    };
    // steps whose rotation is non-trivial (all of them for the reference's parameter choices): ONE batched
    // moddown of their 2 polynomials each, ONE batched modup of their c1, then one fused rotation kernel per step
    // accumulating straight into `outer`
    std::vector<int32_t> hoisted, plain;
    for (int32_t i = 1; i < b; i++)
        if (mine(i)) (rot_out[i] != 0 ? hoisted : plain).push_back(i);
    const bool contiguous = !hoisted.empty() && hoisted.back() - hoisted.front() + 1 == (int32_t) hoisted.size();
    if (contiguous) {
        const u32 nh = (u32) hoisted.size();
        auto down = ev.pool.get(size_t(nh) * 2 * LN * 8);  // [nh][2][Ql][N]
        u64 *dn = reinterpret_cast<u64 *>(down->p);
        ev.ok(do_moddown(c, L, dn, LN, outs[hoisted.front()], QlP * N, 2 * nh, nullptr, 0, 0, ev.st));
        auto digits = ev.pool.get(size_t(nh) * L->h.beta * QlP * N * 8);
        u64 *dg = reinterpret_cast<u64 *>(digits->p);
        ev.ok(do_modup_batch(c, L, dg, dn + LN, 2 * LN, nh, ev.st));
        auto tmp = ev.pool.get(LN * 8);
        for (u32 e = 0; e < nh; e++) {
            const int32_t i = hoisted[e];
            const uint32_t k = find_automorphism_index_2n_complex(rot_out[i], 2 * (uint32_t) N);
            // first += automorphism(c0)   (rotate_c0 + add_two_poly_inplace)
            ev.cu(gather_limbs(reinterpret_cast<u64 *>(tmp->p), dn + size_t(e) * 2 * LN, ev.perm_table(k), (u32) Ql, 1, 0, 0, N, ev.st),
                  "rotate_c0");
            count_launches(1);
            add_to_first(reinterpret_cast<const u64 *>(tmp->p));
            FastRotArgs fa{};
            fa.up = dg + size_t(e) * L->h.beta * QlP * N;
            fa.evk = ev.evk_of(ev.galois_key(k), L);
            fa.perm = ev.perm_table(k);
            fa.c0 = nullptr;
            fa.out = outer.data();
            fa.beta = (u32) L->h.beta;
            fa.size_Ql = (u32) Ql;
            fa.size_QlP = (u32) QlP;
            fa.size_QP = (u32) c->size_QP;
            fa.accumulate = 1;
            fa.auto_index = k;
            fa.logN = c->logN;
            ev.cu(fast_rotation_ext(c->T.mods, fa, ev.p_mod_q(Ql), c->prime_map(Ql), N, ev.st), "fast_rotation_ext");
            count_launches(1);
        }
    } else {
        for (int32_t i : hoisted) {
            Ct x = ev.wrap_copy(outs[i], Ql, 2, true, outer.scale, outer.noise);
            Ct down = ev.keyswitch_down(x);
            const uint32_t k = find_automorphism_index_2n_complex(rot_out[i], 2 * (uint32_t) N);
            auto tmp = ev.pool.get(LN * 8);
            ev.cu(gather_limbs(reinterpret_cast<u64 *>(tmp->p), down.data(), ev.perm_table(k), (u32) Ql, 1, 0, 0, N, ev.st), "rotate_c0");
            count_launches(1);
            add_to_first(reinterpret_cast<const u64 *>(tmp->p));
            auto d2 = ev.fast_rotation_precompute(down);
            Ct r = ev.fast_rotation_ext(down, rot_out[i], d2, false);
            ev.add_inplace(outer, r, false);
        }
    }
    for (int32_t i : plain) {  // rotation by zero: c0 goes down alone, c1 is added in the extended basis
        Ct f = ev.alloc(Ql, 1, false);
        ev.ok(do_moddown(c, L, f.data(), 0, outs[i], 0, 1, nullptr, 0, 0, ev.st));
        add_to_first(f.data());
        const size_t off = QlP * N;
        ev.cu(poly_eltwise(c->T.mods, ELT_ADD, outer.data() + off, 0, outer.data() + off, 0, outs[i] + off, 0, (u32) QlP, 1,
                           c->prime_map(Ql), N, ev.st),
              "add_two_poly_inplace_ext");
        count_launches(1);
    }
    if (peers) {
        // partial sums -> sums: the rows of the block [outer.c0 | outer.c1 | first] (2 QlP + Ql rows) are dealt to the ranks;
        // every rank sends each peer that peer's rows of its partial sums (all-to-all into the peer's slot for this rank),
        // adds up the slots of its own rows, and the summed rows go to everybody (all-gather); modular addition is exact, so
        // the order of the summands does not change a bit
        char *base = static_cast<char *>(shard.bases[shard.rank]);
        const u32 R = (u32) (2 * QlP + Ql), G = shard.world;
        const size_t row = N * 8, slot_rows = (R + G - 1) / G;
        const size_t slots_off = arena_off + shard.sums_bytes, total_off = slots_off + shard.slots_bytes;
        auto row_lo = [&](u32 r) { return (u32) (size_t(R) * r / G); };
        const u64 *po = outer.data(), *pf = first.data();
        auto part = [&](u32 r0, u32 r1, size_t dst_off, PeerSeg (&sg)[kPeerSegs]) {  // rows [r0, r1) of the partial block
            const u32 a1 = std::min<u32>(r1, 2 * (u32) QlP);
            size_t at = dst_off;
            if (r0 < a1) {
                sg[0] = PeerSeg{reinterpret_cast<const char *>(po) + r0 * row, at, (a1 - r0) * row, 1, 0, 0};
                at += (a1 - r0) * row;
            }
            const u32 b0 = std::max<u32>(r0, 2 * (u32) QlP);
            if (b0 < r1) sg[1] = PeerSeg{reinterpret_cast<const char *>(pf) + (b0 - 2 * QlP) * row, at, (r1 - b0) * row, 1, 0, 0};
        };
        PeerSeg segs[kMaxRanks][kPeerSegs] = {};
        for (u32 r = 0; r < G; r++)
            if (r != shard.rank) part(row_lo(r), row_lo(r + 1), slots_off + shard.rank * slot_rows * row, segs[r]);
        const u32 r0 = row_lo(shard.rank), r1 = row_lo(shard.rank + 1);
        {   // the own partial rows into the own slot
            PeerSeg own[kPeerSegs] = {};
            part(r0, r1, slots_off + shard.rank * slot_rows * row, own);
            for (const PeerSeg &sg : own)
                if (sg.len) ev.cu(cudaMemcpyAsync(base + sg.dst_off, sg.src, sg.len, cudaMemcpyDeviceToDevice, ev.st), "own partial rows");
        }
        ev.ok(peer_all_to_all(shard.bases, shard.flags, shard.counter, shard.world, shard.rank, segs, ev.st));
        ev.ok(peer_row_sum(c->T.mods, reinterpret_cast<u64 *>(base + total_off + r0 * row), reinterpret_cast<const u64 *>(base + slots_off),
                           slot_rows * N, G, r0, r1 - r0, (u32) QlP, c->prime_map(Ql), N, ev.st));
        const size_t off = total_off + r0 * row, len = (r1 - r0) * row;
        ev.ok(peer_exchange_strided(base, shard.bases, nullptr, shard.flags, shard.counter, shard.world, shard.rank, &off, &len, 1, 1, 0, 0,
                                    ev.st));
        auto view = [&](size_t at, size_t bytes) {
            auto blk = std::make_shared<Pool::Block>();
            blk->p = base + at;
            blk->bytes = bytes;
            return blk;
        };
        outer.buf = view(total_off, 2 * QlP * row);  // the partial blocks are released in stream order
        first.buf = view(total_off + 2 * QlP * row, Ql * row);
    }
    Ct result = ev.keyswitch_down(outer);
    ev.cu(poly_eltwise(c->T.mods, ELT_ADD, result.data(), 0, result.data(), 0, first.data(), 0, (u32) Ql, 1, PrimeMap{0, 0},
                       N, ev.st),
          "add_two_poly_inplace");
    count_launches(1);
    return result;
}

// EvalCoeffsToSlots (src/bootstrap.cu:1157-1405)
Ct Bootstrapper::coeffs_to_slots(const Ct &in) {
    const FFTParams &p = enc;
    const uint32_t M = 2 * (uint32_t) ev.c->N;
    const int32_t flagRem = p.remCollapse != 0 ? 1 : 0, stop = flagRem ? 0 : -1;
    Ct result = in;
    for (int32_t s = p.levelBudget - 1; s > stop; s--) {
        const int32_t sh = 1 << ((s - flagRem) * p.layersCollapse + p.remCollapse);
        std::vector<int32_t> rot_in(p.g), rot_out(p.b);
        for (int32_t j = 0; j < p.g; j++) rot_in[j] = reduce_rotation((j - int32_t((p.numRotations + 1) / 2) + 1) * sh, slots);
        for (int32_t i = 0; i < p.b; i++) rot_out[i] = reduce_rotation((p.g * i) * sh, M / 4);
        if (s != p.levelBudget - 1) ev.mod_reduce(result, 1);
        result = bsgs_stage(result, A_enc[s], rot_in, rot_out, p.g, p.b, p.numRotations);
    }
    if (flagRem) {
        std::vector<int32_t> rot_in(p.gRem), rot_out(p.bRem);
        for (int32_t j = 0; j < p.gRem; j++) rot_in[j] = reduce_rotation(j - int32_t((p.numRotationsRem + 1) / 2) + 1, slots);
        for (int32_t i = 0; i < p.bRem; i++) rot_out[i] = reduce_rotation(p.gRem * i, M / 4);
        ev.mod_reduce(result, 1);
        result = bsgs_stage(result, A_enc[stop], rot_in, rot_out, p.gRem, p.bRem, p.numRotationsRem);
    }
    return result;
}

// EvalSlotsToCoeffs (src/bootstrap.cu:1407-1655)
Ct Bootstrapper::slots_to_coeffs(const Ct &in) {
    const FFTParams &p = dec;
    const uint32_t M = 2 * (uint32_t) ev.c->N;
    const int32_t flagRem = p.remCollapse != 0 ? 1 : 0;
    Ct result = in;
    for (int32_t s = 0; s < p.levelBudget - flagRem; s++) {
        const int32_t sh = 1 << (s * p.layersCollapse);
        std::vector<int32_t> rot_in(p.g), rot_out(p.b);
        for (int32_t j = 0; j < p.g; j++) rot_in[j] = reduce_rotation((j - int32_t((p.numRotations + 1) / 2) + 1) * sh, M / 4);
        for (int32_t i = 0; i < p.b; i++) rot_out[i] = reduce_rotation((p.g * i) * sh, M / 4);
        if (s != 0) ev.mod_reduce(result, 1);
        result = bsgs_stage(result, A_dec[s], rot_in, rot_out, p.g, p.b, p.numRotations);
    }
    if (flagRem) {
        const int32_t s = p.levelBudget - flagRem, sh = 1 << (s * p.layersCollapse);
        std::vector<int32_t> rot_in(p.gRem), rot_out(p.bRem);
        for (int32_t j = 0; j < p.gRem; j++) rot_in[j] = reduce_rotation((j - int32_t((p.numRotationsRem + 1) / 2) + 1) * sh, M / 4);
        for (int32_t i = 0; i < p.bRem; i++) rot_out[i] = reduce_rotation((p.gRem * i) * sh, M / 4);
        ev.mod_reduce(result, 1);
        result = bsgs_stage(result, A_dec[s], rot_in, rot_out, p.gRem, p.bRem, p.numRotationsRem);
    }
    return result;
}

// ApplyDoubleAngleIterations (src/bootstrap.cu:1657-1668)
void Bootstrapper::double_angle(Ct &x, uint32_t numIt) {
    const int32_t r = (int32_t) numIt;
    for (int32_t j = 1; j < r + 1; j++) {
        x = ev.square(x);
        Ct y = x;
        ev.add_auto(x, y);
        const double scalar = -1.0 / std::pow((2.0 * M_PI), std::pow(2.0, j - r));
        ev.add_const(x, scalar);
    }
}

// AdjustCiphertext (src/bootstrap.cu:1131-1155)
void Bootstrapper::adjust_ciphertext(Ct &x, double corr) {
    const double targetSF = ev.sf[0], sourceSF = x.scale;
    const double modToDrop = static_cast<double>(ev.c->primes[x.Ql - 1]);
    const double adjustmentFactor = (targetSF / sourceSF) * (modToDrop / sourceSF) * std::pow(2, -corr);
    ev.mult_const(x, adjustmentFactor);
    ev.mod_reduce(x, 1);
    x.scale = targetSF;
}

// EvalBootstrap with numIterations > 1 ("Meta-BTS", src/bootstrap.cu:858-910): bootstrap, measure the bootstrapping error
// at 2^precision times the message against the input, bootstrap the error and subtract it.  Same call sequence as the
// reference, including what it bootstraps first (the input itself, not the scaled-up copy) and its early return.
Ct Bootstrapper::eval(const Ct &in, uint32_t numIterations, uint32_t precision) {
    if (numIterations <= 1) return eval(in);
    if (precision >= 32) throw EvalError(B200FHE_ERR_ARG, "precision must be below 32 (the reference shifts a 32-bit one)");
    const size_t initSizeQ = in.Ql;
    const uint64_t powerOfTwoModulus = uint64_t(1) << precision;
    Ct ctScaledUp = in;
    ev.mult_by_integer(ctScaledUp, powerOfTwoModulus);
    Ct ctInitialBootstrap = eval(in, numIterations - 1, precision);
    ev.mod_reduce(ctInitialBootstrap, 1);
    ev.mult_by_integer(ctInitialBootstrap, powerOfTwoModulus);
    Ct ctBootstrappedScaledDown = ctInitialBootstrap;
    const size_t bootstrappingSizeQ = ctBootstrappedScaledDown.Ql;
    if (bootstrappingSizeQ <= initSizeQ) return in;
    ev.mod_switch_level(ctBootstrappedScaledDown, bootstrappingSizeQ - initSizeQ);
    Ct ctBootstrappingError = ctBootstrappedScaledDown;
    ev.sub_auto(ctBootstrappingError, ctScaledUp);
    Ct ctBootstrappedError = eval(ctBootstrappingError);
    ev.mod_reduce(ctBootstrappedError, 1);
    Ct finalCiphertext = ctInitialBootstrap;
    ev.sub_auto(finalCiphertext, ctBootstrappedError);
    ev.mult_const(finalCiphertext, static_cast<double>(1) / powerOfTwoModulus);
    return finalCiphertext;
}

// A batch of ciphertexts (Ct::batch > 1: independent ciphertexts with identical bookkeeping, e.g. the channels of a layer)
// goes through EvalBootstrap as one: every element-wise step, key switch, rescale and polynomial evaluation is one launch
// set over all of them (grids several times larger, where the few-limb launches of one ciphertext are latency-bound); only
// the linear transforms, whose kernels take one ciphertext, run per element.  Each element's limbs are the ones it would get alone.
template <class F>
static Ct each_of_batch(Evaluator &ev, const Ct &x, F f) {
    if (x.batch == 1) return f(x);
    std::vector<Ct> parts = ev.unpack(x);
    for (auto &p : parts) p = f(p);
    return ev.pack(parts);
}

// EvalBootstrap, numIterations == 1 (src/bootstrap.cu:843-1128)
Ct Bootstrapper::eval(const Ct &in) {
    Context *c = ev.c;
    const uint32_t N = (uint32_t) c->N, M = 2 * N;
    const size_t initSizeQ = in.Ql;
    const double qDouble = static_cast<double>(c->primes[0]);
    const double powP = std::pow(2, 59);
    const int32_t deg = (int32_t) std::round(std::log2(qDouble / powP));
    if (deg > static_cast<int32_t>(correction))
        throw EvalError(B200FHE_ERR_ARG, "Degree [" + std::to_string(deg) + "] must be less than or equal to the correction factor [" +
                                             std::to_string(correction) + "].");
    const uint32_t corr = correction - deg;
    const double post = std::pow(2, static_cast<double>(deg));
    const double pre = 1. / post;
    const uint64_t scalar = (uint64_t) std::llround(post);

    // raise the modulus
    Ct raised = in;
    ev.mod_reduce(raised, raised.noise - 1);
    adjust_ciphertext(raised, corr);
    raised = ev.raise_mod(raised);

    const std::vector<double> &coefficients = coefficients_uniform();
    const double k = K_UNIFORM;
    const double constantEvalMult = pre * (1.0 / (k * N));
    ev.mult_const(raised, constantEvalMult);
    const double lo = -1, hi = 1;

    Ct ctxtDec;
    if (slots == M / 4) {
        // fully packed
        ev.mod_reduce(raised, 1);
        Ct ctxtEnc = each_of_batch(ev, raised, [&](const Ct &p) { return coeffs_to_slots(p); });
        raised = Ct();
        Ct conj = ev.conj_fused(ctxtEnc);
        Ct ctxtEncI = ctxtEnc;
        ev.sub_auto(ctxtEncI, conj);
        ev.add_auto(ctxtEnc, conj);
        conj = Ct();
        ev.mult_by_monomial(ctxtEncI, 3 * M / 4);
        if (ctxtEnc.noise == 2) {
            ev.mod_reduce(ctxtEnc, 1);
            ev.mod_reduce(ctxtEncI, 1);
        }
        const bool peers = shard.world > 1 && !shard.virtual_ranks && ctxtEnc.batch == 1;
        if (peers && (shard.rank ^ 1u) < shard.world) {
            // limb-sharded bootstrap: this rank evaluates ONE half (even ranks the real one, odd ranks the imaginary one) and
            // swaps results with its partner r ^ 1; the halves' bookkeeping is identical step by step, so the half that
            // arrives carries the level, scale and NoiseScaleDeg of the half computed here
            const bool imag = shard.rank & 1u;
            Ct mine = imag ? ctxtEncI : ctxtEnc;
            ctxtEnc = Ct();
            ctxtEncI = Ct();
            mine = ev.chebyshev_series(mine, coefficients, lo, hi);
            ev.mod_reduce(mine, 1);
            double_angle(mine, R_UNIFORM);
            PeerSeg segs[kMaxRanks][kPeerSegs] = {};
            const size_t bytes = size_t(2) * mine.limbs * N * 8;
            segs[shard.rank ^ 1u][0] = PeerSeg{reinterpret_cast<const char *>(mine.data()), shard.evalmod_off, bytes, 1, 0, 0};
            ev.ok(peer_all_to_all(shard.bases, shard.flags, shard.counter, shard.world, shard.rank, segs, ev.st));
            count_launches(1);
            Ct other = ev.wrap_copy(reinterpret_cast<const u64 *>(static_cast<const char *>(shard.bases[shard.rank]) + shard.evalmod_off),
                                    mine.Ql, 2, false, mine.scale, mine.noise);
            ctxtEnc = imag ? other : mine;
            ctxtEncI = imag ? mine : other;
        } else {
            // approximate modular reduction of the real and the imaginary half as ONE batch-2 ciphertext: their
            // bookkeeping is identical step by step, so every launch of EvalMod covers both (half the launches,
            // twice the grid)
            const u32 B = ctxtEnc.batch;
            std::vector<Ct> hs;
            if (B == 1) {
                hs = {ctxtEnc, ctxtEncI};
            } else {
                std::vector<Ct> re = ev.unpack(ctxtEnc), im = ev.unpack(ctxtEncI);
                hs = re;
                hs.insert(hs.end(), im.begin(), im.end());
            }
            Ct both = ev.pack(hs);
            hs.clear();
            ctxtEnc = Ct();
            ctxtEncI = Ct();
            both = ev.chebyshev_series(both, coefficients, lo, hi);
            ev.mod_reduce(both, 1);
            double_angle(both, R_UNIFORM);
            std::vector<Ct> halves = ev.unpack(both);
            if (B == 1) {
                ctxtEnc = halves[0];
                ctxtEncI = halves[1];
            } else {
                ctxtEnc = ev.pack(std::vector<Ct>(halves.begin(), halves.begin() + B));
                ctxtEncI = ev.pack(std::vector<Ct>(halves.begin() + B, halves.end()));
            }
            if (peers) {  // the rank without partner still takes part in the exchange (its barrier)
                PeerSeg segs[kMaxRanks][kPeerSegs] = {};
                ev.ok(peer_all_to_all(shard.bases, shard.flags, shard.counter, shard.world, shard.rank, segs, ev.st));
                count_launches(1);
            }
        }
        ev.mult_by_monomial(ctxtEncI, M / 4);
        ev.add_auto(ctxtEnc, ctxtEncI);
        ctxtEncI = Ct();
        ev.mult_by_integer(ctxtEnc, scalar);
        ev.mod_reduce(ctxtEnc, 1);
        ctxtDec = each_of_batch(ev, ctxtEnc, [&](const Ct &p) { return slots_to_coeffs(p); });
    } else {
        // sparsely packed: partial sums first
        for (uint32_t j = 1; j < N / (2 * slots); j <<= 1) {
            Ct temp = ev.rotate_fused(raised, (int32_t) (j * slots));
            ev.add_auto(raised, temp);
        }
        ev.mod_reduce(raised, 1);
        Ct ctxtEnc = each_of_batch(ev, raised, [&](const Ct &p) { return coeffs_to_slots(p); });
        raised = Ct();
        Ct conj = ev.conj_fused(ctxtEnc);
        ev.add_auto(ctxtEnc, conj);
        conj = Ct();
        if (ctxtEnc.noise == 2) ev.mod_reduce(ctxtEnc, 1);
        ctxtEnc = ev.chebyshev_series(ctxtEnc, coefficients, lo, hi);
        ev.mod_reduce(ctxtEnc, 1);
        double_angle(ctxtEnc, R_UNIFORM);
        ev.mult_by_integer(ctxtEnc, scalar);
        ev.mod_reduce(ctxtEnc, 1);
        ctxtDec = each_of_batch(ev, ctxtEnc, [&](const Ct &p) { return slots_to_coeffs(p); });
        Ct rotated = ev.rotate_fused(ctxtDec, (int32_t) slots);
        ev.add_auto(ctxtDec, rotated);
    }
    const uint64_t corFactor = (uint64_t) 1 << std::llround(corr);
    ev.mult_by_integer(ctxtDec, corFactor);
    if (ctxtDec.Ql <= initSizeQ) return in;  // no level gained: the reference hands the input back
    return ctxtDec;
}

}  // namespace b200fhe
