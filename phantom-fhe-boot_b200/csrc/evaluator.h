// evaluator.h — host side of the CKKS evaluator the bootstrap is built from (FLEXIBLEAUTO bookkeeping of the
// reference: src/evaluate.cu:2154-4036, include/evaluate.cuh:262-470) over the sm_100a kernels.
//
// Design (not the reference's): ciphertexts are views on blocks of a stream-ordered pool — no cudaMallocAsync,
// no deep copy per operation (the reference copies a whole ciphertext in every helper); copies happen only when a
// shared block is about to be written (copy-on-write).  Every arithmetic step is canonical mod q, so results are
// bit-identical to the reference whatever the fusion.  All work of one Evaluator goes to ONE stream.
#pragma once
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "context.h"
#include "evalk.cuh"

namespace b200fhe {

struct EvalError : std::runtime_error {
    int code;
    EvalError(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

// ---- block pool: cudaMalloc only when no free block of the size class exists; blocks return on last release.
// Single stream => reuse in stream order is safe without events.
class Pool {
  public:
    struct Block {
        void *p = nullptr;
        size_t bytes = 0;
    };
    ~Pool();
    std::shared_ptr<Block> get(size_t bytes);
    size_t bytes_reserved() const { return reserved_; }
    void trim();  // cudaFree every idle block
    // exchange the whole state (idle blocks, reservation) with another pool; blocks in flight return to the state they
    // came from.  A captured CUDA graph runs on a pool of its own: its baked addresses are never handed to anybody else.
    void swap_state(Pool &o) {
        std::swap(sh_, o.sh_);
        std::swap(reserved_, o.reserved_);
    }

  private:
    struct Shared {
        std::multimap<size_t, void *> idle;
        bool alive = true;
    };
    std::shared_ptr<Shared> sh_ = std::make_shared<Shared>();
    size_t reserved_ = 0;
};

class Evaluator;

// A pending affine combination  sum_i s_i (.) block_i  (+ kappa on polynomial 0): what a chain of EvalMultConst /
// EvalAdd / EvalSub / EvalAddConst / MultByInteger / ModSwitchLevel leaves behind until somebody needs the limbs.
// Everything is exact arithmetic mod q, so evaluating the chain in one pass gives the reference's bits.
struct Lazy {
    struct Term {
        std::shared_ptr<Pool::Block> buf;
        size_t limbs;              // limb count of the block's layout (polynomial stride = limbs * N)
        std::vector<uint64_t> s;   // per-limb multiplier (at least Ql entries of the owner)
    };
    std::vector<Term> terms;
    std::vector<uint64_t> kappa;   // empty = none
    mutable std::shared_ptr<Pool::Block> done;  // set by the first owner that materialises; shared by all copies
};

// A ciphertext (or extended ciphertext / single polynomial) in NTT form: [polys][limbs][N]
struct Ct {
    mutable std::shared_ptr<Pool::Block> buf;
    mutable std::shared_ptr<const Lazy> lazy;  // non-null: value = the combination, buf unused (never for ext())
    Evaluator *ev = nullptr;
    size_t Ql = 0;      // ciphertext primes q_0..q_{Ql-1}; reference chain_index = |Q| - Ql + 1
    size_t limbs = 0;   // Ql, or Ql + |P| for the extended basis (KeySwitchExt / EvalFastRotationExt results)
    u32 polys = 0;
    u32 batch = 1;      // independent ciphertexts with identical bookkeeping, laid out [batch][polys][limbs][N]
    double scale = 1.0;
    u32 noise = 1;      // NoiseScaleDeg
    u64 *data() const;  // evaluates a pending combination first
    bool ext() const { return limbs != Ql; }
    bool empty() const { return !buf && !lazy; }
    u32 npoly() const { return batch * polys; }
};

struct LongDiv {
    std::vector<double> q, r;
};
// src/util.cu:88-300 (Degree, LongDivisionChebyshev, ComputeDegreesPS), same floating-point operation order
uint32_t poly_degree(const std::vector<double> &c);
LongDiv long_division_chebyshev(const std::vector<double> &f, const std::vector<double> &g);
void compute_degrees_ps(uint32_t n, uint32_t &k, uint32_t &m);
uint32_t reduce_rotation(int32_t index, uint32_t slots);                 // src/util.cu:355
uint32_t find_automorphism_index_2n_complex(int32_t i, uint32_t m);     // src/util.cu:908

class Evaluator {
  public:
    Evaluator(Context *ctx, cudaStream_t stream);
    Context *c;
    cudaStream_t st;
    Pool pool;
    std::vector<double> sf, sf_big;  // m_scalingFactorsReal / m_scalingFactorsRealBig (include/ciphertext.h:320-366)

    // keys (device pointers owned by the caller)
    bool have_relin = false;
    std::vector<const uint64_t *> relin;                       // dnum digit pointers
    std::map<uint32_t, std::vector<const uint64_t *>> galois;  // automorphism index -> digit pointers

    size_t chain_index(const Ct &x) const { return c->size_Q - x.Ql + 1; }
    double level_sf(const Ct &x) const { return sf[c->size_Q - x.Ql]; }  // m_scalingFactorsReal[chain_index - 1]

    Ct alloc(size_t Ql, u32 polys, bool ext, u32 batch = 1);
    Ct pack(const std::vector<Ct> &xs);            // same level / scale / NoiseScaleDeg -> one batched ciphertext
    std::vector<Ct> unpack(const Ct &x);
    Ct wrap_copy(const u64 *dev, size_t Ql, u32 polys, bool ext, double scale, u32 noise);  // D2D copy in
    void own(Ct &x);  // copy-on-write: make x the only owner of its block
    // element-wise update of x: returns where to READ x from; x.data() is where to WRITE (a fresh block when the
    // old one is shared, so no copy is ever made just to be overwritten)
    const u64 *begin_update(Ct &x);
    // pending affine combinations (see Lazy): x := w (.) x, a := a +- b, x := x +- kappa on polynomial 0
    std::shared_ptr<Lazy> lazy_of(const Ct &x) const;
    void lazy_scale(Ct &x, const std::vector<uint64_t> &w);
    void lazy_add(Ct &a, const Ct &b, bool sub);
    void lazy_add_const(Ct &x, const std::vector<uint64_t> &k, bool sub);
    void materialize(const Ct &x);

    // --- level / scale management
    void mod_reduce(Ct &x, size_t levels);          // EvalModReduceInPlace
    void mod_switch_level(Ct &x, size_t levels);    // ModSwitchLevelInPlace
    std::vector<uint64_t> element_for_mult(const Ct &x, double operand) const;    // GetElementForEvalMult
    std::vector<uint64_t> element_for_add(const Ct &x, double operand) const;     // GetElementForEvalAddOrSub
    void mult_const_core(Ct &x, double scalar);     // EvalMultConstInplaceCore
    void mult_const(Ct &x, double scalar);          // EvalMultConstInplace
    void add_const(Ct &x, double operand);          // EvalAddConstInPlaceWrap
    void adjust_levels_and_depth(Ct &a, Ct &b);     // AdjustLevelsAndDepthInPlace
    void add_inplace(Ct &a, const Ct &b, bool sub);
    void add_auto(Ct &a, const Ct &b);              // EvalAddAutoInplace
    void sub_auto(Ct &a, const Ct &b);              // EvalSubAutoInplace
    Ct mult_auto(const Ct &a, const Ct &b);         // EvalMultAuto
    Ct square(const Ct &a);                         // EvalSquare
    void mult_by_integer(Ct &x, uint64_t v);        // MultByIntegerInPlace
    void mult_by_monomial(Ct &x, uint32_t power);   // MultByMonomialInPlace
    Ct raise_mod(const Ct &x);                      // RaiseMod
    Ct rotate_fused(const Ct &x, int32_t index);    // EvalRotateFused
    Ct conj_fused(const Ct &x);                     // EvalConjFused
    // EvalRotateFused of one ciphertext by R indices with the modup shared; dst [R][2][Ql][N] (device)
    void rotate_fused_many(const Ct &x, const std::vector<int32_t> &indices, u64 *dst);

    // --- SEAL-style surface (src/evaluate.cu:95-2149; evalseal.cu): scale bookkeeping by the caller, NoiseScaleDeg kept
    std::map<uint32_t, std::vector<const uint64_t *>> seal_galois;  // Galois element -> key digits (PhantomGaloisKey)
    void negate(Ct &x);                                                    // negate_inplace
    void add_plain(Ct &x, const u64 *pt, double pt_scale, bool sub);       // add_plain_inplace / sub_plain_inplace
    void multiply_plain(Ct &x, const u64 *pt, double pt_scale);            // multiply_plain_inplace (CKKS: NTT form)
    Ct add_many(const std::vector<Ct> &xs);                                // add_many
    Ct multiply(const Ct &a, const Ct &b);                                 // multiply_inplace: 3 polynomials
    Ct relinearize(const Ct &x3);                                          // relinearize_inplace
    Ct rescale_to_next(const Ct &x);                                       // rescale_to_next (2 or 3 polynomials)
    Ct mod_switch_to_next(const Ct &x);                                    // mod_switch_to_next: drop the last prime
    Ct apply_galois(const Ct &x, uint32_t galois_elt);                     // apply_galois_inplace
    Ct rotate(const Ct &x, int step);                                      // rotate_inplace (NAF over the keys present)
    Ct hoisting(const Ct &x, const std::vector<int> &steps);               // hoisting_inplace: sum of rotations

    // --- hoisting in the extended basis
    // ranges: only these limb runs (first, count) of Q_l P of every digit (a limb-sharded stage reads no other rows)
    std::shared_ptr<Pool::Block> fast_rotation_precompute(const Ct &x, const std::vector<std::pair<u32, u32>> *ranges = nullptr);  // [beta][QlP][N]
    Ct fast_rotation_ext(const Ct &x, int32_t index, const std::shared_ptr<Pool::Block> &digits, bool add_first);
    // the same for a list of (non-zero) rotation indices: one launch per 64 rotations, digits read once
    // ranges: limb runs (first, count) of Q_l P to compute, one launch each (a limb-sharded stage computes the rank's own
    // rows only; the other rows of the results are left as allocated); nullptr = all limbs
    std::vector<Ct> fast_rotations_ext(const Ct &x, const std::vector<int32_t> &indices,
                                       const std::shared_ptr<Pool::Block> &digits, bool add_first,
                                       const std::vector<std::pair<u32, u32>> *ranges = nullptr);
    Ct keyswitch_ext(const Ct &x);                  // KeySwitchExt
    Ct keyswitch_down(Ct &x);                       // KeySwitchDown            (x's P limbs are clobbered)
    Ct keyswitch_down_first(Ct &x);                 // KeySwitchDownFirstElement (1 polynomial)
    void mult_ext(Ct &x, const u64 *pt);            // EvalMultExtInPlace (plaintext [QlP][N], same level)
    void add_ext(Ct &a, const Ct &b);               // EvalAddExtInPlace

    // --- polynomial evaluation
    typedef std::vector<std::shared_ptr<Ct>> CtVec;  // shared handles: the reference mutates T[i] through aliases
    Ct linear_wsum(CtVec &cts, const std::vector<double> &w);   // EvalLinearWSumMutable
    Ct chebyshev_series(const Ct &x, const std::vector<double> &coeffs, double a, double b);  // EvalChebyshevSeries
    Ct chebyshev_ps(const Ct &x, const std::vector<double> &coeffs, double a, double b);
    Ct chebyshev_linear(const Ct &x, const std::vector<double> &coeffs, double a, double b);
    Ct inner_chebyshev_ps(const std::vector<double> &coeffs, uint32_t k, uint32_t m, CtVec &T, CtVec &T2);
    Ct ps_combine(const std::vector<double> &f2, uint32_t k, uint32_t m, CtVec &T, CtVec &T2, bool top);

    // helpers
    DeviceLevel *level(size_t Ql);
    EvkPtrs evk_of(const std::vector<const uint64_t *> &digits, DeviceLevel *L) const;
    const std::vector<const uint64_t *> &galois_key(uint32_t auto_index) const;
    const u32 *perm_table(uint32_t auto_index);  // cached device permutation of an automorphism
    LimbScalars scalars(const std::vector<uint64_t> &w, size_t Ql) const;  // + Shoup companions (prime(t) of Q_l P)
    const LimbScalars &p_mod_q(size_t Ql);
    void cu(cudaError_t e, const char *what) const;
    void ok(int rc) const;

  private:
    std::map<uint32_t, std::unique_ptr<DevBuf>> perms_;
    std::map<size_t, LimbScalars> pmodq_;
    std::map<uint32_t, std::unique_ptr<DevBuf>> monomials_;  // power -> NTT(+-x^power) over all |Q| limbs
};

// the evaluator behind an opaque C-ABI handle (capi_eval.cu)
Evaluator *eval_of(void *eval_handle);

}  // namespace b200fhe
