// rns_shim.cu — the CKKS branches of DRNSTool::modup / moddown / moddown_from_NTT (src/rns_bconv.cu:530,711,791),
// DRNSTool::divide_and_round_q_last_ntt (src/rns.cu:1160), key_switch_inner_prod and keyswitch_inplace
// (src/eval_key_switch.cu:87,112) on libb200fhe.so.  The reference's own definitions stay in the build under the names
// below (shim/Makefile renames them in the compiled objects, the sources are untouched) and serve the BFV / BGV branches,
// which are outside the CKKS hot path.
#include "shim_common.h"

using namespace phantom;
using b200shim::ck;
using b200shim::ctx_for_tool;

// the reference's compiled definitions, renamed (Itanium ABI: `this` first, references as pointers)
extern "C" {
void b200ref_DRNSTool_modup(const DRNSTool *self, uint64_t *dst, const uint64_t *cks, const DNTTTable *t, const scheme_type *sch,
                            const cudaStream_t *s);
void b200ref_DRNSTool_moddown(const DRNSTool *self, uint64_t *ct_i, uint64_t *cx_i, const DNTTTable *t, const scheme_type *sch,
                              const cudaStream_t *s);
void b200ref_DRNSTool_moddown_from_NTT(const DRNSTool *self, uint64_t *ct_i, uint64_t *cx_i, const DNTTTable *t,
                                       const scheme_type *sch, const cudaStream_t *s);
void b200ref_DRNSTool_divide_and_round_q_last_ntt(const DRNSTool *self, uint64_t *src, size_t cipher_size, const DNTTTable *t,
                                                  uint64_t *dst, const cudaStream_t *s);
void b200ref_key_switch_inner_prod(uint64_t *p_cx, const uint64_t *p_t_mod_up, const uint64_t *const *rlk, const DRNSTool *tool,
                                   const DModulus *modulus_QP, size_t reduction_threshold, const cudaStream_t *s);
void b200ref_keyswitch_inplace(const PhantomContext *context, PhantomCiphertext *encrypted, uint64_t *c2,
                               const PhantomRelinKey *relin_keys, bool is_relin, const cudaStream_t *s);
}

namespace {
// the composite kernels of libb200fhe.so cover N = 2^12 .. 2^17; smaller rings keep the reference's op-by-op sequence,
// whose transforms still come from ntt_shim.cu
bool composite_ok(const DRNSTool &tool) { return b200shim::level() >= 2 && tool.n() >= 4096 && tool.n() <= 131072 && tool.size_P() >= 1; }
}  // namespace

namespace phantom {

void DRNSTool::modup(uint64_t *dst, const uint64_t *cks, const DNTTTable &ntt_tables, const scheme_type &scheme,
                     const cudaStream_t &stream) const {
    if (scheme != scheme_type::ckks || !composite_ok(*this)) return b200ref_DRNSTool_modup(this, dst, cks, &ntt_tables, &scheme, &stream);
    ck(b200fhe_modup(ctx_for_tool(ntt_tables, *this, stream), base_Ql_.size(), dst, cks, (void *) stream));
}

// CKKS: ct = (cx - NTT(bconv(INTT(cx[P])))) * P^-1, the same value moddown_from_NTT produces (the reference only
// fuses the last step differently, src/rns_bconv.cu:741-752 against :836-838)
void DRNSTool::moddown(uint64_t *ct_i, uint64_t *cx_i, const DNTTTable &ntt_tables, const scheme_type &scheme,
                       const cudaStream_t &stream) const {
    if (scheme != scheme_type::ckks || !composite_ok(*this)) return b200ref_DRNSTool_moddown(this, ct_i, cx_i, &ntt_tables, &scheme, &stream);
    ck(b200fhe_moddown_from_ntt(ctx_for_tool(ntt_tables, *this, stream), base_Ql_.size(), ct_i, cx_i, (void *) stream));
}

void DRNSTool::moddown_from_NTT(uint64_t *ct_i, uint64_t *cx_i, const DNTTTable &ntt_tables, const scheme_type &scheme,
                                const cudaStream_t &stream) const {
    if (scheme != scheme_type::ckks || !composite_ok(*this))
        return b200ref_DRNSTool_moddown_from_NTT(this, ct_i, cx_i, &ntt_tables, &scheme, &stream);
    ck(b200fhe_moddown_from_ntt(ctx_for_tool(ntt_tables, *this, stream), base_Ql_.size(), ct_i, cx_i, (void *) stream));  // may alias
}

// CKKS only in the reference too (rescale_to_next -> mod_switch_scale_to_next, src/evaluate.cu:1591-1628).  The reference
// leaves the last limb of src in coefficient form behind; src here is left untouched (its caller passes a scratch copy).
void DRNSTool::divide_and_round_q_last_ntt(uint64_t *src, size_t cipher_size, const DNTTTable &rns_tables, uint64_t *dst,
                                           const cudaStream_t &stream) const {
    if (!composite_ok(*this)) return b200ref_DRNSTool_divide_and_round_q_last_ntt(this, src, cipher_size, &rns_tables, dst, &stream);
    ck(b200fhe_rescale(ctx_for_tool(rns_tables, *this, stream), base_Ql_.size(), dst, src, cipher_size, (void *) stream));
}

// rlk: the DEVICE array of digit pointers (PhantomRelinKey::public_keys_ptr()), consumed as it is
void key_switch_inner_prod(uint64_t *p_cx, const uint64_t *p_t_mod_up, const uint64_t *const *rlk, const DRNSTool &rns_tool,
                           const DModulus *modulus_QP, size_t reduction_threshold, const cudaStream_t &stream) {
    // The tool does not name its DNTTTable, but modulus_QP is the table's own modulus array
    // (context.gpu_rns_tables().modulus() at every call site), and modup has registered that table before any inner product.
    b200fhe_ctx *c = composite_ok(rns_tool) ? b200shim::ctx_by_moduli(modulus_QP) : nullptr;
    if (!c) return b200ref_key_switch_inner_prod(p_cx, p_t_mod_up, rlk, &rns_tool, modulus_QP, reduction_threshold, &stream);
    // 128-bit lazy accumulation: no intermediate reductions, so reduction_threshold is not needed
    ck(b200fhe_inner_prod_devkeys(c, rns_tool.base_Ql().size(), p_cx, p_t_mod_up, rlk, (void *) stream));
}

void keyswitch_inplace(const PhantomContext &context, PhantomCiphertext &encrypted, uint64_t *c2, const PhantomRelinKey &relin_keys,
                       bool is_relin, const cudaStream_t &stream) {
    auto &key_parms = context.get_context_data(0).parms();
    if (key_parms.scheme() != scheme_type::ckks)
        return b200ref_keyswitch_inplace(&context, &encrypted, c2, &relin_keys, is_relin, &stream);
    const uint32_t levelsDropped = encrypted.chain_index() - 1;  // src/eval_key_switch.cu:149
    auto &rns_tool = context.get_context_data(1 + levelsDropped).gpu_rns_tool();
    if (!composite_ok(rns_tool)) return b200ref_keyswitch_inplace(&context, &encrypted, c2, &relin_keys, is_relin, &stream);
    // modup -> inner product -> 2 x moddown -> add to (c0, c1): one call, no temporaries allocated
    ck(b200fhe_keyswitch_devkeys(ctx_for_tool(context.gpu_rns_tables(), rns_tool, stream), rns_tool.base_Ql().size(),
                                 encrypted.data(), c2, relin_keys.public_keys_ptr(), (void *) stream));
}

}  // namespace phantom
