// enc_shim.cu — PhantomCKKSEncoder::decode_internal (include/ckks.h:120, src/ckks.cu:456-521) on b200fhe_decode.
// Not on the timed path, but bootstrapping_example.cu judges its result through it, and the reference's own CRT
// composition (include/uintmath.cuh:326-336: a carry chain split across separate asm statements) is miscompiled by
// nvcc 12.9 -O3 for more than two limbs: the pure reference build prints garbage there.
#include "shim_common.h"

using b200shim::ck;

extern "C" void b200ref_PhantomCKKSEncoder_decode_internal(PhantomCKKSEncoder *self, const PhantomContext *context,
                                                           const PhantomPlaintext *plain, std::vector<cuDoubleComplex> *destination,
                                                           const cudaStream_t *stream);

namespace {
std::mutex g_mu;
std::map<b200fhe_ctx *, b200fhe_encoder *> g_enc;
}  // namespace

void PhantomCKKSEncoder::decode_internal(const PhantomContext &context, const PhantomPlaintext &plain,
                                         std::vector<cuDoubleComplex> &destination, const cudaStream_t &stream) {
    auto &parms = context.get_context_data(context.get_first_index()).parms();
    const size_t N = parms.poly_modulus_degree(), size_Q = parms.coeff_modulus().size();
    if (parms.scheme() != phantom::scheme_type::ckks || plain.chain_index() == 0 || N < 4096 || N > 131072 || slots_ != N / 2 ||
        parms.special_modulus_size() == 0)
        return b200ref_PhantomCKKSEncoder_decode_internal(this, &context, &plain, &destination, &stream);
    auto &context_data = context.get_context_data(plain.chain_index());
    if (plain.scale() <= 0 || (static_cast<int>(log2(plain.scale())) >= context_data.total_coeff_modulus_bit_count()))
        throw std::invalid_argument("scale out of bounds");  // src/ckks.cu:468-472
    b200fhe_ctx *c = b200shim::ctx_for_tool(context.gpu_rns_tables(), context.get_context_data(1).gpu_rns_tool(), stream);
    b200fhe_encoder *enc;
    {
        std::lock_guard<std::mutex> g(g_mu);
        auto it = g_enc.find(c);
        if (it == g_enc.end()) {
            b200fhe_encoder *e = nullptr;
            ck(b200fhe_encoder_create(c, &e));
            it = g_enc.emplace(c, e).first;
        }
        enc = it->second;
    }
    destination.resize(slots_);
    static_assert(sizeof(cuDoubleComplex) == 2 * sizeof(double), "interleaved (re, im)");
    ck(b200fhe_decode(enc, plain.data(), size_Q - (plain.chain_index() - 1), 0, plain.scale(),
                      reinterpret_cast<double *>(destination.data()), (void *) stream));
}
