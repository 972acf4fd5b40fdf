// ntt_shim.cu — replaces the reference's src/ntt/*.cu (fntt_2d, intt_2d, ntt_modup, ntt_moddown, ntt_1d,
// ntt_keyswitch_old): every launcher of include/ntt.cuh:157-226 keeps its signature and calls the negacyclic transform of
// libb200fhe.so.  The reference's launcher variants differ only in which prime a limb uses; here that is one
// b200fhe_limb_range per call.
#include "shim_common.h"

using b200shim::ck;
using b200shim::ctx_of;

namespace {

// limbs [start, start + limbs); limb t >= from uses prime base + (t - from); limbs [e0, e1) skipped.
// from == 0 means "no remap" in the C ABI, so a remap that covers limb 0 is expressed by shifting the buffer instead.
struct Call {
    uint64_t *out;
    const uint64_t *in;
    b200fhe_limb_range r;
};
Call make(uint64_t *out, const uint64_t *in, size_t n, size_t start, size_t limbs, size_t from, size_t base, size_t e0 = 0,
          size_t e1 = 0) {
    Call c{out, in, {start, limbs, from, base, e0, e1}};
    if (from <= start && base != from) {  // every limb of the range is remapped: t -> t + (base - from), identity map
        const ptrdiff_t shift = (ptrdiff_t) base - (ptrdiff_t) from;
        c.out = reinterpret_cast<uint64_t *>(reinterpret_cast<uintptr_t>(out) - shift * (ptrdiff_t) (n * 8));
        c.in = reinterpret_cast<const uint64_t *>(reinterpret_cast<uintptr_t>(in) - shift * (ptrdiff_t) (n * 8));
        c.r = {start + shift, limbs, 0, 0, e1 > e0 ? e0 + shift : 0, e1 > e0 ? e1 + shift : 0};
    } else if (from >= start + limbs || base == from) {
        c.r.special_from = c.r.special_base = 0;  // no limb of the range is remapped
    }
    return c;
}
void forward(const DNTTTable &t, const Call &c, const cudaStream_t &s) {
    ck(b200fhe_ntt_forward_ex(ctx_of(t, s), c.out, c.in, &c.r, (void *) s));
}
void inverse(const DNTTTable &t, const Call &c, const uint64_t *scale, const uint64_t *scale_shoup, const cudaStream_t &s) {
    ck(b200fhe_ntt_inverse_ex(ctx_of(t, s), c.out, c.in, &c.r, scale, scale_shoup, (void *) s));
}
void need(bool ok, const char *what) {
    if (!ok) throw std::invalid_argument(std::string("b200shim: ") + what);
}

}  // namespace

// ------------------------------------------------------------------------------------------ forward, src/ntt/fntt_2d.cu
void nwt_2d_radix8_forward_inplace(uint64_t *inout, const DNTTTable &t, size_t coeff_modulus_size, size_t start_modulus_idx,
                                   const cudaStream_t &stream) {  // fntt_2d.cu:620
    forward(t, make(inout, inout, t.n(), start_modulus_idx, coeff_modulus_size, 0, 0), stream);
}

// limb coeff_modulus_size - 1 uses prime total_modulus_size - 1 (fntt_2d.cu:226)
void nwt_2d_radix8_forward_inplace_include_temp_mod(uint64_t *inout, const DNTTTable &t, size_t coeff_modulus_size,
                                                    size_t start_modulus_idx, size_t total_modulus_size,
                                                    const cudaStream_t &stream) {  // fntt_2d.cu:658
    need(start_modulus_idx == 0 && coeff_modulus_size >= 1, "include_temp_mod expects start_modulus_idx == 0");
    forward(t, make(inout, inout, t.n(), 0, coeff_modulus_size, coeff_modulus_size - 1, total_modulus_size - 1), stream);
}

// the last size_P limbs of the range use the last size_P primes of the table (fntt_2d.cu:434-436)
void nwt_2d_radix8_forward_inplace_include_special_mod(uint64_t *inout, const DNTTTable &t, size_t coeff_modulus_size,
                                                       size_t start_modulus_idx, size_t size_QP, size_t size_P,
                                                       const cudaStream_t &stream) {  // fntt_2d.cu:695
    need(size_P <= coeff_modulus_size && size_P <= size_QP, "include_special_mod: size_P exceeds the range");
    forward(t, make(inout, inout, t.n(), start_modulus_idx, coeff_modulus_size, start_modulus_idx + coeff_modulus_size - size_P,
                    size_QP - size_P),
            stream);
}

void nwt_2d_radix8_forward_inplace_include_special_mod_exclude_range(uint64_t *inout, const DNTTTable &t,
                                                                     size_t coeff_modulus_size, size_t start_modulus_idx,
                                                                     size_t size_QP, size_t size_P, size_t excluded_range_start,
                                                                     size_t excluded_range_end,
                                                                     const cudaStream_t &stream) {  // ntt_modup.cu:610
    need(size_P <= coeff_modulus_size && size_P <= size_QP, "exclude_range: size_P exceeds the range");
    forward(t, make(inout, inout, t.n(), start_modulus_idx, coeff_modulus_size, start_modulus_idx + coeff_modulus_size - size_P,
                    size_QP - size_P, excluded_range_start, excluded_range_end),
            stream);
}

// ct = (cx - NTT(delta)) * PInv (ntt_moddown.cu:209-214); delta is clobbered like in the reference
void nwt_2d_radix8_forward_inplace_fuse_moddown(uint64_t *ct, const uint64_t *cx, const uint64_t *bigPInv_mod_q,
                                                const uint64_t *bigPInv_mod_q_shoup, uint64_t *delta, const DNTTTable &t,
                                                size_t coeff_modulus_size, size_t start_modulus_idx,
                                                const cudaStream_t &stream) {  // ntt_moddown.cu:222
    ck(b200fhe_ntt_forward_fuse_moddown(ctx_of(t, stream), ct, cx, bigPInv_mod_q, bigPInv_mod_q_shoup, delta, coeff_modulus_size,
                                        start_modulus_idx, (void *) stream));
}

// out[limb] = NTT of in[limb] under prime `modulus_index` for every limb of the range (ntt_keyswitch_old.cu:36-46; BGV/BFV
// plaintext lift: one limb per call in every reference caller)
void nwt_2d_radix8_forward_modup_fuse(uint64_t *out, const uint64_t *in, size_t modulus_index, const DNTTTable &t,
                                      size_t coeff_modulus_size, size_t start_modulus_idx, const cudaStream_t &stream) {
    for (size_t i = 0; i < coeff_modulus_size; i++) {
        const size_t limb = start_modulus_idx + i;
        forward(t, make(out + limb * t.n(), in + limb * t.n(), t.n(), 0, 1, 0, modulus_index), stream);
    }
}

// ------------------------------------------------------------------------------------------ inverse, src/ntt/intt_2d.cu
void nwt_2d_radix8_backward_inplace(uint64_t *inout, const DNTTTable &t, size_t coeff_modulus_size, size_t start_modulus_idx,
                                    const cudaStream_t &stream) {  // intt_2d.cu:724
    inverse(t, make(inout, inout, t.n(), start_modulus_idx, coeff_modulus_size, 0, 0), nullptr, nullptr, stream);
}

void nwt_2d_radix8_backward_inplace_scale(uint64_t *inout, const DNTTTable &t, size_t coeff_modulus_size,
                                          size_t start_modulus_idx, const uint64_t *scale, const uint64_t *scale_shoup,
                                          const cudaStream_t &stream) {  // intt_2d.cu:759, scale[limb] (:308)
    inverse(t, make(inout, inout, t.n(), start_modulus_idx, coeff_modulus_size, 0, 0), scale, scale_shoup, stream);
}

void nwt_2d_radix8_backward_inplace_include_special_mod(uint64_t *inout, const DNTTTable &t, size_t coeff_modulus_size,
                                                        size_t start_modulus_idx, size_t size_QP, size_t size_P,
                                                        const cudaStream_t &stream) {  // intt_2d.cu:796
    need(size_P <= coeff_modulus_size && size_P <= size_QP, "include_special_mod: size_P exceeds the range");
    inverse(t, make(inout, inout, t.n(), start_modulus_idx, coeff_modulus_size, start_modulus_idx + coeff_modulus_size - size_P,
                    size_QP - size_P),
            nullptr, nullptr, stream);
}

// The reference indexes N^-1 by the limb and the scale by the prime here (intt_2d.cu:646-656,719), which only agree when
// the temporary modulus is the table's last row at the limb's own index — the one way it is called (src/evaluate.cu:614).
void nwt_2d_radix8_backward_inplace_include_temp_mod_scale(uint64_t *inout, const DNTTTable &t, size_t coeff_modulus_size,
                                                           size_t start_modulus_idx, size_t total_modulus_size,
                                                           const uint64_t *scale, const uint64_t *scale_shoup,
                                                           const cudaStream_t &stream) {  // intt_2d.cu:836
    need(start_modulus_idx == 0 && coeff_modulus_size == total_modulus_size,
         "include_temp_mod_scale is defined for coeff_modulus_size == total_modulus_size, start 0");
    inverse(t, make(inout, inout, t.n(), 0, coeff_modulus_size, 0, 0), scale, scale_shoup, stream);
}

// ------------------------------------------------------------------------------------------ out of place, ntt_modup.cu
void nwt_2d_radix8_backward(uint64_t *out, const uint64_t *in, const DNTTTable &t, size_t coeff_modulus_size,
                            size_t start_modulus_idx, const cudaStream_t &stream) {  // ntt_modup.cu:320
    inverse(t, make(out, in, t.n(), start_modulus_idx, coeff_modulus_size, 0, 0), nullptr, nullptr, stream);
}

void nwt_2d_radix8_backward_scale(uint64_t *out, const uint64_t *in, const DNTTTable &t, size_t coeff_modulus_size,
                                  size_t start_modulus_idx, const uint64_t *scale, const uint64_t *scale_shoup,
                                  const cudaStream_t &stream) {  // ntt_modup.cu:356, scale[limb] (:310)
    inverse(t, make(out, in, t.n(), start_modulus_idx, coeff_modulus_size, 0, 0), scale, scale_shoup, stream);
}

// ------------------------------------------------------------------------------------------ 1-D, src/ntt/ntt_1d.cu
// N <= 2048, one CTA per limb.  The twiddle arrays are the caller's; the transform uses the context's own, generated from
// the same primes (see ctx_of_moduli).
static void fwd_1d(uint64_t *inout, const DModulus *modulus, size_t dim, size_t size, size_t start, const cudaStream_t &s) {
    ck(b200fhe_ntt_forward(b200shim::ctx_of_moduli(modulus, start + size, dim, s), inout, size, start, 0, 1, 0, (void *) s));
}
static void inv_1d(uint64_t *inout, const DModulus *modulus, size_t dim, size_t size, size_t start, const cudaStream_t &s) {
    ck(b200fhe_ntt_inverse(b200shim::ctx_of_moduli(modulus, start + size, dim, s), inout, size, start, 0, 1, 0, (void *) s));
}
void fnwt_1d(uint64_t *inout, const uint64_t *, const uint64_t *, const DModulus *modulus, size_t dim,
             size_t coeff_modulus_size, size_t start_modulus_idx, const cudaStream_t &stream) {  // ntt_1d.cu:146
    fwd_1d(inout, modulus, dim, coeff_modulus_size, start_modulus_idx, stream);
}
void fnwt_1d_opt(uint64_t *inout, const uint64_t *, const uint64_t *, const DModulus *modulus, size_t dim,
                 size_t coeff_modulus_size, size_t start_modulus_idx, const cudaStream_t &stream) {  // ntt_1d.cu:166
    fwd_1d(inout, modulus, dim, coeff_modulus_size, start_modulus_idx, stream);
}
// `scalar` is N^-1 in every reference caller (the y side already carries it through itwiddle[1], host/ntt.cu:53-55)
void inwt_1d(uint64_t *inout, const uint64_t *, const uint64_t *, const DModulus *modulus, const uint64_t *, const uint64_t *,
             size_t dim, size_t coeff_modulus_size, size_t start_modulus_idx, const cudaStream_t &stream) {  // ntt_1d.cu:256
    inv_1d(inout, modulus, dim, coeff_modulus_size, start_modulus_idx, stream);
}
void inwt_1d_opt(uint64_t *inout, const uint64_t *, const uint64_t *, const DModulus *modulus, const uint64_t *,
                 const uint64_t *, size_t dim, size_t coeff_modulus_size, size_t start_modulus_idx,
                 const cudaStream_t &stream) {  // ntt_1d.cu:276
    inv_1d(inout, modulus, dim, coeff_modulus_size, start_modulus_idx, stream);
}
