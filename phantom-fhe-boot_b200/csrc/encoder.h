// encoder.h — CKKS encoder / decoder on the sm_100a NTT.
// Replaces PhantomCKKSEncoder::encode / encode_ext / encode_sparse / encode_sparse_ext / decode
// (src/ckks.cu:45-669, include/ckks.h:175-345), the special FFT (src/fft.cu) and DRNSBase::decompose_array /
// compose_array (src/rns_base.cu:49-260).
//
// The floating-point part is written with the reference's butterfly expressions (cuCmul / cuCadd / cuCsub, the same
// root table construction, the same final scaling) so that every double — and therefore every residue after rounding —
// is the one the reference produces; the stage schedule (one launch per stage here) does not change the data flow.
#pragma once
#include <cuComplex.h>

#include <map>
#include <memory>
#include <vector>

#include "context.h"
#include "evalk.cuh"

namespace b200fhe {

struct FftTables {  // tables of one special FFT size (n slots, m = 4n roots)
    uint32_t n = 0, logn = 0, m = 0;
    DevBuf roots;  // [m] cuDoubleComplex: powers of the primitive m-th root (util::ComplexRoots, fft.cu:15-58)
    DevBuf group;  // [n/2] u32: 5^i mod m
};

class Encoder {
  public:
    explicit Encoder(Context *ctx);
    Context *c;
    size_t slots;  // N/2

    // values: n complex numbers on the host (n <= slots).  sparse == false: zero-padded to `slots`
    // (encode_internal[_ext]); sparse == true: n must be a power of two, the polynomial lives on every
    // (slots/n)-th coefficient (encode_sparse_internal[_ext]).  ext: limbs of Q_l followed by the special primes.
    // out: device, [Ql (+|P|)][N], NTT form.  Returns a B200FHE_* status.
    int encode(const cuDoubleComplex *values, size_t n, size_t size_Ql, double scale, bool ext, bool sparse, u64 *out,
               cudaStream_t st);
    // encode_sparse (ext == false) of `batch` messages, values [batch][n] on the host -> out [batch][Ql][N] on the device
    // lift != nullptr: every plaintext is also multiplied limb-wise by lift->w and rescaled by its last prime before the
    // NTT (inv_q_last: device, [size_Ql - 1], q_last^-1 mod q_j); out is then [batch][size_Ql - 1][N]
    int encode_sparse_batch(const cuDoubleComplex *values, size_t n, size_t batch, size_t size_Ql, double scale, u64 *out,
                            cudaStream_t st, const LimbScalars *lift = nullptr, const u64 *inv_q_last = nullptr,
                            bool compact = false);
    // compact (n <= 1024): out is [batch][limbs][2n]; entry k of a limb is the value of ALL N / 2n NTT-form coefficients
    // [k * N/2n, (k+1) * N/2n) of the plaintext (the transform of a polynomial in X^(N/2n) repeats, ntt.cuh)
    // plain: device, [limbs][N] NTT form over q_0..q_{limbs-1} (ext: Q_l then P); out: `slots` complex numbers (host)
    int decode(const u64 *plain, size_t size_Ql, bool ext, double scale, cuDoubleComplex *out, cudaStream_t st);

  private:
    const FftTables &tables(uint32_t n, bool carried_group);
    std::map<uint32_t, std::unique_ptr<FftTables>> tabs_;
    DevBuf buf_in_, buf_tmp_, buf_res_, buf_batch_in_, buf_batch_res_, buf_batch_aux_;
};

}  // namespace b200fhe
