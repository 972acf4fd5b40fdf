// hostref_shim.cpp — OUR flat C interface over the reference's SEAL-derived HOST code
// (/root/reference/src/host/*.cu, compiled unmodified with g++ by oracle/Makefile into
// oracle/_ref/libphantom_hostref.so).  TEST INFRASTRUCTURE ONLY: used to pin oracle/fhe_oracle.c's
// primes, NTT tables and base-conversion constants against the reference's own generator and to
// produce tests/golden/*.npz (tests/golden/make_golden.py).  Never loaded by the product.
#include <cstdint>
#include <cstring>
#include <vector>

#include "host/modulus.h"
#include "host/ntt.h"
#include "host/rns.h"

using namespace phantom;
using namespace phantom::arith;
using namespace phantom::util;

extern "C" {

// CoeffModulus::Create (src/host/modulus.cu:80)
int href_coeff_modulus_create(size_t N, const int *bits, size_t count, uint64_t *out) {
    try {
        std::vector<int> b(bits, bits + count);
        auto m = CoeffModulus::Create(N, b);
        for (size_t i = 0; i < count; i++) out[i] = m[i].value();
        return 0;
    } catch (...) { return -1; }
}

// Modulus::const_ratio (src/host/modulus.cu:29-41)
int href_const_ratio(uint64_t q, uint64_t *ratio2) {
    try {
        Modulus m(q);
        ratio2[0] = m.const_ratio()[0];
        ratio2[1] = m.const_ratio()[1];
        return 0;
    } catch (...) { return -1; }
}

// arith::NTT (src/host/ntt.cu:11-56)
int href_ntt_tables(int logN, uint64_t q, uint64_t *root, uint64_t *tw, uint64_t *tw_shoup, uint64_t *itw,
                    uint64_t *itw_shoup, uint64_t *ninv, uint64_t *ninv_shoup) {
    try {
        Modulus m(q);
        NTT t(logN, m);
        size_t N = size_t(1) << logN;
        *root = t.get_root();
        std::memcpy(tw, t.get_from_root_powers().data(), N * 8);
        std::memcpy(tw_shoup, t.get_from_root_powers_shoup().data(), N * 8);
        std::memcpy(itw, t.get_from_inv_root_powers().data(), N * 8);
        std::memcpy(itw_shoup, t.get_from_inv_root_powers_shoup().data(), N * 8);
        *ninv = t.inv_degree_modulo();
        *ninv_shoup = t.inv_degree_modulo_shoup();
        return 0;
    } catch (...) { return -1; }
}

// RNSBase(ibase).QHatInvModq + BaseConverter(ibase, obase).QHatModp (src/host/rns.cu:438-458)
int href_base_converter(const uint64_t *ibase, size_t ni, const uint64_t *obase, size_t no, uint64_t *hat_inv,
                        uint64_t *hat_inv_shoup, uint64_t *QHatModp /* [no][ni] */) {
    try {
        std::vector<Modulus> vi, vo;
        for (size_t i = 0; i < ni; i++) vi.emplace_back(ibase[i]);
        for (size_t j = 0; j < no; j++) vo.emplace_back(obase[j]);
        RNSBase bi(vi), bo(vo);
        BaseConverter conv(bi, bo);
        std::memcpy(hat_inv, bi.QHatInvModq(), ni * 8);
        std::memcpy(hat_inv_shoup, bi.QHatInvModq_shoup(), ni * 8);
        for (size_t j = 0; j < no; j++) std::memcpy(QHatModp + j * ni, conv.QHatModp(j), ni * 8);
        return 0;
    } catch (...) { return -1; }
}
}
