// hostmath.h — host-side number theory and table generation of the product library.
// Semantics follow the reference's SEAL-derived host code so every table is bit-identical:
//   primes      src/host/numth.cu:207-233 (get_primes), src/host/modulus.cu:80-111 (CoeffModulus::Create)
//   psi         src/host/numth.cu:309-332 (minimal primitive 2N-th root)
//   NTT tables  src/host/ntt.cu:11-56 (bit-reversed powers, Shoup companions, N^-1 folded into itw[1])
//   RNS consts  src/rns.cu:60-195 (DRNSTool ctor), src/host/rns.cu:438-458 (BaseConverter)
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace b200fhe {
namespace host {

typedef unsigned __int128 u128_t;

inline uint64_t mulmod(uint64_t a, uint64_t b, uint64_t q) { return (uint64_t) (((u128_t) a * b) % q); }
uint64_t powmod(uint64_t a, uint64_t e, uint64_t q);
inline uint64_t invmod(uint64_t a, uint64_t q) { return powmod(a % q, q - 2, q); }
inline uint64_t shoup(uint64_t w, uint64_t q) { return (uint64_t) (((u128_t) w << 64) / q); }
bool is_prime(uint64_t n);
void barrett_ratio(uint64_t q, uint64_t &r0, uint64_t &r1);

std::vector<uint64_t> get_primes(size_t ntt_size, int bit_size, size_t count);
std::vector<uint64_t> coeff_modulus_create(size_t N, const std::vector<int> &bit_sizes);
uint64_t minimal_primitive_root(uint64_t degree, uint64_t q);
uint32_t reverse_bits(uint32_t x, int bits);

struct PrimeTables {  // one prime, reference DNTTTable layout
    std::vector<uint64_t> tw, tw_shoup, itw, itw_shoup;
    uint64_t ninv, ninv_shoup;
};
PrimeTables ntt_tables(int logN, uint64_t q);

// constants of one level (|Ql| = size_Ql) of the hybrid key-switching RNS tool
struct LevelConsts {
    size_t size_Ql, size_P, size_Q, alpha, beta;
    std::vector<uint64_t> partQlHatInv, partQlHatInv_shoup;      // [Ql]
    std::vector<std::vector<uint64_t>> modup_mat;                 // per digit [obase][ibase]
    std::vector<uint64_t> bigP_mod_q, bigP_mod_q_shoup;           // [Ql]
    std::vector<uint64_t> bigPInv_mod_q, bigPInv_mod_q_shoup;     // [Ql]
    std::vector<uint64_t> PHatInv_mod_p, PHatInv_mod_p_shoup;     // [P]
    std::vector<uint64_t> PHat_mod_q;                             // [Ql][P]
    std::vector<uint64_t> inv_q_last_mod_q, inv_q_last_mod_q_shoup;  // [Ql-1]
    size_t digit_size(size_t d) const { return d == beta - 1 ? size_Ql - alpha * (beta - 1) : alpha; }
};
LevelConsts level_consts(const std::vector<uint64_t> &primes_QP, size_t size_P, size_t size_Ql);

}  // namespace host
}  // namespace b200fhe
