// shim_common.h — shared pieces of the C++ drop-in shim: the reference's own symbols (include/ntt.cuh:157-226,
// include/rns.cuh:156-163, include/evaluate.cuh:27-34, include/bootstrap.cuh:113 of alexlee838/phantom-fhe-boot),
// re-implemented over the C ABI of libb200fhe.so (include/b200fhe.h).  Compiled against the reference's headers and
// linked with the reference's remaining, unmodified translation units (shim/Makefile); reference callers such as
// test/ntt_test.cu and bootstrapping/bootstrapping_example.cu compile and link unchanged.
#pragma once
#include <complex>
#include <cstdint>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

// Inside the reference tree the shim would be a friend of the classes it re-implements members of; built out of tree it
// reads their private fields (layout is unchanged: access specifiers do not move members within one access block order
// the reference's compiler already fixed, and every class here is compiled with the same definition on both sides).
#define private public
#include "phantom.h"
#include "bootstrap.cuh"
#undef private

#include "b200fhe.h"

namespace b200shim {

// C-ABI status -> the exception type the reference throws at the same place (error convention)
inline void ck(int rc) {
    if (rc == B200FHE_OK) return;
    const std::string msg = std::string("b200fhe: ") + b200fhe_last_error();
    if (rc == B200FHE_ERR_ARG) throw std::invalid_argument(msg);
    if (rc == B200FHE_ERR_UNSUPPORTED) throw std::logic_error(msg);
    throw std::runtime_error(msg);
}

// The b200fhe context that stands for a DNTTTable: built once from the table's own device arrays
// (b200fhe_ctx_create_from_tables), one per (table, host thread) because composite calls on a context share a
// workspace and the reference's evaluator is safe from concurrent host threads on their own streams.
b200fhe_ctx *ctx_of(const DNTTTable &t, cudaStream_t s);
// the same for the 1-D entry points, which pass raw arrays: keyed by the DModulus array
b200fhe_ctx *ctx_of_moduli(const DModulus *modulus_dev, size_t count, size_t dim, cudaStream_t s);
// a DRNSTool call on `t`: fixes the number of special primes of the context (once) and checks the tool's Q_l primes
b200fhe_ctx *ctx_for_tool(const DNTTTable &t, const phantom::DRNSTool &tool, cudaStream_t s);
// the context of the table whose modulus array this is, if one was registered on this thread (else nullptr)
b200fhe_ctx *ctx_by_moduli(const DModulus *modulus_dev);
// how much of the path the shim takes over (environment variable B200SHIM_LEVEL, read once; diagnostic):
//   1 = transforms only (ntt_shim.cu), 2 = + modup / moddown / rescale / key switch (rns_shim.cu),
//   3 = + EvalBootstrap (boot_shim.cu); default 3.  The decoder replacement (enc_shim.cu) is always on.
int level();
// drop every cached context (tables about to be freed and re-created at the same addresses)
void forget_all();

}  // namespace b200shim

extern "C" void b200shim_forget_all(void);
extern "C" const char *b200shim_version(void);
