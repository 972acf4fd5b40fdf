// shim_registry.cu — DNTTTable -> b200fhe_ctx.  The reference hands every launcher a DNTTTable (device arrays only);
// the context that owns our interleaved copies of those tables is found here.
#include <chrono>
#include <cstdlib>
#include <thread>
#include <unordered_map>

#include "shim_common.h"

namespace b200shim {
namespace {

struct Key {
    const void *tw, *mod;
    uint64_t n, size;
    std::thread::id tid;
    bool operator==(const Key &o) const { return tw == o.tw && mod == o.mod && n == o.n && size == o.size && tid == o.tid; }
};
struct KeyHash {
    size_t operator()(const Key &k) const {
        size_t h = std::hash<const void *>()(k.tw) * 1000003u ^ std::hash<const void *>()(k.mod);
        h = h * 1000003u ^ std::hash<uint64_t>()(k.n * 131 + k.size);
        return h * 1000003u ^ std::hash<std::thread::id>()(k.tid);
    }
};
struct Entry {
    b200fhe_ctx *ctx = nullptr;
    std::vector<uint64_t> primes;
    std::chrono::steady_clock::time_point checked;
    bool special_set = false;
    std::vector<const void *> tools_ok;  // DRNSTools whose Q_l primes were compared with the context's
};

std::mutex g_mu;
std::unordered_map<Key, Entry, KeyHash> g_map;

int log2_exact(uint64_t n) {
    int l = 0;
    while ((uint64_t(1) << l) < n) l++;
    if ((uint64_t(1) << l) != n) throw std::invalid_argument("b200shim: polynomial degree is not a power of two");
    return l;
}

// the moduli live on the device only (DModulus = {value, ratio0, ratio1}): read them on the caller's stream
std::vector<uint64_t> read_primes(const DModulus *mod_dev, size_t count, cudaStream_t s) {
    std::vector<uint64_t> raw(3 * count), out(count);
    if (cudaMemcpyAsync(raw.data(), mod_dev, count * sizeof(DModulus), cudaMemcpyDeviceToHost, s) != cudaSuccess ||
        cudaStreamSynchronize(s) != cudaSuccess)
        throw std::runtime_error(std::string("b200shim: reading the moduli of a DNTTTable failed: ") +
                                 cudaGetErrorString(cudaGetLastError()));
    static_assert(sizeof(DModulus) == 24, "DModulus layout");
    for (size_t i = 0; i < count; i++) out[i] = raw[3 * i];
    return out;
}

// A table freed and re-created at the same addresses with other primes would hit a stale entry: the primes are
// re-read at most every few milliseconds (one small copy + a stream synchronisation) and the entry rebuilt on a change.
constexpr auto kRecheck = std::chrono::milliseconds(5);

}  // namespace

b200fhe_ctx *ctx_of(const DNTTTable &t, cudaStream_t s) {
    Key k{t.twiddle(), t.modulus(), t.n(), t.size(), std::this_thread::get_id()};
    if (!k.tw || !k.mod || !k.n || !k.size) throw std::invalid_argument("b200shim: empty DNTTTable");
    std::unique_lock<std::mutex> g(g_mu);
    auto it = g_map.find(k);
    const auto now = std::chrono::steady_clock::now();
    if (it != g_map.end()) {
        Entry &e = it->second;
        if (now - e.checked < kRecheck) return e.ctx;
        if (read_primes(t.modulus(), t.size(), s) == e.primes) {
            e.checked = std::chrono::steady_clock::now();
            return e.ctx;
        }
        b200fhe_ctx_destroy(e.ctx);  // the stream was just synchronised: nothing of ours is in flight on it
        g_map.erase(it);
    }
    Entry e;
    e.primes = read_primes(t.modulus(), t.size(), s);
    int dev = 0;
    cudaGetDevice(&dev);
    ck(b200fhe_ctx_create_from_tables(log2_exact(t.n()), e.primes.data(), e.primes.size(), 0, t.twiddle(), t.twiddle_shoup(),
                                      t.itwiddle(), t.itwiddle_shoup(), t.n_inv_mod_q(), t.n_inv_mod_q_shoup(), (void *) s,
                                      &e.ctx));
    e.checked = std::chrono::steady_clock::now();
    b200fhe_ctx *c = e.ctx;
    g_map.emplace(k, std::move(e));
    return c;
}

b200fhe_ctx *ctx_for_tool(const DNTTTable &t, const phantom::DRNSTool &tool, cudaStream_t s) {
    b200fhe_ctx *c = ctx_of(t, s);
    Key k{t.twiddle(), t.modulus(), t.n(), t.size(), std::this_thread::get_id()};
    std::unique_lock<std::mutex> g(g_mu);
    Entry &e = g_map.at(k);
    if (!e.special_set) {
        if (tool.size_QP() != t.size()) throw std::invalid_argument("b200shim: DRNSTool and DNTTTable disagree on |QP|");
        ck(b200fhe_ctx_set_special(c, tool.size_P()));
        e.special_set = true;
    }
    for (const void *p : e.tools_ok)
        if (p == &tool) return c;
    const auto &base = tool.host_base_Ql().base();
    const size_t Ql = tool.host_base_Ql().size();
    if (Ql > e.primes.size() - tool.size_P()) throw std::invalid_argument("b200shim: |Q_l| exceeds |Q|");
    for (size_t i = 0; i < Ql; i++)
        if (base[i].value() != e.primes[i]) throw std::invalid_argument("b200shim: the DRNSTool's Q_l is not a prefix of the table's primes");
    e.tools_ok.push_back(&tool);
    return c;
}

// 1-D entry points (fnwt_1d / inwt_1d, include/ntt.cuh:157-171) get raw twiddle arrays, forward and inverse in separate
// calls, so the context is built from the primes alone: b200fhe_ctx_create generates the same tables the reference's host
// NTT class does (same minimal 2N-th root, checked against the reference's host code in tests/test_golden.py).
b200fhe_ctx *ctx_of_moduli(const DModulus *mod_dev, size_t count, size_t dim, cudaStream_t s) {
    Key k{nullptr, mod_dev, dim, count, std::this_thread::get_id()};
    if (!mod_dev || !count || !dim) throw std::invalid_argument("b200shim: empty modulus array");
    std::unique_lock<std::mutex> g(g_mu);
    auto it = g_map.find(k);
    const auto now = std::chrono::steady_clock::now();
    if (it != g_map.end()) {
        Entry &e = it->second;
        if (now - e.checked < kRecheck) return e.ctx;
        if (read_primes(mod_dev, count, s) == e.primes) {
            e.checked = std::chrono::steady_clock::now();
            return e.ctx;
        }
        b200fhe_ctx_destroy(e.ctx);
        g_map.erase(it);
    }
    Entry e;
    e.primes = read_primes(mod_dev, count, s);
    int dev = 0;
    cudaGetDevice(&dev);
    ck(b200fhe_ctx_create(log2_exact(dim), e.primes.data(), e.primes.size(), 0, dev, &e.ctx));
    e.checked = std::chrono::steady_clock::now();
    b200fhe_ctx *c = e.ctx;
    g_map.emplace(k, std::move(e));
    return c;
}

b200fhe_ctx *ctx_by_moduli(const DModulus *mod_dev) {
    std::unique_lock<std::mutex> g(g_mu);
    const auto tid = std::this_thread::get_id();
    for (auto &kv : g_map)
        if (kv.first.mod == mod_dev && kv.first.tw && kv.first.tid == tid && kv.second.special_set) return kv.second.ctx;
    return nullptr;
}

int level() {
    static const int l = [] {
        const char *e = getenv("B200SHIM_LEVEL");
        const int v = e ? atoi(e) : 3;
        return v < 1 ? 1 : (v > 3 ? 3 : v);
    }();
    return l;
}

void forget_all() {
    std::unique_lock<std::mutex> g(g_mu);
    cudaDeviceSynchronize();
    for (auto &kv : g_map) b200fhe_ctx_destroy(kv.second.ctx);
    g_map.clear();
}

}  // namespace b200shim

extern "C" void b200shim_forget_all(void) { b200shim::forget_all(); }
extern "C" const char *b200shim_version(void) { return "b200shim 0.1 over libb200fhe.so"; }
