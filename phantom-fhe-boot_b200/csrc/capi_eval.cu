// capi_eval.cu — C ABI of the evaluator and bootstrap layers (include/b200fhe.h, second half).
#include <cstring>

#include "../../include/b200fhe.h"
#include "bootstrapper.h"
#include "launch.cuh"
#include "dnn.h"
#include "encoder.h"
#include "keygen.h"

using namespace b200fhe;

namespace {
struct EvalBox {
    std::unique_ptr<Evaluator> own;  // set when created through b200fhe_eval_create
    Evaluator *ev = nullptr;
};
struct CtBox {
    Ct ct;
};
struct BootBox {
    std::unique_ptr<Bootstrapper> bt;
    EvalBox evbox;
};

template <class F>
int guarded(F &&f) {
    try {
        f();
        return B200FHE_OK;
    } catch (const EvalError &e) {
        return fail(e.code, "%s", e.what());
    } catch (const std::bad_alloc &) {
        return fail(B200FHE_ERR_ALLOC, "out of host memory");
    } catch (const std::exception &e) {
        return fail(B200FHE_ERR_ARG, "%s", e.what());
    }
}

Evaluator *E(b200fhe_eval *h) {
    if (!h) throw EvalError(B200FHE_ERR_ARG, "null evaluator");
    return reinterpret_cast<EvalBox *>(h)->ev;
}
Ct &C(b200fhe_ct *h) {
    if (!h) throw EvalError(B200FHE_ERR_ARG, "null ciphertext");
    return reinterpret_cast<CtBox *>(h)->ct;
}
const Ct &C(const b200fhe_ct *h) {
    if (!h) throw EvalError(B200FHE_ERR_ARG, "null ciphertext");
    return reinterpret_cast<const CtBox *>(h)->ct;
}
void put(b200fhe_ct **out, Ct &&x) {
    if (!out) throw EvalError(B200FHE_ERR_ARG, "null output");
    *out = reinterpret_cast<b200fhe_ct *>(new CtBox{std::move(x)});
}
Bootstrapper *B(b200fhe_boot *h) {
    if (!h) throw EvalError(B200FHE_ERR_ARG, "null bootstrapper");
    return reinterpret_cast<BootBox *>(h)->bt.get();
}
void need_device(Context *c) {
    if (!c) throw EvalError(B200FHE_ERR_ARG, "null context");
    if (c->host_only) throw EvalError(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    if (c->size_P == 0) throw EvalError(B200FHE_ERR_ARG, "context has no special primes (size_P == 0)");
}
}  // namespace

Evaluator *b200fhe::eval_of(void *ev) { return ev ? reinterpret_cast<EvalBox *>(ev)->ev : nullptr; }


extern "C" {

int b200fhe_eval_create(b200fhe_ctx *ctx, void *stream, b200fhe_eval **out) {
    return guarded([&] {
        auto *c = reinterpret_cast<Context *>(ctx);
        need_device(c);
        if (!out) throw EvalError(B200FHE_ERR_ARG, "null output");
        auto *box = new EvalBox();
        try {
            box->own = std::make_unique<Evaluator>(c, reinterpret_cast<cudaStream_t>(stream));
        } catch (...) {
            delete box;
            throw;
        }
        box->ev = box->own.get();
        *out = reinterpret_cast<b200fhe_eval *>(box);
    });
}

int b200fhe_eval_destroy(b200fhe_eval *ev) {
    auto *box = reinterpret_cast<EvalBox *>(ev);
    if (box && box->own) delete box;  // the evaluator of a bootstrapper dies with it
    return B200FHE_OK;
}

int b200fhe_eval_scaling_factors(b200fhe_eval *ev, int which, double *out, size_t capacity, size_t *count) {
    return guarded([&] {
        const auto &v = which == 0 ? E(ev)->sf : E(ev)->sf_big;
        if (count) *count = v.size();
        if (out) {
            if (capacity < v.size()) throw EvalError(B200FHE_ERR_ARG, "capacity too small");
            std::memcpy(out, v.data(), v.size() * sizeof(double));
        }
    });
}

int b200fhe_eval_set_relin_key(b200fhe_eval *ev, const uint64_t *const *digits, size_t n) {
    return guarded([&] {
        if (!digits || n == 0) throw EvalError(B200FHE_ERR_ARG, "null evaluation key");
        E(ev)->relin.assign(digits, digits + n);
        E(ev)->have_relin = true;
    });
}

int b200fhe_eval_set_galois_key(b200fhe_eval *ev, uint32_t auto_index, const uint64_t *const *digits, size_t n) {
    return guarded([&] {
        if (!digits || n == 0 || !(auto_index & 1)) throw EvalError(B200FHE_ERR_ARG, "bad Galois key");
        E(ev)->galois[auto_index].assign(digits, digits + n);
    });
}

uint32_t b200fhe_auto_index(int logN, int32_t rotation) {
    return find_automorphism_index_2n_complex(rotation, 2u << logN);
}

int b200fhe_ct_import(b200fhe_eval *ev, const uint64_t *dev, size_t size_Ql, uint32_t polys, int extended, double scale,
                      uint32_t noise, b200fhe_ct **out) {
    return guarded([&] {
        if (!dev || polys == 0) throw EvalError(B200FHE_ERR_ARG, "bad argument");
        put(out, E(ev)->wrap_copy((const u64 *) dev, size_Ql, polys, extended != 0, scale, noise));
    });
}

int b200fhe_ct_export(b200fhe_eval *ev, const b200fhe_ct *ct, uint64_t *dev_out, size_t capacity_words) {
    return guarded([&] {
        const Ct &x = C(ct);
        Evaluator *e = E(ev);
        const size_t words = size_t(x.polys) * x.limbs * e->c->N;
        if (!dev_out || capacity_words < words) throw EvalError(B200FHE_ERR_ARG, "capacity too small");
        e->cu(cudaMemcpyAsync(dev_out, x.data(), words * 8, cudaMemcpyDeviceToDevice, e->st), "export");
    });
}

int b200fhe_ct_describe(const b200fhe_ct *ct, b200fhe_ct_desc *out) {
    return guarded([&] {
        const Ct &x = C(ct);
        if (!out) throw EvalError(B200FHE_ERR_ARG, "null output");
        out->size_Ql = x.Ql;
        out->limbs = x.limbs;
        out->polys = x.polys;
        out->noise_scale_deg = x.noise;
        out->scale = x.scale;
    });
}

int b200fhe_ct_clone(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out) {
    return guarded([&] {
        (void) E(ev);
        Ct y = C(ct);  // shares the block; copy-on-write keeps value semantics
        put(out, std::move(y));
    });
}

int b200fhe_ct_free(b200fhe_ct *ct) {
    delete reinterpret_cast<CtBox *>(ct);
    return B200FHE_OK;
}

int b200fhe_eval_mod_reduce(b200fhe_eval *ev, b200fhe_ct *ct, size_t levels) {
    return guarded([&] { E(ev)->mod_reduce(C(ct), levels); });
}
int b200fhe_eval_mod_switch_level(b200fhe_eval *ev, b200fhe_ct *ct, size_t levels) {
    return guarded([&] { E(ev)->mod_switch_level(C(ct), levels); });
}
int b200fhe_eval_mult_const(b200fhe_eval *ev, b200fhe_ct *ct, double scalar) {
    return guarded([&] { E(ev)->mult_const(C(ct), scalar); });
}
int b200fhe_eval_add_const(b200fhe_eval *ev, b200fhe_ct *ct, double operand) {
    return guarded([&] { E(ev)->add_const(C(ct), operand); });
}
int b200fhe_eval_mult_by_integer(b200fhe_eval *ev, b200fhe_ct *ct, uint64_t v) {
    return guarded([&] { E(ev)->mult_by_integer(C(ct), v); });
}
int b200fhe_eval_mult_by_monomial(b200fhe_eval *ev, b200fhe_ct *ct, uint32_t power) {
    return guarded([&] { E(ev)->mult_by_monomial(C(ct), power); });
}
int b200fhe_eval_add_auto(b200fhe_eval *ev, b200fhe_ct *a, const b200fhe_ct *b) {
    return guarded([&] { E(ev)->add_auto(C(a), C(b)); });
}
int b200fhe_eval_sub_auto(b200fhe_eval *ev, b200fhe_ct *a, const b200fhe_ct *b) {
    return guarded([&] { E(ev)->sub_auto(C(a), C(b)); });
}
int b200fhe_eval_mult_auto(b200fhe_eval *ev, const b200fhe_ct *a, const b200fhe_ct *b, b200fhe_ct **out) {
    return guarded([&] { put(out, E(ev)->mult_auto(C(a), C(b))); });
}
int b200fhe_eval_square(b200fhe_eval *ev, const b200fhe_ct *a, b200fhe_ct **out) {
    return guarded([&] { put(out, E(ev)->square(C(a))); });
}
int b200fhe_eval_raise_mod(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out) {
    return guarded([&] { put(out, E(ev)->raise_mod(C(ct))); });
}
int b200fhe_eval_rotate_fused(b200fhe_eval *ev, const b200fhe_ct *ct, int32_t index, b200fhe_ct **out) {
    return guarded([&] { put(out, E(ev)->rotate_fused(C(ct), index)); });
}
int b200fhe_eval_conj_fused(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out) {
    return guarded([&] { put(out, E(ev)->conj_fused(C(ct))); });
}
int b200fhe_eval_chebyshev(b200fhe_eval *ev, const b200fhe_ct *ct, const double *coeffs, size_t n, double a, double b,
                           b200fhe_ct **out) {
    return guarded([&] {
        if (!coeffs || n == 0) throw EvalError(B200FHE_ERR_ARG, "The coefficients vector can not be empty");
        put(out, E(ev)->chebyshev_series(C(ct), std::vector<double>(coeffs, coeffs + n), a, b));
    });
}

int b200fhe_eval_fast_rotations_ext(b200fhe_eval *ev, const b200fhe_ct *ct, const int32_t *indices, size_t n,
                                    int add_first, b200fhe_ct **outs) {
    return guarded([&] {
        if (!indices || !outs) throw EvalError(B200FHE_ERR_ARG, "null argument");
        Evaluator *e = E(ev);
        auto digits = e->fast_rotation_precompute(C(ct));
        std::vector<Ct> r = e->fast_rotations_ext(C(ct), std::vector<int32_t>(indices, indices + n), digits, add_first != 0);
        for (size_t i = 0; i < n; i++) put(outs + i, std::move(r[i]));
    });
}
int b200fhe_eval_keyswitch_ext(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out) {
    return guarded([&] { put(out, E(ev)->keyswitch_ext(C(ct))); });
}
int b200fhe_eval_keyswitch_down(b200fhe_eval *ev, const b200fhe_ct *ct, int first_only, b200fhe_ct **out) {
    return guarded([&] {
        Ct x = C(ct);  // moddown transforms the P limbs in place: work on a private copy
        put(out, first_only ? E(ev)->keyswitch_down_first(x) : E(ev)->keyswitch_down(x));
    });
}
int b200fhe_eval_mult_ext(b200fhe_eval *ev, b200fhe_ct *ct, const uint64_t *plain_dev, double plain_scale) {
    return guarded([&] {
        if (!plain_dev) throw EvalError(B200FHE_ERR_ARG, "null plaintext");
        Ct &x = C(ct);
        E(ev)->mult_ext(x, (const u64 *) plain_dev);
        x.scale = x.scale * plain_scale;
        x.noise += 1;
    });
}
int b200fhe_eval_add_ext(b200fhe_eval *ev, b200fhe_ct *a, const b200fhe_ct *b) {
    return guarded([&] { E(ev)->add_ext(C(a), C(b)); });
}
size_t b200fhe_eval_pool_bytes(b200fhe_eval *ev) {
    return ev ? reinterpret_cast<EvalBox *>(ev)->ev->pool.bytes_reserved() : 0;
}

// ------------------------------------------------------------------------------------------- bootstrap
int b200fhe_boot_create(b200fhe_ctx *ctx, void *stream, uint32_t budget_enc, uint32_t budget_dec, uint32_t dim1_enc,
                        uint32_t dim1_dec, uint32_t num_slots, uint32_t correction_factor, b200fhe_boot **out) {
    return guarded([&] {
        auto *c = reinterpret_cast<Context *>(ctx);
        // a host-only context may carry the setup logic (parameters, rotation indices, stage shapes); every
        // compute entry then fails in Evaluator::alloc with "no CPU fallback"
        if (!c) throw EvalError(B200FHE_ERR_ARG, "null context");
        if (c->size_P == 0) throw EvalError(B200FHE_ERR_ARG, "context has no special primes (size_P == 0)");
        if (!out) throw EvalError(B200FHE_ERR_ARG, "null output");
        auto box = std::make_unique<BootBox>();
        box->bt = std::make_unique<Bootstrapper>(c, reinterpret_cast<cudaStream_t>(stream), budget_enc, budget_dec,
                                                 dim1_enc, dim1_dec, num_slots, correction_factor);
        box->evbox.ev = &box->bt->ev;
        *out = reinterpret_cast<b200fhe_boot *>(box.release());
    });
}

int b200fhe_boot_destroy(b200fhe_boot *bt) {
    delete reinterpret_cast<BootBox *>(bt);
    return B200FHE_OK;
}

b200fhe_eval *b200fhe_boot_evaluator(b200fhe_boot *bt) {
    return bt ? reinterpret_cast<b200fhe_eval *>(&reinterpret_cast<BootBox *>(bt)->evbox) : nullptr;
}

int b200fhe_boot_params(b200fhe_boot *bt, int which, int32_t *out9) {
    return guarded([&] {
        if (!out9) throw EvalError(B200FHE_ERR_ARG, "null output");
        const FFTParams &p = which == 0 ? B(bt)->enc : B(bt)->dec;
        const int32_t v[9] = {p.levelBudget, p.layersCollapse, p.remCollapse, p.numRotations, p.b,
                              p.g,           p.numRotationsRem, p.bRem,       p.gRem};
        std::memcpy(out9, v, sizeof v);
    });
}

int b200fhe_boot_rotation_indices(b200fhe_boot *bt, int32_t *out, size_t capacity, size_t *count) {
    return guarded([&] {
        const auto v = B(bt)->rotation_indices();
        if (count) *count = v.size();
        if (out) {
            if (capacity < v.size()) throw EvalError(B200FHE_ERR_ARG, "capacity too small");
            std::memcpy(out, v.data(), v.size() * sizeof(int32_t));
        }
    });
}

int b200fhe_boot_stage_info(b200fhe_boot *bt, int which, size_t stage, size_t *n_stages, size_t *n_diagonals,
                            size_t *size_Ql) {
    return guarded([&] {
        Bootstrapper *b = B(bt);
        if (n_stages) *n_stages = b->stage_count(which);
        if (n_diagonals) *n_diagonals = b->stage_size(which, stage);
        if (size_Ql) *size_Ql = b->stage_level(which, stage);
    });
}

int b200fhe_boot_set_diagonal(b200fhe_boot *bt, int which, size_t stage, size_t index, const uint64_t *dev,
                              size_t size_Ql, double scale) {
    return guarded([&] {
        if (!dev) throw EvalError(B200FHE_ERR_ARG, "null plaintext");
        B(bt)->set_diagonal(which, stage, index, (const u64 *) dev, size_Ql, scale);
    });
}

size_t b200fhe_boot_peer_arena_bytes(b200fhe_boot *bt) {
    return bt ? reinterpret_cast<BootBox *>(bt)->bt->peer_arena_bytes() : 0;
}

int b200fhe_boot_set_peer_shard(b200fhe_boot *bt, uint32_t world, uint32_t rank, void *const *arena_bases, size_t arena_bytes,
                                uint32_t *counter, int virtual_ranks) {
    return guarded([&] {
        Bootstrapper *b = B(bt);
        PeerShard s{};
        if (world <= 1) {  // back to the single-GPU stage
            b->shard = s;
            return;
        }
        if (world > 9 || rank >= world) throw EvalError(B200FHE_ERR_ARG, "limb-sharded bootstrap: 2 .. 9 ranks");
        s.world = world;
        s.rank = rank;
        s.virtual_ranks = virtual_ranks != 0;
        if (!s.virtual_ranks) {
            const size_t half = b->peer_arena_half_bytes();
            if (!arena_bases || !counter) throw EvalError(B200FHE_ERR_ARG, "null arena table / counter");
            if (arena_bytes < b->peer_arena_bytes()) throw EvalError(B200FHE_ERR_ARG, "arena smaller than b200fhe_boot_peer_arena_bytes()");
            for (uint32_t g = 0; g < world; g++) {
                if (!arena_bases[g]) throw EvalError(B200FHE_ERR_ARG, "null arena base");
                s.bases[g] = arena_bases[g];
                s.flags[g] = reinterpret_cast<uint64_t *>(static_cast<char *>(arena_bases[g]) + b->peer_arena_bytes() - 256);
            }
            s.counter = counter;
            s.half_bytes = half;
            s.sums_bytes = b->peer_sums_bytes();
            s.slots_bytes = b->peer_slots_bytes();
            s.evalmod_off = 2 * half;
        }
        b->shard = s;
    });
}

// ---- EvalBootstrap of one fixed input buffer as a CUDA graph --------------------------------------------------------------
// A bootstrap is ~1400 dependent launches on up to 5 streams; issuing them costs the host ~20 ms, which is what limits the
// limb-sharded bootstrap (15 ms of GPU work per rank on 8 GPUs) and any caller whose host thread is busy.  Captured once
// (relaxed stream capture of the bootstrapper's stream; the digit lanes fork and join with events, so they are part of the
// graph) and replayed, the host cost is one cudaGraphLaunch.  The graph bakes addresses: it runs on a block pool of its
// own that is never handed to anybody else, the context's workspaces are pinned (an operation that would have to grow
// them fails instead of moving them), and its input is the buffer of the ciphertext given at capture.
struct BootGraph {
    Bootstrapper *bt = nullptr;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    Pool pool;   // every block the captured bootstrap uses
    Ct in, out;
    bool pinned = false;
    ~BootGraph() {
        if (pinned) bt->ev.c->pin_workspaces(false);
        if (exec) cudaGraphExecDestroy(exec);
        if (graph) cudaGraphDestroy(graph);
        out = Ct();
        in = Ct();
    }
};

int b200fhe_boot_graph_capture(b200fhe_boot *bt_, const b200fhe_ct *ct, b200fhe_boot_graph **out) {
    return guarded([&] {
        Bootstrapper *bt = B(bt_);
        if (!out) throw EvalError(B200FHE_ERR_ARG, "null output");
        Evaluator &ev = bt->ev;
        Context *c = ev.c;
        auto g = std::make_unique<BootGraph>();
        g->bt = bt;
        g->in = C(ct);
        g->in.data();  // a pending combination is evaluated before the capture
        if (bt->shard.world > 1 && !bt->shard.virtual_ranks && ((bt->stage_count(0) + bt->stage_count(1)) & 1))
            throw EvalError(B200FHE_ERR_UNSUPPORTED, "limb-sharded graph needs an even number of linear-transform stages (arena halves)");
        // one eager run: level tables, permutation tables, monomials, kernel attributes and workspaces reach their final state
        bt->eval(g->in);
        ev.cu(cudaStreamSynchronize(ev.st), "warm-up bootstrap");
        ev.pool.trim();
        c->pin_workspaces(true);
        g->pinned = true;
        ev.pool.swap_state(g->pool);
        const bool keep_pdl = getenv("B200FHE_GRAPH_PDL") && getenv("B200FHE_GRAPH_PDL")[0] == '1';
        if (!keep_pdl) pdl_suppress(true);
        // the legacy default stream cannot be captured: the capture runs on a stream of its own, the graph is launched on ev.st
        cudaStream_t user = ev.st, cap = nullptr;
        cudaError_t e = cudaStreamCreateWithFlags(&cap, cudaStreamNonBlocking);
        std::string err;
        if (e == cudaSuccess) e = cudaStreamBeginCapture(cap, cudaStreamCaptureModeRelaxed);
        if (e == cudaSuccess) {
            ev.st = cap;
            try {
                g->out = bt->eval(g->in);
                g->out.data();
            } catch (const std::exception &x) {
                err = x.what();
            }
            ev.st = user;
            e = cudaStreamEndCapture(cap, &g->graph);
        }
        if (cap) cudaStreamDestroy(cap);
        if (!keep_pdl) pdl_suppress(false);
        ev.pool.swap_state(g->pool);
        if (!err.empty()) throw EvalError(B200FHE_ERR_CUDA, "bootstrap under stream capture: " + err);
        ev.cu(e, "stream capture of the bootstrap");
        ev.cu(cudaGraphInstantiate(&g->exec, g->graph, 0), "cudaGraphInstantiate");
        *out = reinterpret_cast<b200fhe_boot_graph *>(g.release());
    });
}

int b200fhe_boot_graph_launch(b200fhe_boot_graph *g_) {
    return guarded([&] {
        auto *g = reinterpret_cast<BootGraph *>(g_);
        if (!g || !g->exec) throw EvalError(B200FHE_ERR_ARG, "null graph");
        g->bt->ev.cu(cudaGraphLaunch(g->exec, g->bt->ev.st), "cudaGraphLaunch");
        count_launches(1);
    });
}

int b200fhe_boot_graph_io(b200fhe_boot_graph *g_, uint64_t **input_limbs, b200fhe_ct **output) {
    return guarded([&] {
        auto *g = reinterpret_cast<BootGraph *>(g_);
        if (!g) throw EvalError(B200FHE_ERR_ARG, "null graph");
        if (input_limbs) *input_limbs = reinterpret_cast<uint64_t *>(g->in.data());
        if (output) {
            Ct o = g->out;
            put(output, std::move(o));
        }
    });
}

int b200fhe_boot_graph_destroy(b200fhe_boot_graph *g_) {
    auto *g = reinterpret_cast<BootGraph *>(g_);
    if (g) {
        cudaStreamSynchronize(g->bt->ev.st);
        delete g;
    }
    return B200FHE_OK;
}

int b200fhe_boot_eval(b200fhe_boot *bt, const b200fhe_ct *ct, b200fhe_ct **out) {
    return guarded([&] { put(out, B(bt)->eval(C(ct))); });
}
int b200fhe_boot_eval_iter(b200fhe_boot *bt, const b200fhe_ct *ct, uint32_t num_iterations, uint32_t precision, b200fhe_ct **out) {
    return guarded([&] { put(out, B(bt)->eval(C(ct), num_iterations, precision)); });
}
int b200fhe_boot_coeffs_to_slots(b200fhe_boot *bt, const b200fhe_ct *ct, b200fhe_ct **out) {
    return guarded([&] { put(out, B(bt)->coeffs_to_slots(C(ct))); });
}
int b200fhe_boot_slots_to_coeffs(b200fhe_boot *bt, const b200fhe_ct *ct, b200fhe_ct **out) {
    return guarded([&] { put(out, B(bt)->slots_to_coeffs(C(ct))); });
}

}  // extern "C"

// ------------------------------------------------------------------------------------------- encoder
extern "C" int b200fhe_encoder_create(b200fhe_ctx *ctx, b200fhe_encoder **out) {
    return guarded([&] {
        auto *c = reinterpret_cast<Context *>(ctx);
        if (!c || !out) throw EvalError(B200FHE_ERR_ARG, "null argument");
        if (c->host_only) throw EvalError(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
        *out = reinterpret_cast<b200fhe_encoder *>(new Encoder(c));
    });
}
extern "C" void b200fhe_encoder_destroy(b200fhe_encoder *enc) { delete reinterpret_cast<Encoder *>(enc); }
extern "C" size_t b200fhe_encoder_slots(const b200fhe_encoder *enc) {
    return enc ? reinterpret_cast<const Encoder *>(enc)->slots : 0;
}
extern "C" int b200fhe_encode(b200fhe_encoder *enc, const double *values_re_im, size_t n, size_t size_Ql, double scale,
                              int ext, int sparse, uint64_t *out_dev, void *stream) {
    if (!enc) return fail(B200FHE_ERR_ARG, "null encoder");
    return reinterpret_cast<Encoder *>(enc)->encode(reinterpret_cast<const cuDoubleComplex *>(values_re_im), n, size_Ql, scale,
                                                    ext != 0, sparse != 0, (u64 *) out_dev,
                                                    reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int b200fhe_encode_sparse_batch(b200fhe_encoder *enc, const double *values_re_im, size_t n, size_t batch,
                                           size_t size_Ql, double scale, uint64_t *out_dev, void *stream) {
    if (!enc) return fail(B200FHE_ERR_ARG, "null encoder");
    return reinterpret_cast<Encoder *>(enc)->encode_sparse_batch(reinterpret_cast<const cuDoubleComplex *>(values_re_im), n, batch,
                                                                 size_Ql, scale, (u64 *) out_dev,
                                                                 reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int b200fhe_decode(b200fhe_encoder *enc, const uint64_t *plain_dev, size_t size_Ql, int ext, double scale,
                              double *out_re_im, void *stream) {
    if (!enc) return fail(B200FHE_ERR_ARG, "null encoder");
    return reinterpret_cast<Encoder *>(enc)->decode((const u64 *) plain_dev, size_Ql, ext != 0, scale,
                                                    reinterpret_cast<cuDoubleComplex *>(out_re_im),
                                                    reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int b200fhe_boot_precompute(b200fhe_boot *bt, b200fhe_encoder *enc) {
    return guarded([&] {
        if (!enc) throw EvalError(B200FHE_ERR_ARG, "null encoder");
        B(bt)->precompute(*reinterpret_cast<Encoder *>(enc));
    });
}
extern "C" int b200fhe_boot_get_diagonal(b200fhe_boot *bt, int which, size_t stage, size_t index, const uint64_t **dev,
                                         size_t *size_Ql, double *scale) {
    return guarded([&] {
        Bootstrapper *b = B(bt);
        const auto &A = which == 0 ? b->A_enc : b->A_dec;
        if (stage >= A.size() || index >= A[stage].size()) throw EvalError(B200FHE_ERR_ARG, "diagonal index out of range");
        const Diagonal &d = A[stage][index];
        if (dev) *dev = (const uint64_t *) d.p;
        if (size_Ql) *size_Ql = d.Ql;
        if (scale) *scale = d.scale;
    });
}

// ------------------------------------------------------------------------------------------- key generation
namespace {
KeyGen *K(b200fhe_keygen *h) {
    if (!h) throw EvalError(B200FHE_ERR_ARG, "null key generator");
    return reinterpret_cast<KeyGen *>(h);
}
}  // namespace
extern "C" int b200fhe_keygen_create(b200fhe_ctx *ctx, const uint8_t *seed32, void *stream, b200fhe_keygen **out) {
    return guarded([&] {
        if (!out) throw EvalError(B200FHE_ERR_ARG, "null output");
        *out = reinterpret_cast<b200fhe_keygen *>(
            new KeyGen(reinterpret_cast<Context *>(ctx), seed32, reinterpret_cast<cudaStream_t>(stream)));
    });
}
extern "C" int b200fhe_keygen_create_reference_prng(b200fhe_ctx *ctx, uint64_t master_seed, void *stream, b200fhe_keygen **out) {
    return guarded([&] {
        if (!out) throw EvalError(B200FHE_ERR_ARG, "null output");
        *out = reinterpret_cast<b200fhe_keygen *>(
            new KeyGen(reinterpret_cast<Context *>(ctx), master_seed, reinterpret_cast<cudaStream_t>(stream), KeyGen::RefPrng{}));
    });
}
extern "C" void b200fhe_keygen_destroy(b200fhe_keygen *kg) { delete reinterpret_cast<KeyGen *>(kg); }
extern "C" const uint64_t *b200fhe_keygen_secret(const b200fhe_keygen *kg) {
    return kg ? (const uint64_t *) reinterpret_cast<const KeyGen *>(kg)->secret() : nullptr;
}
extern "C" int b200fhe_keygen_set_secret(b200fhe_keygen *kg, const uint64_t *secret_dev) {
    return guarded([&] { K(kg)->set_secret((const u64 *) secret_dev); });
}
extern "C" int b200fhe_keygen_get_secret(b200fhe_keygen *kg, uint64_t *secret_dev_out) {
    return guarded([&] {
        KeyGen *k = K(kg);
        if (!secret_dev_out) throw EvalError(B200FHE_ERR_ARG, "null output");
        if (cudaMemcpyAsync(secret_dev_out, k->secret(), k->c->size_QP * k->c->N * 8, cudaMemcpyDeviceToDevice, k->st) != cudaSuccess)
            throw EvalError(B200FHE_ERR_CUDA, "copy of the secret key failed");
    });
}
extern "C" int b200fhe_keygen_relin_key(b200fhe_keygen *kg, b200fhe_evk **out) {
    return guarded([&] {
        if (!out) throw EvalError(B200FHE_ERR_ARG, "null output");
        *out = reinterpret_cast<b200fhe_evk *>(K(kg)->relin_key().release());
    });
}
extern "C" int b200fhe_keygen_galois_key(b200fhe_keygen *kg, uint32_t auto_index, b200fhe_evk **out) {
    return guarded([&] {
        if (!out) throw EvalError(B200FHE_ERR_ARG, "null output");
        *out = reinterpret_cast<b200fhe_evk *>(K(kg)->galois_key(auto_index).release());
    });
}
extern "C" size_t b200fhe_evk_digits(const b200fhe_evk *key, const uint64_t **ptrs, size_t capacity) {
    if (!key) return 0;
    const auto p = reinterpret_cast<const SwitchKey *>(key)->ptrs();
    if (ptrs)
        for (size_t i = 0; i < p.size() && i < capacity; i++) ptrs[i] = p[i];
    return p.size();
}
extern "C" void b200fhe_evk_free(b200fhe_evk *key) { delete reinterpret_cast<SwitchKey *>(key); }
extern "C" int b200fhe_keygen_public_key(b200fhe_keygen *kg, uint64_t *pk_dev_out) {
    return guarded([&] { K(kg)->public_key((u64 *) pk_dev_out); });
}
extern "C" int b200fhe_encrypt_symmetric(b200fhe_keygen *kg, const uint64_t *plain_dev, size_t size_Ql, uint64_t *ct_dev_out) {
    return guarded([&] { K(kg)->encrypt_symmetric((const u64 *) plain_dev, size_Ql, (u64 *) ct_dev_out); });
}
extern "C" int b200fhe_encrypt_asymmetric(b200fhe_keygen *kg, const uint64_t *pk_dev, const uint64_t *plain_dev,
                                          size_t size_Ql, uint64_t *ct_dev_out) {
    return guarded([&] { K(kg)->encrypt_asymmetric((const u64 *) pk_dev, (const u64 *) plain_dev, size_Ql, (u64 *) ct_dev_out); });
}
extern "C" int b200fhe_decrypt(b200fhe_keygen *kg, const uint64_t *ct_dev, size_t size_Ql, uint32_t polys,
                               uint64_t *plain_dev_out) {
    return guarded([&] { K(kg)->decrypt((const u64 *) ct_dev, size_Ql, polys, (u64 *) plain_dev_out); });
}
extern "C" int b200fhe_keygen_sample(b200fhe_keygen *kg, int kind, uint64_t *out_dev, size_t limbs, size_t size_Ql) {
    return guarded([&] { K(kg)->sample(kind, (u64 *) out_dev, limbs, size_Ql); });
}
extern "C" int b200fhe_chacha20_block(b200fhe_ctx *ctx, const uint32_t *key8, uint32_t counter, const uint32_t *nonce3,
                                      uint32_t *out16_host) {
    auto *c = reinterpret_cast<Context *>(ctx);
    if (!c || !key8 || !nonce3 || !out16_host) return fail(B200FHE_ERR_ARG, "null argument");
    if (c->host_only) return fail(B200FHE_ERR_CUDA, "context has no device (host-only); no CPU fallback");
    return chacha20_block_device(key8, counter, nonce3, out16_host);
}

// ------------------------------------------------------------------------------------------- batched ciphertexts
extern "C" int b200fhe_ct_pack(b200fhe_eval *ev, const b200fhe_ct *const *cts, size_t n, b200fhe_ct **out) {
    return guarded([&] {
        if (!cts || n == 0) throw EvalError(B200FHE_ERR_ARG, "nothing to pack");
        std::vector<Ct> v;
        for (size_t i = 0; i < n; i++) v.push_back(C(cts[i]));
        put(out, E(ev)->pack(v));
    });
}
extern "C" int b200fhe_ct_unpack(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **outs, size_t capacity, size_t *count) {
    return guarded([&] {
        const Ct &x = C(ct);
        if (count) *count = x.batch;
        if (!outs) return;
        if (capacity < x.batch) throw EvalError(B200FHE_ERR_ARG, "capacity too small");
        std::vector<Ct> v = E(ev)->unpack(x);
        for (size_t i = 0; i < v.size(); i++) outs[i] = reinterpret_cast<b200fhe_ct *>(new CtBox{std::move(v[i])});
    });
}

// ------------------------------------------------------------------------------------------- SEAL-style surface
extern "C" int b200fhe_eval_set_seal_galois_key(b200fhe_eval *ev, uint32_t galois_elt, const uint64_t *const *digits, size_t n) {
    return guarded([&] {
        if (!digits || n == 0 || !(galois_elt & 1)) throw EvalError(B200FHE_ERR_ARG, "bad Galois key");
        E(ev)->seal_galois[galois_elt].assign(digits, digits + n);
    });
}
extern "C" int b200fhe_eval_negate(b200fhe_eval *ev, b200fhe_ct *ct) { return guarded([&] { E(ev)->negate(C(ct)); }); }
extern "C" int b200fhe_eval_add_plain(b200fhe_eval *ev, b200fhe_ct *ct, const uint64_t *plain_dev, double plain_scale, int sub) {
    return guarded([&] { E(ev)->add_plain(C(ct), (const u64 *) plain_dev, plain_scale, sub != 0); });
}
extern "C" int b200fhe_eval_multiply_plain(b200fhe_eval *ev, b200fhe_ct *ct, const uint64_t *plain_dev, double plain_scale) {
    return guarded([&] { E(ev)->multiply_plain(C(ct), (const u64 *) plain_dev, plain_scale); });
}
extern "C" int b200fhe_eval_add_many(b200fhe_eval *ev, const b200fhe_ct *const *cts, size_t n, b200fhe_ct **out) {
    return guarded([&] {
        if (!cts) throw EvalError(B200FHE_ERR_ARG, "null list");
        std::vector<Ct> v;
        for (size_t i = 0; i < n; i++) v.push_back(C(cts[i]));
        put(out, E(ev)->add_many(v));
    });
}
extern "C" int b200fhe_eval_multiply(b200fhe_eval *ev, const b200fhe_ct *a, const b200fhe_ct *b, b200fhe_ct **out) {
    return guarded([&] { put(out, E(ev)->multiply(C(a), C(b))); });
}
extern "C" int b200fhe_eval_relinearize(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out) {
    return guarded([&] { put(out, E(ev)->relinearize(C(ct))); });
}
extern "C" int b200fhe_eval_rescale_to_next(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out) {
    return guarded([&] { put(out, E(ev)->rescale_to_next(C(ct))); });
}
extern "C" int b200fhe_eval_mod_switch_to_next(b200fhe_eval *ev, const b200fhe_ct *ct, b200fhe_ct **out) {
    return guarded([&] { put(out, E(ev)->mod_switch_to_next(C(ct))); });
}
extern "C" int b200fhe_eval_apply_galois(b200fhe_eval *ev, const b200fhe_ct *ct, uint32_t galois_elt, b200fhe_ct **out) {
    return guarded([&] { put(out, E(ev)->apply_galois(C(ct), galois_elt)); });
}
extern "C" int b200fhe_eval_rotate(b200fhe_eval *ev, const b200fhe_ct *ct, int step, b200fhe_ct **out) {
    return guarded([&] { put(out, E(ev)->rotate(C(ct), step)); });
}
extern "C" int b200fhe_eval_hoisting(b200fhe_eval *ev, const b200fhe_ct *ct, const int *steps, size_t n, b200fhe_ct **out) {
    return guarded([&] {
        if (!steps) throw EvalError(B200FHE_ERR_ARG, "null steps");
        put(out, E(ev)->hoisting(C(ct), std::vector<int>(steps, steps + n)));
    });
}
extern "C" int b200fhe_keygen_galois_key_seal(b200fhe_keygen *kg, uint32_t galois_elt, b200fhe_evk **out) {
    return guarded([&] {
        if (!out) throw EvalError(B200FHE_ERR_ARG, "null output");
        if (!kg) throw EvalError(B200FHE_ERR_ARG, "null key generator");
        *out = reinterpret_cast<b200fhe_evk *>(reinterpret_cast<KeyGen *>(kg)->galois_key_seal(galois_elt).release());
    });
}

// ------------------------------------------------------------------------------------------- DNN layers
namespace {
DNN *D(b200fhe_dnn *h) {
    if (!h) throw EvalError(B200FHE_ERR_ARG, "null DNN");
    return reinterpret_cast<DNN *>(h);
}
TensorCT tensor_in(const b200fhe_ct *const *cts, const b200fhe_tensor_desc *d) {
    if (!cts || !d) throw EvalError(B200FHE_ERR_ARG, "null tensor");
    TensorCT t;
    t.width = d->width;
    t.slotstr = d->slotstr;
    t.num_ch = d->num_ch;
    const size_t n = d->num_cts ? d->num_cts : d->num_ch;
    for (size_t i = 0; i < n; i++) t.ctks.push_back(C(cts[i]));
    return t;
}
void tensor_out(TensorCT &&t, b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *d) {
    if (!outs || capacity < t.ctks.size()) throw EvalError(B200FHE_ERR_ARG, "output capacity too small");
    for (size_t i = 0; i < t.ctks.size(); i++) outs[i] = reinterpret_cast<b200fhe_ct *>(new CtBox{std::move(t.ctks[i])});
    if (d) {
        d->width = t.width;
        d->slotstr = t.slotstr;
        d->num_ch = t.num_ch;
        d->num_cts = t.ctks.size();
    }
}
void merge_indices(std::vector<int32_t> &v, const int32_t *inout, size_t *count) {
    if (!count) throw EvalError(B200FHE_ERR_ARG, "null count");
    if (*count && !inout) throw EvalError(B200FHE_ERR_ARG, "null index list");
    if (*count) v.assign(inout, inout + *count);
}
void store_indices(const std::vector<int32_t> &v, int32_t *inout, size_t capacity, size_t *count) {
    *count = v.size();
    if (inout) {
        if (capacity < v.size()) throw EvalError(B200FHE_ERR_ARG, "capacity too small");
        std::memcpy(inout, v.data(), v.size() * sizeof(int32_t));
    }
}
}  // namespace

extern "C" int b200fhe_dnn_create(b200fhe_eval *ev, b200fhe_encoder *enc, double scale, b200fhe_dnn **out) {
    return guarded([&] {
        if (!enc || !out) throw EvalError(B200FHE_ERR_ARG, "null argument");
        *out = reinterpret_cast<b200fhe_dnn *>(new DNN(E(ev), reinterpret_cast<Encoder *>(enc), scale));
    });
}
extern "C" int b200fhe_dnn_set_relin_key(b200fhe_dnn *d, const uint64_t *const *digits, size_t n) {
    return guarded([&] {
        if (!digits || n == 0) throw EvalError(B200FHE_ERR_ARG, "null evaluation key");
        D(d)->set_relin_key(digits, n);
    });
}
extern "C" int b200fhe_dnn_set_galois_key(b200fhe_dnn *d, uint32_t auto_index, const uint64_t *const *digits, size_t n) {
    return guarded([&] {
        if (!digits || n == 0 || !(auto_index & 1)) throw EvalError(B200FHE_ERR_ARG, "bad Galois key");
        D(d)->set_galois_key(auto_index, digits, n);
    });
}
extern "C" int b200fhe_dnn_set_lanes(b200fhe_dnn *d, b200fhe_boot *bt, size_t n) {
    return guarded([&] { D(d)->set_lanes(*B(bt), n); });
}
extern "C" int b200fhe_dnn_set_batch(b200fhe_dnn *d, size_t n) {
    return guarded([&] { D(d)->set_batch(n); });
}
extern "C" int b200fhe_dnn_set_peer_ranks(b200fhe_dnn *d_, uint32_t world, uint32_t rank, void *const *arena_bases, size_t arena_bytes,
                                          uint32_t *counter) {
    return guarded([&] {
        auto *d = reinterpret_cast<DNN *>(d_);
        if (!d) throw EvalError(B200FHE_ERR_ARG, "null dnn");
        DNN::Ranks r{};
        if (world > 1) {
            if (world > 9 || rank >= world) throw EvalError(B200FHE_ERR_ARG, "rank-sharded layers: 2 .. 9 ranks");
            if (!arena_bases || !counter || arena_bytes < 512 || (arena_bytes - 256) / 2 % 16)
                throw EvalError(B200FHE_ERR_ARG, "rank-sharded layers: arena = 2 halves (multiple of 16 bytes each) + 256 bytes of flags");
            r.world = world;
            r.rank = rank;
            r.half_bytes = (arena_bytes - 256) / 2;
            for (uint32_t g = 0; g < world; g++) {
                if (!arena_bases[g]) throw EvalError(B200FHE_ERR_ARG, "null arena base");
                r.bases[g] = arena_bases[g];
                r.flags[g] = reinterpret_cast<uint64_t *>(static_cast<char *>(arena_bases[g]) + 2 * r.half_bytes);
            }
            r.counter = counter;
        }
        d->ranks = r;
    });
}
extern "C" void b200fhe_dnn_destroy(b200fhe_dnn *d) { delete reinterpret_cast<DNN *>(d); }
extern "C" int b200fhe_dnn_add_rotation_indices(int32_t *inout, size_t capacity, size_t *count, size_t width, size_t kernel_h,
                                                size_t slotstr) {
    return guarded([&] {
        std::vector<int32_t> v;
        merge_indices(v, inout, count);
        DNN::add_rotation_indices(v, width, kernel_h, slotstr);
        store_indices(v, inout, capacity, count);
    });
}
extern "C" int b200fhe_dnn_add_avgpool_rotations(int32_t *inout, size_t capacity, size_t *count, size_t width, size_t slotstr) {
    return guarded([&] {
        std::vector<int32_t> v;
        merge_indices(v, inout, count);
        DNN::add_avgpool_rotations(v, width, slotstr);
        store_indices(v, inout, capacity, count);
    });
}
extern "C" int b200fhe_dnn_enc_tensor(b200fhe_dnn *d, b200fhe_keygen *kg, const uint64_t *pk_dev, const double *image, size_t width,
                                      size_t num_ch, b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *out_desc) {
    return guarded([&] {
        if (!pk_dev || !image) throw EvalError(B200FHE_ERR_ARG, "null argument");
        tensor_out(D(d)->enc_tensor(image, width, num_ch, *K(kg), (const u64 *) pk_dev), outs, capacity, out_desc);
    });
}
extern "C" int b200fhe_dnn_dec_tensor(b200fhe_dnn *d, b200fhe_keygen *kg, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc,
                                      double *out) {
    return guarded([&] {
        if (!out) throw EvalError(B200FHE_ERR_ARG, "null output");
        D(d)->dec_tensor(tensor_in(in, desc), *K(kg), out);
    });
}
extern "C" int b200fhe_dnn_conv(b200fhe_dnn *d, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc, const double *weight,
                                size_t kernel_h, size_t out_ch, size_t stride, b200fhe_ct **outs, size_t capacity,
                                b200fhe_tensor_desc *out_desc) {
    return guarded([&] {
        if (!weight) throw EvalError(B200FHE_ERR_ARG, "null weight");
        tensor_out(D(d)->conv(tensor_in(in, desc), weight, kernel_h, out_ch, stride), outs, capacity, out_desc);
    });
}
extern "C" int b200fhe_dnn_batch_norm(b200fhe_dnn *d, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc,
                                      const double *weight, const double *bias, const double *mean, const double *var,
                                      b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *out_desc) {
    return guarded([&] {
        if (!weight || !bias || !mean || !var) throw EvalError(B200FHE_ERR_ARG, "null parameter");
        tensor_out(D(d)->batch_norm(tensor_in(in, desc), weight, bias, mean, var), outs, capacity, out_desc);
    });
}
extern "C" int b200fhe_dnn_add(b200fhe_dnn *d, const b200fhe_ct *const *a, const b200fhe_tensor_desc *a_desc,
                               const b200fhe_ct *const *b, const b200fhe_tensor_desc *b_desc, b200fhe_ct **outs, size_t capacity,
                               b200fhe_tensor_desc *out_desc) {
    return guarded([&] { tensor_out(D(d)->add(tensor_in(a, a_desc), tensor_in(b, b_desc)), outs, capacity, out_desc); });
}
extern "C" int b200fhe_dnn_relu(b200fhe_dnn *d, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc, int a, int b, int deg,
                                b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *out_desc) {
    return guarded([&] { tensor_out(D(d)->relu(tensor_in(in, desc), a, b, deg), outs, capacity, out_desc); });
}
extern "C" int b200fhe_dnn_relu_deg_two(b200fhe_dnn *d, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc,
                                        b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *out_desc) {
    return guarded([&] { tensor_out(D(d)->relu_deg_two(tensor_in(in, desc)), outs, capacity, out_desc); });
}
extern "C" int b200fhe_dnn_sign(b200fhe_dnn *d, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc, size_t k,
                                b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *out_desc) {
    return guarded([&] { tensor_out(D(d)->sign(tensor_in(in, desc), k), outs, capacity, out_desc); });
}
extern "C" int b200fhe_dnn_relu_composite(b200fhe_dnn *d, b200fhe_boot *bt, const b200fhe_ct *const *in,
                                          const b200fhe_tensor_desc *desc, b200fhe_ct **outs, size_t capacity,
                                          b200fhe_tensor_desc *out_desc) {
    return guarded([&] { tensor_out(D(d)->relu_composite(tensor_in(in, desc), *B(bt)), outs, capacity, out_desc); });
}
extern "C" int b200fhe_dnn_bootstrap(b200fhe_dnn *d, b200fhe_boot *bt, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc,
                                     b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *out_desc) {
    return guarded([&] { tensor_out(D(d)->bootstrap(tensor_in(in, desc), *B(bt)), outs, capacity, out_desc); });
}
extern "C" int b200fhe_dnn_avgpool_fullcon(b200fhe_dnn *d, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc,
                                           const double *weight, size_t T, size_t t, const double *bias, b200fhe_ct **outs,
                                           size_t capacity, b200fhe_tensor_desc *out_desc) {
    return guarded([&] {
        if (!weight || !bias) throw EvalError(B200FHE_ERR_ARG, "null parameter");
        tensor_out(D(d)->avgpool_fullcon(tensor_in(in, desc), weight, T, t, bias), outs, capacity, out_desc);
    });
}
extern "C" int b200fhe_dnn_softmax(b200fhe_dnn *d, b200fhe_boot *bt, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc,
                                   size_t BoundB, size_t BoundR, size_t GumbelLambda, size_t GoldSchmidtD, b200fhe_ct **outs,
                                   size_t capacity, b200fhe_tensor_desc *out_desc) {
    return guarded([&] {
        tensor_out(D(d)->softmax(tensor_in(in, desc), *B(bt), BoundB, BoundR, GumbelLambda, GoldSchmidtD), outs, capacity, out_desc);
    });
}

// ---- ResNet-20 driver (Resnet/models/model_resnet20.cu)
extern "C" size_t b200fhe_resnet20_param_count(size_t in_ch, size_t base_ch, size_t blocks, size_t classes) {
    return resnet20_param_count(in_ch, base_ch, blocks, classes);
}
extern "C" int b200fhe_resnet20_rotation_indices(size_t width, int32_t *out, size_t capacity, size_t *count) {
    return guarded([&] {
        if (!count) throw EvalError(B200FHE_ERR_ARG, "null count");
        store_indices(resnet20_rotation_indices(width), out, capacity, count);
    });
}
extern "C" int b200fhe_resnet20_infer(b200fhe_dnn *d, b200fhe_boot *bt, const b200fhe_ct *const *in, const b200fhe_tensor_desc *desc,
                                      const double *params, size_t nparams, size_t base_ch, size_t blocks, size_t classes,
                                      b200fhe_ct **outs, size_t capacity, b200fhe_tensor_desc *out_desc, double *layer_ms4) {
    return guarded([&] {
        if (!params) throw EvalError(B200FHE_ERR_ARG, "null parameters");
        tensor_out(resnet20_infer(*D(d), *B(bt), tensor_in(in, desc), params, nparams, base_ch, blocks, classes, layer_ms4), outs,
                   capacity, out_desc);
    });
}
