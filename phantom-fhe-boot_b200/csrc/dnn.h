// dnn.h — encrypted CNN layers over the evaluator and the bootstrapper.
// Replaces phantom::DNN / TensorCT (include/dnn.cuh:33-80, src/dnn.cu): Conv, BatchNorm, Add, Relu, ReluDegTwo, Sign,
// ReluComposite, BootStrap, AvgPoolFullCon, SoftMax, the rotation-index helpers, EncTensor / DecTensor.
//
// A tensor is one ciphertext per channel; channel k holds a width x width image on a (width << slotstr)^2 grid of
// sparse slots, pixel (i, j) in slot ((i * L + j) << slotstr), L = width << slotstr.  Every layer produces the
// reference's ciphertext limbs bit for bit; what differs is the schedule:
//   * Conv: the reference key-switches input channel k once per (output channel, kernel tap) — out_ch * in_ch * kh^2
//     full key switches.  A rotated input does not depend on the output channel, so here channel k is rotated kh^2
//     times ONCE (modup shared by the taps, one batched moddown), its level adjustment (the rescale EvalMultAutoInplace
//     applies to a NoiseScaleDeg-2 input) is done once, and each output channel is ONE plaintext-matrix product
//     sum_{k,tap} rot[k][tap] (.) pt[h][k][tap] with 128-bit accumulation (bsgs_matmul) instead of in_ch * kh^2
//     multiply passes + add passes.  The plaintexts of an output channel are encoded into one block and brought to the
//     product's level as one batched ciphertext.
//   * element-wise layers (BatchNorm, Add, the affine parts of AvgPoolFullCon) stay pending affine combinations.
#pragma once
#include <functional>

#include "bootstrapper.h"
#include "encoder.h"
#include "keygen.h"

namespace b200fhe {

struct TensorCT {  // include/dnn.cuh:33-39
    std::vector<Ct> ctks;
    size_t width = 0, slotstr = 0, num_ch = 0;
};

// EvalChebyshevCoefficients (src/evaluate.cu:3585-3609), same summation order
std::vector<double> chebyshev_coefficients(const std::function<double(double)> &f, double a, double b, uint32_t degree);

class DNN {
  public:
    DNN(Evaluator *ev, Encoder *enc, double scale);
    Evaluator *ev;
    Encoder *enc;
    double scale;
    // The reference's DNN owns a relinearisation key (RelinKeyGen -> mul_key_) and a rotation-key set (BuildGaloisKey ->
    // gk_) that are DIFFERENT key material from the bootstrapper's (FHECKKSRNS::mul_key / galois_keys), although both
    // work on one context.  Keys set here are used by the layers' own multiplications and rotations; the bootstraps
    // inside ReluComposite / BootStrap keep using the evaluator's keys.  Unset: the evaluator's keys serve both.
    bool have_relin = false;
    std::vector<const uint64_t *> relin;
    std::map<uint32_t, std::vector<const uint64_t *>> galois;

    // AddRotationIndicesTo / AddAvgPoolRotationsTo (src/dnn.cu:283-330): append the indices not yet present
    static void add_rotation_indices(std::vector<int32_t> &v, size_t width, size_t kernel_h, size_t slotstr);
    static void add_avgpool_rotations(std::vector<int32_t> &v, size_t width, size_t slotstr);

    // image: [width][width][num_ch] row-major; one fresh ciphertext per channel at the top level (EncTensor)
    TensorCT enc_tensor(const double *image, size_t width, size_t num_ch, KeyGen &kg, const u64 *pk);
    // out: [width][width][num_ch] (DecTensor): decrypt, decode, first width^2 slots of the period
    void dec_tensor(const TensorCT &t, KeyGen &kg, double *out);

    // weight: [kh][kh][in_ch][out_ch] row-major, indexed weight[i][j][k][h] exactly as the reference does
    TensorCT conv(const TensorCT &in, const double *weight, size_t kh, size_t out_ch, size_t stride);
    TensorCT batch_norm(const TensorCT &in, const double *weight, const double *bias, const double *mean, const double *var);
    TensorCT add(const TensorCT &a, const TensorCT &b);
    TensorCT relu(const TensorCT &in, int a, int b, int deg);
    TensorCT relu_deg_two(const TensorCT &in);
    TensorCT sign(const TensorCT &in, size_t k);
    TensorCT relu_composite(const TensorCT &in, Bootstrapper &bt);
    TensorCT bootstrap(const TensorCT &in, Bootstrapper &bt);
    // weight: [T][t] row-major (T outputs, t = in.num_ch inputs), bias [T]
    TensorCT avgpool_fullcon(const TensorCT &in, const double *weight, size_t T, size_t t, const double *bias);
    TensorCT softmax(const TensorCT &in, Bootstrapper &bt, size_t BoundB, size_t BoundR, size_t GumbelLambda, size_t GoldSchmidtD);

    // ---- lanes: independent channels of ReluComposite / BootStrap run concurrently on `n` extra streams.
    // One sparse bootstrap fills a fraction of the B200 (its launches are partial waves of 148 SMs), and a layer has
    // 16-64 independent channels.  A lane is a private Context (workspace), Evaluator (pool), Bootstrapper and stream
    // that shares the key and diagonal MEMORY of `bt` (pointers only); channel i runs on lane i % n.  The arithmetic
    // of a channel is the same whichever lane runs it, so the limbs do not change.  n == 0 switches lanes off.
    void set_lanes(Bootstrapper &bt, size_t n);
    size_t lane_count() const { return lanes_.size(); }
    // ---- channel batches: up to `n` channels with identical bookkeeping (the channels of a layer) go through ReluComposite /
    // BootStrap as ONE batched ciphertext (Ct::batch), so every key switch, rescale and polynomial step is one launch set over
    // all of them; with lanes, each lane takes batches of its channels.  Same limbs per channel.  n <= 1: one channel at a time.
    void set_batch(size_t n) { batch_ = n < 1 ? 1 : n; }
    size_t batch() const { return batch_; }
    // ---- ranks: ONE image on several GPUs (no counterpart in the reference).  Every rank runs the network on the same
    // encrypted input with the same keys; in ReluComposite / BootStrap (95 % of a ResNet-20) rank r takes the channels
    // [n r / G, n (r + 1) / G) of the layer -- through its lanes and batches as usual -- and the result ciphertexts are
    // all-gathered over NVLink peer memory (peer_exchange_strided: one store kernel that doubles as the barrier) into a
    // symmetric arena, two halves used by alternate layers.  The channels of a layer have identical bookkeeping, so a rank
    // knows level, scale and NoiseScaleDeg of the ciphertexts that arrive from its own results.  Layers with fewer channels
    // than ranks, or channels whose bookkeeping differs, run replicated.  Conv: rank r takes a run of INPUT channels
    // (rotations, plaintext encodes and products shrink by the number of ranks) and the partial sums of every output channel
    // are reduced across the ranks (all-to-all into per-sender slots, row-sum kernel, all-gather).  The limbs are the
    // single-GPU limbs.
    struct Ranks {
        u32 world = 1, rank = 0;
        void *bases[kMaxRanks] = {};
        uint64_t *flags[kMaxRanks] = {};
        uint32_t *counter = nullptr;
        size_t half_bytes = 0;
        u32 parity = 0;
    } ranks;
    ~DNN();
    DNN(const DNN &) = delete;
    DNN &operator=(const DNN &) = delete;
    Ct relu_composite_one(const Ct &in, Bootstrapper &bt);  // one channel of ReluComposite on this object's evaluator

  private:
    struct Lane;
    std::vector<std::unique_ptr<Lane>> lanes_;
    size_t batch_ = 4;
    // f on the channels idx of `in`, in batches of at most batch_, on (d, b)'s evaluator; `src` != null: the channel data is
    // copied in from these pointers (lanes: the channels live in another evaluator's pool); results into ys[idx[k]]
    static void run_batched(DNN &d, Bootstrapper &b, const TensorCT &in, const std::vector<size_t> &idx, const u64 *const *src,
                            size_t batch, const std::function<Ct(DNN &, Bootstrapper &, const Ct &)> &f, std::vector<Ct> &ys);
    // f(lane DNN, lane bootstrapper, channel copy) -> result on the lane; results are copied into this evaluator's pool
    TensorCT run_on_lanes(const TensorCT &in, const std::function<Ct(DNN &, Bootstrapper &, const Ct &)> &f);
    TensorCT run_here(const TensorCT &in, Bootstrapper &bt, const std::function<Ct(DNN &, Bootstrapper &, const Ct &)> &f);
    TensorCT run_local(const TensorCT &in, Bootstrapper &bt, const std::function<Ct(DNN &, Bootstrapper &, const Ct &)> &f);
    TensorCT run_on_ranks(const TensorCT &in, Bootstrapper &bt, const std::function<Ct(DNN &, Bootstrapper &, const Ct &)> &f);
    struct KeyScope {  // the layers' own keys on the evaluator for the lifetime of the object
        DNN &d;
        explicit KeyScope(DNN &dnn) : d(dnn) { swap(); }
        ~KeyScope() { swap(); }
        void swap() {
            if (d.have_relin) {
                std::swap(d.ev->relin, d.relin);
                std::swap(d.ev->have_relin, d.have_relin_saved_);
            }
            if (d.have_galois_) std::swap(d.ev->galois, d.galois);
        }
    };
    bool have_relin_saved_ = true;  // what ev->have_relin becomes while the scope is active
    bool have_galois_ = false;

  public:
    void set_relin_key(const uint64_t *const *digits, size_t n) {
        relin.assign(digits, digits + n);
        have_relin = true;
    }
    void set_galois_key(uint32_t auto_index, const uint64_t *const *digits, size_t n) {
        galois[auto_index].assign(digits, digits + n);
        have_galois_ = true;
    }

  private:
    static TensorCT like(const TensorCT &in) {
        TensorCT r;
        r.width = in.width;
        r.slotstr = in.slotstr;
        r.num_ch = in.num_ch;
        return r;
    }
};

// Resnet/models/model_resnet20.cu: PrePareResNet20's rotation set and ResNet20_infer (resnet20.cu)
size_t resnet20_param_count(size_t in_ch, size_t base_ch, size_t blocks, size_t classes);
std::vector<int32_t> resnet20_rotation_indices(size_t width);
TensorCT resnet20_infer(DNN &m, Bootstrapper &bt, const TensorCT &input, const double *params, size_t nparams, size_t base_ch,
                        size_t blocks, size_t classes, double *layer_ms);

}  // namespace b200fhe
