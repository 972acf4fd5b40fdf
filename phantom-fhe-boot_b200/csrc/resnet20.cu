// resnet20.cu — encrypted ResNet-20 inference driver over the DNN layers.
// Replaces ResNet20_infer / PrePareResNet20 (Resnet/models/model_resnet20.cu:3-153): conv1-bn1-relu, three stages of
// `blocks` residual blocks (stages 2 and 3 open with a stride-2 block whose skip path is a 1x1 stride-2 convolution +
// batch norm), average pool + fully connected.  The reference reads its weights from .npy files that are not in the
// tree (Resnet/models/weight_loader.cu); here they arrive as ONE flat array in the order the reference loads them,
// each convolution already in the layers' [kh][kw][in][out] layout (LoadWeight4D's target layout) and each batch norm
// as weight, bias, running_mean, running_var.
#include <chrono>

#include "../../include/b200fhe.h"
#include "dnn.h"

namespace b200fhe {

namespace {
struct Cursor {
    const double *p;
    size_t left;
    const double *take(size_t n) {
        if (n > left) throw EvalError(B200FHE_ERR_ARG, "resnet20: parameter array too short");
        const double *r = p;
        p += n;
        left -= n;
        return r;
    }
};
struct Clock {
    cudaStream_t st;
    double *acc;  // [4]: conv, batch norm + add, relu (sign polynomials + bootstraps), pool + fully connected
    std::chrono::steady_clock::time_point t0;
    void start() {
        if (!acc) return;
        cudaStreamSynchronize(st);
        t0 = std::chrono::steady_clock::now();
    }
    void stop(int which) {
        if (!acc) return;
        cudaStreamSynchronize(st);
        acc[which] += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
};
}  // namespace

size_t resnet20_param_count(size_t in_ch, size_t base_ch, size_t blocks, size_t classes) {
    size_t n = 9 * in_ch * base_ch + 4 * base_ch;
    size_t ch = base_ch;
    for (int stage = 0; stage < 3; stage++) {
        const size_t out = stage == 0 ? ch : 2 * ch;
        for (size_t b = 0; b < blocks; b++) {
            const size_t cin = b == 0 ? ch : out;
            n += 9 * cin * out + 4 * out + 9 * out * out + 4 * out;
            if (stage > 0 && b == 0) n += cin * out + 4 * out;
        }
        ch = out;
    }
    return n + classes * ch + classes;
}

// PrePareResNet20's rotation set (model_resnet20.cu:3-15) for a width x width input
std::vector<int32_t> resnet20_rotation_indices(size_t width) {
    std::vector<int32_t> v;
    DNN::add_rotation_indices(v, width, 3, 0);
    DNN::add_rotation_indices(v, width / 2, 3, 1);
    DNN::add_rotation_indices(v, width / 4, 3, 2);
    DNN::add_avgpool_rotations(v, width / 4, 2);
    return v;
}

TensorCT resnet20_infer(DNN &m, Bootstrapper &bt, const TensorCT &input, const double *params, size_t nparams, size_t base_ch,
                        size_t blocks, size_t classes, double *layer_ms) {
    if (nparams != resnet20_param_count(input.num_ch, base_ch, blocks, classes))
        throw EvalError(B200FHE_ERR_ARG, "resnet20: parameter count does not match the architecture");
    Cursor cur{params, nparams};
    Clock clk{m.ev->st, layer_ms, {}};
    if (layer_ms)
        for (int i = 0; i < 4; i++) layer_ms[i] = 0;
    auto conv = [&](const TensorCT &x, size_t kh, size_t out, size_t stride) {
        const double *w = cur.take(kh * kh * x.num_ch * out);
        clk.start();
        TensorCT y = m.conv(x, w, kh, out, stride);
        clk.stop(0);
        return y;
    };
    auto bn = [&](const TensorCT &x) {
        const size_t c = x.num_ch;
        const double *w = cur.take(c), *b = cur.take(c), *mean = cur.take(c), *var = cur.take(c);
        clk.start();
        TensorCT y = m.batch_norm(x, w, b, mean, var);
        clk.stop(1);
        return y;
    };
    auto relu = [&](const TensorCT &x) {
        clk.start();
        TensorCT y = m.relu_composite(x, bt);
        clk.stop(2);
        return y;
    };
    TensorCT x = relu(bn(conv(input, 3, base_ch, 1)));
    size_t ch = base_ch;
    for (int stage = 0; stage < 3; stage++) {
        const size_t out = stage == 0 ? ch : 2 * ch;
        for (size_t b = 0; b < blocks; b++) {
            const bool down = stage > 0 && b == 0;
            TensorCT skip = x;
            TensorCT y = relu(bn(conv(x, 3, out, down ? 2 : 1)));
            y = bn(conv(y, 3, out, 1));
            if (down) skip = bn(conv(skip, 1, out, 2));
            clk.start();
            y = m.add(y, skip);
            clk.stop(1);
            x = relu(y);
        }
        ch = out;
    }
    const double *fw = cur.take(classes * ch), *fb = cur.take(classes);
    clk.start();
    TensorCT r = m.avgpool_fullcon(x, fw, classes, ch, fb);
    clk.stop(3);
    return r;
}

}  // namespace b200fhe
