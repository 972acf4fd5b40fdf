// bootstrapper.h — CKKS bootstrapping (FHECKKSRNS, src/bootstrap.cu / include/bootstrap.cuh of the reference) over
// the Evaluator: ModRaise -> CoeffsToSlots -> EvalMod (Chebyshev + double-angle) -> SlotsToCoeffs.
#pragma once
#include "evaluator.h"

namespace b200fhe {
class Encoder;
}

namespace b200fhe {

// GetCollapsedFFTParams (src/util.cu:768-817), CKKS_BOOT_PARAMS order
struct FFTParams {
    int32_t levelBudget, layersCollapse, remCollapse, numRotations, b, g, numRotationsRem, bRem, gRem;
};
FFTParams collapsed_fft_params(uint32_t slots, uint32_t levelBudget, uint32_t dim1);

// one pre-encoded plaintext diagonal in the extended basis Q_l P ([QlP][N], NTT form), caller-owned device memory
struct Diagonal {
    const u64 *p = nullptr;
    size_t Ql = 0;
    double scale = 0;
};

// The linear-transform stages of ONE bootstrap spread over `world` GPUs by the limbs of Q_l P (no counterpart in the
// reference).  Every rank runs the whole bootstrap on the same input (same keys, same diagonals, same bits); in a stage
// the two HBM-bound kernels -- all hoisted giant-step rotations and the plaintext-matrix product, 27 % of a bootstrap on
// one GPU -- compute only the rank's own limb rows (both are limb-wise: a row needs the same row of the digits, the key
// and the diagonals, nothing else), and the rows of the b inner sums are exchanged once per stage over NVLink peer memory
// (peer_exchange_strided: one store kernel that doubles as the barrier).  The inner sums live in a symmetric arena,
// two halves used alternately: a peer's rows of stage s + 1 arrive in the other half while this rank may still be
// reading stage s, and nobody reaches stage s + 2 before everybody has passed the exchange of stage s + 1.
// Between the stages, the approximate modular reduction of the real and the imaginary half (one batch-2 ciphertext on one
// GPU) is split between partner ranks r and r ^ 1: even ranks evaluate the real half, odd ranks the imaginary one, and the
// partners swap results (peer_all_to_all); a last rank without partner keeps the batch.
struct PeerShard {
    u32 world = 1, rank = 0;
    bool virtual_ranks = false;   // one GPU computes every rank's rows in turn, no arena, no exchange (tests the range plumbing)
    void *bases[kMaxRanks] = {};          // the arena of every rank, mapped into this process (entry [rank] is the own one)
    uint64_t *flags[kMaxRanks] = {};      // flag rows (uint64_t[world]) of every rank
    uint32_t *counter = nullptr;  // uint32_t[4], zero-initialised, own device memory
    size_t half_bytes = 0;        // size of one half of the arena = sums_bytes + slots_bytes + the summed rows
    size_t sums_bytes = 0, slots_bytes = 0;
    size_t evalmod_off = 0;       // after the halves: where the partner rank stores its half of EvalMod ([2][Q][N])
    u32 parity = 0;
};

class Bootstrapper {
  public:
    Bootstrapper(Context *ctx, cudaStream_t stream, uint32_t budgetEnc, uint32_t budgetDec, uint32_t dim1Enc,
                 uint32_t dim1Dec, uint32_t numSlots, uint32_t correctionFactor);
    Evaluator ev;
    uint32_t slots, correction;
    uint32_t ctor_args[6];  // budgetEnc, budgetDec, dim1Enc, dim1Dec, numSlots, correctionFactor as given (for clones)
    FFTParams enc, dec;
    std::vector<std::vector<Diagonal>> A_enc, A_dec;  // [stage][index]  m_U0hatTPreFFT / m_U0PreFFT

    std::vector<int32_t> rotation_indices() const;      // FindBootstrapRotationIndices
    size_t stage_count(int which) const { return which == 0 ? A_enc.size() : A_dec.size(); }
    size_t stage_size(int which, size_t stage) const;   // number of diagonals of a stage
    size_t stage_level(int which, size_t stage) const;  // |Q_l| the stage's plaintexts are encoded at
    void set_diagonal(int which, size_t stage, size_t idx, const u64 *dev, size_t Ql, double scale);
    // EvalBootstrapSetup(precompute = true): compute and encode every diagonal into device memory owned by this object
    // (bootsetup.cu); an alternative to passing pre-encoded diagonals in with set_diagonal
    void precompute(Encoder &encoder);

    Ct eval(const Ct &in);                // EvalBootstrap (numIterations == 1)
    Ct eval(const Ct &in, uint32_t numIterations, uint32_t precision);  // EvalBootstrap, Meta-BTS iterations
    Ct coeffs_to_slots(const Ct &in);     // EvalCoeffsToSlots
    Ct slots_to_coeffs(const Ct &in);     // EvalSlotsToCoeffs
    void double_angle(Ct &x, uint32_t numIt);  // ApplyDoubleAngleIterations
    void adjust_ciphertext(Ct &x, double correction);

    PeerShard shard;
    size_t peer_sums_bytes() const, peer_slots_bytes() const, peer_total_bytes() const;
    size_t peer_arena_half_bytes() const;
    size_t peer_arena_bytes() const;       // two halves, the EvalMod swap block, the flag rows

  private:
    std::vector<std::unique_ptr<DevBuf>> owned_;  // diagonals produced by precompute(), one block per stage
    // one collapsed-FFT level: hoisted giant-step rotations, plaintext-matrix product, baby-step rotations
    Ct bsgs_stage(const Ct &in, const std::vector<Diagonal> &A, const std::vector<int32_t> &rot_in,
                  const std::vector<int32_t> &rot_out, int32_t g, int32_t b, int32_t numRot);
};

}  // namespace b200fhe
