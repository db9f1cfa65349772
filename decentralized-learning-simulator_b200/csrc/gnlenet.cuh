// GNLeNet (reference: dasklearn/models/cifar10.py:103-142) — constants, workspace carving and the
// kernel-launch interface shared by the train / grad / eval entry points.
#pragma once
#include "common.cuh"

namespace dls {
namespace gnl {

constexpr int P = DLS_GNLENET_P;
// flat parameter offsets in parameters() order
constexpr int OFF_W1 = 0, OFF_B1 = 2400, OFF_G1W = 2432, OFF_G1B = 2464;
constexpr int OFF_W2 = 2496, OFF_B2 = 28096, OFF_G2W = 28128, OFF_G2B = 28160;
constexpr int OFF_W3 = 28192, OFF_B3 = 79392, OFF_G3W = 79456, OFF_G3B = 79520;
constexpr int OFF_FCW = 79584, OFF_FCB = 85344;

constexpr int IMG = 32, IMG_C = 3, IMG_BYTES = IMG * IMG * IMG_C;
constexpr int C1 = 32, H1 = 32, PH1 = 15;     // conv1 out 32x32x32 -> pool -> 15x15x32
constexpr int C2 = 32, H2 = 15, PH2 = 7;      // conv2 out 15x15x32 -> pool -> 7x7x32
constexpr int C3 = 64, H3 = 7, PH3 = 3;       // conv3 out 7x7x64   -> pool -> 3x3x64
constexpr int NCLS = 10, FEAT = 576;
constexpr float GN_EPS = 1e-5f;

constexpr int A1_SZ = PH1 * PH1 * C1;   // 7200
constexpr int C2_SZ = H2 * H2 * C2;     // 7200
constexpr int A2_SZ = PH2 * PH2 * C2;   // 1568
constexpr int C3_SZ = H3 * H3 * C3;     // 3136
constexpr int A3_SZ = PH3 * PH3 * C3;   // 576
constexpr int GNP_SZ = 3 * (C1 + C2 + C3);   // per-slot partials: GN gamma, GN beta, conv bias gradient (384)

// bf16 plane runs (tensor-core path): RUN = (H-1)*(H+2) + H pixels, padded to a multiple of 8
constexpr int RUN15 = 14 * 17 + 15, RUNP15 = 256;     // 15x15 maps (a1, dc2): 253 pixels, plane = 4096 B
constexpr int RUN7 = 6 * 9 + 7, RUNP7 = 64;           // 7x7 maps (a2, dc3): 61 pixels, plane = 1024 B
constexpr int RUN14 = 13 * 16 + 14, RUNP14 = 224;     // 14x14 maps (FEMNIST a1, dc2): 222 pixels
__host__ __device__ constexpr int plane_runp(int h) { return h == 15 ? RUNP15 : (h == 14 ? RUNP14 : RUNP7); }
constexpr int A1P_SZ = 4 * RUNP15 * 8, A2P_SZ = 4 * RUNP7 * 8, DC2P_SZ = 4 * RUNP15 * 8, DC3P_SZ = 8 * RUNP7 * 8;   // bf16 elements per sample

constexpr int W2T_SZ = 25 * 32 * 32, W3T_SZ = 25 * 32 * 64;
constexpr int TC_PREP_ELEMS = 2 * 25600 + 2 * 51200 + 2560;   // bf16 UMMA-layout weight copies per model (tc_conv.cu)
constexpr int PREP_SZ = 2 * W2T_SZ + 2 * W3T_SZ;   // W2t, W2d, W3t, W3d per model

// Optimizer update fused into the conv2 / conv3 weight-gradient epilogues (train waves, tensor-core path): the
// gradient of those 76 800 weights never goes to memory; torch.optim.SGD semantics as in sgd_update_kernel.
struct SgdFuse {
    float *X = nullptr;                  // stacked parameters (nullptr: not fused, gradients go to G)
    int64_t ldx = 0;
    float *M = nullptr;                  // momentum rows (nullptr when mu == 0)
    int64_t ldm = 0;
    const int32_t *task_row = nullptr;   // [n_groups] row of X
    const int32_t *task_peer = nullptr;  // [n_groups] row of M
    const uint8_t *first = nullptr;      // [n_groups] 1: the node has no momentum buffer yet (nullptr: none is first)
    float lr = 0.f, mu = 0.f, wd = 0.f;
};

// All activations are NHWC fp32, one contiguous record per sample slot.
struct Workspace {
    // forward (kept for backward)
    float *p1, *a1, *c2, *a2, *c3, *a3;
    float *st1, *st2, *st3;              // [slot][4] = mean0, rstd0, mean1, rstd1
    uint8_t *idx1, *idx2, *idx3;         // arg-max position inside the 3x3 window (0..8)
    // backward
    float *da3, *dc3, *da2, *dc2, *da1, *dp1, *gnp;
    float *G;                            // [n_groups][ldg] gradient rows (train) — may point to user memory
    int64_t ldg;
    float *prep;                         // [n_models][PREP_SZ]
    void *prep_tc;                       // [n_models][TC_PREP_ELEMS] bf16, only when precision == 1
    // precision 1: bf16 channel-chunk planes, the layout the tensor-core convolutions consume, written by the
    // producing kernels and staged with bulk (TMA) copies.  Per sample: [C/8 planes][RUNP pixels][8 channels];
    // pixel index r = y * (H + 2) + x (the padded grid's run of H rows; the two gap pixels after a row are zero).
    uint16_t *a1p, *a2p, *dc2p, *dc3p, *zplane;
    SgdFuse sgd;                         // set by dls_train_wave for the duration of a step
    int precision;                       // 0 = fp32 CUDA cores, 1 = bf16 tcgen05 convolutions
    float *glog;                         // [n_groups][2] loss sum, correct count (eval) / loss (train)
    int32_t *desc_i32;                   // scratch for generated descriptors (eval)
};

struct Pass {
    int n_groups, gs, n_models;
    const int32_t *gmodel;     // [n_groups] model index, nullptr = identity
    const int32_t *mrow;       // [n_models] row of X
    const int32_t *img_idx;    // [n_groups*gs] global image id, -1 = empty slot
    const int32_t *gcount;     // [n_groups] active slots per group (mean-loss divisor)
    const uint8_t *images;
    const int64_t *labels;
    const float *X;
    int64_t ldx;
};

__device__ __forceinline__ const float *model_row(const Pass &p, int g, int &m) {
    m = p.gmodel ? p.gmodel[g] : g;
    return p.X + (int64_t)p.mrow[m] * p.ldx;
}

int64_t workspace_bytes(int n_groups, int gs, int n_models, bool backward, bool own_G, int precision = 0);
// Carves `base`; returns bytes used.
int64_t carve(Workspace &ws, void *base, int n_groups, int gs, int n_models, bool backward, bool own_G,
              int precision = 0);

// tensor-core (tcgen05) convolutions, tc_conv.cu
int prep_weights_tc(const Pass &p, Workspace &ws, bool backward, cudaStream_t st);
int conv2_fwd_tc(const Pass &p, Workspace &ws, cudaStream_t st);
int conv3_fwd_tc(const Pass &p, Workspace &ws, cudaStream_t st);
int conv2_dgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st);
int conv3_dgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st);
int conv2_wgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st);
int prep_w1_tc(const Pass &p, Workspace &ws, cudaStream_t st);
int conv1_fwd_tc(const Pass &p, Workspace &ws, bool save, cudaStream_t st);
int conv1_wgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st);
int conv3_wgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st);

// FEMNIST conv2 (32 -> 64 channels, 14x14 maps) on the same kernels (tc_conv.cu).  `prep_tc`: [n_models][TC_PREP_ELEMS]
// bf16 (only the 32->64 slots are used); planes: [slot][C/8][RUNP14][8] bf16; c2 / da1: fp32 NHWC, no bias added.
int fem_prep_conv2_tc(const Pass &p, void *prep_tc, int w_off, bool backward, cudaStream_t st);
int fem_conv2_fwd_tc(const Pass &p, const uint16_t *a1p, float *c2, const void *prep_tc, cudaStream_t st);
int fem_conv2_dgrad_tc(const Pass &p, const uint16_t *dc2p, float *da1, const void *prep_tc, cudaStream_t st);
int fem_conv2_wgrad_tc(const Pass &p, const uint16_t *a1p, const uint16_t *dc2p, const uint16_t *zeros, float *G,
                       int64_t ldg, int w_off, cudaStream_t st);

int prep_weights(const Pass &p, Workspace &ws, bool backward, cudaStream_t st);
int forward(const Pass &p, Workspace &ws, bool save_for_backward, cudaStream_t st);
// mode 0: train (writes dlogits-derived grads: needs backward buffers); mode 1: eval
int classifier(const Pass &p, Workspace &ws, int mode, float *loss_out, float *out_logits,
               int logits_stride_model, cudaStream_t st);
// `lane` selects the side stream the weight-gradient kernels run on (two half-waves may be in flight)
int backward(const Pass &p, Workspace &ws, cudaStream_t st, int lane = 0);

// Streams owned by the library (created once per device): which = 2*lane is the weight-gradient side stream
// of a lane, which = 2*lane + 1 the main stream of lanes >= 1 of a split train wave (lane 0 runs on the caller's).
constexpr int MAX_LANES = 8;
struct SideStream { cudaStream_t s = nullptr; cudaEvent_t fork = nullptr, join = nullptr; };
int side_stream(SideStream **out, int which);

}  // namespace gnl
}  // namespace dls
