// dls_eval_wave_noalloc — the test wave for callers that cannot allocate device memory.
//
// Every other wave entry point takes a caller-owned workspace, which assumes the caller has a
// device allocator at hand.  Some users of this library do not: a broker plugin bound through a
// foreign-function layer (ctypes / cgo / JNI stubs that only pass the model, image and result
// buffers through), a monitoring thread that wants an accuracy reading while the training
// workspace is owned by the solver thread, or a latency-sensitive caller that must not trigger
// cudaMalloc's device-wide synchronisation in the middle of a round.  For them this file keeps
// the evaluation scratch (activations of one chunk of samples, re-laid conv weights, per-group
// partial results) in a fixed array that is part of the loaded module.
//
// Cost and limits: the array is NOALLOC_SCRATCH_BYTES (32 MiB) of global memory on EVERY GPU
// that loads the library, held until the process exits.  It bounds a call to at most
// DLS_NOALLOC_MAX_MODELS models and DLS_NOALLOC_MAX_TEST test images; larger requests are
// refused with DLS_E_TOO_BIG (use dls_eval_wave with your own workspace instead).
//
// Sharing: all calls on a device use the same array and it keeps whatever the previous call left
// in it.  The parts that are accumulated into or read before being fully written (the per-group
// result table and the descriptor block) are cleared at the start of every call.  Two calls that
// overlap on the GPU (different streams, or different host threads) race on the array — serialise
// them.
//
// Set-up: the address of the array is resolved once per device, by the first call made while that
// device is current (cudaGetSymbolAddress).  Make that first call outside of any stream capture.
#include "gnlenet.cuh"

#define DLS_NOALLOC_MAX_MODELS 16
#define DLS_NOALLOC_MAX_TEST 4096

namespace dls {

constexpr int64_t NOALLOC_SCRATCH_BYTES = 32ll << 20;
constexpr int NOALLOC_GS = 32;            // samples per group (same grouping as dls_eval_wave)
constexpr int NOALLOC_CHUNK_GROUPS = 8;   // 256 sample slots in flight per pass
constexpr int NOALLOC_MAX_DEVICES = 64;

// In host code `g_noalloc_scratch` is only a handle for the symbol; its device address comes from
// cudaGetSymbolAddress below and is what the kernels receive.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];
static void *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};

static int noalloc_scratch(void **out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return DLS_E_TOO_BIG;
    if (!g_noalloc_ptr[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);   // per device: the module is loaded per context
        if (e != cudaSuccess) return (int)e;
        if (!p) return DLS_E_ARG;
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}

// (model, chunk-of-32-samples) descriptors for groups [g0, g0 + n_groups)
static __global__ void noalloc_desc_kernel(int32_t *gmodel, int32_t *img_idx, int32_t *gcount, int g0,
                                           int n_groups, int gs, int gpm, int n_test)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_groups * gs) return;
    const int gl = i / gs, j = i - gl * gs, g = g0 + gl;
    const int m = g / gpm, chunk = g - m * gpm;
    const int sample = chunk * gs + j;
    img_idx[i] = sample < n_test ? sample : -1;
    if (j == 0) {
        gmodel[gl] = m;
        const int rem = n_test - chunk * gs;
        gcount[gl] = rem < gs ? (rem > 0 ? rem : 0) : gs;
    }
}

// per-model totals, summed in group order (deterministic, fp64)
static __global__ void noalloc_reduce_kernel(const float *glog_all, int gpm, int n_models, int32_t *out_correct,
                                             double *out_loss)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_models) return;
    double ls = 0.0;
    int hs = 0;
    for (int k = 0; k < gpm; ++k) {
        ls += (double)glog_all[((int64_t)m * gpm + k) * 2];
        hs += (int)glog_all[((int64_t)m * gpm + k) * 2 + 1];
    }
    out_loss[m] = ls;
    out_correct[m] = hs;
}

}  // namespace dls

using namespace dls;

extern "C" int dls_eval_wave_noalloc(int32_t arch, const float *X, int64_t ldx, const int32_t *model_rows,
                                     int32_t n_models, const uint8_t *test_images,
                                     const int64_t *test_labels, int32_t n_test, int32_t *out_correct,
                                     double *out_loss_sum, void *stream)
{
    if (arch != DLS_ARCH_GNLENET) return DLS_E_SHAPE;
    if (!X || !model_rows || !test_images || !test_labels || !out_correct || !out_loss_sum || n_models < 0 ||
        n_test < 0)
        return DLS_E_ARG;
    if (!aligned16(X) || ldx % 4) return DLS_E_ALIGN;
    if (ldx < gnl::P) return DLS_E_SHAPE;
    if (n_models > DLS_NOALLOC_MAX_MODELS || n_test > DLS_NOALLOC_MAX_TEST) return DLS_E_TOO_BIG;
    if (n_models == 0 || n_test == 0) return 0;

    const int gpm = (n_test + NOALLOC_GS - 1) / NOALLOC_GS;
    const int64_t total_groups = (int64_t)n_models * gpm;
    const int chunk = (int)(total_groups < NOALLOC_CHUNK_GROUPS ? total_groups : NOALLOC_CHUNK_GROUPS);
    const int64_t ws_bytes = gnl::workspace_bytes(chunk, NOALLOC_GS, n_models, false, false);
    const int64_t glog_bytes = total_groups * 2 * (int64_t)sizeof(float);
    if (ws_bytes + glog_bytes + 256 > NOALLOC_SCRATCH_BYTES) return DLS_E_TOO_BIG;

    void *scratch = nullptr;
    int rc = noalloc_scratch(&scratch);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;

    gnl::Workspace ws;
    const int64_t used = gnl::carve(ws, scratch, chunk, NOALLOC_GS, n_models, false, false);
    float *glog_all = reinterpret_cast<float *>(reinterpret_cast<char *>(scratch) + used);
    // leftovers of an earlier call must not leak into this one: clear what is read-modify-written
    cudaError_t e = cudaMemsetAsync(glog_all, 0, (size_t)glog_bytes, st);
    if (e != cudaSuccess) return (int)e;
    const int64_t desc_count = (int64_t)chunk * NOALLOC_GS + 2 * (int64_t)chunk + 16;
    e = cudaMemsetAsync(ws.desc_i32, 0xff, (size_t)desc_count * sizeof(int32_t), st);   // -1 = empty slot
    if (e != cudaSuccess) return (int)e;

    int32_t *d_img_idx = ws.desc_i32;
    int32_t *d_gmodel = ws.desc_i32 + (int64_t)chunk * NOALLOC_GS;
    int32_t *d_gcount = d_gmodel + chunk;
    gnl::Pass p;
    p.gs = NOALLOC_GS; p.n_models = n_models; p.gmodel = d_gmodel; p.mrow = model_rows;
    p.img_idx = d_img_idx; p.gcount = d_gcount; p.images = test_images; p.labels = test_labels;
    p.X = X; p.ldx = ldx; p.n_groups = chunk;
    if ((rc = gnl::prep_weights(p, ws, false, st))) return rc;
    for (int64_t g0 = 0; g0 < total_groups; g0 += chunk) {
        const int ng = (int)((total_groups - g0) < chunk ? (total_groups - g0) : chunk);
        p.n_groups = ng;
        noalloc_desc_kernel<<<(ng * NOALLOC_GS + 255) / 256, 256, 0, st>>>(d_gmodel, d_img_idx, d_gcount, (int)g0,
                                                                            ng, NOALLOC_GS, gpm, n_test);
        DLS_LAUNCH_CHECK();
        ws.glog = glog_all + g0 * 2;
        if ((rc = gnl::forward(p, ws, false, st))) return rc;
        if ((rc = gnl::classifier(p, ws, 1, nullptr, nullptr, 0, st))) return rc;
    }
    noalloc_reduce_kernel<<<(n_models + 127) / 128, 128, 0, st>>>(glog_all, gpm, n_models, out_correct,
                                                                  out_loss_sum);
    DLS_LAUNCH_CHECK();
    return 0;
}
