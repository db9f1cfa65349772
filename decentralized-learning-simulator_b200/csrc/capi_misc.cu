#include "common.cuh"
namespace dls {
int64_t g_launch_count = 0;

static DeviceInfo g_dev_info[MAX_DEVICES];

int device_info(const DeviceInfo **out, int *dev_out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= MAX_DEVICES) return DLS_E_TOO_BIG;
    DeviceInfo &di = g_dev_info[dev];
    if (di.sms == 0) {
        int sms = 0, smem = 0;
        if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return (int)e;
        if ((e = cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev)) != cudaSuccess) return (int)e;
        di.smem_optin = smem;
        di.sms = sms;
    }
    *out = &di;
    if (dev_out) *dev_out = dev;
    return 0;
}

static int g_tune[TUNE_COUNT] = {};
int tune_value(int key) { return (key >= 0 && key < TUNE_COUNT) ? g_tune[key] : 0; }

int num_sms()
{
    const DeviceInfo *di = nullptr;
    return device_info(&di) == 0 ? di->sms : 0;
}
}  // namespace dls
extern "C" int dls_version(void) { return 100; }
extern "C" const char *dls_build_info(void) {
    return "dls_b200 sm_100a nvcc " __DATE__ " " __TIME__;
}
extern "C" int64_t dls_launch_count(void) { return dls::g_launch_count; }
extern "C" int dls_tune(int32_t key, int32_t value)
{
    if (key < 0 || key >= dls::TUNE_COUNT) return DLS_E_ARG;
    dls::g_tune[key] = value;
    return 0;
}

// ------------------------------------------------------------------------------------------
// gradient_update wave (dasklearn/functions.py:84-86 -> ModelTrainer.gradient_update,
// dasklearn/model_trainer.py:133-143): ONE step of a FRESH torch.optim.SGD with a supplied gradient.
// A fresh optimizer has no momentum buffer, so buf = g (+ wd * p) and p <- p - lr * buf whatever the
// momentum setting; the node's persistent momentum state is not touched.  The reference's
// `p.add_(buf, alpha=-lr)` rounds once (fused multiply-add): fmaf(-lr, buf, p) reproduces its bits.
// ------------------------------------------------------------------------------------------
namespace dls {
__global__ void __launch_bounds__(256)
gradient_update_kernel(float *__restrict__ X, int64_t ldx, const int32_t *__restrict__ p_rows,
                       const int32_t *__restrict__ g_rows, const int32_t *__restrict__ out_rows, int64_t P4,
                       int64_t P, float lr, float wd)
{
    const int r = blockIdx.y;
    const float *p = X + (int64_t)p_rows[r] * ldx;
    const float *g = X + (int64_t)g_rows[r] * ldx;
    float *o = X + (int64_t)out_rows[r] * ldx;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 pv = reinterpret_cast<const float4 *>(p)[i], gv = reinterpret_cast<const float4 *>(g)[i];
        float4 b = gv;
        if (wd != 0.f) { b.x = fmaf(wd, pv.x, gv.x); b.y = fmaf(wd, pv.y, gv.y); b.z = fmaf(wd, pv.z, gv.z); b.w = fmaf(wd, pv.w, gv.w); }
        float4 ov;
        ov.x = fmaf(-lr, b.x, pv.x); ov.y = fmaf(-lr, b.y, pv.y); ov.z = fmaf(-lr, b.z, pv.z); ov.w = fmaf(-lr, b.w, pv.w);
        if (4 * i + 3 < P) {
            reinterpret_cast<float4 *>(o)[i] = ov;
        } else {                                              // ragged tail of the row
            const float t[4] = {ov.x, ov.y, ov.z, ov.w};
            for (int k = 0; k < 4 && 4 * i + k < P; ++k) o[4 * i + k] = t[k];
        }
    }
}
}  // namespace dls

extern "C" int dls_gradient_update(float *X, int64_t ldx, const int32_t *p_rows, const int32_t *g_rows,
                                   const int32_t *out_rows, int32_t n_rows, int64_t P, float lr, float wd,
                                   void *stream)
{
    if (!X || !p_rows || !g_rows || !out_rows || n_rows < 0 || P <= 0 || ldx < P) return DLS_E_ARG;
    if (!dls::aligned16(X) || (ldx & 3)) return DLS_E_ALIGN;      // the kernel reads and writes float4
    if (n_rows == 0) return 0;
    const int64_t P4 = (P + 3) / 4;
    int sms = dls::num_sms();
    if (sms <= 0) sms = 148;
    dim3 grid((unsigned)((P4 + 255) / 256 < sms ? (P4 + 255) / 256 : sms), (unsigned)n_rows);
    dls::gradient_update_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(X, ldx, p_rows, g_rows, out_rows, P4, P, lr, wd);
    DLS_LAUNCH_CHECK();
    return 0;
}
