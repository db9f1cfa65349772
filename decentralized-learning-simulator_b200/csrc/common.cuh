// Shared helpers for the dls_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/dls_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "dls_b200 kernels are written for sm_100a (Blackwell B200) only"
#endif

namespace dls {

extern int64_t g_launch_count;   // bumped by every kernel launch this library makes

inline int check_last(cudaError_t e) { return e == cudaSuccess ? 0 : (int)e; }

#define DLS_LAUNCH_CHECK()                                   \
    do {                                                     \
        ::dls::g_launch_count++;                             \
        cudaError_t e__ = cudaGetLastError();                \
        if (e__ != cudaSuccess) return (int)e__;             \
    } while (0)

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---- per-device state (a C-ABI caller may drive several GPUs from one process) ----
constexpr int MAX_DEVICES = 64;
struct DeviceInfo { int sms = 0; int smem_optin = 0; };
// Properties of the CURRENT device, queried once per device (capi_misc.cu).
int device_info(const DeviceInfo **out, int *dev_out = nullptr);
// Number of SMs of the current device (148 on a B200); 0 if the query fails.
int num_sms();

// Experiment knobs set through dls_tune() (diagnostic entry point; 0 = automatic choice).
enum TuneKey { TUNE_AGG_TW4 = 0, TUNE_AGG_STAGES = 1, TUNE_LANES = 2, TUNE_CONV_UNGROUPED = 3, TUNE_DEBUG = 4, TUNE_NO_SGD_FUSE = 5, TUNE_PUSH_CTAS = 6, TUNE_PUSH_MODE = 7, TUNE_COUNT = 8 };
int tune_value(int key);

// One-time, per-device opt-in to `bytes` of dynamic shared memory for `kernel`
// (cudaFuncSetAttribute applies to the device that is current when it is called).
struct PerDeviceOnce { bool done[MAX_DEVICES] = {}; };
template <typename K>
inline int ensure_dyn_smem(K kernel, int bytes, PerDeviceOnce &once) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= MAX_DEVICES) return DLS_E_TOO_BIG;
    if (!once.done[dev]) {
        e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        if (e != cudaSuccess) return (int)e;
        once.done[dev] = true;
    }
    return 0;
}

__device__ __forceinline__ float4 ldg_stream4(const float *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide sum with a fixed reduction tree (deterministic); `red` has >= 32 floats.
__device__ __forceinline__ float block_sum(float v, float *red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float t = (threadIdx.x < nw) ? red[threadIdx.x] : 0.f;
    if (warp == 0) {
        t = warp_sum(t);
        if (lane == 0) red[0] = t;
    }
    __syncthreads();
    t = red[0];
    return t;
}

// Two block-wide sums with one pair of barriers (fixed reduction tree; `red` has >= 64 floats, blockDim <= 1024).
__device__ __forceinline__ void block_sum2(float &a, float &b, float *red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    a = warp_sum(a);
    b = warp_sum(b);
    __syncthreads();
    if (lane == 0) { red[warp * 2] = a; red[warp * 2 + 1] = b; }
    __syncthreads();
    float sa = 0.f, sb = 0.f;
    for (int w = 0; w < nw; ++w) { sa += red[w * 2]; sb += red[w * 2 + 1]; }
    a = sa;
    b = sb;
}

}  // namespace dls
