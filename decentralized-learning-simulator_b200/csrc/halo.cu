// Halo push for the multi-GPU aggregate wave: every rank writes the rows its neighbours on other GPUs need
// straight into THEIR halo rows over NVLink (peer-mapped memory, plain 16-byte stores), one kernel, no staging
// copy and no receive side.  The reference analogue is a pickled nn.Module sent broker to broker over ZeroMQ
// (dasklearn/broker.py:207-219).
//
// Stores to a peer are posted (fire and forget), so a grid sized to the SM count keeps many 16-byte stores in
// flight per SM; loads come from local HBM.  The kernel ends with a system-scope fence; the caller follows it
// with a cross-GPU barrier before any rank reads its halo.
#include "common.cuh"

namespace dls {

// grid = (column blocks, transfers); transfer e copies X[src_rows[e], 0:P) to the row at dst_ptrs[e].
__global__ void __launch_bounds__(256)
push_rows_kernel(const float *__restrict__ X, int64_t ldx, const int32_t *__restrict__ src_rows,
                 const uint64_t *__restrict__ dst_ptrs, int64_t P4)
{
    const int e = blockIdx.y;
    const float4 *src = reinterpret_cast<const float4 *>(X + (int64_t)src_rows[e] * ldx);
    float4 *dst = reinterpret_cast<float4 *>(dst_ptrs[e]);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    // four independent 16-byte loads in flight per thread before the stores are issued
    for (; c + 3 * stride < P4; c += 4 * stride) {
        const float4 a = ldg_stream4(reinterpret_cast<const float *>(src + c));
        const float4 b = ldg_stream4(reinterpret_cast<const float *>(src + c + stride));
        const float4 d = ldg_stream4(reinterpret_cast<const float *>(src + c + 2 * stride));
        const float4 f = ldg_stream4(reinterpret_cast<const float *>(src + c + 3 * stride));
        dst[c] = a; dst[c + stride] = b; dst[c + 2 * stride] = d; dst[c + 3 * stride] = f;
    }
    for (; c < P4; c += stride) dst[c] = ldg_stream4(reinterpret_cast<const float *>(src + c));
    __threadfence_system();
}

// 32-byte accesses (one full sector per lane, 1 KB per warp instruction), four in flight per thread
__global__ void __launch_bounds__(256)
push_rows_v8_kernel(const float *__restrict__ X, int64_t ldx, const int32_t *__restrict__ src_rows,
                    const uint64_t *__restrict__ dst_ptrs, int64_t P8)
{
    const int e = blockIdx.y;
    const float *src = X + (int64_t)src_rows[e] * ldx;
    float *dst = reinterpret_cast<float *>(dst_ptrs[e]);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; c < P8; c += 4 * stride) {
        uint32_t v[4][8];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (c + u * stride < P8)
                asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(v[u][0]), "=r"(v[u][1]), "=r"(v[u][2]), "=r"(v[u][3]), "=r"(v[u][4]), "=r"(v[u][5]),
                               "=r"(v[u][6]), "=r"(v[u][7]) : "l"(src + (c + u * stride) * 8));
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (c + u * stride < P8)
                asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                             :: "l"(dst + (c + u * stride) * 8), "r"(v[u][0]), "r"(v[u][1]), "r"(v[u][2]), "r"(v[u][3]),
                                "r"(v[u][4]), "r"(v[u][5]), "r"(v[u][6]), "r"(v[u][7]) : "memory");
    }
    __threadfence_system();
}

// Persistent form: the work items (transfer e, 32 KB column block b) are numbered e-major and CTA c takes items
// c, c + gridDim.x, ...: at any moment the whole GPU works on a short window of the transfer list, so the ORDER of
// the list decides which peers are written at the same time (the caller sorts it by (destination - rank) mod N so
// that no GPU is the target of all the others at once), and the number of CTAs bounds the stores in flight.
constexpr int PUSH_BLOCK8 = 256 * 4;                      // 32-byte units per work item
__global__ void __launch_bounds__(256)
push_rows_persistent_kernel(const float *__restrict__ X, int64_t ldx, const int32_t *__restrict__ src_rows,
                            const uint64_t *__restrict__ dst_ptrs, int64_t P8, int bx, int64_t n_items)
{
    for (int64_t w = blockIdx.x; w < n_items; w += gridDim.x) {
        const int e = (int)(w / bx), b = (int)(w - (int64_t)e * bx);
        const float *src = X + (int64_t)src_rows[e] * ldx;
        float *dst = reinterpret_cast<float *>(dst_ptrs[e]);
        const int64_t c0 = (int64_t)b * PUSH_BLOCK8 + threadIdx.x;
        uint32_t v[4][8];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (c0 + u * 256 < P8)
                asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(v[u][0]), "=r"(v[u][1]), "=r"(v[u][2]), "=r"(v[u][3]), "=r"(v[u][4]), "=r"(v[u][5]),
                               "=r"(v[u][6]), "=r"(v[u][7]) : "l"(src + (c0 + u * 256) * 8));
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (c0 + u * 256 < P8)
                asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                             :: "l"(dst + (c0 + u * 256) * 8), "r"(v[u][0]), "r"(v[u][1]), "r"(v[u][2]), "r"(v[u][3]),
                                "r"(v[u][4]), "r"(v[u][5]), "r"(v[u][6]), "r"(v[u][7]) : "memory");
    }
    __threadfence_system();
}

}  // namespace dls

extern "C" int dls_push_rows(const float *X, int64_t ldx, const int32_t *src_rows, const uint64_t *dst_ptrs,
                             int32_t n_transfers, int64_t P, void *stream)
{
    if (!X || !src_rows || !dst_ptrs || n_transfers < 0 || P <= 0 || ldx < P) return DLS_E_ARG;
    if (!dls::aligned16(X) || (ldx & 3)) return DLS_E_ALIGN;
    if (n_transfers == 0) return 0;
    if (n_transfers > 65535) return DLS_E_TOO_BIG;
    const int64_t P4 = (P + 3) / 4;                       // rows are padded to a multiple of 4 floats (ld % 4 == 0)
    int sms = dls::num_sms();
    if (sms <= 0) sms = 148;
    // enough CTAs to cover the SMs several times over even for a handful of transfers
    int bx = (int)((P4 + 4 * 256 - 1) / (4 * 256));
    const int per_sm = dls::tune_value(dls::TUNE_PUSH_CTAS) > 0 ? dls::tune_value(dls::TUNE_PUSH_CTAS) : 8;
    const int want = (per_sm * sms + n_transfers - 1) / n_transfers;
    if (bx > want) bx = want;
    if (bx < 1) bx = 1;
    dim3 grid((unsigned)bx, (unsigned)n_transfers);
    // Long rows (FEMNIST: 6.8 MB) are exchanged in column ranges that overlap with the aggregate of the previous range
    // (partition.py): the persistent form with ONE CTA per SM moves the same 700 GB/s per GPU (scripts/push_bw.py) and
    // leaves room on every SM for the aggregate kernel's CTA, where thousands of short-lived CTAs would run first.
    const bool long_rows = P >= (1 << 18) && dls::tune_value(dls::TUNE_PUSH_MODE) == 0;
    if ((long_rows || dls::tune_value(dls::TUNE_PUSH_MODE) == 2) && (ldx & 7) == 0 && ((uintptr_t)X & 31) == 0) {
        const int64_t P8 = (P + 7) / 8;
        const int pbx = (int)((P8 + dls::PUSH_BLOCK8 - 1) / dls::PUSH_BLOCK8);
        const int64_t n_items = (int64_t)pbx * n_transfers;
        int64_t ctas = (int64_t)(long_rows && dls::tune_value(dls::TUNE_PUSH_CTAS) <= 0 ? 1 : per_sm) * sms;
        if (ctas > n_items) ctas = n_items;
        dls::push_rows_persistent_kernel<<<(unsigned)ctas, 256, 0, (cudaStream_t)stream>>>(X, ldx, src_rows, dst_ptrs, P8, pbx,
                                                                                        n_items);
    } else if (dls::tune_value(dls::TUNE_PUSH_MODE) == 1 && (ldx & 7) == 0 && ((uintptr_t)X & 31) == 0)
        dls::push_rows_v8_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(X, ldx, src_rows, dst_ptrs, (P + 7) / 8);
    else
        dls::push_rows_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(X, ldx, src_rows, dst_ptrs, P4);
    DLS_LAUNCH_CHECK();
    return 0;
}
