// reconstruct_from_chunks wave (reference: dasklearn/functions.py:142-146 -> ChunkManager.reconstruct_model,
// dasklearn/simulation/conflux/chunk_manager.py:34-53): per chunk index the mean over a VARIABLE set of
// contributor models of the same column range, `torch.mean(torch.stack(pieces), dim=0)`.
//
// Bit-exactness needs torch's CPU summation order (aten/src/ATen/native/cpu/SumKernel.cpp, AVX2 dispatch,
// vector width 8).  The stack [k, n] is reduced over its OUTER dimension:
//   * columns below the last multiple of 32 of the piece (vectorized_outer_sum -> multi_row_sum): the
//     contributors are added one after the other starting from 0, with a four-level cascade every 16 rows;
//   * the remaining (< 32) columns (row_sum): four interleaved partial sums over contributors 4i+q, each
//     with the same cascade, remainder contributors into partial 0, then ((p0 + p1) + p2) + p3;
//   * pieces shorter than 8 columns (scalar_outer_sum): the same split at the last multiple of 4.
// The sum is then divided (not multiplied by a reciprocal) by float(k).
// HBM-bound: every contributor element is read once, every output element written once.
#include "common.cuh"

namespace dls {

// Cascade sum of x[0], x[stride], ... (count terms): multi_row_sum for one column.
__device__ __forceinline__ float cascade_sum(const float *const *rows, int first, int stride, int count, int col)
{
    // level_power = max(4, ceil_log2(count) / 4): 4 for every count < 2^20
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    int i = 0;
    while (i + 16 <= count) {
#pragma unroll 4
        for (int j = 0; j < 16; ++j, ++i) a0 = __fadd_rn(a0, __ldg(rows[first + i * stride] + col));
        a1 = __fadd_rn(a1, a0); a0 = 0.f;
        if ((i & (15 << 4)) == 0) {
            a2 = __fadd_rn(a2, a1); a1 = 0.f;
            if ((i & (15 << 8)) == 0) { a3 = __fadd_rn(a3, a2); a2 = 0.f; }
        }
    }
    for (; i < count; ++i) a0 = __fadd_rn(a0, __ldg(rows[first + i * stride] + col));
    a0 = __fadd_rn(a0, a1);
    a0 = __fadd_rn(a0, a2);
    a0 = __fadd_rn(a0, a3);
    return a0;
}

constexpr int CHUNK_MAX_SRC = 1024;       // contributors of one segment kept in shared memory

// grid = (column tiles, segments); block = 256 threads, 4 columns per thread (blockDim apart: coalesced).
__global__ void __launch_bounds__(256)
chunk_mean_kernel(const float *__restrict__ X, int64_t ldx, float *__restrict__ Y, int64_t ldy,
                  const int32_t *__restrict__ seg_ptr, const int32_t *__restrict__ seg_src,
                  const int32_t *__restrict__ seg_out, const int64_t *__restrict__ seg_col0,
                  const int32_t *__restrict__ seg_len)
{
    __shared__ const float *rows[CHUNK_MAX_SRC];
    const int s = blockIdx.y;
    const int e0 = seg_ptr[s], k = seg_ptr[s + 1] - e0;
    const int n = seg_len[s];
    const int64_t c0 = seg_col0[s];
    for (int i = threadIdx.x; i < k; i += blockDim.x) rows[i] = X + (int64_t)seg_src[e0 + i] * ldx + c0;
    __syncthreads();
    const int nb = n >= 8 ? (n & ~31) : (n & ~3);
    float *out = Y + (int64_t)seg_out[s] * ldy + c0;
    const float kf = (float)k;
    const int base = blockIdx.x * 1024 + threadIdx.x;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int c = base + u * 256;
        if (c >= n) continue;
        float sum;
        if (c < nb) {
            sum = cascade_sum(rows, 0, 1, k, c);
        } else {
            const int g = k >> 2;
            float p0 = cascade_sum(rows, 0, 4, g, c);
            const float p1 = cascade_sum(rows, 1, 4, g, c);
            const float p2 = cascade_sum(rows, 2, 4, g, c);
            const float p3 = cascade_sum(rows, 3, 4, g, c);
            for (int i = 4 * g; i < k; ++i) p0 = __fadd_rn(p0, __ldg(rows[i] + c));
            sum = __fadd_rn(__fadd_rn(__fadd_rn(p0, p1), p2), p3);
        }
        out[c] = __fdiv_rn(sum, kf);
    }
}

}  // namespace dls

extern "C" int dls_chunk_mean_f32(const float *X, int64_t ldx, float *Y, int64_t ldy, const int32_t *seg_ptr,
                                  const int32_t *seg_src_rows, const int32_t *seg_out_row, const int64_t *seg_col0,
                                  const int32_t *seg_len, int32_t n_segs, int32_t max_len, int32_t max_contributors,
                                  void *stream)
{
    if (!X || !Y || !seg_ptr || !seg_src_rows || !seg_out_row || !seg_col0 || !seg_len || n_segs < 0 || max_len < 0)
        return DLS_E_ARG;
    if (max_contributors > dls::CHUNK_MAX_SRC) return DLS_E_TOO_BIG;
    if (n_segs == 0 || max_len == 0) return 0;
    if (n_segs > 65535) return DLS_E_TOO_BIG;
    dim3 grid((unsigned)((max_len + 1023) / 1024), (unsigned)n_segs);
    dls::chunk_mean_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(X, ldx, Y, ldy, seg_ptr, seg_src_rows, seg_out_row,
                                                                    seg_col0, seg_len);
    DLS_LAUNCH_CHECK();
    return 0;
}
