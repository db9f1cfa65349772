// Aggregate wave: sparse mixing-matrix product Y = W . X over the stacked parameter tensor.
//
// Replaces FedAvg.aggregate (reference: dasklearn/gradient_aggregation/fedavg.py:12-26) for all
// aggregate tasks of a wavefront.  Arithmetic contract (bit-exact with the reference):
//   acc = x_first * 0;  for e in list order: acc = fadd_rn(acc, fmul_rn(w_e, x_e))
// i.e. separate multiply and add, never an FMA, summation in the task's `models` list order.
//
// HBM-bound: algorithmic traffic is 2*N*P*4 bytes (every X row read once, every Y row written
// once).  Every X row is consumed by fan-in many output rows, so the re-reads must be served on
// chip.  Two variants:
//   gather  (1): one CTA = one output row x one column tile; CTAs are ordered column-tile-major
//                so that the N source-row segments of the current column tiles stay L2-resident
//                while all rows consume them.  Re-reads cost L2 bandwidth.
//   staged  (2): one CTA = one column tile of ALL source rows staged in shared memory once
//                (cp.async, double buffered); every output row is formed from shared memory, so
//                HBM and L2 see each X element exactly once.
#include "common.cuh"

namespace dls {

constexpr int AGG_THREADS = 256;
constexpr int AGG_UNROLL = 2;                                  // float4 per thread per row
constexpr int AGG_TILE4 = AGG_THREADS * AGG_UNROLL;            // float4 columns per CTA tile
constexpr int AGG_MAX_FAN_SMEM = 64;                           // neighbour entries staged per pass

__device__ __forceinline__ void acc_term(float4 &a, const float w, const float4 x) {
    a.x = __fadd_rn(a.x, __fmul_rn(w, x.x));
    a.y = __fadd_rn(a.y, __fmul_rn(w, x.y));
    a.z = __fadd_rn(a.z, __fmul_rn(w, x.z));
    a.w = __fadd_rn(a.w, __fmul_rn(w, x.w));
}

// grid.x = n_col_tiles * n_rows, column tile is the slow index.
__global__ void __launch_bounds__(AGG_THREADS)
agg_gather_kernel(const float *__restrict__ X, int64_t ldx, float *__restrict__ Y, int64_t ldy,
                  const int32_t *__restrict__ rowptr, const int32_t *__restrict__ colidx,
                  const float *__restrict__ wval, const int32_t *__restrict__ src_rows,
                  const int32_t *__restrict__ out_rows, int n_rows, int64_t P4)
{
    __shared__ int32_t s_col[AGG_MAX_FAN_SMEM];
    __shared__ float s_w[AGG_MAX_FAN_SMEM];

    const int row = blockIdx.x % n_rows;
    const int64_t tile = blockIdx.x / n_rows;
    const int beg = rowptr[row], end = rowptr[row + 1];
    const int64_t orow = out_rows ? out_rows[row] : row;
    const int64_t c0 = tile * AGG_TILE4 + threadIdx.x;          // first float4 column of this thread

    float4 acc[AGG_UNROLL];
    bool first = true;
    for (int base = beg; base < end; base += AGG_MAX_FAN_SMEM) {
        const int n = min(AGG_MAX_FAN_SMEM, end - base);
        __syncthreads();
        if (threadIdx.x < n) {
            const int32_t cidx = colidx[base + threadIdx.x];
            s_col[threadIdx.x] = src_rows ? src_rows[cidx] : cidx;
            s_w[threadIdx.x] = wval[base + threadIdx.x];
        }
        __syncthreads();
        int j = 0;
        // neighbours in groups of 4: 4*AGG_UNROLL independent 128-bit loads in flight per thread
        for (; j + 4 <= n; j += 4) {
            float4 v[4][AGG_UNROLL];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float *xr = X + (int64_t)s_col[j + q] * ldx;
#pragma unroll
                for (int u = 0; u < AGG_UNROLL; ++u) {
                    const int64_t c = c0 + (int64_t)u * AGG_THREADS;
                    v[q][u] = (c < P4) ? ldg_stream4(xr + 4 * c) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            if (first) {
#pragma unroll
                for (int u = 0; u < AGG_UNROLL; ++u) {
                    acc[u].x = __fmul_rn(v[0][u].x, 0.f); acc[u].y = __fmul_rn(v[0][u].y, 0.f);
                    acc[u].z = __fmul_rn(v[0][u].z, 0.f); acc[u].w = __fmul_rn(v[0][u].w, 0.f);
                }
                first = false;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float w = s_w[j + q];
#pragma unroll
                for (int u = 0; u < AGG_UNROLL; ++u) acc_term(acc[u], w, v[q][u]);
            }
        }
        for (; j < n; ++j) {
            const float *xr = X + (int64_t)s_col[j] * ldx;
            const float w = s_w[j];
            float4 v[AGG_UNROLL];
#pragma unroll
            for (int u = 0; u < AGG_UNROLL; ++u) {
                const int64_t c = c0 + (int64_t)u * AGG_THREADS;
                v[u] = (c < P4) ? ldg_stream4(xr + 4 * c) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            if (first) {
#pragma unroll
                for (int u = 0; u < AGG_UNROLL; ++u) {
                    acc[u].x = __fmul_rn(v[u].x, 0.f); acc[u].y = __fmul_rn(v[u].y, 0.f);
                    acc[u].z = __fmul_rn(v[u].z, 0.f); acc[u].w = __fmul_rn(v[u].w, 0.f);
                }
                first = false;
            }
#pragma unroll
            for (int u = 0; u < AGG_UNROLL; ++u) acc_term(acc[u], w, v[u]);
        }
    }
    if (first) {   // empty model list: the reference would fail; define the result as zeros
#pragma unroll
        for (int u = 0; u < AGG_UNROLL; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float *yr = Y + orow * ldy;
#pragma unroll
    for (int u = 0; u < AGG_UNROLL; ++u) {
        const int64_t c = c0 + (int64_t)u * AGG_THREADS;
        if (c < P4) *reinterpret_cast<float4 *>(yr + 4 * c) = acc[u];
    }
}

// ------------------------------------------------------------------------------------------
// Staged variant.  CTA (persistent, grid = #SMs * k) loops over column tiles of TW4 float4
// columns; for each tile it copies the segment [tile] of every source row into shared memory
// with cp.async (16-byte, coalesced per row), double buffered, then every warp forms output
// rows from shared memory: lane l of a row group handles float4 column l.
// Shared layout: buf[stage][src_row][TW4] float4.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" :: "n"(N));
}

template <int TW4>   // float4 columns per tile: 8 (128 B per row segment), 16 or 32
__global__ void __launch_bounds__(512)
agg_staged_kernel(const float *__restrict__ X, int64_t ldx, float *__restrict__ Y, int64_t ldy,
                  const int32_t *__restrict__ rowptr, const int32_t *__restrict__ colidx,
                  const float *__restrict__ wval, const int32_t *__restrict__ src_rows,
                  const int32_t *__restrict__ out_rows, int n_rows, int n_src, int64_t P4, int n_tiles)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4 *buf = reinterpret_cast<float4 *>(smem_raw);            // [2][n_src][TW4]
    const int stage_elems = n_src * TW4;
    constexpr int ROWS_PER_PASS = 512 / TW4;                       // output rows formed per pass
    const int lane_c = threadIdx.x % TW4;                          // float4 column inside the tile
    const int rsub = threadIdx.x / TW4;

    auto issue = [&](int tile, int stage) {
        float4 *dst = buf + stage * stage_elems;
        const int64_t c0 = (int64_t)tile * TW4;
        for (int i = threadIdx.x; i < stage_elems; i += 512) {
            const int r = i / TW4, c = i % TW4;
            const int64_t xr = src_rows ? __ldg(src_rows + r) : r;
            if (c0 + c < P4) cp_async16(dst + i, X + xr * ldx + 4 * (c0 + c));
        }
        cp_async_commit();
    };

    int tile = blockIdx.x;
    int stage = 0;
    if (tile < n_tiles) issue(tile, 0);
    for (; tile < n_tiles; tile += gridDim.x, stage ^= 1) {
        const int next = tile + gridDim.x;
        if (next < n_tiles) { issue(next, stage ^ 1); cp_async_wait<1>(); }
        else cp_async_wait<0>();
        __syncthreads();
        const float4 *src = buf + stage * stage_elems;
        const int64_t c = (int64_t)tile * TW4 + lane_c;
        for (int r = rsub; r < n_rows; r += ROWS_PER_PASS) {
            const int beg = __ldg(rowptr + r), end = __ldg(rowptr + r + 1);
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            if (beg < end) {
                const float4 x0 = src[__ldg(colidx + beg) * TW4 + lane_c];
                acc.x = __fmul_rn(x0.x, 0.f); acc.y = __fmul_rn(x0.y, 0.f);
                acc.z = __fmul_rn(x0.z, 0.f); acc.w = __fmul_rn(x0.w, 0.f);
            }
            for (int e = beg; e < end; ++e) {
                const float w = __ldg(wval + e);
                const float4 x = src[__ldg(colidx + e) * TW4 + lane_c];
                acc_term(acc, w, x);
            }
            if (c < P4) {
                const int64_t orow = out_rows ? out_rows[r] : r;
                *reinterpret_cast<float4 *>(Y + orow * ldy + 4 * c) = acc;
            }
        }
        __syncthreads();   // everyone done reading `stage` before it is refilled
    }
}

// Scalar fallback for unaligned rows (ld % 4 != 0 or unaligned base): same arithmetic.
__global__ void __launch_bounds__(256)
agg_scalar_kernel(const float *__restrict__ X, int64_t ldx, float *__restrict__ Y, int64_t ldy,
                  const int32_t *__restrict__ rowptr, const int32_t *__restrict__ colidx,
                  const float *__restrict__ wval, const int32_t *__restrict__ src_rows,
                  const int32_t *__restrict__ out_rows, int n_rows, int64_t P)
{
    const int row = blockIdx.y;
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= P) return;
    const int beg = rowptr[row], end = rowptr[row + 1];
    float acc = 0.f;
    auto src = [&](int e) -> int64_t { const int32_t ci = colidx[e]; return src_rows ? src_rows[ci] : ci; };
    if (beg < end) acc = __fmul_rn(X[src(beg) * ldx + c], 0.f);
    for (int e = beg; e < end; ++e)
        acc = __fadd_rn(acc, __fmul_rn(wval[e], X[src(e) * ldx + c]));
    const int64_t orow = out_rows ? out_rows[row] : row;
    Y[orow * ldy + c] = acc;
}

__global__ void __launch_bounds__(256)
copy_rows_kernel(const float *__restrict__ X, int64_t ldx, float *__restrict__ Y, int64_t ldy,
                 const int32_t *__restrict__ src_rows, const int32_t *__restrict__ dst_rows,
                 int64_t P4)
{
    const int r = blockIdx.y;
    const float *xs = X + (int64_t)(src_rows ? src_rows[r] : r) * ldx;   // ldx == 0: broadcast
    float *yd = Y + (int64_t)(dst_rows ? dst_rows[r] : r) * ldy;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < P4;
         c += (int64_t)gridDim.x * blockDim.x)
        reinterpret_cast<float4 *>(yd)[c] = reinterpret_cast<const float4 *>(xs)[c];
}

static int g_num_sms = 0;
static int g_max_smem_optin = 0;
static int device_props() {
    if (g_num_sms == 0) {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return (int)e;
        e = cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return (int)e;
        e = cudaDeviceGetAttribute(&g_max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        if (e != cudaSuccess) return (int)e;
    }
    return 0;
}

template <int TW4>
static int launch_staged(const float *X, int64_t ldx, float *Y, int64_t ldy, const int32_t *rowptr,
                         const int32_t *colidx, const float *wval, const int32_t *src_rows,
                         const int32_t *out_rows, int n_rows, int n_src, int64_t P4, cudaStream_t st)
{
    const size_t smem = (size_t)2 * n_src * TW4 * sizeof(float4);
    cudaError_t e = cudaFuncSetAttribute(agg_staged_kernel<TW4>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    const int n_tiles = (int)((P4 + TW4 - 1) / TW4);
    const int grid = min(n_tiles, g_num_sms);
    agg_staged_kernel<TW4><<<grid, 512, smem, st>>>(X, ldx, Y, ldy, rowptr, colidx, wval, src_rows,
                                                     out_rows, n_rows, n_src, P4, n_tiles);
    DLS_LAUNCH_CHECK();
    return 0;
}

}  // namespace dls

using namespace dls;

extern "C" int dls_agg_spmm_f32(const float *X, int64_t ldx, float *Y, int64_t ldy,
                                const int32_t *rowptr, const int32_t *colidx, const float *wval,
                                const int32_t *src_rows, const int32_t *out_rows, int32_t n_rows,
                                int32_t n_src_rows, int64_t P, int32_t variant, void *stream)
{
    if (!X || !Y || !rowptr || !colidx || !wval || n_rows < 0 || P < 0 || ldx < P || ldy < P)
        return DLS_E_ARG;
    if (n_rows == 0 || P == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int rc = device_props();
    if (rc) return rc;
    const bool vec_ok = aligned16(X) && aligned16(Y) && (ldx % 4 == 0) && (ldy % 4 == 0) &&
                        (((P + 3) / 4) * 4 <= ldx) && (((P + 3) / 4) * 4 <= ldy);
    if (!vec_ok) {
        if (variant == 2) return DLS_E_ALIGN;
        dim3 grid((unsigned)((P + 255) / 256), (unsigned)n_rows);
        agg_scalar_kernel<<<grid, 256, 0, st>>>(X, ldx, Y, ldy, rowptr, colidx, wval, src_rows, out_rows, n_rows, P);
        DLS_LAUNCH_CHECK();
        return 0;
    }
    const int64_t P4 = (P + 3) / 4;
    // staged variant needs 2 stages of all source rows x >= 8 float4 columns in shared memory
    const int64_t smem_budget = g_max_smem_optin - 1024;
    int tw4 = 0;
    if (n_src_rows > 0) {
        if ((int64_t)2 * n_src_rows * 32 * 16 <= smem_budget) tw4 = 32;
        else if ((int64_t)2 * n_src_rows * 16 * 16 <= smem_budget) tw4 = 16;
        else if ((int64_t)2 * n_src_rows * 8 * 16 <= smem_budget) tw4 = 8;
    }
    if (variant == 2 && tw4 == 0) return DLS_E_SHAPE;
    const bool use_staged = (variant == 2);   // auto keeps the gather variant until measured otherwise
    if (use_staged) {
        if (tw4 == 32) return launch_staged<32>(X, ldx, Y, ldy, rowptr, colidx, wval, src_rows, out_rows, n_rows, n_src_rows, P4, st);
        if (tw4 == 16) return launch_staged<16>(X, ldx, Y, ldy, rowptr, colidx, wval, src_rows, out_rows, n_rows, n_src_rows, P4, st);
        return launch_staged<8>(X, ldx, Y, ldy, rowptr, colidx, wval, src_rows, out_rows, n_rows, n_src_rows, P4, st);
    }
    const int64_t n_tiles = (P4 + AGG_TILE4 - 1) / AGG_TILE4;
    const int64_t grid = n_tiles * n_rows;
    if (grid > 2147483647LL) return DLS_E_TOO_BIG;
    agg_gather_kernel<<<(unsigned)grid, AGG_THREADS, 0, st>>>(X, ldx, Y, ldy, rowptr, colidx, wval,
                                                              src_rows, out_rows, n_rows, P4);
    DLS_LAUNCH_CHECK();
    return 0;
}

static int copy_rows_impl(const float *X, int64_t ldx, float *Y, int64_t ldy,
                          const int32_t *src_rows, const int32_t *dst_rows, int32_t n_rows,
                          int64_t P, void *stream, bool bcast)
{
    if (!X || !Y || n_rows < 0 || P < 0) return DLS_E_ARG;
    if (n_rows == 0 || P == 0) return 0;
    if (!aligned16(X) || !aligned16(Y) || ldy % 4 || ((P + 3) / 4) * 4 > ldy) return DLS_E_ALIGN;
    if (!bcast && (ldx % 4 || ((P + 3) / 4) * 4 > ldx)) return DLS_E_ALIGN;
    const int64_t P4 = (P + 3) / 4;
    const int64_t gx = (P4 + 255) / 256;
    dim3 grid((unsigned)(gx < 64 ? gx : 64), (unsigned)n_rows);
    copy_rows_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(X, ldx, Y, ldy, src_rows, dst_rows, P4);
    DLS_LAUNCH_CHECK();
    return 0;
}

extern "C" int dls_broadcast_row(const float *src, float *Y, int64_t ldy, const int32_t *dst_rows,
                                 int32_t n_rows, int64_t P, void *stream)
{
    // the source vector must be readable up to the next multiple of 4 floats
    return copy_rows_impl(src, 0, Y, ldy, nullptr, dst_rows, n_rows, P, stream, true);
}

extern "C" int dls_copy_rows(const float *X, int64_t ldx, float *Y, int64_t ldy,
                             const int32_t *src_rows, const int32_t *dst_rows, int32_t n_rows,
                             int64_t P, void *stream)
{
    return copy_rows_impl(X, ldx, Y, ldy, src_rows, dst_rows, n_rows, P, stream, false);
}
