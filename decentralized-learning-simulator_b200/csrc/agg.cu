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

// acc = fadd_rn(acc, fmul_rn(w, x)) on four lanes.  (The packed mul.rn.f32x2 / add.rn.f32x2 forms of
// sm_100a were tried: ptxas fuses the pair into one FFMA2, i.e. a single rounding, which breaks
// bit-exactness with the reference — keep the scalar intrinsics.)
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

// Packed CSR entry kept in shared memory: byte offset of the source row inside a stage + weight.
struct __align__(8) AggEntry { uint32_t off; float w; };

constexpr int STG_THREADS = 1024;
constexpr int STG_MAX_STAGES = 8;

template <int TW4, bool CSR_SMEM>   // TW4 = float4 columns per tile (4, 8 or 16)
__global__ void __launch_bounds__(STG_THREADS, 1)
agg_staged_kernel(const float *__restrict__ X, int64_t ldx, float *__restrict__ Y, int64_t ldy,
                  const int32_t *__restrict__ rowptr, const int32_t *__restrict__ colidx,
                  const float *__restrict__ wval, const int32_t *__restrict__ src_rows,
                  const int32_t *__restrict__ out_rows, int n_rows, int n_src, int n_ent, int64_t P4,
                  int n_tiles, int n_stages, const AggEntry *__restrict__ g_ent)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int stage_bytes = n_src * TW4 * 16;
    unsigned char *stages = smem_raw;                                        // [n_stages][n_src][TW4] float4
    int32_t *s_rowptr = reinterpret_cast<int32_t *>(smem_raw + (size_t)n_stages * stage_bytes);
    AggEntry *s_ent = reinterpret_cast<AggEntry *>(s_rowptr + ((n_rows + 1 + 3) & ~3));
    constexpr int ROWS_PER_PASS = STG_THREADS / TW4;
    const int lane_c = threadIdx.x % TW4;
    const int rsub = threadIdx.x / TW4;

    for (int i = threadIdx.x; i <= n_rows; i += STG_THREADS) s_rowptr[i] = rowptr[i];
    if (CSR_SMEM) {
        for (int i = threadIdx.x; i < n_ent; i += STG_THREADS) {
            AggEntry en;
            en.off = (uint32_t)colidx[i] * (TW4 * 16);
            en.w = wval[i];
            s_ent[i] = en;
        }
    }
    const AggEntry *ent = CSR_SMEM ? s_ent : g_ent;

    auto issue = [&](int tile, int stage) {
        if (tile < n_tiles) {
            unsigned char *dst = stages + (size_t)stage * stage_bytes;
            const int64_t c0 = (int64_t)tile * TW4;
            for (int i = threadIdx.x; i < n_src * TW4; i += STG_THREADS) {
                const int r = i / TW4, c = i % TW4;
                const int64_t xr = src_rows ? __ldg(src_rows + r) : r;
                if (c0 + c < P4) cp_async16(dst + (size_t)i * 16, X + xr * ldx + 4 * (c0 + c));
            }
        }
        cp_async_commit();          // always commit so that group counting stays uniform
    };

    // tiles of this CTA: blockIdx.x, blockIdx.x + grid, ...
    const int first = blockIdx.x;
    for (int s = 0; s < n_stages - 1; ++s) issue(first + s * gridDim.x, s);
    int it = 0;
    for (int tile = first; tile < n_tiles; tile += gridDim.x, ++it) {
        // tile `it` has landed once at most n_stages-2 younger groups are still pending
        switch (n_stages) {
            case 2: cp_async_wait<0>(); break;
            case 3: cp_async_wait<1>(); break;
            case 4: cp_async_wait<2>(); break;
            case 5: cp_async_wait<3>(); break;
            case 6: cp_async_wait<4>(); break;
            case 7: cp_async_wait<5>(); break;
            default: cp_async_wait<6>(); break;
        }
        __syncthreads();            // data visible to all; everyone finished computing tile it-1
        issue(tile + (n_stages - 1) * gridDim.x, (it + n_stages - 1) % n_stages);
        const unsigned char *src = stages + (size_t)(it % n_stages) * stage_bytes + lane_c * 16;
        const int64_t c = (int64_t)tile * TW4 + lane_c;
        for (int r = rsub; r < n_rows; r += ROWS_PER_PASS) {
            const int beg = s_rowptr[r], end = s_rowptr[r + 1];
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            int e = beg;
            if (beg < end) {
                const float4 x0 = *reinterpret_cast<const float4 *>(src + ent[beg].off);
                acc.x = __fmul_rn(x0.x, 0.f); acc.y = __fmul_rn(x0.y, 0.f);
                acc.z = __fmul_rn(x0.z, 0.f); acc.w = __fmul_rn(x0.w, 0.f);
            }
            for (; e + 4 <= end; e += 4) {
                AggEntry en[4];
                float4 xv[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) en[q] = ent[e + q];
#pragma unroll
                for (int q = 0; q < 4; ++q) xv[q] = *reinterpret_cast<const float4 *>(src + en[q].off);
#pragma unroll
                for (int q = 0; q < 4; ++q) acc_term(acc, en[q].w, xv[q]);
            }
            for (; e < end; ++e) {
                const AggEntry en = ent[e];
                acc_term(acc, en.w, *reinterpret_cast<const float4 *>(src + en.off));
            }
            if (c < P4) {
                const int64_t orow = out_rows ? out_rows[r] : r;
                *reinterpret_cast<float4 *>(Y + orow * ldy + 4 * c) = acc;
            }
        }
    }
    cp_async_wait<0>();
}

// ------------------------------------------------------------------------------------------
// Staged variant specialised for a UNIFORM fan-in known at compile time (every aggregate of a
// D-PSGD wave on a k-regular topology has k+1 models): no row pointers, fully unrolled neighbour
// loop, per-thread source pointers for the tile copies computed once.  Same arithmetic.
// ------------------------------------------------------------------------------------------
template <int TW4, int FAN>
__global__ void __launch_bounds__(STG_THREADS, 1)
agg_staged_uniform_kernel(const float *__restrict__ X, int64_t ldx, float *__restrict__ Y, int64_t ldy,
                          const int32_t *__restrict__ colidx, const float *__restrict__ wval,
                          const int32_t *__restrict__ src_rows, const int32_t *__restrict__ out_rows,
                          int n_rows, int n_src, int64_t P4, int n_tiles, int n_stages)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int stage_bytes = n_src * TW4 * 16;
    unsigned char *stages = smem_raw;
    AggEntry *s_ent = reinterpret_cast<AggEntry *>(smem_raw + (size_t)n_stages * stage_bytes);
    constexpr int ROWS_PER_PASS = STG_THREADS / TW4;
    constexpr int MAX_CP = 8;                                    // copies per thread per tile handled from registers
    const int lane_c = threadIdx.x % TW4;
    const int rsub = threadIdx.x / TW4;
    constexpr int FANP = (FAN + 1) & ~1;                         // entries per row padded to 16-byte pairs
    for (int i = threadIdx.x; i < n_rows * FANP; i += STG_THREADS) {
        const int r = i / FANP, q = i - r * FANP;
        AggEntry en;
        en.off = 0; en.w = 0.f;
        if (q < FAN) { en.off = (uint32_t)colidx[r * FAN + q] * (TW4 * 16); en.w = wval[r * FAN + q]; }
        s_ent[i] = en;
    }
    // this thread copies elements i = tid + k*1024 of every tile: source row r = i / TW4, column lane_c
    const float *cp_src[MAX_CP];
    const int n_cp = (n_src * TW4 + STG_THREADS - 1) / STG_THREADS;
#pragma unroll
    for (int k = 0; k < MAX_CP; ++k) {
        const int r = rsub + k * ROWS_PER_PASS;
        const int64_t xr = (r < n_src) ? (src_rows ? (int64_t)__ldg(src_rows + r) : (int64_t)r) : 0;
        cp_src[k] = X + xr * ldx + 4 * lane_c;
    }
    const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(stages);

    auto issue = [&](int tile, int stage) {
        if (tile < n_tiles) {
            const int64_t c0 = (int64_t)tile * TW4;
            if (c0 + lane_c < P4) {
                const uint32_t dst = smem_base + (uint32_t)stage * stage_bytes + threadIdx.x * 16;
#pragma unroll
                for (int k = 0; k < MAX_CP; ++k)
                    if (k < n_cp && rsub + k * ROWS_PER_PASS < n_src)
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;"
                                     :: "r"(dst + k * (STG_THREADS * 16)), "l"(cp_src[k] + 4 * c0));
            }
        }
        cp_async_commit();
    };

    const int first = blockIdx.x;
    for (int s = 0; s < n_stages - 1; ++s) issue(first + s * gridDim.x, s);
    int it = 0, cur = 0;
    for (int tile = first; tile < n_tiles; tile += gridDim.x, ++it) {
        switch (n_stages) {
            case 2: cp_async_wait<0>(); break;
            case 3: cp_async_wait<1>(); break;
            case 4: cp_async_wait<2>(); break;
            case 5: cp_async_wait<3>(); break;
            case 6: cp_async_wait<4>(); break;
            case 7: cp_async_wait<5>(); break;
            default: cp_async_wait<6>(); break;
        }
        __syncthreads();
        int nxt = cur - 1;
        if (nxt < 0) nxt = n_stages - 1;                         // buffer of tile it-1 == (it + n_stages - 1) % n_stages
        issue(tile + (n_stages - 1) * gridDim.x, nxt);
        const unsigned char *src = stages + (size_t)cur * stage_bytes + lane_c * 16;
        const int64_t c = (int64_t)tile * TW4 + lane_c;
        for (int r = rsub; r < n_rows; r += ROWS_PER_PASS) {
            const uint4 *er = reinterpret_cast<const uint4 *>(s_ent + r * FANP);
            AggEntry en[FANP];
            float4 xv[FAN];
#pragma unroll
            for (int q = 0; q < FANP / 2; ++q) {
                const uint4 pr = er[q];
                en[2 * q].off = pr.x; en[2 * q].w = __uint_as_float(pr.y);
                en[2 * q + 1].off = pr.z; en[2 * q + 1].w = __uint_as_float(pr.w);
            }
#pragma unroll
            for (int q = 0; q < FAN; ++q) xv[q] = *reinterpret_cast<const float4 *>(src + en[q].off);
            float4 acc;
            acc.x = __fmul_rn(xv[0].x, 0.f); acc.y = __fmul_rn(xv[0].y, 0.f);
            acc.z = __fmul_rn(xv[0].z, 0.f); acc.w = __fmul_rn(xv[0].w, 0.f);
#pragma unroll
            for (int q = 0; q < FAN; ++q) acc_term(acc, en[q].w, xv[q]);
            if (c < P4) {
                const int64_t orow = out_rows ? __ldg(out_rows + r) : r;
                *reinterpret_cast<float4 *>(Y + orow * ldy + 4 * c) = acc;
            }
        }
        if (++cur == n_stages) cur = 0;
    }
    cp_async_wait<0>();
}

// ------------------------------------------------------------------------------------------
// Staged variant for a uniform fan-in AND one weight for the whole wave (a D-PSGD aggregate wave without
// `weights`: every entry is float(1 / (k+1)), fedavg.py:14-15).  The CSR shrinks to 16-bit source indices
// (FAN rounded up to 8 per row), which is what lets 1024 source rows x fan-in 11 fit next to three stages.
// Same arithmetic: acc = x_first * 0; acc = fadd_rn(acc, fmul_rn(w, x)) in list order.
// ------------------------------------------------------------------------------------------
template <int TW4, int FAN>
__global__ void __launch_bounds__(STG_THREADS, 1)
agg_staged_u16_kernel(const float *__restrict__ X, int64_t ldx, float *__restrict__ Y, int64_t ldy,
                      const int32_t *__restrict__ colidx, const float *__restrict__ wval,
                      const int32_t *__restrict__ src_rows, const int32_t *__restrict__ out_rows,
                      int n_rows, int n_src, int64_t P4, int n_tiles, int n_stages)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int stage_bytes = n_src * TW4 * 16;
    const float w = __ldg(wval);                                 // the wave's one weight
    unsigned char *stages = smem_raw;
    uint16_t *s_idx = reinterpret_cast<uint16_t *>(smem_raw + (size_t)n_stages * stage_bytes);
    constexpr int ROWS_PER_PASS = STG_THREADS / TW4;
    constexpr int MAX_CP = 8;
    constexpr int FANP = (FAN + 7) & ~7;                         // 16-bit entries per row, 16-byte groups
    const int lane_c = threadIdx.x % TW4;
    const int rsub = threadIdx.x / TW4;
    for (int i = threadIdx.x; i < n_rows * FANP; i += STG_THREADS) {
        const int r = i / FANP, q = i - r * FANP;
        s_idx[i] = q < FAN ? (uint16_t)colidx[r * FAN + q] : (uint16_t)0;
    }
    const float *cp_src[MAX_CP];
    const int n_cp = (n_src * TW4 + STG_THREADS - 1) / STG_THREADS;
#pragma unroll
    for (int k = 0; k < MAX_CP; ++k) {
        const int r = rsub + k * ROWS_PER_PASS;
        const int64_t xr = (r < n_src) ? (src_rows ? (int64_t)__ldg(src_rows + r) : (int64_t)r) : 0;
        cp_src[k] = X + xr * ldx + 4 * lane_c;
    }
    const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(stages);

    auto issue = [&](int tile, int stage) {
        if (tile < n_tiles) {
            const int64_t c0 = (int64_t)tile * TW4;
            if (c0 + lane_c < P4) {
                const uint32_t dst = smem_base + (uint32_t)stage * stage_bytes + threadIdx.x * 16;
#pragma unroll
                for (int k = 0; k < MAX_CP; ++k)
                    if (k < n_cp && rsub + k * ROWS_PER_PASS < n_src)
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;"
                                     :: "r"(dst + k * (STG_THREADS * 16)), "l"(cp_src[k] + 4 * c0));
            }
        }
        cp_async_commit();
    };

    const int first = blockIdx.x;
    for (int s = 0; s < n_stages - 1; ++s) issue(first + s * gridDim.x, s);
    int it = 0, cur = 0;
    for (int tile = first; tile < n_tiles; tile += gridDim.x, ++it) {
        switch (n_stages) {
            case 2: cp_async_wait<0>(); break;
            case 3: cp_async_wait<1>(); break;
            case 4: cp_async_wait<2>(); break;
            case 5: cp_async_wait<3>(); break;
            case 6: cp_async_wait<4>(); break;
            case 7: cp_async_wait<5>(); break;
            default: cp_async_wait<6>(); break;
        }
        __syncthreads();
        int nxt = cur - 1;
        if (nxt < 0) nxt = n_stages - 1;
        issue(tile + (n_stages - 1) * gridDim.x, nxt);
        const unsigned char *src = stages + (size_t)cur * stage_bytes + lane_c * 16;
        const int64_t c = (int64_t)tile * TW4 + lane_c;
        for (int r = rsub; r < n_rows; r += ROWS_PER_PASS) {
            const uint4 *er = reinterpret_cast<const uint4 *>(s_idx + r * FANP);
            uint32_t pk[FANP / 2];
#pragma unroll
            for (int q = 0; q < FANP / 8; ++q) {
                const uint4 pr = er[q];
                pk[4 * q] = pr.x; pk[4 * q + 1] = pr.y; pk[4 * q + 2] = pr.z; pk[4 * q + 3] = pr.w;
            }
            float4 xv[FAN];
#pragma unroll
            for (int q = 0; q < FAN; ++q) {
                const uint32_t idx = (q & 1) ? (pk[q >> 1] >> 16) : (pk[q >> 1] & 0xffffu);
                xv[q] = *reinterpret_cast<const float4 *>(src + idx * (TW4 * 16));
            }
            float4 acc;
            acc.x = __fmul_rn(xv[0].x, 0.f); acc.y = __fmul_rn(xv[0].y, 0.f);
            acc.z = __fmul_rn(xv[0].z, 0.f); acc.w = __fmul_rn(xv[0].w, 0.f);
#pragma unroll
            for (int q = 0; q < FAN; ++q) acc_term(acc, w, xv[q]);
            if (c < P4) {
                const int64_t orow = out_rows ? __ldg(out_rows + r) : r;
                *reinterpret_cast<float4 *>(Y + orow * ldy + 4 * c) = acc;
            }
        }
        if (++cur == n_stages) cur = 0;
    }
    cp_async_wait<0>();
}

// CSR entries in global memory for waves whose CSR does not fit in shared memory next to the stages.
template <int TW4>
__global__ void agg_pack_entries_kernel(const int32_t *__restrict__ colidx, const float *__restrict__ wval,
                                        AggEntry *__restrict__ out, int n_ent)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_ent) { AggEntry en; en.off = (uint32_t)colidx[i] * (TW4 * 16); en.w = wval[i]; out[i] = en; }
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

struct StagedPlan { int tw4; int stages; bool csr_smem; size_t smem; int sms; };

// Largest tile width that still leaves >= 3 stages (2 loads in flight while one tile is consumed).
static bool plan_staged(const DeviceInfo &di, int n_rows, int n_src, int n_ent, StagedPlan &pl, int u16_fan = 0)
{
    pl.sms = di.sms;
    const int64_t budget = di.smem_optin - 256;
    // u16_fan > 0: the uniform-weight kernel (16-bit indices, fan-in padded to 8, no row pointers)
    const int64_t rowptr_bytes = u16_fan ? 0 : (int64_t)((n_rows + 1 + 3) & ~3) * 4;
    const int64_t csr_bytes = u16_fan ? (int64_t)n_rows * ((u16_fan + 7) & ~7) * 2
                                      : ((int64_t)n_ent + n_rows) * 8;   // room for the uniform kernel's pair padding
    const int e_tw = tune_value(TUNE_AGG_TW4), e_st = tune_value(TUNE_AGG_STAGES);   // dls_tune() overrides (0 = auto)
    const int cand[3] = {16, 8, 4};
    for (int pass = 0; pass < 2; ++pass) {
        const int min_stages = pass == 0 ? 3 : 2;
        for (int ci = 0; ci < 3; ++ci) {
            const int tw4 = cand[ci];
            if (e_tw && e_tw != tw4) continue;
            const int64_t stage = (int64_t)n_src * tw4 * 16;
            for (int csr = 1; csr >= 0; --csr) {
                const int64_t fixed = rowptr_bytes + (csr ? csr_bytes : 0);
                int64_t st = (budget - fixed) / stage;
                if (st > STG_MAX_STAGES) st = STG_MAX_STAGES;
                if (e_st && e_st <= st) st = e_st;
                if (st >= min_stages) {
                    pl.tw4 = tw4; pl.stages = (int)st; pl.csr_smem = csr != 0;
                    pl.smem = (size_t)(st * stage + fixed);
                    return true;
                }
            }
        }
    }
    return false;
}

template <int TW4, bool CSR_SMEM>
static int launch_staged_t(const StagedPlan &pl, const float *X, int64_t ldx, float *Y, int64_t ldy,
                           const int32_t *rowptr, const int32_t *colidx, const float *wval,
                           const int32_t *src_rows, const int32_t *out_rows, int n_rows, int n_src, int n_ent,
                           int64_t P4, const AggEntry *g_ent, cudaStream_t st)
{
    cudaError_t e = cudaFuncSetAttribute(agg_staged_kernel<TW4, CSR_SMEM>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
    if (e != cudaSuccess) return (int)e;
    const int n_tiles = (int)((P4 + TW4 - 1) / TW4);
    const int rounds = (n_tiles + pl.sms - 1) / pl.sms;
    const int grid = (n_tiles + rounds - 1) / rounds;          // same critical path, fewer idle tails
    agg_staged_kernel<TW4, CSR_SMEM><<<grid, STG_THREADS, pl.smem, st>>>(
        X, ldx, Y, ldy, rowptr, colidx, wval, src_rows, out_rows, n_rows, n_src, n_ent, P4, n_tiles, pl.stages,
        g_ent);
    DLS_LAUNCH_CHECK();
    return 0;
}

template <int TW4, int FAN>
static int launch_uniform_t(const StagedPlan &pl, const float *X, int64_t ldx, float *Y, int64_t ldy,
                            const int32_t *colidx, const float *wval, const int32_t *src_rows,
                            const int32_t *out_rows, int n_rows, int n_src, int64_t P4, cudaStream_t st)
{
    if ((n_src * TW4 + STG_THREADS - 1) / STG_THREADS > 8) return DLS_E_SHAPE;
    cudaError_t e = cudaFuncSetAttribute(agg_staged_uniform_kernel<TW4, FAN>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
    if (e != cudaSuccess) return (int)e;
    const int n_tiles = (int)((P4 + TW4 - 1) / TW4);
    const int rounds = (n_tiles + pl.sms - 1) / pl.sms;
    const int grid = (n_tiles + rounds - 1) / rounds;
    agg_staged_uniform_kernel<TW4, FAN><<<grid, STG_THREADS, pl.smem, st>>>(
        X, ldx, Y, ldy, colidx, wval, src_rows, out_rows, n_rows, n_src, P4, n_tiles, pl.stages);
    DLS_LAUNCH_CHECK();
    return 0;
}

template <int TW4>
static int launch_uniform_fan(int fan, const StagedPlan &pl, const float *X, int64_t ldx, float *Y, int64_t ldy,
                              const int32_t *colidx, const float *wval, const int32_t *src_rows,
                              const int32_t *out_rows, int n_rows, int n_src, int64_t P4, cudaStream_t st)
{
    switch (fan) {
#define DLS_FAN_CASE(F) \
    case F: return launch_uniform_t<TW4, F>(pl, X, ldx, Y, ldy, colidx, wval, src_rows, out_rows, n_rows, n_src, P4, st);
        DLS_FAN_CASE(2) DLS_FAN_CASE(3) DLS_FAN_CASE(4) DLS_FAN_CASE(5) DLS_FAN_CASE(6) DLS_FAN_CASE(7)
        DLS_FAN_CASE(8) DLS_FAN_CASE(9) DLS_FAN_CASE(10) DLS_FAN_CASE(11) DLS_FAN_CASE(12) DLS_FAN_CASE(13)
        DLS_FAN_CASE(14) DLS_FAN_CASE(15) DLS_FAN_CASE(16)
#undef DLS_FAN_CASE
        default: return DLS_E_SHAPE;
    }
}

template <int TW4, int FAN>
static int launch_u16_t(const StagedPlan &pl, const float *X, int64_t ldx, float *Y, int64_t ldy,
                        const int32_t *colidx, const float *w, const int32_t *src_rows, const int32_t *out_rows, int n_rows,
                        int n_src, int64_t P4, cudaStream_t st)
{
    if ((n_src * TW4 + STG_THREADS - 1) / STG_THREADS > 8) return DLS_E_SHAPE;
    cudaError_t e = cudaFuncSetAttribute(agg_staged_u16_kernel<TW4, FAN>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
    if (e != cudaSuccess) return (int)e;
    const int n_tiles = (int)((P4 + TW4 - 1) / TW4);
    const int rounds = (n_tiles + pl.sms - 1) / pl.sms;
    const int grid = (n_tiles + rounds - 1) / rounds;
    agg_staged_u16_kernel<TW4, FAN><<<grid, STG_THREADS, pl.smem, st>>>(
        X, ldx, Y, ldy, colidx, w, src_rows, out_rows, n_rows, n_src, P4, n_tiles, pl.stages);
    DLS_LAUNCH_CHECK();
    return 0;
}

template <int TW4>
static int launch_u16_fan(int fan, const StagedPlan &pl, const float *X, int64_t ldx, float *Y, int64_t ldy,
                          const int32_t *colidx, const float *w, const int32_t *src_rows, const int32_t *out_rows,
                          int n_rows, int n_src, int64_t P4, cudaStream_t st)
{
    switch (fan) {
#define DLS_FAN_CASE(F) \
    case F: return launch_u16_t<TW4, F>(pl, X, ldx, Y, ldy, colidx, w, src_rows, out_rows, n_rows, n_src, P4, st);
        DLS_FAN_CASE(2) DLS_FAN_CASE(3) DLS_FAN_CASE(4) DLS_FAN_CASE(5) DLS_FAN_CASE(6) DLS_FAN_CASE(7)
        DLS_FAN_CASE(8) DLS_FAN_CASE(9) DLS_FAN_CASE(10) DLS_FAN_CASE(11) DLS_FAN_CASE(12) DLS_FAN_CASE(13)
        DLS_FAN_CASE(14) DLS_FAN_CASE(15) DLS_FAN_CASE(16)
#undef DLS_FAN_CASE
        default: return DLS_E_SHAPE;
    }
}

static int launch_u16(int fan, const StagedPlan &pl, const float *X, int64_t ldx, float *Y, int64_t ldy,
                      const int32_t *colidx, const float *w, const int32_t *src_rows, const int32_t *out_rows, int n_rows,
                      int n_src, int64_t P4, cudaStream_t st)
{
    if (pl.tw4 == 16) return launch_u16_fan<16>(fan, pl, X, ldx, Y, ldy, colidx, w, src_rows, out_rows, n_rows, n_src, P4, st);
    if (pl.tw4 == 8) return launch_u16_fan<8>(fan, pl, X, ldx, Y, ldy, colidx, w, src_rows, out_rows, n_rows, n_src, P4, st);
    if (pl.tw4 == 4) return launch_u16_fan<4>(fan, pl, X, ldx, Y, ldy, colidx, w, src_rows, out_rows, n_rows, n_src, P4, st);
    return DLS_E_SHAPE;
}

static int launch_uniform(int fan, const StagedPlan &pl, const float *X, int64_t ldx, float *Y, int64_t ldy,
                          const int32_t *colidx, const float *wval, const int32_t *src_rows,
                          const int32_t *out_rows, int n_rows, int n_src, int64_t P4, cudaStream_t st)
{
    if (pl.tw4 == 16) return launch_uniform_fan<16>(fan, pl, X, ldx, Y, ldy, colidx, wval, src_rows, out_rows, n_rows, n_src, P4, st);
    if (pl.tw4 == 8) return launch_uniform_fan<8>(fan, pl, X, ldx, Y, ldy, colidx, wval, src_rows, out_rows, n_rows, n_src, P4, st);
    if (pl.tw4 == 4) return launch_uniform_fan<4>(fan, pl, X, ldx, Y, ldy, colidx, wval, src_rows, out_rows, n_rows, n_src, P4, st);
    return DLS_E_SHAPE;
}

static int launch_staged(const StagedPlan &pl, const float *X, int64_t ldx, float *Y, int64_t ldy,
                         const int32_t *rowptr, const int32_t *colidx, const float *wval,
                         const int32_t *src_rows, const int32_t *out_rows, int n_rows, int n_src, int n_ent,
                         int64_t P4, cudaStream_t st)
{
#define DLS_STAGED_CASE(TW)                                                                                       \
    if (pl.tw4 == TW) {                                                                                           \
        if (pl.csr_smem)                                                                                          \
            return launch_staged_t<TW, true>(pl, X, ldx, Y, ldy, rowptr, colidx, wval, src_rows, out_rows,       \
                                             n_rows, n_src, n_ent, P4, nullptr, st);                              \
        return DLS_E_SHAPE;                                                                                       \
    }
    DLS_STAGED_CASE(16)
    DLS_STAGED_CASE(8)
    DLS_STAGED_CASE(4)
#undef DLS_STAGED_CASE
    return DLS_E_SHAPE;
}

}  // namespace dls

using namespace dls;

extern "C" int dls_agg_spmm_f32(const float *X, int64_t ldx, float *Y, int64_t ldy,
                                const int32_t *rowptr, const int32_t *colidx, const float *wval,
                                const int32_t *src_rows, const int32_t *out_rows, int32_t n_rows,
                                int32_t n_src_rows, int32_t nnz, int64_t P, int32_t variant, void *stream)
{
    if (!X || !Y || !rowptr || !colidx || !wval || n_rows < 0 || nnz < 0 || P < 0 || ldx < P || ldy < P)
        return DLS_E_ARG;
    if (n_rows == 0 || P == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const DeviceInfo *di = nullptr;
    int rc = device_info(&di);
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
    // the staged variant keeps >= 2 tiles of ALL source rows (and normally the CSR) in shared memory
    StagedPlan pl;
    bool can_stage = false;
    const int n_ent = nnz;
    if (n_src_rows > 0 && n_src_rows <= 65535 && nnz >= 0 && (variant == 2 || variant == 3 || variant == 0))
        can_stage = plan_staged(*di, n_rows, n_src_rows, n_ent, pl) && pl.csr_smem;
    if (variant == 4) {
        // the caller asserts a uniform fan-in AND that every entry carries the weight wval[0]
        if (n_src_rows <= 0 || n_src_rows > 65535 || nnz % n_rows != 0) return DLS_E_SHAPE;
        const int fan = nnz / n_rows;
        if (fan < 2 || fan > 16 || !plan_staged(*di, n_rows, n_src_rows, n_ent, pl, fan) || !pl.csr_smem) return DLS_E_SHAPE;
        return launch_u16(fan, pl, X, ldx, Y, ldy, colidx, wval, src_rows, out_rows, n_rows, n_src_rows, P4, st);
    }
    if (variant == 3) {
        // the caller asserts a uniform fan-in: rowptr[r] == r * (nnz / n_rows)
        if (!can_stage || nnz % n_rows != 0) return DLS_E_SHAPE;
        return launch_uniform(nnz / n_rows, pl, X, ldx, Y, ldy, colidx, wval, src_rows, out_rows, n_rows,
                              n_src_rows, P4, st);
    }
    if (variant == 2 && !can_stage) return DLS_E_SHAPE;
    const bool use_staged = (variant == 2);   // auto keeps the gather variant until measured otherwise
    if (use_staged)
        return launch_staged(pl, X, ldx, Y, ldy, rowptr, colidx, wval, src_rows, out_rows, n_rows, n_src_rows,
                             n_ent, P4, st);
    const int64_t n_tiles = (P4 + AGG_TILE4 - 1) / AGG_TILE4;
    const int64_t grid = n_tiles * n_rows;
    if (grid > 2147483647LL) return DLS_E_TOO_BIG;
    agg_gather_kernel<<<(unsigned)grid, AGG_THREADS, 0, st>>>(X, ldx, Y, ldy, rowptr, colidx, wval,
                                                              src_rows, out_rows, n_rows, P4);
    DLS_LAUNCH_CHECK();
    return 0;
}

// Would `variant` (2 = staged, 3 = staged with uniform fan-in) accept a wave of this shape on the current
// device?  1 = yes, 0 = no (the gather variant 1 always works), negative = error.  Lets the caller pick a
// variant up front instead of probing with a launch.
extern "C" int dls_agg_variant_fits(int32_t variant, int32_t n_rows, int32_t n_src_rows, int32_t nnz)
{
    if (n_rows < 0 || n_src_rows < 0 || nnz < 0) return DLS_E_ARG;
    if (variant == 0 || variant == 1) return 1;
    if (variant != 2 && variant != 3 && variant != 4) return DLS_E_ARG;
    if (n_rows == 0) return 1;
    const DeviceInfo *di = nullptr;
    int rc = device_info(&di);
    if (rc) return rc < 0 ? rc : -rc;
    StagedPlan pl;
    if (n_src_rows <= 0 || n_src_rows > 65535) return 0;
    if (variant == 4) {
        if (nnz % n_rows != 0) return 0;
        const int fan = nnz / n_rows;
        if (fan < 2 || fan > 16) return 0;
        if (!plan_staged(*di, n_rows, n_src_rows, nnz, pl, fan) || !pl.csr_smem) return 0;
        return (n_src_rows * pl.tw4 + STG_THREADS - 1) / STG_THREADS > 8 ? 0 : 1;
    }
    if (!(plan_staged(*di, n_rows, n_src_rows, nnz, pl) && pl.csr_smem)) return 0;
    if (variant == 3) {
        if (nnz % n_rows != 0) return 0;
        const int fan = nnz / n_rows;
        if (fan < 2 || fan > 16) return 0;
        if ((n_src_rows * pl.tw4 + STG_THREADS - 1) / STG_THREADS > 8) return 0;
    }
    return 1;
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
