// FEMNIST fc1 (3136 -> 512, 6.4 MB of fp32 weights per node) on tensor cores (tcgen05, bf16 operands, fp32 TMEM
// accumulators).  The layer is bound by streaming the node's weights from HBM, so every kernel reads them exactly
// once, as fp32 straight from the parameter row, converting to bf16 on the way into shared memory:
//   forward        h1[s][n]  = relu(b[n] + sum_k W[n][k] a2[s][k])      D[n, s], A = W tile (K-major),  B = a2
//   data gradient  da2[s][k] = sum_n W[n][k] dh1[s][n]                  D[k, s], A = W^T tile (MN-major), B = dh1
//   weight grad.   dW[n][k]  = sum_s dh1[s][n] a2[s][k]                 D[n, k], A = dh1^T, B = a2 (both MN-major)
// The small operands exist as bf16 in global memory in ONE layout each that serves both of their uses (the 16-byte
// unit = 8 consecutive k (or n) of one sample; units of consecutive samples are adjacent):
//   a2k [node][k/8][NSP][8]    K-major B of the forward, MN-major B of the weight gradient
//   dh1k[node][n/8][NSP][8]    K-major B of the data gradient, MN-major A of the weight gradient
// (NSP = the batch rounded up to 16; rows of missing samples are zero), staged by bulk asynchronous copies.
// The weight-gradient kernel applies the optimizer step in its epilogue (train waves): the 1.6 M fc1 gradients
// never go to memory, X (and the momentum row) are read and written once.
#include "gnlenet.cuh"
#include "tc_common.cuh"

namespace dls {
namespace fem {

using gnl::Pass;
using gnl::model_row;
using gnl::SgdFuse;

constexpr int FEAT = 3136, HID = 512, NK8 = FEAT / 8;

__device__ __forceinline__ void ld_global_v8(const void *p, float (&r)[8]) {
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7])
                 : "l"(p) : "memory");
}

// Weight streaming (forward, data gradient): warps 0-3 load fp32 weights (256-bit loads, eight per thread in
// flight), convert and store the bf16 UMMA layout; warp 4 issues the MMAs (and the bulk copies of B).  Small CTAs,
// several per SM, because a converter warp waits out one DRAM latency per chunk.  Item -> (row, 32-byte unit)
// mapping: the 8 lanes of a quarter-warp take 8 DIFFERENT rows — the destination (unit * rows + row) * 16 puts equal
// row mod 8 on the same banks, and a 128-bit store is processed per quarter-warp — so the stores are conflict-free
// and a warp instruction reads 8 rows x 128 contiguous bytes.
// Tried and dropped (measured on the 128-node FEMNIST wave, whole round): rows fetched by bulk (TMA-engine) copies —
// about 60 SM cycles per copy whatever its size, 256-byte rows move at 4 B/clk per SM (12.8 ms against 11.7 ms);
// 16-byte LDGSTS copies into a private fp32 ring (two half-used sectors per item, 12.0 ms); L2 prefetch of the next
// chunk (slower).  A probe that only reads the matrix in this pattern reaches 6.3 TB/s (scripts/probes/dram_runs.cu).
constexpr int FC_THREADS = 160, FC_CONV = 128, FC_CONV_WARPS = 4;

// ------------------------------------------------------------------------------------------
// forward: one CTA = one node x 128 output features; K streamed in chunks of 64
// ------------------------------------------------------------------------------------------
constexpr int FW_KC = 64, FW_NB = 2, FW_A_ST = (FW_KC / 8) * 128 * 16;
template <int NSP> struct FwCfg {
    static constexpr int B_ST = (FW_KC / 8) * NSP * 16, STAGE = FW_A_ST + B_ST, SMEM = FW_NB * STAGE + 128;
};

template <int NSP>
__global__ void __launch_bounds__(FC_THREADS, 4)
fc1_fwd_tc_kernel(Pass p, const uint16_t *__restrict__ a2k, float *__restrict__ h1, int w_off, int b_off)
{
    using C = FwCfg<NSP>;
    constexpr int NCHUNK = FEAT / FW_KC;                                 // 49
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *sS = smem;
    uint64_t *bars = reinterpret_cast<uint64_t *>(sS + FW_NB * C::STAGE);
    uint64_t *bfull = bars, *bempty = bfull + FW_NB, *done = bempty + FW_NB;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(done + 1);
    const int t = threadIdx.x, warp = tc::uniform_warp_idx(), lane = t & 31;
    const int mt = blockIdx.x, g = blockIdx.y;
    if (p.gcount[g] <= 0) return;
    int m;
    const float *row = model_row(p, g, m);
    const float *W = row + w_off + (int64_t)mt * 128 * FEAT;
    if (t == 0) {
        for (int i = 0; i < FW_NB; ++i) { tc::mbar_init(&bfull[i], FC_CONV_WARPS + 1); tc::mbar_init(&bempty[i], 1); }
        tc::mbar_init(done, 1);
        tc::mbar_fence_init();
    }
    if (warp == 4) tc::tmem_alloc(tmem_slot, 64);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp == 4) {
        const bool leader = tc::elect_one();
        const uint32_t idesc = tc::instr_desc_bf16(128, NSP, 0, 0);
        auto fetch_b = [&](int c) {
            const int sb = c % FW_NB;
            tc::mbar_expect_tx(&bfull[sb], (uint32_t)C::B_ST);
            tc::bulk_g2s(tc::smem_u32(sS + sb * C::STAGE + FW_A_ST), a2k + ((int64_t)g * NK8 + c * (FW_KC / 8)) * (NSP * 8),
                         (uint32_t)C::B_ST, &bfull[sb]);
        };
        if (leader)
            for (int c = 0; c < FW_NB && c < NCHUNK; ++c) fetch_b(c);
        __syncwarp();
        for (int c = 0; c < NCHUNK; ++c) {
            const int b = c % FW_NB;
            tc::mbar_wait(&bfull[b], (uint32_t)((c / FW_NB) & 1));
            tc::fence_after_sync();
            const uint32_t a_base = tc::smem_u32(sS + b * C::STAGE), b_base = a_base + FW_A_ST;
            if (leader) {
#pragma unroll
                for (int ks = 0; ks < FW_KC / 16; ++ks)
                    tc::mma_f16(tmem, tc::smem_desc(a_base + ks * 2 * 2048, 2048, 128),
                                tc::smem_desc(b_base + ks * 2 * NSP * 16, NSP * 16, 128), idesc, (c | ks) != 0 ? 1u : 0u);
                tc::mma_commit(&bempty[b]);
                if (c + FW_NB < NCHUNK) {                               // B of chunk c + NB goes where these MMAs read
                    tc::mbar_wait(&bempty[b], (uint32_t)((c / FW_NB) & 1));
                    fetch_b(c + FW_NB);
                }
            }
            __syncwarp();
        }
        if (leader) tc::mma_commit(done);
        __syncwarp();
    } else {
        auto item_n = [](int i) { return ((i >> 6) << 3) | (i & 7); };
        auto item_k8 = [](int i) { return (((i >> 5) & 1) << 2) | ((i >> 3) & 3); };
        for (int c = 0; c < NCHUNK; ++c) {
            const int sb = c % FW_NB;
            const float *src0 = W + c * FW_KC;
            float v[8][8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int i = t + j * FC_CONV;
                tc::ld_global_nc_v8(src0 + (int64_t)item_n(i) * FEAT + item_k8(i) * 8, v[j]);
            }
            if (c >= FW_NB) tc::mbar_wait(&bempty[sb], (uint32_t)(((c / FW_NB) - 1) & 1));
            uint8_t *sA = sS + sb * C::STAGE;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int i = t + j * FC_CONV;
                *reinterpret_cast<uint4 *>(sA + (item_k8(i) * 128 + item_n(i)) * 16) = tc::pack8_bf16(v[j]);
            }
            tc::fence_async_smem();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&bfull[sb]);
        }
        tc::mbar_wait(done, 0);
        tc::fence_after_sync();
        uint32_t r[NSP];
#pragma unroll
        for (int c0 = 0; c0 < NSP; c0 += 16)
            tc::tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, *reinterpret_cast<uint32_t(*)[16]>(&r[c0]));
        tc::tmem_ld_wait();
        const int n = mt * 128 + warp * 32 + lane;
        const float bias = row[b_off + n];
#pragma unroll
        for (int s = 0; s < NSP; ++s) {
            if (s < p.gs && p.img_idx[g * p.gs + s] >= 0) {
                const float x = __uint_as_float(r[s]) + bias;
                h1[(int64_t)(g * p.gs + s) * HID + n] = x > 0.f ? x : 0.f;
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 4) tc::tmem_dealloc(tmem, 64);
}

// ------------------------------------------------------------------------------------------
// data gradient: one CTA = one node x 128 input features (k); the reduction runs over the 512 rows of W
// ------------------------------------------------------------------------------------------
constexpr int DG_NC = 64, DG_NB = 2, DG_A_ST = 16 * DG_NC * 16, DG_TILES = (FEAT + 127) / 128;
template <int NSP> struct DgCfg {
    static constexpr int B_ALL = (HID / 8) * NSP * 16, SMEM = B_ALL + DG_NB * DG_A_ST + 128;
};

template <int NSP>
__global__ void __launch_bounds__(FC_THREADS, 3)
fc1_dgrad_tc_kernel(Pass p, const uint16_t *__restrict__ dh1k, float *__restrict__ da2, int w_off)
{
    using C = DgCfg<NSP>;
    constexpr int NCHUNK = HID / DG_NC;                                  // 8
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *sB = smem, *sA0 = smem + C::B_ALL;
    uint64_t *bars = reinterpret_cast<uint64_t *>(sA0 + DG_NB * DG_A_ST);
    uint64_t *bfull = bars, *bempty = bfull + DG_NB, *done = bempty + DG_NB, *bbar = done + 1;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bbar + 1);
    const int t = threadIdx.x, warp = tc::uniform_warp_idx(), lane = t & 31;
    const int kt = blockIdx.x, g = blockIdx.y, k0 = kt * 128;
    if (p.gcount[g] <= 0) return;
    int m;
    const float *row = model_row(p, g, m);
    const float *W = row + w_off + k0;
    const int kvalid8 = (FEAT - k0 < 128 ? FEAT - k0 : 128) / 8;         // 16-byte units of a row segment that exist
    if (t == 0) {
        for (int i = 0; i < DG_NB; ++i) { tc::mbar_init(&bfull[i], FC_CONV_WARPS); tc::mbar_init(&bempty[i], 1); }
        tc::mbar_init(done, 1);
        tc::mbar_init(bbar, 1);
        tc::mbar_fence_init();
    }
    if (warp == 4) tc::tmem_alloc(tmem_slot, 64);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp == 4) {
        const bool leader = tc::elect_one();
        if (leader) {
            tc::mbar_expect_tx(bbar, (uint32_t)C::B_ALL);
            tc::bulk_g2s(tc::smem_u32(sB), dh1k + (int64_t)g * (HID / 8) * (NSP * 8), (uint32_t)C::B_ALL, bbar);
        }
        __syncwarp();
        tc::mbar_wait(bbar, 0);
        const uint32_t idesc = tc::instr_desc_bf16(128, NSP, 1, 0);
        const uint32_t b_base = tc::smem_u32(sB);
        for (int c = 0; c < NCHUNK; ++c) {
            const int b = c % DG_NB;
            tc::mbar_wait(&bfull[b], (uint32_t)((c / DG_NB) & 1));
            tc::fence_after_sync();
            const uint32_t a_base = tc::smem_u32(sA0 + b * DG_A_ST);
            if (leader) {
#pragma unroll
                for (int ks = 0; ks < DG_NC / 16; ++ks)
                    tc::mma_f16(tmem, tc::smem_desc(a_base + ks * 256, 128, DG_NC * 16),
                                tc::smem_desc(b_base + (c * (DG_NC / 16) + ks) * 2 * NSP * 16, NSP * 16, 128), idesc,
                                (c | ks) != 0 ? 1u : 0u);
                tc::mma_commit(&bempty[b]);
            }
            __syncwarp();
        }
        if (leader) tc::mma_commit(done);
        __syncwarp();
    } else {
        auto item_n = [](int i) { return ((i >> 7) << 3) | (i & 7); };
        auto item_m8 = [](int i) { return (((i >> 5) & 3) << 2) | ((i >> 3) & 3); };
        for (int c = 0; c < NCHUNK; ++c) {
            const int sb = c % DG_NB;
            const float *src0 = W + (int64_t)c * DG_NC * FEAT;
            float v[8][8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int i = t + j * FC_CONV, m8 = item_m8(i);
#pragma unroll
                for (int e = 0; e < 8; ++e) v[j][e] = 0.f;
                if (m8 < kvalid8) tc::ld_global_nc_v8(src0 + (int64_t)item_n(i) * FEAT + m8 * 8, v[j]);
            }
            if (c >= DG_NB) tc::mbar_wait(&bempty[sb], (uint32_t)(((c / DG_NB) - 1) & 1));
            uint8_t *sA = sA0 + sb * DG_A_ST;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int i = t + j * FC_CONV;
                *reinterpret_cast<uint4 *>(sA + (item_m8(i) * DG_NC + item_n(i)) * 16) = tc::pack8_bf16(v[j]);
            }
            tc::fence_async_smem();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&bfull[sb]);
        }
        tc::mbar_wait(done, 0);
        tc::fence_after_sync();
        uint32_t r[NSP];
#pragma unroll
        for (int c0 = 0; c0 < NSP; c0 += 16)
            tc::tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, *reinterpret_cast<uint32_t(*)[16]>(&r[c0]));
        tc::tmem_ld_wait();
        const int k = k0 + warp * 32 + lane;
        if (k < FEAT) {
#pragma unroll
            for (int s = 0; s < NSP; ++s)
                if (s < p.gs && p.img_idx[g * p.gs + s] >= 0) da2[(int64_t)(g * p.gs + s) * FEAT + k] = __uint_as_float(r[s]);
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 4) tc::tmem_dealloc(tmem, 64);
}

// ------------------------------------------------------------------------------------------
// weight gradient (+ optimizer step): one CTA = one node x 128 output features; 25 tiles of 128 columns (k)
// ------------------------------------------------------------------------------------------
// the 25 column tiles of a (node, 128 features) pair are split over WG_ZS CTAs: 512 CTAs of 25 tiles are 1.15 waves
// of the 444 resident CTAs (the second wave is a tail of 68), 2560 CTAs of 5 tiles are 5.8
constexpr int WG_THREADS_FC = 160, WG_TN = 128, WG_TILES = (FEAT + WG_TN - 1) / WG_TN, WG_NST = 3, WG_ZS = 5, WG_TPC = WG_TILES / WG_ZS;
static_assert(WG_TPC * WG_ZS == WG_TILES, "tiles split evenly");
template <int NSP> struct WgCfg {
    static constexpr int A_BYTES = 16 * NSP * 16, B_ST = (WG_TN / 8) * NSP * 16, SMEM = A_BYTES + WG_NST * B_ST + 128;
};

template <int NSP, bool MOM>
__global__ void __launch_bounds__(WG_THREADS_FC, 3)
fc1_wgrad_tc_kernel(Pass p, const uint16_t *__restrict__ a2k, const uint16_t *__restrict__ dh1k,
                    const float *__restrict__ dh1, float *__restrict__ G, int64_t ldg, SgdFuse sgd, int w_off, int b_off)
{
    using C = WgCfg<NSP>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *sA = smem, *sB0 = smem + C::A_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(sB0 + WG_NST * C::B_ST);
    uint64_t *bfull = bars, *bempty = bars + WG_NST, *tfull = bars + 2 * WG_NST, *tempty = tfull + 1, *abar = tempty + 1;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(abar + 1);
    const int t = threadIdx.x, warp = tc::uniform_warp_idx(), lane = t & 31;
    const int mt = blockIdx.x, g = blockIdx.y, j0 = blockIdx.z * WG_TPC, j1 = j0 + WG_TPC;
    if (p.gcount[g] <= 0) return;
    if (t == 0) {
        for (int i = 0; i < WG_NST; ++i) { tc::mbar_init(&bfull[i], 1); tc::mbar_init(&bempty[i], 1); }
        tc::mbar_init(tfull, 1); tc::mbar_init(tempty, 128); tc::mbar_init(abar, 1);
        tc::mbar_fence_init();
    }
    if (warp == 4) tc::tmem_alloc(tmem_slot, 128);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp == 4) {
        const bool leader = tc::elect_one();
        auto fetch = [&](int j) {                                        // tile j of a2k -> stage j % WG_NST
            const int cols8 = (FEAT - j * WG_TN < WG_TN ? FEAT - j * WG_TN : WG_TN) / 8;
            const uint32_t bytes = (uint32_t)(cols8 * NSP * 16);
            tc::mbar_expect_tx(&bfull[(j - j0) % WG_NST], bytes);
            tc::bulk_g2s(tc::smem_u32(sB0 + ((j - j0) % WG_NST) * C::B_ST), a2k + ((int64_t)g * NK8 + j * (WG_TN / 8)) * (NSP * 8),
                         bytes, &bfull[(j - j0) % WG_NST]);
        };
        if (leader) {
            tc::mbar_expect_tx(abar, (uint32_t)C::A_BYTES);
            tc::bulk_g2s(tc::smem_u32(sA), dh1k + ((int64_t)g * (HID / 8) + mt * 16) * (NSP * 8), (uint32_t)C::A_BYTES, abar);
            for (int j = j0; j < j0 + WG_NST && j < j1; ++j) fetch(j);
        }
        __syncwarp();
        tc::mbar_wait(abar, 0);
        const uint32_t a_base = tc::smem_u32(sA);
        for (int j = j0; j < j1; ++j) {
            const int b = (j - j0) % WG_NST;
            tc::mbar_wait(&bfull[b], (uint32_t)(((j - j0) / WG_NST) & 1));
            if (j > j0) tc::mbar_wait(tempty, (uint32_t)((j - j0 - 1) & 1));
            tc::fence_after_sync();
            const int cols = FEAT - j * WG_TN < WG_TN ? FEAT - j * WG_TN : WG_TN;
            const uint32_t idesc = tc::instr_desc_bf16(128, cols, 1, 1);
            const uint32_t b_base = tc::smem_u32(sB0 + b * C::B_ST);
            if (leader) {
#pragma unroll
                for (int ks = 0; ks < NSP / 16; ++ks)
                    tc::mma_f16(tmem, tc::smem_desc(a_base + ks * 256, 128, NSP * 16), tc::smem_desc(b_base + ks * 256, 128, NSP * 16),
                                idesc, ks != 0 ? 1u : 0u);
                tc::mma_commit(tfull);
                tc::mma_commit(&bempty[b]);
                if (j + WG_NST < j1) {
                    tc::mbar_wait(&bempty[b], (uint32_t)(((j - j0) / WG_NST) & 1));
                    fetch(j + WG_NST);
                }
            }
            __syncwarp();
        }
    } else {
        const int n = mt * 128 + warp * 32 + lane;
        if (blockIdx.z == 0) {   // fc1 bias gradient from the fp32 dh1 rows (sample order)
            float bsum = 0.f;
            for (int s = 0; s < p.gs; ++s)
                if (p.img_idx[g * p.gs + s] >= 0) bsum += dh1[(int64_t)(g * p.gs + s) * HID + n];
            G[(int64_t)g * ldg + b_off + n] = bsum;
        }
        const bool fused = sgd.X != nullptr;
        float *xr = fused ? sgd.X + (int64_t)sgd.task_row[g] * sgd.ldx + w_off + (int64_t)n * FEAT
                          : G + (int64_t)g * ldg + w_off + (int64_t)n * FEAT;
        float *mr = MOM ? sgd.M + (int64_t)sgd.task_peer[g] * sgd.ldm + w_off + (int64_t)n * FEAT : nullptr;   // MOM implies fused
        const bool fst = (fused && sgd.first) ? (sgd.first[g] != 0) : false;
        const float lr = sgd.lr, mu = sgd.mu, wd = sgd.wd;
        for (int j = j0; j < j1; ++j) {
            const int cols = FEAT - j * WG_TN < WG_TN ? FEAT - j * WG_TN : WG_TN;
            tc::mbar_wait(tfull, (uint32_t)((j - j0) & 1));
            tc::fence_after_sync();
            if constexpr (!MOM) {
                // no momentum row: 64 columns (256 B of the row) requested per pass, consumed in two TMEM reads
                for (int c0 = 0; c0 < cols; c0 += 64) {
                    const int k = j * WG_TN + c0;
                    float pv[64];
                    if (fused) {
#pragma unroll
                        for (int q = 0; q < 8; ++q) ld_global_v8(xr + k + q * 8, *reinterpret_cast<float(*)[8]>(&pv[q * 8]));
                    }
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t r[32];
                        tc::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(c0 + 32 * h), r);
                        tc::tmem_ld_wait();
                        if (h == 1 && c0 + 64 >= cols) {                 // accumulator free for the next tile
                            tc::fence_before_sync();
                            tc::mbar_arrive(tempty);
                        }
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            float gv = __uint_as_float(r[i]);
                            if (fused) {
                                const float x = pv[32 * h + i];
                                if (wd != 0.f) gv = fmaf(wd, x, gv);
                                gv = fmaf(-lr, gv, x);
                            }
                            pv[32 * h + i] = gv;
                        }
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            tc::st_global_v8(xr + k + 32 * h + q * 8, pv[32 * h + q * 8], pv[32 * h + q * 8 + 1],
                                             pv[32 * h + q * 8 + 2], pv[32 * h + q * 8 + 3], pv[32 * h + q * 8 + 4],
                                             pv[32 * h + q * 8 + 5], pv[32 * h + q * 8 + 6], pv[32 * h + q * 8 + 7]);
                    }
                }
            } else {
            for (int c0 = 0; c0 < cols; c0 += 32) {
                const int k = j * WG_TN + c0;
                float pv[32], mv[32];
                {
#pragma unroll
                    for (int q = 0; q < 4; ++q) ld_global_v8(xr + k + q * 8, *reinterpret_cast<float(*)[8]>(&pv[q * 8]));
                    if (!fst) {
#pragma unroll
                        for (int q = 0; q < 4; ++q) ld_global_v8(mr + k + q * 8, *reinterpret_cast<float(*)[8]>(&mv[q * 8]));
                    }
                }
                uint32_t r[32];
                tc::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
                tc::tmem_ld_wait();
                if (c0 + 32 >= cols) {                                   // accumulator free for the next tile
                    tc::fence_before_sync();
                    tc::mbar_arrive(tempty);
                }
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    float gv = __uint_as_float(r[i]);
                    if (wd != 0.f) gv = fmaf(wd, pv[i], gv);
                    const float bv = fst ? gv : __fadd_rn(__fmul_rn(mv[i], mu), gv);
                    mv[i] = bv;
                    pv[i] = fmaf(-lr, bv, pv[i]);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    tc::st_global_v8(xr + k + q * 8, pv[q * 8], pv[q * 8 + 1], pv[q * 8 + 2], pv[q * 8 + 3], pv[q * 8 + 4],
                                     pv[q * 8 + 5], pv[q * 8 + 6], pv[q * 8 + 7]);
                    tc::st_global_v8(mr + k + q * 8, mv[q * 8], mv[q * 8 + 1], mv[q * 8 + 2], mv[q * 8 + 3], mv[q * 8 + 4],
                                     mv[q * 8 + 5], mv[q * 8 + 6], mv[q * 8 + 7]);
                }
            }
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 4) tc::tmem_dealloc(tmem, 128);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
int fc_nsp(int gs) { return (gs + 15) / 16 * 16; }

template <int NSP>
static int fwd_launch(const Pass &p, const uint16_t *a2k, float *h1, int w_off, int b_off, cudaStream_t st)
{
    static PerDeviceOnce attr;
    int rc = ensure_dyn_smem(fc1_fwd_tc_kernel<NSP>, FwCfg<NSP>::SMEM, attr);
    if (rc) return rc;
    fc1_fwd_tc_kernel<NSP><<<dim3(HID / 128, p.n_groups), FC_THREADS, FwCfg<NSP>::SMEM, st>>>(p, a2k, h1, w_off, b_off);
    DLS_LAUNCH_CHECK();
    return 0;
}
template <int NSP>
static int dgrad_launch(const Pass &p, const uint16_t *dh1k, float *da2, int w_off, cudaStream_t st)
{
    static PerDeviceOnce attr;
    int rc = ensure_dyn_smem(fc1_dgrad_tc_kernel<NSP>, DgCfg<NSP>::SMEM, attr);
    if (rc) return rc;
    fc1_dgrad_tc_kernel<NSP><<<dim3(DG_TILES, p.n_groups), FC_THREADS, DgCfg<NSP>::SMEM, st>>>(p, dh1k, da2, w_off);
    DLS_LAUNCH_CHECK();
    return 0;
}
template <int NSP>
static int wgrad_launch(const Pass &p, const uint16_t *a2k, const uint16_t *dh1k, const float *dh1, float *G, int64_t ldg,
                        const SgdFuse &sgd, int w_off, int b_off, cudaStream_t st)
{
    static PerDeviceOnce attr, attr_m;
    const dim3 grid(HID / 128, p.n_groups, WG_ZS);
    if (sgd.X && sgd.M && sgd.mu != 0.f) {
        int rc = ensure_dyn_smem(fc1_wgrad_tc_kernel<NSP, true>, WgCfg<NSP>::SMEM, attr_m);
        if (rc) return rc;
        fc1_wgrad_tc_kernel<NSP, true><<<grid, WG_THREADS_FC, WgCfg<NSP>::SMEM, st>>>(p, a2k, dh1k, dh1, G, ldg, sgd, w_off, b_off);
    } else {
        int rc = ensure_dyn_smem(fc1_wgrad_tc_kernel<NSP, false>, WgCfg<NSP>::SMEM, attr);
        if (rc) return rc;
        fc1_wgrad_tc_kernel<NSP, false><<<grid, WG_THREADS_FC, WgCfg<NSP>::SMEM, st>>>(p, a2k, dh1k, dh1, G, ldg, sgd, w_off, b_off);
    }
    DLS_LAUNCH_CHECK();
    return 0;
}

#define FC_DISPATCH(fn, ...)                                   \
    switch (fc_nsp(p.gs)) {                                    \
    case 16: return fn<16>(__VA_ARGS__);                       \
    case 32: return fn<32>(__VA_ARGS__);                       \
    case 48: return fn<48>(__VA_ARGS__);                       \
    case 64: return fn<64>(__VA_ARGS__);                       \
    default: return DLS_E_SHAPE;                               \
    }

int fc1_fwd_tc(const Pass &p, const uint16_t *a2k, float *h1, int w_off, int b_off, cudaStream_t st)
{
    FC_DISPATCH(fwd_launch, p, a2k, h1, w_off, b_off, st)
}
int fc1_dgrad_tc(const Pass &p, const uint16_t *dh1k, float *da2, int w_off, cudaStream_t st)
{
    FC_DISPATCH(dgrad_launch, p, dh1k, da2, w_off, st)
}
int fc1_wgrad_tc(const Pass &p, const uint16_t *a2k, const uint16_t *dh1k, const float *dh1, float *G, int64_t ldg,
                 const SgdFuse &sgd, int w_off, int b_off, cudaStream_t st)
{
    FC_DISPATCH(wgrad_launch, p, a2k, dh1k, dh1, G, ldg, sgd, w_off, b_off, st)
}

}  // namespace fem
}  // namespace dls
