// Tensor-core (tcgen05) implicit-GEMM 5x5 convolutions for GNLeNet, bf16 operands, fp32 TMEM
// accumulators (precision 1).
//
// Implicit GEMM without im2col: the activations of SB samples are staged ONCE in shared memory as
// bf16 "channel-chunk planes"  plane[c/8][padded pixel][8 channels]  (16 bytes per pixel).  In the
// no-swizzle K-major UMMA layout a row of the A operand is 16 bytes and 8 consecutive rows are a
// 128-byte core matrix — which is exactly 8 consecutive pixels of a plane.  So the A operand of
// tap (ky,kx) for 128 consecutive output-pixel origins is just the plane address shifted by
// (ky*PP + kx) pixels: 25 taps x CIN/16 MMAs (M=128, N=COUT, K=16) accumulate one output tile in
// TMEM, reading the same staged bytes.  Origins are enumerated over the PADDED grid (PP = H+4
// wide), so rows whose origin falls in the padding are computed and discarded (H=15: 62 % useful).
//   A desc: start = plane(2h) + (tile*128 + ky*PP + kx)*16, LBO = plane stride, SBO = 128
//   B desc: weights as [k/8][COUT][8] bf16 (k = tap*CIN + ci), LBO = COUT*16, SBO = 128
// The same kernel computes the data gradient (input = dY, weights transposed + tap-flipped).
#include "gnlenet.cuh"
#include "tc_common.cuh"
#include <cuda_bf16.h>

namespace dls {
namespace gnl {

// bf16 UMMA-layout weight copies per model (elements):
//   W2u [100][32][8] | W2ud [100][32][8] | W3u [100][64][8] | W3ud [200][32][8]
constexpr int TCO_W2U = 0, TCO_W2UD = 25600, TCO_W3U = 51200, TCO_W3UD = 102400;

__device__ __forceinline__ uint4 pack_bf16x8(const float4 a, const float4 b) {
    __nv_bfloat162 q0 = __floats2bfloat162_rn(a.x, a.y), q1 = __floats2bfloat162_rn(a.z, a.w);
    __nv_bfloat162 q2 = __floats2bfloat162_rn(b.x, b.y), q3 = __floats2bfloat162_rn(b.z, b.w);
    uint4 v;
    v.x = *reinterpret_cast<uint32_t *>(&q0); v.y = *reinterpret_cast<uint32_t *>(&q1);
    v.z = *reinterpret_cast<uint32_t *>(&q2); v.w = *reinterpret_cast<uint32_t *>(&q3);
    return v;
}

constexpr int PREP_ROW = 801;                 // padded source row (16 output channels x 800 weights) in shared memory
constexpr int PREP_SMEM = 16 * PREP_ROW * 4;

// One CTA = 16 output channels of conv2 (parts 0-1) or conv3 (parts 2-5) of one model: their 12800 fp32
// weights are read once, coalesced, into shared memory and every 16-byte bf16 operand entry (8 k-values of
// one n) is gathered from there and stored whole.
//   W2u / W3u  : entry (kc = tap*4 + ci/8, n = co)          values e: W[co][ci = 8*(kc&3) + e][tap]
//   W2ud / W3ud: entry (kc = tap'*(NCO/8) + co/8, n = ci)    values e: W[co = 8*(kc % (NCO/8)) + e][ci][24 - tap']
// GROUPED layouts (the default): the taps of one kernel row are stacked along N in three groups, kx = (0,1), (2,3)
// and (4), so that one MMA with N = 2*NN columns evaluates two taps on the same shifted window (see
// conv_tc3_kernel).  Per kernel row ky: [group 0: chunk][j*NN + n][8], [group 1: ...], [group 2: chunk][n][8].
__device__ __forceinline__ int64_t grouped_entry(int tap, int chunk, int nck, int nn, int n) {
    const int ky = tap / 5, kx = tap - ky * 5, grp = kx >> 1, j = kx & 1;
    const int ng = (grp == 2 ? 1 : 2) * nn;
    return (int64_t)ky * (nck * 5 * nn) + grp * (nck * 2 * nn) + chunk * ng + j * nn + n;
}

__global__ void __launch_bounds__(256)
prep_weights_tc_kernel(Pass p, __nv_bfloat16 *prep, int backward, const int32_t *model_active, int grouped,
                       int part0, int off_w2, int off_w3)
{
    extern __shared__ __align__(16) float wsrc[];                 // [16][PREP_ROW]
    const int m = blockIdx.y, part = part0 + blockIdx.x, t = threadIdx.x;
    if (model_active && model_active[m] == 0) return;
    const bool l3 = part >= 2;
    const int co0 = (l3 ? part - 2 : part) * 16, nco = l3 ? 64 : 32;
    const float *row = p.X + (int64_t)p.mrow[m] * p.ldx + (l3 ? off_w3 : off_w2) + co0 * 800;
    __nv_bfloat16 *dst_u = prep + (int64_t)m * TC_PREP_ELEMS + (l3 ? TCO_W3U : TCO_W2U);
    __nv_bfloat16 *dst_d = prep + (int64_t)m * TC_PREP_ELEMS + (l3 ? TCO_W3UD : TCO_W2UD);
    for (int i = t; i < 16 * 200; i += 256) {                     // 16 rows of 800 floats, float4 loads
        const int r = i / 200, c4 = i - r * 200;
        const float4 v = __ldg(reinterpret_cast<const float4 *>(row + r * 800) + c4);
        float *d = wsrc + r * PREP_ROW + c4 * 4;
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    }
    __syncthreads();
    const bool do_d = !l3 || backward;
    for (int i = t; i < (do_d ? 3200 : 1600); i += 256) {
        float v[8];
        int64_t entry;
        __nv_bfloat16 *dst;
        if (i < 1600) {
            const int nl = i & 15, kc = i >> 4, tap = kc >> 2, ci0 = (kc & 3) * 8;
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = wsrc[nl * PREP_ROW + (ci0 + e) * 25 + tap];
            entry = grouped ? grouped_entry(tap, kc & 3, 4, nco, co0 + nl) : (int64_t)kc * nco + co0 + nl;
            dst = dst_u;
        } else {
            const int j = i - 1600, n = j & 31, tg = j >> 5, cg = tg & 1, tap = tg >> 1;
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = wsrc[(cg * 8 + e) * PREP_ROW + n * 25 + (24 - tap)];
            entry = grouped ? grouped_entry(tap, co0 / 8 + cg, nco / 8, 32, n)
                            : (int64_t)(tap * (nco / 8) + co0 / 8 + cg) * 32 + n;
            dst = dst_d;
        }
        *reinterpret_cast<uint4 *>(dst + entry * 8) = pack_bf16x8(make_float4(v[0], v[1], v[2], v[3]),
                                                                  make_float4(v[4], v[5], v[6], v[7]));
    }
}

// Staging loop with U independent 32-byte global loads in flight per thread (the copies are
// latency bound otherwise).  `loc(i, src, dst)` maps item i to its fp32 source (nullptr = zeros) and
// its 16-byte shared-memory destination.
template <int U, typename Loc>
__device__ __forceinline__ void stage_items(int first, int n_items, int stride, Loc loc, int dup_off = 0,
                                            int dup_off2 = 0, int dup_off3 = 0) {
    for (int base = first; base < n_items; base += U * stride) {
        float4 a[U], b[U];
        uint8_t *dst[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * stride;
            dst[u] = nullptr;
            a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            b[u] = a[u];
            if (i < n_items) {
                const float *src;
                loc(i, src, dst[u]);
                if (src) {                                        // one 256-bit load: a full 32-byte sector per lane
                    float r[8];
                    tc::ld_global_nc_v8(src, r);
                    a[u] = make_float4(r[0], r[1], r[2], r[3]);
                    b[u] = make_float4(r[4], r[5], r[6], r[7]);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (dst[u]) {
                const uint4 v = pack_bf16x8(a[u], b[u]);
                *reinterpret_cast<uint4 *>(dst[u]) = v;
                if (dup_off) *reinterpret_cast<uint4 *>(dst[u] + dup_off) = v;     // shifted copies
                if (dup_off2) *reinterpret_cast<uint4 *>(dst[u] + dup_off2) = v;
                if (dup_off3) *reinterpret_cast<uint4 *>(dst[u] + dup_off3) = v;
            }
    }
}

// Padded plane geometry.  The pads are SHARED: a row's two right pad pixels are the next row's two left
// pad pixels (pitch H + 2), and a sample's two bottom pad rows are the next sample's two top pad rows
// (sample stride (H + 2) rows).  All pads are zero, so every shifted window still reads zeros outside
// the image, while the origins (window top-left corners) of a sample shrink from (H+4)^2 to (H-1)*PP + H:
// 253 for H = 15 — two M = 128 tiles per sample — and 61 for H = 7.
template <int H>
struct PlaneGeom {
    static constexpr int PP = H + 2;                       // row pitch
    static constexpr int NPS = (H + 2) * PP;               // sample stride (pixels)
    static constexpr int NORIG = (H - 1) * PP + H;         // origins of one sample
    static constexpr int EXTENT = (H + 4) * PP + 2;        // pixels a sample's windows can touch
    static constexpr int TAIL = 4 * PP + 4;                // furthest tap shift
};

template <int CIN, int COUT, int H, int SB>
struct ConvTcCfg {
    using G = PlaneGeom<H>;
    static constexpr int PP = G::PP, NPS = G::NPS, NORIG = G::NORIG, TAIL = G::TAIL;
    // M tiles of a chunk: either every sample gets its own ceil(NORIG / 128) tiles, or the chunk's origins
    // are enumerated contiguously across samples — whichever needs fewer tiles
    static constexpr int TPS = (NORIG + 127) / 128;
    static constexpr int TILES_GLOBAL = ((SB - 1) * NPS + NORIG + 127) / 128;
    static constexpr bool PER_SAMPLE = SB * TPS <= TILES_GLOBAL;
    static constexpr int REACH = PER_SAMPLE ? (SB - 1) * NPS + TPS * 128 + TAIL : TILES_GLOBAL * 128 + TAIL;
    static constexpr int FULL = (SB - 1) * NPS + G::EXTENT;
    static constexpr int NPIX = (((REACH > FULL ? REACH : FULL) + 7) / 8) * 8;
    static constexpr int PLANE = NPIX * 16;                // bytes
    static constexpr int NCH = CIN / 8;
    static constexpr int KCH = 25 * NCH;                   // 8-element K chunks
    static constexpr int B_BYTES = KCH * COUT * 16;
    static constexpr int A_BYTES = NCH * PLANE;
    static constexpr int SMEM = B_BYTES + 2 * A_BYTES + COUT * 4 + 128;
    static constexpr int TMEM_COLS = (2 * COUT <= 64) ? 64 : 128;
    __host__ __device__ static constexpr int tiles(int ns) {
        return PER_SAMPLE ? ns * TPS : ((ns - 1) * NPS + NORIG + 127) / 128;
    }
    __host__ __device__ static constexpr int tile_start(int tile) {      // first origin (pixel index) of a tile
        return PER_SAMPLE ? (tile / TPS) * NPS + (tile % TPS) * 128 : tile * 128;
    }
};

// Warp-specialised, persistent over the samples of a node (or half a node):
//   warps 0-3  epilogue: TMEM -> registers -> (+bias) -> global, one TMEM lane quadrant each
//   warp  4    MMA issuer (one elected lane)
//   warps 5-7  producers: fp32 NHWC activations -> bf16 planes of chunk c+1 while chunk c is consumed
// Shared memory: the node's weights (loaded once per CTA) + two activation buffers of SB samples.
// mbarriers: full[b]/empty[b] per activation buffer, tfull[t]/tempty[t] per TMEM accumulator buffer.
constexpr int CT_THREADS = 256;
constexpr int CT_PRODUCERS = CT_THREADS - 160;

template <int CIN, int COUT, int H, int SB, bool BIAS>
__global__ void __launch_bounds__(CT_THREADS, 1)
conv_tc_kernel(const float *__restrict__ in, float *__restrict__ out, Pass p,
               const __nv_bfloat16 *__restrict__ wprep, int wprep_off, int bias_off, int spc)
{
    using C = ConvTcCfg<CIN, COUT, H, SB>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *sB = smem;
    uint8_t *sA0 = smem + C::B_BYTES;
    float *sbias = reinterpret_cast<float *>(sA0 + 2 * C::A_BYTES);
    uint64_t *bars = reinterpret_cast<uint64_t *>(sbias + COUT);       // full[2] empty[2] tfull[2] tempty[2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 8);

    const int t = threadIdx.x, warp = tc::uniform_warp_idx();
    const int parts = (p.gs + spc - 1) / spc;
    const int g = blockIdx.x / parts, part = blockIdx.x - g * parts;
    if (p.gcount[g] <= 0) return;
    const int s_begin = part * spc;
    int s_end = s_begin + spc;
    if (s_end > p.gs) s_end = p.gs;
    // empty slots (img_idx < 0) sit at the tail of a group: stop at the last active sample
    while (s_end > s_begin && p.img_idx[g * p.gs + s_end - 1] < 0) --s_end;
    if (s_end <= s_begin) return;
    const int n_chunks = (s_end - s_begin + SB - 1) / SB;
    int m;
    const float *row = model_row(p, g, m);

    {
        const uint4 *src = reinterpret_cast<const uint4 *>(wprep + (int64_t)m * TC_PREP_ELEMS + wprep_off);
        uint4 *dst = reinterpret_cast<uint4 *>(sB);
        for (int i = t; i < C::B_BYTES / 16; i += CT_THREADS) dst[i] = __ldg(src + i);
        uint4 *za = reinterpret_cast<uint4 *>(sA0);
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        for (int i = t; i < 2 * C::A_BYTES / 16; i += CT_THREADS) za[i] = z;
        if (t < COUT) sbias[t] = BIAS ? row[bias_off + t] : 0.f;
    }
    if (t == 0) {
        tc::mbar_init(&bars[0], CT_PRODUCERS); tc::mbar_init(&bars[1], CT_PRODUCERS);
        tc::mbar_init(&bars[2], 1); tc::mbar_init(&bars[3], 1);
        tc::mbar_init(&bars[4], 1); tc::mbar_init(&bars[5], 1);
        tc::mbar_init(&bars[6], 128); tc::mbar_init(&bars[7], 128);
        tc::mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    auto chunk_samples = [&](int c) { const int r = (s_end - s_begin) - c * SB; return r < SB ? r : SB; };
    auto chunk_tiles = [&](int c) { return C::tiles(chunk_samples(c)); };

    if (warp >= 5) {
        // ------------------------------ producers ------------------------------
        const int pt = t - 160;
        for (int c = 0; c < n_chunks; ++c) {
            const int b = c & 1;
            if (c >= 2) tc::mbar_wait(&bars[2 + b], (uint32_t)(((c >> 1) - 1) & 1));
            uint8_t *sA = sA0 + b * C::A_BYTES;
            const int ns = chunk_samples(c), slot0 = g * p.gs + s_begin + c * SB;
            stage_items<8>(pt, ns * H * H * C::NCH, CT_PRODUCERS, [&](int i, const float *&src, uint8_t *&dst) {
                const int ch = i % C::NCH, q = i / C::NCH, s = q / (H * H), pix = q - s * (H * H);
                const int y = pix / H, x = pix - y * H;
                src = in + (int64_t)(slot0 + s) * (H * H * CIN) + pix * CIN + ch * 8;
                dst = sA + ch * C::PLANE + (s * C::NPS + (y + 2) * C::PP + (x + 2)) * 16;
            });
            tc::fence_async_smem();
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(tc::smem_u32(&bars[b])) : "memory");
        }
    } else if (warp == 4) {
        // ------------------------------ MMA issuer ------------------------------
        // The whole warp walks the loop (warp-uniform control flow); one elected lane issues.
        const bool leader = tc::elect_one();
        const uint32_t idesc = tc::instr_desc_bf16(128, COUT, 0, 0);
        const uint64_t bd0 = tc::smem_desc(tc::smem_u32(sB), COUT * 16, 128);
        int gt = 0;
        for (int c = 0; c < n_chunks; ++c) {
            const int b = c & 1;
            tc::mbar_wait(&bars[b], (uint32_t)((c >> 1) & 1));
            tc::fence_after_sync();
            const uint32_t a_base = tc::smem_u32(sA0 + b * C::A_BYTES);
            const int nt = chunk_tiles(c);
            for (int tile = 0; tile < nt; ++tile, ++gt) {
                const int tb = gt & 1;
                if (gt >= 2) { tc::mbar_wait(&bars[6 + tb], (uint32_t)(((gt >> 1) - 1) & 1)); tc::fence_after_sync(); }
                const uint32_t d = tmem + (uint32_t)(tb * COUT);
                const uint64_t ad_tile = tc::smem_desc(a_base + (uint32_t)C::tile_start(tile) * 16u, C::PLANE, 128);
                if (leader) {
#pragma unroll
                    for (int tap = 0; tap < 25; ++tap) {
#pragma unroll
                        for (int h = 0; h < CIN / 16; ++h) {
                            const uint64_t ad = ad_tile + (uint64_t)((((tap / 5) * C::PP + (tap % 5)) * 16 + 2 * h * C::PLANE) >> 4);
                            const uint64_t bd = bd0 + (uint64_t)(((tap * C::NCH + 2 * h) * (COUT * 16)) >> 4);
                            tc::mma_f16(d, ad, bd, idesc, (tap | h) != 0 ? 1u : 0u);
                        }
                    }
                    tc::mma_commit(&bars[4 + tb]);
                }
                __syncwarp();
            }
            if (leader) tc::mma_commit(&bars[2 + b]);      // activation buffer b is free once these MMAs finish
            __syncwarp();
        }
    } else {
        // ------------------------------ epilogue (warps 0-3) ------------------------------
        int gt = 0;
        for (int c = 0; c < n_chunks; ++c) {
            const int ns = chunk_samples(c), nt = chunk_tiles(c), slot0 = g * p.gs + s_begin + c * SB;
            for (int tile = 0; tile < nt; ++tile, ++gt) {
                const int tb = gt & 1;
                tc::mbar_wait(&bars[4 + tb], (uint32_t)((gt >> 1) & 1));
                tc::fence_after_sync();
                uint32_t r[COUT];
#pragma unroll
                for (int c0 = 0; c0 < COUT; c0 += 32)
                    tc::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(tb * COUT + c0),
                                  *reinterpret_cast<uint32_t(*)[32]>(&r[c0]));
                tc::tmem_ld_wait();
                tc::fence_before_sync();
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(tc::smem_u32(&bars[6 + tb])) : "memory");
                const int o = C::tile_start(tile) + t;      // origin handled by this thread (TMEM lane t)
                const int s = o / C::NPS, loc = o - s * C::NPS, oy = loc / C::PP, ox = loc - oy * C::PP;
                if (s < ns && oy < H && ox < H) {
                    float *dst = out + (int64_t)(slot0 + s) * (H * H * COUT) + (oy * H + ox) * COUT;
#pragma unroll
                    for (int j = 0; j < COUT; j += 8)
                        tc::st_global_v8(dst + j, __uint_as_float(r[j]) + sbias[j], __uint_as_float(r[j + 1]) + sbias[j + 1],
                                         __uint_as_float(r[j + 2]) + sbias[j + 2], __uint_as_float(r[j + 3]) + sbias[j + 3],
                                         __uint_as_float(r[j + 4]) + sbias[j + 4], __uint_as_float(r[j + 5]) + sbias[j + 5],
                                         __uint_as_float(r[j + 6]) + sbias[j + 6], __uint_as_float(r[j + 7]) + sbias[j + 7]);
                }
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, C::TMEM_COLS);
}

// ------------------------------------------------------------------------------------------
// Grouped-tap form of the same convolution (the default).  A tcgen05.mma with N <= 32 costs 44 SM cycles
// whatever it computes (the 4 KB A window is read from shared memory at 128 B/clk), N = 64 costs 48 and
// N = 96 only 56 (scripts/probe_mma_rate.py), so the taps of a kernel row are stacked along N: the weights of
// taps kx = 0,1,2 form one B operand of 3*COUT columns and those of kx = 3,4 one of 2*COUT columns.  All
// column blocks of an MMA see the SAME shifted window, so block j of the accumulator holds the tap's product
// for the window one/two pixels to the LEFT of where the output needs it:
//     out[q] = D[q][block 0] + D[q + 1][block 1] + D[q + 2][block 2]
// (block j of the second MMA, whose window starts at kx = 3, accumulates into the same columns: tap 3 has
// displacement 0, tap 4 displacement 1).  Per tile 5 x CIN/16 x 2 MMAs (104 cycles per pair) replace
// 25 x CIN/16 (5 x 44 = 220 cycles): 2.1x fewer tensor-pipe cycles for COUT = 32.  The displaced sum is two
// warp shuffles per value in the epilogue; the two lanes at the end of a warp take their neighbours from the
// next warp through a 384-byte exchange in shared memory, and the last two origins of a tile are completed by
// warp 0 when the next tile (the following 128 origins) arrives.
// ------------------------------------------------------------------------------------------
template <int CIN, int COUT, int H, int SB, int NST>
struct ConvTc3Cfg : ConvTcCfg<CIN, COUT, H, SB> {
    using Base = ConvTcCfg<CIN, COUT, H, SB>;
    static constexpr int NB = 2 * COUT;                               // accumulator columns of one tile
    static constexpr int TB_STRIDE = 2 * COUT;                         // column stride between the two accumulators
    static constexpr int TMEM_COLS = 2 * TB_STRIDE;
    static constexpr int NCG = COUT / 32;                              // channel groups of 32
    static constexpr int RING = 8;                                     // boundary ring: tiles in flight between epilogue and finisher
    static constexpr int RQ = 64;                                      // floats per (quadrant, channel group): r1(lane 0), partial(lane 31)
    static constexpr int RING_FLOATS = RING * 4 * NCG * RQ;
    static constexpr int SMEM = Base::B_BYTES + NST * Base::A_BYTES + RING_FLOATS * 4 + 320;
    static constexpr int RUN = Base::NORIG;                            // pixels of a sample's plane run
    static constexpr int RUNP = plane_runp(H);                         // plane stride (pixels) in global memory
    static_assert(SMEM <= 227 * 1024, "shared memory budget");
    static_assert(Base::TPS * 128 >= Base::NORIG + 1 || !Base::PER_SAMPLE, "last valid origin needs its neighbour in the tile");
    static_assert(Base::TILES_GLOBAL * 128 >= (SB - 1) * Base::NPS + Base::NORIG + 1 || Base::PER_SAMPLE,
                  "last valid origin needs its neighbour in the tile");
};

// warps 0-3 / 8-11: the two epilogue groups, 4: MMA issuer, 5-7: producers (5: bulk copies), 12: boundary finisher
constexpr int CT3_THREADS = 416;
constexpr int CT3_PRODUCERS = 96;

// PLANES: the input is already a set of bf16 channel-chunk planes in global memory (`in_planes`, written by the
// producing kernel); one elected thread of warp 5 stages a chunk with bulk asynchronous copies (TMA engine, one
// per sample and plane, completing on the stage's mbarrier) NST - 1 chunks ahead.  Otherwise three producer
// warps convert fp32 NHWC activations (`in`) as in conv_tc_kernel.
//
// The output is the convolution WITHOUT its bias (the consumers, the GroupNorm kernels, add it on load).
//
// Epilogue.  TMEM is read at 64 B/clk and shared memory is saturated by the MMAs' operand reads, so the epilogue
// warps touch shared memory only with stores: a quadrant's lanes 0..29 finish their displaced sums with warp
// shuffles and store their pixels; what the boundary pixels (lanes 30, 31) need — their partial sums and the
// next quadrant's r2(lane 0), r2(lane 1), r1(lane 0) — goes into a ring in shared memory, and a separate
// FINISHER warp completes those pixels one channel per lane, off the critical path.
template <int CIN, int COUT, int H, int SB, bool PLANES, int NST>
__global__ void __launch_bounds__(CT3_THREADS, 1)
conv_tc3_kernel(const float *__restrict__ in, const uint16_t *__restrict__ in_planes, float *__restrict__ out, Pass p,
                const __nv_bfloat16 *__restrict__ wprep, int wprep_off, int spc)
{
    using C = ConvTc3Cfg<CIN, COUT, H, SB, NST>;
    constexpr int NCK = CIN / 8;                                       // 8-element K chunks per tap
    constexpr int NCG = C::NCG, RING = C::RING, RQ = C::RQ;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *sB = smem;
    uint8_t *sA0 = smem + C::B_BYTES;
    float *ring = reinterpret_cast<float *>(sA0 + NST * C::A_BYTES);   // [RING][4 quadrants][NCG][RQ]
    uint64_t *bars = reinterpret_cast<uint64_t *>(ring + C::RING_FLOATS);
    uint64_t *full = bars, *empty = bars + NST, *tfull = bars + 2 * NST, *tempty = tfull + 2;
    uint64_t *rfull = tempty + 2, *rempty = rfull + RING, *wbar = rempty + RING;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(wbar + 1);

    const int t = threadIdx.x, warp = tc::uniform_warp_idx(), lane = t & 31;
    const int parts = (p.gs + spc - 1) / spc;
    const int g = blockIdx.x / parts, part = blockIdx.x - g * parts;
    if (p.gcount[g] <= 0) return;
    const int s_begin = part * spc;
    int s_end = s_begin + spc;
    if (s_end > p.gs) s_end = p.gs;
    while (s_end > s_begin && p.img_idx[g * p.gs + s_end - 1] < 0) --s_end;
    if (s_end <= s_begin) return;
    const int n_chunks = (s_end - s_begin + SB - 1) / SB;
    int m;
    model_row(p, g, m);

    {
        uint4 *za = reinterpret_cast<uint4 *>(sA0);
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        for (int i = t; i < NST * C::A_BYTES / 16; i += CT3_THREADS) za[i] = z;
    }
    if (t == 0) {
        tc::mbar_init(wbar, 1);
        for (int i = 0; i < NST; ++i) { tc::mbar_init(&full[i], PLANES ? 1 : CT3_PRODUCERS); tc::mbar_init(&empty[i], 1); }
        tc::mbar_init(&tfull[0], 1); tc::mbar_init(&tfull[1], 1);
        tc::mbar_init(&tempty[0], 128); tc::mbar_init(&tempty[1], 128);
        for (int i = 0; i < RING; ++i) { tc::mbar_init(&rfull[i], 4); tc::mbar_init(&rempty[i], 1); }
        tc::mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    auto chunk_samples = [&](int c) { const int r = (s_end - s_begin) - c * SB; return r < SB ? r : SB; };
    auto chunk_tiles = [&](int c) { return C::tiles(chunk_samples(c)); };
    // output row of origin o for channel group cg; nullptr when o is a padding origin
    auto pixel_ptr = [&](int o, int ns, int slot0, int cg) -> float * {
        const int s = o / C::NPS, loc = o - s * C::NPS, oy = loc / C::PP, ox = loc - oy * C::PP;
        return (s < ns && oy < H && ox < H) ? out + (int64_t)(slot0 + s) * (H * H * COUT) + (oy * H + ox) * COUT + cg * 32
                                            : nullptr;
    };

    if (warp >= 5 && warp < 8) {
        // ------------------------------ producers ------------------------------
        if (PLANES) {
            if (warp == 5 && tc::elect_one()) {
                for (int c = 0; c < n_chunks; ++c) {
                    const int b = c % NST;
                    if (c >= NST) tc::mbar_wait(&empty[b], (uint32_t)(((c / NST) - 1) & 1));
                    const int ns = chunk_samples(c), slot0 = g * p.gs + s_begin + c * SB;
                    const uint32_t dst0 = tc::smem_u32(sA0 + b * C::A_BYTES) + (2 * C::PP + 2) * 16;
                    tc::mbar_expect_tx(&full[b], (uint32_t)(ns * C::NCH * C::RUN * 16));
                    for (int s = 0; s < ns; ++s) {
                        const uint16_t *src = in_planes + (int64_t)(slot0 + s) * (C::NCH * C::RUNP * 8);
#pragma unroll
                        for (int ch = 0; ch < C::NCH; ++ch)
                            tc::bulk_g2s(dst0 + (uint32_t)(ch * C::PLANE + s * C::NPS * 16), src + ch * C::RUNP * 8,
                                         C::RUN * 16, &full[b]);
                    }
                }
            }
        } else {
            const int pt = t - 160;
            for (int c = 0; c < n_chunks; ++c) {
                const int b = c % NST;
                if (c >= NST) tc::mbar_wait(&empty[b], (uint32_t)(((c / NST) - 1) & 1));
                uint8_t *sA = sA0 + b * C::A_BYTES;
                const int ns = chunk_samples(c), slot0 = g * p.gs + s_begin + c * SB;
                stage_items<8>(pt, ns * H * H * C::NCH, CT3_PRODUCERS, [&](int i, const float *&src, uint8_t *&dst) {
                    const int ch = i % C::NCH, q = i / C::NCH, s = q / (H * H), pix = q - s * (H * H);
                    const int y = pix / H, x = pix - y * H;
                    src = in + (int64_t)(slot0 + s) * (H * H * CIN) + pix * CIN + ch * 8;
                    dst = sA + ch * C::PLANE + (s * C::NPS + (y + 2) * C::PP + (x + 2)) * 16;
                });
                tc::fence_async_smem();
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(tc::smem_u32(&full[b])) : "memory");
            }
        }
    } else if (warp == 4) {
        // ------------------------------ MMA issuer ------------------------------
        const bool leader = tc::elect_one();
        if (leader) {                                                  // the node's weights: one bulk copy (TMA engine)
            tc::mbar_expect_tx(wbar, (uint32_t)C::B_BYTES);
            tc::bulk_g2s(tc::smem_u32(sB), wprep + (int64_t)m * TC_PREP_ELEMS + wprep_off, (uint32_t)C::B_BYTES, wbar);
        }
        __syncwarp();
        tc::mbar_wait(wbar, 0);
        const uint32_t idesc2 = tc::instr_desc_bf16(128, 2 * COUT, 0, 0);
        const uint32_t idesc1 = tc::instr_desc_bf16(128, COUT, 0, 0);
        const uint64_t bdA = tc::smem_desc(tc::smem_u32(sB), 2 * COUT * 16, 128);                           // kx = 0, 1
        const uint64_t bdB = tc::smem_desc(tc::smem_u32(sB) + NCK * 2 * COUT * 16, 2 * COUT * 16, 128);     // kx = 2, 3
        const uint64_t bdC = tc::smem_desc(tc::smem_u32(sB) + NCK * 4 * COUT * 16, COUT * 16, 128);         // kx = 4
        int gt = 0;
        for (int c = 0; c < n_chunks; ++c) {
            const int b = c % NST;
            tc::mbar_wait(&full[b], (uint32_t)((c / NST) & 1));
            tc::fence_after_sync();
            const uint32_t a_base = tc::smem_u32(sA0 + b * C::A_BYTES);
            const int nt = chunk_tiles(c);
            for (int tile = 0; tile < nt; ++tile, ++gt) {
                const int tb = gt & 1;
                if (gt >= 2) { tc::mbar_wait(&tempty[tb], (uint32_t)(((gt >> 1) - 1) & 1)); tc::fence_after_sync(); }
                const uint32_t d = tmem + (uint32_t)(tb * C::TB_STRIDE);
                const uint64_t ad_tile = tc::smem_desc(a_base + (uint32_t)C::tile_start(tile) * 16u, C::PLANE, 128);
                if (leader) {
#pragma unroll
                    for (int ky = 0; ky < 5; ++ky) {
#pragma unroll
                        for (int h = 0; h < CIN / 16; ++h) {
                            const uint64_t ad0 = ad_tile + (uint64_t)(((ky * C::PP) * 16 + 2 * h * C::PLANE) >> 4);   // window of kx = 0
                            const uint32_t brow = (uint32_t)(ky * (NCK * 5 * COUT) * 16);   // bytes of one kernel row
                            const uint64_t oA = (uint64_t)((brow + 2 * h * (2 * COUT) * 16) >> 4);
                            const uint64_t oC = (uint64_t)((brow + 2 * h * COUT * 16) >> 4);
                            tc::mma_f16(d, ad0, bdA + oA, idesc2, (ky | h) != 0 ? 1u : 0u);
                            tc::mma_f16(d, ad0 + 2, bdB + oA, idesc2, 1u);                  // window of kx = 2: +2 pixels = +32 B
                            tc::mma_f16(d, ad0 + 4, bdC + oC, idesc1, 1u);                  // kx = 4
                        }
                    }
                    tc::mma_commit(&tfull[tb]);
                }
                __syncwarp();
            }
            if (leader) tc::mma_commit(&empty[b]);
            __syncwarp();
        }
    } else if (warp == 12) {
        // ------------------------------ boundary finisher ------------------------------
        // lane = channel.  For tile gt: the pixels of lanes 30 / 31 of quadrants 0-2, and (when tile gt continues
        // tile gt - 1) the previous tile's quadrant-3 pixels, which need this tile's quadrant-0 vectors.
        int gt = 0;
        for (int c = 0; c < n_chunks; ++c) {
            const int ns = chunk_samples(c), nt = chunk_tiles(c), slot0 = g * p.gs + s_begin + c * SB;
            for (int tile = 0; tile < nt; ++tile, ++gt) {
                const int sl = gt % RING;
                tc::mbar_wait(&rfull[sl], (uint32_t)((gt / RING) & 1));
                const bool cont = tile > 0 && C::tile_start(tile) == C::tile_start(tile - 1) + 128;
                const float *rs = ring + sl * (4 * NCG * RQ);
                const float *rp = ring + ((gt + RING - 1) % RING) * (4 * NCG * RQ);       // previous tile's slot
                // all shared-memory loads of the tile first (their latency is long while the MMAs stream operands),
                // then the sums and the stores
                float va[NCG][4], vb[NCG][4];
#pragma unroll
                for (int cg = 0; cg < NCG; ++cg) {
#pragma unroll
                    for (int q = 0; q < 3; ++q) {
                        va[cg][q] = rs[(q * NCG + cg) * RQ + 32 + lane];
                        vb[cg][q] = rs[((q + 1) * NCG + cg) * RQ + lane];
                    }
                    va[cg][3] = cont ? rp[(3 * NCG + cg) * RQ + 32 + lane] : 0.f;
                    vb[cg][3] = rs[(0 * NCG + cg) * RQ + lane];
                }
#pragma unroll
                for (int cg = 0; cg < NCG; ++cg) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        if (q == 3 && !cont) continue;
                        const int o = q < 3 ? C::tile_start(tile) + q * 32 + 31 : C::tile_start(tile) - 1;
                        float *d31 = pixel_ptr(o, ns, slot0, cg);
                        if (d31) d31[lane] = va[cg][q] + vb[cg][q];
                    }
                }
                __syncwarp();
                // slots are released in order: the previous tile's once this tile has used it, and this tile's own
                // right away when it is the last of its chunk (its quadrant-3 boundary pixels are padding)
                if (lane == 0) {
                    if (tile > 0) tc::mbar_arrive(&rempty[(gt + RING - 1) % RING]);
                    if (tile == nt - 1) tc::mbar_arrive(&rempty[sl]);
                }
            }
        }
    } else {
        // ------------------------------ epilogue: two groups of four warps ------------------------------
        // Group 0 (warps 0-3) drains the tiles with even running index (TMEM accumulator 0), group 1 (warps 8-11)
        // the odd ones.  Lane quadrant = warp % 4.
        const int grp = warp >> 3, wq = warp & 3;
        int gt = 0;
        for (int c = 0; c < n_chunks; ++c) {
            const int ns = chunk_samples(c), nt = chunk_tiles(c), slot0 = g * p.gs + s_begin + c * SB;
            for (int tile = 0; tile < nt; ++tile, ++gt) {
                const int tb = gt & 1;
                if (tb != grp) continue;
                const int k = gt >> 1, sl = gt % RING;
                tc::mbar_wait(&tfull[tb], (uint32_t)(k & 1));
                tc::fence_after_sync();
                bool ring_ok = gt < RING;                              // slot free? (waited for lazily, once per tile)
#pragma unroll
                for (int cg = 0; cg < NCG; ++cg) {
                    uint32_t r0[32], r1[32];
                    const uint32_t ta = tmem + ((uint32_t)(wq * 32) << 16) + (uint32_t)(tb * C::TB_STRIDE + cg * 32);
                    tc::tmem_ld32(ta, r0);
                    tc::tmem_ld32(ta + COUT, r1);
                    tc::tmem_ld_wait();
                    if (cg == NCG - 1) {                           // the accumulator may be overwritten now
                        tc::fence_before_sync();
                        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(tc::smem_u32(&tempty[tb])) : "memory");
                    }
                    // displaced sum inside the warp: one shuffle per value (lane 31's is skipped by the predicate)
                    float sum[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        float acc = __uint_as_float(r0[j]);
                        asm volatile("{\n\t.reg .pred p;\n\t.reg .f32 v;\n\t"
                                     "shfl.sync.down.b32 v|p, %1, 1, 0x1f, 0xffffffff;\n\t@p add.f32 %0, %0, v;\n\t}"
                                     : "+f"(acc) : "f"(__uint_as_float(r1[j])));
                        sum[j] = acc;
                    }
                    if (lane == 0 || lane == 31) {
                        if (!ring_ok) { tc::mbar_wait(&rempty[sl], (uint32_t)(((gt / RING) - 1) & 1)); ring_ok = true; }
                        float4 *rq = reinterpret_cast<float4 *>(ring + ((sl * 4 + wq) * NCG + cg) * RQ);
                        if (lane == 0) {
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                rq[j] = make_float4(__uint_as_float(r1[4 * j]), __uint_as_float(r1[4 * j + 1]),
                                                    __uint_as_float(r1[4 * j + 2]), __uint_as_float(r1[4 * j + 3]));
                        } else {
#pragma unroll
                            for (int j = 0; j < 8; ++j) rq[8 + j] = make_float4(sum[4 * j], sum[4 * j + 1], sum[4 * j + 2], sum[4 * j + 3]);
                        }
                    }
                    if (lane < 31) {
                        float *dst = pixel_ptr(C::tile_start(tile) + wq * 32 + lane, ns, slot0, cg);
                        if (dst) {
#pragma unroll
                            for (int j = 0; j < 32; j += 8)
                                tc::st_global_v8(dst + j, sum[j], sum[j + 1], sum[j + 2], sum[j + 3], sum[j + 4], sum[j + 5],
                                                 sum[j + 6], sum[j + 7]);
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&rfull[sl]);             // this quadrant's boundary data of the tile is in the ring
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, C::TMEM_COLS);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
template <int CIN, int COUT, int H, int SB, bool BIAS, int NSTP = 3>
static int launch_conv_tc(const float *in, const uint16_t *in_planes, float *out, const Pass &p, const __nv_bfloat16 *wprep,
                          int wprep_off, int bias_off, cudaStream_t st)
{
    using C = ConvTcCfg<CIN, COUT, H, SB>;
    // a CTA serves a whole node when there are enough nodes to fill the GPU, else half a node
    int spc = p.gs;
    if (p.n_groups < num_sms() && p.gs > SB) spc = ((p.gs / 2 + SB - 1) / SB) * SB;
    const int parts = (p.gs + spc - 1) / spc;
    if (!tune_value(TUNE_CONV_UNGROUPED)) {
        if (in_planes) {
            constexpr int NST = NSTP;
            using C3 = ConvTc3Cfg<CIN, COUT, H, SB, NST>;
            static PerDeviceOnce attr3p;
            int rc3 = ensure_dyn_smem(conv_tc3_kernel<CIN, COUT, H, SB, true, NST>, C3::SMEM, attr3p);
            if (rc3) return rc3;
            conv_tc3_kernel<CIN, COUT, H, SB, true, NST><<<p.n_groups * parts, CT3_THREADS, C3::SMEM, st>>>(
                in, in_planes, out, p, wprep, wprep_off, spc);
            DLS_LAUNCH_CHECK();
            return 0;
        }
        using C3 = ConvTc3Cfg<CIN, COUT, H, SB, 2>;
        static PerDeviceOnce attr3;
        int rc3 = ensure_dyn_smem(conv_tc3_kernel<CIN, COUT, H, SB, false, 2>, C3::SMEM, attr3);
        if (rc3) return rc3;
        conv_tc3_kernel<CIN, COUT, H, SB, false, 2><<<p.n_groups * parts, CT3_THREADS, C3::SMEM, st>>>(
            in, nullptr, out, p, wprep, wprep_off, spc);
        DLS_LAUNCH_CHECK();
        return 0;
    }
    static PerDeviceOnce attr;
    int rc = ensure_dyn_smem(conv_tc_kernel<CIN, COUT, H, SB, BIAS>, C::SMEM, attr);
    if (rc) return rc;
    conv_tc_kernel<CIN, COUT, H, SB, BIAS><<<p.n_groups * parts, CT_THREADS, C::SMEM, st>>>(in, out, p, wprep, wprep_off,
                                                                                   bias_off, spc);
    DLS_LAUNCH_CHECK();
    return 0;
}

int prep_weights_tc(const Pass &p, Workspace &ws, bool backward, cudaStream_t st)
{
    static PerDeviceOnce attr;
    int rc = ensure_dyn_smem(prep_weights_tc_kernel, PREP_SMEM, attr);
    if (rc) return rc;
    dim3 grid(6, p.n_models);
    const int32_t *active = p.gmodel ? nullptr : p.gcount;
    prep_weights_tc_kernel<<<grid, 256, PREP_SMEM, st>>>(p, reinterpret_cast<__nv_bfloat16 *>(ws.prep_tc), backward ? 1 : 0,
                                                 active, tune_value(TUNE_CONV_UNGROUPED) ? 0 : 1, 0, OFF_W2, OFF_W3);
    DLS_LAUNCH_CHECK();
    if (ws.zplane) {                              // a plane run of zeros (dY of the missing samples of a partial chunk)
        cudaError_t e = cudaMemsetAsync(ws.zplane, 0, RUNP15 * 16, st);
        if (e != cudaSuccess) return (int)e;
    }
    return prep_w1_tc(p, ws, st);
}

int conv2_fwd_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    return launch_conv_tc<32, 32, 15, 1, true>(ws.a1, tune_value(TUNE_CONV_UNGROUPED) ? nullptr : ws.a1p, ws.c2, p, reinterpret_cast<const __nv_bfloat16 *>(ws.prep_tc),
                                               TCO_W2U, OFF_B2, st);
}
int conv3_fwd_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    return launch_conv_tc<32, 64, 7, 3, true>(ws.a2, tune_value(TUNE_CONV_UNGROUPED) ? nullptr : ws.a2p, ws.c3, p, reinterpret_cast<const __nv_bfloat16 *>(ws.prep_tc),
                                              TCO_W3U, OFF_B3, st);
}
int conv3_dgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    return launch_conv_tc<64, 32, 7, 3, false>(ws.dc3, tune_value(TUNE_CONV_UNGROUPED) ? nullptr : ws.dc3p, ws.da2, p, reinterpret_cast<const __nv_bfloat16 *>(ws.prep_tc),
                                               TCO_W3UD, 0, st);
}
int conv2_dgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    return launch_conv_tc<32, 32, 15, 1, false>(ws.dc2, tune_value(TUNE_CONV_UNGROUPED) ? nullptr : ws.dc2p, ws.da1, p, reinterpret_cast<const __nv_bfloat16 *>(ws.prep_tc),
                                                TCO_W2UD, 0, st);
}


// ==========================================================================================
// Weight gradient on tensor cores.
//   dW[co][ci][tap] = sum over samples s and output pixels q of dY[s][q][co] * in[s][q + shift(tap)][ci]
// As a GEMM per tap: D[co, ci] = A^T-form with the PIXEL index as the reduction (K) dimension.
// Both operands are staged as bf16 channel-chunk planes  plane[c/8][pixel][8]  in which the M/N
// index (channel) is the contiguous one, i.e. they are "MN-major" UMMA operands:
//   A = dY   (M = 64 output channels; conv2 has 32 -> planes 4..7 are zero), LBO = 128, SBO = plane
//   B = in   (N = 32 input channels), start shifted by the tap, LBO = 128, SBO = plane
// Pixels are enumerated over the padded grid (like the forward kernel); dY is zero at origins that
// fall into the padding, so those terms vanish.  One CTA = one node x one half of the 25 taps
// (13 x 32 fp32 TMEM columns); it walks over all samples of the node, double buffering the
// staged planes: warps 1-7 stage chunk c+1 while one thread of warp 0 issues the MMAs of chunk c.
// ==========================================================================================
// NIC = shifted copies of the INPUT planes (by 0, 2, 4 pixels) stacked along N: one MMA with N = 32 * NIC columns
// then covers NIC horizontal tap positions and the dY operand (4 KB) is read from shared memory once for all of
// them — 44 SM cycles for N = 32, 48 for N = 64, 56 for N = 96 (scripts/probe_mma_rate.py).
template <int COUT, int H, int SB, bool PAIR = false, bool QUAD = false, int NIC = 1>
struct WgradTcCfg {
    static constexpr int CIN = 32;
    using G = PlaneGeom<H>;
    static constexpr int PP = G::PP, NPS = G::NPS;
    // dY is zero at every origin past the last valid one of the chunk's last sample, so the reduction
    // stops there (H = 15, SB = 1: 253 origins -> 16 K steps)
    static constexpr int NVALID = (SB - 1) * NPS + G::NORIG;
    static constexpr int DMAX = QUAD ? PP + 1 : 1;                  // largest shift of a dY copy
    static constexpr int KPIX = ((NVALID + DMAX + 15) / 16) * 16;  // reduction length per chunk (+ the shifted copies)
    static constexpr int TAIL = ((G::TAIL + 7) / 8) * 8;
    static constexpr int FULL = (((SB - 1) * NPS + G::EXTENT + 7) / 8) * 8;
    static constexpr int NPD = KPIX, NPI = (KPIX + TAIL > FULL) ? KPIX + TAIL : FULL;
    static constexpr int PLANE_D = NPD * 16, PLANE_I = NPI * 16;
    static constexpr int MROWS = QUAD ? 4 * COUT : (PAIR ? 2 * COUT : 64);   // accumulator rows of one MMA
    static constexpr int DPLANES = MROWS / 8;                      // 8-channel dY planes (PAIR / QUAD: shifted copies)
    static constexpr int IPLANES = 4 * NIC;
    static constexpr int BUF = DPLANES * PLANE_D + IPLANES * PLANE_I;
    static constexpr int SMEM = 2 * BUF + 128;
    static_assert(SMEM <= 227 * 1024, "shared memory budget");
    static constexpr int TAPS0 = 13;                                // taps of CTA half 0; half 1 gets 12
};

constexpr int WG_THREADS = 512;
constexpr int WG_PRODUCERS = WG_THREADS - 32;

// PAIR (COUT == 32): instead of four zero planes, rows 32..63 of the M = 64 operand hold a second
// copy of dY shifted by one pixel.  sum_p dY[p-1] * in[p + s] = sum_q dY[q] * in[q + s + 1], so the
// upper half of the accumulator is the weight gradient of the NEXT tap in x: one MMA yields taps
// (ky,kx) and (ky,kx+1), 15 MMAs per K step cover all 25 taps (480 TMEM columns) and one CTA serves
// a whole node.
// QUAD (COUT == 32): FOUR copies of dY — shifted by 0, 1, PP and PP + 1 pixels — fill an M = 128 operand, so one
// MMA with the input shifted by (2a*PP + 2b) yields the 2x2 block of taps (2a + i, 2b + j): nine MMAs per K step
// cover a 6x6 tap grid (the 5x5 kernel plus eleven discarded entries) instead of fifteen M = 64 MMAs.
// PLANES: both operands already exist as bf16 channel-chunk planes in global memory (`in_pl`: the block input,
// `dy_pl`: the gradient w.r.t. the conv output, `zeros`: >= one plane run of zeros); one elected thread stages a
// chunk with bulk asynchronous copies — the shifted copies of dY are the SAME global plane copied to shifted
// destinations — and only the chunks that hold live samples are visited.
template <int COUT, int H, int SB, bool PAIR, bool QUAD, bool PLANES>
__global__ void __launch_bounds__(WG_THREADS, 1)
conv_wgrad_tc_kernel(const float *__restrict__ in, const float *__restrict__ dY, const uint16_t *__restrict__ in_pl,
                     const uint16_t *__restrict__ dy_pl, const uint16_t *__restrict__ zeros, Pass p, Workspace ws, int w_off)
{
    constexpr int NIC = PLANES ? (QUAD ? 2 : 3) : 1;         // input copies (bulk-copy staging only)
    using C = WgradTcCfg<COUT, H, SB, PAIR, QUAD, NIC>;
    static_assert(!QUAD || (PAIR && COUT == 32), "the quad variant is the four-copy form of the pair trick");
    static_assert(!PAIR || (C::NPD >= C::NVALID + C::DMAX), "shifted copies need spare plane entries");
    static_assert(C::NPI >= C::FULL && C::NPI >= C::KPIX + C::TAIL, "input planes must hold every padded pixel of the chunk");
    static_assert(PAIR || COUT == 64, "without the pair trick the operand must have 64 real rows");
    constexpr int NMMA = QUAD ? 9 : (PAIR ? 15 : C::TAPS0);   // accumulator tiles (32 columns each) per CTA
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + 2 * C::BUF);   // full[2], empty[2], done
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 5);
    const int t = threadIdx.x, warp = tc::uniform_warp_idx(), lane = t & 31;
    const int g = PAIR ? blockIdx.x : (blockIdx.x >> 1), half = PAIR ? 0 : (blockIdx.x & 1);
    if (p.gcount[g] <= 0) return;
    const int tap0 = half ? C::TAPS0 : 0, ntap = PAIR ? NMMA : (half ? 25 - C::TAPS0 : C::TAPS0);
    const int gs = p.gs;
    int n_live = gs;                                          // live samples are a prefix of the group
    if (PLANES) { while (n_live > 0 && p.img_idx[g * gs + n_live - 1] < 0) --n_live; }
    const int n_chunks = ((PLANES ? n_live : gs) + SB - 1) / SB;
    __shared__ uint8_t s_live[DLS_MAX_GROUP];                 // sample slot holds an image (no dependent global loads
    if (t < gs) s_live[t] = p.img_idx[g * gs + t] >= 0;       // in the staging loops)

    {   // zero both buffers once: padding borders, K tail and (conv2) the four unused dY planes stay zero
        uint4 *z = reinterpret_cast<uint4 *>(smem);
        const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
        for (int i = t; i < 2 * C::BUF / 16; i += WG_THREADS) z[i] = zero;
    }
    if (t == 0) {
        tc::mbar_init(&bars[0], PLANES ? 1 : WG_PRODUCERS); tc::mbar_init(&bars[1], PLANES ? 1 : WG_PRODUCERS);
        tc::mbar_init(&bars[2], 1); tc::mbar_init(&bars[3], 1); tc::mbar_init(&bars[4], 1);
        tc::mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (PLANES && warp >= 1) {
        // ------------------------------ producer: bulk copies ------------------------------
        if (warp == 1 && tc::elect_one()) {
            constexpr int RUN = C::G::NORIG, RUNP = plane_runp(H), NCOPY = QUAD ? 4 : (PAIR ? 2 : 1);
            constexpr uint32_t TX = (uint32_t)(SB * ((COUT / 8) * NCOPY + 4 * NIC) * RUN * 16);
            for (int c = 0; c < n_chunks; ++c) {
                const int b = c & 1;
                if (c >= 2) tc::mbar_wait(&bars[2 + b], (uint32_t)(((c >> 1) - 1) & 1));
                const uint32_t sD = tc::smem_u32(smem + b * C::BUF), sI = sD + C::DPLANES * C::PLANE_D;
                tc::mbar_expect_tx(&bars[b], TX);
#pragma unroll
                for (int s = 0; s < SB; ++s) {
                    const int sl = c * SB + s;
                    const bool live = sl < n_live;
                    const uint16_t *dsrc = live ? dy_pl + (int64_t)(g * gs + sl) * ((COUT / 8) * RUNP * 8) : zeros;
                    const uint16_t *isrc = live ? in_pl + (int64_t)(g * gs + sl) * (4 * RUNP * 8) : zeros;
#pragma unroll
                    for (int ch = 0; ch < COUT / 8; ++ch) {
                        const uint16_t *src = live ? dsrc + ch * RUNP * 8 : zeros;
#pragma unroll
                        for (int j = 0; j < NCOPY; ++j) {
                            const int shift = j == 0 ? 0 : (j == 1 ? 1 : (j == 2 ? C::PP : C::PP + 1));   // pixels
                            tc::bulk_g2s(sD + (uint32_t)((j * (COUT / 8) + ch) * C::PLANE_D + (s * C::NPS + shift) * 16), src,
                                         RUN * 16, &bars[b]);
                        }
                    }
#pragma unroll
                    for (int j = 0; j < NIC; ++j)            // copy j reads as the input shifted by 2j pixels
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch)
                            tc::bulk_g2s(sI + (uint32_t)((j * 4 + ch) * C::PLANE_I + (s * C::NPS + 2 * C::PP + 2 - 2 * j) * 16),
                                         live ? isrc + ch * RUNP * 8 : zeros, RUN * 16, &bars[b]);
                }
            }
        }
    } else if (warp >= 1) {
        // ------------------------------ producers ------------------------------
        const int pt = t - 32;
        for (int c = 0; c < n_chunks; ++c) {
            const int b = c & 1;
            if (c >= 2) tc::mbar_wait(&bars[2 + b], (uint32_t)(((c >> 1) - 1) & 1));   // MMAs of chunk c-2 done
            uint8_t *sD = smem + b * C::BUF;
            uint8_t *sI = sD + C::DPLANES * C::PLANE_D;
            stage_items<8>(pt, C::NVALID * (COUT / 8), WG_PRODUCERS, [&](int i, const float *&src, uint8_t *&dst) {
                const int pix = i / (COUT / 8), ch = i - pix * (COUT / 8);
                const int s = pix / C::NPS, loc = pix - s * C::NPS, oy = loc / C::PP, ox = loc - oy * C::PP;
                const int sl = c * SB + s;
                const bool ok = sl < gs && oy < H && ox < H && s_live[sl < gs ? sl : 0];
                src = ok ? dY + (int64_t)(g * gs + sl) * (H * H * COUT) + (oy * H + ox) * COUT + ch * 8 : nullptr;
                dst = sD + ch * C::PLANE_D + pix * 16;
            }, PAIR ? (COUT / 8) * C::PLANE_D + 16 : 0, QUAD ? 2 * (COUT / 8) * C::PLANE_D + C::PP * 16 : 0,
               QUAD ? 3 * (COUT / 8) * C::PLANE_D + (C::PP + 1) * 16 : 0);
            stage_items<4>(pt, SB * H * H * 4, WG_PRODUCERS, [&](int i, const float *&src, uint8_t *&dst) {
                const int ch = i & 3, q = i >> 2, s = q / (H * H), pix = q - s * (H * H);
                const int y = pix / H, x = pix - y * H;
                const int sl = c * SB + s;
                const bool ok = sl < gs && s_live[sl < gs ? sl : 0];
                src = ok ? in + (int64_t)(g * gs + sl) * (H * H * 32) + pix * 32 + ch * 8 : nullptr;
                dst = sI + ch * C::PLANE_I + (s * C::NPS + (y + 2) * C::PP + (x + 2)) * 16;
            });
            tc::fence_async_smem();
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(tc::smem_u32(&bars[b])) : "memory");
        }
    } else {
        // ------------------------------ MMA issuer ------------------------------
        // warp-uniform control flow, one elected lane issues (no ELECT/BRA.U.ANY waterfall per MMA)
        const bool leader = tc::elect_one();
        const uint32_t idesc = tc::instr_desc_bf16(C::MROWS, 32, 1, 1);
        const uint32_t idesc_n2 = tc::instr_desc_bf16(C::MROWS, 64, 1, 1), idesc_n3 = tc::instr_desc_bf16(C::MROWS, 96, 1, 1);
        (void)idesc_n2; (void)idesc_n3;
        uint32_t acc = 0;
        uint32_t tshift[NMMA];                       // tap shift in 16-byte units (registers: loops are unrolled)
#pragma unroll
        for (int ti = 0; ti < NMMA; ++ti) {
            if (QUAD) {
                tshift[ti] = (uint32_t)(2 * (ti / 3) * C::PP + 2 * (ti % 3));   // (ky, kx) = (0|2|4, 0|2|4)
            } else if (PAIR) {
                tshift[ti] = (uint32_t)((ti / 3) * C::PP + 2 * (ti % 3));       // (ky, kx = 0,2,4)
            } else {
                const int tap = tap0 + (ti < ntap ? ti : 0);
                tshift[ti] = (uint32_t)((tap / 5) * C::PP + (tap % 5));
            }
        }
        for (int c = 0; c < n_chunks; ++c) {
            const int b = c & 1;
            tc::mbar_wait(&bars[b], (uint32_t)((c >> 1) & 1));
            tc::fence_after_sync();
            const uint32_t d_base = tc::smem_u32(smem + b * C::BUF);
            const uint64_t ad0 = tc::smem_desc(d_base, 128, C::PLANE_D);
            const uint64_t bd0 = tc::smem_desc(d_base + C::DPLANES * C::PLANE_D, 128, C::PLANE_I);
            if (leader) {
#pragma unroll 1
                for (int j = 0; j < C::KPIX / 16; ++j) {
                    const uint64_t ad = ad0 + (uint64_t)(j * 16);
                    const uint64_t bdj = bd0 + (uint64_t)(j * 16);
                    if (NIC == 1) {
#pragma unroll
                        for (int ti = 0; ti < NMMA; ++ti)
                            if (ti < ntap) tc::mma_f16(tmem + (uint32_t)(ti * 32), ad, bdj + tshift[ti], idesc, acc);
                    } else if (QUAD) {
                        // accumulator tile ti = 3a + b <-> taps (2a + i, 2b + j'); b = 0, 1 in one N = 64 MMA, b = 2 alone
#pragma unroll
                        for (int a = 0; a < 3; ++a) {
                            tc::mma_f16(tmem + (uint32_t)(3 * a * 32), ad, bdj + (uint64_t)(2 * a * C::PP), idesc_n2, acc);
                            tc::mma_f16(tmem + (uint32_t)((3 * a + 2) * 32), ad, bdj + (uint64_t)(2 * a * C::PP + 4), idesc, acc);
                        }
                    } else {
                        // accumulator tile ti = 3 ky + kxp <-> taps (ky, 2 kxp + copy): the three kxp in one N = 96 MMA
#pragma unroll
                        for (int ky = 0; ky < 5; ++ky)
                            tc::mma_f16(tmem + (uint32_t)(3 * ky * 32), ad, bdj + (uint64_t)(ky * C::PP), idesc_n3, acc);
                    }
                    acc = 1;
                }
                tc::mma_commit(&bars[2 + b]);          // buffer b may be refilled once these MMAs are done
            }
            __syncwarp();
        }
        if (leader) tc::mma_commit(&bars[4]);
    }
    __syncwarp();
    // ------------------------------ epilogue: TMEM -> gradient row ------------------------------
    // The accumulators are transposed through shared memory (the staging buffers are dead once the last MMA
    // has completed) into the layout of the gradient row, [co][ci][tap] with a padded pitch of 801 floats per
    // output channel, 32 output channels at a time, and leave the SM as coalesced stores.
    static_assert(PAIR, "the transposed epilogue is written for the shifted-copy forms");
    float *G = ws.G + (int64_t)g * ws.ldg + w_off;
    float *sT = reinterpret_cast<float *>(smem);                   // [TR][801]
    constexpr int TR = (32 * 801 * 4 <= 2 * C::BUF) ? 32 : 16;     // output channels per pass
    static_assert(TR * 801 * 4 <= 2 * C::BUF, "transpose tile must fit in the staging buffers");
    if (warp < 4) {
        tc::mbar_wait(&bars[4], 0);
        tc::fence_after_sync();
    }
    const int mrow = warp * 32 + lane;                             // M = 128: accumulator row = TMEM lane (warps 0-3)
    for (int half = 0; half < COUT / TR; ++half) {
        if (warp < 4) {
            const int cp = mrow / COUT, co = mrow % COUT;          // shifted copy and output channel of this row
            // tcgen05.ld is warp-wide: the branch around it is warp-uniform (the warp's 32 channels overlap this
            // pass or not); only the shared-memory stores are per lane
            if (((warp * 32) % COUT) / 32 == (half * TR) / 32) {
                for (int ti = 0; ti < NMMA; ++ti) {
                    int ky, kx;
                    if (QUAD) { ky = 2 * (ti / 3) + (cp >> 1); kx = 2 * (ti % 3) + (cp & 1); }
                    else      { ky = ti / 3;                   kx = 2 * (ti % 3) + cp; }
                    if (ky < 5 && kx < 5) {                        // warp-uniform: a warp holds one copy
                        uint32_t r[32];
                        tc::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(ti * 32), r);
                        tc::tmem_ld_wait();
                        if (co / TR == half) {
                            float *dst = sT + (co % TR) * 801 + ky * 5 + kx;
#pragma unroll
                            for (int ci = 0; ci < 32; ++ci) dst[ci * 25] = __uint_as_float(r[ci]);
                        }
                    }
                }
            }
        }
        __syncthreads();
        if (ws.sgd.X) {
            // fused optimizer step on these TR x 800 weights (g += wd p; buf = g on the node's first step, else
            // buf = mu buf + g; p -= lr buf — the arithmetic of sgd_update_kernel)
            const SgdFuse &o = ws.sgd;
            float *xr = o.X + (int64_t)o.task_row[g] * o.ldx + w_off + half * TR * 800;
            float *mr = (o.M && o.mu != 0.f) ? o.M + (int64_t)o.task_peer[g] * o.ldm + w_off + half * TR * 800 : nullptr;
            const bool fst = o.first ? (o.first[g] != 0) : false;
            constexpr int U = 10;                                  // independent loads in flight per thread
            for (int i0 = t; i0 < TR * 800; i0 += U * WG_THREADS) {
                float pv[U], mv[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int i = i0 + u * WG_THREADS;
                    pv[u] = i < TR * 800 ? xr[i] : 0.f;
                    mv[u] = (mr && !fst && i < TR * 800) ? mr[i] : 0.f;
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int i = i0 + u * WG_THREADS;
                    if (i >= TR * 800) break;
                    const int co = i / 800, rem = i - co * 800;
                    float gv = sT[co * 801 + rem];
                    if (o.wd != 0.f) gv = fmaf(o.wd, pv[u], gv);
                    if (mr) {
                        const float bv = fst ? gv : __fadd_rn(__fmul_rn(mv[u], o.mu), gv);
                        mr[i] = bv;
                        gv = bv;
                    }
                    xr[i] = fmaf(-o.lr, gv, pv[u]);
                }
            }
        } else {
            float *Gh = G + half * TR * 800;
            for (int i = t; i < TR * 800; i += WG_THREADS) {
                const int co = i / 800, rem = i - co * 800;
                Gh[i] = sT[co * 801 + rem];
            }
        }
        __syncthreads();
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 512);
}


template <int COUT, int H, int SB, bool PAIR, bool QUAD = false>
static int launch_wgrad_tc(const float *in, const float *dY, const uint16_t *in_pl, const uint16_t *dy_pl, const Pass &p,
                           Workspace &ws, int w_off, cudaStream_t st)
{
    using C = WgradTcCfg<COUT, H, SB, PAIR, QUAD>;
    if (in_pl && dy_pl) {
        using CP = WgradTcCfg<COUT, H, SB, PAIR, QUAD, QUAD ? 2 : 3>;
        static PerDeviceOnce attrp;
        int rc = ensure_dyn_smem(conv_wgrad_tc_kernel<COUT, H, SB, PAIR, QUAD, true>, CP::SMEM, attrp);
        if (rc) return rc;
        conv_wgrad_tc_kernel<COUT, H, SB, PAIR, QUAD, true><<<p.n_groups * (PAIR ? 1 : 2), WG_THREADS, CP::SMEM, st>>>(
            in, dY, in_pl, dy_pl, ws.zplane, p, ws, w_off);
        DLS_LAUNCH_CHECK();
        return 0;
    }
    static PerDeviceOnce attr;
    int rc = ensure_dyn_smem(conv_wgrad_tc_kernel<COUT, H, SB, PAIR, QUAD, false>, C::SMEM, attr);
    if (rc) return rc;
    conv_wgrad_tc_kernel<COUT, H, SB, PAIR, QUAD, false><<<p.n_groups * (PAIR ? 1 : 2), WG_THREADS, C::SMEM, st>>>(
        in, dY, nullptr, nullptr, nullptr, p, ws, w_off);
    DLS_LAUNCH_CHECK();
    return 0;       // the bias gradient comes from the per-sample partials of gn_block_bwd (gn_param_reduce)
}

int conv2_wgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    const bool pl = !tune_value(TUNE_CONV_UNGROUPED);
    return launch_wgrad_tc<32, 15, 1, true, true>(ws.a1, ws.dc2, pl ? ws.a1p : nullptr, pl ? ws.dc2p : nullptr, p, ws, OFF_W2, st);
}
int conv3_wgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    const bool pl = !tune_value(TUNE_CONV_UNGROUPED);
    return launch_wgrad_tc<64, 7, 2, true>(ws.a2, ws.dc3, pl ? ws.a2p : nullptr, pl ? ws.dc3p : nullptr, p, ws, OFF_W3, st);
}

// ------------------------------------------------------------------------------------------
// FEMNIST conv2 (32 -> 64, 14x14): the weight tensor has the shape of GNLeNet's conv3, so the bf16 copies use
// the same slots of the prep record; two M = 128 tiles per sample (222 origins).
// ------------------------------------------------------------------------------------------
int fem_prep_conv2_tc(const Pass &p, void *prep_tc, int w_off, bool backward, cudaStream_t st)
{
    static PerDeviceOnce attr;
    int rc = ensure_dyn_smem(prep_weights_tc_kernel, PREP_SMEM, attr);
    if (rc) return rc;
    const int32_t *active = p.gmodel ? nullptr : p.gcount;
    prep_weights_tc_kernel<<<dim3(4, p.n_models), 256, PREP_SMEM, st>>>(p, reinterpret_cast<__nv_bfloat16 *>(prep_tc),
                                                                     backward ? 1 : 0, active, 1, 2, 0, w_off);
    DLS_LAUNCH_CHECK();
    return 0;
}
int fem_conv2_fwd_tc(const Pass &p, const uint16_t *a1p, float *c2, const void *prep_tc, cudaStream_t st)
{
    if (!a1p || tune_value(TUNE_CONV_UNGROUPED)) return DLS_E_ARG;
    return launch_conv_tc<32, 64, 14, 1, false, 3>(nullptr, a1p, c2, p, reinterpret_cast<const __nv_bfloat16 *>(prep_tc),
                                                   TCO_W3U, 0, st);
}
int fem_conv2_dgrad_tc(const Pass &p, const uint16_t *dc2p, float *da1, const void *prep_tc, cudaStream_t st)
{
    if (!dc2p || tune_value(TUNE_CONV_UNGROUPED)) return DLS_E_ARG;
    return launch_conv_tc<64, 32, 14, 1, false, 2>(nullptr, dc2p, da1, p, reinterpret_cast<const __nv_bfloat16 *>(prep_tc),
                                                   TCO_W3UD, 0, st);
}
int fem_conv2_wgrad_tc(const Pass &p, const uint16_t *a1p, const uint16_t *dc2p, const uint16_t *zeros, float *G,
                       int64_t ldg, int w_off, cudaStream_t st)
{
    if (!a1p || !dc2p || !zeros) return DLS_E_ARG;
    Workspace w{};
    w.G = G; w.ldg = ldg; w.zplane = const_cast<uint16_t *>(zeros);
    return launch_wgrad_tc<64, 14, 1, true>(nullptr, nullptr, a1p, dc2p, p, w, w_off, st);
}

}  // namespace gnl
}  // namespace dls
