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
__global__ void __launch_bounds__(256)
prep_weights_tc_kernel(Pass p, __nv_bfloat16 *prep, int backward, const int32_t *model_active)
{
    extern __shared__ __align__(16) float wsrc[];                 // [16][PREP_ROW]
    const int m = blockIdx.y, part = blockIdx.x, t = threadIdx.x;
    if (model_active && model_active[m] == 0) return;
    const bool l3 = part >= 2;
    const int co0 = (l3 ? part - 2 : part) * 16, nco = l3 ? 64 : 32;
    const float *row = p.X + (int64_t)p.mrow[m] * p.ldx + (l3 ? OFF_W3 : OFF_W2) + co0 * 800;
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
            entry = (int64_t)kc * nco + co0 + nl;
            dst = dst_u;
        } else {
            const int j = i - 1600, n = j & 31, tg = j >> 5, cg = tg & 1, tap = tg >> 1;
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = wsrc[(cg * 8 + e) * PREP_ROW + n * 25 + (24 - tap)];
            entry = (int64_t)(tap * (nco / 8) + co0 / 8 + cg) * 32 + n;
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
// host side
// ------------------------------------------------------------------------------------------
template <int CIN, int COUT, int H, int SB, bool BIAS>
static int launch_conv_tc(const float *in, float *out, const Pass &p, const __nv_bfloat16 *wprep, int wprep_off,
                          int bias_off, cudaStream_t st)
{
    using C = ConvTcCfg<CIN, COUT, H, SB>;
    static PerDeviceOnce attr;
    int rc = ensure_dyn_smem(conv_tc_kernel<CIN, COUT, H, SB, BIAS>, C::SMEM, attr);
    if (rc) return rc;
    // a CTA serves a whole node when there are enough nodes to fill the GPU, else half a node
    int spc = p.gs;
    if (p.n_groups < num_sms() && p.gs > SB) spc = ((p.gs / 2 + SB - 1) / SB) * SB;
    const int parts = (p.gs + spc - 1) / spc;
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
                                                 active);
    DLS_LAUNCH_CHECK();
    return prep_w1_tc(p, ws, st);
}

int conv2_fwd_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    return launch_conv_tc<32, 32, 15, 1, true>(ws.a1, ws.c2, p, reinterpret_cast<const __nv_bfloat16 *>(ws.prep_tc),
                                               TCO_W2U, OFF_B2, st);
}
int conv3_fwd_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    return launch_conv_tc<32, 64, 7, 3, true>(ws.a2, ws.c3, p, reinterpret_cast<const __nv_bfloat16 *>(ws.prep_tc),
                                              TCO_W3U, OFF_B3, st);
}
int conv3_dgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    return launch_conv_tc<64, 32, 7, 3, false>(ws.dc3, ws.da2, p, reinterpret_cast<const __nv_bfloat16 *>(ws.prep_tc),
                                               TCO_W3UD, 0, st);
}
int conv2_dgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    return launch_conv_tc<32, 32, 15, 1, false>(ws.dc2, ws.da1, p, reinterpret_cast<const __nv_bfloat16 *>(ws.prep_tc),
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
template <int COUT, int H, int SB, bool PAIR = false, bool QUAD = false>
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
    static constexpr int BUF = DPLANES * PLANE_D + 4 * PLANE_I;
    static constexpr int SMEM = 2 * BUF + 128;
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
template <int COUT, int H, int SB, bool PAIR, bool QUAD>
__global__ void __launch_bounds__(WG_THREADS, 1)
conv_wgrad_tc_kernel(const float *__restrict__ in, const float *__restrict__ dY, Pass p, Workspace ws, int w_off)
{
    using C = WgradTcCfg<COUT, H, SB, PAIR, QUAD>;
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
    const int gs = p.gs, n_chunks = (gs + SB - 1) / SB;
    __shared__ uint8_t s_live[DLS_MAX_GROUP];                 // sample slot holds an image (no dependent global loads
    if (t < gs) s_live[t] = p.img_idx[g * gs + t] >= 0;       // in the staging loops)

    {   // zero both buffers once: padding borders, K tail and (conv2) the four unused dY planes stay zero
        uint4 *z = reinterpret_cast<uint4 *>(smem);
        const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
        for (int i = t; i < 2 * C::BUF / 16; i += WG_THREADS) z[i] = zero;
    }
    if (t == 0) {
        tc::mbar_init(&bars[0], WG_PRODUCERS); tc::mbar_init(&bars[1], WG_PRODUCERS);
        tc::mbar_init(&bars[2], 1); tc::mbar_init(&bars[3], 1); tc::mbar_init(&bars[4], 1);
        tc::mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp >= 1) {
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
#pragma unroll
                    for (int ti = 0; ti < NMMA; ++ti)
                        if (ti < ntap) tc::mma_f16(tmem + (uint32_t)(ti * 32), ad, bdj + tshift[ti], idesc, acc);
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
        float *Gh = G + half * TR * 800;
        for (int i = t; i < TR * 800; i += WG_THREADS) {
            const int co = i / 800, rem = i - co * 800;
            Gh[i] = sT[co * 801 + rem];
        }
        __syncthreads();
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 512);
}


template <int COUT, int H, int SB, bool PAIR, bool QUAD = false>
static int launch_wgrad_tc(const float *in, const float *dY, const Pass &p, Workspace &ws, int w_off, int b_off,
                           cudaStream_t st)
{
    using C = WgradTcCfg<COUT, H, SB, PAIR, QUAD>;
    static PerDeviceOnce attr;
    int rc = ensure_dyn_smem(conv_wgrad_tc_kernel<COUT, H, SB, PAIR, QUAD>, C::SMEM, attr);
    if (rc) return rc;
    conv_wgrad_tc_kernel<COUT, H, SB, PAIR, QUAD><<<p.n_groups * (PAIR ? 1 : 2), WG_THREADS, C::SMEM, st>>>(in, dY, p,
                                                                                                            ws, w_off);
    DLS_LAUNCH_CHECK();
    (void)b_off;      // the bias gradient comes from the per-sample partials of gn_block_bwd (gn_param_reduce)
    return 0;
}

int conv2_wgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    return launch_wgrad_tc<32, 15, 1, true, true>(ws.a1, ws.dc2, p, ws, OFF_W2, OFF_B2, st);
}
int conv3_wgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    return launch_wgrad_tc<64, 7, 2, true>(ws.a2, ws.dc3, p, ws, OFF_W3, OFF_B3, st);
}

}  // namespace gnl
}  // namespace dls
