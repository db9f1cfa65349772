// conv1 (3 -> 32, 5x5, pad 2, 32x32 images) on tensor cores, fused with maxpool 3/2 + GroupNorm +
// ReLU (block 1 of GNLeNet, dasklearn/models/cifar10.py:120-123), and its weight gradient.
//
// K = 75 is too thin for a channel-plane layout (3 channels), so the image is "x-expanded" in
// shared memory: an entry (py, ox) = the 15 values (kx, c) of padded row py that output column ox needs
// (+1 zero), as two 8-element halves j.  For a fixed kernel row ky the A operand of a run of output
// pixels is then the expansion shifted by ky rows: one K = 16 MMA per kernel row.  The forward kernel
// stores the entries in four parity planes (stride-2 windows for the fused pooling), the weight
// gradient interleaves the halves (`E2[py][j][ox]`); both are described at their kernels below.
#include "gnlenet.cuh"
#include "tc_common.cuh"
#include <cuda_bf16.h>

namespace dls {
namespace gnl {

constexpr int E_NP = 36 * 32;                 // entries per plane (py, ox)
constexpr int E_PLANE = E_NP * 16;            // bytes
constexpr int C1_W_BYTES = 10 * 32 * 16;      // W1u: [ky*2 + j][co][8] bf16
constexpr int TCO_W1U = 153600;               // offset inside the per-model bf16 prep block (after conv2/3)

// W1u[kc = ky*2+j][n = co][e] = W1[co][c][ky][kx] with 8j+e = kx*3+c (< 15), else 0
__global__ void __launch_bounds__(256)
prep_w1_tc_kernel(Pass p, __nv_bfloat16 *prep, const int32_t *model_active)
{
    const int m = blockIdx.x;
    if (model_active && model_active[m] == 0) return;
    const float *row = p.X + (int64_t)p.mrow[m] * p.ldx;
    __nv_bfloat16 *dst = prep + (int64_t)m * TC_PREP_ELEMS + TCO_W1U;
    for (int i = threadIdx.x; i < 2560; i += 256) {
        const int e = i & 7, n = (i >> 3) & 31, kc = i >> 8, ky = kc >> 1, j = kc & 1;
        const int q = 8 * j + e;
        float v = 0.f;
        if (q < 15) { const int kx = q / 3, c = q - kx * 3; v = row[OFF_W1 + n * 75 + c * 25 + ky * 5 + kx]; }
        dst[i] = __float2bfloat16_rn(v);
    }
}

// The image is first converted once to bf16 rows in shared memory, padded like the convolution
// (2 zero pixels left, >= 2 right): row[6 + x*3 + c] = bf16(byte / 255).
constexpr int ROW16 = 116;                       // bf16 elements per padded row (58 32-bit words)
constexpr int S16_BYTES = 32 * ROW16 * 2;

__device__ __forceinline__ void convert_rows(const uint8_t *simg, const __nv_bfloat16 *lut, uint32_t *s16, int t,
                                             int nthreads)
{
    for (int i = t; i < 32 * (ROW16 / 2); i += nthreads) {
        const int row = i / (ROW16 / 2), w = i - row * (ROW16 / 2), e0 = 2 * w - 6;
        const uint8_t *rp = simg + row * 96;
        const uint32_t lo = (e0 >= 0 && e0 < 96) ? __bfloat16_as_ushort(lut[rp[e0]]) : 0u;
        const uint32_t hi = (e0 + 1 >= 0 && e0 + 1 < 96) ? __bfloat16_as_ushort(lut[rp[e0 + 1]]) : 0u;
        s16[i] = lo | (hi << 16);
    }
}

// ------------------------------------------------------------------------------------------
// conv1 + maxpool as NINE shifted convolutions (the forward kernel in use).
//
// A pooled pixel (py, px) is the maximum over conv pixels (2py + dy, 2px + dx), dy, dx in 0..2.  For a
// fixed (dy, dx) those conv pixels form a stride-2 sub-grid; with the expanded image split into four
// parity planes (row parity x column parity, 16 entries per half row) that sub-grid is again a plain
// shifted window, so tile T(dy, dx) — M = 128 rows = 8 pooled rows x 16 pooled columns — is five MMAs
// (one per ky).  TMEM lane m of every tile is the SAME pooled pixel: the 3x3 maximum and its arg-max
// are nine register compares per channel, in (dy, dx) order like the scan of the reference's
// max_pool2d, with no shared-memory ring, no cross-lane traffic and no barriers between tiles.
// The price is 90 MMAs per image instead of 40 (conv pixels shared by windows are recomputed); the
// tensor pipe has the room, the CUDA cores do not.
//   warps 0-3 / 4-7: pooled rows 0-7 / 8-14, one TMEM lane quadrant each: drain, pool, GroupNorm, store
//   warp 8         : MMA issuer; tiles flow through a ring of eight 32-column TMEM accumulators
// ------------------------------------------------------------------------------------------
constexpr int EP_PLANE = 20 * 16 * 16;        // parity plane: 18 half-rows of 16 entries (+2 rows junk lanes may read)
constexpr int C1P_POOL = 256, C1P_PRODUCERS = 96;
constexpr int C1P_THREADS = C1P_POOL + 32 + C1P_PRODUCERS;     // warps 0-7 pool, warp 8 issuer, warps 9-11 producers
constexpr int C1P_SLOTS = 8;                   // TMEM accumulator ring (32 columns each)
constexpr int C1P_STAGE = 8 * EP_PLANE + C1_W_BYTES;           // one staged image: parity planes + the node's W1
constexpr int C1P_CAP = 256;                  // sample slots per CTA (descriptor table in shared memory)
constexpr int C1P_SMEM = 2 * C1P_STAGE + S16_BYTES + 2 * 3072 + 512 + 4 * 96 * 4 + 4 * 16 * 4 +
                         (4 + 2 * C1P_SLOTS) * 8 + 64 + 4 * C1P_CAP * 4;

// E entries go into parity planes: plane ((pr&1)*2 + (ox&1))*2 + j, entry (pr>>1)*16 + (ox>>1); an entry holds
// the 8 consecutive bf16 of padded row pr-2 that start at element ox*3 + 8j (the 16th value of a pixel is a zero).
// Producer-side versions for exactly C1P_PRODUCERS = 96 threads with loop-invariant index math.
// The pad words of the bf16 rows are zeroed once per CTA (s16 is producer-private scratch).
__device__ __forceinline__ void c1p_zero_row_pads(uint32_t *s16, int pt)
{
    for (int i = pt; i < 32 * 10; i += C1P_PRODUCERS) {
        const int row = i / 10, k = i - row * 10;
        s16[row * (ROW16 / 2) + (k < 3 ? k : 48 + k)] = 0u;              // words 0-2 and 51-57
    }
}
// thread -> data word w (two bytes -> two bf16) of rows r0, r0+2, ...: sixteen independent chains
__device__ __forceinline__ void c1p_convert_rows(const uint8_t *simg, const __nv_bfloat16 *lut, uint32_t *s16, int pt)
{
    const int w = pt % 48, r0 = pt / 48;
    const unsigned short *lut16 = reinterpret_cast<const unsigned short *>(lut);
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int row = r0 + 2 * k;
        const uint32_t two = *reinterpret_cast<const unsigned short *>(simg + row * 96 + 2 * w);
        s16[row * (ROW16 / 2) + 3 + w] = (uint32_t)lut16[two & 0xFFu] | ((uint32_t)lut16[two >> 8] << 16);
    }
}
// thread -> (entry column ox, half j) and a third of the 36 padded rows; 2 passes x 12 rows
__device__ __forceinline__ void c1p_build_expanded(const uint32_t *s16, uint8_t *sE, int pt)
{
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
        const int sl = pt + C1P_PRODUCERS * pass;
        const int combo = sl & 63, third = sl >> 6;
        const int oxh = combo & 15, j = (combo >> 4) & 1, cpar = combo >> 5;
        const int off = (2 * oxh + cpar) * 3 + 8 * j;
        const uint32_t sh = (off & 1) ? 16u : 0u, mask_w = j ? 0x0000FFFFu : 0xFFFFFFFFu;
        const uint32_t *col = s16 + (off >> 1);
        uint8_t *dcol = sE + (cpar * 2 + j) * EP_PLANE + oxh * 16;
#pragma unroll
        for (int k = 0; k < 12; ++k) {
            const int pr = third * 12 + k, y = pr - 2;                    // 12 is even: pr & 1 == k & 1
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (y >= 0 && y < 32) {
                const uint32_t *rw = col + y * (ROW16 / 2);
                const uint32_t w0 = rw[0], w1 = rw[1], w2 = rw[2], w3 = rw[3], w4 = rw[4];
                v.x = __funnelshift_r(w0, w1, sh); v.y = __funnelshift_r(w1, w2, sh);
                v.z = __funnelshift_r(w2, w3, sh); v.w = __funnelshift_r(w3, w4, sh) & mask_w;
            }
            *reinterpret_cast<uint4 *>(dcol + (k & 1) * 4 * EP_PLANE + (pr >> 1) * 256) = v;
        }
    }
}

__device__ __forceinline__ void pool_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ void c1p_producer_bar() { asm volatile("bar.sync 2, 96;" ::: "memory"); }

// Persistent: a CTA owns a contiguous range of sample slots (two CTAs per SM, ranges differ by at most
// one sample, so the SMs are balanced at sample granularity) and pipelines them: the producers expand
// image i+1 into the other stage while the issuer and the pool warps work on image i.
__global__ void __launch_bounds__(C1P_THREADS, 2)
conv1_pool_tc_kernel(Pass p, Workspace ws, int save_bwd)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *stage0 = smem;                                                // [2][planes | W1]
    uint32_t *s16 = reinterpret_cast<uint32_t *>(smem + 2 * C1P_STAGE);    // producers' scratch
    uint8_t *simg = reinterpret_cast<uint8_t *>(s16) + S16_BYTES;
    __nv_bfloat16 *lut = reinterpret_cast<__nv_bfloat16 *>(simg + 2 * 3072);          // simg: two raw images
    float *params = reinterpret_cast<float *>(simg + 2 * 3072 + 512);      // [4][bias 32 | gamma 32 | beta 32]
    float *red = params + 4 * 96;                                          // [2 parities][2 passes][16]
    uint64_t *bars = reinterpret_cast<uint64_t *>(red + 4 * 16);           // full[2] empty[2] tfull[S] tempty[S]
    uint64_t *tfull = bars + 4, *tempty = tfull + C1P_SLOTS;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + C1P_SLOTS);
    int32_t *tab_slot = reinterpret_cast<int32_t *>(tmem_slot + 16);       // active samples of this CTA, compacted
    int32_t *tab_id = tab_slot + C1P_CAP, *tab_m = tab_id + C1P_CAP, *tab_row = tab_m + C1P_CAP;
    __shared__ int n_active_s;

    const int t = threadIdx.x, warp = tc::uniform_warp_idx(), lane = t & 31;
    const int n_slots = p.n_groups * p.gs;
    const int s_begin = (int)(((int64_t)blockIdx.x * n_slots) / gridDim.x);
    const int s_end = (int)(((int64_t)(blockIdx.x + 1) * n_slots) / gridDim.x);
    // descriptor table: no dependent global loads inside the pipelined loops
    if (warp == 0) {
        int n = 0;
        for (int base = s_begin; base < s_end; base += 32) {
            const int sl = base + lane;
            const int id = sl < s_end ? p.img_idx[sl] : -1;
            const unsigned mask = __ballot_sync(0xffffffffu, id >= 0);
            if (id >= 0) {
                const int pos = n + __popc(mask & ((1u << lane) - 1u));
                const int g = sl / p.gs;
                const int mm = p.gmodel ? p.gmodel[g] : g;
                tab_slot[pos] = sl; tab_id[pos] = id; tab_m[pos] = mm; tab_row[pos] = p.mrow[mm];
            }
            n += __popc(mask);
        }
        if (lane == 0) n_active_s = n;
    }

    for (int i = t; i < 256; i += C1P_THREADS) lut[i] = __float2bfloat16_rn(__fdiv_rn((float)i, 255.f));
    if (t == 0) {
        tc::mbar_init(&bars[0], C1P_PRODUCERS); tc::mbar_init(&bars[1], C1P_PRODUCERS);
        tc::mbar_init(&bars[2], 1); tc::mbar_init(&bars[3], 1);
        for (int i = 0; i < C1P_SLOTS; ++i) { tc::mbar_init(&tfull[i], 1); tc::mbar_init(&tempty[i], 128); }
        tc::mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, 32 * C1P_SLOTS);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    const int n_active = n_active_s;

    if (warp >= 9) {
        // ------------------------------ producers ------------------------------
        // All global traffic is asynchronous (cp.async): the raw bytes of image it+1 are in flight while
        // image it is expanded, W1 / bias / gamma / beta of image it land during its expansion.
        const int pt = t - (C1P_POOL + 32);
        auto fetch_image = [&](int k, int buf) {
            if (k < n_active) {
                const uint8_t *src = p.images + (int64_t)tab_id[k] * IMG_BYTES;
                uint8_t *dst = simg + buf * 3072;
                tc::cp_async16(dst + pt * 16, src + pt * 16);
                tc::cp_async16(dst + (pt + C1P_PRODUCERS) * 16, src + (pt + C1P_PRODUCERS) * 16);
            }
            tc::cp_async_commit();
        };
        fetch_image(0, 0);
        c1p_zero_row_pads(s16, pt);
        for (int it = 0; it < n_active; ++it) {
            const int b = it & 1;
            const int m = tab_m[it];
            const float *row = p.X + (int64_t)tab_row[it] * p.ldx;
            uint8_t *sE = stage0 + b * C1P_STAGE;
            if (it >= 2) tc::mbar_wait(&bars[2 + b], (uint32_t)(((it >> 1) - 1) & 1));      // MMAs of image it-2 done
            {
                const uint8_t *wsrc = reinterpret_cast<const uint8_t *>(reinterpret_cast<const __nv_bfloat16 *>(ws.prep_tc) +
                                                                         (int64_t)m * TC_PREP_ELEMS + TCO_W1U);
                uint8_t *wdst = sE + 8 * EP_PLANE;
                for (int i = pt; i < C1_W_BYTES / 16; i += C1P_PRODUCERS) tc::cp_async16(wdst + i * 16, wsrc + i * 16);
                if (pt < 24) {                                           // bias | gamma | beta: 3 x 128 bytes
                    const int off = pt < 8 ? OFF_B1 : (pt < 16 ? OFF_G1W : OFF_G1B);
                    tc::cp_async16(params + (it & 3) * 96 + pt * 4, row + off + (pt & 7) * 4);
                }
                tc::cp_async_commit();
            }
            fetch_image(it + 1, (it + 1) & 1);
            tc::cp_async_wait<2>();                                      // image it has landed
            c1p_producer_bar();
            c1p_convert_rows(simg + b * 3072, lut, s16, pt);
            c1p_producer_bar();
            c1p_build_expanded(s16, sE, pt);
            tc::cp_async_wait<1>();                                      // W1 and the parameters of image it
            tc::fence_async_smem();
            tc::mbar_arrive(&bars[b]);
        }
        tc::cp_async_wait<0>();
    } else if (warp == 8) {
        // ------------------------------ MMA issuer ------------------------------
        const bool leader = tc::elect_one();
        const uint32_t idesc = tc::instr_desc_bf16(128, 32, 0, 0);
        for (int it = 0; it < n_active; ++it) {
            const int b = it & 1;
            tc::mbar_wait(&bars[b], (uint32_t)((it >> 1) & 1));
            tc::fence_after_sync();
            const uint32_t e_base = tc::smem_u32(stage0 + b * C1P_STAGE);
            const uint64_t ad0 = tc::smem_desc(e_base, EP_PLANE, 128);
            const uint64_t bd0 = tc::smem_desc(e_base + 8 * EP_PLANE, 32 * 16, 128);
#pragma unroll
            for (int seq = 0; seq < 18; ++seq) {
                const int mt = seq & 1, tt = seq >> 1, dy = tt / 3, dx = tt % 3;
                const int gseq = it * 18 + seq, sl = gseq & (C1P_SLOTS - 1);
                if (gseq >= C1P_SLOTS) {
                    tc::mbar_wait(&tempty[sl], (uint32_t)(((gseq / C1P_SLOTS) - 1) & 1));
                    tc::fence_after_sync();
                }
                if (leader) {
#pragma unroll
                    for (int ky = 0; ky < 5; ++ky) {
                        const int sy = dy + ky;
                        const int plane = ((sy & 1) * 2 + (dx & 1)) * 2;
                        const int entry = (8 * mt + (sy >> 1)) * 16 + (dx >> 1);
                        tc::mma_f16(tmem + (uint32_t)(sl * 32), ad0 + (uint64_t)((plane * EP_PLANE + entry * 16) >> 4),
                                    bd0 + (uint64_t)(ky * 2 * 32), idesc, ky > 0 ? 1u : 0u);
                    }
                    tc::mma_commit(&tfull[sl]);
                }
                __syncwarp();
            }
            if (leader) tc::mma_commit(&bars[2 + b]);          // the stage is free once these MMAs are done
            __syncwarp();
        }
    } else {
        // ------------------------------ pool / normalise (warps 0-7) ------------------------------
        const int mt = warp >> 2, q = warp & 3;
        const int mrow = q * 32 + lane, py = 8 * mt + (mrow >> 4), px = mrow & 15;
        const bool valid = py < 15 && px < 15;
        const int pp = py * 15 + px;
        for (int it = 0; it < n_active; ++it) {
            const int slot = tab_slot[it];
            const float *prm = params + (it & 3) * 96;
            float *red1 = red + (it & 1) * 32, *red2 = red1 + 16;
            tc::mbar_wait(&bars[it & 1], (uint32_t)((it >> 1) & 1));      // acquire the producers' parameter writes
            // The bias is added after the maximum: rounding is monotone, so the maximum VALUE is the one of
            // max(x + b); only an exact tie created by that rounding could pick a later window position.
            float best[32];
            uint32_t bi[8];              // arg-max position (0..8) per channel, one byte each (idx1 layout)
#pragma unroll
            for (int w = 0; w < 8; ++w) bi[w] = 0u;
#pragma unroll
            for (int tt = 0; tt < 9; ++tt) {
                const int gseq = it * 18 + 2 * tt + mt, sl = gseq & (C1P_SLOTS - 1);
                tc::mbar_wait(&tfull[sl], (uint32_t)((gseq / C1P_SLOTS) & 1));
                tc::fence_after_sync();
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    uint32_t r[16];
                    tc::tmem_ld16(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(sl * 32 + hf * 16), r);
                    tc::tmem_ld_wait();
                    if (hf == 1) { tc::fence_before_sync(); tc::mbar_arrive(&tempty[sl]); }
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        const int co = hf * 16 + c;
                        const float x = __uint_as_float(r[c]);
                        if (tt == 0) {
                            best[co] = x;
                        } else if (x > best[co]) {                    // strict >: the first maximum wins
                            best[co] = x;
                            bi[co >> 2] = __byte_perm(bi[co >> 2], (uint32_t)tt,
                                                      (0x3210u & ~(0xFu << (4 * (co & 3)))) | (0x4u << (4 * (co & 3))));
                        }
                    }
                }
            }
#pragma unroll
            for (int co = 0; co < 32; ++co) best[co] += prm[co];
            // pooled values and arg-max bytes go out first (256-bit stores: one full sector per lane)
            if (valid && save_bwd) {
                float *p1 = ws.p1 + (int64_t)slot * A1_SZ + pp * 32;
#pragma unroll
                for (int c8 = 0; c8 < 4; ++c8)
                    tc::st_global_v8(p1 + c8 * 8, best[c8 * 8], best[c8 * 8 + 1], best[c8 * 8 + 2], best[c8 * 8 + 3],
                                     best[c8 * 8 + 4], best[c8 * 8 + 5], best[c8 * 8 + 6], best[c8 * 8 + 7]);
                tc::st_global_v8u(ws.idx1 + (int64_t)slot * A1_SZ + pp * 32, bi);
            }

            // GroupNorm(2, 32): two passes over the registers (mean, then centred sum of squares)
            float s0 = 0.f, s1 = 0.f;
            if (valid) {
#pragma unroll
                for (int co = 0; co < 16; ++co) { s0 += best[co]; s1 += best[16 + co]; }
            }
            s0 = warp_sum(s0); s1 = warp_sum(s1);
            if (lane == 0) { red1[warp * 2] = s0; red1[warp * 2 + 1] = s1; }
            pool_bar();
            float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) { sum0 += red1[w * 2]; sum1 += red1[w * 2 + 1]; }
            const float inv_n = 1.f / 3600.f;
            const float mean0 = sum0 * inv_n, mean1 = sum1 * inv_n;
            float q0 = 0.f, q1 = 0.f;
#pragma unroll
            for (int co = 0; co < 16; ++co) {                           // centre in place
                best[co] -= mean0; best[16 + co] -= mean1;
                q0 = fmaf(best[co], best[co], q0); q1 = fmaf(best[16 + co], best[16 + co], q1);
            }
            if (!valid) { q0 = 0.f; q1 = 0.f; }
            q0 = warp_sum(q0); q1 = warp_sum(q1);
            if (lane == 0) { red2[warp * 2] = q0; red2[warp * 2 + 1] = q1; }
            pool_bar();
            float v0 = 0.f, v1 = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) { v0 += red2[w * 2]; v1 += red2[w * 2 + 1]; }
            const float rstd0 = 1.f / sqrtf(v0 * inv_n + GN_EPS), rstd1 = 1.f / sqrtf(v1 * inv_n + GN_EPS);

            if (valid) {
                // the block output leaves as bf16 channel-chunk planes (what conv2 forward and the conv2 weight
                // gradient stage with bulk copies): entry r = py * 17 + px of plane c8, 16 bytes; the thread of a
                // row's last pixel also writes the two zero gap pixels that follow it
                uint4 *a1p = reinterpret_cast<uint4 *>(ws.a1p + (int64_t)slot * A1P_SZ) + (py * 17 + px);
#pragma unroll
                for (int c8 = 0; c8 < 4; ++c8) {
                    float o[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const int co = c8 * 8 + e;
                        const float v = best[co] * (co < 16 ? rstd0 : rstd1) * prm[32 + co] + prm[64 + co];
                        o[e] = v > 0.f ? v : 0.f;
                    }
                    a1p[c8 * RUNP15] = tc::pack8_bf16(o);
                    if (ws.a1) tc::st_global_v8(ws.a1 + (int64_t)slot * A1_SZ + pp * 32 + c8 * 8, o[0], o[1], o[2], o[3], o[4], o[5], o[6], o[7]);
                    if (px == 14 && py < 14) {
                        a1p[c8 * RUNP15 + 1] = make_uint4(0u, 0u, 0u, 0u);
                        a1p[c8 * RUNP15 + 2] = make_uint4(0u, 0u, 0u, 0u);
                    }
                }
            }
            if (save_bwd && t == 0) {
                float *st = ws.st1 + (int64_t)slot * 4;
                st[0] = mean0; st[1] = rstd0; st[2] = mean1; st[3] = rstd1;
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 32 * C1P_SLOTS);
}

// ------------------------------------------------------------------------------------------
// conv1 weight gradient on tensor cores.
//   dW1[co][c][ky][kx] = sum over samples and conv-output positions o of dc1[o][co] * P[o + (ky,kx)][c]
// where dc1 is the gradient w.r.t. the conv1 output: the pooled gradient dp1 routed to the arg-max
// position of every 3x3/2 window (overlapping windows add up).  Reduction (K) = the 1024 positions
// of an image; A = the x-expanded image E2[py][j][ox] (MN-major: the 16 chunks (ky, j) of a pixel are
// 512 bytes apart, ky >= 5 are don't-care rows), B = dc1 as dense bf16 channel planes.
// One CTA per node; warps 1-15 stage sample s+1 (expand image, scatter dp1) while warp 0 issues the
// 64 MMAs (M=128, N=32, K=16) of sample s.
// ------------------------------------------------------------------------------------------
constexpr int DC_PLANE = 1024 * 16;
constexpr int C1W_BUF = 2 * E_PLANE + 4 * DC_PLANE;
constexpr int C1W_NB = 2;                     // operand buffers per CTA.  Tried: 1 buffer and two CTAs per SM (two nodes
                                              // interleaving on one SM): 3 % slower end to end (64-register cap)
constexpr int C1W_SMEM = C1W_NB * C1W_BUF + 2 * 3072 + 512 + 128 + S16_BYTES + DLS_MAX_GROUP * 4;
constexpr int C1W_THREADS = 512, C1W_PRODUCERS = 480;

// E for the weight gradient: the two 8-element halves of a row are INTERLEAVED, E2[py][j][ox], so that the
// (ky, j) operand chunks of one pixel are a uniform 512 bytes apart and a single M = 128 MMA covers all of them.
__device__ __forceinline__ void build_expanded_il(const uint32_t *s16, uint8_t *sE, int t, int nthreads)
{
    for (int i = t; i < 36 * 32 * 2; i += nthreads) {
        const int ox = i & 31, j = (i >> 5) & 1, py = i >> 6;
        const int y = py - 2;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (y >= 0 && y < 32) {
            const int off = ox * 3 + 8 * j;
            const uint32_t *rw = s16 + y * (ROW16 / 2) + (off >> 1);
            const uint32_t w0 = rw[0], w1 = rw[1], w2 = rw[2], w3 = rw[3], w4 = rw[4];
            const uint32_t sh = (off & 1) ? 16u : 0u;
            v.x = __funnelshift_r(w0, w1, sh); v.y = __funnelshift_r(w1, w2, sh);
            v.z = __funnelshift_r(w2, w3, sh); v.w = __funnelshift_r(w3, w4, sh);
            if (j) v.w &= 0x0000FFFFu;           // element 15 of the (kx, c) vector is the zero pad
        }
        *reinterpret_cast<uint4 *>(sE + i * 16) = v;          // ((py*2 + j)*32 + ox) == i
    }
}

__device__ __forceinline__ void producer_sync() { asm volatile("bar.sync 1, 480;" ::: "memory"); }

__global__ void __launch_bounds__(C1W_THREADS, C1W_NB == 1 ? 2 : 1)
conv1_wgrad_tc_kernel(Pass p, Workspace ws)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *simg = smem + C1W_NB * C1W_BUF;
    __nv_bfloat16 *lut = reinterpret_cast<__nv_bfloat16 *>(simg + 2 * 3072);               // simg: two raw images
    uint64_t *bars = reinterpret_cast<uint64_t *>(simg + 2 * 3072 + 512);   // full[2], empty[2], done
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 5);
    uint32_t *s16 = reinterpret_cast<uint32_t *>(simg + 2 * 3072 + 512 + 128);
    int32_t *s_ids = reinterpret_cast<int32_t *>(reinterpret_cast<uint8_t *>(s16) + S16_BYTES);   // image id per sample
    const int t = threadIdx.x, warp = tc::uniform_warp_idx(), lane = t & 31;
    const int g = blockIdx.x, gs = p.gs;
    if (p.gcount[g] <= 0) return;
    if (t < gs) s_ids[t] = p.img_idx[g * gs + t];          // no dependent global loads inside the sample loop
    {
        uint4 *z = reinterpret_cast<uint4 *>(smem);
        const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
        for (int i = t; i < C1W_NB * C1W_BUF / 16; i += C1W_THREADS) z[i] = zero;
        for (int i = t; i < 256; i += C1W_THREADS) lut[i] = __float2bfloat16_rn(__fdiv_rn((float)i, 255.f));
    }
    if (t == 0) {
        tc::mbar_init(&bars[0], C1W_PRODUCERS); tc::mbar_init(&bars[1], C1W_PRODUCERS);
        tc::mbar_init(&bars[2], 1); tc::mbar_init(&bars[3], 1); tc::mbar_init(&bars[4], 1);
        tc::mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, 64);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp >= 1) {
        const int pt = t - 32;
        // Software pipeline over the samples of the node: the raw image bytes of sample s+1 (cp.async into a
        // double buffer) and its pooled-gradient values / arg-max bytes (registers) are requested while
        // sample s is expanded and scattered, so no global-memory latency sits between the producer barriers.
        constexpr int MAXIT = 5;                            // ceil(8*8*32 / 480) windows of one colour per thread
        float dv[4][MAXIT];
        int ii[4][MAXIT];                                   // arg-max position 0..8, -1 = no window
        auto fetch_image = [&](int s) {
            if (s < gs) {
                const int id = s_ids[s];
                if (id >= 0 && pt < IMG_BYTES / 16)
                    tc::cp_async16(simg + (s & 1) * 3072 + pt * 16, p.images + (int64_t)id * IMG_BYTES + pt * 16);
            }
            tc::cp_async_commit();
        };
        auto fetch_grad = [&](int s) {
            const bool live = s < gs && s_ids[s < gs ? s : 0] >= 0;
            const float *dp = ws.dp1 + (int64_t)(g * gs + (live ? s : 0)) * A1_SZ;
            const uint8_t *ix = ws.idx1 + (int64_t)(g * gs + (live ? s : 0)) * A1_SZ;
#pragma unroll
            for (int colour = 0; colour < 4; ++colour) {
                const int cy = colour >> 1, cx = colour & 1;
                const int ny = (15 - cy + 1) / 2, nx = (15 - cx + 1) / 2;
#pragma unroll
                for (int it = 0; it < MAXIT; ++it) {
                    const int i = pt + it * C1W_PRODUCERS;
                    dv[colour][it] = 0.f;
                    ii[colour][it] = -1;
                    if (live && i < ny * nx * 32) {
                        const int co = i & 31, w = i >> 5, py = 2 * (w / nx) + cy, px = 2 * (w % nx) + cx;
                        const int pp = py * 15 + px;
                        dv[colour][it] = __ldg(dp + pp * 32 + co);
                        ii[colour][it] = __ldg(ix + pp * 32 + co);
                    }
                }
            }
        };
        fetch_image(0);
        fetch_grad(0);
        for (int s = 0; s < gs; ++s) {
            const int b = s % C1W_NB;
            const int id = s_ids[s];
            if (s >= C1W_NB) tc::mbar_wait(&bars[2 + b], (uint32_t)(((s / C1W_NB) - 1) & 1));
            uint8_t *sE = smem + b * C1W_BUF;
            uint8_t *sD = sE + 2 * E_PLANE;
            {
                uint4 *z = reinterpret_cast<uint4 *>(sD);
                const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
                for (int i = pt; i < 4 * DC_PLANE / 16; i += C1W_PRODUCERS) z[i] = zero;
            }
            tc::cp_async_wait<0>();                         // this sample's image bytes have landed
            producer_sync();
            if (id >= 0) {
                convert_rows(simg + (s & 1) * 3072, lut, s16, pt, C1W_PRODUCERS);
                producer_sync();
                fetch_image(s + 1);
                build_expanded_il(s16, sE, pt, C1W_PRODUCERS);
                // windows of one colour (py%2, px%2) never touch the same conv position, so the
                // read-modify-write below is race free within a pass and the sum order is fixed
#pragma unroll
                for (int colour = 0; colour < 4; ++colour) {
                    const int cy = colour >> 1, cx = colour & 1;
                    const int nx = (15 - cx + 1) / 2;
#pragma unroll
                    for (int it = 0; it < MAXIT; ++it) {
                        if (ii[colour][it] >= 0) {
                            const int i = pt + it * C1W_PRODUCERS;
                            const int co = i & 31, w = i >> 5, py = 2 * (w / nx) + cy, px = 2 * (w % nx) + cx;
                            const int pos = ii[colour][it];
                            const int o = (2 * py + pos / 3) * 32 + 2 * px + pos % 3;
                            __nv_bfloat16 *c = reinterpret_cast<__nv_bfloat16 *>(sD + (co >> 3) * DC_PLANE + o * 16 + (co & 7) * 2);
                            *c = __float2bfloat16_rn(__bfloat162float(*c) + dv[colour][it]);
                        }
                    }
                    producer_sync();
                }
            } else {
                fetch_image(s + 1);
            }
            fetch_grad(s + 1);                              // in flight across the next sample's zero fill / expansion
            tc::fence_async_smem();
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(tc::smem_u32(&bars[b])) : "memory");
        }
        tc::cp_async_wait<0>();
    } else {
        const bool leader = tc::elect_one();          // warp-uniform loop, one elected lane issues
        const uint32_t idesc = tc::instr_desc_bf16(128, 32, 1, 1);
        uint32_t acc = 0;
        for (int s = 0; s < gs; ++s) {
            const int b = s % C1W_NB;
            tc::mbar_wait(&bars[b], (uint32_t)((s / C1W_NB) & 1));
            tc::fence_after_sync();
            const uint32_t e_base = tc::smem_u32(smem + b * C1W_BUF);
            const uint64_t ad0 = tc::smem_desc(e_base, 128, 512);          // 16 M chunks (ky, j), 512 bytes apart
            const uint64_t bd0 = tc::smem_desc(e_base + 2 * E_PLANE, 128, DC_PLANE);
            if (leader) {
#pragma unroll 8
                for (int k = 0; k < 64; ++k) {                              // 16 pixels: row k/2, half k%2
                    tc::mma_f16(tmem, ad0 + (uint64_t)((k >> 1) * 64 + (k & 1) * 16), bd0 + (uint64_t)(k * 16), idesc, acc);
                    acc = 1;
                }
                tc::mma_commit(&bars[2 + b]);
            }
            acc = 1;
            __syncwarp();
        }
        if (leader) tc::mma_commit(&bars[4]);
    }
    __syncwarp();
    if (warp < 4) {
        tc::mbar_wait(&bars[4], 0);
        tc::fence_after_sync();
        float *G = ws.G + (int64_t)g * ws.ldg + OFF_W1;
        const int mrow = warp * 32 + lane;                 // accumulator row = TMEM lane: chunk (ky, j) x 8 elements
        const int ch = mrow >> 3, e = mrow & 7, ky = ch >> 1, q = 8 * (ch & 1) + e;
        uint32_t r[32];
        tc::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16), r);
        tc::tmem_ld_wait();
        if (ky < 5 && q < 15) {
            const int kx = q / 3, c = q - kx * 3;
#pragma unroll
            for (int co = 0; co < 32; ++co) G[co * 75 + c * 25 + ky * 5 + kx] = __uint_as_float(r[co]);
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 64);
}


int conv1_wgrad_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    static PerDeviceOnce attr;
    int rc = ensure_dyn_smem(conv1_wgrad_tc_kernel, C1W_SMEM, attr);
    if (rc) return rc;
    conv1_wgrad_tc_kernel<<<p.n_groups, C1W_THREADS, C1W_SMEM, st>>>(p, ws);
    DLS_LAUNCH_CHECK();
    return 0;         // conv1's bias gradient: per-sample partials of gn_block_bwd, summed by gn_param_reduce
}

int prep_w1_tc(const Pass &p, Workspace &ws, cudaStream_t st)
{
    const int32_t *active = p.gmodel ? nullptr : p.gcount;
    prep_w1_tc_kernel<<<p.n_models, 256, 0, st>>>(p, reinterpret_cast<__nv_bfloat16 *>(ws.prep_tc), active);
    DLS_LAUNCH_CHECK();
    return 0;
}

int conv1_fwd_tc(const Pass &p, Workspace &ws, bool save, cudaStream_t st)
{
    static PerDeviceOnce attr;
    int rc = ensure_dyn_smem(conv1_pool_tc_kernel, C1P_SMEM, attr);
    if (rc) return rc;
    const int n_slots = p.n_groups * p.gs;
    int sms = num_sms();
    if (sms <= 0) sms = 148;
    int grid = n_slots < 2 * sms ? n_slots : 2 * sms;              // two persistent CTAs per SM
    if ((n_slots + grid - 1) / grid > C1P_CAP) grid = (n_slots + C1P_CAP - 1) / C1P_CAP;
    Workspace w2 = ws;
    if (!tune_value(TUNE_CONV_UNGROUPED)) w2.a1 = nullptr;          // only the bf16 planes a1p are consumed downstream
    conv1_pool_tc_kernel<<<grid, C1P_THREADS, C1P_SMEM, st>>>(p, w2, save ? 1 : 0);
    DLS_LAUNCH_CHECK();
    return 0;
}

}  // namespace gnl
}  // namespace dls
