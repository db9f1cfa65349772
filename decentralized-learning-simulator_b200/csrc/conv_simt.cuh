// CUDA-core (fp32) 5x5 convolution building blocks shared by the GNLeNet and FEMNIST kernels.
#pragma once
#include "gnlenet.cuh"

namespace dls {
namespace gnl {

// ------------------------------------------------------------------------------------------
// generic 5x5 / pad 2 convolution, NHWC, weights [tap][CIN][COUT]; one CTA per sample slot.
// lane = output channel (COUT/32 per lane), warp = output rows {w, w+8}.  Used for conv2/conv3
// forward and (with the transposed + flipped weight copies) for their data gradients.
// ------------------------------------------------------------------------------------------
template <int CIN, int COUT, int H, bool BIAS>
__global__ void __launch_bounds__(256)
conv5_kernel(const float *__restrict__ in, float *__restrict__ out, Pass p, const float *prep,
             int prep_stride, int prep_off, int bias_off)
{
    constexpr int HP = H + 4;
    constexpr int RPW = (H + 7) / 8;
    constexpr int NCO = COUT / 32;
    extern __shared__ __align__(16) float ins[];   // [HP][HP][CIN]
    const int slot = blockIdx.x, t = threadIdx.x;
    if (p.img_idx[slot] < 0) return;
    const int g = slot / p.gs;
    int m;
    const float *row = model_row(p, g, m);
    const float *wt = prep + (int64_t)m * prep_stride + prep_off;
    const float *src = in + (int64_t)slot * (H * H * CIN);
    for (int i = t; i < HP * HP * CIN; i += 256) {
        const int c = i % CIN, pp = i / CIN, yy = pp / HP, xx = pp - yy * HP;
        const bool inside = (yy >= 2) && (yy < H + 2) && (xx >= 2) && (xx < H + 2);
        ins[i] = inside ? src[((yy - 2) * H + (xx - 2)) * CIN + c] : 0.f;
    }
    __syncthreads();

    const int lane = t & 31, w = t >> 5;
    float acc[RPW][H][NCO];
#pragma unroll
    for (int n = 0; n < NCO; ++n) {
        const float b = BIAS ? row[bias_off + lane + 32 * n] : 0.f;
#pragma unroll
        for (int r = 0; r < RPW; ++r)
#pragma unroll
            for (int x = 0; x < H; ++x) acc[r][x][n] = b;
    }
    int yr[RPW];
#pragma unroll
    for (int r = 0; r < RPW; ++r) yr[r] = min(w + 8 * r, H - 1);

    for (int ky = 0; ky < 5; ++ky) {
#pragma unroll 2
        for (int ci = 0; ci < CIN; ++ci) {
            float rv[RPW][HP];
#pragma unroll
            for (int r = 0; r < RPW; ++r) {
                const float *ip = ins + ((yr[r] + ky) * HP) * CIN + ci;
#pragma unroll
                for (int xx = 0; xx < HP; ++xx) rv[r][xx] = ip[xx * CIN];
            }
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) {
                float wv[NCO];
#pragma unroll
                for (int n = 0; n < NCO; ++n)
                    wv[n] = __ldg(wt + ((ky * 5 + kx) * CIN + ci) * COUT + lane + 32 * n);
#pragma unroll
                for (int r = 0; r < RPW; ++r)
#pragma unroll
                    for (int x = 0; x < H; ++x)
#pragma unroll
                        for (int n = 0; n < NCO; ++n)
                            acc[r][x][n] = fmaf(rv[r][x + kx], wv[n], acc[r][x][n]);
            }
        }
    }
    float *dst = out + (int64_t)slot * (H * H * COUT);
#pragma unroll
    for (int r = 0; r < RPW; ++r) {
        const int y = w + 8 * r;
        if (y < H) {
#pragma unroll
            for (int x = 0; x < H; ++x)
#pragma unroll
                for (int n = 0; n < NCO; ++n) dst[(y * H + x) * COUT + lane + 32 * n] = acc[r][x][n];
        }
    }
}

// ------------------------------------------------------------------------------------------
// weight gradient of a 5x5 convolution with CIN = 32: dW[co][ci][tap] = sum over the group's
// samples and pixels of dY[pix][co] * in[pix shifted by tap][ci];  also the bias gradient.
// grid = (n_groups, COUT/16); warp w owns output channels {2w, 2w+1} of the chunk, lane = ci.
// Written straight into the gradient row in the reference's [co][ci][5][5] layout.
// ------------------------------------------------------------------------------------------
template <int COUT, int H>
__global__ void __launch_bounds__(256)
conv_wgrad_kernel(const float *__restrict__ in, const float *__restrict__ dY, Pass p, Workspace ws,
                  int w_off, int b_off)
{
    constexpr int CIN = 32, HP = H + 4;
    extern __shared__ __align__(16) float smw[];
    float *ins = smw;                        // [HP][HP][32]
    float *dys = smw + HP * HP * CIN;        // [H*H][16]
    const int g = blockIdx.x, chunk = blockIdx.y, t = threadIdx.x, lane = t & 31, w = t >> 5;
    if (p.gcount[g] <= 0) return;
    for (int i = t; i < HP * HP * CIN; i += 256) ins[i] = 0.f;
    float acc0[25], acc1[25];
#pragma unroll
    for (int k = 0; k < 25; ++k) { acc0[k] = 0.f; acc1[k] = 0.f; }
    float bs0 = 0.f, bs1 = 0.f;
    const int gs = p.gs;
    for (int s = 0; s < gs; ++s) {
        const int slot = g * gs + s;
        if (p.img_idx[slot] < 0) continue;
        __syncthreads();
        const float *src = in + (int64_t)slot * (H * H * CIN);
        for (int i = t; i < H * H * CIN; i += 256) {
            const int c = i & 31, pix = i >> 5, y = pix / H, x = pix - y * H;
            ins[((y + 2) * HP + x + 2) * CIN + c] = src[i];
        }
        const float *dsrc = dY + (int64_t)slot * (H * H * COUT) + chunk * 16;
        for (int i = t; i < H * H * 16; i += 256) dys[i] = dsrc[(i >> 4) * COUT + (i & 15)];
        __syncthreads();
        int y = 0, x = 0;
        for (int pix = 0; pix < H * H; ++pix) {
            const float2 d = *reinterpret_cast<const float2 *>(dys + pix * 16 + 2 * w);
            const float *ip = ins + (y * HP + x) * CIN + lane;
#pragma unroll
            for (int ky = 0; ky < 5; ++ky)
#pragma unroll
                for (int kx = 0; kx < 5; ++kx) {
                    const float v = ip[(ky * HP + kx) * CIN];
                    acc0[ky * 5 + kx] = fmaf(d.x, v, acc0[ky * 5 + kx]);
                    acc1[ky * 5 + kx] = fmaf(d.y, v, acc1[ky * 5 + kx]);
                }
            if (++x == H) { x = 0; ++y; }
        }
        for (int pix = lane; pix < H * H; pix += 32) {
            bs0 += dys[pix * 16 + 2 * w];
            bs1 += dys[pix * 16 + 2 * w + 1];
        }
    }
    float *G = ws.G + (int64_t)g * ws.ldg;
    const int co0 = chunk * 16 + 2 * w;
    float *g0 = G + w_off + (co0 * CIN + lane) * 25;
    float *g1 = g0 + CIN * 25;
#pragma unroll
    for (int k = 0; k < 25; ++k) { g0[k] = acc0[k]; g1[k] = acc1[k]; }
    bs0 = warp_sum(bs0);
    bs1 = warp_sum(bs1);
    if (lane == 0) { G[b_off + co0] = bs0; G[b_off + co0 + 1] = bs1; }
}

}  // namespace gnl
}  // namespace dls
