// GNLeNet forward / backward kernels, fp32 CUDA-core path (precision 0).
//
// Reference semantics (dasklearn/models/cifar10.py:119-142, SURVEY §8 addendum 1):
//   block 1: conv5x5(3->32,p2)  -> maxpool3/2 -> GN(2,32) -> ReLU
//   block 2: conv5x5(32->32,p2) -> GN(2,32)   -> maxpool3/2 -> ReLU
//   block 3: conv5x5(32->64,p2) -> GN(2,64)   -> maxpool3/2 -> ReLU
//   flatten (C,H,W order) -> Linear(576,10) -> log_softmax ; loss = NLL mean over the batch
// All activations are NHWC fp32 records per sample slot in the caller's workspace.  Every
// reduction uses a fixed order (no float atomics), so results are run-to-run deterministic.
#include "gnlenet.cuh"
#include "conv_simt.cuh"
#include "tc_common.cuh"
#include <math.h>

namespace dls {
namespace gnl {

// ------------------------------------------------------------------------------------------
// workspace
// ------------------------------------------------------------------------------------------
static inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

struct Carver {
    char *base;
    int64_t off;
    template <typename T> T *take(int64_t n) {
        off = align_up(off, 256);
        T *p = base ? reinterpret_cast<T *>(base + off) : nullptr;
        off += n * (int64_t)sizeof(T);
        return p;
    }
};

int64_t carve(Workspace &ws, void *base, int n_groups, int gs, int n_models, bool backward, bool own_G,
              int precision)
{
    Carver c{reinterpret_cast<char *>(base), 0};
    const int64_t S = (int64_t)n_groups * gs;
    ws.a1 = c.take<float>(S * A1_SZ);
    ws.c2 = c.take<float>(S * C2_SZ);
    ws.a2 = c.take<float>(S * A2_SZ);
    ws.c3 = c.take<float>(S * C3_SZ);
    ws.a3 = c.take<float>(S * A3_SZ);
    ws.glog = c.take<float>((int64_t)n_groups * 2);
    ws.prep = c.take<float>((int64_t)n_models * (PREP_SZ + 2400));
    ws.desc_i32 = c.take<int32_t>(S + 2 * (int64_t)n_groups + 16);
    ws.precision = precision;
    ws.prep_tc = precision == 1 ? (void *)c.take<uint16_t>((int64_t)n_models * TC_PREP_ELEMS) : nullptr;
    ws.a1p = ws.a2p = ws.dc2p = ws.dc3p = ws.zplane = nullptr;
    if (precision == 1) {
        ws.zplane = c.take<uint16_t>(RUNP15 * 8);
        ws.a1p = c.take<uint16_t>(S * A1P_SZ);
        ws.a2p = c.take<uint16_t>(S * A2P_SZ);
        if (backward) {
            ws.dc2p = c.take<uint16_t>(S * DC2P_SZ);
            ws.dc3p = c.take<uint16_t>(S * DC3P_SZ);
        }
    }
    ws.p1 = ws.st1 = ws.st2 = ws.st3 = nullptr;
    ws.idx1 = ws.idx2 = ws.idx3 = nullptr;
    ws.da3 = ws.dc3 = ws.da2 = ws.dc2 = ws.da1 = ws.dp1 = ws.gnp = nullptr;
    if (backward) {
        ws.p1 = c.take<float>(S * A1_SZ);
        ws.st1 = c.take<float>(S * 4);
        ws.st2 = c.take<float>(S * 4);
        ws.st3 = c.take<float>(S * 4);
        ws.idx1 = c.take<uint8_t>(S * A1_SZ);
        ws.idx2 = c.take<uint8_t>(S * A2_SZ);
        ws.idx3 = c.take<uint8_t>(S * A3_SZ);
        ws.da3 = c.take<float>(S * A3_SZ);
        ws.dc3 = c.take<float>(S * C3_SZ);
        ws.da2 = c.take<float>(S * A2_SZ);
        ws.dc2 = c.take<float>(S * C2_SZ);
        ws.da1 = c.take<float>(S * A1_SZ);
        ws.dp1 = c.take<float>(S * A1_SZ);
        ws.gnp = c.take<float>(S * GNP_SZ);
        if (own_G) {
            ws.ldg = align_up(P, 64);
            ws.G = c.take<float>((int64_t)n_groups * ws.ldg);
        }
    }
    return align_up(c.off, 256);
}

int64_t workspace_bytes(int n_groups, int gs, int n_models, bool backward, bool own_G, int precision)
{
    Workspace ws;
    return carve(ws, nullptr, n_groups, gs, n_models, backward, own_G, precision);
}

// ------------------------------------------------------------------------------------------
// weight preparation: per model, conv weights re-laid as [tap][cin][cout] so that the lane
// index (= output channel) is contiguous; the dgrad copies are transposed and tap-flipped.
//   prep layout (floats): W1t[75][32] | W2t[25][32][32] | W2d[25][32][32] | W3t[25][32][64] | W3d[25][64][32]
// ------------------------------------------------------------------------------------------
constexpr int PREP_STRIDE = PREP_SZ + 2400;
constexpr int PO_W1T = 0, PO_W2T = 2400, PO_W2D = PO_W2T + W2T_SZ, PO_W3T = PO_W2D + W2T_SZ,
              PO_W3D = PO_W3T + W3T_SZ;

__global__ void __launch_bounds__(256)
prep_weights_kernel(Pass p, float *prep, int backward, const int32_t *model_active)
{
    const int m = blockIdx.y;
    if (model_active && model_active[m] == 0) return;
    const float *row = p.X + (int64_t)p.mrow[m] * p.ldx;
    float *dst = prep + (int64_t)m * PREP_STRIDE;
    const int total = backward ? PREP_STRIDE : (PO_W3T + W3T_SZ);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        float v;
        if (i < PO_W2T) {                       // W1t[k][co], k = (ky*3+ci)*5+kx
            const int co = i & 31, k = i >> 5;
            const int kx = k % 5, ci = (k / 5) % 3, ky = k / 15;
            v = row[OFF_W1 + co * 75 + ci * 25 + ky * 5 + kx];
        } else if (i < PO_W2D) {                // W2t[tap][ci][co]
            const int j = i - PO_W2T, co = j & 31, ci = (j >> 5) & 31, tap = j >> 10;
            v = row[OFF_W2 + co * 800 + ci * 25 + tap];
        } else if (i < PO_W3T) {                // W2d[tap'][co][ci] = W2[co][ci][24-tap']
            const int j = i - PO_W2D, ci = j & 31, co = (j >> 5) & 31, tap = j >> 10;
            v = row[OFF_W2 + co * 800 + ci * 25 + (24 - tap)];
        } else if (i < PO_W3D) {                // W3t[tap][ci][co64]
            const int j = i - PO_W3T, co = j & 63, ci = (j >> 6) & 31, tap = j >> 11;
            v = row[OFF_W3 + co * 800 + ci * 25 + tap];
        } else {                                // W3d[tap'][co64][ci] = W3[co][ci][24-tap']
            const int j = i - PO_W3D, ci = j & 31, co = (j >> 5) & 63, tap = j >> 11;
            v = row[OFF_W3 + co * 800 + ci * 25 + (24 - tap)];
        }
        dst[i] = v;
    }
}

// ------------------------------------------------------------------------------------------
// block 1 forward: conv1 + maxpool + GroupNorm + ReLU, one CTA per sample slot
// ------------------------------------------------------------------------------------------
constexpr int K1_SMEM_FLOATS = 3888 + 2400 + 32 + 8448 + 7200 + 32;
constexpr int K1_SMEM_BYTES = K1_SMEM_FLOATS * 4 + 7200;

__global__ void __launch_bounds__(256)
conv1_pool_gn_relu_kernel(Pass p, Workspace ws, int save_bwd)
{
    extern __shared__ __align__(16) float sm[];
    float *img = sm;              // [36][36][3] zero padded, value/255
    float *wsm = img + 3888;      // [75][32]
    float *bsm = wsm + 2400;      // [32]
    float *ct = bsm + 32;         // conv tile [32][33][8]
    float *ps = ct + 8448;        // pooled [225][32]
    float *red = ps + 7200;       // [32]
    uint8_t *is = reinterpret_cast<uint8_t *>(red + 32);   // [225][32]

    const int slot = blockIdx.x, t = threadIdx.x;
    const int id = p.img_idx[slot];
    if (id < 0) return;
    const int g = slot / p.gs;
    int m;
    const float *row = model_row(p, g, m);
    const float *w1t = ws.prep + (int64_t)m * PREP_STRIDE + PO_W1T;

    for (int i = t; i < 3888; i += 256) img[i] = 0.f;
    for (int i = t; i < 2400; i += 256) wsm[i] = w1t[i];
    if (t < 32) bsm[t] = row[OFF_B1 + t];
    __syncthreads();
    const uint8_t *src = p.images + (int64_t)id * IMG_BYTES;
    for (int i = t; i < IMG_BYTES; i += 256) {
        const int y = i / 96, r = i - y * 96;          // r = x*3 + c
        img[((y + 2) * 36 + 2) * 3 + r] = __fdiv_rn((float)src[i], 255.f);
    }
    __syncthreads();

    const int cs = t & 7, y = t >> 3;
    for (int cg = 0; cg < 4; ++cg) {
        const int co = cg * 8 + cs;
        float acc[32];
        const float b = bsm[co];
#pragma unroll
        for (int x = 0; x < 32; ++x) acc[x] = b;
        for (int ky = 0; ky < 5; ++ky) {
#pragma unroll
            for (int ci = 0; ci < 3; ++ci) {
                float r[36];
                const float *ip = img + ((y + ky) * 36) * 3 + ci;
#pragma unroll
                for (int xx = 0; xx < 36; ++xx) r[xx] = ip[xx * 3];
#pragma unroll
                for (int kx = 0; kx < 5; ++kx) {
                    const float w = wsm[((ky * 3 + ci) * 5 + kx) * 32 + co];
#pragma unroll
                    for (int x = 0; x < 32; ++x) acc[x] = fmaf(r[x + kx], w, acc[x]);
                }
            }
        }
#pragma unroll
        for (int x = 0; x < 32; ++x) ct[(y * 33 + x) * 8 + cs] = acc[x];
        __syncthreads();
        for (int o = t; o < PH1 * PH1 * 8; o += 256) {
            const int c = o & 7, pp = o >> 3, py = pp / PH1, px = pp - py * PH1;
            float best = -INFINITY;
            int bi = 0;
#pragma unroll
            for (int dy = 0; dy < 3; ++dy)
#pragma unroll
                for (int dx = 0; dx < 3; ++dx) {
                    const float v = ct[((2 * py + dy) * 33 + 2 * px + dx) * 8 + c];
                    if (v > best) { best = v; bi = dy * 3 + dx; }   // first maximum wins
                }
            ps[pp * 32 + cg * 8 + c] = best;
            is[pp * 32 + cg * 8 + c] = (uint8_t)bi;
        }
        __syncthreads();
    }

    // GroupNorm(2, 32) over 16 channels x 225 pixels per group, then ReLU
    const int ch = t & 31, grp = ch >> 4;
    float s = 0.f;
    for (int e = t; e < A1_SZ; e += 256) s += ps[e];
    const float sum0 = block_sum(grp == 0 ? s : 0.f, red);
    const float sum1 = block_sum(grp == 1 ? s : 0.f, red);
    const float inv_n = 1.f / 3600.f;
    const float mean0 = sum0 * inv_n, mean1 = sum1 * inv_n;
    const float mean = grp ? mean1 : mean0;
    float q = 0.f;
    for (int e = t; e < A1_SZ; e += 256) { const float d = ps[e] - mean; q = fmaf(d, d, q); }
    float q0 = grp == 0 ? q : 0.f, q1 = grp == 1 ? q : 0.f;
    block_sum2(q0, q1, red);
    const float rstd0 = 1.f / sqrtf(q0 * inv_n + GN_EPS), rstd1 = 1.f / sqrtf(q1 * inv_n + GN_EPS);
    const float rstd = grp ? rstd1 : rstd0;
    const float gam = row[OFF_G1W + ch], bet = row[OFF_G1B + ch];
    float *a1 = ws.a1 + (int64_t)slot * A1_SZ;
    for (int e = t; e < A1_SZ; e += 256) {
        const float v = (ps[e] - mean) * rstd * gam + bet;
        a1[e] = v > 0.f ? v : 0.f;
    }
    if (save_bwd) {
        float *p1 = ws.p1 + (int64_t)slot * A1_SZ;
        uint8_t *i1 = ws.idx1 + (int64_t)slot * A1_SZ;
        for (int e = t; e < A1_SZ; e += 256) { p1[e] = ps[e]; i1[e] = is[e]; }
        if (t == 0) {
            float *st = ws.st1 + (int64_t)slot * 4;
            st[0] = mean0; st[1] = rstd0; st[2] = mean1; st[3] = rstd1;
        }
    }
}

// ------------------------------------------------------------------------------------------
// blocks 2/3 forward tail: GroupNorm(2, C) -> maxpool 3/2 -> ReLU, one CTA per sample slot
// ------------------------------------------------------------------------------------------
template <int C, int H, int PH>
__global__ void __launch_bounds__(256)
gn_pool_relu_kernel(const float *__restrict__ cpre, float *__restrict__ aout, float *__restrict__ stats,
                    uint8_t *__restrict__ idx, Pass p, int gw_off, int gb_off, int save_bwd, int cbias_off,
                    uint16_t *__restrict__ planes)
{
    constexpr int SZ = H * H * C;
    __shared__ __align__(16) float cs[SZ];
    __shared__ __align__(16) float ps[PH == 7 ? PH * PH * C : 4];   // pooled output, staged for the bf16 plane stores
    __shared__ float red[64];
    const int slot = blockIdx.x, t = threadIdx.x;
    if (p.img_idx[slot] < 0) return;
    const int g = slot / p.gs;
    int m;
    const float *row = model_row(p, g, m);
    const float *src = cpre + (int64_t)slot * SZ;
    // The sample is read once with 16-byte loads, all of them in flight together (the kernel is a
    // stream over c: memory-level parallelism is what it needs), and stays in registers for both
    // statistics passes.  A thread's four consecutive channels belong to one GroupNorm group.
    static_assert(SZ % 4 == 0 && C % 4 == 0 && (C / 2) % 4 == 0, "float4 path");
    constexpr int N4 = SZ / 4, NE4 = (N4 + 255) / 256;
    const int c0 = (4 * t) % C, grp = c0 >= (C / 2);              // 1024 % C == 0: fixed channels per thread
    float4 v[NE4];
    // cbias_off >= 0: `cpre` is the convolution without its bias (tensor-core path); it is added here
    const float4 cb = cbias_off >= 0 ? *reinterpret_cast<const float4 *>(row + cbias_off + c0) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < NE4; ++k) {
        const int e4 = t + k * 256;
        v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (e4 < N4) {
            const float4 x = __ldg(reinterpret_cast<const float4 *>(src) + e4);
            v[k] = make_float4(x.x + cb.x, x.y + cb.y, x.z + cb.z, x.w + cb.w);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < NE4; ++k) s += (v[k].x + v[k].y) + (v[k].z + v[k].w);
    float sum0 = grp == 0 ? s : 0.f, sum1 = grp == 1 ? s : 0.f;
    block_sum2(sum0, sum1, red);
    const float inv_n = 1.f / (float)(SZ / 2);
    const float mean0 = sum0 * inv_n, mean1 = sum1 * inv_n;
    const float mean = grp ? mean1 : mean0;
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < NE4; ++k) {
        if (t + k * 256 < N4) {
            const float dx = v[k].x - mean, dy = v[k].y - mean, dz = v[k].z - mean, dw = v[k].w - mean;
            q = fmaf(dx, dx, q); q = fmaf(dy, dy, q); q = fmaf(dz, dz, q); q = fmaf(dw, dw, q);
        }
    }
    float q0 = grp == 0 ? q : 0.f, q1 = grp == 1 ? q : 0.f;
    block_sum2(q0, q1, red);
    const float rstd0 = 1.f / sqrtf(q0 * inv_n + GN_EPS), rstd1 = 1.f / sqrtf(q1 * inv_n + GN_EPS);
    const float rstd = grp ? rstd1 : rstd0;
    const float4 gam = *reinterpret_cast<const float4 *>(row + gw_off + c0);
    const float4 bet = *reinterpret_cast<const float4 *>(row + gb_off + c0);
#pragma unroll
    for (int k = 0; k < NE4; ++k) {
        const int e4 = t + k * 256;
        if (e4 < N4) {
            float4 o;
            o.x = (v[k].x - mean) * rstd * gam.x + bet.x; o.y = (v[k].y - mean) * rstd * gam.y + bet.y;
            o.z = (v[k].z - mean) * rstd * gam.z + bet.z; o.w = (v[k].w - mean) * rstd * gam.w + bet.w;
            reinterpret_cast<float4 *>(cs)[e4] = o;
        }
    }
    __syncthreads();
    float *dst = aout + (int64_t)slot * (PH * PH * C);
    for (int o = t; o < PH * PH * C; o += 256) {
        const int c = o % C, pp = o / C, py = pp / PH, px = pp - py * PH;
        float best = -INFINITY;
        int bi = 0;
#pragma unroll
        for (int dy = 0; dy < 3; ++dy)
#pragma unroll
            for (int dx = 0; dx < 3; ++dx) {
                const float v = cs[((2 * py + dy) * H + 2 * px + dx) * C + c];
                if (v > best) { best = v; bi = dy * 3 + dx; }
            }
        dst[o] = best > 0.f ? best : 0.f;
        if (PH == 7) ps[o] = best > 0.f ? best : 0.f;
        if (save_bwd) idx[(int64_t)slot * (PH * PH * C) + o] = (uint8_t)bi;
    }
    if (PH == 7 && planes) {
        // the block output also leaves as bf16 channel-chunk planes [C/8][RUNP7][8] (what conv3 forward and the conv3
        // weight gradient stage with bulk copies): entry r = py * 9 + px; the two gap pixels after a row are zero
        __syncthreads();
        uint4 *pl = reinterpret_cast<uint4 *>(planes + (int64_t)slot * (C / 8) * RUNP7 * 8);
        for (int e = t; e < PH * PH * (C / 8); e += 256) {
            const int c8 = e % (C / 8), pp = e / (C / 8), py = pp / PH, px = pp - py * PH;
            float o[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = ps[pp * C + c8 * 8 + k];
            pl[c8 * RUNP7 + py * (PH + 2) + px] = tc::pack8_bf16(o);
        }
        for (int e = t; e < (PH - 1) * 2 * (C / 8); e += 256) {
            const int c8 = e % (C / 8), gq = e / (C / 8), py = gq >> 1;
            pl[c8 * RUNP7 + py * (PH + 2) + PH + (gq & 1)] = make_uint4(0u, 0u, 0u, 0u);
        }
    }
    if (save_bwd && t == 0) {
        float *st = stats + (int64_t)slot * 4;
        st[0] = mean0; st[1] = rstd0; st[2] = mean1; st[3] = rstd1;
    }
}

// ------------------------------------------------------------------------------------------
// classifier: Linear(576,10) + log_softmax + NLL.  One CTA per group (minibatch).
//   mode 0 (train): writes dW_fc, db_fc into the gradient row, da3 for every slot, mean loss
//   mode 1 (eval):  per-group loss sum and number of argmax hits (+ optional log-probs)
// a3 is NHWC [hw][c]; the reference flattens C,H,W so feature f = c*9 + hw.
// ------------------------------------------------------------------------------------------
constexpr int CLS_THREADS = 576;       // one thread per classifier input feature in the weight-gradient loop

__global__ void __launch_bounds__(CLS_THREADS)
classifier_kernel(Pass p, Workspace ws, int mode, float *loss_out, float *out_logits, int logits_stride_model,
                  int logits_group0)
{
    __shared__ float wfc[NCLS * FEAT];
    __shared__ float bfc[NCLS];
    __shared__ float lg[DLS_MAX_GROUP][NCLS];     // logits -> dlogits
    __shared__ float lossv[DLS_MAX_GROUP];
    __shared__ int hit[DLS_MAX_GROUP];
    const int g = blockIdx.x, t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int cnt = p.gcount[g];
    if (cnt <= 0) return;
    int m;
    const float *row = model_row(p, g, m);
    for (int i = t; i < NCLS * FEAT; i += CLS_THREADS) wfc[i] = row[OFF_FCW + i];
    if (t < NCLS) bfc[t] = row[OFF_FCB + t];
    __syncthreads();
    const int gs = p.gs;
    for (int s = w; s < gs; s += CLS_THREADS / 32) {
        const int slot = g * gs + s;
        if (p.img_idx[slot] < 0) { if (lane == 0) { lossv[s] = 0.f; hit[s] = 0; } continue; }
        const float *a3 = ws.a3 + (int64_t)slot * A3_SZ;
        float part[NCLS];
#pragma unroll
        for (int o = 0; o < NCLS; ++o) part[o] = 0.f;
        for (int i = lane; i < A3_SZ; i += 32) {
            const int hw = i >> 6, c = i & 63, f = c * 9 + hw;
            const float a = a3[i];
#pragma unroll
            for (int o = 0; o < NCLS; ++o) part[o] = fmaf(a, wfc[o * FEAT + f], part[o]);
        }
#pragma unroll
        for (int o = 0; o < NCLS; ++o) part[o] = warp_sum(part[o]) + bfc[o];
        if (lane == 0) {
            float mx = part[0];
            int am = 0;
#pragma unroll
            for (int o = 1; o < NCLS; ++o) if (part[o] > mx) { mx = part[o]; am = o; }
            float se = 0.f;
#pragma unroll
            for (int o = 0; o < NCLS; ++o) se += expf(part[o] - mx);
            const float lse = mx + logf(se);
            const int y = (int)p.labels[p.img_idx[slot]];
            float logp_y = 0.f;
#pragma unroll
            for (int o = 0; o < NCLS; ++o) {
                const float logp = part[o] - lse;
                if (o == y) logp_y = logp;
                if (mode == 0) lg[s][o] = (expf(logp) - (o == y ? 1.f : 0.f)) / (float)cnt;
                else lg[s][o] = logp;
            }
            lossv[s] = -logp_y;
            hit[s] = (am == y) ? 1 : 0;
        }
    }
    __syncthreads();
    if (t == 0) {
        float ls = 0.f;
        int hs = 0;
        for (int s = 0; s < gs; ++s) if (p.img_idx[g * gs + s] >= 0) { ls += lossv[s]; hs += hit[s]; }
        if (mode == 0) { if (loss_out) loss_out[g] = ls / (float)cnt; }
        else { ws.glog[(int64_t)g * 2] = ls; ws.glog[(int64_t)g * 2 + 1] = (float)hs; }
    }
    if (mode == 1) {
        if (out_logits) {
            // eval groups are (model, chunk of gs test samples); img_idx is the test sample id
            for (int i = t; i < gs * NCLS; i += CLS_THREADS) {
                const int s = i / NCLS, o = i - s * NCLS;
                const int id = p.img_idx[g * gs + s];
                if (id >= 0) out_logits[(int64_t)m * logits_stride_model + (int64_t)id * NCLS + o] = lg[s][o];
            }
        }
        return;
    }
    // ---- train: gradients of the classifier and of its input ----
    float *G = ws.G + (int64_t)g * ws.ldg;
    for (int i = t; i < A3_SZ; i += CLS_THREADS) {          // i indexes a3 NHWC; f = c*9+hw
        const int hw = i >> 6, c = i & 63, f = c * 9 + hw;
        float acc[NCLS];
#pragma unroll
        for (int o = 0; o < NCLS; ++o) acc[o] = 0.f;
        for (int s = 0; s < gs; ++s) {
            const int slot = g * gs + s;
            if (p.img_idx[slot] < 0) continue;
            const float a = ws.a3[(int64_t)slot * A3_SZ + i];
#pragma unroll
            for (int o = 0; o < NCLS; ++o) acc[o] = fmaf(lg[s][o], a, acc[o]);
        }
#pragma unroll
        for (int o = 0; o < NCLS; ++o) G[OFF_FCW + o * FEAT + f] = acc[o];
    }
    if (t < NCLS) {
        float b = 0.f;
        for (int s = 0; s < gs; ++s) if (p.img_idx[g * gs + s] >= 0) b += lg[s][t];
        G[OFF_FCB + t] = b;
    }
    for (int j = t; j < gs * A3_SZ; j += CLS_THREADS) {
        const int s = j / A3_SZ, i = j - s * A3_SZ;
        const int slot = g * gs + s;
        if (p.img_idx[slot] < 0) continue;
        const int hw = i >> 6, c = i & 63, f = c * 9 + hw;
        float d = 0.f;
#pragma unroll
        for (int o = 0; o < NCLS; ++o) d = fmaf(lg[s][o], wfc[o * FEAT + f], d);
        ws.da3[(int64_t)slot * A3_SZ + i] = d;
    }
}

// ------------------------------------------------------------------------------------------
// backward of {GroupNorm -> maxpool -> ReLU} (POOL = true, blocks 2/3) or {GroupNorm -> ReLU}
// (POOL = false, block 1 after the pool).  One CTA per sample slot.
//   in : da   gradient w.r.t. the block output  [PH*PH*C]  (POOL) / [H*H*C] (!POOL)
//        aout block output (ReLU mask), idx arg-max positions, cpre the GroupNorm input
//   out: dc   gradient w.r.t. the GroupNorm input [H*H*C]; per-slot dgamma/dbeta partials
// Max-pool backward sends each output gradient to the FIRST maximum of its window and sums
// overlapping windows in output-scan order (the order PyTorch's CPU kernel accumulates in).
// ------------------------------------------------------------------------------------------
template <int C, int H, int PH, bool POOL>
__global__ void __launch_bounds__(256, (POOL && H == 15) ? 4 : 4)
gn_block_bwd_kernel(const float *__restrict__ da, const float *__restrict__ aout,
                    const uint8_t *__restrict__ idx, const float *__restrict__ cpre,
                    const float *__restrict__ stats, float *__restrict__ dc, float *__restrict__ gnp,
                    Pass p, int gw_off, int gnp_off, int cbias_off, uint16_t *__restrict__ planes)
{
    constexpr int SZ = H * H * C;
    constexpr int OSZ = POOL ? PH * PH * C : SZ;
    constexpr int RUNP = H == 15 ? RUNP15 : RUNP7;
    __shared__ float dg[SZ];
    __shared__ float red[32];
    const int slot = blockIdx.x, t = threadIdx.x;
    if (p.img_idx[slot] < 0) return;
    const int g = slot / p.gs;
    int m;
    const float *row = model_row(p, g, m);
    const float *das = da + (int64_t)slot * OSZ;
    const float *as = aout + (int64_t)slot * OSZ;
    const float *cp = cpre + (int64_t)slot * SZ;
    const float *st = stats + (int64_t)slot * 4;
    const int ch = t % C, grp = ch >= (C / 2);
    const float mean = st[grp * 2], rstd = st[grp * 2 + 1];
    const float gam = row[gw_off + ch];
    const float cb = cbias_off >= 0 ? row[cbias_off + ch] : 0.f;    // `cpre` without the conv bias (tensor-core path)

    // Every global load of the sample is issued up front — the pooled gradients with their ReLU masks and arg-max
    // bytes, and the pre-norm activations — so the kernel pays ONE memory latency instead of one per phase.
    constexpr int NE = (SZ + 255) / 256;
    constexpr int NP = POOL ? (OSZ + 255) / 256 : 1;
    float xh[NE];
    float pv[NP];
    int ppos[NP], pcol[NP];
    if (POOL) {
        const uint8_t *is = idx + (int64_t)slot * OSZ;
#pragma unroll
        for (int k = 0; k < NP; ++k) {
            const int o = t + k * 256;
            pv[k] = 0.f; ppos[k] = 0; pcol[k] = -1;
            if (o < OSZ) {
                const int c = o % C, pp = o / C, py = pp / PH, px = pp - py * PH;
                const int li = is[o];
                pv[k] = as[o] > 0.f ? das[o] : 0.f;
                ppos[k] = ((2 * py + li / 3) * H + 2 * px + li % 3) * C + c;
                pcol[k] = (py & 1) * 2 + (px & 1);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const int e = t + k * 256;
        xh[k] = e < SZ ? cp[e] : 0.f;
    }
    if (POOL) {
        // route each pooled gradient to the arg-max position of its 3x3/2 window.  Windows of one colour
        // (py%2, px%2) are disjoint, so four passes add without races and in a fixed order.
        for (int e = t; e < SZ; e += 256) dg[e] = 0.f;
        __syncthreads();
#pragma unroll
        for (int colour = 0; colour < 4; ++colour) {
#pragma unroll
            for (int k = 0; k < NP; ++k)
                if (pcol[k] == colour && pv[k] != 0.f) dg[ppos[k]] += pv[k];
            __syncthreads();
        }
    }
    // One pass over the sample: x-hat and the incoming gradient stay in registers between the statistics
    // and the output pass.  Block 1 (no pooling in front of the ReLU) re-derives the ReLU mask from x-hat instead
    // of reading the activations back.
    const float bet = POOL ? 0.f : row[gw_off + C + ch];
    float s1 = 0.f, s2 = 0.f, dgam = 0.f, dbet = 0.f;     // 256 % C == 0: a thread only sees channel `ch`
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const int e = t + k * 256;
        if (e < SZ) {
            const float xhat = ((xh[k] + cb) - mean) * rstd;
            const float d = POOL ? dg[e] : ((xhat * gam + bet) > 0.f ? das[e] : 0.f);
            xh[k] = xhat;
            if (!POOL) dg[e] = d;                          // own element: no barrier needed before the second pass
            const float dxh = d * gam;
            s1 += dxh;
            s2 = fmaf(dxh, xhat, s2);
            dgam = fmaf(d, xhat, dgam);
            dbet += d;
        }
    }
    // the four group sums in one reduction: warp shuffles, then the eight warp partials in a fixed order
    float r0 = warp_sum(grp == 0 ? s1 : 0.f), r1 = warp_sum(grp == 1 ? s1 : 0.f);
    float r2 = warp_sum(grp == 0 ? s2 : 0.f), r3 = warp_sum(grp == 1 ? s2 : 0.f);
    if ((t & 31) == 0) { red[(t >> 5) * 4] = r0; red[(t >> 5) * 4 + 1] = r1; red[(t >> 5) * 4 + 2] = r2; red[(t >> 5) * 4 + 3] = r3; }
    __syncthreads();
    float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { a0 += red[w * 4]; a1 += red[w * 4 + 1]; b0 += red[w * 4 + 2]; b1 += red[w * 4 + 3]; }
    const float inv_n = 1.f / (float)(SZ / 2);
    const float m1 = (grp ? a1 : a0) * inv_n, m2 = (grp ? b1 : b0) * inv_n;
    float *dst = dc + (int64_t)slot * SZ;
    float dbias = 0.f;                                     // sum of dc per channel = conv bias gradient partial
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const int e = t + k * 256;
        if (e < SZ) {
            const float xhat = xh[k];
            const float v = rstd * (dg[e] * gam - m1 - xhat * m2);
            if (planes) dg[e] = v;                         // own element: staged for the plane stores below
            else dst[e] = v;
            dbias += v;
        }
    }
    if (POOL && planes) {
        // precision 1: the gradient w.r.t. the conv output leaves ONLY as bf16 channel-chunk planes [C/8][RUNP][8]
        // (entry r = y * (H + 2) + x, gap pixels zero): both its consumers, the data-gradient and the weight-gradient
        // convolutions, stage it with bulk copies
        __syncthreads();
        uint4 *pl = reinterpret_cast<uint4 *>(planes + (int64_t)slot * (C / 8) * RUNP * 8);
        for (int e = t; e < H * H * (C / 8); e += 256) {
            const int c8 = e % (C / 8), pix = e / (C / 8), y = pix / H, x = pix - y * H;
            float o[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = dg[pix * C + c8 * 8 + k];
            pl[c8 * RUNP + y * (H + 2) + x] = tc::pack8_bf16(o);
        }
        for (int e = t; e < (H - 1) * 2 * (C / 8); e += 256) {
            const int c8 = e % (C / 8), gq = e / (C / 8), y = gq >> 1;
            pl[c8 * RUNP + y * (H + 2) + H + (gq & 1)] = make_uint4(0u, 0u, 0u, 0u);
        }
    }
    // per-channel parameter-gradient partials of this sample: combine the 256/C threads of a channel
    // in a fixed order (dg is dead now: reuse it as scratch after a barrier)
    __syncthreads();
    dg[t] = dgam;
    dg[256 + t] = dbet;
    dg[512 + t] = dbias;
    __syncthreads();
    if (t < C) {
        float sg = 0.f, sb = 0.f, sc = 0.f;
#pragma unroll
        for (int j = 0; j < 256 / C; ++j) { sg += dg[t + j * C]; sb += dg[256 + t + j * C]; sc += dg[512 + t + j * C]; }
        float *o = gnp + (int64_t)slot * GNP_SZ + gnp_off;
        o[t] = sg;
        o[C + t] = sb;
        o[2 * C + t] = sc;
    }
}

// ------------------------------------------------------------------------------------------
// conv1 weight gradient.  The gradient w.r.t. the conv1 output is non-zero only at the arg-max
// position of every pooling window, so dW1[co] = sum_p dp1[p][co] * patch(argmax(p, co)).
// grid = (n_groups, 4); warp = output channel, lanes cover the 75 (ci,ky,kx) weights.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
conv1_wgrad_kernel(Pass p, Workspace ws)
{
    __shared__ float img[3888];
    __shared__ float dps[PH1 * PH1 * 8];
    __shared__ uint8_t ids[PH1 * PH1 * 8];
    const int g = blockIdx.x, chunk = blockIdx.y, t = threadIdx.x, lane = t & 31, w = t >> 5;
    if (p.gcount[g] <= 0) return;
    for (int i = t; i < 3888; i += 256) img[i] = 0.f;
    int off[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int e = min(lane + 32 * k, 74);
        const int ci = e / 25, r = e - ci * 25, ky = r / 5, kx = r - ky * 5;
        off[k] = (ky * 36 + kx) * 3 + ci;
    }
    float acc[3] = {0.f, 0.f, 0.f};
    float bsum = 0.f;
    const int gs = p.gs;
    for (int s = 0; s < gs; ++s) {
        const int slot = g * gs + s;
        const int id = p.img_idx[slot];
        if (id < 0) continue;
        __syncthreads();
        const uint8_t *src = p.images + (int64_t)id * IMG_BYTES;
        for (int i = t; i < IMG_BYTES; i += 256) {
            const int y = i / 96, r = i - y * 96;
            img[((y + 2) * 36 + 2) * 3 + r] = __fdiv_rn((float)src[i], 255.f);
        }
        const float *dp = ws.dp1 + (int64_t)slot * A1_SZ + chunk * 8;
        const uint8_t *ip = ws.idx1 + (int64_t)slot * A1_SZ + chunk * 8;
        for (int i = t; i < PH1 * PH1 * 8; i += 256) {
            dps[i] = dp[(i >> 3) * 32 + (i & 7)];
            ids[i] = ip[(i >> 3) * 32 + (i & 7)];
        }
        __syncthreads();
        int py = 0, px = 0;
        for (int pp = 0; pp < PH1 * PH1; ++pp) {
            const float d = dps[pp * 8 + w];
            const int ii = ids[pp * 8 + w];
            const int Y = 2 * py + ii / 3, X = 2 * px + ii % 3;
            const float *bp = img + (Y * 36 + X) * 3;
#pragma unroll
            for (int k = 0; k < 3; ++k) acc[k] = fmaf(d, bp[off[k]], acc[k]);
            bsum += d;
            if (++px == PH1) { px = 0; ++py; }
        }
    }
    float *G = ws.G + (int64_t)g * ws.ldg;
    const int co = chunk * 8 + w;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int e = lane + 32 * k;
        if (e < 75) G[OFF_W1 + co * 75 + e] = acc[k];
    }
    if (lane == 0) G[OFF_B1 + co] = bsum;
}

// Sum the per-sample partials (GroupNorm gamma/beta and conv bias gradients) of a group in slot order.
__global__ void __launch_bounds__(GNP_SZ)
gn_param_reduce_kernel(Pass p, Workspace ws)
{
    const int g = blockIdx.x, t = threadIdx.x;
    if (p.gcount[g] <= 0) return;
    float s = 0.f;
    for (int k = 0; k < p.gs; ++k) {
        const int slot = g * p.gs + k;
        if (p.img_idx[slot] >= 0) s += ws.gnp[(int64_t)slot * GNP_SZ + t];
    }
    int dst;                                  // per layer: [gamma C][beta C][conv bias C]
    if (t < 32) dst = OFF_G1W + t;
    else if (t < 64) dst = OFF_G1B + (t - 32);
    else if (t < 96) dst = OFF_B1 + (t - 64);
    else if (t < 128) dst = OFF_G2W + (t - 96);
    else if (t < 160) dst = OFF_G2B + (t - 128);
    else if (t < 192) dst = OFF_B2 + (t - 160);
    else if (t < 256) dst = OFF_G3W + (t - 192);
    else if (t < 320) dst = OFF_G3B + (t - 256);
    else dst = OFF_B3 + (t - 320);
    ws.G[(int64_t)g * ws.ldg + dst] = s;
}

// ------------------------------------------------------------------------------------------
// host-side launch sequences
// ------------------------------------------------------------------------------------------
static bool g_attr_done[MAX_DEVICES] = {};      // per device: cudaFuncSetAttribute applies to the current device
static int set_attrs()
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= MAX_DEVICES) return DLS_E_TOO_BIG;
    if (g_attr_done[dev]) return 0;
#define SET_SMEM(k, bytes)                                                                   \
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)); \
    if (e != cudaSuccess) return (int)e;
    SET_SMEM(conv1_pool_gn_relu_kernel, K1_SMEM_BYTES);
    SET_SMEM((conv5_kernel<32, 32, 15, true>), 19 * 19 * 32 * 4);
    SET_SMEM((conv5_kernel<32, 32, 15, false>), 19 * 19 * 32 * 4);
    SET_SMEM((conv5_kernel<32, 64, 7, true>), 11 * 11 * 32 * 4);
    SET_SMEM((conv5_kernel<64, 32, 7, false>), 11 * 11 * 64 * 4);
    SET_SMEM((conv_wgrad_kernel<32, 15>), (19 * 19 * 32 + 225 * 16) * 4);
    SET_SMEM((conv_wgrad_kernel<64, 7>), (11 * 11 * 32 + 49 * 16) * 4);
#undef SET_SMEM
    g_attr_done[dev] = true;
    return 0;
}

int prep_weights(const Pass &p, Workspace &ws, bool backward, cudaStream_t st)
{
    dim3 grid(8, p.n_models);
    // in train passes model index == group index, so gcount doubles as the "model active" mask
    const int32_t *active = p.gmodel ? nullptr : p.gcount;
    if (ws.precision == 1) return prep_weights_tc(p, ws, backward, st);   // bf16 UMMA layouts only
    prep_weights_kernel<<<grid, 256, 0, st>>>(p, ws.prep, backward ? 1 : 0, active);
    DLS_LAUNCH_CHECK();
    return 0;
}

int forward(const Pass &p, Workspace &ws, bool save, cudaStream_t st)
{
    int rc = set_attrs();
    if (rc) return rc;
    const int S = p.n_groups * p.gs;
    const int sv = save ? 1 : 0;
    // the grouped tensor-core convolutions leave the conv bias to the GroupNorm kernels
    const bool nobias = ws.precision == 1 && !tune_value(TUNE_CONV_UNGROUPED);
    if (ws.precision == 1) {
        if ((rc = conv1_fwd_tc(p, ws, save, st))) return rc;
    } else {
        conv1_pool_gn_relu_kernel<<<S, 256, K1_SMEM_BYTES, st>>>(p, ws, sv);
        DLS_LAUNCH_CHECK();
    }
    if (ws.precision == 1) {
        if ((rc = conv2_fwd_tc(p, ws, st))) return rc;
    } else {
        conv5_kernel<32, 32, 15, true><<<S, 256, 19 * 19 * 32 * 4, st>>>(ws.a1, ws.c2, p, ws.prep, PREP_STRIDE, PO_W2T, OFF_B2);
        DLS_LAUNCH_CHECK();
    }
    gn_pool_relu_kernel<32, 15, 7><<<S, 256, 0, st>>>(ws.c2, ws.a2, ws.st2, ws.idx2, p, OFF_G2W, OFF_G2B, sv, nobias ? OFF_B2 : -1, nobias ? ws.a2p : nullptr);
    DLS_LAUNCH_CHECK();
    if (ws.precision == 1) {
        if ((rc = conv3_fwd_tc(p, ws, st))) return rc;
    } else {
        conv5_kernel<32, 64, 7, true><<<S, 256, 11 * 11 * 32 * 4, st>>>(ws.a2, ws.c3, p, ws.prep, PREP_STRIDE, PO_W3T, OFF_B3);
        DLS_LAUNCH_CHECK();
    }
    gn_pool_relu_kernel<64, 7, 3><<<S, 256, 0, st>>>(ws.c3, ws.a3, ws.st3, ws.idx3, p, OFF_G3W, OFF_G3B, sv, nobias ? OFF_B3 : -1, nullptr);
    DLS_LAUNCH_CHECK();
    return 0;
}

int classifier(const Pass &p, Workspace &ws, int mode, float *loss_out, float *out_logits,
               int logits_stride_model, cudaStream_t st)
{
    if (p.gs > DLS_MAX_GROUP) return DLS_E_SHAPE;
    classifier_kernel<<<p.n_groups, CLS_THREADS, 0, st>>>(p, ws, mode, loss_out, out_logits, logits_stride_model, 0);
    DLS_LAUNCH_CHECK();
    return 0;
}

// The weight-gradient kernels only feed the optimizer, so in tensor-core mode they run on a side
// stream next to the data-gradient chain (fork/join with events): with one CTA per node, 256 nodes
// fill 1.73 waves of the 148 SMs, and the other stream's kernels use the SMs an unfinished wave
// leaves idle.  The side stream is created once per device and joined before this call returns.
static SideStream g_side[MAX_DEVICES][2 * MAX_LANES];   // per device: [2*lane] weight-gradient side stream, [2*lane+1] the lane's main stream

int side_stream(SideStream **out, int which)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= MAX_DEVICES || which < 0 || which >= 2 * MAX_LANES) return DLS_E_TOO_BIG;
    SideStream &ss = g_side[dev][which];
    if (!ss.s) {
        if ((e = cudaStreamCreateWithFlags(&ss.s, cudaStreamNonBlocking)) != cudaSuccess) return (int)e;
        if ((e = cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming)) != cudaSuccess) return (int)e;
        if ((e = cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming)) != cudaSuccess) return (int)e;
    }
    *out = &ss;
    return 0;
}

#define DLS_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return (int)e__; } while (0)

int backward(const Pass &p, Workspace &ws, cudaStream_t st, int lane)
{
    int rc = set_attrs();
    if (rc) return rc;
    const int S = p.n_groups * p.gs;
    if (ws.precision == 1) {
        SideStream *ss = nullptr;
        if ((rc = side_stream(&ss, 2 * lane))) return rc;
        const bool nobias = !tune_value(TUNE_CONV_UNGROUPED);     // conv outputs carry no bias (conv_tc3_kernel)
        gn_block_bwd_kernel<64, 7, 3, true><<<S, 256, 0, st>>>(ws.da3, ws.a3, ws.idx3, ws.c3, ws.st3, ws.dc3,
                                                              ws.gnp, p, OFF_G3W, 192, nobias ? OFF_B3 : -1, nobias ? ws.dc3p : nullptr);
        DLS_LAUNCH_CHECK();
        DLS_CUDA(cudaEventRecord(ss->fork, st));
        DLS_CUDA(cudaStreamWaitEvent(ss->s, ss->fork, 0));
        if ((rc = conv3_wgrad_tc(p, ws, ss->s))) return rc;
        if ((rc = conv3_dgrad_tc(p, ws, st))) return rc;
        gn_block_bwd_kernel<32, 15, 7, true><<<S, 256, 0, st>>>(ws.da2, ws.a2, ws.idx2, ws.c2, ws.st2, ws.dc2,
                                                               ws.gnp, p, OFF_G2W, 96, nobias ? OFF_B2 : -1, nobias ? ws.dc2p : nullptr);
        DLS_LAUNCH_CHECK();
        DLS_CUDA(cudaEventRecord(ss->fork, st));
        DLS_CUDA(cudaStreamWaitEvent(ss->s, ss->fork, 0));
        if ((rc = conv2_wgrad_tc(p, ws, ss->s))) return rc;
        if ((rc = conv2_dgrad_tc(p, ws, st))) return rc;
        gn_block_bwd_kernel<32, 15, 15, false><<<S, 256, 0, st>>>(ws.da1, ws.a1, nullptr, ws.p1, ws.st1, ws.dp1,
                                                                 ws.gnp, p, OFF_G1W, 0, -1, nullptr);
        DLS_LAUNCH_CHECK();
        if ((rc = conv1_wgrad_tc(p, ws, st))) return rc;
        gn_param_reduce_kernel<<<p.n_groups, GNP_SZ, 0, st>>>(p, ws);
        DLS_LAUNCH_CHECK();
        DLS_CUDA(cudaEventRecord(ss->join, ss->s));
        DLS_CUDA(cudaStreamWaitEvent(st, ss->join, 0));
        return 0;
    }
    // block 3
    gn_block_bwd_kernel<64, 7, 3, true><<<S, 256, 0, st>>>(ws.da3, ws.a3, ws.idx3, ws.c3, ws.st3, ws.dc3,
                                                          ws.gnp, p, OFF_G3W, 192, -1, nullptr);
    DLS_LAUNCH_CHECK();
    if (ws.precision == 1) {
        if ((rc = conv3_wgrad_tc(p, ws, st))) return rc;
    } else {
        conv_wgrad_kernel<64, 7><<<dim3(p.n_groups, 4), 256, (11 * 11 * 32 + 49 * 16) * 4, st>>>(
            ws.a2, ws.dc3, p, ws, OFF_W3, OFF_B3);
        DLS_LAUNCH_CHECK();
    }
    if (ws.precision == 1) {
        if ((rc = conv3_dgrad_tc(p, ws, st))) return rc;
    } else {
        conv5_kernel<64, 32, 7, false><<<S, 256, 11 * 11 * 64 * 4, st>>>(ws.dc3, ws.da2, p, ws.prep, PREP_STRIDE, PO_W3D, 0);
        DLS_LAUNCH_CHECK();
    }
    // block 2
    gn_block_bwd_kernel<32, 15, 7, true><<<S, 256, 0, st>>>(ws.da2, ws.a2, ws.idx2, ws.c2, ws.st2, ws.dc2,
                                                           ws.gnp, p, OFF_G2W, 96, -1, nullptr);
    DLS_LAUNCH_CHECK();
    if (ws.precision == 1) {
        if ((rc = conv2_wgrad_tc(p, ws, st))) return rc;
    } else {
        conv_wgrad_kernel<32, 15><<<dim3(p.n_groups, 2), 256, (19 * 19 * 32 + 225 * 16) * 4, st>>>(
            ws.a1, ws.dc2, p, ws, OFF_W2, OFF_B2);
        DLS_LAUNCH_CHECK();
    }
    if (ws.precision == 1) {
        if ((rc = conv2_dgrad_tc(p, ws, st))) return rc;
    } else {
        conv5_kernel<32, 32, 15, false><<<S, 256, 19 * 19 * 32 * 4, st>>>(ws.dc2, ws.da1, p, ws.prep, PREP_STRIDE, PO_W2D, 0);
        DLS_LAUNCH_CHECK();
    }
    // block 1 (GroupNorm -> ReLU after the pool), then the sparse conv1 weight gradient
    gn_block_bwd_kernel<32, 15, 15, false><<<S, 256, 0, st>>>(ws.da1, ws.a1, nullptr, ws.p1, ws.st1, ws.dp1,
                                                             ws.gnp, p, OFF_G1W, 0, -1, nullptr);
    DLS_LAUNCH_CHECK();
    if (ws.precision == 1) {
        if ((rc = conv1_wgrad_tc(p, ws, st))) return rc;
    } else {
        conv1_wgrad_kernel<<<dim3(p.n_groups, 4), 256, 0, st>>>(p, ws);
        DLS_LAUNCH_CHECK();
    }
    gn_param_reduce_kernel<<<p.n_groups, GNP_SZ, 0, st>>>(p, ws);
    DLS_LAUNCH_CHECK();
    return 0;
}

}  // namespace gnl
}  // namespace dls
