// FEMNIST CNN train / eval waves.  precision 0: fp32 CUDA cores throughout.  precision 1: conv2 (94 % of the
// FLOPs) forward / data gradient / weight gradient on tensor cores (tc_conv.cu: bf16 operands, fp32 accumulate),
// fed with bf16 channel-chunk planes that conv1 and the pool backward write; everything else fp32.
//
// Architecture (reference: dasklearn/models/femnist.py:45-85; the reference itself cannot import
// this model — its dataset module is missing — so parity is against the CPU restatement in
// oracle/oracle.py, "parity unpinned"):
//   conv5x5(1->32,p2) -> ReLU -> maxpool 2/2 -> conv5x5(32->64,p2) -> ReLU -> maxpool 2/2
//   -> flatten (C,H,W) 3136 -> Linear 512 -> ReLU -> Linear 62 ; loss = cross entropy (mean)
// Flat parameter offsets (parameters() order): conv1.w 0 | conv1.b 800 | conv2.w 832 | conv2.b 52032 |
//   fc1.w 52096 [512][3136] | fc1.b 1657728 | fc2.w 1658240 [62][512] | fc2.b 1689984 ; P = 1690046.
// The per-node fc1 weight (6.4 MB) makes this model HBM-bound: every step streams it for the
// forward, the data gradient, the weight gradient and the optimizer update.
#include "gnlenet.cuh"
#include "conv_simt.cuh"
#include "tc_common.cuh"
#include <math.h>

namespace dls {
namespace fem {

using gnl::Pass;
using gnl::model_row;
using gnl::SgdFuse;

// fc1 on tensor cores (tc_fc.cu)
int fc_nsp(int gs);
int fc1_fwd_tc(const Pass &p, const uint16_t *a2k, float *h1, int w_off, int b_off, cudaStream_t st);
int fc1_dgrad_tc(const Pass &p, const uint16_t *dh1k, float *da2, int w_off, cudaStream_t st);
int fc1_wgrad_tc(const Pass &p, const uint16_t *a2k, const uint16_t *dh1k, const float *dh1, float *G, int64_t ldg,
                 const SgdFuse &sgd, int w_off, int b_off, cudaStream_t st);

constexpr int P = DLS_FEMNIST_P;
constexpr int OFF_W1 = 0, OFF_B1 = 800, OFF_W2 = 832, OFF_B2 = 52032, OFF_FC1W = 52096, OFF_FC1B = 1657728,
              OFF_FC2W = 1658240, OFF_FC2B = 1689984;
constexpr int IMG_BYTES = 784, NCLS = 62, FEAT = 3136, HID = 512;
constexpr int A1_SZ = 196 * 32, C2_SZ = 196 * 64, A2_SZ = 49 * 64;
constexpr int PREP_STRIDE = 800 + 2 * 51200;          // W1t[25][32] | W2t[25][32][64] | W2d[25][64][32]
constexpr int PO_W1T = 0, PO_W2T = 800, PO_W2D = 800 + 51200;

struct Ws {
    float *a1, *c2, *a2, *h1, *lg;                      // forward (lg = logits -> dlogits)
    uint8_t *idx1, *idx2;
    float *dh1, *da2, *dc2, *da1;                      // backward
    float *prep, *glog, *G;
    int64_t ldg;
    int32_t *desc_i32;
    // precision 1: bf16 planes [slot][C/8][RUNP14][8] (pixel r = y * 16 + x, gap pixels zero) + bf16 weight copies
    uint16_t *a1p, *dc2p, *zplane;
    void *prep_tc;
    // fc1 operands (tc_fc.cu): a2k [node][392][nsp][8], dh1k [node][64][nsp][8] bf16, rows of missing samples zero
    uint16_t *a2k, *dh1k;
    int nsp;
    SgdFuse sgd;                                        // optimizer step fused into the fc1 weight gradient (train waves)
    int tc;
};
constexpr int A1P_SZ = 4 * gnl::RUNP14 * 8, DC2P_SZ = 8 * gnl::RUNP14 * 8;      // bf16 elements per sample

static inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }
struct Carver {
    char *base; int64_t off;
    template <typename T> T *take(int64_t n) {
        off = align_up(off, 256);
        T *p = base ? reinterpret_cast<T *>(base + off) : nullptr;
        off += n * (int64_t)sizeof(T);
        return p;
    }
};

static int64_t carve(Ws &ws, void *base, int n_groups, int gs, int n_models, bool backward, bool own_G)
{
    Carver c{reinterpret_cast<char *>(base), 0};
    const int64_t S = (int64_t)n_groups * gs;
    ws.tc = 0;
    ws.a1p = c.take<uint16_t>(S * A1P_SZ);               // sized for either precision
    ws.prep_tc = c.take<uint16_t>((int64_t)n_models * gnl::TC_PREP_ELEMS);
    ws.zplane = c.take<uint16_t>(gnl::RUNP15 * 8);
    ws.dc2p = backward ? c.take<uint16_t>(S * DC2P_SZ) : nullptr;
    ws.nsp = fc_nsp(gs);
    ws.a2k = c.take<uint16_t>((int64_t)n_groups * (FEAT / 8) * ws.nsp * 8);
    ws.dh1k = backward ? c.take<uint16_t>((int64_t)n_groups * (HID / 8) * ws.nsp * 8) : nullptr;
    ws.sgd = SgdFuse();
    ws.a1 = c.take<float>(S * A1_SZ);
    ws.c2 = c.take<float>(S * C2_SZ);
    ws.a2 = c.take<float>(S * A2_SZ);
    ws.h1 = c.take<float>(S * HID);
    ws.lg = c.take<float>(S * 64);
    ws.glog = c.take<float>((int64_t)n_groups * 2);
    ws.prep = c.take<float>((int64_t)n_models * PREP_STRIDE);
    ws.desc_i32 = c.take<int32_t>(S + 2 * (int64_t)n_groups + 16);
    ws.idx1 = ws.idx2 = nullptr;
    ws.dh1 = ws.da2 = ws.dc2 = ws.da1 = ws.G = nullptr;
    ws.ldg = 0;
    if (backward) {
        ws.idx1 = c.take<uint8_t>(S * A1_SZ);
        ws.idx2 = c.take<uint8_t>(S * A2_SZ);
        ws.dh1 = c.take<float>(S * HID);
        ws.da2 = c.take<float>(S * A2_SZ);
        ws.dc2 = c.take<float>(S * C2_SZ);
        ws.da1 = c.take<float>(S * A1_SZ);
        if (own_G) {
            ws.ldg = align_up(P, 64);
            ws.G = c.take<float>((int64_t)n_groups * ws.ldg);
        }
    }
    return align_up(c.off, 256);
}

// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
prep_kernel(Pass p, float *prep, const int32_t *model_active, int limit)
{
    const int m = blockIdx.y;
    if (model_active && model_active[m] == 0) return;
    const float *row = p.X + (int64_t)p.mrow[m] * p.ldx;
    float *dst = prep + (int64_t)m * PREP_STRIDE;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < limit; i += gridDim.x * blockDim.x) {
        float v;
        if (i < PO_W2T) { const int co = i & 31, tap = i >> 5; v = row[OFF_W1 + co * 25 + tap]; }
        else if (i < PO_W2D) { const int j = i - PO_W2T, co = j & 63, ci = (j >> 6) & 31, tap = j >> 11;
                               v = row[OFF_W2 + co * 800 + ci * 25 + tap]; }
        else { const int j = i - PO_W2D, ci = j & 31, co = (j >> 5) & 63, tap = j >> 11;
               v = row[OFF_W2 + co * 800 + ci * 25 + (24 - tap)]; }
        dst[i] = v;
    }
}

// conv1 (1->32) + ReLU + maxpool 2/2, one CTA per sample slot; a1 is [196 pooled pixels][32]
__global__ void __launch_bounds__(256)
conv1_relu_pool_kernel(Pass p, Ws ws, int save_bwd)
{
    __shared__ float img[32 * 32];
    __shared__ float wsm[25 * 32];
    __shared__ float bsm[32];
    __shared__ __align__(16) __nv_bfloat16 spl[A1P_SZ];    // the sample's bf16 planes (tensor-core conv2)
    const int slot = blockIdx.x, t = threadIdx.x;
    const int id = p.img_idx[slot];
    if (id < 0) return;
    if (ws.tc)
        for (int i = t; i < A1P_SZ / 8; i += 256) reinterpret_cast<uint4 *>(spl)[i] = make_uint4(0u, 0u, 0u, 0u);
    int m;
    const float *row = model_row(p, slot / p.gs, m);
    const float *w1t = ws.prep + (int64_t)m * PREP_STRIDE + PO_W1T;
    for (int i = t; i < 1024; i += 256) img[i] = 0.f;
    for (int i = t; i < 800; i += 256) wsm[i] = w1t[i];
    if (t < 32) bsm[t] = row[OFF_B1 + t];
    __syncthreads();
    const uint8_t *src = p.images + (int64_t)id * IMG_BYTES;
    for (int i = t; i < IMG_BYTES; i += 256) {
        const int y = i / 28, x = i - y * 28;
        img[(y + 2) * 32 + x + 2] = __fdiv_rn((float)src[i], 255.f);
    }
    __syncthreads();
    for (int o = t; o < A1_SZ; o += 256) {
        const int co = o & 31, pp = o >> 5, py = pp / 14, px = pp - py * 14;
        float best = -INFINITY;
        int bi = 0;
#pragma unroll
        for (int d = 0; d < 4; ++d) {
            const int y = 2 * py + (d >> 1), x = 2 * px + (d & 1);
            float acc = bsm[co];
#pragma unroll
            for (int ky = 0; ky < 5; ++ky)
#pragma unroll
                for (int kx = 0; kx < 5; ++kx) acc = fmaf(img[(y + ky) * 32 + x + kx], wsm[(ky * 5 + kx) * 32 + co], acc);
            acc = acc > 0.f ? acc : 0.f;
            if (acc > best) { best = acc; bi = d; }
        }
        ws.a1[(int64_t)slot * A1_SZ + o] = best;
        if (save_bwd) ws.idx1[(int64_t)slot * A1_SZ + o] = (uint8_t)bi;
        if (ws.tc) spl[((co >> 3) * gnl::RUNP14 + py * 16 + px) * 8 + (co & 7)] = __float2bfloat16_rn(best);
    }
    if (ws.tc) {                                         // planes leave as coalesced 16-byte stores (gap pixels: zero)
        __syncthreads();
        uint4 *dst = reinterpret_cast<uint4 *>(ws.a1p + (int64_t)slot * A1P_SZ);
        for (int i = t; i < 4 * gnl::RUNP14; i += 256) dst[i] = reinterpret_cast<const uint4 *>(spl)[i];
    }
}

// ReLU + maxpool 2/2 of conv2's output; writes a2 in the reference's flatten order [c][7][7]
__global__ void __launch_bounds__(256)
relu_pool2_kernel(Pass p, Ws ws, int save_bwd)
{
    __shared__ __align__(16) __nv_bfloat16 sk[A2_SZ];      // the sample's features in flatten order, bf16 (tensor-core fc1)
    const int slot = blockIdx.x, t = threadIdx.x;
    const int g = slot / p.gs, sidx = slot - g * p.gs;
    uint4 *a2k = ws.tc ? reinterpret_cast<uint4 *>(ws.a2k) + ((int64_t)g * (FEAT / 8)) * ws.nsp + sidx : nullptr;
    if (p.img_idx[slot] < 0) {
        if (a2k && p.gcount[g] > 0)
            for (int e8 = t; e8 < FEAT / 8; e8 += 256) a2k[(int64_t)e8 * ws.nsp] = make_uint4(0u, 0u, 0u, 0u);
        return;
    }
    // results are staged in shared memory in flatten order [c][49] and leave as coalesced 16-byte stores (written
    // straight from the channel-fastest compute order they are 4-byte stores 196 bytes apart)
    __shared__ __align__(16) float sf[A2_SZ];
    __shared__ __align__(16) uint8_t si[A2_SZ];
    const float *__restrict__ src = ws.c2 + (int64_t)slot * C2_SZ;
    int m;
    const float *row = model_row(p, slot / p.gs, m);
    const float bias = ws.tc ? row[OFF_B2 + (t & 63)] : 0.f;   // the tensor-core convolution leaves the bias to its consumer
    constexpr int U = 4;
    for (int o0 = t; o0 < A2_SZ; o0 += U * 256) {            // c = o & 63 is the same for all of a thread's outputs
        float v[U][4];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int o = o0 + u * 256, pp = o >> 6, py = pp / 7, px = pp - py * 7;
#pragma unroll
            for (int d = 0; d < 4; ++d)
                v[u][d] = o < A2_SZ ? __ldg(src + ((2 * py + (d >> 1)) * 14 + 2 * px + (d & 1)) * 64 + (o & 63)) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int o = o0 + u * 256, c = o & 63, pp = o >> 6;
            if (o >= A2_SZ) break;
            float best = -INFINITY;
            int bi = 0;
#pragma unroll
            for (int d = 0; d < 4; ++d) {
                float x = v[u][d] + bias;
                x = x > 0.f ? x : 0.f;
                if (x > best) { best = x; bi = d; }
            }
            sf[c * 49 + pp] = best;
            si[c * 49 + pp] = (uint8_t)bi;
            if (a2k) sk[c * 49 + pp] = __float2bfloat16_rn(best);
        }
    }
    __syncthreads();
    float4 *a2o = reinterpret_cast<float4 *>(ws.a2 + (int64_t)slot * A2_SZ);
    for (int i = t; i < A2_SZ / 4; i += 256) a2o[i] = reinterpret_cast<const float4 *>(sf)[i];
    if (save_bwd) {
        uint4 *io = reinterpret_cast<uint4 *>(ws.idx2 + (int64_t)slot * A2_SZ);
        for (int i = t; i < A2_SZ / 16; i += 256) io[i] = reinterpret_cast<const uint4 *>(si)[i];
    }
    if (a2k)
        for (int e8 = t; e8 < FEAT / 8; e8 += 256) a2k[(int64_t)e8 * ws.nsp] = reinterpret_cast<const uint4 *>(sk)[e8];
}

// ------------------------------------------------------------------------------------------
// dense layers, one node (group) per blockIdx.x, weights W[N][K] row-major taken from the node's row
// ------------------------------------------------------------------------------------------
// out[s][n] = act(b[n] + sum_k in[s][k] * W[n][k]);  grid = (groups, ceil(N/32)); 8 warps x 4 neurons,
// lane = sample (and lane+32 when the batch is larger than 32)
template <int K, int N, bool RELU>
__global__ void __launch_bounds__(256)
dense_fwd_kernel(const float *__restrict__ in, float *__restrict__ out, int out_stride, Pass p, int w_off, int b_off)
{
    constexpr int KT = 64;
    __shared__ float in_s[KT][65];
    __shared__ __align__(16) float w_s[32][KT];
    const int g = blockIdx.x, n0 = blockIdx.y * 32, t = threadIdx.x, lane = t & 31, w = t >> 5;
    if (p.gcount[g] <= 0) return;
    int m;
    const float *row = model_row(p, g, m);
    const int gs = p.gs;
    float acc[4][2];
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[j][0] = acc[j][1] = 0.f;
    for (int k0 = 0; k0 < K; k0 += KT) {
        __syncthreads();
        for (int i = t; i < gs * KT; i += 256) {
            const int s = i / KT, k = i - s * KT;
            const int slot = g * gs + s;
            in_s[k][s] = (k0 + k < K && p.img_idx[slot] >= 0) ? in[(int64_t)slot * K + k0 + k] : 0.f;
        }
        for (int i = t; i < 32 * KT; i += 256) {
            const int n = i / KT, k = i - n * KT;
            w_s[n][k] = (n0 + n < N && k0 + k < K) ? row[w_off + (int64_t)(n0 + n) * K + k0 + k] : 0.f;
        }
        __syncthreads();
#pragma unroll 4
        for (int k = 0; k < KT; k += 4) {
            float a0[4], a1[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) { a0[q] = in_s[k + q][lane]; a1[q] = in_s[k + q][lane + 32]; }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 wv = *reinterpret_cast<const float4 *>(&w_s[4 * w + j][k]);
                acc[j][0] = fmaf(a0[0], wv.x, acc[j][0]); acc[j][0] = fmaf(a0[1], wv.y, acc[j][0]);
                acc[j][0] = fmaf(a0[2], wv.z, acc[j][0]); acc[j][0] = fmaf(a0[3], wv.w, acc[j][0]);
                acc[j][1] = fmaf(a1[0], wv.x, acc[j][1]); acc[j][1] = fmaf(a1[1], wv.y, acc[j][1]);
                acc[j][1] = fmaf(a1[2], wv.z, acc[j][1]); acc[j][1] = fmaf(a1[3], wv.w, acc[j][1]);
            }
        }
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int s = lane + 32 * h;
        if (s >= gs || p.img_idx[g * gs + s] < 0) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + 4 * w + j;
            if (n < N) {
                float v = acc[j][h] + row[b_off + n];
                if (RELU) v = v > 0.f ? v : 0.f;
                out[(int64_t)(g * gs + s) * out_stride + n] = v;
            }
        }
    }
}

// din[s][k] = sum_n dout[s][n] * W[n][k]  (optionally masked by act[s][k] > 0);
// grid = (groups, ceil(K/128)); thread = (k, half of the samples)
template <int K, int N, bool MASK>
__global__ void __launch_bounds__(256)
dense_dgrad_kernel(const float *__restrict__ dout, int dout_stride, float *__restrict__ din,
                   const float *__restrict__ act, Pass p, int w_off, uint16_t *dk = nullptr, int nsp = 0)
{
    constexpr int NT = 64;
    __shared__ __align__(16) float d_s[DLS_MAX_GROUP][NT];
    const int g = blockIdx.x, t = threadIdx.x, k = blockIdx.y * 128 + (t & 127), half = t >> 7;
    if (p.gcount[g] <= 0) return;
    int m;
    const float *row = model_row(p, g, m);
    const int gs = p.gs, s0 = half * 32;
    float acc[32];
#pragma unroll
    for (int s = 0; s < 32; ++s) acc[s] = 0.f;
    for (int nb = 0; nb < N; nb += NT) {
        __syncthreads();
        for (int i = t; i < gs * NT; i += 256) {
            const int s = i / NT, n = i - s * NT;
            const int slot = g * gs + s;
            d_s[s][n] = (nb + n < N && p.img_idx[slot] >= 0) ? dout[(int64_t)slot * dout_stride + nb + n] : 0.f;
        }
        __syncthreads();
        if (k < K && s0 < gs) {
#pragma unroll 2
            for (int n = 0; n < NT; n += 4) {
                float wv[4];
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    wv[q] = (nb + n + q < N) ? __ldg(row + w_off + (int64_t)(nb + n + q) * K + k) : 0.f;
#pragma unroll
                for (int s = 0; s < 32; ++s) {
                    const float4 d = *reinterpret_cast<const float4 *>(&d_s[s0 + s][n]);
                    acc[s] = fmaf(d.x, wv[0], acc[s]); acc[s] = fmaf(d.y, wv[1], acc[s]);
                    acc[s] = fmaf(d.z, wv[2], acc[s]); acc[s] = fmaf(d.w, wv[3], acc[s]);
                }
            }
        }
    }
    if (k < K) {
#pragma unroll
        for (int s = 0; s < 32; ++s) {
            const int sl = s0 + s;
            if (sl >= gs) break;
            const int slot = g * gs + sl;
            const bool live = p.img_idx[slot] >= 0;
            float v = acc[s];
            if (MASK) v = (live && act[(int64_t)slot * K + k] > 0.f) ? v : 0.f;
            if (live) din[(int64_t)slot * K + k] = v;
            if (dk)                                        // bf16 unit layout [node][k/8][nsp][8] (tc_fc.cu)
                reinterpret_cast<__nv_bfloat16 *>(dk)[(((int64_t)g * (K / 8) + (k >> 3)) * nsp + sl) * 8 + (k & 7)] =
                    __float2bfloat16_rn(live ? v : 0.f);
        }
    }
}

// dW[n][k] = sum_s dout[s][n] * in[s][k], db[n] = sum_s dout[s][n], straight into the gradient row;
// grid = (groups, ceil(N/32), ceil(K/128)); thread = (k, 16 neurons)
template <int K, int N>
__global__ void __launch_bounds__(256)
dense_wgrad_kernel(const float *__restrict__ dout, int dout_stride, const float *__restrict__ in, Pass p,
                   float *G, int64_t ldg, int w_off, int b_off)
{
    __shared__ __align__(16) float d_s[DLS_MAX_GROUP][32];
    __shared__ float in_s[DLS_MAX_GROUP][128];
    const int g = blockIdx.x, n0 = blockIdx.y * 32, k0 = blockIdx.z * 128, t = threadIdx.x;
    if (p.gcount[g] <= 0) return;
    const int gs = p.gs;
    for (int i = t; i < gs * 32; i += 256) {
        const int s = i >> 5, n = i & 31, slot = g * gs + s;
        d_s[s][n] = (n0 + n < N && p.img_idx[slot] >= 0) ? dout[(int64_t)slot * dout_stride + n0 + n] : 0.f;
    }
    for (int i = t; i < gs * 128; i += 256) {
        const int s = i >> 7, k = i & 127, slot = g * gs + s;
        in_s[s][k] = (k0 + k < K && p.img_idx[slot] >= 0) ? in[(int64_t)slot * K + k0 + k] : 0.f;
    }
    __syncthreads();
    const int k = t & 127, nh = (t >> 7) * 16;
    float acc[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) acc[j] = 0.f;
    for (int s = 0; s < gs; ++s) {
        const float a = in_s[s][k];
#pragma unroll
        for (int j = 0; j < 16; j += 4) {
            const float4 d = *reinterpret_cast<const float4 *>(&d_s[s][nh + j]);
            acc[j] = fmaf(d.x, a, acc[j]); acc[j + 1] = fmaf(d.y, a, acc[j + 1]);
            acc[j + 2] = fmaf(d.z, a, acc[j + 2]); acc[j + 3] = fmaf(d.w, a, acc[j + 3]);
        }
    }
    float *Gr = G + (int64_t)g * ldg;
    if (k0 + k < K) {
#pragma unroll
        for (int j = 0; j < 16; ++j)
            if (n0 + nh + j < N) Gr[w_off + (int64_t)(n0 + nh + j) * K + k0 + k] = acc[j];
    }
    if (blockIdx.z == 0 && t < 32 && n0 + t < N) {
        float b = 0.f;
        for (int s = 0; s < gs; ++s) b += d_s[s][t];
        Gr[b_off + n0 + t] = b;
    }
}

// cross entropy over the 62 logits of every sample of a group; mode 0: logits -> dlogits / count and the
// mean loss; mode 1: per-group loss sum and hit count
__global__ void __launch_bounds__(64)
ce_kernel(Pass p, Ws ws, int mode, float *loss_out, float *out_logits, int logits_stride_model)
{
    __shared__ float lossv[DLS_MAX_GROUP];
    __shared__ int hit[DLS_MAX_GROUP];
    const int g = blockIdx.x, s = threadIdx.x;
    const int cnt = p.gcount[g];
    if (cnt <= 0) return;
    const int gs = p.gs;
    lossv[s] = 0.f; hit[s] = 0;
    if (s < gs) {
        const int slot = g * gs + s;
        const int id = p.img_idx[slot];
        if (id >= 0) {
            float *lg = ws.lg + (int64_t)slot * 64;
            float mx = lg[0];
            int am = 0;
            for (int o = 1; o < NCLS; ++o) if (lg[o] > mx) { mx = lg[o]; am = o; }
            float se = 0.f;
            for (int o = 0; o < NCLS; ++o) se += expf(lg[o] - mx);
            const float lse = mx + logf(se);
            const int y = (int)p.labels[id];
            lossv[s] = lse - lg[y];
            hit[s] = (am == y);
            if (mode == 1 && out_logits) {
                const int m = p.gmodel ? p.gmodel[g] : g;
                for (int o = 0; o < NCLS; ++o) out_logits[(int64_t)m * logits_stride_model + (int64_t)id * NCLS + o] = lg[o];
            }
            if (mode == 0)
                for (int o = 0; o < NCLS; ++o) lg[o] = (expf(lg[o] - lse) - (o == y ? 1.f : 0.f)) / (float)cnt;
        }
    }
    __syncthreads();
    if (s == 0) {
        float ls = 0.f; int hs = 0;
        for (int k = 0; k < gs; ++k) { ls += lossv[k]; hs += hit[k]; }
        if (mode == 0) { if (loss_out) loss_out[g] = ls / (float)cnt; }
        else { ws.glog[(int64_t)g * 2] = ls; ws.glog[(int64_t)g * 2 + 1] = (float)hs; }
    }
}

// backward of ReLU + maxpool 2/2 (block 2): da2 is in flatten order [c][49]; dc2 is [196][64]
__global__ void __launch_bounds__(256)
pool2_bwd_kernel(Pass p, Ws ws)
{
    const int slot = blockIdx.x, t = threadIdx.x;
    if (p.img_idx[slot] < 0) return;
    if (ws.tc) {
        // bf16 planes: one 16-byte entry (8 channels of a pixel) per item, gap pixels zeroed.  The sample's masked
        // pooled gradients and arg-max bytes are fetched once, coalesced, into shared memory (every pooled value is
        // needed by the four pixels of its window, and the [c][49] layout is 196 bytes apart across channels)
        __shared__ float sg[A2_SZ];
        __shared__ uint8_t si[A2_SZ];
        for (int i = t; i < A2_SZ; i += 256) {
            const int64_t b = (int64_t)slot * A2_SZ + i;
            const float a = ws.a2[b], d = ws.da2[b];
            sg[i] = a > 0.f ? d : 0.f;
            si[i] = ws.idx2[b];
        }
        __syncthreads();
        uint4 *dst = reinterpret_cast<uint4 *>(ws.dc2p + (int64_t)slot * DC2P_SZ);
        for (int i = t; i < 196 * 8; i += 256) {
            const int ch = i / 196, pix = i - ch * 196, y = pix / 14, x = pix - y * 14;
            const int sel = (y & 1) * 2 + (x & 1);
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int o = (ch * 8 + e) * 49 + (y >> 1) * 7 + (x >> 1);
                v[e] = (int)si[o] == sel ? sg[o] : 0.f;
            }
            dst[ch * gnl::RUNP14 + y * 16 + x] = tc::pack8_bf16(v);
        }
        for (int i = t; i < 8 * 26; i += 256) {
            const int pl = i / 26, r = i - pl * 26;
            dst[pl * gnl::RUNP14 + (r >> 1) * 16 + 14 + (r & 1)] = make_uint4(0u, 0u, 0u, 0u);
        }
        return;
    }
    for (int e = t; e < C2_SZ; e += 256) {
        const int c = e & 63, pix = e >> 6, y = pix / 14, x = pix - y * 14;
        const int o = c * 49 + (y >> 1) * 7 + (x >> 1);
        const int64_t b = (int64_t)slot * A2_SZ + o;
        const bool hit = (int)ws.idx2[b] == ((y & 1) * 2 + (x & 1)) && ws.a2[b] > 0.f;
        ws.dc2[(int64_t)slot * C2_SZ + e] = hit ? ws.da2[b] : 0.f;
    }
}

// conv2 bias gradient (tensor-core path): every pooled position routes its gradient to exactly one conv2 pixel, so
// db2[c] = sum over samples and pooled positions of da2 where the pooled activation is positive.  grid = (nodes, 8):
// warp = channel, lane = pooled position (two passes over the 49), samples in slot order; fixed summation order.
__global__ void __launch_bounds__(256)
bias2_grad_kernel(Pass p, Ws ws)
{
    const int g = blockIdx.x, c = blockIdx.y * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (p.gcount[g] <= 0) return;
    float acc = 0.f;
    for (int s = 0; s < p.gs; ++s) {
        const int slot = g * p.gs + s;
        if (p.img_idx[slot] < 0) continue;
        const int64_t b = (int64_t)slot * A2_SZ + c * 49;
        float v = ws.a2[b + lane] > 0.f ? ws.da2[b + lane] : 0.f;
        if (lane + 32 < 49) v += ws.a2[b + lane + 32] > 0.f ? ws.da2[b + lane + 32] : 0.f;
        acc += v;
    }
    acc = warp_sum(acc);
    if (lane == 0) ws.G[(int64_t)g * ws.ldg + OFF_B2 + c] = acc;
}

// conv1 weight/bias gradient: the gradient reaches conv1 only at the arg-max of each 2x2 window and only where the
// pooled activation is positive.  One CTA per node; lane = output channel, the 8 warps share the 196 pooled pixels
// of a sample.  For a pooled pixel the 32 channels read the image at only four places (their arg-max positions
// (dy, dx)), and with a row pitch of 34 floats those four addresses — offsets 0, 1, 34, 35 — lie in four different
// banks: every one of the 25 tap reads is a single conflict-free (multicast) wavefront for all 32 channels, and the
// gradients / arg-max bytes are plain coalesced 128-byte loads from [pixel][channel], requested one whole sample
// ahead (a sample's accumulation takes about one DRAM latency; five-pixel batches left the kernel latency bound at
// 480 us per 128-node launch, pixel-per-lane forms were bound by bank conflicts and 32-sector loads at 435 us; this
// form: 210 us).  Per-warp partial sums are combined once, in a fixed order.
constexpr int C1W_WARPS = 8;
__global__ void __launch_bounds__(32 * C1W_WARPS)
conv1_wgrad_kernel(Pass p, Ws ws)
{
    constexpr int IP = 34;
    __shared__ float img[2][32 * IP];
    __shared__ float part[C1W_WARPS][26][32];
    const int g = blockIdx.x, t = threadIdx.x, lane = t & 31, w = t >> 5;
    if (p.gcount[g] <= 0) return;
    for (int i = t; i < 2 * 32 * IP; i += 32 * C1W_WARPS) (&img[0][0])[i] = 0.f;
    float acc[25], bsum = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) acc[k] = 0.f;
    auto load_img = [&](int s) -> uint32_t {
        const int id = s < p.gs ? p.img_idx[g * p.gs + s] : -1;
        return (id >= 0 && t < 196) ? reinterpret_cast<const uint32_t *>(p.images + (int64_t)id * IMG_BYTES)[t] : 0u;
    };
    // the warp's pooled pixels are pp = w + 8 k, k = 0..24; a whole sample's gradients and arg-max bytes are requested
    // one SAMPLE ahead (both loads unconditional): the accumulation of a sample takes about as long as one DRAM latency
    constexpr int NK = (196 + C1W_WARPS - 1) / C1W_WARPS;  // 25
    float dn[NK];
    uint32_t in[(NK + 15) / 16];                           // 2 bits per pixel
    auto request = [&](int s) {
        const int slot = g * p.gs + s;
        const bool live = s < p.gs && p.img_idx[slot] >= 0;
        const float *a1s = ws.a1 + (int64_t)slot * A1_SZ + lane, *da1s = ws.da1 + (int64_t)slot * A1_SZ + lane;
        const uint8_t *ixs = ws.idx1 + (int64_t)slot * A1_SZ + lane;
#pragma unroll
        for (int q = 0; q < (NK + 15) / 16; ++q) in[q] = 0u;
#pragma unroll
        for (int k = 0; k < NK; ++k) {
            const int pp = w + C1W_WARPS * k;
            dn[k] = 0.f;
            if (live && pp < 196) {
                const float a = a1s[pp * 32], dd = da1s[pp * 32];
                const uint32_t ix = ixs[pp * 32];
                dn[k] = a > 0.f ? dd : 0.f;
                in[k >> 4] |= (ix & 3u) << (2 * (k & 15));
            }
        }
    };
    uint32_t px4 = load_img(0);
    request(0);
    __syncthreads();
    for (int s = 0; s < p.gs; ++s) {
        float *im = img[s & 1];
        if (t < 196) {
            const int y = (4 * t) / 28, x = 4 * t - y * 28;
#pragma unroll
            for (int e = 0; e < 4; ++e) im[(y + 2) * IP + x + 2 + e] = __fdiv_rn((float)((px4 >> (8 * e)) & 0xffu), 255.f);
        }
        float dc[NK];
        uint32_t ic[(NK + 15) / 16];
#pragma unroll
        for (int k = 0; k < NK; ++k) dc[k] = dn[k];
#pragma unroll
        for (int q = 0; q < (NK + 15) / 16; ++q) ic[q] = in[q];
        __syncthreads();                                   // image s visible; everyone is done with image s - 1
        px4 = load_img(s + 1);
        request(s + 1);
#pragma unroll
        for (int k = 0; k < NK; ++k) {
            const int pp = w + C1W_WARPS * k;
            if (pp >= 196) break;                          // warp-uniform
            const int py = pp / 14, pxx = pp - py * 14;
            const uint32_t ii = (ic[k >> 4] >> (2 * (k & 15))) & 3u;
            const float *src = im + (2 * py + (int)(ii >> 1)) * IP + 2 * pxx + (int)(ii & 1u);
            const float d = dc[k];
#pragma unroll
            for (int ky = 0; ky < 5; ++ky)
#pragma unroll
                for (int kx = 0; kx < 5; ++kx) acc[ky * 5 + kx] = fmaf(d, src[ky * IP + kx], acc[ky * 5 + kx]);
            bsum += d;
        }
    }
#pragma unroll
    for (int k = 0; k < 25; ++k) part[w][k][lane] = acc[k];
    part[w][25][lane] = bsum;
    __syncthreads();
    float *G = ws.G + (int64_t)g * ws.ldg;
    for (int i = t; i < 26 * 32; i += 32 * C1W_WARPS) {
        const int k = i >> 5, co = i & 31;
        float v = 0.f;
#pragma unroll
        for (int ww = 0; ww < C1W_WARPS; ++ww) v += part[ww][k][co];
        if (k < 25) G[OFF_W1 + co * 25 + k] = v;
        else G[OFF_B1 + co] = v;
    }
}

// ------------------------------------------------------------------------------------------
static int set_attrs()
{
    static bool done[MAX_DEVICES] = {};                  // per device, like gnlenet.cu
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= MAX_DEVICES) return DLS_E_TOO_BIG;
    if (done[dev]) return 0;
#define SET_SMEM(k, bytes) \
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)); if (e != cudaSuccess) return (int)e;
    SET_SMEM((gnl::conv5_kernel<32, 64, 14, true>), 18 * 18 * 32 * 4);
    SET_SMEM((gnl::conv5_kernel<64, 32, 14, false>), 18 * 18 * 64 * 4);
    SET_SMEM((gnl::conv_wgrad_kernel<64, 14>), (18 * 18 * 32 + 196 * 16) * 4);
#undef SET_SMEM
    done[dev] = true;
    return 0;
}

static int forward(const Pass &p, Ws &ws, bool save, cudaStream_t st)
{
    int rc = set_attrs();
    if (rc) return rc;
    const int S = p.n_groups * p.gs, sv = save ? 1 : 0;
    // tensor-core path: only conv1's transposed weights are needed in fp32 (conv2's operands are the bf16 copies)
    prep_kernel<<<dim3(ws.tc ? 1 : 8, p.n_models), 256, 0, st>>>(p, ws.prep, p.gmodel ? nullptr : p.gcount,
                                                              ws.tc ? PO_W2T : PREP_STRIDE);
    DLS_LAUNCH_CHECK();
    if (ws.tc && (rc = gnl::fem_prep_conv2_tc(p, ws.prep_tc, OFF_W2, save, st))) return rc;
    if (ws.tc && ws.nsp != p.gs) {                        // sample rows gs..nsp-1 of the fc1 operands: zero
        cudaError_t e = cudaMemsetAsync(ws.a2k, 0, (int64_t)p.n_groups * (FEAT / 8) * ws.nsp * 16, st);
        if (e == cudaSuccess && ws.dh1k) e = cudaMemsetAsync(ws.dh1k, 0, (int64_t)p.n_groups * (HID / 8) * ws.nsp * 16, st);
        if (e != cudaSuccess) return (int)e;
    }
    conv1_relu_pool_kernel<<<S, 256, 0, st>>>(p, ws, sv);
    DLS_LAUNCH_CHECK();
    if (ws.tc) {
        if ((rc = gnl::fem_conv2_fwd_tc(p, ws.a1p, ws.c2, ws.prep_tc, st))) return rc;
    } else {
        gnl::conv5_kernel<32, 64, 14, true><<<S, 256, 18 * 18 * 32 * 4, st>>>(ws.a1, ws.c2, p, ws.prep, PREP_STRIDE, PO_W2T, OFF_B2);
        DLS_LAUNCH_CHECK();
    }
    relu_pool2_kernel<<<S, 256, 0, st>>>(p, ws, sv);
    DLS_LAUNCH_CHECK();
    if (ws.tc) {
        if ((rc = fc1_fwd_tc(p, ws.a2k, ws.h1, OFF_FC1W, OFF_FC1B, st))) return rc;
    } else {
        dense_fwd_kernel<FEAT, HID, true><<<dim3(p.n_groups, HID / 32), 256, 0, st>>>(ws.a2, ws.h1, HID, p, OFF_FC1W, OFF_FC1B);
        DLS_LAUNCH_CHECK();
    }
    dense_fwd_kernel<HID, NCLS, false><<<dim3(p.n_groups, 2), 256, 0, st>>>(ws.h1, ws.lg, 64, p, OFF_FC2W, OFF_FC2B);
    DLS_LAUNCH_CHECK();
    return 0;
}

static int backward(const Pass &p, Ws &ws, cudaStream_t st)
{
    const int S = p.n_groups * p.gs;
    dense_wgrad_kernel<HID, NCLS><<<dim3(p.n_groups, 2, HID / 128), 256, 0, st>>>(ws.lg, 64, ws.h1, p, ws.G, ws.ldg, OFF_FC2W, OFF_FC2B);
    DLS_LAUNCH_CHECK();
    dense_dgrad_kernel<HID, NCLS, true><<<dim3(p.n_groups, HID / 128), 256, 0, st>>>(ws.lg, 64, ws.dh1, ws.h1, p, OFF_FC2W,
                                                                                     ws.tc ? ws.dh1k : nullptr, ws.nsp);
    DLS_LAUNCH_CHECK();
    if (ws.tc) {
        // the data gradient reads the weights the fused optimizer step of the weight gradient overwrites: it goes first
        int rc;
        if ((rc = fc1_dgrad_tc(p, ws.dh1k, ws.da2, OFF_FC1W, st))) return rc;
        if ((rc = fc1_wgrad_tc(p, ws.a2k, ws.dh1k, ws.dh1, ws.G, ws.ldg, ws.sgd, OFF_FC1W, OFF_FC1B, st))) return rc;
    } else {
        dense_wgrad_kernel<FEAT, HID><<<dim3(p.n_groups, HID / 32, (FEAT + 127) / 128), 256, 0, st>>>(ws.dh1, HID, ws.a2, p, ws.G, ws.ldg, OFF_FC1W, OFF_FC1B);
        DLS_LAUNCH_CHECK();
        dense_dgrad_kernel<FEAT, HID, false><<<dim3(p.n_groups, (FEAT + 127) / 128), 256, 0, st>>>(ws.dh1, HID, ws.da2, nullptr, p, OFF_FC1W);
        DLS_LAUNCH_CHECK();
    }
    pool2_bwd_kernel<<<S, 256, 0, st>>>(p, ws);
    DLS_LAUNCH_CHECK();
    if (ws.tc) {
        int rc;
        cudaError_t e = cudaMemsetAsync(ws.zplane, 0, gnl::RUNP15 * 16, st);
        if (e != cudaSuccess) return (int)e;
        bias2_grad_kernel<<<dim3(p.n_groups, 8), 256, 0, st>>>(p, ws);
        DLS_LAUNCH_CHECK();
        if ((rc = gnl::fem_conv2_wgrad_tc(p, ws.a1p, ws.dc2p, ws.zplane, ws.G, ws.ldg, OFF_W2, st))) return rc;
        if ((rc = gnl::fem_conv2_dgrad_tc(p, ws.dc2p, ws.da1, ws.prep_tc, st))) return rc;
    } else {
        gnl::Workspace gw{};                   // the shared conv kernels only use the gradient-row fields
        gw.G = ws.G; gw.ldg = ws.ldg;
        gnl::conv_wgrad_kernel<64, 14><<<dim3(p.n_groups, 4), 256, (18 * 18 * 32 + 196 * 16) * 4, st>>>(ws.a1, ws.dc2, p, gw, OFF_W2, OFF_B2);
        DLS_LAUNCH_CHECK();
        gnl::conv5_kernel<64, 32, 14, false><<<S, 256, 18 * 18 * 64 * 4, st>>>(ws.dc2, ws.da1, p, ws.prep, PREP_STRIDE, PO_W2D, 0);
        DLS_LAUNCH_CHECK();
    }
    conv1_wgrad_kernel<<<p.n_groups, 32 * C1W_WARPS, 0, st>>>(p, ws);
    DLS_LAUNCH_CHECK();
    return 0;
}

int64_t train_workspace_bytes(int n_tasks, int batch, bool own_G)
{
    Ws ws;
    return carve(ws, nullptr, n_tasks, batch, n_tasks, true, own_G);
}

int step(const Pass &p, void *workspace, float *G, int64_t ldg, float *loss_out, float **G_used, int64_t *ldg_used,
         int precision, const SgdFuse *sgd, cudaStream_t st)
{
    Ws ws;
    carve(ws, workspace, p.n_groups, p.gs, p.n_models, true, G == nullptr);
    ws.tc = precision == 1;
    if (sgd && ws.tc) ws.sgd = *sgd;
    if (G) { ws.G = G; ws.ldg = ldg; }
    int rc;
    if ((rc = forward(p, ws, true, st))) return rc;
    ce_kernel<<<p.n_groups, 64, 0, st>>>(p, ws, 0, loss_out, nullptr, 0);
    DLS_LAUNCH_CHECK();
    if ((rc = backward(p, ws, st))) return rc;
    if (G_used) *G_used = ws.G;
    if (ldg_used) *ldg_used = ws.ldg;
    return 0;
}

constexpr int EVAL_GS = 32, EVAL_CHUNK_GROUPS = 512;

__global__ void eval_desc_kernel(int32_t *gmodel, int32_t *img_idx, int32_t *gcount, int g0, int n_groups, int gs,
                                 int gpm, int n_test)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_groups * gs) return;
    const int gl = i / gs, j = i - gl * gs, g = g0 + gl, m = g / gpm, chunk = g - m * gpm;
    const int sample = chunk * gs + j;
    img_idx[i] = sample < n_test ? sample : -1;
    if (j == 0) {
        gmodel[gl] = m;
        const int rem = n_test - chunk * gs;
        gcount[gl] = rem < gs ? (rem > 0 ? rem : 0) : gs;
    }
}

__global__ void eval_reduce_kernel(const float *glog_all, int gpm, int n_models, int32_t *out_correct, double *out_loss)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_models) return;
    double ls = 0.0; int hs = 0;
    for (int k = 0; k < gpm; ++k) { ls += (double)glog_all[((int64_t)m * gpm + k) * 2]; hs += (int)glog_all[((int64_t)m * gpm + k) * 2 + 1]; }
    out_loss[m] = ls; out_correct[m] = hs;
}

int64_t eval_workspace_bytes(int n_models, int n_test)
{
    const int gpm = (n_test + EVAL_GS - 1) / EVAL_GS;
    const int64_t total = (int64_t)n_models * gpm;
    const int chunk = (int)(total < EVAL_CHUNK_GROUPS ? total : EVAL_CHUNK_GROUPS);
    Ws ws;
    return carve(ws, nullptr, chunk > 0 ? chunk : 1, EVAL_GS, n_models, false, false) + total * 8 + 1024;
}

int eval(const float *X, int64_t ldx, const int32_t *model_rows, int n_models, const uint8_t *test_images,
         const int64_t *test_labels, int n_test, int32_t *out_correct, double *out_loss_sum, float *out_logits,
         int precision, void *workspace, cudaStream_t st)
{
    const int gpm = (n_test + EVAL_GS - 1) / EVAL_GS;
    const int64_t total = (int64_t)n_models * gpm;
    const int chunk = (int)(total < EVAL_CHUNK_GROUPS ? total : EVAL_CHUNK_GROUPS);
    Ws ws;
    const int64_t used = carve(ws, workspace, chunk, EVAL_GS, n_models, false, false);
    ws.tc = precision == 1;
    float *glog_all = reinterpret_cast<float *>(reinterpret_cast<char *>(workspace) + used);
    int32_t *d_img_idx = ws.desc_i32, *d_gmodel = ws.desc_i32 + (int64_t)chunk * EVAL_GS, *d_gcount = d_gmodel + chunk;
    Pass p;
    p.gs = EVAL_GS; p.n_models = n_models; p.gmodel = d_gmodel; p.mrow = model_rows; p.img_idx = d_img_idx;
    p.gcount = d_gcount; p.images = test_images; p.labels = test_labels; p.X = X; p.ldx = ldx; p.n_groups = chunk;
    int rc;
    for (int64_t g0 = 0; g0 < total; g0 += chunk) {
        const int ng = (int)((total - g0) < chunk ? (total - g0) : chunk);
        p.n_groups = ng;
        eval_desc_kernel<<<(ng * EVAL_GS + 255) / 256, 256, 0, st>>>(d_gmodel, d_img_idx, d_gcount, (int)g0, ng, EVAL_GS, gpm, n_test);
        DLS_LAUNCH_CHECK();
        ws.glog = glog_all + g0 * 2;
        if ((rc = forward(p, ws, false, st))) return rc;
        ce_kernel<<<ng, 64, 0, st>>>(p, ws, 1, nullptr, out_logits, n_test * NCLS);
        DLS_LAUNCH_CHECK();
    }
    eval_reduce_kernel<<<(n_models + 127) / 128, 128, 0, st>>>(glog_all, gpm, n_models, out_correct, out_loss_sum);
    DLS_LAUNCH_CHECK();
    return 0;
}

}  // namespace fem
}  // namespace dls
