// dls_tc_probe — runs a caller-described sequence of tcgen05.mma instructions on caller-provided
// shared-memory operand images and returns the accumulator tile from TMEM.  It pins down the
// descriptor encodings (tests/test_gpu_tc_probe.py) that the tensor-core convolution kernels use.
#include "common.cuh"
#include "tc_common.cuh"

namespace dls {

__global__ void __launch_bounds__(128)
tc_probe_kernel(const uint8_t *__restrict__ a_img, int a_bytes, const uint8_t *__restrict__ b_img, int b_bytes,
                const uint64_t *__restrict__ a_desc, const uint64_t *__restrict__ b_desc, int n_mma,
                uint32_t idesc, int ncols, float *__restrict__ out, int repeat, long long *__restrict__ cycles_out)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base;
    uint8_t *sa = smem;
    uint8_t *sb = smem + ((a_bytes + 1023) & ~1023);
    const int t = threadIdx.x, warp = t >> 5;
    for (int i = t; i < a_bytes; i += 128) sa[i] = a_img[i];
    for (int i = t; i < b_bytes; i += 128) sb[i] = b_img[i];
    if (t == 0) { tc::mbar_init(&bar, 1); tc::mbar_fence_init(); }
    if (warp == 0) tc::tmem_alloc(&tmem_base, 512);
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tm = tmem_base;
    long long t0 = 0;
    if (t == 0) {
        const uint32_t abase = tc::smem_u32(sa) >> 4, bbase = tc::smem_u32(sb) >> 4;
        uint64_t ad[8], bd[8];
        for (int i = 0; i < 8; ++i) { ad[i] = a_desc[i % n_mma] + abase; bd[i] = b_desc[i % n_mma] + bbase; }
        for (int i = 0; i < n_mma; ++i)
            tc::mma_f16(tm, a_desc[i] + abase, b_desc[i] + bbase, idesc, i > 0 ? 1u : 0u);
        t0 = clock64();
        // throughput probe: re-issue the first (up to 8) MMAs `repeat` times into spare columns
        for (int r = 0; r < repeat; ++r) {
#pragma unroll
            for (int i = 0; i < 8; ++i) tc::mma_f16(tm + 256, ad[i], bd[i], idesc, 1u);
        }
        tc::mma_commit(&bar);
    }
    tc::mbar_wait(&bar, 0);
    tc::fence_after_sync();
    if (t == 0 && cycles_out) *cycles_out = clock64() - t0;
    uint32_t r[32];
    for (int c0 = 0; c0 < ncols; c0 += 32) {
        tc::tmem_ld32(tm + ((uint32_t)(warp * 32) << 16) + c0, r);
        tc::tmem_ld_wait();
        for (int j = 0; j < 32 && c0 + j < ncols; ++j) out[(size_t)t * ncols + c0 + j] = __uint_as_float(r[j]);
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tm, 512);
}

// Contention probe: warp 4 issues `n_mma` MMAs (N = 96 columns at TMEM column 256) while warps 0-3 run `ld_loops`
// rounds of three 32-column TMEM loads (and `shfl_per_round` warp shuffles) each.  cycles[0] = MMA stream,
// cycles[1] = warp 0's load loop.
__global__ void __launch_bounds__(160)
tc_contention_kernel(int n_mma, int ld_loops, int shfl_per_round, long long *__restrict__ cycles, float *__restrict__ sink)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int t = threadIdx.x, warp = t >> 5;
    for (int i = t; i < 32768; i += 160) reinterpret_cast<uint32_t *>(smem)[i] = 0u;
    if (t == 0) { tc::mbar_init(&bar, 1); tc::mbar_fence_init(); }
    if (warp == 0) tc::tmem_alloc(&tmem_base, 512);
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tm = tmem_base;
    if (warp == 4) {
        if (t == 128) {
            const uint32_t base = tc::smem_u32(smem);
            const uint64_t ad = tc::smem_desc(base + 16, 4096, 128), bd = tc::smem_desc(base + 65536, 1536, 128);
            const uint32_t idesc = tc::instr_desc_bf16(128, 96, 0, 0);
            const long long t0 = clock64();
            for (int i = 0; i < n_mma; ++i) tc::mma_f16(tm + 256, ad + (uint64_t)((i & 7) * 2), bd, idesc, 1u);
            tc::mma_commit(&bar);
            tc::mbar_wait(&bar, 0);
            cycles[0] = clock64() - t0;
        }
    } else {
        const long long t0 = clock64();
        float acc = 0.f;
        for (int r = 0; r < ld_loops; ++r) {
            uint32_t a[32], b[32], c[32];
            const uint32_t ta = tm + ((uint32_t)(warp * 32) << 16);
            if (shfl_per_round == -1) {            // shape probe: 16-column loads, six per round (same bytes as 3 x 32)
                uint32_t q0[16], q1[16], q2[16], q3[16], q4[16], q5[16];
                tc::tmem_ld16(ta, q0); tc::tmem_ld16(ta + 16, q1); tc::tmem_ld16(ta + 32, q2);
                tc::tmem_ld16(ta + 48, q3); tc::tmem_ld16(ta + 64, q4); tc::tmem_ld16(ta + 80, q5);
                tc::tmem_ld_wait();
                acc += __uint_as_float(q0[r & 15]) + __uint_as_float(q1[r & 15]) + __uint_as_float(q2[r & 15]) +
                       __uint_as_float(q3[r & 15]) + __uint_as_float(q4[r & 15]) + __uint_as_float(q5[r & 15]);
                continue;
            }
            if (shfl_per_round == -2) {            // one warp only does the loads (others idle): per-warp bandwidth
                if (warp != 0) break;
            }
            tc::tmem_ld32(ta, a); tc::tmem_ld32(ta + 32, b); tc::tmem_ld32(ta + 64, c);
            tc::tmem_ld_wait();
            float v = __uint_as_float(a[r & 31]) + __uint_as_float(b[r & 31]) + __uint_as_float(c[r & 31]);
            for (int j = 0; j < shfl_per_round; ++j) v += __shfl_down_sync(0xffffffffu, v, 1);
            acc += v;
        }
        if (t == 0) cycles[1] = clock64() - t0;
        if (acc == 12345.f) sink[t] = acc;
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tm, 512);
}

}  // namespace dls

using namespace dls;

extern "C" int dls_tc_contention(int32_t n_mma, int32_t ld_loops, int32_t shfl_per_round, long long *cycles, float *sink,
                                 void *stream)
{
    if (!cycles || !sink || n_mma < 0 || ld_loops < 0) return DLS_E_ARG;
    const int smem = 132 * 1024;
    cudaError_t e = cudaFuncSetAttribute(tc_contention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    tc_contention_kernel<<<1, 160, smem, (cudaStream_t)stream>>>(n_mma, ld_loops, shfl_per_round, cycles, sink);
    DLS_LAUNCH_CHECK();
    return 0;
}

// a_desc / b_desc: descriptors with start-address field RELATIVE to the operand image (the kernel
// adds the image's shared-memory address); out: [128 TMEM lanes][ncols] floats, ncols <= 256.
extern "C" int dls_tc_probe(const uint8_t *a_img, int32_t a_bytes, const uint8_t *b_img, int32_t b_bytes,
                            const uint64_t *a_desc, const uint64_t *b_desc, int32_t n_mma, uint32_t idesc,
                            int32_t ncols, float *out, int32_t repeat, long long *cycles_out, void *stream)
{
    if (!a_img || !b_img || !a_desc || !b_desc || !out || n_mma <= 0 || ncols <= 0 || ncols > 256) return DLS_E_ARG;
    const size_t smem = ((size_t)(a_bytes + 1023) & ~(size_t)1023) + (size_t)b_bytes + 1024;
    if (smem > 200 * 1024) return DLS_E_TOO_BIG;
    cudaError_t e = cudaFuncSetAttribute(tc_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    tc_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(a_img, a_bytes, b_img, b_bytes, a_desc, b_desc, n_mma,
                                                            idesc, ncols, out, repeat, cycles_out);
    DLS_LAUNCH_CHECK();
    return 0;
}
