// How fast can 128-thread CTAs stream a [rows][3136] fp32 matrix when every CTA owns 128 rows and reads them in
// column chunks — i.e. in runs of RUN contiguous bytes per row, rows 12544 B apart (the access pattern of the FEMNIST
// fc1 kernels)?  Prints GB/s per run length.   nvcc -arch=sm_100a -O3 -o dram_runs dram_runs.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void ld8(const float *p, float (&r)[8]) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7]) : "l"(p));
}

// lanes_per_row * 32 B = run length; a CTA pass = 128 threads x 8 loads = 32 KB = (1024 / lanes_per_row) row pieces
template <int LPR>
__global__ void __launch_bounds__(128) stream_kernel(const float *X, float *out, int row_floats, int rows_per_cta)
{
    const int t = threadIdx.x;
    const float *base = X + (int64_t)blockIdx.x * rows_per_cta * row_floats;
    float acc = 0.f;
    const int run_floats = LPR * 8;
    const int chunks = row_floats / run_floats;             // column chunks of the row
    for (int c = 0; c < chunks; ++c) {
        // all rows_per_cta rows x one run: items = rows_per_cta * LPR, 8 per thread per pass
        for (int i0 = 0; i0 < rows_per_cta * LPR; i0 += 128 * 8) {
            float v[8][8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int i = i0 + j * 128 + t;
                ld8(base + (int64_t)(i / LPR) * row_floats + c * run_floats + (i % LPR) * 8, v[j]);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j)
#pragma unroll
                for (int e = 0; e < 8; ++e) acc += v[j][e];
        }
    }
    if (acc == 123.456f) out[0] = acc;
}

template <int LPR> void run(const float *X, float *out, int rows, int row_floats, int ctas_rows)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = rows / ctas_rows;
    stream_kernel<LPR><<<grid, 128>>>(X, out, row_floats, ctas_rows);
    cudaEventRecord(e0);
    for (int it = 0; it < 3; ++it) stream_kernel<LPR><<<grid, 128>>>(X, out, row_floats, ctas_rows);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const int used = (row_floats / (LPR * 8)) * LPR * 8;
    printf("run %5d B: %.1f us, %.0f GB/s  (%s)\n", LPR * 32, ms / 3 * 1e3, (double)rows * used * 4 / (ms / 3 * 1e-3) / 1e9,
           cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    const int row_floats = 3136, rows = 128 * 512;          // 822 MB
    float *X, *out;
    cudaMalloc(&X, (size_t)rows * row_floats * 4);
    cudaMalloc(&out, 16);
    cudaMemset(X, 0, (size_t)rows * row_floats * 4);
    run<8>(X, out, rows, row_floats, 128);
    run<16>(X, out, rows, row_floats, 128);
    run<32>(X, out, rows, row_floats, 128);
    run<64>(X, out, rows, row_floats, 128);
    run<392>(X, out, rows, row_floats, 128);
    // the same with 64-row CTAs (twice as many CTAs)
    run<8>(X, out, rows, row_floats, 64);
    run<32>(X, out, rows, row_floats, 64);
    return 0;
}
