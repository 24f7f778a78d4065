// What can HBM3e sustain for a 1 : 2 read : write stream (int16 in, f32 out — the shape of the
// convert+features kernel) compared with a 1 : 1 copy?  Grid-stride, 16-byte loads, streaming stores.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/hbm_ratio_probe.bin tools/hbm_ratio_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int W>   // W 16-byte stores per 16-byte load
__global__ void stream(const uint4 *__restrict__ in, uint4 *__restrict__ out, size_t n)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint4 v = in[i];
#pragma unroll
        for (int k = 0; k < W; k++) {
            v.x += k;
            __stcs(out + i * W + k, v);
        }
    }
}

template <int W>
static void run(const char *name, const uint4 *in, uint4 *out, size_t n)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 3; i++) stream<W><<<148 * 16, 256>>>(in, out, n);
    cudaEventRecord(a);
    const int reps = 20;
    for (int i = 0; i < reps; i++) stream<W><<<148 * 16, 256>>>(in, out, n);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    const double bytes = (double)n * 16 * (1 + W) * reps;
    printf("%-28s %8.1f GB/s  (%.3f ms per pass, %.0f MB read + %.0f MB written)\n", name, bytes / (ms * 1e-3) / 1e9,
           ms / reps, n * 16 / 1e6, n * 16.0 * W / 1e6);
}

int main()
{
    const size_t n = 655360000 / 16;          // the bench's 655 MB of PCM
    uint4 *in, *out;
    cudaMalloc(&in, n * 16);
    cudaMalloc(&out, n * 16 * 2);
    cudaMemset(in, 1, n * 16);
    run<1>("1 : 1 copy", in, out, n);
    run<2>("1 : 2 (int16 in, f32 out)", in, out, n);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
