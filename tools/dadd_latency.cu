// Micro-benchmark: latency of a dependent DADD chain and of the f32->f64 widening on sm_100a.
#include <cstdio>
#include <cstdint>
__global__ void k(const float* in, double* out, long long* cyc, int n) {
    double s = in[0];
    double q[8];
    for (int j = 0; j < 8; j++) q[j] = in[j + 1];
    long long t0 = clock64();
    for (int i = 0; i < n; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) s += q[j];          // 8 dependent DADDs
    }
    long long t1 = clock64();
    float f = in[threadIdx.x % 8];
    double acc = 0;
    for (int i = 0; i < n; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) { f = f * 1.0001f; acc += (double)(f * f); }   // FMUL + F2F + dependent DADD
    }
    long long t2 = clock64();
    out[threadIdx.x] = s + acc;
    if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; }
}
int main() {
    float h[16]; for (int i = 0; i < 16; i++) h[i] = 1.0f + i * 0.001f;
    float* d; double* o; long long* c; cudaMalloc(&d, 64); cudaMalloc(&o, 8 * 1024); cudaMalloc(&c, 16);
    cudaMemcpy(d, h, 64, cudaMemcpyHostToDevice);
    for (int threads : {32, 128, 512}) {
        k<<<1, threads>>>(d, o, c, 2000);
        long long r[2]; cudaMemcpy(r, c, 16, cudaMemcpyDeviceToHost);
        printf("threads %4d: dependent DADD %.1f cycles each; FMUL+F2F+DADD step %.1f cycles each (err=%s)\n", threads,
               r[0] / 16000.0, r[1] / 16000.0, cudaGetErrorString(cudaGetLastError()));
    }
}
