// Tuning aid: what a pure streaming READ reaches on this GPU for the launch shapes wt_select_kernel can use
// (persistent grid, float4 loads, U of them in flight per thread).  nvcc -arch=sm_100a -O3 -o tools/bin/stream_probe
#include <cstdio>
#include <cuda_runtime.h>

template <int U>
__global__ void __launch_bounds__(1024) read_kernel(const float4* __restrict__ x, long long nvec, float* out) {
    float acc = 0.f;
    const long long per = (nvec + gridDim.x - 1) / gridDim.x;
    const long long b = per * blockIdx.x, e = min(nvec, b + per);
    for (long long v = b + threadIdx.x; v < e; v += (long long)blockDim.x * U) {
        float4 r[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v + (long long)u * blockDim.x < e) r[u] = __ldg(&x[v + (long long)u * blockDim.x]);
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v + (long long)u * blockDim.x < e) acc += r[u].x + r[u].y + r[u].z + r[u].w;
    }
    if (acc == 123.456f) *out = acc;
}

template <int U>
void run(const float4* x, long long nvec, float* out, int grid, int block, const char* label) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) read_kernel<U><<<grid, block>>>(x, nvec, out);
    cudaEventRecord(e0);
    const int reps = 20;
    for (int i = 0; i < reps; ++i) read_kernel<U><<<grid, block>>>(x, nvec, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%-28s U=%2d grid %4d x %4d: %7.1f us  %6.2f TB/s\n", label, U, grid, block, ms / reps * 1e3,
           (double)nvec * 16 / (ms / reps * 1e-3) / 1e12);
}

int main() {
    for (long long mb : {45LL, 124LL, 604LL}) {
        const long long nvec = mb * 1000000 / 16;
        float4* x; float* out;
        cudaMalloc(&x, nvec * 16 + 1024 * 1024 * 512);  // + a 512 MB flush region
        cudaMalloc(&out, 4);
        cudaMemset(x, 0, nvec * 16);
        char label[64];
        snprintf(label, sizeof label, "%lld MB", mb);
        run<4>(x, nvec, out, 296, 512, label);
        run<8>(x, nvec, out, 296, 512, label);
        run<4>(x, nvec, out, 592, 512, label);
        run<4>(x, nvec, out, 148, 1024, label);
        run<8>(x, nvec, out, 148, 1024, label);
        run<16>(x, nvec, out, 148, 1024, label);
        run<4>(x, nvec, out, 148 * 8, 256, label);
        run<8>(x, nvec, out, 148 * 8, 256, label);
        cudaFree(x); cudaFree(out);
    }
    return 0;
}
