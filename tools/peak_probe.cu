// peak_probe.cu — measured FP32 / FP64 FMA throughput of the device (the denominators of the SIMT roofline).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/peak_probe tools/peak_probe.cu && tools/peak_probe
// Each thread runs 8 independent FMA chains; 2 flop per FMA.  Prints one JSON line.
#include <cstdio>
#include <cuda_runtime.h>

template <typename T>
__global__ void __launch_bounds__(256) fma_kernel(T *out, int iters, T a, T b)
{
    T x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            x0 = x0 * a + b; x1 = x1 * a + b; x2 = x2 * a + b; x3 = x3 * a + b;
            x4 = x4 * a + b; x5 = x5 * a + b; x6 = x6 * a + b; x7 = x7 * a + b;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

template <typename T>
double measure(int n_sm, int iters)
{
    const int blocks = n_sm * 8, threads = 256;
    T *out;
    cudaMalloc(&out, sizeof(T) * blocks * threads);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(e0);
        fma_kernel<T><<<blocks, threads>>>(out, iters, (T) 1.0000001, (T) 1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * 64.0 * iters * (double) blocks * threads;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaFree(out);
    return best;
}

int main()
{
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, 0) != cudaSuccess) {
        std::printf("{\"error\": \"no CUDA device\"}\n");
        return 1;
    }
    const double f32 = measure<float>(p.multiProcessorCount, 4000);
    const double f64 = measure<double>(p.multiProcessorCount, 2000);
    std::printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_mhz_max\": %d, \"fp32_fma_tflops\": %.2f, \"fp64_fma_tflops\": %.2f, "
                "\"fp32_nominal_tflops\": %.2f}\n", p.name, p.multiProcessorCount, p.clockRate / 1000, f32, f64,
                p.multiProcessorCount * 128 * 2 * (p.clockRate / 1e6) / 1e3);
    return 0;
}
