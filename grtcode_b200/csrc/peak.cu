// Measured FMA throughput of the non-tensor FP64 and FP32 pipes: the roofline
// denominators of the Voigt sweep.  MEASURED_PEAKS.json carries only HBM and
// bf16 tensor figures, so the library measures these itself.
#include "common.cuh"

namespace grt {

template <typename T>
__global__ void __launch_bounds__(256) fma_chain_kernel(T *out, int const iters, T const a, T const b)
{
    T x[8];
#pragma unroll
    for (int k=0; k<8; ++k) x[k] = (T)(threadIdx.x + k);
    for (int it=0; it<iters; ++it)
    {
#pragma unroll
        for (int r=0; r<8; ++r)
        {
#pragma unroll
            for (int k=0; k<8; ++k) x[k] = fma(x[k], a, b);
        }
    }
    T s = 0;
#pragma unroll
    for (int k=0; k<8; ++k) s += x[k];
    if (s == (T)123456789) out[0] = s;   // never true: keeps the chain alive
}

template <typename T>
static int measure(int device, cudaStream_t st, double *tflops)
{
    T *out;
    GRT_CUDA(cudaMalloc(&out, sizeof(T)));
    int const blocks = sm_count(device)*8;
    int const iters = 4096;
    cudaEvent_t e0, e1;
    GRT_CUDA(cudaEventCreate(&e0));
    GRT_CUDA(cudaEventCreate(&e1));
    double best = 0.;
    for (int rep=0; rep<4; ++rep)
    {
        GRT_CUDA(cudaEventRecord(e0, st));
        fma_chain_kernel<T><<<blocks, 256, 0, st>>>(out, iters, (T)0.999999, (T)1e-7);
        GRT_CUDA(cudaEventRecord(e1, st));
        GRT_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        GRT_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        double const flops = 2.0*64.0*(double)iters*256.0*(double)blocks;
        double const tf = flops/((double)ms*1e-3)/1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    GRT_CUDA(cudaEventDestroy(e0));
    GRT_CUDA(cudaEventDestroy(e1));
    GRT_CUDA(cudaFree(out));
    *tflops = best;
    return 0;
}

}  // namespace grt

extern "C" int grt_measure_fma_peak(int device, double *fp64_tflops, double *fp32_tflops)
{
    cudaStream_t st;
    GRT_TRY(grt::stream_of(device, &st));
    GRT_TRY(grt::measure<double>(device, st, fp64_tflops));
    GRT_TRY(grt::measure<float>(device, st, fp32_tflops));
    return 0;
}
