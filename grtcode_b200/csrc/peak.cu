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

// ---- self-tests of the arithmetic shortcuts of the sweep (voigt.cuh) -----------------
#include "voigt.cuh"

namespace grt {

// div_exact vs IEEE division and widen_positive vs (double) on pseudo-random operands in
// the ranges the sweep produces: a in [1e-19, 1e2], b in [1, 1e16].
__global__ void selftest_kernel(unsigned long long n, unsigned long long seed,
                                unsigned long long *bad)
{
    unsigned long long i = (unsigned long long)blockIdx.x*blockDim.x + threadIdx.x;
    unsigned long long local = 0;
    for (; i<n; i+=(unsigned long long)gridDim.x*blockDim.x)
    {
        unsigned long long z = seed + i*0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30))*0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27))*0x94D049BB133111EBull;
        z ^= z >> 31;
        unsigned const lo = (unsigned)z;
        unsigned const hi = (unsigned)(z >> 32);
        // random mantissas, exponents spread over the ranges above
        float const a = __uint_as_float(((64u + lo%70u) << 23) | (hi & 0x7fffffu));
        float const b = __uint_as_float(((127u + (lo >> 8)%54u) << 23) | ((hi >> 9) & 0x7fffffu) |
                                        ((lo >> 20) & 0x1ffu));
        float const q1 = div_exact(a, b);
        float const q2 = __fdiv_rn(a, b);
        if (__float_as_uint(q1) != __float_as_uint(q2)) ++local;
        if (widen_positive(q2) != (double)q2) ++local;
        // the packed (FFMA2) sequence, the second half fed with the swapped-mantissa pair
        float const a2 = __uint_as_float(((64u + hi%70u) << 23) | (lo & 0x7fffffu));
        float const b2 = __uint_as_float(((127u + (hi >> 8)%54u) << 23) | ((lo >> 9) & 0x7fffffu));
        float p1, p2;
        unpack2(div_exact2(pack2(a, a2), pack2(b, b2)), p1, p2);
        if (__float_as_uint(p1) != __float_as_uint(q2)) ++local;
        if (__float_as_uint(p2) != __float_as_uint(__fdiv_rn(a2, b2))) ++local;
        unpack2(div_exact2_seeded(pack2(a, a2), pack2(b, b2)), p1, p2);
        if (__float_as_uint(p1) != __float_as_uint(q2)) ++local;
        if (__float_as_uint(p2) != __float_as_uint(__fdiv_rn(a2, b2))) ++local;
        // XU-free |x|: random (t2, v, REPWID) of the sweep's ranges, and exact ties (a double
        // half-way between two neighbouring floats, times one)
        {
            double const v0 = 1. + (double)(lo >> 4)*1e-4;                    // 1 .. 26844 cm-1
            double const dv = ((double)(hi >> 8) - 8388608.)*3e-6 + 1e-9;     // +-25 cm-1
            float const rw = __uint_as_float(((127u + 4u + (lo & 7u)) << 23) | (hi & 0x7fffffu));
            double const rwd = widen_positive(rw);
            double const t2 = v0 + dv;
            float const want = __double2float_rn(__dmul_rn(fabs(__dsub_rn(t2, v0)), rwd));
            if (want >= 64.f && want < 3e7f &&
                __float_as_uint(abs_abscissa_alu(t2, v0, rwd)) != __float_as_uint(want)) ++local;
            float const fl = __uint_as_float(((127u + 6u + lo%18u) << 23) | (hi & 0x7fffffu));
            float const fn = __uint_as_float(__float_as_uint(fl) + 1u);
            double const mid = 0.5*((double)fl + (double)fn);
            double const near_mid = __hiloint2double(__double2hiint(mid), __double2loint(mid) + (int)(lo & 3u) - 1);
            if (__float_as_uint(abs_abscissa_alu(mid, 0., 1.)) != __float_as_uint(__double2float_rn(mid))) ++local;
            if (__float_as_uint(abs_abscissa_alu(near_mid, 0., 1.)) !=
                __float_as_uint(__double2float_rn(near_mid))) ++local;
            if (__float_as_uint(abs_abscissa_alu((double)fn, 0., 1.)) != __float_as_uint(fn)) ++local;
        }
        // x*x + yq must round twice (RFM_voigt.c:166,170): the packed add of scalar products
        float s1, s2;
        unpack2(add2(mul2_unfused(pack2(a, a2), pack2(a, a2)), pack2(b, b2)), s1, s2);
        if (__float_as_uint(s1) != __float_as_uint(__fadd_rn(__fmul_rn(a, a), b))) ++local;
        if (__float_as_uint(s2) != __float_as_uint(__fadd_rn(__fmul_rn(a2, a2), b2))) ++local;
    }
    if (local) atomicAdd(bad, local);
}

}  // namespace grt

extern "C" int grt_selftest_arithmetic(int device, unsigned long long n, unsigned long long seed,
                                       unsigned long long *mismatches)
{
    cudaStream_t st;
    GRT_TRY(grt::stream_of(device, &st));
    unsigned long long *d;
    GRT_CUDA(cudaMalloc(&d, sizeof(*d)));
    GRT_CUDA(cudaMemsetAsync(d, 0, sizeof(*d), st));
    grt::selftest_kernel<<<grt::sm_count(device)*8, 256, 0, st>>>(n, seed, d);
    GRT_CUDA(cudaGetLastError());
    GRT_CUDA(cudaMemcpyAsync(mismatches, d, sizeof(*d), cudaMemcpyDeviceToHost, st));
    GRT_CUDA(cudaStreamSynchronize(st));
    GRT_CUDA(cudaFree(d));
    return 0;
}


// ---- self-test: the per-line Voigt evaluation (hoisted constants, packed CPF12, exact
// division without the slow path) against the straight restatement voigt_core -----------
namespace grt {

__global__ void selftest_voigt_kernel(unsigned long long n, unsigned long long seed,
                                      unsigned long long *bad)
{
    unsigned long long i = (unsigned long long)blockIdx.x*blockDim.x + threadIdx.x;
    unsigned long long local = 0;
    for (; i<n; i+=(unsigned long long)gridDim.x*blockDim.x)
    {
        unsigned long long z = seed + i*0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30))*0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27))*0x94D049BB133111EBull;
        z ^= z >> 31;
        // Y log-uniform over [1e-7, 70], x uniform over [-1.2 XLIM0, 1.2 XLIM0]
        float const u = (float)(z & 0xffffffu)*(1.f/16777216.f);
        float const w = (float)((z >> 24) & 0xffffffu)*(1.f/16777216.f);
        float const y = exp2f(-23.f + 29.1f*u);
        float xlim0, xlim1;
        voigt_limits(y, &xlim0, &xlim1);
        float x = (2.f*w - 1.f)*1.2f*xlim0;
        if ((z >> 48) & 1) x *= 0.1f;                 // half of the samples near the centre
        VoigtLine line;
        voigt_line_init(line, y, xlim0, xlim1);
        double const a = voigt_line_eval(line, x);
        double const b = voigt_core(x, y, xlim0, xlim1);
        if (__double_as_longlong(a) != __double_as_longlong(b)) ++local;
    }
    if (local) atomicAdd(bad, local);
}

}  // namespace grt

extern "C" int grt_selftest_voigt(int device, unsigned long long n, unsigned long long seed,
                                  unsigned long long *mismatches)
{
    cudaStream_t st;
    GRT_TRY(grt::stream_of(device, &st));
    unsigned long long *d;
    GRT_CUDA(cudaMalloc(&d, sizeof(*d)));
    GRT_CUDA(cudaMemsetAsync(d, 0, sizeof(*d), st));
    grt::selftest_voigt_kernel<<<grt::sm_count(device)*8, 256, 0, st>>>(n, seed, d);
    GRT_CUDA(cudaGetLastError());
    GRT_CUDA(cudaMemcpyAsync(mismatches, d, sizeof(*d), cudaMemcpyDeviceToHost, st));
    GRT_CUDA(cudaStreamSynchronize(st));
    GRT_CUDA(cudaFree(d));
    return 0;
}
