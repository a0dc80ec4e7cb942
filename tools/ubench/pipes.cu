// Micro-benchmark of B200 issue/pipe behaviour for the instruction mix of the wing sweep:
// cycles per loop iteration (one warp's view, W warps per SM sub-partition) for bodies made
// of independent DFMA / FFMA / FFMA2 / F2F.F32.F64 / MUFU.RCP / LDS.64 / IMAD streams.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu ; run: ./pipes
#include <cstdio>
#include <cuda_runtime.h>

template <int ND, int NF, int NF2, int NC, int NM, int NL, int NI>
__global__ void body(double *out, long long *cycles, int iters, double a, float b)
{
    __shared__ double sm[1024];
    sm[threadIdx.x & 1023] = a;
    __syncthreads();
    double d[8]; float f[8]; unsigned long long p[8]; float c[8]; float m[8]; double l[8]; int ii[8];
#pragma unroll
    for (int k=0; k<8; ++k) { d[k] = a + k; f[k] = b + k; p[k] = (unsigned long long)(threadIdx.x + k)*0x3f8000013f800001ull; c[k] = b; m[k] = b + 2.f + k; l[k] = 0.; ii[k] = threadIdx.x + k; }
    long long t0 = clock64();
    for (int it=0; it<iters; ++it)
    {
#pragma unroll
        for (int r=0; r<8; ++r)
        {
            if (r < ND) d[r] = fma(d[r], a, a);
            if (r < NF) f[r] = fmaf(f[r], b, b);
            if (r < NF2) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(p[r]) : "l"(0x3f8000003f800000ull));
            if (r < NC) { c[r] = __double2float_rn(d[(r + 4) & 7] + (double)c[r]*0.); }
            if (r < NM) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(m[r]));
            if (r < NL) l[r] += sm[(threadIdx.x + 32*r + it) & 1023];
            if (r < NI) ii[r] = ii[r]*3 + it;
        }
        if (ND > 8) {
#pragma unroll
            for (int r=0; r<ND-8; ++r) d[r] = fma(d[r], a, a);
        }
        if (NF > 8) {
#pragma unroll
            for (int r=0; r<NF-8; ++r) f[r & 7] = fmaf(f[r & 7], b, b);
        }
    }
    long long t1 = clock64();
    double s = 0.;
#pragma unroll
    for (int k=0; k<8; ++k) s += d[k] + f[k] + (double)p[k] + c[k] + m[k] + l[k] + ii[k];
    if (s == 1.2345) out[0] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
}

template <int ND, int NF, int NF2, int NC, int NM, int NL, int NI>
void run(char const *name)
{
    double *out; long long *cyc;
    cudaMalloc(&out, 8); cudaMalloc(&cyc, 8);
    int const iters = 20000;
    for (int warps_per_smsp : {1, 2, 4, 6, 8})
    {
        int threads = 128*warps_per_smsp;
        if (threads > 1024) continue;
        body<ND, NF, NF2, NC, NM, NL, NI><<<148, threads>>>(out, cyc, iters, 1.0000001, 1.0000001f);
        cudaDeviceSynchronize();
        body<ND, NF, NF2, NC, NM, NL, NI><<<148, threads>>>(out, cyc, iters, 1.0000001, 1.0000001f);
        cudaDeviceSynchronize();
        long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        int const n = (ND > 8 ? ND : ND) + NF + NF2 + NC + NM + NL + NI;
        printf("%-34s warps/SMSP %d: %7.2f cycles/iter/warp-set  (%d instr/iter/warp, %.2f issue/clk/SMSP)\n",
               name, warps_per_smsp, (double)h/iters, n, (double)n*warps_per_smsp/((double)h/iters));
    }
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    run<8,0,0,0,0,0,0>("8 DFMA");
    run<0,8,0,0,0,0,0>("8 FFMA");
    run<0,0,8,0,0,0,0>("8 FFMA2");
    run<0,0,0,0,8,0,0>("8 MUFU.RCP");
    run<8,0,0,8,0,0,0>("8 DADD+8 F2F.F32.F64 (dep)");
    run<0,0,0,0,0,8,0>("8 LDS.64+DADD");
    run<0,0,0,0,0,0,8>("8 IMAD");
    run<8,8,0,0,0,0,0>("8 DFMA + 8 FFMA");
    run<8,0,0,0,0,0,8>("8 DFMA + 8 IMAD");
    run<8,0,8,0,0,0,0>("8 DFMA + 8 FFMA2");
    run<0,8,8,0,0,0,0>("8 FFMA + 8 FFMA2");
    run<8,0,0,0,8,0,0>("8 DFMA + 8 MUFU");
    run<0,8,0,0,8,0,0>("8 FFMA + 8 MUFU");
    run<0,8,0,0,8,0,8>("8 FFMA + 8 MUFU + 8 IMAD");
    run<8,8,8,0,8,0,8>("8 DFMA+8 FFMA+8 FFMA2+8 MUFU+8 IMAD");
    run<8,8,8,8,8,8,8>("all seven x8");
    return 0;
}
