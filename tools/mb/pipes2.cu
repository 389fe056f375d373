// microbenchmark 2: packed-fp32 (FFMA2, sm_100 fma.rn.f32x2) and packed-half issue rates, alone and paired
// with the other instruction kinds of the node test.  Prints warp-instructions per cycle per SM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define N_ACC 8
#define ITERS 4096
enum { FFMA2, FFMA2_DP4A, FFMA2_FMNMX, FFMA2_LOP, FFMA2_FFMA, HFMA2, HMNMX2, FMUL2, FADD2, NKIND };
const char* names[] = {"ffma2","ffma2+dp4a","ffma2+fmnmx","ffma2+lop3","ffma2+ffma","hfma2","hmnmx2","fmul2","fadd2"};
template<int K> __global__ void __launch_bounds__(256) k(uint32_t* out, uint32_t seed)
{
    uint32_t a[N_ACC]; float f[N_ACC]; unsigned long long p[N_ACC]; uint32_t h[N_ACC];
    for (int i=0;i<N_ACC;i++){ a[i]=seed+threadIdx.x*(i+1); f[i]=(float)a[i]; p[i]=((unsigned long long)a[i]<<32)|a[i]; h[i]=a[i]; }
    uint32_t sel = seed | 0x3210u; float c1 = 1.0001f, c2 = 0.5f;
    unsigned long long c1p, c2p;
    asm("mov.b64 %0, {%1,%2};" : "=l"(c1p) : "f"(c1), "f"(c1));
    asm("mov.b64 %0, {%1,%2};" : "=l"(c2p) : "f"(c2), "f"(c2));
    for (int it=0; it<ITERS; it++) {
#pragma unroll
        for (int i=0;i<N_ACC;i++) {
            if (K==FFMA2||K==FFMA2_DP4A||K==FFMA2_FMNMX||K==FFMA2_LOP||K==FFMA2_FFMA) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(c1p), "l"(c2p));
            if (K==FMUL2) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(c1p));
            if (K==FADD2) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(c1p));
            if (K==FFMA2_DP4A) asm volatile("dp4a.u32.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(sel), "r"(seed));
            if (K==FFMA2_FMNMX) asm volatile("max.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(c2));
            if (K==FFMA2_LOP) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(sel), "r"(seed));
            if (K==FFMA2_FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(c1), "f"(c2));
            if (K==HFMA2) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(h[i]) : "r"(sel), "r"(seed));
            if (K==HMNMX2) asm volatile("max.f16x2 %0, %0, %1;" : "+r"(h[i]) : "r"(sel));
        }
    }
    uint32_t s=0; for (int i=0;i<N_ACC;i++) s += a[i] + (uint32_t)f[i] + (uint32_t)p[i] + (uint32_t)(p[i]>>32) + h[i];
    if (s==0x12345) out[0]=s;
}
template<int K> void run(uint32_t* out, int ninstr)
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
    k<K><<<sms*4,256>>>(out, 0); cudaDeviceSynchronize();
    cudaEventRecord(a); k<K><<<sms*4,256>>>(out, 0); cudaEventRecord(b); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms,a,b);
    double cyc = ms*1e-3*clk*1e3;
    double winstr_per_sm = 4.0*8*ITERS*N_ACC*ninstr;   // 4 CTAs x 8 warps
    printf("%-12s %6.2f warp-instr/clk/SM (%.3f ms)\n", names[K], winstr_per_sm/cyc, ms);
}
int main(){ uint32_t* out; cudaMalloc(&out,4);
    run<FFMA2>(out,1); run<FMUL2>(out,1); run<FADD2>(out,1); run<FFMA2_DP4A>(out,2); run<FFMA2_FMNMX>(out,2); run<FFMA2_LOP>(out,2); run<FFMA2_FFMA>(out,2);
    run<HFMA2>(out,1); run<HMNMX2>(out,1);
    return 0; }
