// microbenchmark: issue rates of the instruction kinds the node test is built from, alone and paired,
// to learn which share a pipe on sm_100a.  Prints warp-instructions per cycle per SM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define N_ACC 8
#define ITERS 4096
enum { PRMT, DP4A, LOP, FFMA, FMNMX, IMADSHL, SELOP, FSETOP, PRMT_FFMA, DP4A_FFMA, DP4A_LOP, FMNMX_FFMA, PRMT_LOP, DP4A_PRMT, SHF, SHF_FFMA, FMNMX3, NKIND };
const char* names[] = {"prmt","dp4a","lop3","ffma","fmnmx","imad.shl","sel","fset","prmt+ffma","dp4a+ffma","dp4a+lop3","fmnmx+ffma","prmt+lop3","dp4a+prmt","shf","shf+ffma","fmnmx3"};
template<int K> __global__ void __launch_bounds__(256) k(uint32_t* out, uint32_t seed)
{
    uint32_t a[N_ACC]; float f[N_ACC];
    for (int i=0;i<N_ACC;i++){ a[i]=seed+threadIdx.x*(i+1); f[i]=(float)a[i]; }
    uint32_t sel = seed | 0x3210u; float c1 = 1.0001f, c2 = 0.5f;
    for (int it=0; it<ITERS; it++) {
#pragma unroll
        for (int i=0;i<N_ACC;i++) {
            if (K==PRMT||K==PRMT_FFMA||K==PRMT_LOP||K==DP4A_PRMT) asm volatile("prmt.b32 %0, %0, %1, 0x7604;" : "+r"(a[i]) : "r"(sel));
            if (K==DP4A||K==DP4A_FFMA||K==DP4A_LOP||K==DP4A_PRMT) asm volatile("dp4a.u32.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(sel), "r"(seed));
            if (K==LOP||K==DP4A_LOP||K==PRMT_LOP) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(sel), "r"(seed));
            if (K==FFMA||K==PRMT_FFMA||K==DP4A_FFMA||K==FMNMX_FFMA||K==SHF_FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(c1), "f"(c2));
            if (K==FMNMX||K==FMNMX_FFMA) asm volatile("max.f32 %0, %0, %1;" : "+f"(f[(i+1)%N_ACC]) : "f"(c2));
            if (K==FMNMX3) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(c2), "f"(c1));
            if (K==IMADSHL) asm volatile("mad.lo.u32 %0, %0, 65536, %1;" : "+r"(a[i]) : "r"(seed));
            if (K==SELOP) asm volatile("{.reg .pred p; setp.ne.u32 p, %1, 0; selp.u32 %0, %0, %2, p;}" : "+r"(a[i]) : "r"(seed), "r"(sel));
            if (K==FSETOP) asm volatile("set.le.f32.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(c1));
            if (K==SHF||K==SHF_FFMA) asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(sel), "r"(seed));
        }
    }
    uint32_t s=0; for (int i=0;i<N_ACC;i++) s += a[i] + (uint32_t)f[i];
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
    run<PRMT>(out,1); run<DP4A>(out,1); run<LOP>(out,1); run<FFMA>(out,1); run<FMNMX>(out,1); run<FMNMX3>(out,1); run<IMADSHL>(out,1); run<SELOP>(out,2); run<FSETOP>(out,1); run<SHF>(out,1);
    run<PRMT_FFMA>(out,2); run<DP4A_FFMA>(out,2); run<DP4A_LOP>(out,2); run<FMNMX_FFMA>(out,2); run<PRMT_LOP>(out,2); run<DP4A_PRMT>(out,2); run<SHF_FFMA>(out,2);
    return 0; }
