// microbenchmark: throughput of lane-divergent node fetches through L1 on sm_100a.
// Every lane fetches a whole "node" of NODE_BYTES at a random node index with K loads of width W.
// Reports SM cycles per warp-level node fetch (per SM), for L1-resident and L2-resident footprints.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("%s: %s\n",#x,cudaGetErrorString(e)); return 1;}}while(0)

__device__ __forceinline__ uint32_t ld128x(const void* p){ uint4 v = __ldg((const uint4*)p); return v.x^v.y^v.z^v.w; }
__device__ __forceinline__ uint32_t ld256x(const void* p){
    uint32_t a,b,c,d,e,f,g,h;
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];":"=r"(a),"=r"(b),"=r"(c),"=r"(d),"=r"(e),"=r"(f),"=r"(g),"=r"(h):"l"(p));
    return a^b^c^d^e^f^g^h;
}
// MODE 0: 8 x 128-bit (128 B node); 1: 4 x 256-bit (128 B); 2: 5 x 128-bit (80 B node); 3: 2 x 256 (64 B node)
// 4: 3 x 256 (96 B node) ; 5: 1 x 128 (16 B) ; 6: 1 x 256
template<int MODE>
__global__ void __launch_bounds__(128) k(const char* __restrict__ base, uint32_t n_nodes, int iters, uint32_t* out, int share)
{
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    if (share > 1) { uint32_t l = threadIdx.x & 31; s = (blockIdx.x * blockDim.x + (threadIdx.x & ~31u) + l / share) * 2654435761u + 12345u; }
    uint32_t acc = 0;
    constexpr int NB = MODE==0||MODE==1 ? 128 : MODE==2 ? 80 : MODE==3 ? 64 : MODE==4 ? 96 : MODE==5 ? 16 : 32;
    for (int it = 0; it < iters; it++) {
        s = s * 1664525u + 1013904223u;
        // dependent chain like a traversal: next index depends on loaded data (acc is always 0 in practice)
        const uint32_t idx = ((s >> 8) + acc) % n_nodes;   // data is all zero: acc stays 0, but the chain is real
        const char* p = base + (size_t)idx * NB;
        if (MODE==0) { for (int j=0;j<8;j++) acc ^= ld128x(p + 16*j); }
        if (MODE==1) { for (int j=0;j<4;j++) acc ^= ld256x(p + 32*j); }
        if (MODE==2) { for (int j=0;j<5;j++) acc ^= ld128x(p + 16*j); }
        if (MODE==3) { for (int j=0;j<2;j++) acc ^= ld256x(p + 32*j); }
        if (MODE==4) { for (int j=0;j<3;j++) acc ^= ld256x(p + 32*j); }
        if (MODE==5) { acc ^= ld128x(p); }
        if (MODE==6) { acc ^= ld256x(p); }
    }
    if (acc) out[0] = acc;
}

template<int MODE> int run(const char* name, const char* buf, size_t bytes, int nb, int ctas, int share)
{
    uint32_t* out; CK(cudaMalloc(&out, 4));
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int iters = 2000;
    uint32_t n_nodes = (uint32_t)(bytes / nb);
    cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
    k<MODE><<<sms*ctas,128>>>(buf, n_nodes, 200, out, share);
    CK(cudaDeviceSynchronize());
    cudaEventRecord(a);
    k<MODE><<<sms*ctas,128>>>(buf, n_nodes, iters, out, share);
    cudaEventRecord(b); CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms,a,b);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    double cyc = ms*1e-3*clk*1e3;                       // SM cycles (at max clock)
    double warp_fetches_per_sm = (double)ctas*4*iters;
    printf("%-22s footprint %8.1f KB ctas/SM %d share %d: %7.1f cyc per warp-node-fetch per SM, %6.2f G lane-nodes/s\n",
           name, bytes/1024.0, ctas, share, cyc/warp_fetches_per_sm, (double)sms*ctas*128*iters/(ms*1e-3)/1e9);
    cudaFree(out); return 0;
}

int main()
{
    char* buf; size_t cap = 64u<<20; CK(cudaMalloc(&buf, cap)); CK(cudaMemset(buf, 0, cap));
    for (size_t bytes : {(size_t)32<<10, (size_t)8<<20, (size_t)48<<20}) {
        for (int share : {1, 4}) {
            run<0>("8x128b (128B node)", buf, bytes, 128, 5, share);
            run<1>("4x256b (128B node)", buf, bytes, 128, 5, share);
            run<2>("5x128b (80B node)",  buf, bytes, 80, 5, share);
            run<4>("3x256b (96B node)",  buf, bytes, 96, 5, share);
            run<3>("2x256b (64B node)",  buf, bytes, 64, 5, share);
            run<5>("1x128b (16B)",       buf, bytes, 16, 5, share);
            run<6>("1x256b (32B)",       buf, bytes, 32, 5, share);
        }
    }
    run<1>("4x256b (128B node)", buf, (size_t)8<<20, 128, 8, 1);
    run<1>("4x256b (128B node)", buf, (size_t)8<<20, 128, 12, 1);
    run<3>("2x256b (64B node)",  buf, (size_t)8<<20, 64, 12, 1);
    return 0;
}
