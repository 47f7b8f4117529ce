// Microbenchmark: random sector gathers from a buffer of a given size (the practical ceiling
// for hash-probe kernels).  Usage: gather_bench   (prints a table)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint64_t mix64(uint64_t x){x^=x>>30;x*=0xbf58476d1ce4e5b9ull;x^=x>>27;x*=0x94d049bb133111ebull;x^=x>>31;return x;}
template<int BYTES, int ILP>
__global__ void gather(const uint4* __restrict__ buf, uint64_t mask_units, uint64_t n, uint32_t* out){
  uint64_t i = (uint64_t)blockIdx.x*blockDim.x+threadIdx.x;
  uint32_t acc=0;
  for(uint64_t k=i*ILP;k<n;k+=(uint64_t)gridDim.x*blockDim.x*ILP){
    uint4 v[ILP][BYTES/16];
    #pragma unroll
    for(int j=0;j<ILP;j++){
      uint64_t u = mix64(k+j) & mask_units;      // unit index (unit = BYTES)
      #pragma unroll
      for(int q=0;q<BYTES/16;q++) v[j][q]=__ldg(buf+u*(BYTES/16)+q);
    }
    #pragma unroll
    for(int j=0;j<ILP;j++)
      #pragma unroll
      for(int q=0;q<BYTES/16;q++) acc+=v[j][q].x^v[j][q].w;
  }
  if(acc==0x12345678) out[0]=acc;
}
template<int BYTES,int ILP> void run(const uint4*buf,uint64_t bytes,uint32_t*out){
  uint64_t units=bytes/BYTES; uint64_t n=1ull<<28;
  cudaEvent_t a,b;cudaEventCreate(&a);cudaEventCreate(&b);
  gather<BYTES,ILP><<<148*8,256>>>(buf,units-1,n/8,out);
  cudaEventRecord(a);
  gather<BYTES,ILP><<<148*8,256>>>(buf,units-1,n,out);
  cudaEventRecord(b);cudaEventSynchronize(b);float ms;cudaEventElapsedTime(&ms,a,b);
  printf("size %6llu MB  unit %3d B  ilp %d : %7.2f G gathers/s  %8.1f GB/s\n",(unsigned long long)(bytes>>20),BYTES,ILP,n/ms/1e6,n*(double)BYTES/ms/1e6);
}
int main(int argc,char**argv){
  if(argc>1){ size_t g=atoi(argv[1]); cudaError_t e=cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity,g); size_t got=0; cudaDeviceGetLimit(&got,cudaLimitMaxL2FetchGranularity); printf("set L2 fetch granularity %zu -> %s, now %zu\n",g,cudaGetErrorString(e),got);} else { size_t got=0; cudaDeviceGetLimit(&got,cudaLimitMaxL2FetchGranularity); printf("default L2 fetch granularity %zu\n",got);}
  uint64_t maxb=8ull<<30; uint4*buf; cudaMalloc(&buf,maxb); cudaMemset(buf,1,maxb); uint32_t*out; cudaMalloc(&out,4);
  for(uint64_t mb: {64ull,256ull,2048ull,8192ull}){
    uint64_t bytes=mb<<20; // power of two needed for mask: round down
    uint64_t p=1; while(p*2<=bytes)p*=2; if(p!=bytes){ // non-pow2: use modulo-free approx by mask of next pow2 (skip)
      continue; }
    run<16,4>(buf,bytes,out); run<32,1>(buf,bytes,out); run<32,4>(buf,bytes,out); run<64,2>(buf,bytes,out); run<128,2>(buf,bytes,out);
  }
  return 0;
}
