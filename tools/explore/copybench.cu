// Exploration only (not product): which launch shape streams HBM best on B200 for send->recv copies?
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o copybench copybench.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("%s: %s\n",#x,cudaGetErrorString(e)); exit(1);} }while(0)

__device__ __forceinline__ uint4 ld_nc(const uint4*p){uint4 r; asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];":"=r"(r.x),"=r"(r.y),"=r"(r.z),"=r"(r.w):"l"(p)); return r;}
__device__ __forceinline__ uint4 ld_sys(const uint4*p){uint4 r; asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];":"=r"(r.x),"=r"(r.y),"=r"(r.z),"=r"(r.w):"l"(p):"memory"); return r;}
__device__ __forceinline__ uint4 ld_plain(const uint4*p){return *p;}
__device__ __forceinline__ void st_na(uint4*p,uint4 v){asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"::"l"(p),"r"(v.x),"r"(v.y),"r"(v.z),"r"(v.w):"memory");}
__device__ __forceinline__ void st_cs(uint4*p,uint4 v){asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};"::"l"(p),"r"(v.x),"r"(v.y),"r"(v.z),"r"(v.w):"memory");}

template<int U,int LD,int TILED,int ST>
__global__ void copyk(const uint4* __restrict__ s, uint4* __restrict__ d, size_t nvec){
  if (TILED){
    const size_t tile=(size_t)U*blockDim.x;
    for(size_t t=blockIdx.x; t*tile<nvec; t+=gridDim.x){
      const size_t base=t*tile+threadIdx.x; uint4 v[U];
      #pragma unroll
      for(int u=0;u<U;++u){size_t i=base+(size_t)u*blockDim.x; if(i<nvec) v[u]= LD==0?ld_nc(s+i): LD==1?ld_sys(s+i):ld_plain(s+i);}
      #pragma unroll
      for(int u=0;u<U;++u){size_t i=base+(size_t)u*blockDim.x; if(i<nvec){ if(ST==0) st_na(d+i,v[u]); else st_cs(d+i,v[u]);} }
    }
  } else {
    const size_t nthr=(size_t)gridDim.x*blockDim.x;
    for(size_t base=(size_t)blockIdx.x*blockDim.x+threadIdx.x; base<nvec; base+=(size_t)U*nthr){
      uint4 v[U];
      #pragma unroll
      for(int u=0;u<U;++u){size_t i=base+(size_t)u*nthr; if(i<nvec) v[u]= LD==0?ld_nc(s+i): LD==1?ld_sys(s+i):ld_plain(s+i);}
      #pragma unroll
      for(int u=0;u<U;++u){size_t i=base+(size_t)u*nthr; if(i<nvec){ if(ST==0) st_na(d+i,v[u]); else st_cs(d+i,v[u]);} }
    }
  }
}

struct Variant{const char*name; void(*fn)(const uint4*,uint4*,size_t); int threads;};
#define V(U,LD,TI,ST,TH) {#U "x" #TH " ld" #LD " tiled" #TI " st" #ST, copyk<U,LD,TI,ST>, TH}

int main(int argc,char**argv){
  int sms; CK(cudaDeviceGetAttribute(&sms,cudaDevAttrMultiProcessorCount,0));
  std::vector<Variant> vs={
    V(8,1,0,0,512), V(8,0,0,0,512), V(8,0,1,0,512), V(4,0,1,0,512), V(4,0,1,0,1024), V(2,0,1,0,1024), V(4,0,1,0,256), V(8,0,1,0,256),
    V(4,0,0,0,256), V(4,2,1,0,512), V(4,0,1,1,512), V(8,0,1,1,256), V(2,0,1,0,512), V(16,0,1,0,128), V(8,0,1,0,128)};
  size_t sizes[]={66918844/16*16, 268435456, 1073741824};
  const size_t ring=(size_t)3<<30; char*a,*b; CK(cudaMalloc(&a,ring)); CK(cudaMalloc(&b,ring)); CK(cudaMemset(a,1,ring)); CK(cudaMemset(b,2,ring));
  cudaEvent_t e0,e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for(size_t S:sizes){
    int nbuf=(int)(ring/S); int iters= nbuf<4?8: (nbuf>40?40:nbuf);
    // reference: cudaMemcpyAsync D2D
    for(int rep=0;rep<2;++rep){ CK(cudaEventRecord(e0)); for(int i=0;i<iters;++i) CK(cudaMemcpyAsync(b+(size_t)(i%nbuf)*S,a+(size_t)(i%nbuf)*S,S,cudaMemcpyDeviceToDevice)); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); }
    float ms; CK(cudaEventElapsedTime(&ms,e0,e1)); printf("S=%zu cudaMemcpy D2D                      %8.1f GB/s\n",S,2.0*S*iters/(ms*1e-3)/1e9);
    for(auto&v:vs){
      int occ=0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ,v.fn,v.threads,0));
      for(int mult: {1,2}){
        int grid=sms*occ*mult; if(mult==2 && occ>=4) continue;
        for(int rep=0;rep<2;++rep){ CK(cudaEventRecord(e0)); for(int i=0;i<iters;++i) v.fn<<<grid,v.threads>>>((const uint4*)(a+(size_t)(i%nbuf)*S),(uint4*)(b+(size_t)(i%nbuf)*S),S/16); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); }
        CK(cudaGetLastError()); CK(cudaEventElapsedTime(&ms,e0,e1));
        printf("S=%zu %-28s occ=%d grid=%5d %8.1f GB/s  (%.1f us)\n",S,v.name,occ,grid,2.0*S*iters/(ms*1e-3)/1e9,ms*1e3/iters);
      }
    }
  }
  return 0;
}
