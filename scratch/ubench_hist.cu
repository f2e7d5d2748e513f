// Throw-away microbenchmark: which histogram-update mechanism is fastest on B200?
//  A) global RED (L2 atomics) SoA planes   B) global RED AoS 32B cell
//  C) smem CAS atomics (1 node)            D) owner-computes non-atomic smem RMW (lane=feature)
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)

constexpr int FS = 22;      // features per node (C3 onethird)
constexpr int NB = 64;      // bins

__device__ __forceinline__ uint32_t hash32(uint32_t x){ x^=x>>16; x*=0x7feb352dU; x^=x>>15; x*=0x846ca68bU; x^=x>>16; return x; }

__global__ void gen(uint8_t* bins_cm, uint8_t* bins_rm, double* y, int* slot, int N, int S){
  int i = blockIdx.x*blockDim.x+threadIdx.x; if(i>=N) return;
  for(int j=0;j<32;j++){ uint8_t b = hash32(i*37u+j*1000003u)%NB; if(j<FS) bins_cm[(size_t)j*N+i]=b; bins_rm[(size_t)i*32+j]=b; }
  y[i] = (double)(hash32(i^0xabcdefu)%1000)/7.0;
  slot[i] = hash32(i*11u+5u)%S;
}

// A: lane=row, loop over features, 3 REDs per update, SoA planes
__global__ void red_soa(const uint8_t* __restrict__ bins, const double* __restrict__ y, const int* __restrict__ slot,
                        unsigned* cnt, double* sum, double* sq, int N){
  for(int i = blockIdx.x*blockDim.x+threadIdx.x; i<N; i+=gridDim.x*blockDim.x){
    double v=y[i]; int s=slot[i]; double vv=v*v;
    #pragma unroll
    for(int j=0;j<FS;j++){
      int b = bins[(size_t)j*N+i];
      size_t c = ((size_t)s*FS+j)*NB+b;
      atomicAdd(&cnt[c],1u); atomicAdd(&sum[c],v); atomicAdd(&sq[c],vv);
    }
  }
}
struct Cell { double sum, sq; unsigned cnt; unsigned pad; };
__global__ void red_aos(const uint8_t* __restrict__ bins, const double* __restrict__ y, const int* __restrict__ slot,
                        Cell* cells, int N){
  for(int i = blockIdx.x*blockDim.x+threadIdx.x; i<N; i+=gridDim.x*blockDim.x){
    double v=y[i]; int s=slot[i]; double vv=v*v;
    #pragma unroll
    for(int j=0;j<FS;j++){
      int b = bins[(size_t)j*N+i];
      Cell* c = &cells[((size_t)s*FS+j)*NB+b];
      atomicAdd(&c->cnt,1u); atomicAdd(&c->sum,v); atomicAdd(&c->sq,vv);
    }
  }
}
// C: smem atomics, single node (slot ignored), flush with RED
__global__ void smem_cas(const uint8_t* __restrict__ bins, const double* __restrict__ y,
                        unsigned* cnt, double* sum, double* sq, int N){
  __shared__ double ssum[FS*NB]; __shared__ double ssq[FS*NB]; __shared__ unsigned scnt[FS*NB];
  for(int k=threadIdx.x;k<FS*NB;k+=blockDim.x){ssum[k]=0;ssq[k]=0;scnt[k]=0;}
  __syncthreads();
  for(int i = blockIdx.x*blockDim.x+threadIdx.x; i<N; i+=gridDim.x*blockDim.x){
    double v=y[i]; double vv=v*v;
    #pragma unroll
    for(int j=0;j<FS;j++){
      int b = bins[(size_t)j*N+i]; int c=j*NB+b;
      atomicAdd(&scnt[c],1u); atomicAdd(&ssum[c],v); atomicAdd(&ssq[c],vv);
    }
  }
  __syncthreads();
  for(int k=threadIdx.x;k<FS*NB;k+=blockDim.x){atomicAdd(&cnt[k],scnt[k]);atomicAdd(&sum[k],ssum[k]);atomicAdd(&sq[k],ssq[k]);}
}
// D: owner-computes. Each warp owns a private [NB][32] histogram (lane = feature). Rows of "its node" are a
// contiguous chunk here (stands in for a node-sorted index list; idx[] indirection included).
template<int UNROLL>
__global__ void owner(const uint8_t* __restrict__ bins_rm, const double* __restrict__ y, const int* __restrict__ idx,
                      unsigned* cnt, double* sum, double* sq, int N, int rows_per_warp){
  extern __shared__ __align__(16) unsigned char smem[];
  const int W = blockDim.x/32, w = threadIdx.x/32, lane = threadIdx.x%32;
  double* ssum = (double*)smem + (size_t)w*NB*32;
  double* ssq  = (double*)smem + (size_t)W*NB*32 + (size_t)w*NB*32;
  unsigned* scnt = (unsigned*)((double*)smem + (size_t)2*W*NB*32) + (size_t)w*NB*32;
  for(int k=lane;k<NB*32;k+=32){ssum[k]=0;ssq[k]=0;scnt[k]=0;}
  __syncwarp();
  long gw = (long)blockIdx.x*W+w;
  long r0 = gw*rows_per_warp, r1 = r0+rows_per_warp; if(r1>N) r1=N;
  for(long base=r0; base<r1; base+=32){
    int myidx = (base+lane<r1)? idx[base+lane] : -1;
    int cntrows = (int)min((long)32, r1-base);
    for(int k=0;k<cntrows;k+=UNROLL){
      int b[UNROLL]; double v[UNROLL];
      #pragma unroll
      for(int u=0;u<UNROLL;u++){
        int r = __shfl_sync(0xffffffffu, myidx, (k+u)&31);
        bool ok = (k+u)<cntrows;
        b[u] = ok ? bins_rm[(size_t)r*32+lane] : -1;
        v[u] = ok ? y[r] : 0.0;
      }
      #pragma unroll
      for(int u=0;u<UNROLL;u++){
        if(b[u]>=0){
          int c = b[u]*32+lane;
          scnt[c] += 1u; ssum[c] += v[u]; ssq[c] += v[u]*v[u];
        }
      }
    }
  }
  __syncwarp();
  // flush: cell (feature=lane<FS, bin)
  for(int k=lane;k<NB*32;k+=32){ int b=k/32, j=k%32; if(j<FS){ int c=j*NB+b; atomicAdd(&cnt[c],scnt[k]); atomicAdd(&sum[c],ssum[k]); atomicAdd(&sq[c],ssq[k]); } }
}

template<class F> float timeit(F f, int reps=5){
  cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); CK(cudaDeviceSynchronize());
  float best=1e30f;
  for(int r=0;r<reps;r++){ cudaEventRecord(a); f(); cudaEventRecord(b); CK(cudaEventSynchronize(b)); float ms; cudaEventElapsedTime(&ms,a,b); if(ms<best) best=ms; }
  return best;
}

int main(){
  const int N = 10'000'000;
  uint8_t *bins_cm,*bins_rm; double* y; int* slot; int* idx;
  CK(cudaMalloc(&bins_cm,(size_t)FS*N)); CK(cudaMalloc(&bins_rm,(size_t)32*N)); CK(cudaMalloc(&y,(size_t)N*8)); CK(cudaMalloc(&slot,(size_t)N*4)); CK(cudaMalloc(&idx,(size_t)N*4));
  { std::vector<int> h(N); for(int i=0;i<N;i++) h[i]=i; CK(cudaMemcpy(idx,h.data(),(size_t)N*4,cudaMemcpyHostToDevice)); }
  int Ss[] = {1, 8, 64, 512, 8192};
  size_t maxcells = (size_t)8192*FS*NB;
  unsigned* cnt; double *sum,*sq; Cell* cells;
  CK(cudaMalloc(&cnt,maxcells*4)); CK(cudaMalloc(&sum,maxcells*8)); CK(cudaMalloc(&sq,maxcells*8)); CK(cudaMalloc(&cells,maxcells*sizeof(Cell)));
  double upd = (double)N*FS;
  for(int S: Ss){
    gen<<<(N+255)/256,256>>>(bins_cm,bins_rm,y,slot,N,S); CK(cudaDeviceSynchronize());
    size_t cellsN=(size_t)S*FS*NB;
    CK(cudaMemset(cnt,0,cellsN*4)); CK(cudaMemset(sum,0,cellsN*8)); CK(cudaMemset(sq,0,cellsN*8)); CK(cudaMemset(cells,0,cellsN*sizeof(Cell)));
    for(int bs: {256}){
      float ms = timeit([&]{ red_soa<<<148*8,bs>>>(bins_cm,y,slot,cnt,sum,sq,N); });
      printf("A red_soa  S=%5d cells=%8zu (%.1f MB): %.3f ms  %.2f Gupd/s\n",S,cellsN,cellsN*20/1e6,ms,upd/ms/1e6);
      ms = timeit([&]{ red_aos<<<148*8,bs>>>(bins_cm,y,slot,cells,N); });
      printf("B red_aos  S=%5d cells=%8zu (%.1f MB): %.3f ms  %.2f Gupd/s\n",S,cellsN,cellsN*32/1e6,ms,upd/ms/1e6);
    }
  }
  CK(cudaMemset(cnt,0,maxcells*4)); CK(cudaMemset(sum,0,maxcells*8)); CK(cudaMemset(sq,0,maxcells*8));
  { float ms = timeit([&]{ smem_cas<<<148*4,256>>>(bins_cm,y,cnt,sum,sq,N); });
    printf("C smem_cas S=1: %.3f ms  %.2f Gupd/s\n",ms,upd/ms/1e6); }
  // verify D against C counts
  std::vector<unsigned> hc(FS*NB), hd(FS*NB);
  CK(cudaMemset(cnt,0,maxcells*4)); smem_cas<<<148*4,256>>>(bins_cm,y,cnt,sum,sq,N); CK(cudaMemcpy(hc.data(),cnt,FS*NB*4,cudaMemcpyDeviceToHost));
  for(int W: {4,5,8}){
    size_t sm = (size_t)W*NB*32*20;
    if(sm>227*1024) continue;
    CK(cudaFuncSetAttribute(owner<1>,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)sm));
    CK(cudaFuncSetAttribute(owner<2>,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)sm));
    CK(cudaFuncSetAttribute(owner<4>,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)sm));
    CK(cudaFuncSetAttribute(owner<8>,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)sm));
    for(int mult: {1,4}){
      int nblk=148*mult; long warps=(long)nblk*W; int rpw=(int)((N+warps-1)/warps);
      float ms;
      ms = timeit([&]{ owner<1><<<nblk,W*32,sm>>>(bins_rm,y,idx,cnt,sum,sq,N,rpw); });
      printf("D owner<1> W=%d blocks=%d smem=%zuKB: %.3f ms  %.2f Gupd/s(22 of 32 lanes useful) %.2f Glane-upd/s\n",W,nblk,sm/1024,ms,upd/ms/1e6,(double)N*32/ms/1e6);
      ms = timeit([&]{ owner<2><<<nblk,W*32,sm>>>(bins_rm,y,idx,cnt,sum,sq,N,rpw); });
      printf("D owner<2> W=%d blocks=%d: %.3f ms  %.2f Gupd/s  %.2f Glane-upd/s\n",W,nblk,ms,upd/ms/1e6,(double)N*32/ms/1e6);
      ms = timeit([&]{ owner<4><<<nblk,W*32,sm>>>(bins_rm,y,idx,cnt,sum,sq,N,rpw); });
      printf("D owner<4> W=%d blocks=%d: %.3f ms  %.2f Gupd/s  %.2f Glane-upd/s\n",W,nblk,ms,upd/ms/1e6,(double)N*32/ms/1e6);
      ms = timeit([&]{ owner<8><<<nblk,W*32,sm>>>(bins_rm,y,idx,cnt,sum,sq,N,rpw); });
      printf("D owner<8> W=%d blocks=%d: %.3f ms  %.2f Gupd/s  %.2f Glane-upd/s\n",W,nblk,ms,upd/ms/1e6,(double)N*32/ms/1e6);
    }
    CK(cudaMemset(cnt,0,maxcells*4)); { int nblk=148; long warps=(long)nblk*W; int rpw=(int)((N+warps-1)/warps); owner<4><<<nblk,W*32,sm>>>(bins_rm,y,idx,cnt,sum,sq,N,rpw); }
    CK(cudaMemcpy(hd.data(),cnt,FS*NB*4,cudaMemcpyDeviceToHost));
    int bad=0; for(int k=0;k<FS*NB;k++) if(hc[k]!=hd[k]) bad++;
    printf("D verify W=%d: %d mismatching cells of %d\n",W,bad,FS*NB);
  }
  return 0;
}
