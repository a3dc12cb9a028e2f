// Probe: DMMA shapes on sm_100a — fragment layout check + throughput. Scratch, not product.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA %s at %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)

__device__ __forceinline__ void mma884(double &c0,double &c1,double a,double b){
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n":"+d"(c0),"+d"(c1):"d"(a),"d"(b));
}
__device__ __forceinline__ void mma1684(double *c,const double*a,double b){
  asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n":"+d"(c[0]),"+d"(c[1]),"+d"(c[2]),"+d"(c[3]):"d"(a[0]),"d"(a[1]),"d"(b));
}
__device__ __forceinline__ void mma1688(double *c,const double*a,const double* b){
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n":"+d"(c[0]),"+d"(c[1]),"+d"(c[2]),"+d"(c[3]):"d"(a[0]),"d"(a[1]),"d"(a[2]),"d"(a[3]),"d"(b[0]),"d"(b[1]));
}
__device__ __forceinline__ void mma16816(double *c,const double*a,const double* b){
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n":"+d"(c[0]),"+d"(c[1]),"+d"(c[2]),"+d"(c[3]):"d"(a[0]),"d"(a[1]),"d"(a[2]),"d"(a[3]),"d"(a[4]),"d"(a[5]),"d"(a[6]),"d"(a[7]),"d"(b[0]),"d"(b[1]),"d"(b[2]),"d"(b[3]));
}

// layout check: A is 16xK row-major, B is Kx8 (B[k][n]), C 16x8
template<int K> __global__ void layout_check(const double*A,const double*B,double*C){
  int lane=threadIdx.x, gid=lane>>2, tig=lane&3;
  double c[4]={0,0,0,0};
  if(K==4){ double a[2]={A[gid*K+tig],A[(gid+8)*K+tig]}; double b=B[tig*8+gid]; mma1684(c,a,b);}
  if(K==8){ double a[4]={A[gid*K+tig],A[(gid+8)*K+tig],A[gid*K+tig+4],A[(gid+8)*K+tig+4]}; double b[2]={B[tig*8+gid],B[(tig+4)*8+gid]}; mma1688(c,a,b);}
  if(K==16){ double a[8]; for(int i=0;i<8;i++) a[i]=A[(gid+8*(i&1))*K+tig+4*(i>>1)]; double b[4]; for(int i=0;i<4;i++) b[i]=B[(tig+4*i)*8+gid]; mma16816(c,a,b);}
  C[gid*8+tig*2]=c[0]; C[gid*8+tig*2+1]=c[1]; C[(gid+8)*8+tig*2]=c[2]; C[(gid+8)*8+tig*2+1]=c[3];
}
__global__ void layout_check884(const double*A,const double*B,double*C){ // A 8x4, B 4x8
  int lane=threadIdx.x, gid=lane>>2, tig=lane&3; double c0=0,c1=0; mma884(c0,c1,A[gid*4+tig],B[tig*8+gid]);
  C[gid*8+tig*2]=c0; C[gid*8+tig*2+1]=c1;
}

template<int SHAPE> __global__ void __launch_bounds__(256) tput(double*out,int iters){
  double a[8],b[4]; for(int i=0;i<8;i++) a[i]=1.0+1e-9*threadIdx.x+i; for(int i=0;i<4;i++) b[i]=1e-9*(i+1);
  double c[8][4]; for(int t=0;t<8;t++) for(int i=0;i<4;i++) c[t][i]=0;
  for(int it=0;it<iters;it++){
    #pragma unroll
    for(int t=0;t<8;t++){
      if(SHAPE==0) { mma884(c[t][0],c[t][1],a[0],b[0]); mma884(c[t][2],c[t][3],a[1],b[1]); }
      if(SHAPE==1) mma1684(c[t],a,b[0]);
      if(SHAPE==2) mma1688(c[t],a,b);
      if(SHAPE==3) mma16816(c[t],a,b);
      if(SHAPE==4) { // plain DFMA: 32 fma per thread per t
        #pragma unroll
        for(int q=0;q<8;q++){ c[t][0]=fma(a[q],b[0],c[t][0]); c[t][1]=fma(a[q],b[1],c[t][1]); c[t][2]=fma(a[q],b[2],c[t][2]); c[t][3]=fma(a[q],b[3],c[t][3]); }
      }
    }
  }
  double s=0; for(int t=0;t<8;t++) for(int i=0;i<4;i++) s+=c[t][i];
  out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}

template<int K> void check(){
  double hA[16*16],hB[16*8],hC[128],ref[128];
  for(int i=0;i<16*K;i++) hA[i]=drand48()-0.5; for(int i=0;i<K*8;i++) hB[i]=drand48()-0.5;
  for(int r=0;r<16;r++) for(int n=0;n<8;n++){ double s=0; for(int k=0;k<K;k++) s+=hA[r*K+k]*hB[k*8+n]; ref[r*8+n]=s; }
  double *dA,*dB,*dC; CK(cudaMalloc(&dA,sizeof hA)); CK(cudaMalloc(&dB,sizeof hB)); CK(cudaMalloc(&dC,sizeof hC));
  CK(cudaMemcpy(dA,hA,sizeof hA,cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB,hB,sizeof hB,cudaMemcpyHostToDevice));
  layout_check<K><<<1,32>>>(dA,dB,dC); CK(cudaDeviceSynchronize()); CK(cudaMemcpy(hC,dC,sizeof hC,cudaMemcpyDeviceToHost));
  double e=0; for(int i=0;i<128;i++) e=fmax(e,fabs(hC[i]-ref[i])); printf("layout m16n8k%d maxerr %.3e\n",K,e);
}
int main(){
  check<4>(); check<8>(); check<16>();
  { double hA[32],hB[32],hC[64],ref[64]; for(int i=0;i<32;i++){hA[i]=drand48()-0.5;hB[i]=drand48()-0.5;}
    for(int r=0;r<8;r++)for(int n=0;n<8;n++){double s=0;for(int k=0;k<4;k++)s+=hA[r*4+k]*hB[k*8+n];ref[r*8+n]=s;}
    double*dA,*dB,*dC; CK(cudaMalloc(&dA,256));CK(cudaMalloc(&dB,256));CK(cudaMalloc(&dC,512));
    CK(cudaMemcpy(dA,hA,256,cudaMemcpyHostToDevice));CK(cudaMemcpy(dB,hB,256,cudaMemcpyHostToDevice));
    layout_check884<<<1,32>>>(dA,dB,dC);CK(cudaDeviceSynchronize());CK(cudaMemcpy(hC,dC,512,cudaMemcpyDeviceToHost));
    double e=0;for(int i=0;i<64;i++)e=fmax(e,fabs(hC[i]-ref[i]));printf("layout m8n8k4 maxerr %.3e\n",e);}
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p,0)); printf("%s SMs=%d\n",p.name,p.multiProcessorCount);
  double*out; int nb=p.multiProcessorCount*4; CK(cudaMalloc(&out,nb*256*8));
  const char* names[5]={"m8n8k4 x2","m16n8k4","m16n8k8","m16n8k16","DFMA"};
  double flop_per_t[5]={2*2.0*8*8*4, 2.0*16*8*4, 2.0*16*8*8, 2.0*16*8*16, 32*2.0*32};
  for(int sh=0;sh<5;sh++){
    int iters=20000; cudaEvent_t e0,e1; cudaEventCreate(&e0);cudaEventCreate(&e1);
    for(int rep=0;rep<2;rep++){
      cudaEventRecord(e0);
      if(sh==0) tput<0><<<nb,256>>>(out,iters); if(sh==1) tput<1><<<nb,256>>>(out,iters); if(sh==2) tput<2><<<nb,256>>>(out,iters);
      if(sh==3) tput<3><<<nb,256>>>(out,iters); if(sh==4) tput<4><<<nb,256>>>(out,iters);
      cudaEventRecord(e1); CK(cudaDeviceSynchronize());
    }
    float ms; cudaEventElapsedTime(&ms,e0,e1);
    double warps=nb*8.0; double fl=warps*iters*8.0*flop_per_t[sh]; if(sh==4) fl=nb*256.0*iters*8.0*64;
    printf("%-10s %.3f ms  %.2f TFLOP/s\n",names[sh],ms,fl/ms*1e-9);
  }
  return 0;
}
