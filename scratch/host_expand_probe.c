// How fast can the host cores replicate a 128 KB block into a 16 GB destination? (decides whether
// a compact D2H + host-side expansion can beat the 50 GB/s PCIe link for the cfg5 Jacobian)
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <omp.h>
#include <immintrin.h>
static void neg_stream(double* dst, const double* src, size_t n){
#ifdef __AVX512F__
  const __m512d sgn=_mm512_set1_pd(-0.0);
  for(size_t i=0;i<n;i+=8){ __m512d v=_mm512_loadu_pd(src+i); _mm512_stream_pd(dst+i,_mm512_xor_pd(v,sgn)); }
#else
  for(size_t i=0;i<n;i++) dst[i]=-src[i];
#endif
}
int main(){
  size_t blk=128*128, copies=64, units=256; size_t total=blk*copies*units; // 2 GB per pass... repeat
  double* dst=(double*)aligned_alloc(4096,total*8); double* src=(double*)aligned_alloc(4096,blk*units*8);
  #pragma omp parallel for
  for(size_t i=0;i<total;i++) dst[i]=0;
  for(size_t i=0;i<blk*units;i++) src[i]=i;
  for(int mode=0;mode<2;mode++) for(int rep=0;rep<3;rep++){
    double t=omp_get_wtime();
    #pragma omp parallel for schedule(static)
    for(size_t job=0;job<units*copies;job++){ size_t u=job/copies; double* d=dst+job*blk; const double* s=src+u*blk;
      if(mode==0){ for(size_t i=0;i<blk;i++) d[i]=-s[i]; } else neg_stream(d,s,blk); }
    t=omp_get_wtime()-t; printf("mode %d threads %d: %.1f GB/s written\n",mode,omp_get_max_threads(),total*8/t/1e9);
  }
  return (int)dst[123];
}
