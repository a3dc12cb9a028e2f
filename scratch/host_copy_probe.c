// Variants of the host-side expansion inner loop: dst = -src for 128 KB blocks replicated 64 times
// (the x-block of a cfg5 Jacobian unit), all cores, streaming stores.  Decides the loop shape of
// copy_run_avx512 in qcb200.cu.   gcc -O3 -march=native -fopenmp host_copy_probe.c -o host_copy_probe
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <omp.h>
#include <immintrin.h>
static void v0(double* d, const double* s, size_t n){ const __m512d g=_mm512_set1_pd(-0.0);
  for(size_t i=0;i<n;i+=8) _mm512_stream_pd(d+i,_mm512_xor_pd(_mm512_loadu_pd(s+i),g)); }
static void v1(double* d, const double* s, size_t n){ const __m512d g=_mm512_set1_pd(-0.0);
  for(size_t i=0;i<n;i+=32){ __m512d a=_mm512_loadu_pd(s+i),b=_mm512_loadu_pd(s+i+8),c=_mm512_loadu_pd(s+i+16),e=_mm512_loadu_pd(s+i+24);
    _mm512_stream_pd(d+i,_mm512_xor_pd(a,g)); _mm512_stream_pd(d+i+8,_mm512_xor_pd(b,g)); _mm512_stream_pd(d+i+16,_mm512_xor_pd(c,g)); _mm512_stream_pd(d+i+24,_mm512_xor_pd(e,g)); } }
static void v2(double* d, const double* s, size_t n){ const __m512d g=_mm512_set1_pd(-0.0);
  for(size_t i=0;i<n;i+=32){ _mm_prefetch((const char*)(s+i+128),_MM_HINT_T0); _mm_prefetch((const char*)(s+i+136),_MM_HINT_T0); _mm_prefetch((const char*)(s+i+144),_MM_HINT_T0); _mm_prefetch((const char*)(s+i+152),_MM_HINT_T0);
    __m512d a=_mm512_loadu_pd(s+i),b=_mm512_loadu_pd(s+i+8),c=_mm512_loadu_pd(s+i+16),e=_mm512_loadu_pd(s+i+24);
    _mm512_stream_pd(d+i,_mm512_xor_pd(a,g)); _mm512_stream_pd(d+i+8,_mm512_xor_pd(b,g)); _mm512_stream_pd(d+i+16,_mm512_xor_pd(c,g)); _mm512_stream_pd(d+i+24,_mm512_xor_pd(e,g)); } }
static void v3(double* d, const double* s, size_t n){ const __m512d g=_mm512_set1_pd(1.0); (void)s;
  for(size_t i=0;i<n;i+=8) _mm512_stream_pd(d+i,g); }
// 256-bit stores (some cores sustain more write-combining buffers' worth with ymm)
static void v4(double* d, const double* s, size_t n){ const __m256d g=_mm256_set1_pd(-0.0);
  for(size_t i=0;i<n;i+=4) _mm256_stream_pd(d+i,_mm256_xor_pd(_mm256_loadu_pd(s+i),g)); }
typedef void (*fn)(double*,const double*,size_t);
int main(int argc,char**argv){
  size_t blk=128*128, copies=64, units=argc>1?atoi(argv[1]):128; size_t total=blk*copies*units;
  double* dst=(double*)aligned_alloc(4096,total*8); double* src=(double*)aligned_alloc(4096,blk*units*8);
  #pragma omp parallel for
  for(size_t i=0;i<total;i++) dst[i]=0;
  for(size_t i=0;i<blk*units;i++) src[i]=i;
  fn f[5]={v0,v1,v2,v3,v4}; const char* nm[5]={"loop1","unroll4","unroll4+prefetch","fill","ymm"};
  for(int sched=0;sched<2;sched++) for(int mode=0;mode<5;mode++){ double best=1e9;
    for(int rep=0;rep<3;rep++){ double t=omp_get_wtime();
      if(sched==0){
        #pragma omp parallel for schedule(static)
        for(size_t job=0;job<units*copies;job++){ size_t u=job/copies; f[mode](dst+job*blk,src+u*blk,blk); }
      } else {   // one unit per work item, as expand_units hands them out
        #pragma omp parallel for schedule(dynamic,1)
        for(size_t u=0;u<units;u++) for(size_t c=0;c<copies;c++) f[mode](dst+(u*copies+c)*blk,src+u*blk,blk);
      }
      t=omp_get_wtime()-t; if(t<best) best=t; }
    printf("%s %-18s threads %d: %.1f GB/s written\n",sched?"unit-items ":"block-items",nm[mode],omp_get_max_threads(),total*8/best/1e9); }
  _mm_sfence();
  return (int)dst[123];
}
