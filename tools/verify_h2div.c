// Exhaustive CPU check of quant8_h2 (mxmoe_b200/csrc/quant_common.cuh): gcc -O2 -o verify_h2div tools/verify_h2div.c -lm;
// ./verify_h2div 0 31744  -> "bad_code=0" (run in 8 ranges of 0xf80 scales in parallel: ~2 min).  fp16 arithmetic is emulated in
// double (products of two fp16 numbers and the sums used here are exact in double; one rounding to fp16 per operation).
// exhaustive check of the NORMALISED fp16 Markstein division:
//   k: s' = s*2^k in [1,2); x' = RN16(x*2^k); r = RN16(RN32(1/s')); q0 = RN16(x'*r); e = RN16(x' - q0*s'); q1 = RN16(q0 + e*r)
//   code = rint(clamp(q1, +-up))  vs  rint(clamp(RN16(RN32(x/s)), +-up))  for up = 127 and 7
#include <stdio.h>
#include <stdint.h>
#include <math.h>
#include <stdlib.h>
static double h2d(uint16_t h){ int s=h>>15, e=(h>>10)&31, m=h&1023; double v; if(e==0) v=ldexp((double)m,-24); else if(e==31) v=m?NAN:INFINITY; else v=ldexp((double)(m+1024),e-25); return s?-v:v; }
static uint16_t d2h(double d){ if(d!=d) return 0x7e00; uint16_t sg = signbit(d)?0x8000:0; double a=fabs(d); if(a==0) return sg;
  if(a>=65520.0) return sg|0x7c00; int e; frexp(a,&e); int E=e-1;
  if(E<-14){ double q=a*16777216.0; double r=nearbyint(q); return sg|(uint16_t)r; }
  double q=ldexp(a, 10-E); double r=nearbyint(q); if(r>=2048.0){ r=1024.0; E++; if(E>15) return sg|0x7c00; }
  return sg | (uint16_t)(((E+15)<<10) + ((int)r-1024)); }
static uint16_t fma16(uint16_t a,uint16_t b,uint16_t c){ return d2h(fma(h2d(a),h2d(b),h2d(c))); }
static uint16_t mul16(uint16_t a,uint16_t b){ return d2h(h2d(a)*h2d(b)); }
int main(int argc,char**argv){
  int lo=atoi(argv[1]), hi=atoi(argv[2]);
  long long total=0,bad_code=0;
  for(int si=lo; si<hi; ++si){ uint16_t s=(uint16_t)si; int Es=(s>>10)&31; if(Es==0||Es==31) continue; // normal s only
    uint16_t fac=(uint16_t)((30-Es)<<10); if(30-Es==0) fac=0x0200; // 2^(15-Es) as fp16 (2^-15 is the denormal 0x0200)
    uint16_t sp=(uint16_t)((15<<10)|(s&1023)); double sd=h2d(s); float sf=(float)sd;
    float r32=1.0f/(float)h2d(sp); uint16_t r=d2h((double)r32); long long bc=0;
    for(int sgn=0; sgn<2; ++sgn) for(int xi=0; xi<0x7c00; ++xi){ uint16_t x=(uint16_t)(xi|(sgn<<15)); float xf=(float)h2d(x); float q32=xf/sf; if(fabsf(q32)>160.f) break;
      uint16_t qref=d2h((double)q32); uint16_t xp=mul16(x,fac); uint16_t q0=mul16(xp,r); uint16_t e=fma16(q0^0x8000,sp,xp); uint16_t q1=fma16(e,r,q0);
      total++;
      double a=h2d(q1), b=h2d(qref); double a8=fmin(fmax(a,-127),127), b8=fmin(fmax(b,-127),127); if(nearbyint(a8)!=nearbyint(b8)) bc++;
      double a4=fmin(fmax(a,-7),7), b4=fmin(fmax(b,-7),7); if(nearbyint(a4)!=nearbyint(b4)) bc++;
    }
    if(bc) printf("s=0x%04x mant=0x%03x bad_code=%lld\n", s, s&1023, bc);
    bad_code+=bc; }
  fprintf(stderr,"range %x-%x total=%lld bad_code=%lld\n",lo,hi,total,bad_code); return 0; }
