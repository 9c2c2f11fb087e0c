// Checks the fdlibm restatements of acosf and atan2f that csrc/skc_common.cuh carries (skc_acosf, skc_atan2f) against the host libm:
// acosf over its whole domain (2.13 G arguments), atan2f on 4e8 random and structured pairs.  gcc -O2 -ffp-contract=off tools/libm_check.c -lm.
// Result on this image (glibc 2.39): 0 differences for both.
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
static inline int32_t fbits(float f){int32_t i;memcpy(&i,&f,4);return i;}
static inline float bitsf(int32_t i){float f;memcpy(&f,&i,4);return f;}
static float my_acosf(float x){
  const float one=1.0f, pi=3.1415925026e+00f, pio2_hi=1.5707962513e+00f, pio2_lo=7.5497894159e-08f,
  pS0=1.6666667163e-01f,pS1=-3.2556581497e-01f,pS2=2.0121252537e-01f,pS3=-4.0055535734e-02f,pS4=7.9153501429e-04f,pS5=3.4793309169e-05f,
  qS1=-2.4033949375e+00f,qS2=2.0209457874e+00f,qS3=-6.8828397989e-01f,qS4=7.7038154006e-02f;
  float z,p,q,r,w,s,c,df; int32_t hx=fbits(x), ix=hx&0x7fffffff;
  if(ix==0x3f800000){ if(hx>0) return 0.0f; else return pi+(float)2.0*pio2_lo; }
  else if(ix>0x3f800000) return (x-x)/(x-x);
  if(ix<0x3f000000){ if(ix<=0x23000000) return pio2_hi+pio2_lo;
    z=x*x; p=z*(pS0+z*(pS1+z*(pS2+z*(pS3+z*(pS4+z*pS5))))); q=one+z*(qS1+z*(qS2+z*(qS3+z*qS4))); r=p/q; return pio2_hi-(x-(pio2_lo-x*r)); }
  else if(hx<0){ z=(one+x)*(float)0.5; p=z*(pS0+z*(pS1+z*(pS2+z*(pS3+z*(pS4+z*pS5))))); q=one+z*(qS1+z*(qS2+z*(qS3+z*qS4))); s=sqrtf(z); r=p/q; w=r*s-pio2_lo; return pi-(float)2.0*(s+w); }
  else { int32_t idf; z=(one-x)*(float)0.5; s=sqrtf(z); df=s; idf=fbits(df); df=bitsf(idf&0xfffff000); c=(z-df*df)/(s+df);
    p=z*(pS0+z*(pS1+z*(pS2+z*(pS3+z*(pS4+z*pS5))))); q=one+z*(qS1+z*(qS2+z*(qS3+z*qS4))); r=p/q; w=r*s+c; return (float)2.0*(df+w); }
}
static float my_atan2f(float y,float x){
  const float tiny=1.0e-30f, zero=0.0f, pi_o_4=7.8539818525e-01f, pi_o_2=1.5707963705e+00f, pi=3.1415927410e+00f, pi_lo=-8.7422776573e-08f;
  float z; int32_t k,m,hx=fbits(x),hy=fbits(y),ix=hx&0x7fffffff,iy=hy&0x7fffffff;
  if(ix>0x7f800000||iy>0x7f800000) return x+y;
  if(hx==0x3f800000) return atanf(y);
  m=((hy>>31)&1)|((hx>>30)&2);
  if(iy==0){ switch(m){case 0: case 1: return y; case 2: return pi+tiny; case 3: return -pi-tiny;} }
  if(ix==0) return (hy<0)? -pi_o_2-tiny: pi_o_2+tiny;
  if(ix==0x7f800000){ if(iy==0x7f800000){ switch(m){case 0: return pi_o_4+tiny; case 1: return -pi_o_4-tiny; case 2: return (float)3.0*pi_o_4+tiny; case 3: return (float)-3.0*pi_o_4-tiny;} }
    else { switch(m){case 0: return zero; case 1: return -zero; case 2: return pi+tiny; case 3: return -pi-tiny;} } }
  if(iy==0x7f800000) return (hy<0)? -pi_o_2-tiny: pi_o_2+tiny;
  k=(iy-ix)>>23;
  if(k>60) z=pi_o_2+(float)0.5*pi_lo; else if(hx<0&&k<-60) z=0.0f; else z=atanf(fabsf(y/x));
  switch(m){ case 0: return z; case 1: { uint32_t zh=fbits(z); return bitsf(zh^0x80000000);} case 2: return pi-(z-pi_lo); default: return (z-pi_lo)-pi; }
}
int main(int argc,char**argv){
  long bad=0,n=0; 
  for(uint32_t u=0;;u++){ float x=bitsf((int32_t)u); if(!(fabsf(x)<=1.0f)){ if(u==0xffffffffu)break; continue;} float a=acosf(x),b=my_acosf(x); n++; if(fbits(a)!=fbits(b)){ if(bad<5)printf("acosf x=%a lib=%a mine=%a\n",x,a,b); bad++; } if(u==0xffffffffu)break; }
  printf("acosf: %ld args, %ld differ\n",n,bad);
  bad=0;n=0; uint64_t s=88172645463325252ull;
  for(long i=0;i<400000000L;i++){ s^=s<<13; s^=s>>7; s^=s<<17; float y,x; 
    if(i&1){ y=bitsf((int32_t)(s&0xffffffff)); x=bitsf((int32_t)(s>>32)); }
    else { double a=(double)(s&0xffffff)/16777216.0*6.283185307179586; double r=exp(((double)((s>>24)&0xffff)/65536.0-0.5)*20.0); y=(float)(r*sin(a)); x=(float)(r*cos(a)); if((s>>60)==0) x=0.f; if((s>>60)==1) y=0.f; }
    float a=atan2f(y,x),b=my_atan2f(y,x); n++; if(fbits(a)!=fbits(b) && !(a!=a && b!=b)){ if(bad<5)printf("atan2f y=%a x=%a lib=%a mine=%a\n",y,x,a,b); bad++; } }
  printf("atan2f: %ld args, %ld differ\n",n,bad);
  return 0;
}
