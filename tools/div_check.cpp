// Host check behind csrc/crt_device.cuh operator/(V3, float) with CRT_DIV_VIA_DOUBLE: a / s in binary32 equals
// RN32 (RN64 (a * RN64 (1 / s))) for every operand class.   g++ -O2 -ffp-contract=off tools/div_check.cpp -o /tmp/div_check && /tmp/div_check
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <cmath>
static inline uint64_t rng(uint64_t &s){ s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; }
int main(){
  uint64_t st=88172645463325252ull; long bad=0,n=0;
  const uint32_t special[]={0,0x80000000u,1,0x80000001u,0x007fffffu,0x00800000u,0x7f7fffffu,0x7f800000u,0xff800000u,0x7fc00000u,0x3f800000u,0xbf800000u,0x3f7fffffu,0x3f800001u,0x00400000u};
  auto check=[&](uint32_t ua,uint32_t us){
    float a,s; memcpy(&a,&ua,4); memcpy(&s,&us,4);
    volatile float q=a/s; const double r=1.0/double(s); volatile float p=float(double(a)*r);
    float qq=q,pp=p; uint32_t x,y; memcpy(&x,&qq,4); memcpy(&y,&pp,4);
    const bool bothnan = (qq!=qq)&&(pp!=pp);
    if(x!=y && !bothnan){ if(bad<10) printf("a=%08x s=%08x q=%08x p=%08x\n",ua,us,x,y); ++bad;} ++n; };
  for(uint32_t a:special) for(uint32_t s:special) check(a,s);
  for(long i=0;i<300000000;++i){ uint64_t v=rng(st); uint32_t ua=uint32_t(v), us=uint32_t(v>>32);
    if((i&7)==0){ us=(us&0x807fffffu)|(((ua>>23)&0xff)<<23); }   // same exponent: quotients near 1
    if((i&15)==1) ua=special[(v>>20)%15];
    if((i&15)==2) us=special[(v>>24)%15];
    if((i&31)==3){ ua&=0x807fffffu; }   // denormal numerator
    check(ua,us); }
  printf("%ld pairs, %ld mismatches\n",n,bad); return bad!=0; }
