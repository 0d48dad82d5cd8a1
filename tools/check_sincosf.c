#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <string.h>
#include <pthread.h>
static const double hpi_inv=0x1.45F306DC9C883p+23, hpi=0x1.921FB54442D18p0;
static const double C0=0x1p0,C1=-0x1.ffffffd0c621cp-2,C2=0x1.55553e1068f19p-5,C3=-0x1.6c087e89a359dp-10,C4=0x1.99343027bf8c3p-16;
static const double S1=-0x1.555545995a603p-3,S2=0x1.1107605230bc4p-7,S3=-0x1.994eb3774cf24p-13;
#ifdef USEFMA
#define FMA(a,b,c) fma(a,b,c)
#else
#define FMA(a,b,c) ((a)*(b)+(c))
#endif
static inline float poly(double x,double x2,int n,int neg){ // neg: table[1] (c coefficients negated)
  if((n&1)==0){ double x3=x*x2; double s1=FMA(x2,S3,S2); double x7=x3*x2; double s=FMA(x3,S1,x); return (float)FMA(x7,s1,s);}
  else { double sg=neg?-1.0:1.0; double x4=x2*x2; double c2=FMA(x2,sg*C4,sg*C3); double c1=FMA(x2,sg*C1,sg*C0); double x6=x4*x2; double c=FMA(x4,sg*C2,c1); return (float)FMA(x6,c2,c);}
}
static inline uint32_t top12(float f){uint32_t u; memcpy(&u,&f,4); return (u>>20)&0x7ff;}
static void mysincos(float y,float*sn,float*cs){
  double x=y;
  if(top12(y)<top12(0x1.921FB6p-1f)){ double x2=x*x; if(top12(y)<top12(0x1p-12f)){*sn=y;*cs=1.0f;return;} *sn=poly(x,x2,0,0); *cs=poly(x,x2,1,0); return;}
  double r=x*hpi_inv; int n=((int32_t)r+0x800000)>>24;
  x=FMA(-(double)n,hpi,x);
  static const double sign[4]={1,-1,-1,1};
  double s=sign[n&3]; int neg=(n&2)!=0;
  *sn=poly(x*s,x*x,n,neg); *cs=poly(x*s,x*x,n^1,neg);
}
typedef struct{uint32_t lo,hi; long bs,bc;}job;
static void* run(void*p){job*j=p; for(uint32_t u=j->lo;u<j->hi;u++){float f; memcpy(&f,&u,4); float s,c; mysincos(f,&s,&c); if(s!=sinf(f))j->bs++; if(c!=cosf(f))j->bc++;} return 0;}
int main(){ float mx=6.2831855f*1.0001f; uint32_t hi; memcpy(&hi,&mx,4); int T=8; pthread_t th[8]; job jb[8];
 for(int t=0;t<T;t++){jb[t].lo=(uint64_t)hi*t/T; jb[t].hi=(uint64_t)hi*(t+1)/T; jb[t].bs=jb[t].bc=0; pthread_create(&th[t],0,run,&jb[t]);}
 long bs=0,bc=0; for(int t=0;t<T;t++){pthread_join(th[t],0); bs+=jb[t].bs; bc+=jb[t].bc;}
 printf("n=%u sin mismatches %ld cos mismatches %ld\n",hi,bs,bc); return 0;}
