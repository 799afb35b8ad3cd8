#include "mif_math.cuh"
#include <cstdio>
#include <random>
using namespace ipkgpu;
int main(){
  mif_log2_entry tab[128]; mif_log2_table(tab);
  std::mt19937_64 g(1); std::uniform_real_distribution<double> U(0,1);
  double maxabs=0, maxrel=0; 
  for(long i=0;i<3000000;i++){ double w; int mode=i%4;
    if(mode==0) w=U(g); else if(mode==1) w=pow(10,-8*U(g)); else if(mode==2) w=1-1e-6*U(g); else w=pow(2,-40*U(g));
    if(w<=0) continue; double a=mif_log2(w,tab), b=log2l((long double)w); double d=fabs(a-b); if(d>maxabs) maxabs=d; if(b!=0){double r=d/fabs(b); if(r>maxrel&&fabs(b)>1e-3) maxrel=r;} }
  printf("log2: max abs err %.3e, max rel err (|log2|>1e-3) %.3e\n",maxabs,maxrel);
  double mr=0; long mism=0,N=3000000;
  for(long i=0;i<N;i++){ float s=-(float)(12*U(g)); double a=mif_exp10(s); long double b=powl(10.0L,(long double)s); double r=fabs((double)((a-b)/b)); if(r>mr) mr=r; if((float)a!=(float)(double)b) mism++; }
  printf("exp10: max rel err %.3e, float mismatches %ld of %ld\n",mr,mism,N);
  for(float s: {-40.f,-100.f,-299.f,0.f,-1e-30f}){ printf("  s=%g  %.17g vs %.17g\n", s, mif_exp10(s), pow(10.0,(double)s)); }
}
