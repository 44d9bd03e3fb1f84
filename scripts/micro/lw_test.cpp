#include <cstring>
#include <cmath>
#include <cstdio>
#include <vector>
#include <random>
struct d2{double x,y;};
#define E0 4
#define NOCT 18
static std::vector<d2> lw;
static double lwlog(double z){ unsigned long long u; memcpy(&u,&z,8); int hi=(int)(u>>32); int idx=(hi>>13)-((1023+E0)<<7); d2 t=lw[idx];
 double r=fma(z,t.x,-1.0); double p=-1.0/6.0; p=fma(p,r,0.2);p=fma(p,r,-0.25);p=fma(p,r,1.0/3.0);p=fma(p,r,-0.5);p=fma(p*r,r,r);return t.y+p;}
static double core(double z,double relerr){ double L=lwlog(z); double r=(1.0/z)*(1+relerr); r=fma(r,fma(-z,r,1.0),r); double r2=r*r;
 double s=8.4175084175084175084e-4; s=fma(s,r2,-5.9523809523809523810e-4); s=fma(s,r2,7.9365079365079365079e-4); s=fma(s,r2,-2.7777777777777777778e-3); s=fma(s,r2,8.3333333333333333333e-2);
 return fma(z-0.5,L,s*r);}
#include <cstring>
int main(){ lw.resize(NOCT*128); for(int o=0;o<NOCT;o++)for(int j=0;j<128;j++){ long double c=ldexpl(1.0L+(j+0.5L)/128.0L,E0+o); double inv=(double)(1.0L/c); lw[o*128+j]={inv,(double)(-logl((long double)inv))};}
 std::mt19937_64 g(1); std::uniform_real_distribution<double> U(log(16.0),log(4.0e6)); double worst=0,worstl=0;
 for(int i=0;i<20000000;i++){ double z=exp(U(g)); if(i<1000) z=16.0+i*1e-3; long double ref=lgammal((long double)z); double got=core(z,(i&1)?1e-6:-1e-6)-z+0.91893853320467274178L;
  double e=fabs((double)(got-ref))/fmax(1.0,fabs((double)ref)); if(e>worst)worst=e; double el=fabs((double)(lwlog(z)-logl(z)))/fabs((double)logl(z)); if(el>worstl)worstl=el;}
 printf("lgamma worst rel %.3g  log worst rel %.3g\n",worst,worstl);}
