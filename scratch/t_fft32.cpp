#include <cmath>
#include <cstdio>
#include <complex>
#define __host__
#define __device__
#define __forceinline__ inline
#include "../fhe_string_b200/csrc/fft32_codelets.cuh"
int main(){
  double r[32],i[32]; std::complex<double> x[32];
  for(int n=0;n<32;n++){ x[n]={sin(n*1.3)+n, cos(n*0.7)*3}; }
  for(int s=0;s<2;s++){
    for(int n=0;n<32;n++){ r[FHS_BR5(n)]=x[n].real(); i[FHS_BR5(n)]=x[n].imag(); }
    if(s==0) fft32_pos(r,i); else fft32_neg(r,i);
    double maxe=0;
    for(int k=0;k<32;k++){ std::complex<long double> acc=0;
      for(int n=0;n<32;n++){ long double a=(s==0?1:-1)*2*M_PIl*n*k/32; acc+=std::complex<long double>(x[n].real(),x[n].imag())*std::complex<long double>(cosl(a),sinl(a)); }
      maxe=fmax(maxe,fabs(r[k]-(double)acc.real())+fabs(i[k]-(double)acc.imag())); }
    printf("sign %d maxerr %g\n", s==0?1:-1, maxe);
  }
}
