// lgamma for positive finite arguments: Stirling series above 8, upward recurrence below.
#pragma once
#include <math.h>
#ifdef __CUDACC__
#define LG_HD __host__ __device__ __forceinline__
#else
#define LG_HD inline
#endif
LG_HD double lgamma_pos(double z) {
    double lp = 0.;
    if (z < 8.) {
        double p = z;
        z += 1.;
        while (z < 8.) { p *= z; z += 1.; }
        lp = log(p);
    }
    const double r = 1. / z, r2 = r * r;
    double s = 6.4102564102564102564e-3;                 //  1/156
    s = fma(s, r2, -1.9175269175269175269e-3);           // -691/360360
    s = fma(s, r2, 8.4175084175084175084e-4);            //  1/1188
    s = fma(s, r2, -5.9523809523809523810e-4);           // -1/1680
    s = fma(s, r2, 7.9365079365079365079e-4);            //  1/1260
    s = fma(s, r2, -2.7777777777777777778e-3);           // -1/360
    s = fma(s, r2, 8.3333333333333333333e-2);            //  1/12
    return (z - 0.5) * log(z) - z + 0.91893853320467274178 + s * r - lp;
}
