#include <cstdio>
#include <cmath>
#include <random>
#include "lgam.h"
int main() {
    std::mt19937_64 rng(1);
    double worst_abs = 0, worst_rel = 0, wa = 0, wr = 0;
    for (int i = 0; i < 20000000; ++i) {
        double u = std::uniform_real_distribution<double>(-7, 12)(rng);
        double z = exp(u);   // 1e-3 .. 1.6e5
        double a = lgamma_pos(z), b = lgamma(z);
        double e = fabs(a - b), r = e / (fabs(b) + 1e-300);
        if (e > worst_abs) { worst_abs = e; wa = z; }
        double scaled = e / fmax(1.0, fabs(b));
        if (scaled > worst_rel) { worst_rel = scaled; wr = z; }
    }
    printf("worst abs err %.3e at z=%g ; worst err/max(1,|lg|) %.3e at z=%g\n", worst_abs, wa, worst_rel, wr);
    return 0;
}
