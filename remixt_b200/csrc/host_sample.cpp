// Host side of the batched M-step: the segment samples of cn_model.py:475-480,
//     np.random.choice(num_segments, size=sample_size, replace=False, p=weights),
// drawn for all fits of a grid at once, one host thread per group of fits.
//
// Every fit owns its random stream (a numpy RandomState = MT19937); the draws must consume that stream exactly as
// numpy's legacy RandomState.choice does, or the fit would leave the trajectory of the reference after its first
// M-step.  numpy freezes the legacy algorithms (stream compatibility of RandomState is guaranteed), so they are
// restated here:
//   * p given:  repeat { x = random_sample(k); zero p at what was found; cdf = cumsum(p) / cdf[-1];
//                        new = searchsorted(cdf, x, 'right'); keep first occurrences in draw order } until `size` found
//               (numpy/random/mtrand.pyx, RandomState.choice, replace=False branch)
//   * p absent: permutation(n)[:size], i.e. arange(n) shuffled by Fisher-Yates from the top with
//               j = random_interval(i) (masked rejection on 32-bit outputs)
//   * random_sample: (a >> 5, b >> 6) of two 32-bit outputs -> (a * 2^26 + b) / 2^53
// tests/test_mstep_cpu.py checks these against numpy itself, draw by draw and state by state.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/remixt_b200.h"

namespace {

struct MT {
    uint32_t* key;   // 624 words
    int pos;
    void gen() {
        const int N = 624, M = 397;
        const uint32_t MATRIX_A = 0x9908b0dfU, UPPER = 0x80000000U, LOWER = 0x7fffffffU;
        int kk;
        uint32_t y;
        for (kk = 0; kk < N - M; kk++) {
            y = (key[kk] & UPPER) | (key[kk + 1] & LOWER);
            key[kk] = key[kk + M] ^ (y >> 1) ^ (-(int32_t)(y & 1) & MATRIX_A);
        }
        for (; kk < N - 1; kk++) {
            y = (key[kk] & UPPER) | (key[kk + 1] & LOWER);
            key[kk] = key[kk + (M - N)] ^ (y >> 1) ^ (-(int32_t)(y & 1) & MATRIX_A);
        }
        y = (key[N - 1] & UPPER) | (key[0] & LOWER);
        key[N - 1] = key[M - 1] ^ (y >> 1) ^ (-(int32_t)(y & 1) & MATRIX_A);
        pos = 0;
    }
    uint32_t next32() {
        if (pos == 624) gen();
        uint32_t y = key[pos++];
        y ^= (y >> 11);
        y ^= (y << 7) & 0x9d2c5680U;
        y ^= (y << 15) & 0xefc60000U;
        y ^= (y >> 18);
        return y;
    }
    double next_double() {
        int32_t a = next32() >> 5, b = next32() >> 6;
        return (a * 67108864.0 + b) / 9007199254740992.0;
    }
    uint64_t interval(uint64_t max) {
        if (max == 0) return 0;
        uint64_t mask = max, value;
        mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16; mask |= mask >> 32;
        if (max <= 0xffffffffULL) {
            while ((value = (next32() & mask)) > max) {}
        } else {
            while ((value = ((((uint64_t)next32()) << 32 | next32()) & mask)) > max) {}
        }
        return value;
    }
};

// status: 0 ok, 1 "probabilities do not sum to 1", 2 "probabilities are not non-negative", 3 "probabilities contain NaN",
// 4 "Fewer non-zero entries in p than size", 5 "Cannot take a larger sample than population when 'replace=False'"
int choice_one(MT& mt, int n, int size, const double* p_in, int64_t* out, std::vector<double>& p, std::vector<double>& cdf,
               std::vector<int64_t>& perm, std::vector<double>& x) {
    if (size > n) return 5;
    if (!p_in) {
        perm.resize(n);
        for (int i = 0; i < n; ++i) perm[i] = i;
        for (int i = n - 1; i >= 1; --i) {
            int64_t j = (int64_t)mt.interval((uint64_t)i);
            std::swap(perm[i], perm[j]);
        }
        for (int k = 0; k < size; ++k) out[k] = perm[k];
        return 0;
    }
    // validation of p as in mtrand.pyx (Kahan sum against sqrt(eps))
    {
        double sum = p_in[0], c = 0.;
        bool neg = p_in[0] < 0., nan = std::isnan(p_in[0]);
        for (int i = 1; i < n; ++i) {
            double y = p_in[i] - c, t = sum + y;
            c = (t - sum) - y;
            sum = t;
            neg |= p_in[i] < 0.;
            nan |= std::isnan(p_in[i]);
        }
        if (nan) return 3;
        if (neg) return 2;
        if (std::fabs(sum - 1.) > 1.4901161193847656e-08) return 1;
        int nz = 0;
        for (int i = 0; i < n; ++i) nz += p_in[i] > 0.;
        if (nz < size) return 4;
    }
    p.assign(p_in, p_in + n);
    cdf.resize(n);
    x.resize(size);
    std::vector<int64_t> cand(size);
    int found = 0;
    while (found < size) {
        const int k = size - found;
        for (int i = 0; i < k; ++i) x[i] = mt.next_double();
        for (int i = 0; i < found; ++i) p[out[i]] = 0.;
        double acc = 0.;
        for (int i = 0; i < n; ++i) { acc += p[i]; cdf[i] = acc; }
        const double last = cdf[n - 1];
        for (int i = 0; i < n; ++i) cdf[i] /= last;
        for (int i = 0; i < k; ++i) cand[i] = std::upper_bound(cdf.begin(), cdf.end(), x[i]) - cdf.begin();
        // first occurrences, in draw order
        for (int i = 0; i < k; ++i) {
            bool dup = false;
            for (int j = 0; j < i && !dup; ++j) dup = cand[j] == cand[i];
            if (!dup) out[found++] = cand[i];
        }
    }
    return 0;
}

}  // namespace

extern "C" int rmx_host_sample_segments(int num_fits, int num_draws, int n, int size, uint32_t* mt_key, int32_t* mt_pos,
                                        const double* weights, const uint8_t* weighted, int64_t* chosen,
                                        int32_t* status, int num_threads) {
    if (num_fits < 0 || num_draws < 0 || n <= 0 || size < 0 || !mt_key || !mt_pos || !chosen || !status) return RMX_ERR_ARG;
    if (num_threads < 1) num_threads = 1;
    if (num_threads > num_fits) num_threads = num_fits > 0 ? num_fits : 1;
    auto work = [&](int t) {
        std::vector<double> p, cdf, x;
        std::vector<int64_t> perm;
        for (int g = t; g < num_fits; g += num_threads) {
            MT mt{mt_key + (size_t)g * 624, mt_pos[g]};
            bool dead = false;
            for (int d = 0; d < num_draws; ++d) {
                const size_t e = (size_t)d * num_fits + g;
                if (dead) { status[e] = -1; continue; }        // never reached: an earlier draw of this fit raised
                const double* w = (weighted && weighted[e] && weights) ? weights + e * (size_t)n : nullptr;
                status[e] = choice_one(mt, n, size, w, chosen + e * (size_t)size, p, cdf, perm, x);
                dead = status[e] != 0;
            }
            mt_pos[g] = mt.pos;
        }
    };
    if (num_threads == 1) {
        work(0);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < num_threads; ++t) pool.emplace_back(work, t);
        for (auto& th : pool) th.join();
    }
    return RMX_OK;
}
