// rp_philox.cuh — counter-based random numbers for the sweep (host + device)
#pragma once

#include "rp_math.cuh"

namespace rp {

// ---------------------------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11) — counter-based, one independent stream per (particle, cycle, pass)
// ---------------------------------------------------------------------------------------------------------------
struct U4 { unsigned int x, y, z, w; };

RP_HD unsigned int mulhi32(unsigned int a, unsigned int b) {
#ifdef __CUDA_ARCH__
    return __umulhi(a, b);
#else
    return static_cast<unsigned int>((static_cast<unsigned long long>(a) * b) >> 32);
#endif
}

RP_HD U4 philox4x32(U4 c, unsigned int k0, unsigned int k1) {
    const unsigned int M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    #pragma unroll
    for (int r = 0; r < 10; r++) {
        unsigned int hi0 = mulhi32(M0, c.x), lo0 = M0 * c.x;
        unsigned int hi1 = mulhi32(M1, c.z), lo1 = M1 * c.z;
        U4 n;
        n.x = hi1 ^ c.y ^ k0; n.y = lo1; n.z = hi0 ^ c.w ^ k1; n.w = lo0;
        c = n;
        k0 += W0; k1 += W1;
    }
    return c;
}

struct Rng {
    U4 ctr;
    unsigned int k0, k1;
    U4 buf;
    int have;    // doubles left in buf (0..2)
    RP_HD void init(unsigned long long seed, unsigned int a, unsigned int b, unsigned int c) {
        k0 = static_cast<unsigned int>(seed); k1 = static_cast<unsigned int>(seed >> 32);
        ctr.x = a; ctr.y = b; ctr.z = c; ctr.w = 0;
        have = 0;
    }
    // uniform in [0, 1) with 53 random bits
    RP_HD double uniform() {
        if (have == 0) { buf = philox4x32(ctr, k0, k1); ctr.w++; have = 2; }
        unsigned int hi = (have == 2) ? buf.x : buf.z, lo = (have == 2) ? buf.y : buf.w;
        have--;
        unsigned long long bits = (static_cast<unsigned long long>(hi) << 32) | lo;
        return static_cast<double>(bits >> 11) * (1.0 / 9007199254740992.0);
    }
    // std::uniform_real_distribution<double>(a, b) semantics: a + (b - a) * u
    RP_HD double uniform(double a, double b) { return a + (b - a) * uniform(); }
};

}  // namespace rp
