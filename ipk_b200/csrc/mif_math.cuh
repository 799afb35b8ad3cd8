// mif_math.cuh -- the two fp64 transcendentals of the MIF0 filter, written for throughput.  sm_100a (host-compilable for tests).
//
// The reference evaluates, per database entry (ipk/src/filter.cpp:20-23,55-58,103-105):
//     x = float(min(pow(10, log_score), 1));   w = double(x) / score_sum;   term = -w * log2(w)
// libdevice's exp10 / log2 carry full special-case handling and cost ~700 SASS instructions per entry together
// (profiles/r01e: k4_mif0 274 thread instructions per entry, fp64 pipe 36 % busy).  The inputs here are confined --
// log_score is a float in (-45, 0], w a normal double in (0, 1] -- so both reduce to one table look-up or one rounding
// plus a short Horner polynomial:
//   mif_exp10 : 10^s = 2^n * exp(r ln 2),  n = rint(s log2 10),  r from a two-part product (exact first part), degree 11.
//               |relative error| < 1e-14, i.e. the rounding to float agrees with a correctly rounded pow in all but
//               ~1e-7 of the cases (those differ by one float ulp).
//   mif_log2  : w = 2^e m,  m in [1,2) bucketed by its top 7 mantissa bits,  r = m / center - 1 (one fma, |r| < 2^-8),
//               log2(w) = (e + log2 center) + log1p(r) / ln 2 (degree 6).  The upper half of the buckets is folded onto
//               e + 1 so that nothing cancels near w = 1.  |absolute error| < 4e-16 (tests/test_mif_math.py).
// Both are far inside the tolerance that libm-vs-libdevice differences already impose on filter values (tests/parity.py).
#pragma once
#include <cstdint>
#include <cmath>
#include <cstring>

#if defined(__CUDACC__)
#define IPK_HD __host__ __device__ __forceinline__
#else
#define IPK_HD inline
#endif

namespace ipkgpu
{
    constexpr int MIF_LOG2_BUCKETS = 128;
    constexpr int MIF_LOG2_FOLD = 53;          // first bucket whose centre exceeds sqrt(2)

    struct mif_log2_entry { double rcp_center, log2_center; };

    /// host: the bucket table of mif_log2 (uploaded once per context)
    inline void mif_log2_table(mif_log2_entry* tab)
    {
        for (int i = 0; i < MIF_LOG2_BUCKETS; ++i)
        {
            const double center = 1.0 + (double(i) + 0.5) / MIF_LOG2_BUCKETS;
            tab[i].rcp_center = 1.0 / center;
            // the value paired with rcp_center must be the log of the centre the ROUNDED reciprocal stands for
            tab[i].log2_center = -std::log2(tab[i].rcp_center) - (i >= MIF_LOG2_FOLD ? 1.0 : 0.0);
        }
    }

    IPK_HD double mif_bits_to_double(uint64_t b)
    {
#if defined(__CUDA_ARCH__)
        return __longlong_as_double((long long)b);
#else
        double d; memcpy(&d, &b, 8); return d;
#endif
    }
    IPK_HD uint64_t mif_double_to_bits(double d)
    {
#if defined(__CUDA_ARCH__)
        return (uint64_t)__double_as_longlong(d);
#else
        uint64_t b; memcpy(&b, &d, 8); return b;
#endif
    }

    /// log2 of a positive NORMAL double
    IPK_HD double mif_log2(double w, const mif_log2_entry* tab)
    {
        const uint64_t b = mif_double_to_bits(w);
        int e = int(b >> 52) - 1023;
        const int i = int(b >> 45) & (MIF_LOG2_BUCKETS - 1);
        const double m = mif_bits_to_double((b & 0x000fffffffffffffull) | 0x3ff0000000000000ull);
        const mif_log2_entry t = tab[i];
        e += i >= MIF_LOG2_FOLD ? 1 : 0;
        const double r = fma(m, t.rcp_center, -1.0);
        double p = -1.0 / 6.0;
        p = fma(p, r, 1.0 / 5.0);
        p = fma(p, r, -1.0 / 4.0);
        p = fma(p, r, 1.0 / 3.0);
        p = fma(p, r, -1.0 / 2.0);
        p = fma(p, r, 1.0);
        return (double(e) + t.log2_center) + (r * 1.4426950408889634074) * p;
    }

    /// 10^s for a float s in (-300, 0]
    IPK_HD double mif_exp10(float sf)
    {
        const double s = double(sf);
        const double L = 3.3219280948873623479;                 // log2(10)
        const double L_hi = 3.3219280838966369629;              // its leading 28 bits: float * L_hi is exact in double
        const double L_lo = L - L_hi;
        const double n = rint(s * L);
        const double r = fma(s, L_hi, -n) + s * L_lo;           // in [-0.5, 0.5] (+ a rounding)
        const double y = r * 0.69314718055994530942;            // r ln 2
        double p = 1.0 / 39916800.0;                            // 1/11!
        p = fma(p, y, 1.0 / 3628800.0);
        p = fma(p, y, 1.0 / 362880.0);
        p = fma(p, y, 1.0 / 40320.0);
        p = fma(p, y, 1.0 / 5040.0);
        p = fma(p, y, 1.0 / 720.0);
        p = fma(p, y, 1.0 / 120.0);
        p = fma(p, y, 1.0 / 24.0);
        p = fma(p, y, 1.0 / 6.0);
        p = fma(p, y, 0.5);
        p = fma(p, y, 1.0);
        p = fma(p, y, 1.0);
        const uint64_t scale = uint64_t(int64_t(n) + 1023) << 52;                // 2^n, n >= -1022
        return p * mif_bits_to_double(scale);
    }
}
