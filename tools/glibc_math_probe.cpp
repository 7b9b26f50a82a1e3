// tools/glibc_math_probe.cpp - host prototype of sparkle_b200/csrc/spk_glibc_math.cuh against the libm of the box (test bench):
//   g++ -O2 -ffp-contract=off -mfma -DFMA=1 -o /tmp/probe tools/glibc_math_probe.cpp -lm; /tmp/probe     (0 differences in 5e7 arguments each)
//   -DFMA=0: the same without the contractions of the x86-64 FMA build of glibc (24 and 2 in 5e7 differ)
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../sparkle_b200/csrc/spk_glibc_tables.h"
using namespace spk::glibc;
#ifndef FMA
#define FMA 1
#endif
static inline double mad(double a, double b, double c) { return FMA ? std::fma(a, b, c) : a * b + c; }
static inline uint32_t asuint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float asfloat(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint64_t asuint64(double f) { uint64_t u; memcpy(&u, &f, 8); return u; }
static inline double asdouble(uint64_t u) { double f; memcpy(&f, &u, 8); return f; }
// table layout in the binary: sign[4], hpi_inv, hpi, c0, c1, s1, c2, s2, c3, s3, c4
struct SC { const double *t; double c0() const {return t[6];} double c1() const {return t[7];} double s1() const {return t[8];}
  double c2() const {return t[9];} double s2() const {return t[10];} double c3() const {return t[11];} double s3() const {return t[12];} double c4() const {return t[13];} };
static inline float sinf_poly(double x, double x2, SC p, int n)
{
    if ((n & 1) == 0) {
        double x3 = x * x2;
        double s1 = mad(x2, p.s3(), p.s2());
        double x7 = x3 * x2;
        double s = mad(x3, p.s1(), x);
        return (float)mad(x7, s1, s);
    } else {
        double x4 = x2 * x2;
        double c2 = mad(x2, p.c4(), p.c3());
        double c1 = mad(x2, p.c1(), p.c0());
        double x6 = x4 * x2;
        double c = mad(x4, p.c2(), c1);
        return (float)mad(x6, c2, c);
    }
}
static inline uint32_t abstop12(float x) { return (asuint(x) >> 20) & 0x7ff; }
static inline double reduce_fast(double x, const double *t, int *np)
{
    double r = x * t[4];
    int n = ((int32_t)r + 0x800000) >> 24;
    *np = n;
    return FMA ? std::fma(-(double)n, t[5], x) : x - n * t[5];
}
float my_sinf(float y)
{
    double x = y; int n;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        double s = x * x;
        if (abstop12(y) < abstop12(0x1p-12f)) return y;
        return sinf_poly(x, s, SC{kSinCos0}, 0);
    }
    x = reduce_fast(x, kSinCos0, &n);
    double s = kSinCos0[n & 3];
    SC p{(n & 2) ? kSinCos1 : kSinCos0};
    return sinf_poly(x * s, x * x, p, n);
}
float my_cosf(float y)
{
    double x = y; int n;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        double x2 = x * x;
        if (abstop12(y) < abstop12(0x1p-12f)) return 1.0f;
        return sinf_poly(x, x2, SC{kSinCos0}, 1);
    }
    x = reduce_fast(x, kSinCos0, &n);
    double s = kSinCos0[n & 3];
    SC p{(n & 2) ? kSinCos1 : kSinCos0};
    return sinf_poly(x * s, x * x, p, n ^ 1);
}
float my_powf(float x, float y)
{
    uint32_t ix = asuint(x);
    uint32_t tmp = ix - 0x3f330000u;
    int i = (tmp >> (23 - 4)) % 16;
    uint32_t top = tmp & 0xff800000u;
    uint32_t iz = ix - top;
    int k = (int32_t)top >> 23;
    double invc = kPowLog2Tab[2 * i], logc = kPowLog2Tab[2 * i + 1];
    double z = (double)asfloat(iz);
    double r = mad(z, invc, -1.0);
    double y0 = logc + (double)k;
    const double *A = kPowLog2Poly;
    double r2 = r * r;
    double yy = mad(A[0], r, A[1]);
    double p = mad(A[2], r, A[3]);
    double r4 = r2 * r2;
    double q = mad(A[4], r, y0);
    q = mad(p, r2, q);
    yy = mad(yy, r4, q);
    double ylogx = (double)y * yy;
    double kd = ylogx + kExp2ShiftScaled;
    uint64_t ki = asuint64(kd);
    kd -= kExp2ShiftScaled;
    double rr = ylogx - kd;
    uint64_t t = kExp2Tab[ki % 32];
    t += ki << (52 - 5);
    double s = asdouble(t);
    const double *C = kExp2Poly;
    double zz = mad(C[0], rr, C[1]);
    double rr2 = rr * rr;
    double out = mad(C[2], rr, 1.0);
    out = mad(zz, rr2, out);
    out = out * s;
    return (float)out;
}
int main()
{
    uint64_t st = 88172645463325252ull;
    auto rnd = [&]() { st ^= st << 13; st ^= st >> 7; st ^= st << 17; return (double)(st >> 11) * (1.0 / 9007199254740992.0); };
    long N = 50000000, bad_s = 0, bad_c = 0, bad_p1 = 0, bad_p2 = 0, bad_s2 = 0;
    for (long k = 0; k < N; k++) {
        float th = (float)(rnd() * 6.283185307179586);
        if (my_sinf(th) != sinf(th)) bad_s++;
        if (my_cosf(th) != cosf(th)) bad_c++;
        float w = (float)((rnd() - 0.5) * 200.0);
        if (my_sinf(w) != sinf(w) || my_cosf(w) != cosf(w)) bad_s2++;
        float x = (float)(rnd() * 1.2 + 1e-6);
        if (my_powf(x, 2.4f) != powf(x, 2.4f)) bad_p1++;
        if (my_powf(x, 0.45f) != powf(x, 0.45f)) bad_p2++;
    }
    printf("FMA=%d  N=%ld  sin(0..2pi) %ld  cos(0..2pi) %ld  sincos(+-100) %ld  pow(x,2.4) %ld  pow(x,0.45) %ld differ\n", FMA, N, bad_s, bad_c, bad_s2, bad_p1, bad_p2);
    return 0;
}
