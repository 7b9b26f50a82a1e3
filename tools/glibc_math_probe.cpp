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
static inline uint32_t abstop12(float x) { return (asuint(x) >> 20) & 0x7ff; }
// the device's sincos_fast (spk_glibc_math.cuh), statement for statement: branch-free, immediate coefficients, the cosine
// polynomial of __sincosf_table[1] as the negated one of table [0]
static inline void sincos_fast(float y, float &sin_out, float &cos_out)
{
    const double x0 = (double)y;
    const double r = x0 * 0x1.45f306dc9c883p+23;
    const int n = ((int32_t)r + 0x800000) >> 24;
    const double x = mad(-(double)n, 0x1.921fb54442d18p+0, x0);
    const double xs = ((n + 1) & 2) ? -x : x;
    const double x2 = x * x;
    const double x3 = xs * x2;
    const double s1 = mad(x2, -0x1.994eb3774cf24p-13, 0x1.1107605230bc4p-7);
    const double x5 = x3 * x2;
    const double s = mad(x3, -0x1.555545995a603p-3, xs);
    const float S = (float)mad(x5, s1, s);
    const double x4 = x2 * x2;
    const double c2 = mad(x2, 0x1.99343027bf8c3p-16, -0x1.6c087e89a359dp-10);
    const double c1 = mad(x2, -0x1.ffffffd0c621cp-2, 1.0);
    const double x6 = x4 * x2;
    const double c = mad(x4, 0x1.55553e1068f19p-5, c1);
    float C = (float)mad(x6, c2, c);
    C = (n & 2) ? -C : C;
    const bool tiny = abstop12(y) < abstop12(0x1p-12f);
    sin_out = tiny ? y : ((n & 1) ? C : S);
    cos_out = tiny ? 1.0f : ((n & 1) ? S : C);
}
float my_sinf(float y) { float s, c; sincos_fast(y, s, c); return s; }
float my_cosf(float y) { float s, c; sincos_fast(y, s, c); return c; }
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
int main(int argc, char **argv)
{
    // the immediates above are __sincosf_table[0]; table [1] is it with the cosine coefficients negated
    const double imm[8] = {1.0, -0x1.ffffffd0c621cp-2, -0x1.555545995a603p-3, 0x1.55553e1068f19p-5, 0x1.1107605230bc4p-7,
                           -0x1.6c087e89a359dp-10, -0x1.994eb3774cf24p-13, 0x1.99343027bf8c3p-16};
    for (int k = 0; k < 8; k++) {
        const bool cosine = k == 0 || k == 1 || k == 3 || k == 5 || k == 7;
        if (kSinCos0[6 + k] != imm[k] || kSinCos1[6 + k] != (cosine ? -imm[k] : imm[k])) { printf("table mismatch at %d\n", k); return 2; }
    }
    uint64_t st = 88172645463325252ull;
    auto rnd = [&]() { st ^= st << 13; st ^= st >> 7; st ^= st << 17; return (double)(st >> 11) * (1.0 / 9007199254740992.0); };
    long N = argc > 1 ? atol(argv[1]) : 50000000, bad_s = 0, bad_c = 0, bad_p1 = 0, bad_p2 = 0, bad_s2 = 0;
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
    return (bad_s + bad_c + bad_s2 + bad_p1 + bad_p2) ? 1 : 0;
}
