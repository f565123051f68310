/* util.cpp — error convention and the UNI random number generator of the host layer. */
#include <cfloat>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "lvb_host.h"

namespace lvbhost
{

static bool g_crash_throws = false;

void set_crash_throws(bool on) { g_crash_throws = on; }

/* src/Error.c:49-62: message to STDOUT behind "FATAL ERROR: ", flush, exit(EXIT_FAILURE) */
void crash(const char *fmt, ...)
{
    char buf[4096];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (g_crash_throws)
        throw LvbError{buf};
    printf("\nFATAL ERROR: %s\n", buf);
    fflush(stdout);
    fflush(stderr);
    exit(EXIT_FAILURE);
}

/* src/Error.c:64-67 */
void lvb_assertion_fail(const char *test, const char *file, int line)
{
    crash("assertion failed at '%s' line %d: %s", file, line, test);
}

/* ---- Marsaglia's UNI, src/RandomNumberGenerator.c:87-198.  A lagged-Fibonacci sequence
 * (lags 97 and 33, subtraction mod 1) combined with an arithmetic sequence mod 16777213/2^24;
 * all in doubles whose values are multiples of 2^-24, so every operation is exact and the
 * stream is reproducible bit for bit. ---- */
static const double kCd = 7654321.0 / 16777216.0;
static const double kCm = 16777213.0 / 16777216.0;

void Uni::rinit(int ijkl)
{
    if (DBL_MANT_DIG < 24)
        crash("FP type unsuitable for uni() random no. generator\n");
    if (ijkl < 0 || ijkl > 900000000)
        crash("rinit: ijkl = %d -- out of range\n\n", ijkl);
    /* seed -> four small seeds (:179-185) */
    const int ij = ijkl / 30082;
    const int kl = ijkl - 30082 * ij;
    int i = ((ij / 177) % 177) + 2;
    int j = (ij % 177) + 2;
    int k = ((kl / 169) % 178) + 1;
    int l = kl % 169;
    if (i <= 0 || i > 178)
        crash("rinit: i = %d -- out of range\n\n", i);
    if (j <= 0 || j > 178)
        crash("rinit: j = %d -- out of range\n\n", j);
    if (k <= 0 || k > 178)
        crash("rinit: k = %d -- out of range\n\n", k);
    if (l < 0 || l > 168)
        crash("rinit: l = %d -- out of range\n\n", l);
    if (i == 1 && j == 1 && k == 1)
        crash("rinit: 1 1 1 not allowed for 1st 3 seeds\n\n");
    /* 97 table entries of 24 random bits each (:107-131) */
    for (int e = 1; e <= 97; e++)
    {
        double s = 0.0, bit = 0.5;
        for (int b = 0; b < 24; b++)
        {
            const int m = (((i * j) % 179) * k) % 179;
            i = j;
            j = k;
            k = m;
            l = (53 * l + 1) % 169;
            if ((l * m) % 64 >= 32)
                s += bit;
            bit *= 0.5;
        }
        u[e] = s;
    }
    u[0] = 0.0;
    c = 362436.0 / 16777216.0;
    ui = 97;
    uj = 33;
    ready = true;
}

double Uni::uni()
{
    lvb_assert(ready);
    double x = u[ui] - u[uj];
    if (x < 0.0)
        x += 1.0;
    u[ui] = x;
    if (--ui == 0)
        ui = 97;
    if (--uj == 0)
        uj = 97;
    c -= kCd;
    if (c < 0.0)
        c += kCm;
    x -= c;
    if (x < 0.0)
        x += 1.0;
    return x;
}

/* src/RandomNumberGenerator.c:235-256: round to nearest, so 0 and upper have half weight */
long Uni::randpint(long upper)
{
    lvb_assert(upper >= 0);
    const double scaled = uni() * (double)upper;
    long r = (long)(scaled + 0.5);
    if (r < 0)
        r = 0;
    else if (r > upper)
        r = upper;
    return r;
}

} // namespace lvbhost
