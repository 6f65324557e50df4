// The identities the sky shader's pow shortcuts rely on (rt_shade.cuh: environment_light), checked against
// the oracle's pow (oracle/rto_math.h, the same binary32 arithmetic as rt_math.cuh):
//   pow(0, 0.35) == 0, pow(1, 0.35) == 1, and pow(x, 500) == +0 for EVERY float 0 <= x <= 0.8.
// `sky_shortcuts_check full` walks all 1.06e9 floats of that range; the default run takes every 97th.
#include <cstdio>
#include <cstdint>
#include <cstring>
#include "../../oracle/rto_math.h"

int main(int argc, char** argv)
{
    const bool full = argc > 1 && argv[1][0] == 'f';
    if (rto_f2u(rto_powf(0.0f, 0.35f)) != 0u) { std::printf("FAIL pow(0,0.35)\n"); return 1; }
    if (rto_powf(1.0f, 0.35f) != 1.0f) { std::printf("FAIL pow(1,0.35) = %.9g\n", rto_powf(1.0f, 0.35f)); return 1; }
    const uint32_t top = rto_f2u(0.8f);
    const uint32_t step = full ? 1u : 97u;
    long long bad = 0, n = 0;
#pragma omp parallel for reduction(+ : bad, n) schedule(static)
    for (long long u = 0; u <= (long long)top; u += step) {
        const float x = rto_u2f((uint32_t)u);
        if (rto_f2u(rto_powf(x, 500.0f)) != 0u) bad++;
        n++;
    }
    if (rto_f2u(rto_powf(0.8f, 500.0f)) != 0u) bad++;
    if (bad) { std::printf("FAIL %lld of %lld values of pow(x,500), x <= 0.8, are not +0\n", bad, n); return 1; }
    std::printf("OK %lld\n", n);
    return 0;
}
