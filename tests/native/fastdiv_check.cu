// Host-side check of rt::FastDiv (rt_device.cuh): exact quotient for every divisor class the launcher can
// produce, over edge values and a pseudo-random sweep of the full 32-bit range.
#include <cstdio>
#include <cstdint>
#include "rt_device.cuh"

int main(int argc, char** argv)
{
    const bool full = argc > 1 && argv[1][0] == 'f';                 // `fastdiv_check full`: every 32-bit dividend (about a minute)
    const uint32_t divisors[] = {1, 2, 3, 5, 7, 8, 10, 17, 60, 64, 100, 135, 240, 255, 256, 257, 480, 800, 1000, 1920, 3840, 4096, 65535, 65536,
                                 65537, 1000003, 0x7fffffffu, 0x80000000u, 0x80000001u, 0xfffffffeu, 0xffffffffu};
    unsigned long long checked = 0;
    for (uint32_t d : divisors) {
        const rt::FastDiv f = rt::make_fastdiv(d);
        uint32_t x = 12345u;
        for (int i = 0; i < 2000000; i++) {
            x = x * 747796405u + 2891336453u;
            uint32_t n = x;
            if (i < 64) n = (uint32_t)i;                                   // small values
            else if (i < 128) n = 0xffffffffu - (uint32_t)(i - 64);        // the top of the range
            else if (i < 4096) { const uint32_t k = (uint32_t)(i - 128) / 3u; n = k * d + (uint32_t)((i - 128) % 3) - 1u; }   // multiples of d and their neighbours
            if (f.div(n) != n / d) { std::printf("FAIL d=%u n=%u got %u want %u\n", d, n, f.div(n), n / d); return 1; }
            checked++;
        }
    }
    // every dividend for a few small divisors the item decode sees most (dispatch counts, tile rows)
    for (uint32_t d : {3u, 64u, 135u, 240u}) {
        const rt::FastDiv f = rt::make_fastdiv(d);
        for (uint64_t n = 0; n <= (full ? 0xffffffffull : 0x3ffffffull); n += 1) { if (f.div((uint32_t)n) != (uint32_t)n / d) { std::printf("FAIL d=%u n=%llu\n", d, (unsigned long long)n); return 1; } checked++; }
    }
    std::printf("OK %llu\n", checked);
    return 0;
}
