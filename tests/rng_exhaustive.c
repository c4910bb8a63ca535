/* TEST INFRASTRUCTURE.  Exhaustive check (all 2^32 inputs) that the device's division-free RNG
 * float conversion, (float)fma(x, 2^-32, x) with x = r * 2^-32, equals the reference's
 * (float)(r / 4294967295.0) (src/utils.cu:228) bit for bit.  Prints the number of mismatches.
 * Build: gcc -O2 -mfma -fopenmp (fma() must be the IEEE fused operation, as on the GPU). */
#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <string.h>

int main(void) {
    unsigned long long bad = 0, bad_double = 0;
#pragma omp parallel for reduction(+ : bad, bad_double) schedule(static)
    for (long long i = 0; i < (1LL << 32); i++) {
        uint32_t r = (uint32_t)i;
        double q = (double)r / 4294967295.0;
        double x = (double)r * 2.3283064365386963e-10;
        double y = fma(x, 2.3283064365386963e-10, x);
        float a = (float)q, b = (float)y;
        uint32_t ua, ub;
        memcpy(&ua, &a, 4);
        memcpy(&ub, &b, 4);
        bad += ua != ub;
        bad_double += q != y;
    }
    printf("float_mismatches %llu double_mismatches %llu\n", bad, bad_double);
    return bad != 0;
}
