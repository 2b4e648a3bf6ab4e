// Host check of crtdev::sincos_0_2pi (csrc/crt_sincos.h, the code the device runs) over every argument the renderer can
// produce: pi * u and 2 pi * u with u = k * 2^-23 (PCG32::uniform, src/core/crt_random.h:34-36).  Prints how often it
// differs from binary64 sin / cos rounded once to binary32 (must be 0) and how often glibc's sinf / cosf do (the
// documented FP difference between the reference and the device).
//   g++ -O2 -ffp-contract=off tools/sincos_check.cpp -o /tmp/sincos_check && /tmp/sincos_check
#include <cmath>
#include <cstdio>
#include <initializer_list>

#include "../chaos-ray-tracing-course-2025_b200/csrc/crt_sincos.h"

int main() {
    const float pi = 3.14159274101257324219f, twopi = 6.28318548202514648438f;
    long n = 0, own_s = 0, own_c = 0, glibc_s = 0, glibc_c = 0;
    for (unsigned m = 0; m < (1u << 23); ++m) {
        union { unsigned u; float f; } v;
        v.u = 0x3f800000u | m;
        const float u = v.f - 1.0f;
        for (float a : { pi * u, twopi * u }) {
            float s, c;
            crtdev::sincos_0_2pi(a, s, c);
            const float rs = float(sin(double(a))), rc = float(cos(double(a)));
            own_s += s != rs, own_c += c != rc;
            glibc_s += sinf(a) != rs, glibc_c += cosf(a) != rc;
            ++n;
        }
    }
    printf("arguments %ld\nsincos_0_2pi vs rounded binary64: sin %ld cos %ld\nglibc sinf/cosf vs rounded binary64: sin %ld cos %ld\n", n, own_s, own_c,
           glibc_s, glibc_c);
    return own_s == 0 && own_c == 0 ? 0 : 1;
}
