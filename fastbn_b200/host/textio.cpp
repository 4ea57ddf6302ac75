// Fast text <-> double conversion for the potential-table files of the `-a 2` path.
// The reference writes "%.7f " per state (the format of dataset/alarm/alarm_1k_pt) and reads every token with stod
// (src/Inference.cpp:129-145). At 10^6 cases x 105 states snprintf / strtod are the whole cost of the -f4 / -o files, so both
// directions get an exact fast path and fall back to libc for everything the fast path does not cover:
//   FormatFixed7  bit-identical to snprintf("%.7f") (round-half-even on the EXACT binary value, like glibc)
//   ParseDouble   bit-identical to strtod for plain decimals with <= 15 significant digits (Clinger's exact fast path)
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "fastbn_host.h"

namespace textio {

// Writes x as "%.7f" (no terminator) and returns the length. out must hold 352 bytes (the libc fallback for huge values writes up to 318).
int FormatFixed7(double x, char *out) {
    if (!(x >= 0.0) || x >= 1048576.0 || (x == 0.0 && std::signbit(x))) return snprintf(out, 352, "%.7f", x);  // nan, negatives, -0, large: libc
    // y = x * 1e7 exactly = p + err (p the rounded product, err its exact error by fma); p < 2^44 so p - floor(p) is exact
    const double p = x * 1e7;
    const double err = std::fma(x, 1e7, -p);
    const double fl = std::floor(p);
    uint64_t n = (uint64_t)fl;
    const double d = (p - fl) - 0.5;  // exact; a non-zero d is a multiple of ulp(p) > |err|
    if (d > 0.0 || (d == 0.0 && (err > 0.0 || (err == 0.0 && (n & 1))))) ++n;
    // d < 0 with d + err >= 0 cannot happen (|err| <= ulp(p)/2 < |d|); p - fl == 0 with err < 0 rounds to n as well
    const uint64_t ip = n / 10000000u;
    uint32_t fp = (uint32_t)(n % 10000000u);
    char tmp[24];
    int len = 0;
    if (ip == 0) {
        out[len++] = '0';
    } else {
        int k = 0;
        for (uint64_t v = ip; v; v /= 10) tmp[k++] = (char)('0' + v % 10);
        while (k) out[len++] = tmp[--k];
    }
    out[len++] = '.';
    for (int i = 6; i >= 0; --i) {
        out[len + i] = (char)('0' + fp % 10);
        fp /= 10;
    }
    return len + 7;
}

// Parses one number starting at p (no leading blanks), like strtod(p, &q). Returns the end of the token (== p when no
// number was found). [p, e) must be followed by a readable terminator byte (the callers parse inside a NUL-terminated or
// newline-terminated buffer), which the strtod fallback relies on.
const char *ParseDouble(const char *p, const char *e, double *out) {
    static const double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                      1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    const char *s = p;
    bool neg = false;
    if (s < e && (*s == '-' || *s == '+')) neg = *s++ == '-';
    uint64_t mant = 0;
    int digits = 0, frac = 0;
    const char *d0 = s;
    while (s < e && *s >= '0' && *s <= '9') {
        if (mant || *s != '0') ++digits;
        if (digits <= 18) mant = mant * 10 + (uint64_t)(*s - '0');
        ++s;
    }
    bool any = s > d0;
    if (s < e && *s == '.') {
        const char *f0 = ++s;
        while (s < e && *s >= '0' && *s <= '9') {
            if (mant || *s != '0') ++digits;
            if (digits <= 18) mant = mant * 10 + (uint64_t)(*s - '0');
            ++s;
            ++frac;
        }
        any = any || s > f0;
    }
    const bool plain = any && (s >= e || (*s != 'e' && *s != 'E' && *s != 'x' && *s != 'X' && *s != 'n' && *s != 'N' && *s != 'i' && *s != 'I'));
    if (plain && digits <= 15 && frac <= 22) {
        double v = (double)mant / kPow10[frac];  // both exact doubles: one correctly rounded division (Clinger)
        *out = neg ? -v : v;
        return s;
    }
    char *q = nullptr;
    *out = strtod(p, &q);
    return q;
}

}  // namespace textio

// C hooks for the tests (tests/test_host_mirror.py compares both fast paths with libc on random and edge inputs)
extern "C" int fbn_host_format_fixed(double x, char *out352) { return textio::FormatFixed7(x, out352); }
extern "C" int fbn_host_parse_double(const char *s, int len, double *out) {
    const char *q = textio::ParseDouble(s, s + len, out);
    return (int)(q - s);
}
