// rust_fmt.h — number formatting of the reference's writers.
// Every number in the reference's TSV/VCF/stdout output goes through Rust's `{}` (Display): for f64 that is the
// shortest digit string that round-trips, never in scientific notation, "NaN" / "inf" / "-inf" for the
// specials and "-0" for negative zero (cellector/src/main.rs:170, 358, 364, 477-495).  std::to_chars without
// a precision gives the shortest round-trip digits; chars_format::fixed keeps it positional.
#pragma once
#include <charconv>
#include <cmath>
#include <cstdint>
#include <string>

namespace clxhost {

inline void append_f64(std::string& out, double v) {
    if (std::isnan(v)) {
        out += "NaN";
        return;
    }
    if (std::isinf(v)) {
        out += v < 0 ? "-inf" : "inf";
        return;
    }
    // shortest round-trip digits and decimal exponent, then positional layout with zero padding (what Rust's
    // flt2dec::to_shortest_str does: 1e21 -> "1000000000000000000000", 1e-7 -> "0.0000001")
    char buf[40];
    const auto r = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::scientific);
    const char* p = buf;
    if (*p == '-') {
        out += '-';
        ++p;
    }
    char digits[24];
    int nd = 0;
    const char* q = p;
    for (; q < r.ptr && *q != 'e'; ++q)
        if (*q != '.') digits[nd++] = *q;
    int exp10 = 0;
    std::from_chars(q + 1 + (q[1] == '+' ? 1 : 0), r.ptr, exp10);
    while (nd > 1 && digits[nd - 1] == '0') --nd;
    const int point = exp10 + 1;  // digits before the decimal point
    if (point <= 0) {
        out += "0.";
        out.append((size_t)(-point), '0');
        out.append(digits, (size_t)nd);
    } else if (point >= nd) {
        out.append(digits, (size_t)nd);
        out.append((size_t)(point - nd), '0');
    } else {
        out.append(digits, (size_t)point);
        out += '.';
        out.append(digits + point, (size_t)(nd - point));
    }
}

inline void append_u64(std::string& out, uint64_t v) {
    char buf[24];
    const auto r = std::to_chars(buf, buf + sizeof(buf), v);
    out.append(buf, r.ptr);
}

inline std::string f64_str(double v) {
    std::string s;
    append_f64(s, v);
    return s;
}

// Rust `x as usize` for f64: saturating, NaN -> 0
inline uint64_t f64_as_usize(double v) {
    if (!(v > 0.0)) return 0;
    if (v >= 18446744073709551615.0) return UINT64_MAX;
    return (uint64_t)v;
}

}  // namespace clxhost
