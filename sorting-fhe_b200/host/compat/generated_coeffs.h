// generated_coeffs.h / generated_doubled_sinc_coeffs.h -- the reference generates these headers at build time
// (utils/generate_cheb_coeffs.cpp:11-58, utils/generate_cheb_doubled_coeffs.cpp:11-51: degree-13011 Chebyshev
// interpolation of Sinc<2N>::scaled_sinc / doubled_sinc on [-1,1], |c| < 1e-8 zeroed, zero tail trimmed, printed at 6
// significant digits).  This stand-in produces the same tables on first use, so that the reference's unmodified
// src/sort_algo.h compiles without running its generator.
#pragma once
#include <array>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

namespace sortfhe_compat {
inline double scaled_sinc(double x, double n) {
    if (std::abs(x) < 1e-10) return 1.0;
    return std::sin(M_PI * n * x) / (M_PI * n * x);
}
inline std::vector<double> cheb_table(double n, bool doubled) {
    const size_t cnt = 13012;
    const double step = M_PI / (double)cnt;
    std::vector<double> fx(cnt), c(cnt);
    for (size_t i = 0; i < cnt; i++) {
        const double x = std::cos(step * (i + 0.5));
        fx[i] = doubled ? scaled_sinc(x, n) + scaled_sinc(x + 0.5, n) : scaled_sinc(x, n);
    }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)cnt; i++) {
        double s = 0.0;
        for (size_t j = 0; j < cnt; j++) s += fx[j] * std::cos(step * (double)i * (j + 0.5));
        c[i] = s * 2.0 / (double)cnt;
    }
    for (auto& v : c) if (std::abs(v) < 1e-8) v = 0.0;
    size_t last = c.size();
    while (last > 0 && c[last - 1] == 0.0) last--;
    c.resize(last);
    char buf[64];
    for (auto& v : c) { snprintf(buf, sizeof buf, "%.6g", v); v = strtod(buf, nullptr); }
    return c;
}
}  // namespace sortfhe_compat

template <std::size_t N> const std::vector<double>& selectCoefficients() {
    static const std::vector<double> t = sortfhe_compat::cheb_table(2.0 * N, false);
    return t;
}
template <std::size_t N> const std::vector<double>& selectDoubledSincCoefficients() {
    static const std::vector<double> t = sortfhe_compat::cheb_table(2.0 * N, true);
    return t;
}
