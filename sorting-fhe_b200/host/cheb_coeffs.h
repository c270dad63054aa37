// cheb_coeffs.h -- Chebyshev coefficient tables for the index indicator.
// The reference generates `generated_doubled_sinc_coeffs.h` / `generated_coeffs.h` at build time
// (reference utils/generate_cheb_doubled_coeffs.cpp:11-51, utils/generate_cheb_coeffs.cpp:11-58): degree-13011
// Chebyshev interpolation (lbcrypto::EvalChebyshevCoefficients, a plain cosine sum) of Sinc<2N>::doubled_sinc /
// scaled_sinc on [-1,1], |c| < 1e-8 zeroed, the zero tail trimmed, values printed at the default ostream
// precision (6 significant digits).  Here the same table is produced on first use and memoised.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <map>
#include <mutex>
#include <vector>
#include "comparison.h"

// lbcrypto::EvalChebyshevCoefficients(func, a, b, degree): coefficients c_0..c_degree (c_0 enters as c_0/2)
inline std::vector<double> EvalChebyshevCoefficients(const std::function<double(double)>& func, double a, double b,
                                                     uint32_t degree) {
    const size_t n = (size_t)degree + 1;
    const double half_span = 0.5 * (b - a), mid = 0.5 * (b + a), step = M_PI / (double)n;
    std::vector<double> fx(n), c(n);
    for (size_t i = 0; i < n; i++) fx[i] = func(std::cos(step * (i + 0.5)) * half_span + mid);
    const double factor = 2.0 / (double)n;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) {
        double s = 0.0;
        for (size_t j = 0; j < n; j++) s += fx[j] * std::cos(step * (double)i * (j + 0.5));
        c[i] = s * factor;
    }
    return c;
}

inline std::vector<double> truncateChebTable(std::vector<double> c) {
    for (auto& v : c) if (std::abs(v) < 1e-8) v = 0.0;
    size_t last = c.size();
    while (last > 0 && c[last - 1] == 0.0) last--;
    c.resize(last);
    char buf[64];
    for (auto& v : c) { snprintf(buf, sizeof buf, "%.6g", v); v = strtod(buf, nullptr); }   // operator<<(double)
    return c;
}

// selectDoubledSincCoefficients<N>() (reference src/sort_algo.h:725-728)
template <int N> const std::vector<double>& selectDoubledSincCoefficients() {
    static const std::vector<double> t = truncateChebTable(EvalChebyshevCoefficients(Sinc<2 * N>::doubled_sinc, -1.0, 1.0, 13011));
    return t;
}
// selectCoefficients<N>() (reference src/sort_algo.h:627-629)
template <int N> const std::vector<double>& selectCoefficients() {
    static const std::vector<double> t = truncateChebTable(EvalChebyshevCoefficients(Sinc<2 * N>::scaled_sinc, -1.0, 1.0, 13011));
    return t;
}
