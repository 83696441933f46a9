// numpy.cumsum(v / total) with chosen entries zeroed, bit for bit, on several host threads (host_cumsum.cpp)
#pragma once
#include <cstdint>

namespace dvf {
// raw[i] = sum_{k <= i} p_k with p_k = v[k] / total (divide) or v[k] (!divide), p_k = 0 for k in the ascending list
// `skip`; raw may alias v when !divide.  Every partial sum carries the rounding of the sequential recurrence.
void host_scaled_cumsum(const double* v, bool divide, double total, const int64_t* skip, int64_t n_skip, double* raw,
                        int64_t n);
}  // namespace dvf
