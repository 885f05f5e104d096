// Gradient bucket planner: the implementation behind picd_b200.parallel.engine.plan_buckets (bound through ctypes; the
// Python body there is only the fallback for a tree without a build and the oracle of tests/test_bucket_planner.py).
//
// The reference relies on torch DDP's C++ Reducer for bucketing (a 1 MiB first bucket so that
// communication starts early, then 25 MiB buckets, assigned in reverse registration order).  Parameters here are
// laid out in the flat arena in that same order, so a bucket is simply a contiguous [start, end) element range.
// Every bucket end is rounded up to a multiple of `unit` elements (lcm(64, 4 * world)): the two-shot / NVLS bucket kernels
// split a bucket into `world` slices of whole 16-byte vectors, which must work for 3, 5, 6 or 7 ranks as well.
#include <cstdint>

#include "picd_b200.h"      // the public prototype of picd_plan_buckets, checked against the definition below

static inline int64_t round_up(int64_t n, int64_t a) { return a <= 1 ? n : (n + a - 1) / a * a; }

extern "C" int picd_plan_buckets(const int64_t* padded_sizes, int n_params, int64_t first_bytes, int64_t cap_bytes,
                                 int elem_bytes, int64_t unit, int64_t* starts, int64_t* ends, int* param_bucket,
                                 int max_buckets) {
    int nb = 0;
    int64_t start = 0, cur = 0, limit = first_bytes;
    int members = 0;
    for (int i = 0; i < n_params; ++i) {
        if (nb >= max_buckets) return -1;
        param_bucket[i] = nb;
        cur += padded_sizes[i];
        ++members;
        if ((cur - start) * elem_bytes >= limit) {
            cur = start + round_up(cur - start, unit);
            starts[nb] = start;
            ends[nb] = cur;
            ++nb;
            start = cur;
            members = 0;
            limit = cap_bytes;
        }
    }
    if (members > 0) {
        if (nb >= max_buckets) return -1;
        cur = start + round_up(cur - start, unit);
        starts[nb] = start;
        ends[nb] = cur;
        ++nb;
    }
    return nb;
}
