// Gradient bucket planner (native twin of picd_b200.parallel.engine.plan_buckets).
//
// The reference relies on torch DDP's C++ Reducer for bucketing (a 1 MiB first bucket so that
// communication starts early, then 25 MiB buckets, assigned in reverse registration order).  Parameters here are
// laid out in the flat arena in that same order, so a bucket is simply a contiguous [start, end) element range.
#include <cstdint>

extern "C" int picd_plan_buckets(const int64_t* padded_sizes, int n_params, int64_t first_bytes, int64_t cap_bytes,
                                 int elem_bytes, int64_t* starts, int64_t* ends, int* param_bucket, int max_buckets) {
    int nb = 0;
    int64_t start = 0, cur = 0, limit = first_bytes;
    int members = 0;
    for (int i = 0; i < n_params; ++i) {
        if (nb >= max_buckets) return -1;
        param_bucket[i] = nb;
        cur += padded_sizes[i];
        ++members;
        if ((cur - start) * elem_bytes >= limit) {
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
        starts[nb] = start;
        ends[nb] = cur;
        ++nb;
    }
    return nb;
}
