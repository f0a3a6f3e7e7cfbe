// Merging sorted k-lists by rank: the device form of the staff baselines' reduction operator
// ("merge two sorted k-lists, keep the first k", benchmarks/bench_2 @0xbc70) and of the
// reference's gather-then-reselect (engine.cpp:289-309), for any number of lists of any length.
// Lists are sorted by (distance asc, id desc) and ids are unique across lists, so the final
// position of an entry is its own index plus the number of entries that precede it in every
// other list -- a binary search per list, no staging buffer, no capacity limit.
#pragma once

namespace dmlp {
namespace k {

/// Entries of the sorted list (d, id)[0..len) that come strictly before (dv, idv).
__device__ __forceinline__ int entries_before(const double *__restrict__ d, const int *__restrict__ id, int len,
                                              double dv, int idv) {
  int lo = 0, hi = len;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    const double dm = d[mid];
    const bool before = dm < dv || (dm == dv && id[mid] > idv);
    if (before) lo = mid + 1; else hi = mid;
  }
  return lo;
}

/// Real entries at the head of a padded list: padding has id < 0 and follows every real entry.
__device__ __forceinline__ int real_entries(const int *__restrict__ id, int len) {
  int lo = 0, hi = len;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (id[mid] >= 0) lo = mid + 1; else hi = mid;
  }
  return lo;
}

}  // namespace k
}  // namespace dmlp
