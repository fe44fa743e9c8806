// Build shim for the reference oracle: pdqsort (orlp/pdqsort, un-vendored CPM
// dependency of the reference, cmake/Dependencies.cmake:11-12) is only used to
// order points during the CAPT build (collision/capt.hh:145).  std::sort yields
// the same ordering up to ties between equal keys, which does not change any
// geometric query result.
#pragma once
#include <algorithm>
template <class It, class Cmp>
inline void pdqsort_branchless(It b, It e, Cmp c) { std::sort(b, e, c); }
template <class It, class Cmp>
inline void pdqsort(It b, It e, Cmp c) { std::sort(b, e, c); }
