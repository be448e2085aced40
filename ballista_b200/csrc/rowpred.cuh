// rowpred.cuh — device side of RowPred (ops.h): a FilterExec predicate evaluated inside the consumer's kernel
// (join build / probe, hash repartition) instead of being materialised by a compaction pass.
#pragma once
#include "ops.h"

namespace b2 {

// ---- inline FilterExec predicate (RowPred): closed range tests, two instructions per Int32 row ----
__device__ __forceinline__ bool in_range_i32(int32_t x, int lo, unsigned width) { return (unsigned)(x - lo) <= width; }
__device__ __forceinline__ bool in_range_f64(double x, double lo, double hi) { return x >= lo && x <= hi; }
__device__ __forceinline__ bool row_pred(const RowPred& p, int64_t i) {
  bool ok = true;
  for (int t = 0; t < p.n; ++t) {
    if (p.is_f64[t])
      ok = ok && in_range_f64(static_cast<const double*>(p.col[t])[i], p.flo[t], p.fhi[t]);
    else
      ok = ok && in_range_i32(static_cast<const int32_t*>(p.col[t])[i], p.ilo[t], p.iwidth[t]);
  }
  return ok;
}

}  // namespace b2
