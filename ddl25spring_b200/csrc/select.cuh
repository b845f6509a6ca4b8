// Register-resident selection networks for the coordinate-wise robust aggregators (K4 median, K5 trimmed mean).
//
// One thread owns one coordinate and holds the KP (power of two) client values in registers.  The network is Batcher's
// odd-even merge sort (543 compare-exchanges at KP = 64, against 672 for the bitonic network used in round 1), emitted as
// straight-line code over compile-time indices (select_nets.cuh; nvcc does not unroll the generating loop nest, the array
// would land in local memory) so every a[i] stays a register.  A compare-exchange is one FMNMX pair on the ALU pipe
// (64 lanes/clk/SM on B200), which is what bounds these kernels - not HBM - so comparators are what to save:
//
//  * median_of<KP>: only ranks KP/2-1 and KP/2 are read afterwards, so every comparator (or half of one) that cannot reach
//    them is pruned at generation time: 828 min/max operations instead of 1344 at KP = 64.  The caller pads K < KP values
//    SYMMETRICALLY (-inf below, +inf above) so that the median always sits at the centre ranks.
//  * window_mean<KP>: full sort, then the mean of ranks [lo, hi) (trimmed mean, general rank windows).
//
// Pipe balance: a compare-exchange written as two FMNMX keeps only the ALU pipe busy (64 lanes/clk/SM; ncu showed it 97 %
// active), so three of every four exchanges compute the max from the min with two exact integer IMADs on the FMA pipe
// (DDL_CB in select_nets.cuh).
// Non-finite inputs: a NaN behaves like +inf (it moves to the top ranks); a minority of NaN rows cannot poison a coordinate.
#pragma once
#include <cuda_runtime.h>
#include "select_nets.cuh"   // generated straight-line networks (scripts/gen_select_nets.py)

namespace ddl {

// a[] padded symmetrically to KP entries.  pair != 0: mean of the two centre ranks (even K); else the lower centre rank.
// `one` must be the integer 1 passed as a RUN-TIME value (a kernel argument): see DDL_CB in select_nets.cuh.
template <int KP>
__device__ __forceinline__ float median_of(float (&a)[KP], int pair, int one) {
  median_net<KP>(a, one, -one);
  const float m0 = a[KP / 2 - 1], m1 = a[KP / 2];
  return pair ? 0.5f * (m0 + m1) : m0;
}

// Symmetric padding value for slot k >= K: the first pad_hi slots after the data are +inf, the rest -inf, with
// pad_lo = (KP - K) / 2 (rounded down) so that rank r of the data becomes rank r + pad_lo of the padded array:
// the median of an even K lands on ranks KP/2-1 and KP/2, the median of an odd K (and the lower median) on KP/2-1.
__device__ __forceinline__ float median_pad(int k, int K, int KP) {
  const int pad_hi = (KP - K + 1) / 2;
  return (k - K < pad_hi) ? __int_as_float(0x7f800000) : __int_as_float(0xff800000);
}

// a[] padded with +inf above K.  Mean of sorted ranks [lo, hi).
template <int KP>
__device__ __forceinline__ float window_mean(float (&a)[KP], int lo, int hi, int one) {
  sort_net<KP>(a, one, -one);
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < KP; ++k) s += (k >= lo && k < hi) ? a[k] : 0.f;
  return s / (float)(hi - lo);
}

}  // namespace ddl
