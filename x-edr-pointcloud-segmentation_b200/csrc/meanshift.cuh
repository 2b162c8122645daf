// meanshift.cuh -- SlidingMeanShift iteration phase on the GPU.
//
// Replaces ps/mean_shift_segmentation.h:217-260 with internal::GetNeighborsShift
// (:97-131), GetWeightedShift (:134-146), FlatKernelWeights (:89-93) and
// PointIndex (.h:54-77, .cc:22-34).  Seeds are mutually independent while the
// labels are read-only, so this is one thread per seed: a gather over the
// 11x11 INDEX window around round(index), summed in the reference's window
// order (delta_col outer, delta_row inner) in fp32, divided by (float)n.
// Threads of a warp are 32 consecutive rows of one column, so their windows
// overlap almost completely and the gathers are served by L1 (one float4 {point, takes part}
// per window cell, packed by meanshift_pack_kernel).
#pragma once
#include "grow_common.cuh"

namespace pcseg {

struct ModeRecord {  // 32 B per cell
  float mx, my, mz;  // shifted seed point
  float irow, icol;  // shifted float index (PointIndex: x = row, y = col)
  float intensity;   // last window count (double in the reference; small ints)
  float valid;       // 1 = still valid after all iterations
  float is_seed;     // 1 = the cell was a seed (unlabeled, no NaN)
};

struct MeanShiftArgs {
  const float* xyz;
  const int* labels;
  int rows, cols;
  int iterations;
  int cell_begin, cell_end;  // this launch's shard of seed cells
  ModeRecord* modes;         // indexed by (cell - cell_begin)
  const float4* packed;      // [N] {x, y, z, takes part ? 1 : 0}: one 16-byte gather per window cell
};

// A window cell takes part iff it is unlabeled and its point has no NaN (:110-118).  Label and point are
// packed into one float4 per cell, so the 605 window cells a seed visits cost one vector load each
// instead of a label load and three scalar point loads.
__global__ void __launch_bounds__(256) meanshift_pack_kernel(const float* __restrict__ xyz, const int* __restrict__ labels,
                                                             int n, float4* packed) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const V3 p = ld3(xyz, i);
  const bool part = labels[i] == kUnlabeled && !vhasnan(p);
  packed[i] = make_float4(p.x, p.y, p.z, part ? 1.0f : 0.0f);
}

constexpr int kMsHalfWindow = 5;      // ps/mean_shift_segmentation.h:39
constexpr int kMsMinCount = 50;       // 0.5 * 5 * 5 * 4, :245-247

__global__ void __launch_bounds__(256) meanshift_modes_kernel(MeanShiftArgs a) {
  const int cell = a.cell_begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= a.cell_end) return;
  const int rows = a.rows, cols = a.cols;
  ModeRecord rec;
  V3 sp = ld3(a.xyz, cell);
  float irow = (float)(cell % rows), icol = (float)(cell / rows);
  float intensity = 1.0f;  // :217
  bool is_seed = (__ldg(a.labels + cell) == kUnlabeled) && !vhasnan(sp);
  bool valid = is_seed;
  for (int it = 0; it < a.iterations && valid; ++it) {
    const int row = (int)roundf(irow);
    const int col = (int)roundf(icol);
    int count = 0;
    V3 shift{0.0f, 0.0f, 0.0f};
    float srow = 0.0f, scol = 0.0f;
    for (int dc = -kMsHalfWindow; dc <= kMsHalfWindow; ++dc) {
      const int cc = col + dc;
      if (cc < 0 || cc >= cols) continue;
#pragma unroll
      for (int dr = -kMsHalfWindow; dr <= kMsHalfWindow; ++dr) {
        const int rc = row + dr;
        if (rc < 0 || rc >= rows) continue;
        const int n = cc * rows + rc;
        const float4 q = __ldg(a.packed + n);
        if (q.w == 0.0f) continue;
        const V3 p{q.x, q.y, q.z};
        const V3 d = vsub(p, sp);
        if (vsqn(d) > 1.0f) continue;  // kSquareDistanceThreshold, :33,124
        shift = vadd(shift, d);
        srow += (float)rc - irow;
        scol += (float)cc - icol;
        ++count;
      }
    }
    if (count < kMsMinCount) {
      valid = false;
      break;
    }
    const float tw = (float)count;
    sp = vadd(sp, vdiv(shift, tw));
    irow += srow / tw;
    icol += scol / tw;
    intensity = tw;
  }
  rec.mx = sp.x; rec.my = sp.y; rec.mz = sp.z;
  rec.irow = irow; rec.icol = icol;
  rec.intensity = intensity;
  rec.valid = valid ? 1.0f : 0.0f;
  rec.is_seed = is_seed ? 1.0f : 0.0f;
  a.modes[cell - a.cell_begin] = rec;
}

}  // namespace pcseg
