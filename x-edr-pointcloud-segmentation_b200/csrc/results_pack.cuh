// results_pack.cuh -- region results of a whole batch, packed on the device.
//
// After the growth kernels every frame holds its own region meta records, inlier
// lists and boundary lists at fixed per-frame strides.  These kernels turn that
// into what the caller's PlanarRegion / ClusterRegion objects need
// (ps/planar_region.h:427-450, ps/cluster_region.h:161-168), contiguous over the
// batch so that each array leaves the device with ONE copy:
//   pack_offsets_kernel  exclusive prefix sums over frames of the five counts
//   pack_bulk_kernel     inlier / boundary index lists (+ the xyz gathers of
//                        CreatePlanarDetectedObjectProto, ps/detected_objects.cc:36-46,
//                        and CreateClusterDetectedObjectProto, ps/detected_objects.h:62-74),
//                        cell -> inlier-position map for the discontinuity output
//   pack_region_kernel   per region: record, discontinuous boundary set as positions in the
//                        inlier list (ps/detected_objects.cc:41-45), PlanarConvexHull output
//                        (ps/algorithms.h:526-549 with patch P1; collision::ConvexHull restated as
//                        Andrew's monotone chain, PoseFromPlane as in grow_planar.cuh), cluster
//                        centroid / extent (extension).
#pragma once
#include "../../include/pcseg_b200.h"
#include "grow_planar.cuh"
#include "grow_cluster.cuh"

namespace pcseg {

struct PackArgs {
  int F, N, rows;
  const float* xyz;                // [F][N][3]
  const FrameStats* stats;         // [F]
  const PlanarMeta* meta_planar;   // [F][rcap_planar]
  int rcap_planar;
  const ClusterMeta* meta_cluster; // [F][rcap_cluster] (nullable: no cluster stage)
  int rcap_cluster;
  const int* inl_planar;           // [F][N]
  const int* inl_cluster;          // [F][N]
  const int* boundary;             // [F][bcap]
  const unsigned char* bflags;     // [F][bcap]
  int bcap;
  int* scratch;                    // [F][scratch_stride] ints: the idle growth queue
  size_t scratch_stride;
  // packed outputs (device)
  int* offsets;                    // [5][F+1]
  pcseg_planar_record* planar;
  int* p_inl;
  float* p_inl_xyz;                // nullable
  int* bnd;
  float* bnd_xyz;                  // nullable
  unsigned char* bnd_disc;
  int* disc_idx;
  int* disc_pos;
  float* hull;                     // nullable
  pcseg_cluster_record* cluster;
  int* c_inl;
  float* c_inl_xyz;                // nullable
  int want_cluster_stats;
  int* err;                        // [1] OR of pack-side errors
};

constexpr int kPackOffPlanar = 0, kPackOffPInl = 1, kPackOffBnd = 2, kPackOffCluster = 3, kPackOffCInl = 4;

// One CTA: offsets[k][f] = sum over frames < f of count k; offsets[k][F] = total.
__global__ void __launch_bounds__(1024) pack_offsets_kernel(PackArgs a) {
  __shared__ int s_warp[33];
  __shared__ int s_carry[5];
  const int t = threadIdx.x;
  if (t < 5) s_carry[t] = 0;
  __syncthreads();
  for (int f0 = 0; f0 < a.F; f0 += 1024) {
    const int f = f0 + t;
    int v[5] = {0, 0, 0, 0, 0};
    if (f < a.F) {
      const FrameStats s = a.stats[f];
      v[0] = s.n_planar; v[1] = s.planar_inliers; v[2] = s.boundary_len;
      v[3] = a.meta_cluster ? s.n_cluster : 0; v[4] = a.meta_cluster ? s.cluster_inliers : 0;
    }
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      int total;
      const int ex = block_scan_excl<1024>(v[k], &total, s_warp);
      const int carry = s_carry[k];
      if (f < a.F) a.offsets[k * (a.F + 1) + f] = carry + ex;
      __syncthreads();
      if (t == 0) s_carry[k] = carry + total;
      __syncthreads();
    }
  }
  if (t < 5) a.offsets[t * (a.F + 1) + a.F] = s_carry[t];
}

// grid (slices, F): index lists, xyz gathers, position map.
__global__ void __launch_bounds__(256) pack_bulk_kernel(PackArgs a) {
  const int f = blockIdx.y;
  const int t0 = blockIdx.x * blockDim.x + threadIdx.x;
  const int stride = gridDim.x * blockDim.x;
  const FrameStats st = a.stats[f];
  const float* xyz = a.xyz + (size_t)f * a.N * 3;
  const int F1 = a.F + 1;
  {
    const int* src = a.inl_planar + (size_t)f * a.N;
    const int base = a.offsets[kPackOffPInl * F1 + f];
    int* posmap = a.scratch + (size_t)f * a.scratch_stride;
    for (int i = t0; i < st.planar_inliers; i += stride) {
      const int cell = src[i];
      a.p_inl[base + i] = cell;
      posmap[cell] = i;  // frame-level position; pack_region_kernel subtracts the region's begin
      if (a.p_inl_xyz) {
        const V3 p = ld3(xyz, cell);
        float* o = a.p_inl_xyz + 3 * (size_t)(base + i);
        o[0] = p.x; o[1] = p.y; o[2] = p.z;
      }
    }
  }
  if (a.meta_cluster) {
    const int* src = a.inl_cluster + (size_t)f * a.N;
    const int base = a.offsets[kPackOffCInl * F1 + f];
    for (int i = t0; i < st.cluster_inliers; i += stride) {
      const int cell = src[i];
      a.c_inl[base + i] = cell;
      if (a.c_inl_xyz) {
        const V3 p = ld3(xyz, cell);
        float* o = a.c_inl_xyz + 3 * (size_t)(base + i);
        o[0] = p.x; o[1] = p.y; o[2] = p.z;
      }
    }
  }
  {
    const int* src = a.boundary + (size_t)f * a.bcap;
    const unsigned char* fl = a.bflags + (size_t)f * a.bcap;
    const int base = a.offsets[kPackOffBnd * F1 + f];
    for (int i = t0; i < st.boundary_len; i += stride) {
      const int cell = src[i];
      a.bnd[base + i] = cell;
      a.bnd_disc[base + i] = fl[i];
      if (a.bnd_xyz) {
        const V3 p = ld3(xyz, cell);
        float* o = a.bnd_xyz + 3 * (size_t)(base + i);
        o[0] = p.x; o[1] = p.y; o[2] = p.z;
      }
    }
  }
}

// Ascending bitonic sort of n2 (power of two) 64-bit keys by the calling CTA.
__device__ __forceinline__ void cta_bitonic_sort(unsigned long long* a, int n2) {
  for (int k = 2; k <= n2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n2; i += blockDim.x) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long x = a[i], y = a[ixj];
          const bool up = (i & k) == 0;
          if ((x > y) == up) { a[i] = y; a[ixj] = x; }
        }
      }
      __syncthreads();
    }
  }
}

// Order-preserving float <-> unsigned (finite values; -0 folded onto +0 by the caller).
__device__ __forceinline__ unsigned f2ord(float v) {
  const unsigned u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned o) {
  return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

constexpr int kPackSortSmem = 4096;  // keys sorted in shared memory (two buffers); longer lists use the frame's scratch
constexpr int kPackBlock = 256;
constexpr int kPackChunk = 32;       // points per thread in a hull filter round
constexpr size_t kPackRegionSmem = 2 * (size_t)kPackSortSmem * sizeof(unsigned long long);

__device__ __forceinline__ int next_pow2(int n) {
  int p = 1;
  while (p < n) p <<= 1;
  return p;
}

// cross((a - o), (b - o)) in double, as the restated collision::ConvexHull (oracle/shims/collision/convex_hull.h).
// Inputs are floats widened to double: the differences and products are exact, so the SIGN of the result
// is exact and the hull does not depend on the order in which triples are tested.
__device__ __forceinline__ double hull_cross(double ox, double oy, double ax, double ay, double bx, double by) {
  return (ax - ox) * (by - oy) - (ay - oy) * (bx - ox);
}
__device__ __forceinline__ double key_x(unsigned long long v) { return (double)ord2f((unsigned)(v >> 32)); }
__device__ __forceinline__ double key_y(unsigned long long v) { return (double)ord2f((unsigned)v); }
__device__ __forceinline__ bool hull_pops(unsigned long long o, unsigned long long a, unsigned long long b) {
  return hull_cross(key_x(o), key_y(o), key_x(a), key_y(a), key_x(b), key_y(b)) <= 0.0;
}

// Ordered compaction of src[0, n) entries with keep(i) into dst; returns the count (all threads).
template <typename Keep>
__device__ __forceinline__ int cta_compact(const unsigned long long* src, int n, unsigned long long* dst, int* s_w, Keep keep) {
  int written = 0;
  for (int c0 = 0; c0 < n; c0 += kPackBlock) {
    const int i = c0 + threadIdx.x;
    const bool k = i < n && keep(i);
    const unsigned long long v = k ? src[i] : 0ull;
    int total;
    const int off = block_scan_excl<kPackBlock>(k ? 1 : 0, &total, s_w);  // barriers inside: src read before dst written
    if (k) dst[written + off] = v;
    written += total;
  }
  __syncthreads();
  return written;
}

// Andrew's monotone chain over sorted unique points pts[0, m) -> hull (counter-clockwise, collinear
// points dropped) in out[]; returns the vertex count.  One thread.
__device__ int hull_chain(const unsigned long long* pts, int m, unsigned long long* out) {
  int k = 0;
  if (m < 3) {
    for (int i = 0; i < m; ++i) out[k++] = pts[i];
    return k;
  }
  for (int i = 0; i < m; ++i) {
    while (k >= 2 && hull_pops(out[k - 2], out[k - 1], pts[i])) --k;
    out[k++] = pts[i];
  }
  for (int i = m - 2, t0 = k + 1; i >= 0; --i) {
    while (k >= t0 && hull_pops(out[k - 2], out[k - 1], pts[i])) --k;
    out[k++] = pts[i];
  }
  return k - 1;  // the start point is repeated at the end
}

// One CTA per frame.
__global__ void __launch_bounds__(kPackBlock) pack_region_kernel(PackArgs a) {
  extern __shared__ __align__(16) unsigned long long s_keys[];  // 2 * kPackSortSmem words (dynamic)
  __shared__ int s_count;
  __shared__ int s_w[kPackBlock / 32 + 1];
  __shared__ double s_red[kPackBlock / 32][3];
  __shared__ float s_mm[kPackBlock / 32][6];
  const int f = blockIdx.x, t = threadIdx.x;
  const int F1 = a.F + 1;
  const FrameStats st = a.stats[f];
  const float* xyz = a.xyz + (size_t)f * a.N * 3;
  int* scratch = a.scratch + (size_t)f * a.scratch_stride;
  const int* posmap = scratch;
  // sort space behind the position map (8-byte aligned) for boundaries longer than the shared buffers
  unsigned long long* gkeys = reinterpret_cast<unsigned long long*>(scratch + ((a.N + 1) & ~1));
  const size_t gcap = (a.scratch_stride - (size_t)((a.N + 1) & ~1)) / 2;  // 64-bit words available
  const int r0 = a.offsets[kPackOffPlanar * F1 + f];
  const int i0 = a.offsets[kPackOffPInl * F1 + f];
  const int b0 = a.offsets[kPackOffBnd * F1 + f];
  const PlanarMeta* meta = a.meta_planar + (size_t)f * a.rcap_planar;
  const int* bnd = a.boundary + (size_t)f * a.bcap;
  const unsigned char* bfl = a.bflags + (size_t)f * a.bcap;

  for (int r = 0; r < st.n_planar; ++r) {
    const PlanarMeta m = meta[r];
    const int nb = m.boundary_count;
    const int n2 = next_pow2(max(nb, 1));
    unsigned long long* keys = s_keys;                 // two buffers of n2 words each
    unsigned long long* keys2 = s_keys + kPackSortSmem;
    bool ok = true;
    unsigned long long* hull_spill = gkeys;             // chain output when it may not fit in a sort buffer
    if (n2 > kPackSortSmem) {
      keys = gkeys;
      keys2 = gkeys + n2;
      hull_spill = gkeys + 2 * (size_t)n2;
      ok = (size_t)4 * n2 <= gcap;  // else: boundary longer than the scratch can sort
    }
    int n_disc = 0, n_hull = 0;
    if (!ok) {
      if (t == 0) atomicOr(a.err, kErrBoundaryCap);
    } else {
      // ---- discontinuous boundary set -> positions in the inlier list, ascending, unique
      // (the reference walks the inlier list and tests set membership, ps/detected_objects.cc:39-46)
      if (t == 0) s_count = 0;
      __syncthreads();
      for (int i = t; i < nb; i += kPackBlock)
        if (bfl[m.boundary_begin + i])
          keys[atomicAdd(&s_count, 1)] = (unsigned long long)(unsigned)(posmap[bnd[m.boundary_begin + i]] - m.inlier_begin);
      __syncthreads();
      const int nf = s_count;
      const int nf2 = next_pow2(max(nf, 1));
      for (int i = nf + t; i < nf2; i += kPackBlock) keys[i] = ~0ull;
      __syncthreads();
      cta_bitonic_sort(keys, nf2);
      n_disc = cta_compact(keys, nf, keys2, s_w, [&](int i) { return i == 0 || keys[i - 1] != keys[i]; });
      for (int i = t; i < n_disc; i += kPackBlock) {
        const int pos = (int)keys2[i];
        const int o = b0 + m.boundary_begin + i;
        a.disc_pos[o] = pos;
        a.disc_idx[o] = a.inl_planar[(size_t)f * a.N + m.inlier_begin + pos];
      }
      __syncthreads();
      // ---- PlanarConvexHull output, ps/algorithms.h:526-549 (patch P1)
      if (a.hull) {
        float plane[4] = {m.plane[0], m.plane[1], m.plane[2], m.plane[3]};
        const PlaneFrame pf = make_plane_frame(plane);
        for (int i = t; i < n2; i += kPackBlock) {
          unsigned long long key = ~0ull;
          if (i < nb) {
            double x, y;
            plane_xy(pf, ld3(xyz, bnd[m.boundary_begin + i]), &x, &y);
            const float fx = (float)x + 0.0f, fy = (float)y + 0.0f;  // plane_xy widens floats; -0 -> +0
            key = ((unsigned long long)f2ord(fx) << 32) | f2ord(fy);
          }
          keys[i] = key;
        }
        __syncthreads();
        cta_bitonic_sort(keys, n2);
        int mcount = cta_compact(keys, nb, keys2, s_w, [&](int i) { return i == 0 || keys[i - 1] != keys[i]; });
        unsigned long long *cur = keys2, *oth = keys;
        // Filter rounds: a vertex of the hull of the whole set is a vertex of the hull of every subset
        // that contains it, so points that are not on the hull of their own run of <= kPackChunk
        // consecutive (sorted) points are dropped, all runs in parallel, until one thread's chain is short.
        while (mcount > 4 * kPackChunk) {
          const int prev = mcount;
          const int per = min(kPackChunk, max(4, (mcount + kPackBlock - 1) / kPackBlock));
          int written = 0;
          for (int tile0 = 0; tile0 < mcount; tile0 += per * kPackBlock) {
            const int c0 = tile0 + t * per, c1 = min(c0 + per, mcount);
            unsigned keepmask = 0u;
            if (c0 < c1) {
              unsigned long long h[2 * kPackChunk];
              const int hk = hull_chain(cur + c0, c1 - c0, h);
              // mark the chunk's points that are hull vertices (h is a subsequence of the chunk in
              // ascending order for the lower part, descending for the upper part)
              for (int q = 0; q < hk; ++q)
                for (int i = c0; i < c1; ++i)
                  if (cur[i] == h[q]) { keepmask |= 1u << (i - c0); break; }
            }
            int total;
            const int off = block_scan_excl<kPackBlock>(__popc(keepmask), &total, s_w);
            int o = written + off;
            for (unsigned mm = keepmask; mm; mm &= mm - 1) oth[o++] = cur[c0 + __ffs(mm) - 1];
            written += total;
          }
          __syncthreads();
          unsigned long long* tmp = cur; cur = oth; oth = tmp;
          mcount = written;
          if (mcount >= prev) break;  // every point is a vertex of its run: nothing more to drop
        }
        // the chain's stack holds at most 2 * mcount entries
        unsigned long long* hull_out = oth;
        if (2 * mcount > kPackSortSmem && n2 <= kPackSortSmem) hull_out = hull_spill;
        if (n2 > kPackSortSmem) hull_out = hull_spill;
        if (n2 <= kPackSortSmem && hull_out == hull_spill && (size_t)2 * mcount > gcap) mcount = 0;  // cannot happen for N >= 8
        if (t == 0) s_count = hull_chain(cur, mcount, hull_out);
        __syncthreads();
        n_hull = s_count;
        oth = hull_out;
        // back to the cloud frame: world_pose_plane * (x, y, 0)
        const V3 n{plane[0], plane[1], plane[2]};
        const V3 tr = vscale(-plane[3], n);
        float* out = a.hull + 3 * (size_t)(b0 + m.boundary_begin);
        for (int i = t; i < n_hull; i += kPackBlock) {
          const V3 pp{ord2f((unsigned)(oth[i] >> 32)), ord2f((unsigned)oth[i]), 0.0f};
          const V3 w = vadd(qrotate(pf.qx, pf.qy, pf.qz, pf.qw, pp), tr);
          out[3 * i] = w.x; out[3 * i + 1] = w.y; out[3 * i + 2] = w.z;
        }
      }
    }
    if (t == 0) {
      pcseg_planar_record rec;
      rec.frame = f; rec.label_id = m.label_id; rec.seed_index = m.seed_index;
      rec.inlier_begin = i0 + m.inlier_begin; rec.inlier_count = m.inlier_count;
      rec.boundary_begin = b0 + m.boundary_begin; rec.boundary_count = nb;
      rec.discontinuous_begin = b0 + m.boundary_begin; rec.discontinuous_count = n_disc;
      rec.hull_begin = b0 + m.boundary_begin; rec.hull_count = n_hull;
      for (int k = 0; k < 4; ++k) rec.plane[k] = m.plane[k];
      for (int k = 0; k < 3; ++k) rec.centroid[k] = m.centroid[k];
      rec.area = m.area;
      for (int k = 0; k < 10; ++k) rec.estimator[k] = m.est[k];
      a.planar[r0 + r] = rec;
    }
    __syncthreads();
  }

  // ---- cluster regions: record + centroid / extent (segmented reduction, one region at a time
  // by the whole CTA for long lists, one warp per region for short ones)
  if (a.meta_cluster) {
    const int cr0 = a.offsets[kPackOffCluster * F1 + f];
    const int ci0 = a.offsets[kPackOffCInl * F1 + f];
    const ClusterMeta* cm = a.meta_cluster + (size_t)f * a.rcap_cluster;
    const int* inl = a.inl_cluster + (size_t)f * a.N;
    const int lane = t & 31, warp = t >> 5;
    for (int r = 0; r < st.n_cluster; ++r) {
      const ClusterMeta m = cm[r];
      pcseg_cluster_record rec;
      rec.frame = f; rec.label_id = m.label_id;
      rec.inlier_begin = ci0 + m.inlier_begin; rec.inlier_count = m.inlier_count;
      for (int k = 0; k < 3; ++k) { rec.centroid[k] = 0.0f; rec.bbox_min[k] = 0.0f; rec.bbox_max[k] = 0.0f; }
      if (a.want_cluster_stats) {
        // Fixed summation order (restated on the CPU by the tests): warp w of 8 takes the contiguous block
        // [w*B, (w+1)*B) of the inlier list, B = ceil(ceil(n/8)/32)*32; lane l adds its elements i = l (mod 32)
        // in ascending order in double; a warp's total = lanes 0..31 added in that order; the region's total =
        // warps 0..7 added in that order; centroid = (float)(total / n).
        const int n = m.inlier_count;
        const int per_warp = ((n + 7) / 8 + 31) & ~31;
        double sx = 0.0, sy = 0.0, sz = 0.0;
        float mnx = FLT_MAX, mny = FLT_MAX, mnz = FLT_MAX, mxx = -FLT_MAX, mxy = -FLT_MAX, mxz = -FLT_MAX;
        const int wb = warp * per_warp, we = min(wb + per_warp, n);
        for (int i = wb + lane; i < we; i += 32) {
          const V3 p = ld3(xyz, inl[m.inlier_begin + i]);
          sx += (double)p.x; sy += (double)p.y; sz += (double)p.z;
          mnx = fminf(mnx, p.x); mny = fminf(mny, p.y); mnz = fminf(mnz, p.z);
          mxx = fmaxf(mxx, p.x); mxy = fmaxf(mxy, p.y); mxz = fmaxf(mxz, p.z);
        }
        // lanes folded in ascending lane order (sequential double adds), then warps in ascending order
        double wx = 0.0, wy = 0.0, wz = 0.0;
        for (int l = 0; l < 32; ++l) {
          wx += __shfl_sync(0xffffffffu, sx, l);
          wy += __shfl_sync(0xffffffffu, sy, l);
          wz += __shfl_sync(0xffffffffu, sz, l);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o));
          mnz = fminf(mnz, __shfl_xor_sync(0xffffffffu, mnz, o)); mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
          mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o)); mxz = fmaxf(mxz, __shfl_xor_sync(0xffffffffu, mxz, o));
        }
        if (lane == 0) {
          s_red[warp][0] = wx; s_red[warp][1] = wy; s_red[warp][2] = wz;
          s_mm[warp][0] = mnx; s_mm[warp][1] = mny; s_mm[warp][2] = mnz;
          s_mm[warp][3] = mxx; s_mm[warp][4] = mxy; s_mm[warp][5] = mxz;
        }
        __syncthreads();
        if (t == 0) {
          double tx = 0.0, ty = 0.0, tz = 0.0;
          for (int w = 0; w < kPackBlock / 32; ++w) { tx += s_red[w][0]; ty += s_red[w][1]; tz += s_red[w][2]; }
          for (int w = 0; w < kPackBlock / 32; ++w) {
            mnx = fminf(mnx, s_mm[w][0]); mny = fminf(mny, s_mm[w][1]); mnz = fminf(mnz, s_mm[w][2]);
            mxx = fmaxf(mxx, s_mm[w][3]); mxy = fmaxf(mxy, s_mm[w][4]); mxz = fmaxf(mxz, s_mm[w][5]);
          }
          rec.centroid[0] = (float)(tx / (double)n); rec.centroid[1] = (float)(ty / (double)n); rec.centroid[2] = (float)(tz / (double)n);
          rec.bbox_min[0] = mnx; rec.bbox_min[1] = mny; rec.bbox_min[2] = mnz;
          rec.bbox_max[0] = mxx; rec.bbox_max[1] = mxy; rec.bbox_max[2] = mxz;
        }
        __syncthreads();
      }
      if (t == 0) a.cluster[cr0 + r] = rec;
    }
  }
}

// ---- MultichannelCloud::TransformInPlace, ps/multichannel_cloud.cc:53-68 ----
// points: Pose3f * p = q * p + t (Eigen quaternion-vector product); normals: rotationMatrix() * n with
// Eigen's Quaternion::toRotationMatrix and the 3-term row.col reduction e0 + (e1 + e2).  24 B/point
// streamed once per array.
struct TransformArgs {
  float* points;   // nullable
  float* normals;  // nullable
  int n;
  float qx, qy, qz, qw, tx, ty, tz;
};

__global__ void __launch_bounds__(256) transform_in_place_kernel(TransformArgs a) {
  // toRotationMatrix (Geometry/Quaternion.h)
  const float tx2 = 2.0f * a.qx, ty2 = 2.0f * a.qy, tz2 = 2.0f * a.qz;
  const float twx = tx2 * a.qw, twy = ty2 * a.qw, twz = tz2 * a.qw;
  const float txx = tx2 * a.qx, txy = ty2 * a.qx, txz = tz2 * a.qx;
  const float tyy = ty2 * a.qy, tyz = tz2 * a.qy, tzz = tz2 * a.qz;
  const float r00 = 1.0f - (tyy + tzz), r01 = txy - twz, r02 = txz + twy;
  const float r10 = txy + twz, r11 = 1.0f - (txx + tzz), r12 = tyz - twx;
  const float r20 = txz - twy, r21 = tyz + twx, r22 = 1.0f - (txx + tyy);
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    if (a.points) {
      float* p = a.points + 3 * (size_t)i;
      const V3 v{p[0], p[1], p[2]};
      const V3 w = vadd(qrotate(a.qx, a.qy, a.qz, a.qw, v), V3{a.tx, a.ty, a.tz});
      p[0] = w.x; p[1] = w.y; p[2] = w.z;
    }
    if (a.normals) {
      float* p = a.normals + 3 * (size_t)i;
      const V3 v{p[0], p[1], p[2]};
      p[0] = r00 * v.x + (r01 * v.y + r02 * v.z);
      p[1] = r10 * v.x + (r11 * v.y + r12 * v.z);
      p[2] = r20 * v.x + (r21 * v.y + r22 * v.z);
    }
  }
}

// ---- stand-alone gather of Create*DetectedObjectProto for one index list ----
__global__ void __launch_bounds__(256) gather_points_kernel(const float* __restrict__ xyz, const int* __restrict__ idx,
                                                            int n, float* out) {
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const V3 p = ld3(xyz, idx[i]);
    out[3 * (size_t)i] = p.x; out[3 * (size_t)i + 1] = p.y; out[3 * (size_t)i + 2] = p.z;
  }
}
// mark[cell] = 1 for the cells of the discontinuous set; flags[k] = mark[idx[k]]
__global__ void __launch_bounds__(256) mark_cells_kernel(const int* __restrict__ cells, int n, unsigned char* mark, unsigned char v) {
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) mark[cells[i]] = v;
}
// One CTA: ordered compaction of the positions k with mark[idx[k]] != 0.
__global__ void __launch_bounds__(1024) positions_of_marked_kernel(const int* __restrict__ idx, int n,
                                                                   const unsigned char* __restrict__ mark,
                                                                   int* positions, int* n_out) {
  __shared__ int s_w[33];
  int written = 0;
  for (int c0 = 0; c0 < n; c0 += 1024) {
    const int i = c0 + threadIdx.x;
    const bool keep = i < n && mark[idx[i]] != 0;
    int total;
    const int off = block_scan_excl<1024>(keep ? 1 : 0, &total, s_w);
    if (keep) positions[written + off] = i;
    written += total;
  }
  if (threadIdx.x == 0) *n_out = written;
}

}  // namespace pcseg
