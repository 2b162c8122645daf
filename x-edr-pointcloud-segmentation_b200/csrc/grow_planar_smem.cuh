// grow_planar_smem.cuh -- shared-memory resident state of the planar growth
// fast path (used by grow_planar_pipe.cuh).
//
// Every structure on the serial chain lives in shared memory instead of L2:
//   * mask   : 1 bit per cell, set <=> labels[cell] == kUnlabeled (a cell that
//              is in the queue, claimed by a region or parked has the bit clear);
//   * ring   : the FIFO's next entries as (tag = queue position, cell, xyz); the
//              xyz of a newly enqueued cell is fetched with cp.async straight
//              into its slot, so the point gather is off the critical path.  The
//              global queue keeps every cell id; an entry whose slot was busy
//              (queue longer than the ring) is re-read from there (tag mismatch);
//   * htab   : open-addressing claim table, atomicMin on (cell << 7 | key), that
//              picks for every neighbour cell the smallest (lane, slot) key;
//   * stage  : per-lane moment products, so that lane j < 10 adds accumulator j
//              over the accepted lanes in FIFO order (sequential fp32 sums).
// Global memory only sees fire-and-forget stores (labels of accepted cells,
// inlier list, global queue).
#pragma once
#include "grow_planar.cuh"

namespace pcseg {

constexpr int kHashSize = 2048;  // >= 16 * 128 claims per step: almost every claim resolves in one round
constexpr unsigned kHashEmpty = 0xffffffffu;

struct PlanarSmem {
  unsigned* mask;   // [mask_words]
  int* q_tag;       // [ring]
  int* q_cell;      // [ring]
  float* q_xyz;     // [ring * 3]
  float* stage;     // [32 * 11]
  unsigned* htab;   // [kHashSize]
  int ring;         // power of two
};

__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

__device__ __forceinline__ bool mask_test(const unsigned* mask, int cell) {
  return (mask[cell >> 5] >> (cell & 31)) & 1u;
}

}  // namespace pcseg
