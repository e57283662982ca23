// Kernel hand-over per time segment instead of per grid.
//
// A producer kernel (persistent conv kernel, fused stack kernel) counts finished work units in global memory, one
// counter per (utterance, segment of `units_per_seg` consecutive units along time): every epilogue warp arrives on a
// shared-memory mbarrier after its last store of the unit, a dedicated signal warp waits for that barrier and executes
// __threadfence + atomicAdd.  A consumer launched behind it with programmatic dependent launch never executes
// griddepcontrol.wait: its loader thread polls, with acquire loads, the counters of the segments that cover the rows
// it is about to read, then fence.proxy.async.global (its bulk copies read through the async proxy).  Work units are
// ordered (utterance, time) in every kernel, so a consumer CTA that lands on an SM the producer's tail has left idle
// starts at the beginning of the data while the producer is still finishing its end.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace ttsb {

constexpr int FOLLOW_SEGS = 32;   // counters per utterance and counting launch
#ifndef FOLLOW_POLL_NS
#define FOLLOW_POLL_NS 64          // back-off between polls of a counter
#endif
#ifndef FOLLOW_TIMEOUT_NS
#define FOLLOW_TIMEOUT_NS 20000000000ull   // a follower that has polled this long traps instead of hanging the GPU
#endif

struct TcFollow {
  const int* done;        // producer's counters [batch][FOLLOW_SEGS]; nullptr = wait for the whole previous grid
  int unit_rows;          // rows of the producer's output one of its work units covers ...
  int row_offset;         // ... row r belongs to unit (r + row_offset) / unit_rows
  int units_per_seg;
  int n_units;            // per utterance
  int mult;               // counts per finished unit (column groups of the persistent conv kernel)
  const int* armed;       // == armed_id once this inference's first kernel has zeroed the counters (follow_armed)
  int armed_id;
  int early;              // the dependents may launch before the check (unique ids; see follow_armed)
};

#ifdef __CUDACC__
// The counters are zeroed by the first kernel of every inference, once the previous inference has completed.
// Programmatic dependent launches cascade — with small grids a whole inference is resident while the previous one
// drains — so a follower that polled early would take the previous inference's final counts for its producer's.
// The first kernel therefore publishes an `armed` word behind the zeroing and every follower waits for it.  Two forms:
//  * early (ordinary launches): every inference has a process-wide unique id; the first kernel stores it after the
//    zeroing and a follower's loader waits for exactly that value before its first poll.  The word holds the previous
//    inference's id until then, and the first kernel overwrites it only after the previous inference has completed,
//    when none of ITS followers can still be waiting.  Nothing holds the dependent launches back: the whole next
//    inference may be resident, weights prefetched, while the previous one drains;
//  * late (stream capture: a replayed graph carries the same id every time): the first kernel's block 0 stores 0
//    BEFORE it lets its dependents launch and 1 after the zeroing; every following CTA waits for 1 before IT lets its
//    dependents launch (one thread checks, __syncthreads, then every thread triggers — a CTA counts as triggered once
//    all its threads have).  A dependent launches when every CTA of its primary has triggered, so no kernel of the next
//    replay can poll before its own first kernel has cleared the word, and by then every follower CTA of the previous
//    replay has passed its check.
// The kernels that do not follow (first conv of the chain, per-conv fallback) wait for whole grids and need neither.
// Watchdog of the poll loops: a producer that never counts (a bug, or a kernel of the chain that was never launched)
// must surface as a CUDA error, not as a device that spins for ever.  Every 4,096 polls (a few ms) the global timer is
// read; FOLLOW_TIMEOUT_NS (20 s: four orders of magnitude above the longest legitimate wait, a whole previous
// inference) after the first reading the thread traps.  Costs nothing on a wait that is already satisfied.
struct FollowPatience {
  uint32_t polls = 0;
  unsigned long long t0 = 0;
  __device__ __forceinline__ void tick() {
    if ((++polls & 4095u) != 0) return;
    unsigned long long now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
    if (t0 == 0) t0 = now;
    else if (now - t0 > FOLLOW_TIMEOUT_NS) __trap();
  }
};

__device__ __forceinline__ void follow_armed(const TcFollow& f) {
  FollowPatience patience;
  for (;;) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(f.armed) : "memory");
    if (v == f.armed_id) break;
    __nanosleep(FOLLOW_POLL_NS);
    patience.tick();
  }
}

struct FollowState {
  int b = -1, seg = -1;   // segments up to `seg` of utterance `b` are known complete
};

// Blocks until the producer has finished every unit that covers rows [a0, a1) of utterance b (rows outside the
// tensor belong to its first / last unit, whose epilogues write the margins).  One thread.
__device__ __forceinline__ void follow_wait(const TcFollow& f, int b, int a0, int a1, FollowState& st) {
  a0 = a0 < 0 ? 0 : a0;
  a1 = a1 < 1 ? 1 : a1;
  int u0 = (a0 + f.row_offset) / f.unit_rows, u1 = (a1 - 1 + f.row_offset) / f.unit_rows;
  u0 = u0 < f.n_units ? u0 : f.n_units - 1;
  u1 = u1 < f.n_units ? u1 : f.n_units - 1;
  const int s0 = u0 / f.units_per_seg, s1 = u1 / f.units_per_seg;
  if (st.b < 0 && f.early) follow_armed(f);   // first poll: the counters are this inference's from here on
  if (st.b != b) {
    st.b = b;
    st.seg = -1;
  }
  if (s1 <= st.seg) return;
  for (int s = s0 > st.seg + 1 ? s0 : st.seg + 1; s <= s1; ++s) {
    const int left = f.n_units - s * f.units_per_seg;
    const int target = (left < f.units_per_seg ? left : f.units_per_seg) * f.mult;
    const int* p = f.done + b * FOLLOW_SEGS + s;
    FollowPatience patience;
    for (;;) {
      int v;
      asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
      if (v >= target) break;
      __nanosleep(FOLLOW_POLL_NS);
      patience.tick();
    }
  }
  st.seg = s1;
  asm volatile("fence.proxy.async.global;" ::: "memory");
}
#endif

}  // namespace ttsb
