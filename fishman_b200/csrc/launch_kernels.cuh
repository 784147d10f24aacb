// launch_kernels.cuh -- __global__ entry points: the generic one-thread-per-item launcher of
// the bodies in dt_kernels.cuh / vd_kernels.cuh, the cooperative flip loop, the fused
// centroid/update kernel, and the small setup / export kernels of the C ABI.
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include "dt_driver.h"
#include "vd_kernels.cuh"
#include "star_kernels.cuh"

namespace fm {
namespace cg = cooperative_groups;

__device__ __forceinline__ void use_device_stamp(DtCtx& c) {
  if (c.stamp_dev) c.stamp = *c.stamp_dev;
}
__device__ __forceinline__ void use_device_stamp(VdCtx&) {}

// A batch of warm Lloyd iterations is queued without the host looking in between
// (fm_iter_end below).  Once an iteration of the batch converged, found a turned-over triangle
// or hit an error (a peer rank that missed a barrier included), counters[FM_T_HALT] is set and
// every later launch of the batch leaves at once.  ONE word, one load: the kernels of the warm
// iteration are chains of dependent gathers, and a gate in front of the first gather is a
// whole extra memory latency per thread (measured: sweep 0.115 -> 0.183 ms with a two-load
// gate ahead of the row load) -- so the sweep issues its row load first and tests the word
// while the row is in flight (KSweepImpl::run_gated), and the kernels of a cold build, which
// never run in a halted batch, are not gated at all.
__device__ __forceinline__ bool fm_halted(const int* counters) {
  return counters && counters[FM_T_HALT] != 0;   // written only between launches: a plain, L1-cacheable load
}
__device__ __forceinline__ bool fm_halted(const DtCtx& c) { return fm_halted(c.counters); }
__device__ __forceinline__ bool fm_halted(const VdCtx& c) { return fm_halted(c.counters); }
template <class K> struct KGated { static const bool value = false; };
// The sharded sweep is gated (its request / boundary lists are not deduplicated and would grow
// with every halted iteration).  The unsharded sweep is NOT: even overlapped with the row load the
// gate cost it 7 of 122 us (A/B on one box, profiles/r02_g_ab_warm.txt), and a sweep in a halted
// batch is harmless -- it refreshes circumcentres and queues violated edges that fm_flip_coop,
// which is gated, then drops (advancing the stamp, so that the marks it leaves are never taken
// for those of a later repair).
template <> struct KGated<KSweepShT<false>> { static const bool value = true; };

// Resident 256-thread blocks per SM the launcher asks ptxas to make room for.  The gather
// kernels are latency / L1 bound; their filter-only forms (K<false>) contain no calls and
// fit 64 registers without spilling, so four blocks hide the gather latency.
template <class K> struct KOcc { static const int value = 1; };
#ifndef FM_OCC_SWEEP
#define FM_OCC_SWEEP 4
#endif
template <> struct KOcc<KSweepT<false>> { static const int value = FM_OCC_SWEEP; };
template <> struct KOcc<KSweepShT<false>> { static const int value = FM_OCC_SWEEP; };
template <> struct KOcc<KLocateVoteT<false>> { static const int value = 4; };
template <> struct KOcc<KCheckT<false>> { static const int value = 4; };

// Generic launcher: one thread per work item running a body from *_kernels.cuh.
template <class K, class Ctx>
__global__ void __launch_bounds__(256, KOcc<K>::value) fm_kernel(Ctx c, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  use_device_stamp(c);
  if (i >= n) return;
  if constexpr (KGated<K>::value) K::run_gated(c, i);
  else K::run(c, i);
}
// K over [0, counters[idx]): the count lives on the device, the host launches for an upper bound
template <class K>
__global__ void __launch_bounds__(256) fm_kernel_counted(DtCtx c, int idx) {
  use_device_stamp(c);
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < c.counters[idx]) K::run(c, i);
}

// Second half of a filtered launch: the exact body over the deferred items (grid-stride;
// the count is only known on the device).
template <class K>
__global__ void __launch_bounds__(128) fm_kernel_deferred(DtCtx c) {
  use_device_stamp(c);   // not gated: in a halted batch it still has to empty the deferred list of the (ungated) sweep
  int m = c.counters[FM_C_SLOW];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) K::run(c, c.slow[i]);
  // the last block to finish empties the list for the next filtered launch
  __syncthreads();
  if (threadIdx.x == 0 && atomicAdd(&c.counters[FM_C_SLOWDONE], 1) == (int)gridDim.x - 1) {
    c.counters[FM_C_SLOW] = 0;
    c.counters[FM_C_SLOWDONE] = 0;
  }
}

// The north-star's warp-cooperative form of the flip check: FOUR lanes per queued
// triangle.  Lanes 0..2 each fetch one vertex record of the triangle and run ONE edge's chain --
// neighbour row word, opposite vertex record, in-circle filter -- so the three chains that a
// thread per triangle walks one after the other run side by side; the triangle's vertices reach
// the three evaluating lanes by shuffles, the verdict (first violated edge in edge order, or
// "defer to the exact kernel") is taken from two ballots by lane 0.  Filter only, like
// KCheckT<false>; selected with FM_CHECK_COOP4=1 for the A/B (results log).
__global__ void __launch_bounds__(256, 4) fm_check_coop4(DtCtx c, int n) {
  use_device_stamp(c);
  const unsigned full = 0xffffffffu;
  const int gt = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = gt >> 2, sub = gt & 3, lane = threadIdx.x & 31, base = lane & ~3;
  const bool live = i < n, work = live && sub < 3;
  int t = live ? qe_tri(c.active[i]) : 0;
  Row r = live ? c.row[t] : Row{I4{0, 0, 0, 0}, I4{0, 0, 0, 0}};
  P4 mine = {0, 0, 0, 0}, opp = {0, 0, 0, 0};
  if (work) {
    mine = c.pts[geti(r.v, sub)];
    const int ne = geti(r.a, sub);
    const int d = ((const int*)c.row)[8 * (size_t)(ne >> 2) + (ne & 3)];
    opp = c.pts[d];
  }
  P3 p[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    p[k].x = __shfl_sync(full, mine.x, base + k);
    p[k].y = __shfl_sync(full, mine.y, base + k);
    p[k].z = __shfl_sync(full, mine.z, base + k);
  }
  int s = -1;
  if (work) s = incircle_fast(p[0], p[1], p[2], P3{opp.x, opp.y, opp.z});
  const unsigned unsure = (__ballot_sync(full, work && s == 0) >> base) & 7u;
  const unsigned viol = (__ballot_sync(full, work && s > 0) >> base) & 7u;
  if (live && sub == 0) {
    int fe = -1;
    bool defer = false;
    for (int k = 0; k < 3 && fe < 0 && !defer; k++) {   // the verdict a thread per triangle reaches edge by edge
      if ((unsure >> k) & 1u) defer = true;
      else if ((viol >> k) & 1u) fe = k;
    }
    if (defer) { defer_item(c, i); return; }
    c.active[i] = qe_make(t, fe);
    if (fe >= 0) {
      u64 key = vote_key(c.stamp, t);
      atomicMin((unsigned long long*)&c.vote[t], (unsigned long long)key);
      atomicMin((unsigned long long*)&c.vote[geti(r.a, fe) >> 2], (unsigned long long)key);
    }
  }
}

// All flip sub-rounds of one insertion round (or of one warm repair) in ONE cooperative
// launch: check -> mark -> flip -> patch over the current queue, grid-wide barrier between
// phases, until the queue is empty.  No host round trip per sub-round; the stamp lives in
// counters[FM_C_STAMP].  The bodies are the same KCheck/KMark/KFlip/KPatch the host-driven loop runs.
__global__ void __launch_bounds__(256) fm_flip_coop(DtCtx c, int* listA, int* listB, int cur, int skip_check) {
  cg::grid_group grid = cg::this_grid();
  const int gsz = gridDim.x * blockDim.x, gid = blockIdx.x * blockDim.x + threadIdx.x;
  int* list[2] = {listA, listB};
  volatile int* vc = c.counters;
  int stamp = vc[FM_C_STAMP];
  if (fm_halted(c)) {
    // Halted batch: the sweep before this launch still ran (it is not gated).  Drop what it queued
    // and move to a fresh stamp: its amark / vote words must not pass for those of a later repair.
    if (gid == 0) { vc[0] = 0; vc[1] = 0; vc[3] = 0; vc[FM_C_STAMP] = stamp + 1; }
    return;
  }
  if (vc[3] > 0) {  // a triangle turned over: leave everything to the rebuild (and stop the batch: the
                    // centroid stage tests the one halt word)
    if (gid == 0) vc[FM_T_HALT] = FM_HALT_REBUILD;
    return;
  }
  int sub = 0, maxq = 0;
  for (;; sub++) {
    int nact = vc[cur];
    if (nact == 0) break;
    if (sub >= (1 << 20)) { if (gid == 0) *c.err |= FM_E_OVERFLOW; break; }
    maxq = nact > maxq ? nact : maxq;
    c.stamp = stamp;
    c.active = list[cur];
    c.n_active = nact;
    c.next_active = list[cur ^ 1];
    c.qcount = c.counters + (cur ^ 1);
    if (!skip_check) {
      for (int i = gid; i < nact; i += gsz) KCheck::run(c, i);
      grid.sync();
    }
    skip_check = 0;
    for (int i = gid; i < nact; i += gsz) KMark::run(c, i);
    grid.sync();
    const int nflip = vc[FM_C_NFLIP];  // the winners, compact (KMark)
    if (nflip == 0) {
      // Nobody flips: every queued triangle passed its tests (a triangle that lost a vote lost it
      // to a winner), so nothing was queued for the next sub-round either -- the usual last,
      // verifying sub-round.  Leave without the two barriers of the flip / patch phases.
      if (gid == 0) c.counters[cur] = 0;
      sub++;
      stamp++;
      break;
    }
    for (int i = gid; i < nflip; i += gsz) KFlip::run(c, i);
    grid.sync();
    for (int i = gid; i < nflip; i += gsz) KPatch::run(c, i);
    if (gid == 0) {  // consumed: every thread read the two counts right after a barrier
      c.counters[cur] = 0;
      c.counters[FM_C_NFLIP] = 0;
    }
    grid.sync();
    cur ^= 1;
    stamp++;
  }
  if (gid == 0) {
    c.counters[FM_C_STAMP] = stamp;
    c.counters[5] += sub;
    if (maxq > c.counters[6]) c.counters[6] = maxq;
  }
}

// ---- cold build by stars (star_kernels.cuh)
template <class K>
__global__ void __launch_bounds__(256) fm_star_kernel(StarCtx s, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) K::run(s, i);
}
// The wrap kernels: 64 threads per block, 2 (2r+1)^2 ints of shared memory per thread (the candidate
// ranges of the block of cells).  Threads are in SORTED order: the lanes of a warp are neighbours on
// the sphere and read the same few cells' records through L1.
#ifndef FM_STAR_BLOCK
#define FM_STAR_BLOCK 64
#endif
#ifndef FM_STAR_MINB
#define FM_STAR_MINB 10   // resident blocks per SM the register budget is set for: 96 registers (20 bytes of spill).  A/B at
                          // 2.6M, star_wrap: 8 blocks (122 registers) 2.21 ms, 10: 2.09, 12 (80 registers, 124 bytes spilled): 2.31
#endif
__global__ void __launch_bounds__(FM_STAR_BLOCK, FM_STAR_MINB) fm_star_all(StarCtx s, int n) {
  extern __shared__ int star_smem[];
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) KStarAllT<false>::run(s, j, star_smem + threadIdx.x, blockDim.x);
}
// the sampling pass (StarCtx::dry): the first block of every `every` blocks
__global__ void __launch_bounds__(FM_STAR_BLOCK) fm_star_sample(StarCtx s, int n, int every) {
  extern __shared__ int star_smem[];
  int j = blockIdx.x * every * blockDim.x + threadIdx.x;
  if (j < n) KStarAllT<false>::run(s, j, star_smem + threadIdx.x, blockDim.x);
}
__global__ void __launch_bounds__(FM_STAR_BLOCK) fm_star_retry(StarCtx s) {
  extern __shared__ int star_smem[];
  const int m = s.sc[s.in_idx];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x)
    KStarRetry::run(s, i, star_smem + threadIdx.x, blockDim.x);
}
__global__ void __launch_bounds__(FM_STAR_BLOCK) fm_star_slow_near(StarCtx s) {
  extern __shared__ int star_smem[];
  const int m = s.sc[FM_SC_SLOW];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x)
    KStarSlowNear::run(s, i, star_smem + threadIdx.x, blockDim.x);
}
__global__ void __launch_bounds__(FM_STAR_BLOCK) fm_star_slow(StarCtx s) {
  extern __shared__ int star_smem[];
  const int m = s.sc[FM_SC_LIST0];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x)
    KStarSlow::run(s, i, star_smem + threadIdx.x, blockDim.x);
}
__global__ void __launch_bounds__(FM_STAR_BLOCK) fm_star_exact(StarCtx s) {
  extern __shared__ int star_smem[];
  const int m = s.sc[FM_SC_EXACT];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x)
    KStarExact::run(s, i, star_smem + threadIdx.x, blockDim.x);
}

__global__ void fm_bump_stamp(int* counters, int by) { counters[FM_C_STAMP] += by; }

// Last launch of a warm Lloyd iteration: what the host used to do after reading the counters
// back -- the convergence test (north-star kernel (d): the displacement maximum is already
// reduced), the "a triangle turned over, rebuild" decision, the statistics -- and the reset of
// the per-iteration counters for the next iteration of the batch.
__device__ __forceinline__ void fm_iter_end_body(int* counters) {
  if (counters[FM_T_HALT]) return;
  if (counters[FM_C_ERR]) { counters[FM_T_HALT] = FM_HALT_ERROR; return; }   // counters kept for the report
  if (counters[3] > 0) { counters[FM_T_HALT] = FM_HALT_REBUILD; return; }    // the centroid kernel did nothing
  const u64 md = *(const u64*)(counters + FM_C_MAXDISP);
  *(u64*)(counters + FM_T_MAXDISP) = md;
  counters[FM_T_ITERS] += 1;
  *(u64*)(counters + FM_T_FLIPS) += (u64)counters[2];
  counters[FM_T_SUB] += counters[5];
  if (counters[6] > counters[FM_T_MAXQ]) counters[FM_T_MAXQ] = counters[6];
  if (__longlong_as_double((long long)md) <= *(const double*)(counters + FM_T_TOL)) counters[FM_T_HALT] = FM_HALT_CONVERGED;
#pragma unroll
  for (int k = 0; k < 15; k++) counters[k] = 0;
}
__global__ void fm_iter_end(int* counters) { fm_iter_end_body(counters); }

// bookkeeping at the end of a host-driven flip sub-round (what thread 0 of fm_flip_coop does)
__global__ void fm_subround_end(int* counters, int cur, int nact) {
  counters[cur] = 0;
  counters[FM_C_NFLIP] = 0;
  counters[FM_C_STAMP] += 1;
  counters[5] += 1;
  if (nact > counters[6]) counters[6] = nact;
}

// Centroid + generator update + max-displacement in one pass (north-star kernels c + d):
// ring triangle ids are staged in shared memory, the per-cell chord displacement is reduced
// with warp shuffles, then once per block through shared memory, and one atomicMax per block
// reaches HBM.
// No block-level reduction: the warps of a block used to meet at a barrier for the shared-memory
// maximum (ncu: stall_barrier 1.2 per issue) -- now every warp sends its own maximum, if it beats
// the one seen so far.  Worth 1 % (uniform) / 2.6 % (x4); the block size itself (32 / 64 / 128
// threads) measures the same (profiles/r02_i_ab_centroid_block.txt).
#ifndef FM_CENTROID_BLOCK
#define FM_CENTROID_BLOCK 64
#endif
#ifndef FM_CENTROID_BLOCK_RHO
#define FM_CENTROID_BLOCK_RHO 64   // with a density: seven evaluations in flight per sub-triangle want ~128 registers
#endif
#ifndef FM_CENTROID_WARPS
#define FM_CENTROID_WARPS 32       // resident warps per SM the register budget is set for (uniform): 64 registers
#endif
#ifndef FM_CENTROID_WARPS_RHO
#define FM_CENTROID_WARPS_RHO 16   // 128 registers
#endif
#ifndef FM_CENTROID_WARPS_TAB
#define FM_CENTROID_WARPS_TAB 18   // density from the table: resident warps per SM the register budget is set for (96 registers, no spill; A/B 16..24 in profiles/r02_ab_centroid_density.txt)
#endif
template <bool UNIFORM, bool TABLE = false>
__global__ void __launch_bounds__(UNIFORM ? FM_CENTROID_BLOCK : FM_CENTROID_BLOCK_RHO,
                                  UNIFORM ? FM_CENTROID_WARPS * 32 / FM_CENTROID_BLOCK
                                          : (TABLE ? FM_CENTROID_WARPS_TAB : FM_CENTROID_WARPS_RHO) * 32 / FM_CENTROID_BLOCK_RHO)
    fm_centroid_update(const VdCtx c, int n) {
  extern __shared__ int ring_smem[];
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  // the two records every cell starts from are requested before the halt word is looked at
  // (a gate in front of the first load is a whole memory latency per warp)
  if (j < n) {
    asm volatile("prefetch.global.L1 [%0];" ::"l"(c.pts + c.cell_begin + j));
    if (c.ringc) asm volatile("prefetch.global.L1 [%0];" ::"l"(c.ringc + 8 * (size_t)(c.cell_begin + j)));
  }
  if (c.abort_flag && *c.abort_flag) return;  // = the halt word (launch_centroid_kernel)
  double disp = 0.0;
  if (j < n) disp = KCentroid::compute<UNIFORM, TABLE>(c, c.cell_begin + j, ring_smem + threadIdx.x, blockDim.x);
  for (int o = 16; o > 0; o >>= 1) disp = fmax(disp, __shfl_xor_sync(0xffffffffu, disp, o));
  // one atomicMax per WARP at most: only a warp that beats the maximum seen so far sends one
  // (the word is read first; after the first few warps almost nobody does)
  if ((threadIdx.x & 31) == 0 && disp > 0.0) {
    const u64 bits = (u64)__double_as_longlong(disp);
    if (bits > *(volatile u64*)c.maxdisp) atomicMax(c.maxdisp, bits);
  }
}

// ---- peer exchange over NVLink (fm_peer_attach): barrier between the ranks of a sharded mesh
// One PeerSync block per rank, inside the IPC-exported arena, mapped by every peer.
//   flag[k]      written by rank k: the last barrier epoch rank k has reached
//   epoch        local: barriers executed so far (every rank executes the same sequence)
//   maxdisp_all  bits of the largest chord displacement over ALL ranks of the current iteration
//   req_count[k] flip requests rank k put into this rank's inbox (sharded sweep), bad[k] = the
//                turned-over triangles rank k counted in its part of the sweep
struct PeerSync {
  int flag[8];
  int epoch;
  int pad[7];
  u64 maxdisp_all;
  u64 pad2[7];
  int req_count[8];
  int bad[8];
};
struct PeerCtx {
  PeerSync* sync[8];  // by rank; sync[rank] is the local block
  int rank, size;
  long long timeout_cycles;  // SM clocks a rank waits for a peer before it sets FM_E_PEER (FM_PEER_TIMEOUT_MS)
};
// phase 0 (before the centroid kernel: every rank has finished reading the generators),
// phase 1 (after it: every rank's stores into this rank's generator array have landed, and the
// displacement maxima have been combined) and phase 2 (after a sharded sweep: every rank's flip
// requests and circumcentres have landed here, and their counts are known); phase 3 is a plain
// barrier (sharded point location).  One warp; lane k talks to rank k.  A rank that does
// not show up within ~2 s of SM clocks sets FM_E_PEER instead of hanging the device.
__device__ __forceinline__ void fm_peer_barrier_body(const PeerCtx& pc, int phase, int* counters) {
  __shared__ int ep;
  const int k = threadIdx.x;
  PeerSync* self = pc.sync[pc.rank];
  if (k == 0) {
    ep = self->epoch + 1;
    self->epoch = ep;
    if (phase == 0) self->maxdisp_all = 0;
  }
  __syncwarp();
  const int e = ep;
  if (k < pc.size) {
    if (phase == 1) {
      u64 mine = *(volatile u64*)(counters + FM_C_MAXDISP);
      atomicMax_system((unsigned long long*)&pc.sync[k]->maxdisp_all, (unsigned long long)mine);
    }
    if (phase == 2) {
      *(volatile int*)&pc.sync[k]->req_count[pc.rank] = *(volatile int*)(counters + FM_C_REQ);
      *(volatile int*)&pc.sync[k]->bad[pc.rank] = *(volatile int*)(counters + 3);
    }
    __threadfence_system();
    *(volatile int*)&pc.sync[k]->flag[pc.rank] = e;
    long long t0 = clock64();
    while (*(volatile int*)&self->flag[k] < e) {
      if (clock64() - t0 > pc.timeout_cycles) {  // the data of this iteration never arrived: nothing may use it
        atomicOr(counters + FM_C_ERR, FM_E_PEER);
        ((volatile int*)counters)[FM_T_HALT] = FM_HALT_ERROR;
        break;
      }
    }
  }
  __syncwarp();
  if (phase == 1 && k == 0) {
    __threadfence_system();
    *(volatile u64*)(counters + FM_C_MAXDISP) = *(volatile u64*)&self->maxdisp_all;
  }
}
__global__ void fm_peer_barrier(PeerCtx pc, int phase, int* counters) {
  if (fm_halted(counters)) return;  // every rank takes the same decision (combined counts / maxima), or it is an error path
  fm_peer_barrier_body(pc, phase, counters);
}

// Sharded mesh: the barrier that closes the centroid stage (phase 1: all ranks' generator stores have
// landed here, the displacement maxima are combined) and fm_iter_end in ONE launch of one warp.
__device__ __forceinline__ void fm_iter_end_body(int* counters);
__global__ void fm_iter_end_peers(PeerCtx pc, int* counters) {
  if (fm_halted(counters)) return;
  fm_peer_barrier_body(pc, 1, counters);
  __syncwarp();
  if (threadIdx.x == 0) fm_iter_end_body(counters);
}

// Sharded point location: the slots this rank's walks ended in, to every rank's loc[]
struct LocFan {
  int* loc[8];      // by rank
  const int* plist; // this rank's slice of the level
  int rank, size;
};
__global__ void __launch_bounds__(256) fm_loc_fanout(LocFan f, int m) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  int p = f.plist[i], v = f.loc[f.rank][p];
  for (int k = 0; k < f.size; k++)
    if (k != f.rank) f.loc[k][p] = v;
}

// Fan-out of a sharded sweep's findings over NVLink: the flip requests to every rank's inbox
// (region `rank`), the circumcentres of the listed boundary triangles to the owners of their
// vertices.  Small: a few thousand requests and O(sqrt N) boundary triangles per iteration.
struct FanCtx {
  int* inbox[8];   // by rank: this rank's region of that rank's inbox ([rank] = req itself)
  P4* cc[8];       // by rank: circumcentre arrays
  const Row* row;
  const int *req, *bnd;
  int* counters;
  int rank, size, blk_per;
  PeerCtx pc;      // the barrier that follows the fan-out runs in the tail of this kernel
};
__global__ void __launch_bounds__(256) fm_sweep_fanout(FanCtx f) {
  if (fm_halted(f.counters)) return;
  __shared__ int last_block;
  const int gsz = gridDim.x * blockDim.x, gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nreq = f.counters[FM_C_REQ], nbnd = f.counters[FM_C_BND];
  for (int i = gid; i < nreq; i += gsz) {
    int msg = f.req[i];
    for (int k = 0; k < f.size; k++)
      if (k != f.rank) f.inbox[k][i] = msg;
  }
  for (int i = gid; i < nbnd; i += gsz) {
    int t = f.bnd[i];
    I4 v = f.row[t].v;
    P4 rec = f.cc[f.rank][t];
    int o0 = v.x / f.blk_per, o1 = v.y / f.blk_per, o2 = v.z / f.blk_per;
    if (o0 != f.rank) f.cc[o0][t] = rec;
    if (o1 != f.rank && o1 != o0) f.cc[o1][t] = rec;
    if (o2 != f.rank && o2 != o0 && o2 != o1) f.cc[o2][t] = rec;
  }
  // The block that finishes last runs the barrier (phase 2: requests and circumcentres have landed
  // everywhere, their counts are known) -- one launch and one launch gap fewer per iteration.
  __threadfence_system();   // this block's peer stores are out before it counts itself done
  __syncthreads();
  if (threadIdx.x == 0) {
    last_block = atomicAdd(f.counters + FM_C_SLOWDONE, 1) == (int)gridDim.x - 1;   // (idle between filtered launches)
    if (last_block) f.counters[FM_C_SLOWDONE] = 0;
  }
  __syncthreads();
  if (last_block && threadIdx.x < 32) fm_peer_barrier_body(f.pc, 2, f.counters);
}

// After a sharded sweep and its barrier: every rank turns ALL requests (its own and its peers')
// into what the unsharded sweep does on the spot -- the wanted edge, the two votes, the queue
// entry -- and adds up the turned-over counts.  Votes are atomicMin and the flip rounds depend
// only on the SET of queued triangles, so every rank goes on to the same flips.
__global__ void __launch_bounds__(256) fm_apply_requests(DtCtx c, PeerCtx pc, const int* inbox, int cap) {
  if (fm_halted(c)) return;
  use_device_stamp(c);
  const PeerSync* self = pc.sync[pc.rank];
  const int gsz = gridDim.x * blockDim.x, gid = blockIdx.x * blockDim.x + threadIdx.x;
  int cnt[8], bad = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) {  // all counts requested at once: eight independent loads, not a chain
    int n = k < pc.size ? *(volatile const int*)&self->req_count[k] : 0;
    bad += k < pc.size ? *(volatile const int*)&self->bad[k] : 0;
    if (n > cap) { if (gid == 0) atomicOr(c.err, FM_E_OVERFLOW); n = cap; }
    cnt[k] = n;
  }
  // ONE index space over the requests of all ranks (a loop per sending rank would run `size`
  // latency chains -- request, row, two atomics -- one after the other)
  int total = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) total += cnt[k];
  for (int i = gid; i < total; i += gsz) {
    int j = i, k = 0;
#pragma unroll
    for (int q = 0; q < 7; q++)
      if (k == q && j >= cnt[q]) { j -= cnt[q]; k = q + 1; }
    KApply::run(c, inbox[(size_t)k * cap + j]);
  }
  if (gid == 0) c.counters[3] = bad;
}

// out[r] = first triangle slot whose smallest vertex is >= r * per, r = 0 .. size (slots are
// sorted by that key right after a cold build): rank r sweeps the slots [out[r], out[r+1]), the
// ones that go with its cells
__global__ void fm_tri_range(int nt, const Row* row, int per, int size, int* out) {
  int r = threadIdx.x;
  if (r > size) return;
  long long key = (long long)r * per;
  int a = 0, b = nt;
  while (a < b) {
    int mid = (a + b) >> 1;
    I4 v = row[mid].v;
    if (min(v.x, min(v.y, v.z)) < key) a = mid + 1; else b = mid;
  }
  out[r] = (r == size) ? nt : a;
}

// ---- spatial renumbering of the triangle slots (after a cold build)
// Slots are handed out in creation order, so neighbours on the sphere end up far apart in
// memory.  Sorting the slots by their smallest vertex index (vertices are in Morton order)
// makes the per-triangle and per-cell passes of the Lloyd loop gather from nearby rows.
__global__ void fm_tri_keys(int nt, const Row* row, unsigned* key, int* idx) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nt) {
    I4 v = row[t].v;
    key[t] = (unsigned)min(v.x, min(v.y, v.z));
    idx[t] = t;
  }
}
__global__ void fm_tri_newid(int nt, const int* order, int* newid) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nt) newid[order[t]] = t;
}
__global__ void fm_tri_permute(int nt, const int* order, const int* newid, const Row* row, Row* row_out) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nt) {
    Row r = row[order[t]];
    I4 a = r.a;
    row_out[t] = Row{r.v, I4{(newid[a.x >> 2] << 2) | (a.x & 3), (newid[a.y >> 2] << 2) | (a.y & 3),
                             (newid[a.z >> 2] << 2) | (a.z & 3), 0}};
  }
}
__global__ void fm_remap_v2t(int n, const int* newid, int* v2t) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v2t[i] = newid[v2t[i]];
}

// ---- insertion levels (dt_kernels.cuh, "insertion levels")
__global__ void fm_level_keys(int n, const P4* pts, int K, unsigned* key, int* idx) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { key[i] = (unsigned)(K - point_class(pts[i].id, K)); idx[i] = i; }
}
// off[L] = first position of level L in the sorted keys (L = 0 .. K+1)
__global__ void fm_level_offsets(int n, const unsigned* sorted_key, int K, int* off) {
  int L = threadIdx.x;
  if (L > K + 1) return;
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (sorted_key[mid] < (unsigned)L) lo = mid + 1; else hi = mid;
  }
  off[L] = lo;
}
__global__ void fm_prev_keys(int n, const P4* pts, int K, int thr, int* prev, int* next_rev) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    bool coarser = point_class(pts[i].id, K) > thr;
    prev[i] = coarser ? i : -1;
    next_rev[n - 1 - i] = coarser ? n - 1 - i : -1;  // max-scan of the reversed order = nearest to the right
  }
}

__global__ void fm_morton_keys(int n, const double* x, const double* y, const double* z, u64* key,
                               int* idx) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { key[i] = morton_key(x[i], y[i], z[i]); idx[i] = i; }
}

// gather into internal order; inv[orig] = internal
__global__ void fm_gather_sorted(int n, const int* orig, const double* x, const double* y,
                                 const double* z, P4* pts, int* inv) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    int o = orig[i];
    pts[i] = P4{x[o], y[o], z[o], (long long)o};
    inv[o] = i;
  }
}

// scatter back to input order
__global__ void fm_scatter_orig(int n, const P4* pts, double* x, double* y, double* z) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    P4 q = pts[i];
    int o = (int)q.id;
    x[o] = q.x; y[o] = q.y; z[o] = q.z;
  }
}

// ---- one rank's block of generators in block order (fm_block_ids / fm_get_generators_block /
// fm_set_generators_block)
__global__ void fm_block_ids_kernel(int cnt, const P4* blk, int* ids) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < cnt) ids[i] = (int)blk[i].id + 1;
}
__global__ void fm_block_unpack(int cnt, const P4* blk, double* x, double* y, double* z) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < cnt) { P4 q = blk[i]; x[i] = q.x; y[i] = q.y; z[i] = q.z; }
}
struct BlockFan {
  P4* dst[8];   // this rank's block inside every rank's generator array ([0] = the local one)
  int n_dst;
};
__global__ void fm_block_pack_fan(int cnt, const double* x, const double* y, const double* z, BlockFan f) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cnt) return;
  P4 q = f.dst[0][i];           // keeps the id
  q.x = x[i]; q.y = y[i]; q.z = z[i];
  for (int k = 0; k < f.n_dst; k++) f.dst[k][i] = q;
}

// ---- measurement support: the fp64 FMA rate of this GPU (fm_fp64_peak).  Eight independent DFMA
// chains per thread (a DFMA's result is usable after ~8 cycles; 8 chains x 8 warps per scheduler
// keep the pipe full), 256 threads x 8 blocks per SM, no memory traffic.
__global__ void __launch_bounds__(256) fm_dfma_chains(int iters, double seed, double* sink) {
  double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double m = 0.999999, b = 1.0e-9;
#pragma unroll 1
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      a0 = fma(a0, m, b); a1 = fma(a1, m, b); a2 = fma(a2, m, b); a3 = fma(a3, m, b);
      a4 = fma(a4, m, b); a5 = fma(a5, m, b); a6 = fma(a6, m, b); a7 = fma(a7, m, b);
    }
  }
  double s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
  if (s == 123.456) *sink = s;   // never true: keeps the chains alive
}

// duplicates share a Morton key, so they sit in one run of equal keys
__global__ void fm_find_duplicates(int n, const u64* key, const P4* pts, int* dup) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int j = i + 1; j < n && key[j] == key[i]; j++)
    if (pts[j].x == pts[i].x && pts[j].y == pts[i].y && pts[j].z == pts[i].z) { atomicMin(dup, i); return; }
}

__global__ void fm_lonlat_to_xyz(int n, const double* lon, const double* lat, double* x, double* y,
                                 double* z) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {  // cartesian_transform, src/sphere_geometry_mod.F90:76-85
    double cl = cos(lat[i]);
    x[i] = cl * cos(lon[i]); y[i] = cl * sin(lon[i]); z[i] = sin(lat[i]);
  }
}

// argmax of p . d ; one packed (ordered value bits | index) atomicMax per warp
__global__ void fm_argmax_dot(int n, const P4* pts, double dx, double dy, double dz, u64* out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  u64 best = 0;
  if (i < n) {
    float v = (float)(pts[i].x * dx + pts[i].y * dy + pts[i].z * dz);
    unsigned b = __float_as_uint(v);
    b = (b & 0x80000000u) ? ~b : (b | 0x80000000u);  // order-preserving map
    best = ((u64)b << 32) | (unsigned)(0x7fffffff - i);  // ties: lowest index
  }
  for (int o = 16; o > 0; o >>= 1) {
    u64 other = __shfl_xor_sync(0xffffffffu, best, o);
    if (other > best) best = other;
  }
  if ((threadIdx.x & 31) == 0 && best) atomicMax(out, best);
}

__global__ void fm_init_loc(int n, int* loc, int b0, int b1, int b2, int b3) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) loc[i] = (i == b0 || i == b1 || i == b2 || i == b3) ? -1 : 0;
}

__global__ void fm_export_tris(int nt, const Row* row, const P4* pts, int* otri, int* oadj) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nt) {
    I4 v = row[t].v, a = row[t].a;
    otri[3 * t] = (int)pts[v.x].id + 1; otri[3 * t + 1] = (int)pts[v.y].id + 1;
    otri[3 * t + 2] = (int)pts[v.z].id + 1;
    oadj[3 * t] = (a.x >> 2) + 1; oadj[3 * t + 1] = (a.y >> 2) + 1; oadj[3 * t + 2] = (a.z >> 2) + 1;
  }
}

__global__ void fm_pack_cc(int nt, const P4* cc, const double* lon, const double* lat, double* xyz,
                           double* lonlat) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nt) {
    if (xyz) { P4 q = cc[t]; xyz[3 * t] = q.x; xyz[3 * t + 1] = q.y; xyz[3 * t + 2] = q.z; }
    if (lonlat) { lonlat[2 * t] = lon[t]; lonlat[2 * t + 1] = lat[t]; }
  }
}

// Edge tables (DV:411-460): cell i (original id) owns the edges to its ring
// neighbours with a larger id, in ring order; ecount is scanned into offsets.
__global__ void fm_edge_count(int n, const int* off, const int* ring_nbr, int* ecount) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    int c = 0;
    for (int j = off[i]; j < off[i + 1]; j++) c += (ring_nbr[j] > i + 1);
    ecount[i] = c;
  }
}
__global__ void fm_edge_fill(int n, const int* off, const int* ring_tri, const int* ring_nbr,
                             const int* eoff, int* de, int* ve_vvt, int* ve_vc) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    int e = eoff[i], b = off[i], d = off[i + 1] - b;
    for (int j = 0; j < d; j++) {
      int nb = ring_nbr[b + j];
      if (nb > i + 1) {
        de[2 * e] = i + 1; de[2 * e + 1] = nb;                       // DVT1 < DVT2, DV:421-424
        ve_vvt[2 * e] = ring_tri[b + j];                             // VVT1 = ring(j),   DV:444
        ve_vvt[2 * e + 1] = ring_tri[b + (j + 1) % d];               // VVT2 = ring(j+1), DV:445
        ve_vc[2 * e] = i + 1; ve_vc[2 * e + 1] = nb;                 // VC1, VC2, DV:446-447
        e++;
      }
    }
  }
}

}  // namespace fm
