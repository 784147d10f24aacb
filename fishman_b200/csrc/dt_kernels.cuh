// dt_kernels.cuh -- per-thread bodies of the Delaunay kernels.
//
// What they replace in the reference (/root/reference/src/delaunay_voronoi_mod.F90,
// "DV"): the serial insertion loop DV:296-321 with its point-in-triangle buckets
// (DV:247-293, DV:884-948), the 1->3 split flip13 (DV:957-1004) and the Lawson
// legalisation validate_DT/flip22 (DV:845-882, DV:1098-1149).  The reference
// inserts one point at a time in index order and re-buckets the points of every
// destroyed triangle; here every triangle that still contains uninserted points
// takes ONE of them per round (chosen by a 64-bit atomicMin), all winners are
// split at once, and the mesh is brought back to Delaunay by rounds of
// conflict-free 2<->2 flips (pairs claimed by atomicMin votes).  Points are not
// kept in buckets: each uninserted point remembers the triangle SLOT it was last
// in and re-locates itself by a short visibility walk (slots are rewritten in
// place by splits and flips, so the walk starts next to the answer).
//
// Every body is a plain function of (context, thread id) that communicates only
// through atomicMin / atomicAdd / atomicExch and stamped words, so the same
// source runs (a) as the CUDA kernels of the product library and (b) thread by
// thread on the host in tests/emu (CPU logic tests only; never shipped).
#pragma once
#if defined(__CUDACC__)
#include <cooperative_groups.h>
#endif
#include "predicates.cuh"

namespace fm {

struct alignas(16) I4 {
  int x, y, z, w;
};

#if defined(__CUDA_ARCH__)
#define FM_ATOMIC_MIN64(p, v) atomicMin((unsigned long long*)(p), (unsigned long long)(v))
#define FM_ATOMIC_ADD32(p, v) atomicAdd((int*)(p), (int)(v))
#define FM_ATOMIC_EXCH32(p, v) atomicExch((int*)(p), (int)(v))
#define FM_ATOMIC_MAX64(p, v) atomicMax((unsigned long long*)(p), (unsigned long long)(v))
#else
template <class T>
inline T fm_host_min(T* p, T v) { T o = *p; if (v < o) *p = v; return o; }
template <class T>
inline T fm_host_max(T* p, T v) { T o = *p; if (v > o) *p = v; return o; }
template <class T>
inline T fm_host_add(T* p, T v) { T o = *p; *p = o + v; return o; }
template <class T>
inline T fm_host_exch(T* p, T v) { T o = *p; *p = v; return o; }
#define FM_ATOMIC_MIN64(p, v) fm_host_min<unsigned long long>((unsigned long long*)(p), (unsigned long long)(v))
#define FM_ATOMIC_MAX64(p, v) fm_host_max<unsigned long long>((unsigned long long*)(p), (unsigned long long)(v))
#define FM_ATOMIC_ADD32(p, v) fm_host_add<int>((int*)(p), (int)(v))
#define FM_ATOMIC_EXCH32(p, v) fm_host_exch<int>((int*)(p), (int)(v))
#endif

// counter++ returning the old value, for the list heads every thread of a launch may bump
// (queue lengths, winner / flip / deferred lists).  On the device the lanes of a warp that arrive
// together send ONE atomic to L2 and share the result (all those counters sit in one cache line:
// a million first votes of an insertion round were a million serialised atomics on one address
// and took longer than the walks that cast them).
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ int fm_count_inc(int* ctr) {
  cooperative_groups::coalesced_group g = cooperative_groups::coalesced_threads();
  int base = 0;
  if (g.thread_rank() == 0) base = atomicAdd(ctr, (int)g.size());
  return g.shfl(base, 0) + (int)g.thread_rank();
}
#define FM_COUNT_INC(p) fm_count_inc((int*)(p))
#else
#define FM_COUNT_INC(p) fm_host_add<int>((int*)(p), 1)
#endif

// One triangle slot: vertex ids and adjacency side by side, 32 bytes = one sector and one
// 256-bit load for the passes that need both (all of them but the ring walk).
struct alignas(32) Row {
  I4 v, a;
};

typedef unsigned long long u64;
static const u64 kNone64 = ~0ULL;

// slots of DtCtx::counters (16 ints; [0..14] are cleared together at the start of a build or
// repair, the stamp survives)
enum {
  FM_C_NFLIP = 7,      // pairs marked for a flip in the current sub-round = entries of DtCtx::fliplist
  FM_C_SLOW = 8,       // deferred items of a filtered launch
  FM_C_SLOWDONE = 9,   // blocks of the deferred kernel that have finished (self-reset)
  FM_C_MAXDISP = 10,   // 64-bit word: bits of the largest chord displacement ([10], [11])
  FM_C_ERR = 12,       // error bits (DtCtx::err points here in the CUDA library)
  FM_C_REQ = 13,       // sharded sweep: flip requests this rank found
  FM_C_NWIN = 13,      // cold build (no sharded sweep runs then): entries of DtCtx::wlist = splits of the round
  FM_C_BND = 14,       // sharded sweep: swept triangles with a vertex owned by another rank
  FM_C_STAMP = 15,     // next unused stamp
  // [16..31]: state of a BATCH of warm Lloyd iterations queued without a host round trip
  // (fm_iter_end, launch_kernels.cuh); the CUDA library allocates 32 ints
  FM_T_HALT = 16,      // sticky: 0 run; FM_HALT_* = why the remaining launches of the batch do nothing
  FM_T_ITERS = 17,     // iterations completed in the batch
  FM_T_MAXDISP = 18,   // 64-bit: displacement word of the last completed iteration ([18], [19])
  FM_T_FLIPS = 20,     // 64-bit: flips of the batch ([20], [21])
  FM_T_SUB = 22,       // flip sub-rounds of the batch
  FM_T_MAXQ = 23,      // largest flip queue of the batch
  FM_T_TOL = 24        // double: convergence tolerance, written by the host when the batch starts ([24], [25])
};
enum { FM_HALT_CONVERGED = 1, FM_HALT_REBUILD = 2, FM_HALT_ERROR = 3 };

// error bits reported through DtCtx::err
enum { FM_E_WALK = 1, FM_E_TOPOLOGY = 2, FM_E_OVERFLOW = 4, FM_E_PEER = 8 };

// Generator record: 32 bytes = one DRAM sector, so that a vertex gather costs one sector
// instead of four (x, y, z, id in separate arrays) and ONE 256-bit load (LDG.E.ENL2.256 on
// sm_100a).  id = original 0-based id (SoS key).
struct alignas(32) P4 {
  double x, y, z;
  long long id;
};

struct DtCtx {
  int n;                       // points
  P4* pts;                     // internal (space-filling-curve) order
  int* v2t;                    // per vertex: SOME incident triangle (ring start), kept by split/flip
  int* ringc;                  // per vertex: 8 cached ring triangle ids (vd_kernels.cuh, "ring cache"), or
                               // null; a flip invalidates the records of its four vertices
  P4* cc;                      // circumcentres per triangle, maintained by sweep/flip when non-null
  Row* row;                    // per triangle slot: v = vertices (x,y,z), CCW seen from outside;
                               // a[k] = (neighbour << 2 | its edge index) across the edge
                               // opposite vertex k (the reference's adjDT(k), DV:656-669)
  int* loc;                    // per point: slot of a triangle near/containing it, -1 = inserted
  const int* plist;            // points taking part in this round (the current insertion level), or
                               // null = all points; thread i of the locate kernels handles plist[i]
  int n_plist;
  int* wlist;                  // the triangles that split this round (KSplitFlag; length counters[FM_C_NWIN])
  int locate_only;             // sharded point location: walk and record loc[], the votes are cast later (KVote)
  const int* prev;             // per point: nearest point to the left (internal order) of a coarser level
  const int* next_rev;         // the same to the right, stored reversed: next_rev[n-1-i] = n-1-j for the
                               // smallest j >= i of a coarser level, or -1 (null = left side only)
  int level_thr;               // coarseness of the current level
  int level_max;               // coarsest class K
  u64* winner;                 // per triangle: min over its points of (priority << 32 | point); all ones between rounds
  int* rank;                   // per triangle: exclusive scan of "splits this round"
  int n_tri;                   // triangles alive at the start of the round
  // flip rounds
  int stamp;                   // global sub-round counter (>= 1)
  const int* stamp_dev;        // when non-null the device-resident stamp (counters[FM_C_STAMP]) overrides
                               // `stamp` at kernel entry: the flip loop advances it on the device
  u64* vote;                   // per triangle: (~stamp << 32 | voter) claimed by atomicMin
  int* flip_edge;              // per WINNING triangle of the sub-round: the edge it flips (KMark -> KFlip, KPatch)
  int* role;                   // per triangle: stamp << 2 | (1 winner, 2 partner)
  u64* mailbox;                // per (triangle, edge): stamp << 32 | new encoded neighbour
  int* amark;                  // per triangle: last stamp it was queued for
  int* active;                 // triangles to check this sub-round; an entry is (slot | (edge + 1) << 29): the check
                               // writes the edge it wants to flip INTO the entry (0: none), so that the mark phase
                               // reads the queue alone for the many entries that want nothing (qe_* below)
  int n_active;
  int* next_active;            // queue for the next sub-round
  int* fliplist;               // the winners of the current sub-round's votes (KMark), compact:
                               // KFlip and KPatch run over these instead of the whole queue
  int* qcount;                 // length of next_active (one of counters[0], counters[1])
  int* counters;               // [0],[1] queue lengths (ping-pong), [2] flips, [3] turned-over
                               // triangles, [4] splits, [5] flip sub-rounds, [6] largest queue,
                               // then see FM_C_* below
  int* slow;                   // items whose fp64 filter was inconclusive (length in counters[8]);
                               // a second launch runs the exact predicates on them
  int* err;
  // sharded warm sweep (fm_peer_attach; all zero otherwise): this rank sweeps the triangle slots
  // [t_begin, t_begin + launch size).  A violated edge becomes a request (t << 2 | edge) in
  // req[], a swept triangle with a vertex outside [blk_begin, blk_end) goes on bnd[]; a small
  // fan-out kernel then copies the requests into EVERY rank's inbox and the circumcentres of the
  // listed triangles to the ranks that own their vertices (launch_kernels.cuh), so that after a
  // barrier all ranks apply the same requests and run the same flip rounds.
  int t_begin;
  int pf_dist;                 // sweep: L2 prefetch distance in triangle slots (0 = none)
  int *req, *bnd;
  int blk_begin, blk_end;      // cells of this rank
};

// Filtered-then-exact, kernel level.  K<false> bodies evaluate only the fp64 filters (no
// calls, no stack frame, half the registers); an item with an inconclusive filter is pushed
// here and redone from scratch by K<true>, whose predicates go on to expansion arithmetic
// and the SoS tie-break.  K<true> alone is always correct (CPU logic tests, cooperative loop).
FM_HD inline void defer_item(const DtCtx& c, int item);

template <bool EXACT>
FM_HD inline int orient_t(const P3& a, const P3& b, const P3& cc, int ia, int ib, int ic) {
  return EXACT ? orient(a, b, cc, ia, ib, ic) : orient_fast(a, b, cc);
}
template <bool EXACT>
FM_HD inline int incircle_t(const P3& a, const P3& b, const P3& cc, const P3& d, int ia, int ib, int ic, int idd) {
  return EXACT ? incircle(a, b, cc, d, ia, ib, ic, idd) : incircle_fast(a, b, cc, d);
}

FM_HD inline P3 ldp(const DtCtx& c, int i) { const P4& q = c.pts[i]; return P3{q.x, q.y, q.z}; }
// point and SoS id with two 16-byte loads of the same sector
FM_HD inline P3 ldv(const DtCtx& c, int i, int& id) {
  P4 q = c.pts[i];
  id = (int)q.id;
  return P3{q.x, q.y, q.z};
}
// 1/sqrt(v), v > 0 and normal: hardware seed (MUFU.RSQ64H, ~2^-22) + ONE cubically convergent
// step  y (1 + e/2 + 3 e^2/8),  e = 1 - v y^2  = 5 DFMA-class instructions instead of the ~20
// of the library call; the remaining relative error (~0.3 e^3) is far below 2^-53.
FM_HD inline double fm_rsqrt(double v) {
#if defined(__CUDA_ARCH__)
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(v));
  double e = __fma_rn(-__dmul_rn(v, y), y, 1.0);
  double p = __fma_rn(0.375, e, 0.5);
  return __fma_rn(y, __dmul_rn(p, e), y);
#else
  return 1.0 / sqrt(v);
#endif
}

// circumcentre of the CCW triangle (a,b,c): normalised (b-a) x (c-a), DV:1335-1344
FM_HD inline P3 circumcentre(const P3& a, const P3& b, const P3& c, bool& degenerate) {
  double e3x = b.x - a.x, e3y = b.y - a.y, e3z = b.z - a.z;
  double e2x = c.x - a.x, e2y = c.y - a.y, e2z = c.z - a.z;
  // cross_product, src/math_mod.F90:12-22.  On the device the multiply-adds are spelled out so
  // that every kernel that calls this (sweep, flip, circumcentre) rounds identically, whatever
  // the compiler would contract: ranks of a sharded mesh then hold bit-identical circumcentres.
#if defined(__CUDA_ARCH__)
  double nx = __fma_rn(e3y, e2z, -__dmul_rn(e3z, e2y));
  double ny = __fma_rn(e3z, e2x, -__dmul_rn(e3x, e2z));
  double nz = __fma_rn(e3x, e2y, -__dmul_rn(e3y, e2x));
#ifndef FM_CC_DIVIDE
  // |n|^-1 by fm_rsqrt and three multiplications instead of sqrt + three IEEE divisions: ~40 fewer
  // instructions and no slow-path call in the sweep (the result differs from the quotient by an ulp
  // or two; every kernel that writes circumcentres comes through here, so they all agree)
  const double nn = __fma_rn(nx, nx, __fma_rn(ny, ny, __dmul_rn(nz, nz)));
  degenerate = nn < 2.220446049250313e-16 * 2.220446049250313e-16;  // "collinear", DV:1339-1341
  const double r = degenerate ? 1.0 : fm_rsqrt(nn);
  return P3{__dmul_rn(nx, r), __dmul_rn(ny, r), __dmul_rn(nz, r)};
#else
  double l = sqrt(__fma_rn(nx, nx, __fma_rn(ny, ny, __dmul_rn(nz, nz))));
  degenerate = l < 2.220446049250313e-16;
  if (degenerate) l = 1.0;
  return P3{nx / l, ny / l, nz / l};
#endif
#else
  double nx = e3y * e2z - e3z * e2y;
  double ny = e3z * e2x - e3x * e2z;
  double nz = e3x * e2y - e3y * e2x;
  double l = sqrt(nx * nx + ny * ny + nz * nz);
  degenerate = l < 2.220446049250313e-16;  // "collinear", DV:1339-1341
  if (degenerate) l = 1.0;
  return P3{nx / l, ny / l, nz / l};
#endif
}
FM_HD inline int enc(int t, int e) { return (t << 2) | e; }
FM_HD inline int geti(const I4& v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }
FM_HD inline void seti(I4& v, int k, int val) {
  if (k == 0) v.x = val; else if (k == 1) v.y = val; else v.z = val;
}
FM_HD inline unsigned hash32(unsigned v) {  // bijective mixer: insertion priority
  v ^= v >> 16; v *= 0x7feb352dU; v ^= v >> 15; v *= 0x846ca68bU; v ^= v >> 16;
  return v;
}

// 63-bit Morton key of a point of the unit sphere (21 bits per axis): the
// internal point order, so that neighbouring generators are neighbours in HBM.
FM_HD inline u64 spread21(u64 v) {
  v &= 0x1fffffULL;
  v = (v | v << 32) & 0x1f00000000ffffULL;
  v = (v | v << 16) & 0x1f0000ff0000ffULL;
  v = (v | v << 8) & 0x100f00f00f00f00fULL;
  v = (v | v << 4) & 0x10c30c30c30c30c3ULL;
  v = (v | v << 2) & 0x1249249249249249ULL;
  return v;
}
FM_HD inline u64 morton_key(double x, double y, double z) {
  const double s = 1048575.5;  // maps [-1,1] onto [0, 2^21)
  double fx = fmin(fmax((x + 1.0) * s, 0.0), 2097151.0);
  double fy = fmin(fmax((y + 1.0) * s, 0.0), 2097151.0);
  double fz = fmin(fmax((z + 1.0) * s, 0.0), 2097151.0);
  return spread21((u64)fx) | (spread21((u64)fy) << 1) | (spread21((u64)fz) << 2);
}

FM_HD inline void defer_item(const DtCtx& c, int item) {
  int pos = FM_COUNT_INC(&c.counters[FM_C_SLOW]);
  c.slow[pos] = item;
}

// queue entries: slot < 2^29 (fm_create refuses more), wanted edge + 1 in bits 29-30
FM_HD inline int qe_tri(int e) { return e & 0x1fffffff; }
FM_HD inline int qe_edge(int e) { return (e >> 29) - 1; }
FM_HD inline int qe_make(int t, int fe) { return t | ((fe + 1) << 29); }
FM_HD inline void queue_active(const DtCtx& c, int t, int stamp, int fe = -1) {
  if (FM_ATOMIC_EXCH32(&c.amark[t], stamp) != stamp) {
    int pos = FM_COUNT_INC(c.qcount);
    c.next_active[pos] = qe_make(t, fe);
  }
}

// ------------------------------------------------------------ point location
// One thread per point.  Replaces the bucket bookkeeping of DV:247-293 and
// DV:303-317: a visibility walk from the remembered slot to the triangle that
// contains the point (three orient tests per step, the edge just crossed is
// not re-tested), then a vote for that triangle.
// ------------------------------------------------------------ insertion levels
// Points are inserted in O(log N) LEVELS of geometrically growing size (a biased randomised
// insertion order): class(p) = min(K, clz(hash(id)) / 2), so a quarter of the points of class
// >= k is of class >= k+1; classes are processed from K down to 0.  Only the points of the
// current level walk and vote, and they start their walk at a triangle incident to the
// nearest ALREADY INSERTED point to their left in the space-filling-curve order (prev[]),
// a few triangles away from the answer.  The reference instead re-locates every uninserted
// point after every insertion (DV:303-317).
FM_HD inline int clz32(unsigned v) {
#if defined(__CUDA_ARCH__)
  return __clz((int)v);
#else
  int n = 0;
  if (v == 0) return 32;
  while (!(v & 0x80000000u)) { v <<= 1; n++; }
  return n;
#endif
}
#ifndef FM_LEVEL_LOG2
#define FM_LEVEL_LOG2 2   // log2 of the growth factor between insertion levels (2: levels grow x4).  x2 levels
                          // (1) need the same number of rounds and a third of the walk steps, and measure
                          // the same: 14.74 against 14.69 ms per build at 2.6M points (with the one-sided
                          // walk start they were slower, 17.1 ms: ~800 walks of > 30 triangles per build)
#endif
FM_HD inline int point_class(long long id, int K) {
  int c = clz32(hash32((unsigned)id ^ 0x9e3779b9u)) / FM_LEVEL_LOG2;
  return c < K ? c : K;
}

// One thread per point of the level: choose the start slot of its walk.
struct KLevelInit {
  FM_HD static void run(const DtCtx& c, int i) {
    int p = c.plist[i];
    if (c.loc[p] < 0) return;  // a bootstrap vertex
    int j = c.prev[p];
    if (c.next_rev) {
      // The curve neighbour on ONE side can be far away in space (the curve jumps between
      // octants); the nearer of the two sides almost never is.  The few points that started far
      // away used to set the duration of the whole locate launch with walks of > 30 triangles.
      int r = c.next_rev[c.n - 1 - p];
      int jr = r >= 0 ? c.n - 1 - r : -1;
      if (j < 0) {
        j = jr;
      } else if (jr >= 0) {
        const P4 q = c.pts[p], a = c.pts[j], b = c.pts[jr];
        if (q.x * b.x + q.y * b.y + q.z * b.z > q.x * a.x + q.y * a.y + q.z * a.z) j = jr;
      }
    }
    // j may itself still be waiting (carried over from its level): then start where ITS walk stands
    c.loc[p] = j < 0 ? 0 : (c.loc[j] < 0 ? c.v2t[j] : c.loc[j]);
  }
};

#if defined(FM_COUNT_STEPS) && !defined(__CUDACC__)
static long long fm_walk_stats[3];  // triangles visited, walks, walks longer than 30 steps
#endif
template <bool EXACT>
struct KLocateVoteT {
  FM_HD static void run(const DtCtx& c, int i) {
    int p = c.plist ? c.plist[i] : i;
    int t = c.loc[p];
    if (t < 0) return;
    int qi;
    P3 q = ldv(c, p, qi);
    int from = -1, steps = 0;
    for (;;) {
      Row r = c.row[t];
      I4 v = r.v, a = r.a;
      int ia, ib, ic;
      P3 pa = ldv(c, v.x, ia), pb = ldv(c, v.y, ib), pc = ldv(c, v.z, ic);
      int next = -1, s = 1;
      // edge opposite vertex k is (v[k+1], v[k+2]); q must be on its left
      if ((a.x >> 2) != from && (s = orient_t<EXACT>(pb, pc, q, ib, ic, qi)) <= 0) next = a.x >> 2;
      else if ((a.y >> 2) != from && (s = orient_t<EXACT>(pc, pa, q, ic, ia, qi)) <= 0) next = a.y >> 2;
      else if ((a.z >> 2) != from && (s = orient_t<EXACT>(pa, pb, q, ia, ib, qi)) <= 0) next = a.z >> 2;
      if (!EXACT && s == 0) {  // inconclusive: the exact kernel resumes the walk from here
        c.loc[p] = t;
        defer_item(c, i);
        return;
      }
      if (next < 0) break;
      from = t;
      t = next;
      if (++steps > (1 << 20)) { *c.err |= FM_E_WALK; return; }
    }
    c.loc[p] = t;
#if defined(FM_COUNT_STEPS) && !defined(__CUDACC__)  // host emulation only: walk statistics
    fm_walk_stats[0] += steps + 1; fm_walk_stats[1] += 1; fm_walk_stats[2] += steps > 30;
#endif
    if (c.locate_only) return;
    u64 key = ((u64)hash32((unsigned)qi) << 32) | (unsigned)p;
    FM_ATOMIC_MIN64(&c.winner[t], key);   // no result wanted: a fire-and-forget reduction
  }
};
typedef KLocateVoteT<true> KLocateVote;

// Sharded point location, second half (every rank, over ALL points of the level): the vote of a
// point for the triangle its walk ended in -- on whichever rank that walk ran.
struct KVote {
  FM_HD static void run(const DtCtx& c, int i) {
    int p = c.plist ? c.plist[i] : i;
    int t = c.loc[p];
    if (t < 0) return;
    u64 key = ((u64)hash32((unsigned)c.pts[p].id) << 32) | (unsigned)p;
    FM_ATOMIC_MIN64(&c.winner[t], key);
  }
};

// One thread per triangle slot alive: 1 if it splits this round (input of the scan that numbers the
// new slots), and the splitting ones go on the winner list -- the split itself then runs over
// that list, not over all slots.
struct KSplitFlag {
  FM_HD static void run(const DtCtx& c, int t) {
    const bool w = c.winner[t] != kNone64;
    c.rank[t] = w ? 1 : 0;
    if (w) c.wlist[FM_COUNT_INC(&c.counters[FM_C_NWIN])] = t;
  }
};

// ------------------------------------------------------------------ 1->3 split
// One thread per entry of the winner list: triangle t won the point p.  (a,b,c) becomes
//   A = (a,b,p) in slot t,  B = (b,c,p) in slot s1,  C = (c,a,p) in slot s1+1,
// the new vertex always last (as DV:977-982), so the edge inherited from the
// parent is edge 2 of each child.  New slots are numbered by the rank of the parent among the
// splitting slots (the scan of KSplitFlag's flags): dense -- the flip rounds that follow work on
// one compact range of fresh slots -- and a function of the input alone, whatever order the
// list was filled in (the ranks of a sharded mesh hold slot-for-slot identical replicas).  A
// neighbour that splits in the same launch is not touched: where ITS child will be follows from
// its rank alone.
struct KSplit {
  FM_HD static int child_slot(const DtCtx& c, int n, int e) {
    // slot of the child of (splitting) triangle n that inherits n's edge e
    int s1 = c.n_tri + 2 * c.rank[n];
    return e == 2 ? n : (e == 0 ? s1 : s1 + 1);
  }
  FM_HD static void run(const DtCtx& c, int i) {
    int t = c.wlist[i];
    u64 w = c.winner[t];
    int p = (int)(w & 0xffffffffULL);
    Row r = c.row[t];
    I4 v = r.v, a = r.a;
    int s1 = c.n_tri + 2 * c.rank[t], s2 = s1 + 1;
    int childOf[3] = {s1, s2, t};  // parent edge k -> child slot
    int outer[3];
    for (int k = 0; k < 3; k++) {
      int ne = geti(a, k), n = ne >> 2, e = ne & 3;
      if (c.winner[n] != kNone64) {
        outer[k] = enc(child_slot(c, n, e), 2);
      } else {
        outer[k] = ne;
        ((int*)&c.row[n].a)[e] = enc(childOf[k], 2);
      }
    }
    // a.w: edges that are Delaunay BY CONSTRUCTION and need no test while the triangle stands (bit k =
    // edge k).  The spokes of a fresh star are: a, b, c, p are in convex position on the sphere and p
    // lies beyond the face (a,b,c), so (a,b,p), (b,c,p), (c,a,p) are the upper faces of a tetrahedron
    // and meet convexly.  Only the edge inherited from the parent (edge 2) can be illegal.
    c.row[t] = Row{I4{v.x, v.y, p, 0}, I4{enc(s1, 1), enc(s2, 0), outer[2], 3}};
    c.row[s1] = Row{I4{v.y, v.z, p, 0}, I4{enc(s2, 1), enc(t, 0), outer[0], 3}};
    c.row[s2] = Row{I4{v.z, v.x, p, 0}, I4{enc(t, 1), enc(s1, 0), outer[1], 3}};
    c.loc[p] = -1;
    c.v2t[p] = t;
    c.v2t[v.z] = s1;  // the only vertex that left slot t
    queue_active(c, t, c.stamp);
    queue_active(c, s1, c.stamp);
    queue_active(c, s2, c.stamp);
    FM_COUNT_INC(&c.counters[4]);
  }
};
// after the split: the winners of the round are consumed (all ones again)
struct KWinnerReset {
  FM_HD static void run(const DtCtx& c, int i) { c.winner[c.wlist[i]] = kNone64; }
};

// ---------------------------------------------------------------- flip rounds
// Sub-round = check -> mark -> flip -> patch.  validate_DT (DV:845-882) tests the
// vertex opposite the new point's triangle with in_circle and flips on INSIDE;
// here every queued triangle tests its three edges and asks for the first
// violated one.  A flip needs both triangles: the pair belongs to the lowest
// voter (atomicMin), the others retry in the next sub-round.
FM_HD inline u64 vote_key(int stamp, int t) {
  return ((u64)(0x7fffffffu - (unsigned)stamp) << 32) | (unsigned)t;
}

template <bool EXACT>
struct KCheckT {
  FM_HD static void run(const DtCtx& c, int i) {
    int t = qe_tri(c.active[i]);
    Row r = c.row[t];
    I4 v = r.v, a = r.a;
#if !defined(FM_CHECK_SEQUENTIAL)
    // All gathers of the tests are requested before the first is consumed (as the sweep does): the
    // opposite-vertex ids together, then the vertex records together -- two dependent levels after
    // the row instead of up to six when each edge's chain waited for the test before it.
    // In a build (c.cc == null: no generator moves while it runs) the edges a.w marks as Delaunay by
    // construction are not tested at all: one test instead of three for a fresh child of a split,
    // two for a triangle a flip made.  A warm repair tests every edge: the marks date from before
    // the generators moved.
    const int skip = c.cc ? 0 : (a.w & 7);
    const int* trow = (const int*)c.row;
    int d0 = (skip & 1) ? -1 : trow[8 * (size_t)(a.x >> 2) + (a.x & 3)];
    int d1 = (skip & 2) ? -1 : trow[8 * (size_t)(a.y >> 2) + (a.y & 3)];
    int d2 = (skip & 4) ? -1 : trow[8 * (size_t)(a.z >> 2) + (a.z & 3)];
    P4 qa = c.pts[v.x], qb = c.pts[v.y], qc = c.pts[v.z];
    const P4 zero = {0, 0, 0, 0};
    P4 q0 = d0 >= 0 ? c.pts[d0] : zero, q1 = d1 >= 0 ? c.pts[d1] : zero, q2 = d2 >= 0 ? c.pts[d2] : zero;
    P3 pa{qa.x, qa.y, qa.z}, pb{qb.x, qb.y, qb.z}, pc{qc.x, qc.y, qc.z};
    int ia = (int)qa.id, ib = (int)qb.id, ic = (int)qc.id;
    int fe = -1, s;
    if (d0 >= 0) {
      s = incircle_t<EXACT>(pa, pb, pc, P3{q0.x, q0.y, q0.z}, ia, ib, ic, (int)q0.id);
      if (!EXACT && s == 0) { defer_item(c, i); return; }
      if (s > 0) fe = 0;
    }
    if (fe < 0 && d1 >= 0) {
      s = incircle_t<EXACT>(pa, pb, pc, P3{q1.x, q1.y, q1.z}, ia, ib, ic, (int)q1.id);
      if (!EXACT && s == 0) { defer_item(c, i); return; }
      if (s > 0) fe = 1;
    }
    if (fe < 0 && d2 >= 0) {
      s = incircle_t<EXACT>(pa, pb, pc, P3{q2.x, q2.y, q2.z}, ia, ib, ic, (int)q2.id);
      if (!EXACT && s == 0) { defer_item(c, i); return; }
      if (s > 0) fe = 2;
    }
#else
    int ia, ib, ic, idd;
    P3 pa = ldv(c, v.x, ia), pb = ldv(c, v.y, ib), pc = ldv(c, v.z, ic);
    int fe = -1;
    for (int k = 0; k < 3; k++) {
      int ne = geti(a, k), n = ne >> 2, e = ne & 3;
      int d = ((const int*)c.row)[8 * (size_t)n + e];
      P3 pd = ldv(c, d, idd);
      int s = incircle_t<EXACT>(pa, pb, pc, pd, ia, ib, ic, idd);
      if (!EXACT && s == 0) { defer_item(c, i); return; }
      if (s > 0) { fe = k; break; }
    }
#endif
    c.active[i] = qe_make(t, fe);   // a coalesced store instead of a scattered one to flip_edge[t]
    if (fe >= 0) {
      u64 key = vote_key(c.stamp, t);
      FM_ATOMIC_MIN64(&c.vote[t], key);
      FM_ATOMIC_MIN64(&c.vote[geti(a, fe) >> 2], key);
    }
  }
};
typedef KCheckT<true> KCheck;

// Warm restart, first sub-round: ONE pass over all triangles that (a) counts triangles that
// turned over when the generators moved (counters[3]; any -> the caller rebuilds), (b) tests
// every edge once -- from its lower-numbered triangle -- and claims violated ones exactly as
// KCheck does, queueing the claimant for the mark/flip/patch phases, and (c) refreshes the
// circumcentre of every triangle (flips refresh theirs).  Replaces validate + queue_all +
// flip_check over T + circumcentre: the three vertices are gathered once.
template <bool EXACT, bool SHARDED>
struct KSweepImpl {
  FM_HD static int slot(const DtCtx& c, int item) { return SHARDED ? c.t_begin + item : item; }
  FM_HD static void run(const DtCtx& c, int item) {
    const int t = slot(c, item);
#if defined(__CUDA_ARCH__)
    // The row load is a cold DRAM access at the head of a chain of three dependent gathers (half
    // of this kernel's stall samples): ask L2 for the row a later wave of blocks will start with.
    if (c.pf_dist > 0 && t + c.pf_dist < c.n_tri) asm volatile("prefetch.global.L2 [%0];" ::"l"(c.row + t + c.pf_dist));
#endif
    body(c, item, t, c.row[t]);
  }
#if defined(__CUDA_ARCH__)
  // as run(), inside a batch of warm iterations: leaves when the batch has halted.  The halt
  // word is loaded AFTER the row load has been issued and both are in flight together; a gate
  // ahead of the row load would put one more memory latency at the head of every thread's chain.
  __device__ static void run_gated(const DtCtx& c, int item) {
    const int t = slot(c, item);
    if (c.pf_dist > 0 && t + c.pf_dist < c.n_tri) asm volatile("prefetch.global.L2 [%0];" ::"l"(c.row + t + c.pf_dist));
    const Row r = c.row[t];
    const int hv = c.counters[FM_T_HALT];   // plain load: written only between launches, and a volatile
                                            // (L2) load of ONE word by five million threads is a hot spot
    if (hv != 0 && r.v.x >= 0) return;   // (the second term only ties the branch to the row load's issue slot)
    body(c, item, t, r);
  }
#endif
  // the row is passed in so that a launcher can have the rows of several items in flight at once
  FM_HD static void body(const DtCtx& c, int item, int t, const Row& r) {
    I4 v = r.v, a = r.a;
    // stage 1: everything that depends only on the two rows -- three vertex records and the
    // three opposite-vertex ids -- is requested before anything is consumed
    P4 qa = c.pts[v.x], qb = c.pts[v.y], qc = c.pts[v.z];
    const int* trow = (const int*)c.row;
    int n0 = a.x >> 2, n1 = a.y >> 2, n2 = a.z >> 2;
    int d0 = n0 > t ? trow[8 * (size_t)n0 + (a.x & 3)] : -1;
    int d1 = n1 > t ? trow[8 * (size_t)n1 + (a.y & 3)] : -1;
    int d2 = n2 > t ? trow[8 * (size_t)n2 + (a.z & 3)] : -1;
    // stage 2: the opposite vertices
    P4 z = {0, 0, 0, 0};
    P4 q0 = d0 >= 0 ? c.pts[d0] : z;
    P4 q1 = d1 >= 0 ? c.pts[d1] : z;
    P4 q2 = d2 >= 0 ? c.pts[d2] : z;
    P3 pa{qa.x, qa.y, qa.z}, pb{qb.x, qb.y, qb.z}, pc{qc.x, qc.y, qc.z};
    int ia = (int)qa.id, ib = (int)qb.id, ic = (int)qc.id;
    int so = orient_t<EXACT>(pa, pb, pc, ia, ib, ic);
    int fe = -1, s = -1;
    bool unsure = !EXACT && so == 0;
    if (!unsure && d0 >= 0) {
      s = incircle_t<EXACT>(pa, pb, pc, P3{q0.x, q0.y, q0.z}, ia, ib, ic, (int)q0.id);
      if (s > 0) fe = 0;
      unsure = !EXACT && s == 0;
    }
    if (!unsure && fe < 0 && d1 >= 0) {
      s = incircle_t<EXACT>(pa, pb, pc, P3{q1.x, q1.y, q1.z}, ia, ib, ic, (int)q1.id);
      if (s > 0) fe = 1;
      unsure = !EXACT && s == 0;
    }
    if (!unsure && fe < 0 && d2 >= 0) {
      s = incircle_t<EXACT>(pa, pb, pc, P3{q2.x, q2.y, q2.z}, ia, ib, ic, (int)q2.id);
      if (s > 0) fe = 2;
      unsure = !EXACT && s == 0;
    }
    if (unsure) { defer_item(c, item); return; }
    if (so < 0) FM_ATOMIC_ADD32(&c.counters[3], 1);
    if (fe >= 0 && SHARDED) {
      c.req[FM_COUNT_INC(&c.counters[FM_C_REQ])] = (t << 2) | fe;
    } else if (fe >= 0) {
      u64 key = vote_key(c.stamp, t);
      FM_ATOMIC_MIN64(&c.vote[t], key);
      FM_ATOMIC_MIN64(&c.vote[geti(a, fe) >> 2], key);
      queue_active(c, t, c.stamp, fe);
    }
    if (c.cc) {
      bool dg;
      P3 cc = circumcentre(pa, pb, pc, dg);
      if (dg) *c.err |= FM_E_TOPOLOGY;
      c.cc[t] = P4{cc.x, cc.y, cc.z, 0};
      if (SHARDED) {
        int lo = v.x < v.y ? (v.x < v.z ? v.x : v.z) : (v.y < v.z ? v.y : v.z);
        int hi = v.x > v.y ? (v.x > v.z ? v.x : v.z) : (v.y > v.z ? v.y : v.z);
        if (lo < c.blk_begin || hi >= c.blk_end)  // a vertex lives on another rank
          c.bnd[FM_COUNT_INC(&c.counters[FM_C_BND])] = t;
      }
    }
  }
};
template <bool EXACT> struct KSweepT : KSweepImpl<EXACT, false> {};    // the whole mesh
template <bool EXACT> struct KSweepShT : KSweepImpl<EXACT, true> {};   // this rank's slots of a sharded mesh
typedef KSweepT<true> KSweep;

// Sharded sweep, second half (every rank, over ALL ranks' requests): what KSweep does on the spot
// when it finds a violated edge -- the wanted edge, the two votes, the queue entry.
struct KApply {
  FM_HD static void run(const DtCtx& c, int msg) {
    int t = msg >> 2, fe = msg & 3;
    u64 key = vote_key(c.stamp, t);
    FM_ATOMIC_MIN64(&c.vote[t], key);
    FM_ATOMIC_MIN64(&c.vote[geti(c.row[t].a, fe) >> 2], key);
    queue_active(c, t, c.stamp, fe);
  }
};

struct KMark {
  FM_HD static void run(const DtCtx& c, int i) {
    const int e = c.active[i];
    const int t = qe_tri(e), fe = qe_edge(e);
    if (fe < 0) return;            // most entries: nothing but the (coalesced) queue was read
    int n = geti(c.row[t].a, fe) >> 2;
    u64 key = vote_key(c.stamp, t);
    if (c.vote[t] == key && c.vote[n] == key) {
      c.flip_edge[t] = fe;         // for the flip and patch phases (winners only)
      c.role[t] = (c.stamp << 2) | 1;
      c.role[n] = (c.stamp << 2) | 2;
      // Only a fraction of the queue flips (measured: 6 of 32 lanes in a late insertion round);
      // one thread per queue entry would run the flip's chain of dependent gathers in nearly
      // empty warps.  The winners go on a compact list instead.
      c.fliplist[FM_COUNT_INC(&c.counters[FM_C_NFLIP])] = t;
    } else {
      queue_active(c, t, c.stamp + 1);  // lost the vote: try again next sub-round
    }
  }
};

// Flip of the edge opposite vertex k of t = (a,b,c) (rotated so that a = v[k]),
// shared with n whose opposite vertex is d:
//   X = (a,b,d) -> slot t      Y = (a,d,c) -> slot n           (flip22, DV:1098-1149)
// Outer edges move:  (t, c's index) -> (t,2)   (t, b's index) -> (n,1)
//                    (n, c's index) -> (t,0)   (n, b's index) -> (n,0)
struct FlipGeom {
  int t, n, k, e;
  int so[4], io[4];  // old (slot, edge) of the four outer edges
  int sn[4], in[4];  // new (slot, edge)
};
FM_HD inline FlipGeom flip_geom(int t, int k, int n, int e) {
  FlipGeom g;
  g.t = t; g.n = n; g.k = k; g.e = e;
  int kb = (k + 1) % 3, kc = (k + 2) % 3;  // b, c in t
  int ec = (e + 1) % 3, eb = (e + 2) % 3;  // n = (d, c, b) from e
  g.so[0] = t; g.io[0] = kc; g.sn[0] = t; g.in[0] = 2;
  g.so[1] = t; g.io[1] = kb; g.sn[1] = n; g.in[1] = 1;
  g.so[2] = n; g.io[2] = ec; g.sn[2] = t; g.in[2] = 0;
  g.so[3] = n; g.io[3] = eb; g.sn[3] = n; g.in[3] = 0;
  return g;
}

struct KFlip {   // over fliplist[0 .. counters[FM_C_NFLIP])
  FM_HD static void run(const DtCtx& c, int i) {
    int t = c.fliplist[i];
    int k = c.flip_edge[t];
    Row rt = c.row[t];
    I4 vt = rt.v, at = rt.a;
    int ne = geti(at, k), n = ne >> 2, e = ne & 3;
    Row rn = c.row[n];
    I4 vn = rn.v, an = rn.a;
    FlipGeom g = flip_geom(t, k, n, e);
    int va = geti(vt, k), vb = geti(vt, (k + 1) % 3), vc = geti(vt, (k + 2) % 3), vd = geti(vn, e);
    int outer[4];
    for (int j = 0; j < 4; j++) {
      int oe = geti(g.so[j] == t ? at : an, g.io[j]);
      int m = oe >> 2, em = oe & 3;
      outer[j] = oe;
      int nw = enc(g.sn[j], g.in[j]);
      if ((c.role[m] >> 2) == c.stamp) {
        c.mailbox[3 * (size_t)m + em] = ((u64)(unsigned)c.stamp << 32) | (unsigned)nw;
      } else {
        ((int*)&c.row[m].a)[em] = nw;
      }
    }
    // a.w: the new diagonal (a,d) is legal by the very test that ordered the flip (edge 1 of X, edge 2 of Y)
    c.row[t] = Row{I4{va, vb, vd, e}, I4{outer[2], enc(n, 2), outer[0], 2}};  // v.w keeps e for KPatch
    c.row[n] = Row{I4{va, vd, vc, 0}, I4{outer[3], outer[1], enc(t, 1), 4}};
    c.v2t[vb] = t;  // b left slot n, c left slot t (a and d only gained a triangle)
    c.v2t[vc] = n;
    if (c.ringc) {  // the rings of all four vertices changed
      c.ringc[8 * (size_t)va] = -1; c.ringc[8 * (size_t)vb] = -1;
      c.ringc[8 * (size_t)vc] = -1; c.ringc[8 * (size_t)vd] = -1;
    }
    if (c.cc) {
      P3 pa = ldp(c, va), pb = ldp(c, vb), pc = ldp(c, vc), pd = ldp(c, vd);
      bool dg1, dg2;
      P3 c1 = circumcentre(pa, pb, pd, dg1), c2 = circumcentre(pa, pd, pc, dg2);
      if (dg1 || dg2) *c.err |= FM_E_TOPOLOGY;
      c.cc[t] = P4{c1.x, c1.y, c1.z, 0};
      c.cc[n] = P4{c2.x, c2.y, c2.z, 0};
    }
    queue_active(c, t, c.stamp + 1);
    queue_active(c, n, c.stamp + 1);
    FM_COUNT_INC(&c.counters[2]);
  }
};

// After all flips of the sub-round: a pair whose outer neighbour flipped too
// finds that neighbour's new (slot, edge) in the mailbox of its OLD (slot, edge).
// The pre-flip indices come back from flip_edge[t] (k) and row[t].v.w (e).
struct KPatch {  // over fliplist[0 .. counters[FM_C_NFLIP])
  FM_HD static void run(const DtCtx& c, int i) {
    int t = c.fliplist[i];
    int k = c.flip_edge[t];
    int n = c.row[t].a.y >> 2;
    int e = c.row[t].v.w;
    FlipGeom g = flip_geom(t, k, n, e);
    for (int j = 0; j < 4; j++) {
      u64 msg = c.mailbox[3 * (size_t)g.so[j] + g.io[j]];
      if ((int)(msg >> 32) == c.stamp) ((int*)&c.row[g.sn[j]].a)[g.in[j]] = (int)(msg & 0xffffffffULL);
    }
  }
};

// ------------------------------------------------------------- warm restart
// After the generators moved (Lloyd), the old connectivity is still a valid
// triangulation as long as no triangle turned over; it is then repaired by flip
// rounds alone instead of being rebuilt.  KValidate counts triangles that are
// not strictly CCW any more (counters[3]).
struct KValidate {
  FM_HD static void run(const DtCtx& c, int t) {
    I4 v = c.row[t].v;
    int ia, ib, ic;
    P3 pa = ldv(c, v.x, ia), pb = ldv(c, v.y, ib), pc = ldv(c, v.z, ic);
    if (orient(pa, pb, pc, ia, ib, ic) <= 0)
      FM_ATOMIC_ADD32(&c.counters[3], 1);
  }
};

}  // namespace fm
