// dt_driver.h -- host-side control of the Delaunay rounds, written once against
// a small "backend" concept so that the CUDA library (fishman_b200.cu) and the
// CPU logic tests (tests/emu) run the SAME round structure:
//
//   Backend::launch<K>(ctx, n)            run body K for thread ids [0, n)
//   Backend::launch_filtered<K>(ctx, n)   K<false> (fp64 filters only) for [0, n), then K<true>
//                                         (exact) for the items K<false> deferred
//   Backend::locate_round(ctx, m)         point location + votes of the m points of the level
//                                         (= launch_filtered<KLocateVoteT>, or its sharded form)
//   Backend::fill64(ptr, value, count)    device memset of 64-bit words
//   Backend::fill32(ptr, value, count)
//   Backend::launch_counted<K>(ctx, idx, upper)  K::run(ctx, i) for i in [0, counters[idx]), a count the host
//                                         only knows an upper bound of
//   Backend::exclusive_scan_async(ptr, n) in place over ints, no host round trip
//   Backend::read_ints(dst, src, n)       device -> host (synchronising)
//   Backend::write_ints(dst, src, n)      host -> device
//   Backend::argmax_dot(ctx, d)           index of the point with the largest p . d
//   Backend::fetch_point(ctx, i)          coordinates of point i on the host
//   Backend::fetch_orig(ctx, i)
//   Backend::write_tris(ctx, t0, cnt, tri, adj)
//   Backend::prev_coarser(ctx, thr, prev) prev[i] = largest j <= i with point_class(j) > thr, or -1; returns
//                                         DtCtx::next_rev (the same to the right, reversed) or null
//   Backend::bump_stamp(stamp)            advance the sub-round stamp past every value in use (host mirror
//                                         *stamp, or the device-resident counters[FM_C_STAMP])
//   Backend::flip_loop(ctx, list, cur, skip_check, stamp, stats, msg, len, nact_hint)
//                                         flip sub-rounds until the queue is empty; nact_hint =
//                                         length of the first queue when the host knows it, else -1
//
// Replaces: the bootstrap get_three_DVT_indices / init_DTs
// (/root/reference/src/delaunay_voronoi_mod.F90:595-694) and the loop structure
// of delaunay_triangulation (DV:226-356).  The reference starts from three random
// points plus three VIRTUAL antipodal vertices that it must delete again at the
// end (DV:772-843); this driver starts from four REAL points whose tetrahedron
// contains the origin, so there is nothing to delete and the triangle count is
// 2N-4 from the first round on.
#pragma once
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "dt_kernels.cuh"

namespace fm {

struct DtStats {
  int rounds, subrounds, flips, max_active;
  long long located;
};

// Chooses the four bootstrap points and writes the four faces.  Returns 0, or 1
// with msg filled when the points do not surround the origin.
template <class B>
int dt_bootstrap(B& be, DtCtx& c, int boot[4], char* msg, size_t msglen) {
  static const double dirs[4][3] = {{1, 1, 1}, {1, -1, -1}, {-1, 1, -1}, {-1, -1, 1}};
  static const int F[4][3] = {{0, 1, 2}, {0, 3, 1}, {1, 3, 2}, {0, 2, 3}};
  P3 p[4];
  int id[4];
  bool ok = false;
  // attempts 0..7: the extreme points along the axes of a regular tetrahedron,
  // first axis-aligned, then in seven fixed pseudo-random rotations (symmetric
  // sets tie along special axes); attempt 8: a, the point most opposite to a,
  // the point farthest from their great circle, the point most opposite to all.
  const int kLast = 8;
  unsigned long long lcg = 0x9E3779B97F4A7C15ULL;
  for (int attempt = 0; attempt <= kLast && !ok; attempt++) {
    bool distinct = true;
    double R[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    if (attempt > 0 && attempt < kLast) {
      double q[4], nq = 0;
      for (int j = 0; j < 4; j++) {
        lcg = lcg * 6364136223846793005ULL + 1442695040888963407ULL;
        q[j] = (double)((lcg >> 11) & 0xfffff) / 524288.0 - 1.0;
        nq += q[j] * q[j];
      }
      nq = sqrt(nq) + 1e-300;
      double w = q[0] / nq, x = q[1] / nq, y = q[2] / nq, z = q[3] / nq;
      R[0][0] = 1 - 2 * (y * y + z * z); R[0][1] = 2 * (x * y - z * w); R[0][2] = 2 * (x * z + y * w);
      R[1][0] = 2 * (x * y + z * w); R[1][1] = 1 - 2 * (x * x + z * z); R[1][2] = 2 * (y * z - x * w);
      R[2][0] = 2 * (x * z - y * w); R[2][1] = 2 * (y * z + x * w); R[2][2] = 1 - 2 * (x * x + y * y);
    }
    for (int k = 0; k < 4; k++) {
      double d[3];
      if (attempt < kLast || k == 0) {
        for (int r = 0; r < 3; r++)
          d[r] = R[r][0] * dirs[k][0] + R[r][1] * dirs[k][1] + R[r][2] * dirs[k][2];
      } else if (k == 1) {
        d[0] = -p[0].x; d[1] = -p[0].y; d[2] = -p[0].z;
      } else if (k == 2) {
        d[0] = p[0].y * p[1].z - p[0].z * p[1].y;
        d[1] = p[0].z * p[1].x - p[0].x * p[1].z;
        d[2] = p[0].x * p[1].y - p[0].y * p[1].x;
        if (d[0] == 0.0 && d[1] == 0.0 && d[2] == 0.0) { d[0] = p[0].y; d[1] = p[0].z; d[2] = p[0].x; }
      } else {
        d[0] = -(p[0].x + p[1].x + p[2].x); d[1] = -(p[0].y + p[1].y + p[2].y);
        d[2] = -(p[0].z + p[1].z + p[2].z);
      }
      boot[k] = be.argmax_dot(c, d);
      if (attempt == kLast && k == 2) {  // the far side of the great circle may hold the better point
        double nd[3] = {-d[0], -d[1], -d[2]};
        int alt = be.argmax_dot(c, nd);
        P3 q1 = be.fetch_point(c, boot[k]), q2 = be.fetch_point(c, alt);
        double v1 = q1.x * d[0] + q1.y * d[1] + q1.z * d[2];
        double v2 = q2.x * nd[0] + q2.y * nd[1] + q2.z * nd[2];
        if (v2 > v1) boot[k] = alt;
      }
      p[k] = be.fetch_point(c, boot[k]);
      id[k] = be.fetch_orig(c, boot[k]);
      for (int j = 0; j < k; j++)
        if (boot[j] == boot[k]) distinct = false;
    }
    if (!distinct) continue;
    if (orient(p[0], p[1], p[2], id[0], id[1], id[2]) < 0) {
      int tb = boot[1]; boot[1] = boot[2]; boot[2] = tb;
      P3 tp = p[1]; p[1] = p[2]; p[2] = tp;
      int ti = id[1]; id[1] = id[2]; id[2] = ti;
    }
    // closed, consistently oriented surface of the tetrahedron (p0,p1,p2,p3):
    // the origin is strictly inside iff every face is CCW seen from outside
    ok = true;
    for (int f = 0; f < 4; f++) {
      int a = F[f][0], b = F[f][1], d = F[f][2];
      if (orient(p[a], p[b], p[d], id[a], id[b], id[d]) <= 0) ok = false;
    }
  }
  if (!ok && c.n <= 64) {
    // tiny sets: try every quadruple (Caratheodory: if the centre is inside the hull of
    // the points, four of them already enclose it)
    P3 q[64];
    int qi[64];
    for (int i = 0; i < c.n; i++) { q[i] = be.fetch_point(c, i); qi[i] = be.fetch_orig(c, i); }
    for (int a = 0; a < c.n && !ok; a++)
      for (int b = a + 1; b < c.n && !ok; b++)
        for (int d = b + 1; d < c.n && !ok; d++)
          for (int e = d + 1; e < c.n && !ok; e++) {
            int s[4] = {a, b, d, e};
            if (orient(q[s[0]], q[s[1]], q[s[2]], qi[s[0]], qi[s[1]], qi[s[2]]) < 0) { s[1] = d; s[2] = b; }
            bool good = true;
            for (int f = 0; f < 4 && good; f++) {
              int u = s[F[f][0]], v = s[F[f][1]], w = s[F[f][2]];
              good = orient(q[u], q[v], q[w], qi[u], qi[v], qi[w]) > 0;
            }
            if (good) { ok = true; for (int k = 0; k < 4; k++) boot[k] = s[k]; }
          }
  }
  if (!ok) {
    snprintf(msg, msglen,
             "bootstrap: no four points enclosing the sphere centre were found (points confined "
             "to a hemisphere?)");
    return 1;
  }
  I4 tri[4], adj[4];
  for (int f = 0; f < 4; f++) tri[f] = I4{boot[F[f][0]], boot[F[f][1]], boot[F[f][2]], 0};
  for (int f = 0; f < 4; f++) {
    int a[3] = {-1, -1, -1};
    for (int k = 0; k < 3; k++) {
      int u = F[f][(k + 1) % 3], v = F[f][(k + 2) % 3];  // edge opposite vertex k
      for (int g = 0; g < 4; g++)
        for (int e = 0; e < 3; e++)
          if (F[g][(e + 1) % 3] == v && F[g][(e + 2) % 3] == u) a[k] = enc(g, e);
    }
    adj[f] = I4{a[0], a[1], a[2], 0};
  }
  be.write_tris(c, 0, 4, tri, adj);
  return 0;
}

// Flip sub-rounds until the queue runs dry, host-driven (one launch per phase).  The
// current queue is list[cur] with its length in counters[cur]; *stamp is the stamp it was
// queued with.  skip_check: the first sub-round's check phase was already done by the
// producer of the queue (KSweep).  This generic form serves the CPU logic tests; the CUDA
// backend runs the same four bodies inside one cooperative kernel (Backend::flip_loop).
template <class B>
int dt_flip_loop_host(B& be, DtCtx& c, int* list[2], int cur, bool skip_check, int* stamp, DtStats* st,
                      char* msg, size_t msglen, int /*nact_hint*/ = -1) {
  int host[8];
  for (;;) {
    be.read_ints(host, c.counters, 8);
    int nact = host[cur];
    st->flips = host[2];
    if (nact == 0) break;
    st->subrounds++;
    if (nact > st->max_active) st->max_active = nact;
    c.stamp = *stamp;
    c.active = list[cur];
    c.n_active = nact;
    c.next_active = list[cur ^ 1];
    c.qcount = c.counters + (cur ^ 1);
    if (!skip_check) be.template launch_filtered<KCheckT>(c, nact);
    skip_check = false;
    be.template launch<KMark>(c, nact);
    be.template launch_counted<KFlip>(c, FM_C_NFLIP, nact);
    be.template launch_counted<KPatch>(c, FM_C_NFLIP, nact);
    int zero = 0;
    be.write_ints(c.counters + cur, &zero, 1);
    be.write_ints(c.counters + FM_C_NFLIP, &zero, 1);
    (*stamp)++;
    cur ^= 1;
    if (*stamp > (1 << 28)) {
      snprintf(msg, msglen, "flip rounds did not terminate");
      return 1;
    }
  }
  return 0;
}

// Number of coarseness classes above the finest one: the coarsest level holds ~<= 256 points.
inline int dt_level_count(int n) {
  int K = 0;
  while (K < 24 / FM_LEVEL_LOG2 && ((long long)256 << (FM_LEVEL_LOG2 * K)) < n) K++;
  return K;
}

// Runs insertion rounds, level by level, until every point is a vertex.  Preconditions: the
// four bootstrap faces are in slots 0..3, loc[] is 0 for every point but the bootstrap ones
// (-1), vote[] is all ones, role/amark/mailbox hold stamps older than *stamp (zero at first
// use).  listA/listB: two int queues of 2n-4 entries.  lvl_order: the points sorted by
// (level, internal index); lvl_off (host, K+2 entries): start of each level in lvl_order,
// level L = class K-L; prev: scratch of n ints; boot: the four bootstrap points.
template <class B>
int dt_run(B& be, DtCtx& c, int* listA, int* listB, const int* lvl_order, const int* lvl_off, int K,
           int* prev, const int boot[4], int* stamp, DtStats* st, char* msg, size_t msglen) {
  c.n_tri = 4;
  c.wlist = c.fliplist;  // n-1 entries: every split of a build fits; the flip rounds use it after the split is done
  be.fill64(c.winner, kNone64, 2 * (size_t)c.n - 4);  // once per build: a round consumes (resets) the winners it used
  memset(st, 0, sizeof *st);
  int zero[15] = {0}, host[8];
  int* list[2] = {listA, listB};
  be.write_ints(c.counters, zero, 15);
  // Every build starts on a stamp newer than anything left in amark / vote / role / mailbox.  A
  // warm repair that found a turned-over triangle has stamped amark[] and vote[] with the
  // current stamp and left (the flip loop returns at once without advancing it); a rebuild on
  // that same stamp would find its freshly split triangles "already queued" and never check them.
  be.bump_stamp(stamp);
  int boot_class[4];
  for (int k = 0; k < 4; k++) boot_class[k] = point_class(be.fetch_orig(c, boot[k]), K);
  c.level_max = K;
  // A level is not driven to completion.  A round inserts at most one point per triangle, so
  // the mesh can only triple per round and a level four times the size of what is already
  // there saturates: most of its points lose their vote.  After kRoundsPerLevel rounds (one)
  // the losers are carried into the next level and vote again there; only the last level runs
  // dry.  In effect every round inserts about as many points as there are triangles, the
  // early levels merely keep the number of WALKING points near that number.  Driving each
  // level dry took 48 rounds at 2.6M points (six per level, most of them nearly empty, each
  // with the fixed cost of a round: seven launches, one host read, two or more flip
  // sub-rounds) against 22 now (log3 of the triangle count is 13).  Measured at 2.6M points,
  // rounds per level 1 / 2 / 3 / 4: 14.8 / 16.0 / 16.9 / 18.1 ms per build (until dry: 19.7).
  static const int kRoundsPerLevel = [] {  // FM_ROUNDS_PER_LEVEL: tuning switch
    const char* e = getenv("FM_ROUNDS_PER_LEVEL");
    int v = e ? atoi(e) : 1;
    return v >= 1 ? v : 1;
  }();
  int carry = 0;  // points of earlier levels still waiting
  for (int L = 0; L <= K; L++) {
    int m = lvl_off[L + 1] - lvl_off[L];
    int remaining = m;
    for (int k = 0; k < 4; k++) remaining -= (boot_class[k] == K - L);
    if (remaining + carry <= 0) continue;
    c.level_thr = K - L;
    if (L > 0 && remaining > 0) {  // start every walk of this level's points next to an already placed point
      c.plist = lvl_order + lvl_off[L];
      c.n_plist = m;
      c.prev = prev;
      c.next_rev = be.prev_coarser(c, K - L, prev);  // fills prev[]; returns the right-hand array or null
      be.template launch<KLevelInit>(c, m);
    }
    // the voters: this level, and with leftovers all levels before it (inserted points leave at once)
    const int lo = carry > 0 ? 0 : L;
    c.plist = lvl_order + lvl_off[lo];
    m = lvl_off[L + 1] - lvl_off[lo];
    c.n_plist = m;
    remaining += carry;
    for (int rounds_here = 0; remaining > 0 && (L == K || rounds_here < kRoundsPerLevel); rounds_here++) {
      st->rounds++;
      be.locate_round(c, m);
      be.template launch<KSplitFlag>(c, c.n_tri);   // flags for the scan + the winner list
      be.exclusive_scan_async(c.rank, c.n_tri);
      // ONE readback per round: the error bits and the number of splits (= length of the winner list)
      be.read_ints(host, c.counters + FM_C_ERR, 2);
      static_assert(FM_C_NWIN == FM_C_ERR + 1, "one readback for both");
      const int nsplit = host[1];
      if (host[0]) {
        snprintf(msg, msglen, "point location failed (error bits %d): duplicate points?", host[0]);
        return 1;
      }
      if (nsplit <= 0) {
        snprintf(msg, msglen, "no triangle accepted a point although %d remain in level %d", remaining, L);
        return 1;
      }
      c.stamp = *stamp;
      c.next_active = list[0];
      c.qcount = c.counters;
      // the splits run over the winner list; its length is on the device, the host knows it too by now
      be.template launch_counted<KSplit>(c, FM_C_NWIN, nsplit);
      // the winners are consumed: all ones again, through the list when the round was small
      if ((long long)nsplit * 16 < c.n_tri) be.template launch_counted<KWinnerReset>(c, FM_C_NWIN, nsplit);
      else be.fill64(c.winner, kNone64, (size_t)c.n_tri);
      {
        int zero1 = 0;
        be.write_ints(c.counters + FM_C_NWIN, &zero1, 1);
      }
      if (getenv("FM_DEBUG_ROUNDS")) {
        fprintf(stderr, "level %d round %d: voters %d pool %d inserted %d triangles %d", L, rounds_here, m,
                remaining, nsplit, c.n_tri);
#if defined(FM_COUNT_STEPS) && !defined(__CUDACC__)
        fprintf(stderr, "  walks %lld visited %lld long %lld", fm_walk_stats[1], fm_walk_stats[0], fm_walk_stats[2]);
        fm_walk_stats[0] = fm_walk_stats[1] = fm_walk_stats[2] = 0;
#endif
        fprintf(stderr, "\n");
      }
      remaining -= nsplit;
      c.n_tri += 2 * nsplit;
      if (be.flip_loop(c, list, 0, false, stamp, st, msg, msglen, 3 * nsplit)) return 1;
    }
    carry = remaining;
  }
  c.plist = nullptr;
  be.read_ints(host, c.counters, 8);
  st->flips = host[2];
  if (host[5]) { st->subrounds = host[5]; st->max_active = host[6]; }
  return 0;
}

// Warm restart: the connectivity of the previous triangulation with moved vertices, repaired
// by flips.  Launches only (no host synchronisation): counters[3] > 0 afterwards means a
// triangle turned over -- the flip loop then did nothing and the caller must rebuild.
template <class B>
int dt_repair_async(B& be, DtCtx& c, int* listA, int* listB, int* stamp, DtStats* st, char* msg,
                    size_t msglen) {
  int zero[15] = {0};
  int* list[2] = {listA, listB};
  memset(st, 0, sizeof *st);
  be.write_ints(c.counters, zero, 15);
  c.stamp = *stamp;
  c.next_active = list[0];
  c.qcount = c.counters;
  be.template launch_filtered<KSweepT>(c, c.n_tri);
  return be.flip_loop(c, list, 0, true, stamp, st, msg, msglen, -1);
}

}  // namespace fm
