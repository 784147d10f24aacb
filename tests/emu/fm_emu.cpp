// fm_emu.cpp -- HOST execution of the kernel bodies in fishman_b200/csrc/
// (dt_kernels.cuh, dt_driver.h, vd_kernels.cuh), thread by thread, for the
// CPU-side logic tests (tests/test_emu_*.py, run with -m "not gpu").
//
// TEST INFRASTRUCTURE ONLY.  This is not a fallback: the product library
// (fishman_b200/libfishman_b200.so) contains no host execution path and fails
// loudly without a CUDA device.  Nothing under fishman_b200/ loads this file.
#include <algorithm>
#include <numeric>
#include <vector>

#include "../../fishman_b200/csrc/dt_driver.h"

using namespace fm;

struct HostBackend {
  template <class K>
  void launch(const DtCtx& c, int n) {
    for (int i = 0; i < n; i++) K::run(c, i);
  }
  template <class K>
  void launch_counted(const DtCtx& c, int idx, int /*upper*/) {
    for (int i = 0, n = c.counters[idx]; i < n; i++) K::run(c, i);
  }
  // both halves of the filtered-then-exact scheme, as on the device
  template <template <bool> class K>
  void launch_filtered(const DtCtx& c, int n) {
    for (int i = 0; i < n; i++) K<false>::run(c, i);
    int m = c.counters[FM_C_SLOW];
    n_deferred += m;
    for (int i = 0; i < m; i++) K<true>::run(c, c.slow[i]);
    c.counters[FM_C_SLOW] = 0;
  }
  long long n_deferred = 0;
  // point location of a level: whole, or (shard_ranks > 1) the way the ranks of a sharded mesh
  // do it -- each rank walks its slice of the level recording loc[] only, then all votes are cast
  int shard_ranks = 1;
  void locate_round(DtCtx& c, int m) {
    if (shard_ranks <= 1) { launch_filtered<KLocateVoteT>(c, m); return; }
    const int* plist = c.plist;
    for (int r = 0; r < shard_ranks; r++) {
      int mb = (int)((long long)m * r / shard_ranks), me = (int)((long long)m * (r + 1) / shard_ranks);
      DtCtx cs = c;
      cs.plist = plist + mb;
      cs.locate_only = 1;
      launch_filtered<KLocateVoteT>(cs, me - mb);
    }
    launch<KVote>(c, m);
  }
  void fill64(u64* p, u64 v, size_t n) { std::fill(p, p + n, v); }
  void fill32(int* p, int v, size_t n) { std::fill(p, p + n, v); }
  void exclusive_scan_async(int* p, int n) {
    int acc = 0;
    for (int i = 0; i < n; i++) { int v = p[i]; p[i] = acc; acc += v; }
  }
  void read_ints(int* dst, const int* src, int n) { memcpy(dst, src, sizeof(int) * n); }
  void write_ints(int* dst, const int* src, int n) { memcpy(dst, src, sizeof(int) * n); }
  int argmax_dot(const DtCtx& c, const double* d) {
    int best = 0;
    double bv = -1e300;
    for (int i = 0; i < c.n; i++) {
      double v = c.pts[i].x * d[0] + c.pts[i].y * d[1] + c.pts[i].z * d[2];
      if (v > bv) { bv = v; best = i; }
    }
    return best;
  }
  P3 fetch_point(const DtCtx& c, int i) { return ldp(c, i); }
  int fetch_orig(const DtCtx& c, int i) { return (int)c.pts[i].id; }
  void write_tris(const DtCtx& c, int t0, int cnt, const I4* tri, const I4* adj) {
    for (int i = 0; i < cnt; i++) {
      c.row[t0 + i] = Row{tri[i], adj[i]};
      c.v2t[tri[i].x] = c.v2t[tri[i].y] = c.v2t[tri[i].z] = t0 + i;
    }
  }
  std::vector<int> next_rev;
  const int* prev_coarser(const DtCtx& c, int thr, int* prev) {
    int last = -1;
    for (int i = 0; i < c.n; i++) {
      if (point_class(c.pts[i].id, c.level_max) > thr) last = i;
      prev[i] = last;
    }
    next_rev.assign(c.n, -1);
    last = -1;
    for (int r = 0; r < c.n; r++) {  // r = n-1-i: the same scan over the reversed order
      if (point_class(c.pts[c.n - 1 - r].id, c.level_max) > thr) last = r;
      next_rev[r] = last;
    }
    return next_rev.data();
  }
  int flip_loop(DtCtx& c, int* list[2], int cur, bool skip_check, int* stamp, DtStats* st, char* msg,
                size_t msglen, int) {
    // as fm_flip_coop on the device: a turned-over triangle leaves everything to the rebuild
    if (emulate_abort && skip_check && c.counters[3] > 0) return 0;
    return dt_flip_loop_host(*this, c, list, cur, skip_check, stamp, st, msg, msglen);
  }
  bool emulate_abort = false;
  bool skip_bump = false;  // test switch: reproduce a rebuild that starts on a used stamp
  void bump_stamp(int* stamp) { if (!skip_bump) *stamp += 2; }
};

static char g_err[512];

extern "C" const char* fme_last_error() { return g_err; }

// One mesh on the host: the arrays of the CUDA handle as std::vectors.
struct EmuMesh {
  int n = 0, T = 0, stamp = 1, err = 0;
  int counters[16] = {0};
  std::vector<int> orig, loc, rank, flip_edge, role, amark, listA, listB, fliplist, v2t, slow;
  std::vector<P4> pts;
  std::vector<Row> row;
  std::vector<u64> winner, vote, mailbox;
  DtCtx c;
  HostBackend be;
  DtStats st;

  int build(int n_, const double* xyz, int do_sort) {
    n = n_; T = 2 * n - 4;
    orig.resize(n);
    std::iota(orig.begin(), orig.end(), 0);
    if (do_sort) {
      std::vector<u64> key(n);
      for (int i = 0; i < n; i++) key[i] = morton_key(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
      std::stable_sort(orig.begin(), orig.end(), [&](int a, int b) { return key[a] < key[b]; });
    }
    pts.resize(n);
    for (int i = 0; i < n; i++) pts[i] = P4{xyz[3 * orig[i]], xyz[3 * orig[i] + 1], xyz[3 * orig[i] + 2], orig[i]};
    row.assign(T, Row());
    loc.assign(n, 0); rank.assign(T + 1, 0); flip_edge.assign(T, 0); role.assign(T, 0); amark.assign(T, 0);
    listA.assign(T, 0); listB.assign(T, 0); fliplist.assign(T / 2 + 1, 0); v2t.assign(n, -1); slow.assign(T + 1, 0);
    winner.assign(T, 0); vote.assign(T, kNone64); mailbox.assign(3 * (size_t)T, 0);
    return run();
  }
  // bootstrap + insertion rounds over the current pts[]; callable again after the points moved
  // (the in-place rebuild of the CUDA handle: nothing but loc[] is reset)
  int run() {
    std::fill(loc.begin(), loc.end(), 0);
    memset(&c, 0, sizeof c);
    c.n = n; c.pts = pts.data(); c.v2t = v2t.data();
    c.row = row.data(); c.loc = loc.data(); c.winner = winner.data();
    c.rank = rank.data(); c.vote = vote.data(); c.flip_edge = flip_edge.data(); c.role = role.data();
    c.mailbox = mailbox.data(); c.amark = amark.data(); c.fliplist = fliplist.data(); c.counters = counters; c.err = &counters[FM_C_ERR]; c.slow = slow.data();
    int boot[4];
    if (dt_bootstrap(be, c, boot, g_err, sizeof g_err)) return 1;
    for (int k = 0; k < 4; k++) loc[boot[k]] = -1;
    int K = dt_level_count(n);
    std::vector<int> order(n), off(K + 2, 0), prev(n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
      return point_class(pts[a].id, K) > point_class(pts[b].id, K);  // level = K - class, ascending
    });
    for (int i = 0; i < n; i++) off[K - point_class(pts[i].id, K) + 1]++;
    for (int L = 0; L <= K; L++) off[L + 1] += off[L];
    if (dt_run(be, c, listA.data(), listB.data(), order.data(), off.data(), K, prev.data(), boot, &stamp, &st,
               g_err, sizeof g_err))
      return 1;
    if (c.n_tri != T) { snprintf(g_err, sizeof g_err, "triangle count %d != 2n-4", c.n_tri); return 1; }
    return 0;
  }
  void export_tris(int* tri_out, int* adj_out) {
    for (int t = 0; t < T; t++)
      for (int k = 0; k < 3; k++) {
        tri_out[3 * t + k] = orig[geti(row[t].v, k)] + 1;
        adj_out[3 * t + k] = (geti(row[t].a, k) >> 2) + 1;
      }
  }
};

// xyz: n x 3 row major.  tri/adj: (2n-4) x 3, 1-based ORIGINAL ids / triangle ids.
// stats: rounds, subrounds, flips, max_active.
extern "C" int fme_triangulate(int n, const double* xyz, int* tri_out, int* adj_out, int* stats,
                               int do_sort) {
  g_err[0] = 0;
  if (n < 4) { snprintf(g_err, sizeof g_err, "need at least 4 points"); return 1; }
  EmuMesh m;
  if (m.build(n, xyz, do_sort)) return 1;
  m.export_tris(tri_out, adj_out);
  if (stats) { stats[0] = m.st.rounds; stats[1] = m.st.subrounds; stats[2] = m.st.flips; stats[3] = m.st.max_active; }
  return 0;
}

// cold build with the point location of every level done in `nranks` slices (sharded form)
extern "C" int fme_triangulate_sharded(int n, const double* xyz, int nranks, int* tri_out, int* adj_out) {
  g_err[0] = 0;
  EmuMesh m;
  m.be.shard_ranks = nranks;
  if (m.build(n, xyz, 1)) return 1;
  m.export_tris(tri_out, adj_out);
  return 0;
}

// Warm repair after the points moved (xyz_new, input order), the way `nranks` ranks of a sharded
// mesh do it (csrc/fishman_b200.cu repair_async): rank r sweeps the slots [T r / nranks,
// T (r+1) / nranks) with KSweepShT into its own request list, then ALL requests are applied
// (KApply) and the flip loop runs -- on one copy of the mesh, because every rank would do exactly
// the same from here on.  nranks = 1 runs the unsharded dt_repair_async.  rank_order permutes the
// order in which the ranks' requests are applied (the result must not depend on it).
// Returns the number of turned-over triangles through *turned (> 0: the caller would rebuild).
extern "C" int fme_repair(int n, const double* xyz_old, const double* xyz_new, int nranks, int reverse_order,
                          int* tri_out, int* adj_out, int* turned, int* n_requests) {
  g_err[0] = 0;
  EmuMesh m;
  if (m.build(n, xyz_old, 1)) return 1;
  for (int i = 0; i < n; i++) {
    int o = m.orig[i];
    m.pts[i].x = xyz_new[3 * o]; m.pts[i].y = xyz_new[3 * o + 1]; m.pts[i].z = xyz_new[3 * o + 2];
  }
  DtCtx& c = m.c;
  int* list[2] = {m.listA.data(), m.listB.data()};
  *n_requests = 0;
  if (nranks <= 1) {
    if (dt_repair_async(m.be, c, list[0], list[1], &m.stamp, &m.st, g_err, sizeof g_err)) return 1;
  } else {
    int zero[15] = {0};
    m.be.write_ints(c.counters, zero, 15);
    c.stamp = m.stamp;
    c.next_active = list[0];
    c.qcount = c.counters;
    std::vector<std::vector<int>> req(nranks);
    std::vector<int> bnd(m.T + 1);
    int bad = 0;
    for (int r = 0; r < nranks; r++) {
      int tb = (int)((long long)m.T * r / nranks), te = (int)((long long)m.T * (r + 1) / nranks);
      req[r].assign(te - tb + 1, 0);
      DtCtx cs = c;
      cs.t_begin = tb; cs.req = req[r].data(); cs.bnd = bnd.data();
      cs.blk_begin = (int)((long long)n * r / nranks); cs.blk_end = (int)((long long)n * (r + 1) / nranks);
      c.counters[FM_C_REQ] = 0; c.counters[FM_C_BND] = 0; c.counters[3] = 0;
      m.be.launch_filtered<KSweepShT>(cs, te - tb);
      req[r].resize(c.counters[FM_C_REQ]);
      bad += c.counters[3];
      *n_requests += (int)req[r].size();
    }
    c.counters[FM_C_REQ] = 0; c.counters[FM_C_BND] = 0;
    c.counters[3] = bad;
    for (int q = 0; q < nranks; q++) {
      int r = reverse_order ? nranks - 1 - q : q;
      if (reverse_order) std::reverse(req[r].begin(), req[r].end());
      for (int msg : req[r]) KApply::run(c, msg);
    }
    if (bad == 0 && m.be.flip_loop(c, list, 0, true, &m.stamp, &m.st, g_err, sizeof g_err, -1)) return 1;
  }
  *turned = c.counters[3];
  if (m.counters[FM_C_ERR]) { snprintf(g_err, sizeof g_err, "error bits %d", m.counters[FM_C_ERR]); return 1; }
  m.export_tris(tri_out, adj_out);
  return 0;
}

// A failed warm repair followed by the in-place rebuild (fm_retriangulate(h, warm=1) when a
// triangle turned over): build(xyz_old), move the points, sweep (which stamps amark[] / vote[]
// and finds the turned-over triangles; the flip loop then leaves at once, as fm_flip_coop does),
// then bootstrap + insertion rounds again on the same arrays.  skip_bump = 1 reproduces a
// rebuild that starts on the stamp the sweep used.
extern "C" int fme_rebuild_after_failed_repair(int n, const double* xyz_old, const double* xyz_new, int skip_bump,
                                               int* tri_out, int* adj_out, int* turned) {
  g_err[0] = 0;
  EmuMesh m;
  if (m.build(n, xyz_old, 1)) return 1;
  for (int i = 0; i < n; i++) {
    int o = m.orig[i];
    m.pts[i].x = xyz_new[3 * o]; m.pts[i].y = xyz_new[3 * o + 1]; m.pts[i].z = xyz_new[3 * o + 2];
  }
  m.be.emulate_abort = true;
  if (dt_repair_async(m.be, m.c, m.listA.data(), m.listB.data(), &m.stamp, &m.st, g_err, sizeof g_err)) return 1;
  *turned = m.c.counters[3];
  m.counters[FM_C_ERR] = 0;
  m.be.skip_bump = skip_bump != 0;
  if (m.run()) return 1;
  if (m.counters[FM_C_ERR]) { snprintf(g_err, sizeof g_err, "error bits %d", m.counters[FM_C_ERR]); return 1; }
  m.export_tris(tri_out, adj_out);
  return 0;
}

// ---- cold build by stars (csrc/star_kernels.cuh) on the host: grid, stars, rows.
// stats: [0] generators retried with a wider block, [1] generators redone with exact predicates,
// [2] stars that failed, [3] triangles not found in their owner's star, [4] owned triangles, [5] sum of degrees,
// [6] cells per axis.  Returns 0 when the build is complete and consistent, 2 when the product would
// fall back to the insertion build, 1 on a usage error.
#include "../../fishman_b200/csrc/star_kernels.cuh"

extern "C" int fme_triangulate_star(int n, const double* xyz, double kappa, int wide, int* tri_out, int* adj_out, int* stats) {
  g_err[0] = 0;
  if (n < 16) { snprintf(g_err, sizeof g_err, "need at least 16 points"); return 1; }
  const int T = 2 * n - 4;
  std::vector<int> orig(n);
  std::iota(orig.begin(), orig.end(), 0);
  {
    std::vector<u64> key(n);
    for (int i = 0; i < n; i++) key[i] = morton_key(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    std::stable_sort(orig.begin(), orig.end(), [&](int a, int b) { return key[a] < key[b]; });
  }
  std::vector<P4> pts(n), spts(n);
  for (int i = 0; i < n; i++) pts[i] = P4{xyz[3 * orig[i]], xyz[3 * orig[i] + 1], xyz[3 * orig[i] + 2], orig[i]};
  StarCtx s;
  memset(&s, 0, sizeof s);
  const double h = sqrt(4.0 * 3.14159265358979323846 / n);
  int G = (int)(2.0 / (kappa * h));
  if (G < 1) G = 1;
  if (G > kStarMaxG) G = kStarMaxG;
  s.n = n; s.G = G; s.half_g = 0.5 * G; s.cell = 2.0 / G;
  std::vector<unsigned> key(n), skey(n);
  std::vector<int> idx(n), sidx(n), colstart((size_t)G * G + 1), colmid((size_t)G * G), ring((size_t)n * kStarCap, 0), deg(n, 0),
      stri((size_t)n * kStarCap, 0), listA(n), listB(n), listE(n), listS(n), v2t(n, -1);
  std::vector<unsigned> own(n, 0);
  std::vector<u64> off(n + 1, 0);
  std::vector<Row> row(T);
  int sc[8] = {0};
  s.pts = pts.data(); s.key = key.data(); s.idx = idx.data();
  for (int i = 0; i < n; i++) KStarKey::run(s, i);
  std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) { return key[a] < key[b]; });
  for (int j = 0; j < n; j++) { sidx[j] = idx[j]; skey[j] = key[idx[j]]; }
  s.skey = skey.data(); s.sidx = sidx.data(); s.spts = spts.data(); s.colstart = colstart.data(); s.colmid = colmid.data();
  for (int j = 0; j < n; j++) KStarGather::run(s, j);
  for (int c = 0; c <= G * G; c++) KStarCols::run(s, c);
  s.ring = ring.data(); s.deg = deg.data(); s.own = own.data(); s.off = off.data(); s.stri = stri.data();
  s.list_exact = listE.data(); s.list_slow = listS.data(); s.sc = sc; s.row = row.data(); s.v2t = v2t.data();
  std::vector<int> rng(4 * kStarListWide + 2 * (2 * kStarMaxR + 1) * (2 * kStarMaxR + 1));
  s.near2 = wide ? 2.7 : 1.0; s.list_cap = wide ? kStarListWide : kStarList;
  int* lists[2] = {listA.data(), listB.data()};
  s.r = 1; s.in_idx = 1; s.out_idx = 0; s.list_out = lists[0];
  for (int j = 0; j < n; j++) KStarAllT<false>::run(s, j, rng.data(), 1);
  int retried = sc[0];
  for (int r = 2; r <= kStarMaxR; r++) {
    s.r = r; s.in_idx = (r & 1) ? 1 : 0; s.out_idx = s.in_idx ^ 1;
    s.list_in = lists[s.in_idx]; s.list_out = lists[s.out_idx];
    sc[s.out_idx] = 0;
    for (int i = 0, m = sc[s.in_idx]; i < m; i++) KStarRetry::run(s, i, rng.data(), 1);
  }
  {
    StarCtx sl = s;
    sl.list_cap = kStarListLong; sl.list_in = lists[0];
    sc[FM_SC_LIST0] = 0;
    std::vector<int> rl(4 * kStarListLong + 2 * (2 * kStarMaxR + 1) * (2 * kStarMaxR + 1));
    for (int i = 0, m = sc[FM_SC_SLOW]; i < m; i++) KStarSlowNear::run(sl, i, rl.data(), 1);
    if (stats) stats[8] = sc[FM_SC_LIST0];
    for (int i = 0, m = sc[FM_SC_LIST0]; i < m; i++) KStarSlow::run(sl, i, rng.data(), 1);
  }
  for (int i = 0, m = sc[FM_SC_EXACT]; i < m; i++) KStarExact::run(s, i, rng.data(), 1);
  off[n] = 0;
  {
    u64 acc = 0;
    for (int i = 0; i <= n; i++) { u64 v = off[i]; off[i] = acc; acc += v; }
  }
  const int owned = (int)(off[n] & 0xffffffffULL), degsum = (int)(off[n] >> 32);
  if (stats) {
    stats[0] = retried; stats[7] = sc[FM_SC_SLOW]; stats[1] = sc[FM_SC_EXACT]; stats[2] = sc[FM_SC_FAIL]; stats[4] = owned; stats[5] = degsum; stats[6] = G;
  }
  if (sc[FM_SC_FAIL] || owned != T || degsum != 3 * T) {
    if (stats) stats[3] = -1;
    snprintf(g_err, sizeof g_err, "stars incomplete: %d failed, %d owned triangles (want %d), degree sum %d (want %d)",
             sc[FM_SC_FAIL], owned, T, degsum, 3 * T);
    return 2;
  }
  for (int i = 0; i < n; i++) KStarSlots::run(s, i);
  if (!sc[FM_SC_BROKEN])
    for (int i = 0; i < n; i++) KStarRows::run(s, i);
  if (stats) stats[3] = sc[FM_SC_BROKEN];
  if (sc[FM_SC_BROKEN]) { snprintf(g_err, sizeof g_err, "stars inconsistent: %d triangles", sc[FM_SC_BROKEN]); return 2; }
  for (int t = 0; t < T; t++)
    for (int k = 0; k < 3; k++) {
      tri_out[3 * t + k] = orig[geti(row[t].v, k)] + 1;
      adj_out[3 * t + k] = (geti(row[t].a, k) >> 2) + 1;
    }
  return 0;
}
