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
  void fill64(u64* p, u64 v, size_t n) { std::fill(p, p + n, v); }
  void fill32(int* p, int v, size_t n) { std::fill(p, p + n, v); }
  int exclusive_scan(int* p, int n) {
    int acc = 0;
    for (int i = 0; i < n; i++) { int v = p[i]; p[i] = acc; acc += v; }
    return acc;
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
  void prev_coarser(const DtCtx& c, int thr, int* prev) {
    int last = -1;
    for (int i = 0; i < c.n; i++) {
      if (point_class(c.pts[i].id, c.level_max) > thr) last = i;
      prev[i] = last;
    }
  }
  int flip_loop(DtCtx& c, int* list[2], int cur, bool skip_check, int* stamp, DtStats* st, char* msg,
                size_t msglen, int) {
    return dt_flip_loop_host(*this, c, list, cur, skip_check, stamp, st, msg, msglen);
  }
};

static char g_err[512];

extern "C" const char* fme_last_error() { return g_err; }

// xyz: n x 3 row major.  tri/adj: (2n-4) x 3, 1-based ORIGINAL ids / triangle ids.
// stats: rounds, subrounds, flips, max_active.
extern "C" int fme_triangulate(int n, const double* xyz, int* tri_out, int* adj_out, int* stats,
                               int do_sort) {
  g_err[0] = 0;
  if (n < 4) { snprintf(g_err, sizeof g_err, "need at least 4 points"); return 1; }
  std::vector<int> orig(n);
  std::iota(orig.begin(), orig.end(), 0);
  if (do_sort) {
    std::vector<u64> key(n);
    for (int i = 0; i < n; i++) key[i] = morton_key(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    std::stable_sort(orig.begin(), orig.end(), [&](int a, int b) { return key[a] < key[b]; });
  }
  std::vector<P4> pts(n);
  for (int i = 0; i < n; i++) pts[i] = P4{xyz[3 * orig[i]], xyz[3 * orig[i] + 1], xyz[3 * orig[i] + 2], orig[i]};
  int T = 2 * n - 4;
  std::vector<Row> row(T);
  std::vector<int> loc(n, 0), rank(T + 1), flip_edge(T), role(T, 0), amark(T, 0), listA(T), listB(T), v2t(n, -1);
  std::vector<u64> winner(T), vote(T, kNone64), mailbox(3 * (size_t)T, 0);
  int counters[16] = {0}, err = 0;
  std::vector<int> slow(T + 1);
  DtCtx c;
  memset(&c, 0, sizeof c);
  c.n = n; c.pts = pts.data(); c.v2t = v2t.data();
  c.row = row.data(); c.loc = loc.data(); c.winner = winner.data();
  c.rank = rank.data(); c.vote = vote.data(); c.flip_edge = flip_edge.data(); c.role = role.data();
  c.mailbox = mailbox.data(); c.amark = amark.data(); c.counters = counters; c.err = &err; c.slow = slow.data();
  HostBackend be;
  int boot[4];
  if (dt_bootstrap(be, c, boot, g_err, sizeof g_err)) return 1;
  for (int k = 0; k < 4; k++) loc[boot[k]] = -1;
  DtStats st;
  int stamp = 1;
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
  for (int t = 0; t < T; t++)
    for (int k = 0; k < 3; k++) {
      tri_out[3 * t + k] = orig[geti(row[t].v, k)] + 1;
      adj_out[3 * t + k] = (geti(row[t].a, k) >> 2) + 1;
    }
  if (stats) { stats[0] = st.rounds; stats[1] = st.subrounds; stats[2] = st.flips; stats[3] = st.max_active; }
  return 0;
}
