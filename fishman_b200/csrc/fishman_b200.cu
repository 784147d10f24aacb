// fishman_b200.cu -- the C ABI of include/fishman_b200.h on top of the sm_100a kernels.
// Device memory, streams and launches only: no host execution path exists here.
#include <stdarg.h>
#include <stdlib.h>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/fishman_b200.h"
#include "launch_kernels.cuh"
#include "mpas_kernels.cuh"
#include "quality_kernels.cuh"

using namespace fm;

static char g_create_err[512];

#define FM_CUDA(call)                                                                   \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      h->fail("CUDA error %s at %s:%d", cudaGetErrorString(e_), __FILE__, __LINE__);    \
      return 1;                                                                         \
    }                                                                                   \
  } while (0)

// ------------------------------------------------------------------- handle
template <class K> struct KName { static const char* get() { return "kernel"; } };
#define FM_KNAME(K, s) template <> struct KName<K> { static const char* get() { return s; } };
FM_KNAME(KLocateVote, "locate_exact")
FM_KNAME(KLocateVoteT<false>, "locate_vote")
FM_KNAME(KWinnerReset, "winner_reset")
FM_KNAME(KSplitFlag, "split_flag")
FM_KNAME(KVote, "vote")
FM_KNAME(KLevelInit, "level_init")
FM_KNAME(KSplit, "split13")
FM_KNAME(KCheck, "check_exact")
FM_KNAME(KCheckT<false>, "flip_check")
FM_KNAME(KMark, "flip_mark")
FM_KNAME(KFlip, "flip22")
FM_KNAME(KPatch, "flip_patch")
FM_KNAME(KSweep, "sweep_exact")
FM_KNAME(KSweepT<false>, "sweep")
FM_KNAME(KSweepShT<true>, "sweep_exact")
FM_KNAME(KSweepShT<false>, "sweep")
FM_KNAME(KValidate, "validate")
FM_KNAME(KCircum, "circumcentre")
FM_KNAME(KDegree, "ring_degree")
FM_KNAME(KRingFill, "ring_fill")

struct ProfRow {
  std::string name;
  long long launches = 0, threads = 0;
  double ms = 0.0;
};
struct ProfPending {
  cudaEvent_t a, b;
  const char* name;
  long long n;
};

struct fm_handle {
  int n = 0, n_pad = 0, n_tri = 0, rank = 0, size = 1;
  int blk_begin = 0, blk_count = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // generators (internal order) and the permutation
  P4* pts = nullptr;
  int *orig = nullptr, *inv = nullptr;
  int* ringc = nullptr;  // ring cache of the Lloyd kernel: 8 ints per vertex
  // insertion levels: points sorted by (level, internal index), level offsets (host), scratch
  int *lvl_order = nullptr, *prev = nullptr, *next_rev = nullptr;
  int lvl_K = 0;
  std::vector<int> lvl_off;
  // triangulation
  Row* rows = nullptr;
  int *loc = nullptr, *rankbuf = nullptr, *flip_edge = nullptr, *role = nullptr, *amark = nullptr;
  int *listA = nullptr, *listB = nullptr, *fliplist = nullptr, *counters = nullptr, *err = nullptr, *v2t = nullptr;
  u64 *winner = nullptr, *vote = nullptr, *mailbox = nullptr, *scratch64 = nullptr;
  void* cub_tmp = nullptr;
  size_t cub_bytes = 0;
  int* pinned = nullptr;
  int stamp = 1;  // host mirror (unused by the device-resident flip loop)
  int coop_blocks = 0, sms = 0;
  bool have_dt = false, have_cc = false, have_vd = false, have_edges = false;
  bool dt_current = false;  // the connectivity is the Delaunay triangulation of the CURRENT generators
  // Voronoi
  P4* cc = nullptr;
  double *cclon = nullptr, *cclat = nullptr;
  int *deg = nullptr, *ring_tri = nullptr, *ring_nbr = nullptr;
  int *ecount = nullptr, *de = nullptr, *ve_vvt = nullptr, *ve_vc = nullptr;
  double* stage = nullptr;  // 3n doubles: input-order staging of fm_get/set_generators
  // MPAS adapter arrays (fm_mpas), allocated on first use
  struct Mpas {
    int max_edges = 0, cap_max_edges = 0;
    int *ring_edge = nullptr, *cells_on_vertex = nullptr, *edges_on_vertex = nullptr, *n_eoe = nullptr, *eoe = nullptr;
    double *xyz_edge = nullptr, *lonlat_edge = nullptr, *edge5 = nullptr;  // edge5: dv, dc, dv1, dv2, angle (5 x E)
    double *area_cell = nullptr, *area_tri = nullptr, *kite_ring = nullptr, *kite_vtx = nullptr, *weights = nullptr;
    double *sums = nullptr, *part = nullptr;
    bool have = false;
  } mp;
  double* qbuf = nullptr;  // fm_quality: 16 histogram words, 6 extrema, 2 sums, then the block partials
  // cold build by stars (star_kernels.cuh): buffers allocated by the first build that takes that path
  struct StarBuf {
    unsigned *key = nullptr, *skey = nullptr, *own = nullptr;
    int *idx = nullptr, *sidx = nullptr, *colstart = nullptr, *colmid = nullptr, *ring = nullptr, *deg = nullptr, *stri = nullptr;
    int *listA = nullptr, *listB = nullptr, *listE = nullptr, *listS = nullptr, *sc = nullptr;
    u64* off = nullptr;
    P4* spts = nullptr;
    void* tmp = nullptr;
    size_t tmp_bytes = 0;
    int G = 0, Gw = 0;
    bool ready = false, unavailable = false;
  } star;
  int star_mode = 0, star_wide_builds = 0;   // 1: the last star build needed the wide mode (irregular set)
  int star_skip = 0;   // cold builds left before the stars are tried again on a handle whose set the sampling pass refused
  long long n_star = 0, n_star_fallback = 0;   // cold builds by stars / attempts that fell back to the insertion build
  double star_irregular = 0.0;                  // fraction of the sampled generators the first pass did not settle cheaply
  int star_last[8] = {0};                       // the last star build: retried, exact, failed, broken, owned, degree sum, flips of the check, turned-over
  // the density function staged as a table for the Lloyd kernel (Density::tab), for the parameters rho_tab_for
  double* rho_tab = nullptr;
  fm_density rho_tab_for = {};
  bool rho_tab_ok = false;
  double rho_tab_err = 0.0;
  // peer exchange: `pts` and a PeerSync block live in one arena that is exported over CUDA IPC
  // and mapped by the other ranks of the mesh (fm_peer_handle / fm_peer_attach)
  char* arena = nullptr;
  PeerSync* psync = nullptr;
  int* inbox = nullptr;       // size regions of inbox_cap flip requests, one per sending rank
  int* bnd = nullptr;         // swept triangles with a vertex on another rank (inbox_cap entries)
  int inbox_cap = 0;
  size_t off_sync = 0, off_inbox = 0, off_cc = 0, off_loc = 0;  // arena layout (the same on every rank)
  bool shard_locate = false;  // peers attached + FM_SHARD_LOCATE=1: split each level's point location over the ranks
                              // (off by default: measured slower than replicated location)
  int sweep_begin = 0, sweep_end = 0;              // triangle slots this rank sweeps (sharded warm repair)
  void* peer_base[8] = {nullptr};
  PeerCtx peer_ctx = {};
  bool peers_attached = false;
  // the warm Lloyd iteration as a CUDA graph (reset, sweep, deferred sweep, flip loop, centroid,
  // readback copy): one launch call per iteration instead of seven
  cudaGraphExec_t warm_graph = nullptr;
  fm_density warm_graph_rho = {};
  bool graph_disabled = false;
  // accounting
  long long n_launch = 0, n_cold = 0, n_warm = 0, tot_rounds = 0, tot_sub = 0, tot_flips = 0;
  int max_active = 0;
  bool profile = false;
  std::vector<ProfRow> prof;
  std::vector<ProfPending> pending;
  std::vector<cudaEvent_t> ev_pool;
  bool cuda_failed = false;
  char errmsg[512] = {0};

  void fail(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(errmsg, sizeof errmsg, fmt, ap);
    va_end(ap);
  }
  void note_cuda(cudaError_t e, const char* what) {
    if (e != cudaSuccess && !cuda_failed) {
      cuda_failed = true;
      fail("CUDA error in %s: %s", what, cudaGetErrorString(e));
    }
  }
  ProfRow& row(const char* name) {
    for (auto& r : prof)
      if (r.name == name) return r;
    prof.push_back(ProfRow());
    prof.back().name = name;
    return prof.back();
  }
  // Per-launch timing without extra synchronisation: a pair of events from a pool brackets
  // each launch on the launching stream; pairs are resolved at the next point where the host
  // waits for the stream anyway (prof_drain).
  cudaEvent_t ev_get() {
    cudaEvent_t e;
    if (!ev_pool.empty()) { e = ev_pool.back(); ev_pool.pop_back(); return e; }
    cudaEventCreate(&e);
    return e;
  }
  void prof_begin() {
    if (!profile) return;
    ProfPending p;
    p.a = ev_get(); p.b = ev_get(); p.name = ""; p.n = 0;
    cudaEventRecord(p.a, stream);
    pending.push_back(p);
  }
  void prof_end(const char* name, long long n) {
    n_launch++;
    if (!profile) return;
    ProfPending& p = pending.back();
    p.name = name; p.n = n;
    cudaEventRecord(p.b, stream);
  }
  void prof_drain() {  // call only after the stream has been synchronised
    for (auto& p : pending) {
      float ms = 0;
      if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
        ProfRow& r = row(p.name);
        r.launches++; r.threads += p.n; r.ms += ms;
      }
      ev_pool.push_back(p.a); ev_pool.push_back(p.b);
    }
    pending.clear();
  }
  cudaError_t sync() {
    cudaError_t e = cudaStreamSynchronize(stream);
    if (profile) prof_drain();
    return e;
  }
  template <class K, class Ctx>
  void launch(const Ctx& c, int n) {
    if (n <= 0) return;
    prof_begin();
    fm_kernel<K, Ctx><<<(n + 255) / 256, 256, 0, stream>>>(c, n);
    note_cuda(cudaGetLastError(), KName<K>::get());
    prof_end(KName<K>::get(), n);
  }
  DtCtx dt_ctx() {
    DtCtx c;
    memset(&c, 0, sizeof c);
    c.n = n; c.pts = pts; c.v2t = v2t; c.ringc = ringc; c.row = rows; c.loc = loc;
    c.winner = winner; c.rank = rankbuf; c.n_tri = n_tri; c.stamp = stamp; c.stamp_dev = counters + FM_C_STAMP;
    c.vote = vote; c.flip_edge = flip_edge; c.role = role; c.mailbox = mailbox; c.amark = amark;
    c.counters = counters; c.qcount = counters; c.err = err; c.slow = rankbuf; c.fliplist = fliplist;
    static const int pf_waves = getenv("FM_PREFETCH_WAVES") ? atoi(getenv("FM_PREFETCH_WAVES")) : 1;
    c.pf_dist = pf_waves * sms * 4 * 256;  // the sweep runs 4 blocks of 256 per SM: one wave ahead
    return c;
  }
  VdCtx vd_ctx(const Density* rho) {
    VdCtx c;
    memset(&c, 0, sizeof c);
    c.n = n; c.n_tri = n_tri; c.pts = pts; c.row = rows;
    c.cc = cc; c.cclon = cclon; c.cclat = cclat; c.v2t = v2t; c.ringc = ringc;
    c.deg = deg; c.ring_tri = ring_tri; c.ring_nbr = ring_nbr;
    if (rho) c.rho = *rho;
    c.cell_begin = blk_begin; c.cell_end = blk_begin + blk_count;
    c.maxdisp = (u64*)(counters + FM_C_MAXDISP); c.err = err; c.counters = counters;
    if (peers_attached)
      for (int k = 0; k < size; k++)
        if (k != rank) c.peer_pts[c.n_peers++] = (P4*)peer_base[k];
    return c;
  }
};

struct CudaBackend {
  fm_handle* h;
  bool in_batch = false;  // inside a batch of warm iterations fm_iter_end has already cleared the counters
  template <class K>
  void launch(const DtCtx& c, int n) { h->launch<K, DtCtx>(c, n); }
  // K over [0, counters[idx]); `upper` bounds the count (threads past it leave after one load)
  template <class K>
  void launch_counted(const DtCtx& c, int idx, int upper) {
    if (upper <= 0) return;
    h->prof_begin();
    fm_kernel_counted<K><<<(upper + 255) / 256, 256, 0, h->stream>>>(c, idx);
    h->note_cuda(cudaGetLastError(), KName<K>::get());
    h->prof_end(KName<K>::get(), upper);
  }
  // filter-only kernel over [0, n), then the exact kernel over what it deferred
  template <template <bool> class K>
  void launch_filtered(const DtCtx& c, int n) {
    if (n <= 0) return;
    h->launch<K<false>, DtCtx>(c, n);
    h->prof_begin();
    fm_kernel_deferred<K<true>><<<h->sms * 4, 128, 0, h->stream>>>(c);
    h->note_cuda(cudaGetLastError(), KName<K<true>>::get());
    h->prof_end(KName<K<true>>::get(), 0);
  }
  // Point location of one level.  On a sharded mesh with attached peers: this rank walks its slice
  // of the level's points (no votes), its results go to every rank's loc[] over NVLink, barrier,
  // then every rank casts all votes -- the triangulation stays replicated and identical.  Small
  // levels are not worth the exchange and are located whole on every rank.
  void locate_round(DtCtx& c, int m);
  void bump_stamp(int*) {  // the stamp lives on the device (counters[FM_C_STAMP]); no host round trip
    fm_bump_stamp<<<1, 1, 0, h->stream>>>(h->counters, 2);
    h->note_cuda(cudaGetLastError(), "bump_stamp");
  }
  void fill64(u64* p, u64 v, size_t n) {
    // only all-ones is ever requested: a byte memset does it
    h->note_cuda(cudaMemsetAsync(p, (int)(v & 0xff), n * sizeof(u64), h->stream), "fill64");
  }
  void fill32(int* p, int v, size_t n) {
    h->note_cuda(cudaMemsetAsync(p, v & 0xff, n * sizeof(int), h->stream), "fill32");
  }
  void exclusive_scan_async(int* p, int n) {
    h->prof_begin();
    size_t bytes = h->cub_bytes;
    h->note_cuda(cub::DeviceScan::ExclusiveSum(h->cub_tmp, bytes, p, p, n, h->stream), "scan");
    h->prof_end("cub_scan", n);
  }
  void read_ints(int* dst, const int* src, int n) {
    h->note_cuda(cudaMemcpyAsync(h->pinned, src, sizeof(int) * n, cudaMemcpyDeviceToHost, h->stream),
                 "read_ints");
    h->note_cuda(h->sync(), "sync");
    memcpy(dst, h->pinned, sizeof(int) * n);
  }
  void write_ints(int* dst, const int* src, int n) {
    // sources are tiny host constants; zero is the only value ever written
    bool allzero = true;
    for (int i = 0; i < n; i++) allzero = allzero && (src[i] == 0);
    if (allzero && in_batch && dst == h->counters) return;
    if (allzero) h->note_cuda(cudaMemsetAsync(dst, 0, sizeof(int) * n, h->stream), "write_ints");
    else {
      h->note_cuda(cudaMemcpyAsync(dst, src, sizeof(int) * n, cudaMemcpyHostToDevice, h->stream), "write_ints");
      h->note_cuda(h->sync(), "sync");
    }
  }
  int argmax_dot(const DtCtx& c, const double* d) {
    double nn = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]) + 1e-300;
    cudaMemsetAsync(h->scratch64 + 1, 0, sizeof(u64), h->stream);
    fm_argmax_dot<<<(c.n + 255) / 256, 256, 0, h->stream>>>(c.n, c.pts, d[0] / nn, d[1] / nn, d[2] / nn,
                                                            h->scratch64 + 1);
    h->n_launch++;
    u64 v = 0;
    h->note_cuda(cudaMemcpyAsync(&v, h->scratch64 + 1, sizeof v, cudaMemcpyDeviceToHost, h->stream), "argmax");
    h->note_cuda(h->sync(), "sync");
    return 0x7fffffff - (int)(v & 0xffffffffULL);
  }
  P4 fetch4(const DtCtx& c, int i) {
    P4 p = {0, 0, 0, 0};
    cudaMemcpyAsync(&p, c.pts + i, sizeof(P4), cudaMemcpyDeviceToHost, h->stream);
    h->note_cuda(h->sync(), "fetch_point");
    return p;
  }
  P3 fetch_point(const DtCtx& c, int i) { P4 q = fetch4(c, i); return P3{q.x, q.y, q.z}; }
  int fetch_orig(const DtCtx& c, int i) { return (int)fetch4(c, i).id; }
  struct MaxOp {
    __host__ __device__ int operator()(int a, int b) const { return a > b ? a : b; }
  };
  const int* prev_coarser(const DtCtx& c, int thr, int* prev) {
    h->prof_begin();
    fm_prev_keys<<<(c.n + 255) / 256, 256, 0, h->stream>>>(c.n, c.pts, c.level_max, thr, prev, h->next_rev);
    size_t bytes = h->cub_bytes;
    h->note_cuda(cub::DeviceScan::InclusiveScan(h->cub_tmp, bytes, prev, prev, MaxOp(), c.n, h->stream), "prev scan");
    bytes = h->cub_bytes;
    h->note_cuda(cub::DeviceScan::InclusiveScan(h->cub_tmp, bytes, h->next_rev, h->next_rev, MaxOp(), c.n, h->stream),
                 "next scan");
    h->prof_end("prev_coarser", c.n);
    return h->next_rev;
  }
  void write_tris(const DtCtx& c, int t0, int cnt, const I4* tri, const I4* adj) {
    std::vector<Row> rows(cnt);
    for (int i = 0; i < cnt; i++) rows[i] = Row{tri[i], adj[i]};
    cudaMemcpyAsync(c.row + t0, rows.data(), sizeof(Row) * cnt, cudaMemcpyHostToDevice, h->stream);
    h->note_cuda(h->sync(), "write_tris");
    for (int i = 0; i < cnt; i++) {
      int t = t0 + i;
      const int v[3] = {tri[i].x, tri[i].y, tri[i].z};
      for (int k = 0; k < 3; k++) cudaMemcpyAsync(c.v2t + v[k], &t, sizeof(int), cudaMemcpyHostToDevice, h->stream);
      h->note_cuda(h->sync(), "write_tris");  // &t must stay alive until the copy is done
    }
  }
  // Flip sub-rounds until the queue is empty.  While the queue is long (known to the host:
  // nact_hint, then one small readback per sub-round) each phase is its own launch with its
  // own register budget and full occupancy; the tail of short queues -- and the whole loop
  // when the length is not known (warm repair) -- runs as ONE cooperative launch with
  // grid-wide barriers, no host round trips.  Asynchronous in the cooperative part.
  int flip_loop(DtCtx& c, int* list[2], int cur, bool skip_check, int*, DtStats*, char*, size_t, int nact_hint) {
    const int kCoopBelow = 49152;
    int nact = nact_hint;
    while (nact > kCoopBelow) {
      c.active = list[cur];
      c.n_active = nact;
      c.next_active = list[cur ^ 1];
      c.qcount = c.counters + (cur ^ 1);
      static const bool coop4 = getenv("FM_CHECK_COOP4") && atoi(getenv("FM_CHECK_COOP4")) != 0;
      if (!skip_check && coop4) {  // A/B: four lanes per queued triangle, then the exact kernel over what was deferred
        h->prof_begin();
        fm_check_coop4<<<(unsigned)((4LL * nact + 255) / 256), 256, 0, h->stream>>>(c, nact);
        h->note_cuda(cudaGetLastError(), "flip_check_coop4");
        h->prof_end("flip_check_coop4", nact);
        h->prof_begin();
        fm_kernel_deferred<KCheckT<true>><<<h->sms * 4, 128, 0, h->stream>>>(c);
        h->note_cuda(cudaGetLastError(), "check_exact");
        h->prof_end("check_exact", 0);
      } else if (!skip_check) launch_filtered<KCheckT>(c, nact);
      skip_check = false;
      launch<KMark>(c, nact);
      launch_counted<KFlip>(c, FM_C_NFLIP, nact);  // winners are queue entries: at most nact (and, being disjoint pairs, T / 2)
      launch_counted<KPatch>(c, FM_C_NFLIP, nact);
      fm_subround_end<<<1, 1, 0, h->stream>>>(c.counters, cur, nact);
      cur ^= 1;
      read_ints(&nact, c.counters + cur, 1);
      if (h->cuda_failed) return 1;
    }
    if (nact == 0) return 0;
    int skip = skip_check ? 1 : 0;
    void* args[] = {&c, &list[0], &list[1], &cur, &skip};
    h->prof_begin();
    h->note_cuda(cudaLaunchCooperativeKernel((void*)fm_flip_coop, dim3(h->coop_blocks), dim3(256), args, 0,
                                             h->stream), "flip_loop");
    h->prof_end("flip_loop", 0);
    return h->cuda_failed ? 1 : 0;
  }
};

static void launch_peer_barrier(fm_handle* h, int phase);
void CudaBackend::locate_round(DtCtx& c, int m) {
  const int kMinPerRank = 16384;
  if (!(h->peers_attached && h->shard_locate) || m < kMinPerRank * h->size) {
    launch_filtered<KLocateVoteT>(c, m);
    return;
  }
  int mb = (int)((long long)m * h->rank / h->size), me = (int)((long long)m * (h->rank + 1) / h->size);
  DtCtx cs = c;
  cs.plist = c.plist + mb;
  cs.locate_only = 1;
  launch_filtered<KLocateVoteT>(cs, me - mb);
  LocFan f = {};
  for (int k = 0; k < h->size; k++)
    f.loc[k] = (int*)(((k == h->rank) ? h->arena : (char*)h->peer_base[k]) + h->off_loc);
  f.plist = cs.plist; f.rank = h->rank; f.size = h->size;
  launch_peer_barrier(h, 3);  // every rank is done with the previous round's votes and splits: loc[] may be overwritten
  h->prof_begin();
  fm_loc_fanout<<<(me - mb + 255) / 256, 256, 0, h->stream>>>(f, me - mb);
  h->note_cuda(cudaGetLastError(), "loc_fanout");
  h->prof_end("loc_fanout", me - mb);
  launch_peer_barrier(h, 3);
  launch<KVote>(c, m);
}

// ---------------------------------------------------------------------- API
template <class T>
static cudaError_t dalloc(T** p, size_t count) {
  return cudaMalloc((void**)p, sizeof(T) * (count ? count : 1));
}

static int create_impl(int32_t n, const double* hx, const double* hy, const double* hz, int lonlat,
                       int32_t rank, int32_t size, fm_handle** out) {
  *out = nullptr;
  g_create_err[0] = 0;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    snprintf(g_create_err, sizeof g_create_err,
             "fishman_b200 needs a CUDA device and found none (%s); there is no CPU path",
             e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    return 1;
  }
  if (n < 4) { snprintf(g_create_err, sizeof g_create_err, "need at least 4 generators, got %d", n); return 1; }
  if (n >= (1 << 28)) {  // adjacency words are (slot << 2 | edge) in an int32 and T = 2n - 4 slots must stay below 2^29
    snprintf(g_create_err, sizeof g_create_err, "at most 2^28 - 1 generators (packed 32-bit adjacency), got %d", n);
    return 1;
  }
  if (size < 1 || rank < 0 || rank >= size) { snprintf(g_create_err, sizeof g_create_err, "bad rank/size"); return 1; }
  fm_handle* h = new fm_handle();
  // staging of the set-up (keys, sort buffers): freed on every way out
  u64 *k0 = nullptr, *k1 = nullptr;
  int *i0 = nullptr, *dup = nullptr, *doff = nullptr;
  auto free_staging = [&]() {
    void* st[] = {k0, k1, i0, dup, doff};
    for (void* q : st)
      if (q) cudaFree(q);
    k0 = k1 = nullptr; i0 = dup = doff = nullptr;
  };
  auto bail = [&](const char* what, cudaError_t err) {
    snprintf(g_create_err, sizeof g_create_err, "fm_create: %s: %s", what,
             err == cudaSuccess ? h->errmsg : cudaGetErrorString(err));
    free_staging();
    fm_destroy(h);
    return 1;
  };
  h->n = n; h->rank = rank; h->size = size;
  int per = (n + size - 1) / size;
  h->n_pad = per * size;
  h->blk_begin = rank * per;
  h->blk_count = h->blk_begin >= n ? 0 : ((h->blk_begin + per <= n) ? per : n - h->blk_begin);
  size_t T = 2 * (size_t)n - 4;
#define A(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return bail(#call, e_); } while (0)
  int dev = 0, sms = 0, per_sm = 0;
  A(cudaGetDevice(&dev));
  A(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  A(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fm_flip_coop, 256, 0));
  h->coop_blocks = sms * (per_sm > 0 ? per_sm : 1);
  if (const char* e = getenv("FM_COOP_BLOCKS")) { int v = atoi(e); if (v > 0 && v <= h->coop_blocks) h->coop_blocks = v; }
  h->sms = sms;
  if (getenv("FM_NO_GRAPH")) h->graph_disabled = true;  // direct launches (profilers that do not follow graphs)
  A(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  A(cudaEventCreate(&h->ev0)); A(cudaEventCreate(&h->ev1));
  A(cudaMallocHost((void**)&h->pinned, 96 * sizeof(int)));
  {  // arena = everything a peer rank touches (exported as ONE IPC allocation): generator records,
     // PeerSync block, flip-request inbox, circumcentres
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    h->inbox_cap = size > 1 ? (int)((2 * T + size - 1) / size) + 64 : 0;
    h->off_sync = up(sizeof(P4) * (size_t)h->n_pad);
    h->off_inbox = up(h->off_sync + sizeof(PeerSync));
    h->off_cc = up(h->off_inbox + sizeof(int) * (size_t)h->inbox_cap * size);
    h->off_loc = up(h->off_cc + sizeof(P4) * T);
    A(cudaMalloc((void**)&h->arena, h->off_loc + sizeof(int) * (size_t)n));
    h->loc = (int*)(h->arena + h->off_loc);
    if (const char* e = getenv("FM_SHARD_LOCATE")) h->shard_locate = atoi(e) != 0;
    h->pts = (P4*)h->arena;
    h->psync = (PeerSync*)(h->arena + h->off_sync);
    h->inbox = (int*)(h->arena + h->off_inbox);
    h->cc = (P4*)(h->arena + h->off_cc);
    A(cudaMemsetAsync(h->psync, 0, sizeof(PeerSync), h->stream));
    if (h->inbox_cap) A(cudaMemsetAsync(h->inbox, 0, sizeof(int) * (size_t)h->inbox_cap * size, h->stream));
    A(dalloc(&h->bnd, (size_t)h->inbox_cap));
  }
  A(dalloc(&h->orig, n)); A(dalloc(&h->inv, n)); A(dalloc(&h->v2t, n));
  if (!getenv("FM_NO_RING_CACHE")) {
    A(cudaMalloc((void**)&h->ringc, 32 * (size_t)n));
    A(cudaMemsetAsync(h->ringc, 0xff, 32 * (size_t)n, h->stream));
  }
  A(dalloc(&h->lvl_order, n)); A(dalloc(&h->prev, n)); A(dalloc(&h->next_rev, n));
  A(dalloc(&h->rows, T));
  A(dalloc(&h->rankbuf, T + 1)); A(dalloc(&h->flip_edge, T));
  A(dalloc(&h->role, T)); A(dalloc(&h->amark, T)); A(dalloc(&h->listA, T)); A(dalloc(&h->listB, T)); A(dalloc(&h->fliplist, T / 2 + 1));
  A(dalloc(&h->counters, 32));
  h->err = h->counters + FM_C_ERR;
  A(dalloc(&h->winner, T)); A(dalloc(&h->vote, T)); A(dalloc(&h->mailbox, 3 * T)); A(dalloc(&h->scratch64, 4));
  A(cudaMemsetAsync(h->vote, 0xff, sizeof(u64) * T, h->stream));
  A(cudaMemsetAsync(h->role, 0, sizeof(int) * T, h->stream));
  A(cudaMemsetAsync(h->amark, 0, sizeof(int) * T, h->stream));
  A(cudaMemsetAsync(h->mailbox, 0, sizeof(u64) * 3 * T, h->stream));
  A(cudaMemsetAsync(h->pts, 0, sizeof(P4) * h->n_pad, h->stream));
  int init_counters[32] = {0};
  init_counters[FM_C_STAMP] = 1;  // first stamp
  A(cudaMemcpyAsync(h->counters, init_counters, sizeof init_counters, cudaMemcpyHostToDevice, h->stream));
  A(cudaStreamSynchronize(h->stream));
  // staging: input order coordinates, keys, sort buffers
  double *ix = nullptr, *iy = nullptr, *iz = nullptr;
  A(dalloc(&h->stage, 3 * (size_t)n));
  ix = h->stage; iy = ix + n; iz = iy + n;
  A(dalloc(&k0, n)); A(dalloc(&k1, n)); A(dalloc(&i0, n)); A(dalloc(&dup, 1));
  size_t sort_bytes = 0, scan_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, k0, k1, i0, h->orig, n, 0, 63, h->stream);
  cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, h->rankbuf, h->rankbuf, (int)T + 1, h->stream);
  size_t tsort_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tsort_bytes, (unsigned*)nullptr, (unsigned*)nullptr, (int*)nullptr,
                                  (int*)nullptr, (int)T, 0, 32, h->stream);
  h->cub_bytes = sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
  if (tsort_bytes > h->cub_bytes) h->cub_bytes = tsort_bytes;
  size_t pscan_bytes = 0;
  cub::DeviceScan::InclusiveScan(nullptr, pscan_bytes, h->prev, h->prev, CudaBackend::MaxOp(), n, h->stream);
  if (pscan_bytes > h->cub_bytes) h->cub_bytes = pscan_bytes;
  A(cudaMalloc(&h->cub_tmp, h->cub_bytes));
  int nb = (n + 255) / 256;
  if (lonlat) {
    A(cudaMemcpyAsync(k0, hx, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));  // lon
    A(cudaMemcpyAsync(k1, hy, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));  // lat
    fm_lonlat_to_xyz<<<nb, 256, 0, h->stream>>>(n, (const double*)k0, (const double*)k1, ix, iy, iz);
  } else {
    A(cudaMemcpyAsync(ix, hx, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    A(cudaMemcpyAsync(iy, hy, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    A(cudaMemcpyAsync(iz, hz, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
  }
  fm_morton_keys<<<nb, 256, 0, h->stream>>>(n, ix, iy, iz, k0, i0);
  size_t bytes = h->cub_bytes;
  A(cub::DeviceRadixSort::SortPairs(h->cub_tmp, bytes, k0, k1, i0, h->orig, n, 0, 63, h->stream));
  fm_gather_sorted<<<nb, 256, 0, h->stream>>>(n, h->orig, ix, iy, iz, h->pts, h->inv);
  int big = 0x7fffffff;
  A(cudaMemcpyAsync(dup, &big, sizeof(int), cudaMemcpyHostToDevice, h->stream));
  fm_find_duplicates<<<nb, 256, 0, h->stream>>>(n, k1, h->pts, dup);
  int hdup = 0, o1 = 0;
  A(cudaMemcpyAsync(&hdup, dup, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  A(cudaStreamSynchronize(h->stream));
  A(cudaGetLastError());
  h->n_launch += 4;
  {  // insertion levels: stable sort of the internal indices by level
    int K = dt_level_count(n);
    h->lvl_K = K;
    h->lvl_off.assign(K + 2, 0);
    unsigned *lk = (unsigned*)k0, *lk2 = (unsigned*)k1;
    A(dalloc(&doff, K + 2));
    fm_level_keys<<<nb, 256, 0, h->stream>>>(n, h->pts, K, lk, i0);
    bytes = h->cub_bytes;
    int key_bits = 1;  // keys run 0..K (K = 12 with levels growing x4, up to 24 with x2)
    while ((1 << key_bits) <= K + 1) key_bits++;
    A(cub::DeviceRadixSort::SortPairs(h->cub_tmp, bytes, lk, lk2, i0, h->lvl_order, n, 0, key_bits, h->stream));
    fm_level_offsets<<<1, 32, 0, h->stream>>>(n, lk2, K, doff);
    A(cudaMemcpyAsync(h->lvl_off.data(), doff, sizeof(int) * (K + 2), cudaMemcpyDeviceToHost, h->stream));
    A(cudaStreamSynchronize(h->stream));
    h->n_launch += 3;
  }
  if (hdup != big) {
    cudaMemcpy(&o1, h->orig + hdup, sizeof(int), cudaMemcpyDeviceToHost);
    snprintf(h->errmsg, sizeof h->errmsg, "DVT %d coincides with another DVT (duplicate generator)", o1 + 1);
    return bail("duplicate check", cudaSuccess);  // DV:286-288
  }
  free_staging();
#undef A
  *out = h;
  return 0;
}

extern "C" int fm_create(int32_t n, const double* x, const double* y, const double* z, fm_handle** out) {
  return create_impl(n, x, y, z, 0, 0, 1, out);
}
extern "C" int fm_create_lonlat(int32_t n, const double* lon, const double* lat, fm_handle** out) {
  return create_impl(n, lon, lat, nullptr, 1, 0, 1, out);
}
extern "C" int fm_create_dist(int32_t n, const double* x, const double* y, const double* z,
                              int32_t rank, int32_t size, fm_handle** out) {
  return create_impl(n, x, y, z, 0, rank, size, out);
}

extern "C" void fm_destroy(fm_handle* h) {
  if (!h) return;
  if (h->stream) cudaStreamSynchronize(h->stream);
  for (int k = 0; k < 8; k++)
    if (h->peer_base[k]) cudaIpcCloseMemHandle(h->peer_base[k]);
  {
    auto& m = h->mp;
    void* mptrs[] = {m.ring_edge, m.cells_on_vertex, m.edges_on_vertex, m.n_eoe, m.eoe, m.xyz_edge, m.lonlat_edge,
                     m.edge5, m.area_cell, m.area_tri, m.kite_ring, m.kite_vtx, m.weights, m.sums, m.part, h->qbuf, h->rho_tab};
    for (void* p : mptrs)
      if (p) cudaFree(p);
  }
  {
    auto& b = h->star;
    void* sptrs[] = {b.key, b.skey, b.own, b.idx, b.sidx, b.colstart, b.colmid, b.ring, b.deg, b.stri, b.listA, b.listB, b.listE, b.listS, b.sc,
                     b.off, b.spts, b.tmp};
    for (void* p : sptrs)
      if (p) cudaFree(p);
  }
  void* ptrs[] = {h->arena, h->bnd, h->ringc, h->orig, h->inv, h->v2t, h->lvl_order, h->prev, h->next_rev, h->rows, h->rankbuf, h->flip_edge,
                  h->role, h->amark, h->listA, h->listB, h->fliplist, h->counters, h->winner, h->vote,
                  h->mailbox, h->scratch64, h->cub_tmp, h->cclon, h->cclat,
                  h->deg, h->ring_tri, h->ring_nbr, h->ecount, h->de, h->ve_vvt, h->ve_vc, h->stage};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  if (h->pinned) cudaFreeHost(h->pinned);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  for (auto& p : h->pending) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
  for (auto e : h->ev_pool) cudaEventDestroy(e);
  if (h->warm_graph) cudaGraphExecDestroy(h->warm_graph);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

extern "C" const char* fm_last_error(const fm_handle* h) { return h ? h->errmsg : g_create_err; }

// One readback for everything the host needs after a batch of launches: the eight counters,
// the error bits and the displacement word.  Synchronises the stream.
struct Readback {
  int counters[32];
  int err;
  int stamp;
  double maxdisp;
  // state of a batch of warm iterations (FM_T_*)
  int halt, iters, sub, maxq;
  long long flips;
  double batch_maxdisp;
};
static cudaError_t readback_enqueue(fm_handle* h) {
  return cudaMemcpyAsync(h->pinned + 16, h->counters, 32 * sizeof(int), cudaMemcpyDeviceToHost, h->stream);
}
static int readback(fm_handle* h, Readback* r, const char* where, bool enqueued = false) {
  if (h->cuda_failed) return 1;
  cudaError_t e = enqueued ? cudaSuccess : readback_enqueue(h);
  if (e == cudaSuccess) e = h->sync();
  if (e != cudaSuccess) { h->cuda_failed = true; h->fail("CUDA error in %s: %s", where, cudaGetErrorString(e)); return 1; }
  memcpy(r->counters, h->pinned + 16, sizeof r->counters);
  r->err = r->counters[FM_C_ERR];
  r->stamp = r->counters[FM_C_STAMP];
  memcpy(&r->maxdisp, r->counters + FM_C_MAXDISP, sizeof(double));
  r->halt = r->counters[FM_T_HALT]; r->iters = r->counters[FM_T_ITERS];
  r->sub = r->counters[FM_T_SUB]; r->maxq = r->counters[FM_T_MAXQ];
  memcpy(&r->flips, r->counters + FM_T_FLIPS, sizeof(long long));
  memcpy(&r->batch_maxdisp, r->counters + FM_T_MAXDISP, sizeof(double));
  if (r->halt) cudaMemsetAsync(h->counters + FM_T_HALT, 0, sizeof(int), h->stream);  // the batch is over
  if (r->err) {
    cudaMemsetAsync(h->err, 0, sizeof(int), h->stream);
    h->fail("%s: device reported error bits %d (%s%s%s%s)", where, r->err,
            (r->err & FM_E_WALK) ? "point walk did not terminate " : "",
            (r->err & FM_E_TOPOLOGY) ? "degenerate triangle / broken ring " : "",
            (r->err & FM_E_OVERFLOW) ? "flip rounds did not terminate " : "",
            (r->err & FM_E_PEER) ? "a peer rank did not reach the barrier" : "");
    return 1;
  }
  if (r->stamp > (1 << 27)) {  // stamps are about to run out: start over with clean words
    size_t T = 2 * (size_t)h->n - 4;
    int one = 1;
    cudaMemsetAsync(h->vote, 0xff, sizeof(u64) * T, h->stream);
    cudaMemsetAsync(h->role, 0, sizeof(int) * T, h->stream);
    cudaMemsetAsync(h->amark, 0, sizeof(int) * T, h->stream);
    cudaMemsetAsync(h->mailbox, 0, sizeof(u64) * 3 * T, h->stream);
    cudaMemcpyAsync(h->counters + FM_C_STAMP, &one, sizeof(int), cudaMemcpyHostToDevice, h->stream);
    cudaStreamSynchronize(h->stream);
  }
  return 0;
}

static void account(fm_handle* h, const Readback& r, int rounds) {
  h->tot_rounds += rounds; h->tot_sub += r.counters[5]; h->tot_flips += r.counters[2];
  if (r.counters[6] > h->max_active) h->max_active = r.counters[6];
}

// Renumbers the triangle slots in spatial order (see fm_tri_keys).  Scratch: the flip
// bookkeeping arrays, which hold nothing between builds; they are re-initialised after.
static int sort_triangles(fm_handle* h) {
  int T = h->n_tri, nb = (T + 255) / 256;
  unsigned *key = (unsigned*)h->flip_edge, *key2 = (unsigned*)h->role;
  int *idx = h->listA, *order = h->listB, *newid = h->amark;
  Row* row2 = (Row*)h->cc;  // 32T bytes; circumcentres are recomputed after every build
  h->prof_begin();
  fm_tri_keys<<<nb, 256, 0, h->stream>>>(T, h->rows, key, idx);
  size_t bytes = h->cub_bytes;
  int bits = 1;
  while ((1LL << bits) < h->n) bits++;
  FM_CUDA(cub::DeviceRadixSort::SortPairs(h->cub_tmp, bytes, key, key2, idx, order, T, 0, bits, h->stream));
  fm_tri_newid<<<nb, 256, 0, h->stream>>>(T, order, newid);
  fm_tri_permute<<<nb, 256, 0, h->stream>>>(T, order, newid, h->rows, row2);
  fm_remap_v2t<<<(h->n + 255) / 256, 256, 0, h->stream>>>(h->n, newid, h->v2t);
  FM_CUDA(cudaMemcpyAsync(h->rows, row2, sizeof(Row) * T, cudaMemcpyDeviceToDevice, h->stream));
  FM_CUDA(cudaMemsetAsync(h->role, 0, sizeof(int) * T, h->stream));
  FM_CUDA(cudaMemsetAsync(h->amark, 0, sizeof(int) * T, h->stream));
  h->prof_end("sort_triangles", T);
  h->n_launch += 4;
  return 0;
}

static int repair_async(fm_handle* h, bool sharded = false, bool in_batch = false);

// What follows every cold build, whichever way the rows were made: the ring cache is void, the slot ranges of
// a sharded mesh are recomputed (slots are sorted by smallest vertex), counters read back.
static int cold_finish(fm_handle* h, int rounds) {
  if (h->ringc) FM_CUDA(cudaMemsetAsync(h->ringc, 0xff, 32 * (size_t)h->n, h->stream));  // every slot was renumbered
  h->sweep_begin = 0; h->sweep_end = h->n_tri;
  if (h->size > 1) {  // the slots that go with this rank's cells (slots are sorted by smallest vertex now)
    fm_tri_range<<<1, 32, 0, h->stream>>>(h->n_tri, h->rows, h->n_pad / h->size, h->size, h->rankbuf);
    int rg[9] = {0};
    FM_CUDA(cudaMemcpyAsync(rg, h->rankbuf, sizeof(int) * (h->size + 1), cudaMemcpyDeviceToHost, h->stream));
    FM_CUDA(cudaStreamSynchronize(h->stream));
    h->sweep_begin = rg[h->rank]; h->sweep_end = rg[h->rank + 1];
    h->n_launch++;
    bool fits = true;  // decided from ALL ranks' ranges, so that every rank decides the same
    for (int k = 0; k < h->size; k++) fits = fits && (rg[k + 1] - rg[k] <= h->inbox_cap);
    if (!fits) {  // a very uneven point set: equal parts instead
      h->sweep_begin = (int)((long long)h->n_tri * h->rank / h->size);
      h->sweep_end = (int)((long long)h->n_tri * (h->rank + 1) / h->size);
    }
    if (h->warm_graph) { cudaGraphExecDestroy(h->warm_graph); h->warm_graph = nullptr; }  // the range is baked in
  }
  Readback r;
  if (readback(h, &r, "delaunay_triangulation")) return 1;
  account(h, r, rounds);
  h->n_cold++;
  h->have_dt = true; h->have_cc = false; h->have_vd = false; h->have_edges = false; h->mp.have = false;
  h->dt_current = true;
  return 0;
}

// ---- cold build by stars (star_kernels.cuh).  Returns 0 = built and verified, 1 = refused (the caller runs
// the insertion build: too few points, FM_COLD_BUILD=flip, a star that did not close, an inconsistency),
// -1 = CUDA error.  Synchronous.
static const int kStarMinPoints = 4096;
static const double kStarWideKappa = 2.5, kStarWideNear = 2.7;   // wide mode: cell of 2.5 spacings, ~20 expected near candidates
static int star_alloc(fm_handle* h) {
  auto& b = h->star;
  if (b.ready) return 0;
  if (b.unavailable) return 1;
  static const double kappa = getenv("FM_STAR_KAPPA") ? atof(getenv("FM_STAR_KAPPA")) : 2.0;
  const int n = h->n;
  const double hh = sqrt(4.0 * 3.14159265358979323846 / n);
  int G = (int)(2.0 / ((kappa > 0.5 ? kappa : 2.0) * hh));
  if (G > kStarMaxG) G = kStarMaxG;
  if (G < 4) { b.unavailable = true; return 1; }
  b.G = G;
  b.Gw = (int)(2.0 / (kStarWideKappa * hh));   // the wide mode's coarser grid (irregular sets)
  if (b.Gw > G) b.Gw = G;
  if (b.Gw < 4) b.Gw = 4;
  size_t sort_bytes = 0, scan_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, b.key, b.skey, b.idx, b.sidx, n, 0, 32, h->stream);
  cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, b.off, b.off, n + 1, h->stream);
  b.tmp_bytes = sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
  bool ok = true;
  auto A = [&](cudaError_t e) { ok = ok && e == cudaSuccess; };
  A(dalloc(&b.key, n)); A(dalloc(&b.skey, n)); A(dalloc(&b.own, n)); A(dalloc(&b.idx, n)); A(dalloc(&b.sidx, n));
  A(dalloc(&b.colstart, (size_t)G * G + 1)); A(dalloc(&b.colmid, (size_t)G * G)); A(dalloc(&b.ring, (size_t)n * kStarCap)); A(dalloc(&b.deg, n));
  A(dalloc(&b.stri, (size_t)n * kStarCap)); A(dalloc(&b.listA, n)); A(dalloc(&b.listB, n)); A(dalloc(&b.listE, n)); A(dalloc(&b.listS, n));
  A(dalloc(&b.sc, 8)); A(dalloc(&b.off, (size_t)n + 1)); A(dalloc(&b.spts, n)); A(cudaMalloc(&b.tmp, b.tmp_bytes));
  if (ok) {
    A(cudaMemsetAsync(b.ring, 0, sizeof(int) * (size_t)n * kStarCap, h->stream));
    A(cudaMemsetAsync(b.stri, 0, sizeof(int) * (size_t)n * kStarCap, h->stream));
    A(cudaMemsetAsync(b.deg, 0, sizeof(int) * (size_t)n, h->stream));
  }
  if (!ok) {  // not enough memory for the ~300 bytes per generator: the insertion build needs none of it
    cudaGetLastError();
    void* sptrs[] = {b.key, b.skey, b.own, b.idx, b.sidx, b.colstart, b.colmid, b.ring, b.deg, b.stri, b.listA, b.listB, b.listE, b.listS, b.sc,
                     b.off, b.spts, b.tmp};
    for (void* p : sptrs)
      if (p) cudaFree(p);
    b = fm_handle::StarBuf();
    b.unavailable = true;
    return 1;
  }
  {  // the widest retry pass wants more than the default 48 KB of dynamic shared memory
    const int want = (int)(sizeof(int) * (2 * (2 * kStarMaxR + 1) * (2 * kStarMaxR + 1) + 4 * kStarListLong) * FM_STAR_BLOCK);
    if (cudaFuncSetAttribute(fm_star_retry, cudaFuncAttributeMaxDynamicSharedMemorySize, want) != cudaSuccess ||
        cudaFuncSetAttribute(fm_star_all, cudaFuncAttributeMaxDynamicSharedMemorySize, want) != cudaSuccess ||
        cudaFuncSetAttribute(fm_star_sample, cudaFuncAttributeMaxDynamicSharedMemorySize, want) != cudaSuccess ||
        cudaFuncSetAttribute(fm_star_slow, cudaFuncAttributeMaxDynamicSharedMemorySize, want) != cudaSuccess ||
        cudaFuncSetAttribute(fm_star_slow_near, cudaFuncAttributeMaxDynamicSharedMemorySize, want) != cudaSuccess ||
        cudaFuncSetAttribute(fm_star_exact, cudaFuncAttributeMaxDynamicSharedMemorySize, want) != cudaSuccess) {
      cudaGetLastError();
      b.unavailable = true;
      return 1;
    }
  }
  b.ready = true;
  return 0;
}

static int triangulate_star(fm_handle* h) {
  const char* mode = getenv("FM_COLD_BUILD");   // read at every build: the tests switch between the two
  const bool off = mode && !strcmp(mode, "flip");
  if (off || h->n < kStarMinPoints) return 1;
  if (h->star_skip > 0) { h->star_skip--; return 1; }   // refused a moment ago: the same generators, moved a little
  if (star_alloc(h)) return 1;
#define S(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { h->cuda_failed = true; \
    h->fail("CUDA error %s at %s:%d", cudaGetErrorString(e_), __FILE__, __LINE__); return -1; } } while (0)
  auto& b = h->star;
  const int n = h->n, T = 2 * n - 4;
  StarCtx s;
  memset(&s, 0, sizeof s);
  s.n = n;
  s.pts = h->pts; s.key = b.key; s.idx = b.idx; s.skey = b.skey; s.sidx = b.sidx; s.spts = b.spts; s.colstart = b.colstart; s.colmid = b.colmid;
  s.ring = b.ring; s.deg = b.deg; s.own = b.own; s.off = b.off; s.stri = b.stri;
  s.list_exact = b.listE; s.list_slow = b.listS; s.sc = b.sc; s.row = h->rows; s.v2t = h->v2t;
  const int nb = (n + 255) / 256;
  cudaStream_t st = h->stream;
  int* lists[2] = {b.listA, b.listB};
  auto smem = [&](int r) { return sizeof(int) * (2 * (2 * r + 1) * (2 * r + 1) + 4 * s.list_cap) * FM_STAR_BLOCK; };
  const size_t smem_full = sizeof(int) * 2 * (2 * kStarMaxR + 1) * (2 * kStarMaxR + 1) * FM_STAR_BLOCK;
  static const double near_scale = getenv("FM_STAR_NEAR") ? atof(getenv("FM_STAR_NEAR")) : 1.0;
  static const double max_frac = getenv("FM_STAR_MAX_IRREGULAR") ? atof(getenv("FM_STAR_MAX_IRREGULAR")) : 0.10;
  static const bool no_wide = !(getenv("FM_STAR_WIDE") && atoi(getenv("FM_STAR_WIDE")) != 0);   // measured: no faster than insertion
  // Two settings.  Regular (lattice-like sets, relaxed meshes, graded or not): cell of 2 spacings, ~7.5 expected
  // near candidates, 16-entry lists.  Wide (irregular sets -- uniform random points want the neighbours within
  // ~2.3 spacings, a third of them more than 16): cell of 2.5 spacings, ~20 expected, 40-entry lists.  The wide
  // one works (same triangles) but measures no faster than insertion + flips on uniform random points -- 14.8
  // against 13.9 ms at 2.6M, 48.4 against 49.3 ms at 10M (profiles/r02_ab_star_wide.txt) -- so it is OFF unless
  // FM_STAR_WIDE=1.  Which setting: 1/64 of the generators (whole blocks, so that the lanes of a warp are
  // neighbours as in the real pass) go through the first pass without recording anything; a setting is taken
  // when at most a tenth of them (wide: 15 %) need the full wrap, a wider block or three rounds of the near
  // list.  A handle remembers that it needed the wide one (and tries the regular one again every fourth build:
  // a Lloyd run makes the set regular).  Neither: insertion build.
  bool chosen = false;
  const int first = (h->star_mode == 1 && (h->star_wide_builds & 3) != 3 && !no_wide) ? 1 : 0;
  for (int attempt = 0; attempt < 2 && !chosen; attempt++) {
    const int mode = attempt == 0 ? first : 1 - first;
    if (mode == 1 && no_wide) continue;
    const int G = mode ? b.Gw : b.G;
    s.G = G; s.half_g = 0.5 * G; s.cell = 2.0 / G;
    s.list_cap = mode ? kStarListWide : kStarList;
    s.near2 = (near_scale > 0.0 ? near_scale : 1.0) * (mode ? kStarWideNear : 1.0);
    // 1. grid: keys, sort by cell, records in sorted order, column starts
    h->prof_begin();
    S(cudaMemsetAsync(b.sc, 0, 8 * sizeof(int), st));
    fm_star_kernel<KStarKey><<<nb, 256, 0, st>>>(s, n);
    int bits = 1;
    while ((1ULL << bits) < (unsigned long long)G * G * G) bits++;
    size_t bytes = b.tmp_bytes;
    S(cub::DeviceRadixSort::SortPairs(b.tmp, bytes, b.key, b.skey, b.idx, b.sidx, n, 0, bits, st));
    fm_star_kernel<KStarGather><<<nb, 256, 0, st>>>(s, n);
    fm_star_kernel<KStarCols><<<(G * G + 1 + 255) / 256, 256, 0, st>>>(s, G * G + 1);
    h->prof_end("star_grid", n);
    h->n_launch += 5;
    // 2. the sample
    s.r = 1; s.in_idx = 1; s.out_idx = 0; s.list_out = lists[0]; s.list_in = lists[1];
    const int every = 64, nblk = (n + FM_STAR_BLOCK - 1) / FM_STAR_BLOCK, sblk = (nblk + every - 1) / every;
    s.dry = 1;
    h->prof_begin();
    fm_star_sample<<<sblk, FM_STAR_BLOCK, smem(1), st>>>(s, n, every);
    h->prof_end("star_sample", sblk * FM_STAR_BLOCK);
    s.dry = 0;
    int sc0[8];
    S(cudaMemcpyAsync(sc0, b.sc, sizeof sc0, cudaMemcpyDeviceToHost, st));
    S(h->sync());
    S(cudaMemsetAsync(b.sc, 0, 8 * sizeof(int), st));
    const double frac = (double)(sc0[FM_SC_RETRIED] + sc0[FM_SC_MULTI]) / ((double)sblk * FM_STAR_BLOCK);
    h->star_irregular = frac;
    if (frac <= (mode ? 1.5 * max_frac : max_frac)) {
      chosen = true;
      if (mode) h->star_wide_builds++; else h->star_wide_builds = 0;
      h->star_mode = mode;
    }
  }
  if (!chosen) { h->n_star_fallback++; h->star_last[2] = -1; h->star_skip = 3; return 1; }
  // 3. stars: all generators with a block of radius 1, then the ones that need more / the exact predicates
  h->prof_begin();
  fm_star_all<<<(n + FM_STAR_BLOCK - 1) / FM_STAR_BLOCK, FM_STAR_BLOCK, smem(1), st>>>(s, n);
  h->prof_end("star_wrap", n);
  h->prof_begin();
  for (int r = 2; r <= kStarMaxR; r++) {
    s.r = r; s.in_idx = (r & 1) ? 1 : 0; s.out_idx = s.in_idx ^ 1;
    s.list_in = lists[s.in_idx]; s.list_out = lists[s.out_idx];
    S(cudaMemsetAsync(b.sc + s.out_idx, 0, sizeof(int), st));
    fm_star_retry<<<h->sms * (r == 2 ? 4 : 1), FM_STAR_BLOCK, smem(r), st>>>(s);   // (few generators get here: a small grid starts and ends quickly)
  }
  {  // the overflowed lists: the list form again with long lists, then (list A, counted in sc[0]) the full wrap
    StarCtx sl = s;
    sl.list_cap = kStarListLong; sl.list_in = lists[0];
    S(cudaMemsetAsync(b.sc + FM_SC_LIST0, 0, sizeof(int), st));
    const size_t smem_long = sizeof(int) * (2 * (2 * kStarMaxR + 1) * (2 * kStarMaxR + 1) + 4 * kStarListLong) * FM_STAR_BLOCK;
    fm_star_slow_near<<<h->sms * 4, FM_STAR_BLOCK, smem_long, st>>>(sl);
    fm_star_slow<<<h->sms * 2, FM_STAR_BLOCK, smem_full, st>>>(sl);
  }
  fm_star_exact<<<h->sms * 2, FM_STAR_BLOCK, smem_full, st>>>(s);
  h->prof_end("star_retry", 0);
  // 3. rows
  h->prof_begin();
  S(cudaMemsetAsync(b.off + n, 0, sizeof(u64), st));
  size_t bytes = b.tmp_bytes;
  S(cub::DeviceScan::ExclusiveSum(b.tmp, bytes, b.off, b.off, n + 1, st));
  fm_star_kernel<KStarSlots><<<nb, 256, 0, st>>>(s, n);
  fm_star_kernel<KStarRows><<<nb, 256, 0, st>>>(s, n);
  h->prof_end("star_rows", n);
  h->n_launch += 8 + (kStarMaxR - 1);
  S(cudaGetLastError());
  int sc[8];
  u64 tot = 0;
  S(cudaMemcpyAsync(sc, b.sc, sizeof sc, cudaMemcpyDeviceToHost, st));
  S(cudaMemcpyAsync(&tot, b.off + n, sizeof tot, cudaMemcpyDeviceToHost, st));
  S(h->sync());
  const int owned = (int)(tot & 0xffffffffULL), degsum = (int)(tot >> 32);
  h->star_last[0] = sc[FM_SC_SLOW]; h->star_last[1] = sc[FM_SC_EXACT]; h->star_last[2] = sc[FM_SC_FAIL]; h->star_last[3] = sc[FM_SC_BROKEN];
  h->star_last[4] = owned; h->star_last[5] = degsum; h->star_last[6] = 0; h->star_last[7] = 0;
  if (sc[FM_SC_FAIL] || sc[FM_SC_BROKEN] || owned != T || degsum != 3 * T) {   // the rows are rewritten by the insertion build
    h->n_star_fallback++;
    return 1;
  }
  h->n_tri = T;
  S(cudaMemsetAsync(h->role, 0, sizeof(int) * (size_t)T, st));
  S(cudaMemsetAsync(h->amark, 0, sizeof(int) * (size_t)T, st));
  // 4. the check: orientation of every triangle and the local Delaunay test of every edge (the warm sweep;
  // it also writes the circumcentres).  A consistent closed surface that passes is THE triangulation.
  fm_bump_stamp<<<1, 1, 0, st>>>(h->counters, 2);
  h->have_dt = true;
  if (h->ringc) S(cudaMemsetAsync(h->ringc, 0xff, 32 * (size_t)n, st));
  h->sweep_begin = 0; h->sweep_end = T;
  if (repair_async(h)) return -1;
  Readback r;
  if (readback(h, &r, "delaunay_triangulation (stars)")) return -1;
  h->star_last[6] = r.counters[2]; h->star_last[7] = r.counters[3];
  if (r.counters[3] > 0 || r.counters[2] > 0) {   // never seen; the insertion build starts from scratch
    h->n_star_fallback++;
    h->have_dt = false;
    return 1;
  }
  h->n_star++;
  if (cold_finish(h, 0)) return -1;
  h->have_cc = true;   // written by the sweep
  return 0;
#undef S
}

// cold build: by stars when the generators allow it (star_kernels.cuh), else bootstrap + insertion rounds; synchronous
static int triangulate_cold(fm_handle* h) {
  {
    int rc = triangulate_star(h);
    if (rc == 0) return 0;
    if (rc < 0) return 1;   // a CUDA error, not a refusal
  }
  CudaBackend be{h};
  DtCtx c = h->dt_ctx();
  c.cc = nullptr;  // circumcentres are computed once at the end, not per flip
  c.ringc = nullptr;  // the ring cache is invalidated as a whole after the build
  int boot[4];
  if (dt_bootstrap(be, c, boot, h->errmsg, sizeof h->errmsg)) return 1;
  fm_init_loc<<<(h->n + 255) / 256, 256, 0, h->stream>>>(h->n, h->loc, boot[0], boot[1], boot[2], boot[3]);
  h->n_launch++;
  DtStats st;
  int rc = dt_run(be, c, h->listA, h->listB, h->lvl_order, h->lvl_off.data(), h->lvl_K, h->prev, boot, &h->stamp, &st,
                  h->errmsg, sizeof h->errmsg);
  if (h->cuda_failed || rc) return 1;
  h->n_tri = c.n_tri;
  if (h->n_tri != 2 * h->n - 4) { h->fail("triangle count %d != 2N-4", h->n_tri); return 1; }
  if (sort_triangles(h)) return 1;
  return cold_finish(h, st.rounds);
}

// warm repair of the existing connectivity, launches only; the caller reads counters[3]
static void launch_peer_barrier(fm_handle* h, int phase);
// sharded: (Lloyd loop on a mesh with attached peers) each rank sweeps its own triangle slots and
// sends what it finds to all ranks; barrier; every rank applies all requests and runs the same
// flip rounds.  Every rank must call this in step.
static int repair_async(fm_handle* h, bool sharded, bool in_batch) {
  CudaBackend be{h};
  be.in_batch = in_batch;
  DtCtx c = h->dt_ctx();
  c.cc = h->cc;
  DtStats st;
  if (!sharded) {
    int rc = dt_repair_async(be, c, h->listA, h->listB, &h->stamp, &st, h->errmsg, sizeof h->errmsg);
    return (h->cuda_failed || rc) ? 1 : 0;
  }
  int zero[15] = {0};
  int* list[2] = {h->listA, h->listB};
  be.write_ints(c.counters, zero, 15);
  c.stamp = h->stamp;
  c.next_active = list[0];
  c.qcount = c.counters;
  DtCtx cs = c;  // the sweep's view: its own range; findings go on the request / boundary lists
  FanCtx f = {};
  cs.t_begin = h->sweep_begin;
  f.blk_per = h->n_pad / h->size;
  cs.blk_begin = h->blk_begin; cs.blk_end = h->blk_begin + f.blk_per;
  for (int k = 0; k < h->size; k++) {
    char* base = (k == h->rank) ? h->arena : (char*)h->peer_base[k];
    f.inbox[k] = (int*)(base + h->off_inbox) + (size_t)h->rank * h->inbox_cap;
    f.cc[k] = (P4*)(base + h->off_cc);
  }
  cs.req = f.inbox[h->rank];  // this rank's own region of its own inbox
  cs.bnd = h->bnd;
  f.row = h->rows; f.req = cs.req; f.bnd = h->bnd; f.counters = h->counters;
  f.rank = h->rank; f.size = h->size; f.pc = h->peer_ctx;
  be.launch_filtered<KSweepShT>(cs, h->sweep_end - h->sweep_begin);
  h->prof_begin();
  fm_sweep_fanout<<<h->sms, 256, 0, h->stream>>>(f);   // + the barrier, in the tail of its last block
  h->note_cuda(cudaGetLastError(), "sweep_fanout");
  h->prof_end("sweep_fanout", 0);
  h->prof_begin();
  fm_apply_requests<<<h->sms * 2, 256, 0, h->stream>>>(c, h->peer_ctx, h->inbox, h->inbox_cap);
  h->note_cuda(cudaGetLastError(), "apply_requests");
  h->prof_end("apply_requests", 0);
  int rc = be.flip_loop(c, list, 0, true, &h->stamp, &st, h->errmsg, sizeof h->errmsg, -1);
  return (h->cuda_failed || rc) ? 1 : 0;
}

static int circumcentres(fm_handle* h, bool lonlat) {
  size_t T = (size_t)h->n_tri;
  if (lonlat && !h->cclon) { FM_CUDA(dalloc(&h->cclon, T)); FM_CUDA(dalloc(&h->cclat, T)); }
  VdCtx c = h->vd_ctx(nullptr);
  if (!lonlat) { c.cclon = nullptr; c.cclat = nullptr; }
  h->launch<KCircum, VdCtx>(c, h->n_tri);
  h->have_cc = true;
  return h->cuda_failed ? 1 : 0;
}

// triangulation of the current generators, synchronous; warm falls back to a rebuild
static int triangulate(fm_handle* h, bool warm, const char* where) {
  if (warm && h->have_dt) {
    if (repair_async(h)) return 1;
    Readback r;
    if (readback(h, &r, where)) return 1;
    if (r.counters[3] == 0) {
      account(h, r, 0);
      h->n_warm++;
      h->have_cc = true; h->have_vd = false; h->have_edges = false; h->mp.have = false;
      h->dt_current = true;
      return 0;
    }
  }
  return triangulate_cold(h);
}

extern "C" int fm_triangulate(fm_handle* h) { return triangulate(h, false, "fm_triangulate"); }
extern "C" int fm_retriangulate(fm_handle* h, int32_t warm) { return triangulate(h, warm != 0, "fm_retriangulate"); }

extern "C" int fm_counts(const fm_handle* h, int32_t* n_tri, int32_t* n_edge) {
  if (n_tri) *n_tri = 2 * h->n - 4;
  if (n_edge) *n_edge = 3 * h->n - 6;
  return 0;
}

// The generators moved since the connectivity was last brought to Delaunay (fm_lloyd ends with a
// centroid update, fm_set_generators replaces them): repair it first, so that what leaves the
// library is always the triangulation / diagram of the CURRENT generators.
static int ensure_current(fm_handle* h, const char* where) {
  if (h->have_dt && !h->dt_current) return triangulate(h, true, where);
  return 0;
}

extern "C" int fm_get_triangles(fm_handle* h, int32_t* tri, int32_t* adj) {
  if (!h->have_dt) { h->fail("fm_get_triangles before fm_triangulate"); return 1; }
  if (ensure_current(h, "fm_get_triangles")) return 1;
  int T = h->n_tri;
  int *dt = nullptr, *da = nullptr;
  FM_CUDA(dalloc(&dt, 3 * (size_t)T));
  FM_CUDA(dalloc(&da, 3 * (size_t)T));
  fm_export_tris<<<(T + 255) / 256, 256, 0, h->stream>>>(T, h->rows, h->pts, dt, da);
  h->n_launch++;
  if (tri) FM_CUDA(cudaMemcpyAsync(tri, dt, sizeof(int) * 3 * T, cudaMemcpyDeviceToHost, h->stream));
  if (adj) FM_CUDA(cudaMemcpyAsync(adj, da, sizeof(int) * 3 * T, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(h->sync());
  cudaFree(dt); cudaFree(da);
  return 0;
}

extern "C" int fm_voronoi(fm_handle* h, int32_t all) {
  if (!h->have_dt) { h->fail("voronoi_diagram before delaunay_triangulation"); return 1; }
  if (ensure_current(h, "fm_voronoi")) return 1;
  size_t T = (size_t)h->n_tri;
  int n = h->n;
  if (circumcentres(h, true)) return 1;
  if (!h->deg) {
    FM_CUDA(dalloc(&h->deg, n + 1));
    FM_CUDA(dalloc(&h->ring_tri, 3 * T));
    FM_CUDA(dalloc(&h->ring_nbr, 3 * T));
  }
  VdCtx c = h->vd_ctx(nullptr);
  h->launch<KDegree, VdCtx>(c, n);
  size_t bytes = h->cub_bytes;
  FM_CUDA(cudaMemsetAsync(h->deg + n, 0, sizeof(int), h->stream));
  FM_CUDA(cub::DeviceScan::ExclusiveSum(h->cub_tmp, bytes, h->deg, h->deg, n + 1, h->stream));
  h->launch<KRingFill, VdCtx>(c, n);
  h->n_launch++;
  Readback r;
  if (readback(h, &r, "fm_voronoi")) return 1;
  h->have_vd = true;
  if (all) {
    size_t E = 3 * (size_t)n - 6;
    if (!h->ecount) {
      FM_CUDA(dalloc(&h->ecount, n + 1));
      FM_CUDA(dalloc(&h->de, 2 * E)); FM_CUDA(dalloc(&h->ve_vvt, 2 * E)); FM_CUDA(dalloc(&h->ve_vc, 2 * E));
    }
    int nb = (n + 255) / 256;
    fm_edge_count<<<nb, 256, 0, h->stream>>>(n, h->deg, h->ring_nbr, h->ecount);
    FM_CUDA(cudaMemsetAsync(h->ecount + n, 0, sizeof(int), h->stream));
    bytes = h->cub_bytes;
    FM_CUDA(cub::DeviceScan::ExclusiveSum(h->cub_tmp, bytes, h->ecount, h->ecount, n + 1, h->stream));
    fm_edge_fill<<<nb, 256, 0, h->stream>>>(n, h->deg, h->ring_tri, h->ring_nbr, h->ecount, h->de,
                                             h->ve_vvt, h->ve_vc);
    h->n_launch += 3;
    int total = 0;
    FM_CUDA(cudaMemcpyAsync(&total, h->ecount + n, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    FM_CUDA(h->sync());
    if ((size_t)total != E) { h->fail("edge count %d != 3N-6 (euler_formula)", total); return 1; }
    h->have_edges = true;
  }
  return 0;
}

extern "C" int fm_get_circumcentres(fm_handle* h, double* xyz, double* lonlat) {
  if (!h->have_vd) { h->fail("fm_get_circumcentres before fm_voronoi"); return 1; }
  int T = h->n_tri;
  double *d3 = nullptr, *d2 = nullptr;
  if (xyz) FM_CUDA(dalloc(&d3, 3 * (size_t)T));
  if (lonlat) FM_CUDA(dalloc(&d2, 2 * (size_t)T));
  fm_pack_cc<<<(T + 255) / 256, 256, 0, h->stream>>>(T, h->cc, h->cclon, h->cclat, d3, d2);
  h->n_launch++;
  if (xyz) FM_CUDA(cudaMemcpyAsync(xyz, d3, sizeof(double) * 3 * T, cudaMemcpyDeviceToHost, h->stream));
  if (lonlat) FM_CUDA(cudaMemcpyAsync(lonlat, d2, sizeof(double) * 2 * T, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(h->sync());
  if (d3) cudaFree(d3);
  if (d2) cudaFree(d2);
  return 0;
}

extern "C" int fm_get_rings(fm_handle* h, int32_t* off, int32_t* ring_tri, int32_t* ring_nbr) {
  if (!h->have_vd) { h->fail("fm_get_rings before fm_voronoi"); return 1; }
  size_t tot = 3 * (size_t)h->n_tri;
  FM_CUDA(cudaMemcpyAsync(off, h->deg, sizeof(int) * (h->n + 1), cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(cudaMemcpyAsync(ring_tri, h->ring_tri, sizeof(int) * tot, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(cudaMemcpyAsync(ring_nbr, h->ring_nbr, sizeof(int) * tot, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(h->sync());
  return 0;
}

extern "C" int fm_get_edges(fm_handle* h, int32_t* de, int32_t* ve_vvt, int32_t* ve_vc) {
  if (!h->have_edges) { h->fail("fm_get_edges before fm_voronoi(all=1)"); return 1; }
  size_t E2 = 2 * (3 * (size_t)h->n - 6);
  if (de) FM_CUDA(cudaMemcpyAsync(de, h->de, sizeof(int) * E2, cudaMemcpyDeviceToHost, h->stream));
  if (ve_vvt) FM_CUDA(cudaMemcpyAsync(ve_vvt, h->ve_vvt, sizeof(int) * E2, cudaMemcpyDeviceToHost, h->stream));
  if (ve_vc) FM_CUDA(cudaMemcpyAsync(ve_vc, h->ve_vc, sizeof(int) * E2, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(h->sync());
  return 0;
}

// ---- MPAS adapter arithmetic on the device: mpas_mesh_adapt MM:181-345 and
// calc_weightsOnEdge MM:442-561 over the ring / edge tables of fm_voronoi(h, 1)
extern "C" int fm_mpas(fm_handle* h, int32_t strict, int32_t* max_edges) {
  if (!h->have_edges) { h->fail("fm_mpas needs fm_voronoi(h, 1) (mpas_mesh_output after voronoi_diagram(all=.true.))"); return 1; }
  auto& m = h->mp;
  size_t n = (size_t)h->n, T = (size_t)h->n_tri, E = 3 * n - 6;
  if (!m.ring_edge) {
    FM_CUDA(dalloc(&m.ring_edge, 3 * T)); FM_CUDA(dalloc(&m.cells_on_vertex, 3 * T)); FM_CUDA(dalloc(&m.edges_on_vertex, 3 * T));
    FM_CUDA(dalloc(&m.n_eoe, E)); FM_CUDA(dalloc(&m.xyz_edge, 3 * E)); FM_CUDA(dalloc(&m.lonlat_edge, 2 * E));
    FM_CUDA(dalloc(&m.edge5, 5 * E)); FM_CUDA(dalloc(&m.area_cell, n)); FM_CUDA(dalloc(&m.area_tri, T));
    FM_CUDA(dalloc(&m.kite_ring, 3 * T)); FM_CUDA(dalloc(&m.kite_vtx, 3 * T)); FM_CUDA(dalloc(&m.sums, 4));
    FM_CUDA(dalloc(&m.part, 3 * ((T + 127) / 128)));
  }
  MpasCtx c;
  memset(&c, 0, sizeof c);
  c.n = h->n; c.n_tri = h->n_tri; c.n_edge = (int)E;
  c.pts = h->pts; c.inv = h->inv; c.row = h->rows; c.cc = h->cc;
  c.off = h->deg; c.ring_tri = h->ring_tri; c.ring_nbr = h->ring_nbr; c.eoff = h->ecount;
  c.de = h->de; c.ve_vvt = h->ve_vvt;
  c.ring_edge = m.ring_edge; c.xyz_edge = m.xyz_edge; c.lonlat_edge = m.lonlat_edge;
  c.dv = m.edge5; c.dc = m.edge5 + E; c.dv1 = m.edge5 + 2 * E; c.dv2 = m.edge5 + 3 * E; c.angle = m.edge5 + 4 * E;
  c.area_cell = m.area_cell; c.area_tri = m.area_tri; c.kite_ring = m.kite_ring; c.kite_vtx = m.kite_vtx;
  c.cells_on_vertex = m.cells_on_vertex; c.edges_on_vertex = m.edges_on_vertex; c.n_eoe = m.n_eoe;
  c.sums = m.sums; c.maxdeg = h->counters + 7; c.err = h->err;
  c.part = m.part; c.ps = (int)((T + 127) / 128);
  FM_CUDA(cudaMemsetAsync(m.sums, 0, 4 * sizeof(double), h->stream));
  FM_CUDA(cudaMemsetAsync(h->counters + 7, 0, sizeof(int), h->stream));
  const int B = 128;
  auto nb = [&](size_t cnt) { return (unsigned)((cnt + B - 1) / B); };
  h->prof_begin(); fm_mpas_ring_edge<<<nb(n), B, 0, h->stream>>>(c); h->prof_end("mpas_ring_edge", n);
  h->prof_begin(); fm_mpas_edges<<<nb(E), B, 0, h->stream>>>(c); h->prof_end("mpas_edges", E);
  h->prof_begin(); fm_mpas_cells<<<nb(n), B, 0, h->stream>>>(c); h->prof_end("mpas_cells", n);
  h->prof_begin(); fm_mpas_vertices<<<nb(T), B, 0, h->stream>>>(c); h->prof_end("mpas_vertices", T);
  // the three sums of the reference's checks, in a fixed order (no floating-point atomics)
  fm_sum_partials<<<1, 256, 0, h->stream>>>(m.part, (int)nb(n), 0, m.sums);
  fm_sum_partials<<<1, 256, 0, h->stream>>>(m.part + 2 * c.ps, (int)nb(T), 0, m.sums + 1);
  fm_sum_partials<<<1, 256, 0, h->stream>>>(m.part + c.ps, (int)nb(n), 0, m.sums + 2);
  h->n_launch += 3;
  Readback r;
  if (readback(h, &r, "fm_mpas")) return 1;
  int me = r.counters[7];
  FM_CUDA(cudaMemsetAsync(h->counters + 7, 0, sizeof(int), h->stream));  // slot 7 is FM_C_NFLIP in the flip rounds
  if (me < 3 || me > 4096) { h->fail("fm_mpas: bad largest cell degree %d", me); return 1; }
  if (me > m.cap_max_edges) {
    if (m.eoe) { cudaFree(m.eoe); cudaFree(m.weights); m.eoe = nullptr; m.weights = nullptr; }
    FM_CUDA(dalloc(&m.eoe, 2 * (size_t)me * E)); FM_CUDA(dalloc(&m.weights, 2 * (size_t)me * E));
    m.cap_max_edges = me;
  }
  m.max_edges = me;
  c.max_edges = me; c.eoe = m.eoe; c.weights = m.weights;
  h->prof_begin(); fm_mpas_weights<<<nb(E), B, 0, h->stream>>>(c); h->prof_end("mpas_weights", E);
  double sums[4];
  FM_CUDA(cudaMemcpyAsync(sums, m.sums, sizeof sums, cudaMemcpyDeviceToHost, h->stream));
  if (readback(h, &r, "fm_mpas")) return 1;
  if (max_edges) *max_edges = me;
  if (strict) {  // the reference's fatal checks: MM:306, MM:320, MM:339, MM:469-475
    const double four_pi = 4.0 * 3.14159265358979323846;
    const char* what = nullptr;
    if (fabs(sums[0] - four_pi) > 1.0e-10) what = "Too large total cell area error!";
    else if (fabs(sums[1] - four_pi) > 1.0e-10) what = "Too large total triangle area error!";
    else if (fabs(sums[2] - four_pi) > 1.0e-10) what = "Too large total kite area error!";
    else if (sums[3] > 1.0e-10) what = "Internal error! (kite fractions of a cell do not sum to 1)";
    if (what) { h->fail("%s", what); return 1; }
  }
  m.have = true;
  return 0;
}

template <class T>
static int d2h(fm_handle* h, T* dst, const T* src, size_t cnt) {
  if (dst) FM_CUDA(cudaMemcpyAsync(dst, src, sizeof(T) * cnt, cudaMemcpyDeviceToHost, h->stream));
  return 0;
}
#define FM_NEED_MPAS(name) if (!h->mp.have) { h->fail(name " before fm_mpas"); return 1; }
extern "C" int fm_get_mpas_edges(fm_handle* h, double* xyz, double* lonlat, double* dv, double* dc, double* dv1,
                                 double* dv2, double* angle) {
  FM_NEED_MPAS("fm_get_mpas_edges");
  size_t E = 3 * (size_t)h->n - 6;
  auto& m = h->mp;
  if (d2h(h, xyz, m.xyz_edge, 3 * E) || d2h(h, lonlat, m.lonlat_edge, 2 * E) || d2h(h, dv, m.edge5, E) ||
      d2h(h, dc, m.edge5 + E, E) || d2h(h, dv1, m.edge5 + 2 * E, E) || d2h(h, dv2, m.edge5 + 3 * E, E) ||
      d2h(h, angle, m.edge5 + 4 * E, E)) return 1;
  FM_CUDA(h->sync());
  return 0;
}
extern "C" int fm_get_mpas_cells(fm_handle* h, int32_t* edges_on_cell, double* area_cell, double* kite_on_cell) {
  FM_NEED_MPAS("fm_get_mpas_cells");
  auto& m = h->mp;
  size_t tot = 3 * (size_t)h->n_tri;
  if (d2h(h, edges_on_cell, m.ring_edge, tot) || d2h(h, area_cell, m.area_cell, (size_t)h->n) ||
      d2h(h, kite_on_cell, m.kite_ring, tot)) return 1;
  FM_CUDA(h->sync());
  return 0;
}
extern "C" int fm_get_mpas_vertices(fm_handle* h, int32_t* cells_on_vertex, int32_t* edges_on_vertex,
                                    double* area_triangle, double* kite_areas_on_vertex) {
  FM_NEED_MPAS("fm_get_mpas_vertices");
  auto& m = h->mp;
  size_t T = (size_t)h->n_tri;
  if (d2h(h, cells_on_vertex, m.cells_on_vertex, 3 * T) || d2h(h, edges_on_vertex, m.edges_on_vertex, 3 * T) ||
      d2h(h, area_triangle, m.area_tri, T) || d2h(h, kite_areas_on_vertex, m.kite_vtx, 3 * T)) return 1;
  FM_CUDA(h->sync());
  return 0;
}
extern "C" int fm_get_mpas_edges_on_edge(fm_handle* h, int32_t* n_edges_on_edge, int32_t* edges_on_edge,
                                         double* weights_on_edge) {
  FM_NEED_MPAS("fm_get_mpas_edges_on_edge");
  auto& m = h->mp;
  size_t E = 3 * (size_t)h->n - 6, row = 2 * (size_t)m.max_edges;
  if (d2h(h, n_edges_on_edge, m.n_eoe, E) || d2h(h, edges_on_edge, m.eoe, row * E) ||
      d2h(h, weights_on_edge, m.weights, row * E)) return 1;
  FM_CUDA(h->sync());
  return 0;
}

static void launch_peer_barrier(fm_handle* h, int phase) {
  h->prof_begin();
  fm_peer_barrier<<<1, 32, 0, h->stream>>>(h->peer_ctx, phase, h->counters);
  h->note_cuda(cudaGetLastError(), "peer_barrier");
  h->prof_end("peer_barrier", 0);
}

static Density make_density(const fm_density* rho);
static Density make_density_h(fm_handle* h, const fm_density* rho);
static int ensure_density_table(fm_handle* h, const fm_density* rho);
static void launch_centroid_kernel(fm_handle* h, const fm_density* rho, bool guarded, bool dry = false);
// Centroid stage of one iteration.  On a sharded mesh with attached peers the kernel runs
// between two barriers: no rank overwrites generators another rank is still reading, and the
// iteration ends when every block has arrived everywhere.
static void launch_centroid(fm_handle* h, const fm_density* rho, bool guarded, bool closing_barrier = true) {
  if (h->peers_attached) launch_peer_barrier(h, 0);
  launch_centroid_kernel(h, rho, guarded);
  if (h->peers_attached && closing_barrier) launch_peer_barrier(h, 1);
}

static void launch_centroid_kernel(fm_handle* h, const fm_density* rho, bool guarded, bool dry) {
  Density d = make_density_h(h, rho);
  VdCtx c = h->vd_ctx(&d);
  // guarded (warm iteration): skipped when the batch has halted -- converged, error, or a triangle
  // turned over (fm_flip_coop / fm_apply_requests' count turns that into the halt word)
  c.abort_flag = guarded ? h->counters + FM_T_HALT : nullptr;
  int cnt = h->blk_count;
  if (dry) {  // fm_quality: every cell, nothing moved, nothing sent
    c.dry = 1; c.n_peers = 0; c.cell_begin = 0; c.cell_end = h->n; c.ringc = nullptr;
    cnt = h->n;
  }
  if (cnt <= 0) return;
  const int B = d.kind == 0 ? FM_CENTROID_BLOCK : FM_CENTROID_BLOCK_RHO;
  size_t smem = sizeof(int) * kRingCap * B;
  h->prof_begin();
  if (d.kind == 0) fm_centroid_update<true><<<(cnt + B - 1) / B, B, smem, h->stream>>>(c, cnt);
  else if (d.tab) fm_centroid_update<false, true><<<(cnt + B - 1) / B, B, smem, h->stream>>>(c, cnt);
  else fm_centroid_update<false><<<(cnt + B - 1) / B, B, smem, h->stream>>>(c, cnt);
  h->note_cuda(cudaGetLastError(), "centroid_update");
  h->prof_end("centroid_update", cnt);
}

static bool same_density(const fm_density* a, const fm_density& b) {
  fm_density d = {};
  if (a) d = *a;
  else { d.cz = 1; d.alpha = 1; d.gamma = 1; }
  return d.kind == b.kind && d.cx == b.cx && d.cy == b.cy && d.cz == b.cz && d.beta == b.beta &&
         d.alpha == b.alpha && d.gamma == b.gamma;
}

static void launch_iter_end(fm_handle* h) {
  h->prof_begin();
  if (h->peers_attached) fm_iter_end_peers<<<1, 32, 0, h->stream>>>(h->peer_ctx, h->counters);  // + the closing barrier
  else fm_iter_end<<<1, 1, 0, h->stream>>>(h->counters);
  h->note_cuda(cudaGetLastError(), "iter_end");
  h->prof_end("iter_end", 0);
}

// Queues one warm iteration (repair + centroid + fm_iter_end) on the stream; the host does not
// look at anything in between (fm_iter_end takes the decisions, see lloyd_run).  Outside
// profiling the sequence is captured once into a CUDA graph and replayed; every pointer and
// size in it is fixed for the life of the handle, the stamp and the queue lengths live on the
// device.  If capture is not possible the launches are issued directly.
static int warm_iteration_enqueue(fm_handle* h, const fm_density* rho) {
  const int launches = h->peers_attached ? 8 : 5;  // sweep, deferred sweep, flip loop, centroid, iter_end (+ fan-out, 1 barrier, apply)
  if (!h->profile && !h->graph_disabled) {
    if (h->warm_graph && !same_density(rho, h->warm_graph_rho)) {
      cudaGraphExecDestroy(h->warm_graph);
      h->warm_graph = nullptr;
    }
    if (!h->warm_graph) {
      cudaGraph_t g = nullptr;
      long long launches_before = h->n_launch;
      cudaError_t e = cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal);
      bool ok = (e == cudaSuccess);
      if (ok) {
        int rc = repair_async(h, h->peers_attached, true);
        launch_centroid(h, rho, true, false);
        launch_iter_end(h);
        e = cudaStreamEndCapture(h->stream, &g);
        ok = (rc == 0 && e == cudaSuccess && g != nullptr && !h->cuda_failed);
      }
      if (ok) ok = cudaGraphInstantiate(&h->warm_graph, g, 0) == cudaSuccess;
      if (g) cudaGraphDestroy(g);
      h->n_launch = launches_before;
      if (!ok) {
        cudaGetLastError();  // clear the capture error; fall back to direct launches for good
        h->cuda_failed = false;
        h->errmsg[0] = 0;
        h->warm_graph = nullptr;
        h->graph_disabled = true;
      } else {
        fm_density d = {};
        if (rho) d = *rho;
        else { d.cz = 1; d.alpha = 1; d.gamma = 1; }
        h->warm_graph_rho = d;
      }
    }
    if (h->warm_graph) {
      h->n_launch += launches;
      cudaError_t e = cudaGraphLaunch(h->warm_graph, h->stream);
      if (e != cudaSuccess) { h->fail("cudaGraphLaunch: %s", cudaGetErrorString(e)); return 1; }
      return 0;
    }
  }
  if (repair_async(h, h->peers_attached, true)) return 1;
  launch_centroid(h, rho, true, false);
  launch_iter_end(h);
  return 0;
}

// Starts a batch of warm iterations: per-iteration counters and the batch state cleared, the
// tolerance where fm_iter_end reads it.  One small host -> device copy, no synchronisation.
static int batch_begin(fm_handle* h, double tol) {
  int* w = h->pinned + 64;  // 32 ints of pinned staging (the copy is asynchronous: the words must stay put)
  memset(w, 0, 32 * sizeof(int));
  memcpy(w + FM_T_TOL, &tol, sizeof(double));
  // [0..14] and [16..31]; the stamp ([15]) lives on
  FM_CUDA(cudaMemcpyAsync(h->counters, w, 15 * sizeof(int), cudaMemcpyHostToDevice, h->stream));
  FM_CUDA(cudaMemcpyAsync(h->counters + 16, w + 16, 16 * sizeof(int), cudaMemcpyHostToDevice, h->stream));
  return 0;
}

// One cold iteration: rebuild, circumcentres, centroid stage, readback.
static int cold_iteration(fm_handle* h, const fm_density* rho, double* maxdisp) {
  Readback r;
  if (triangulate_cold(h)) return 1;
  if (!h->have_cc && circumcentres(h, false)) return 1;   // (the stars' verifying sweep has written them already)
  launch_centroid(h, rho, false);
  if (readback(h, &r, "fm_lloyd")) return 1;
  h->have_cc = false; h->have_vd = false; h->have_edges = false; h->mp.have = false;
  h->dt_current = false;  // the generators moved after the triangulation
  *maxdisp = r.maxdisp;
  return h->cuda_failed ? 1 : 0;
}

// Up to max_iter Lloyd iterations on this rank's block, stopping when the displacement is <= tol.
// Warm iterations (sweep + cooperative flip loop + centroid + fm_iter_end) are queued in batches
// and the host waits ONCE per batch: the convergence test, the statistics and the decision "a
// triangle turned over -> rebuild" are taken on the device by fm_iter_end after every iteration,
// and the remaining launches of a batch that stopped early do nothing.  A turned-over triangle
// stops the batch BEFORE its centroid stage; the host then redoes that iteration cold and goes on.
static int lloyd_run(fm_handle* h, const fm_density* rho, int max_iter, double tol, int warm, int* iters,
                     double* maxdisp) {
  static const int kBatch = [] {
    const char* e = getenv("FM_LLOYD_BATCH");
    int v = e ? atoi(e) : 32;
    return v >= 1 ? v : 1;
  }();
  int it = 0;
  double md = 0.0;
  if (ensure_density_table(h, rho)) return 1;
  while (it < max_iter) {
    if (!(warm && h->have_dt)) {
      if (cold_iteration(h, rho, &md)) return 1;
      it++;
      if (md <= tol) break;
      continue;
    }
    int batch = max_iter - it < kBatch ? max_iter - it : kBatch;
    if (batch_begin(h, tol)) return 1;
    for (int b = 0; b < batch; b++)
      if (warm_iteration_enqueue(h, rho)) return 1;
    Readback r;
    if (readback(h, &r, "fm_lloyd")) return 1;
    h->tot_sub += r.sub; h->tot_flips += r.flips;
    if (r.maxq > h->max_active) h->max_active = r.maxq;
    h->n_warm += r.iters;
    it += r.iters;
    if (r.iters > 0) {
      md = r.batch_maxdisp;
      h->have_cc = false; h->have_vd = false; h->have_edges = false; h->mp.have = false;
      h->dt_current = false;
    }
    if (r.halt == FM_HALT_CONVERGED) break;
    if (r.halt == FM_HALT_REBUILD) {
      if (cold_iteration(h, rho, &md)) return 1;
      it++;
      if (md <= tol) break;
    } else if (r.halt == FM_HALT_ERROR || (r.halt == 0 && r.iters != batch)) {
      h->fail("fm_lloyd: the device stopped a batch of iterations (halt %d after %d of %d)", r.halt, r.iters, batch);
      return 1;
    }
  }
  if (iters) *iters = it;
  *maxdisp = md;
  return 0;
}

extern "C" int fm_lloyd_local(fm_handle* h, const fm_density* rho, int32_t warm, double* max_disp) {
  double md = 0.0;
  if (lloyd_run(h, rho, 1, -1.0, warm, nullptr, &md)) return 1;
  if (max_disp) *max_disp = md;
  return 0;
}

extern "C" int fm_lloyd(fm_handle* h, const fm_density* rho, int32_t max_iter, double tol, int32_t warm,
                        int32_t* iters, double* max_disp) {
  if (h->size != 1 && !h->peers_attached) {
    h->fail("fm_lloyd on a sharded handle without peers: fm_peer_attach first, or use fm_lloyd_local + all-gather");
    return 1;
  }
  double md = 0.0;
  int it = 0;
  if (lloyd_run(h, rho, max_iter, tol, warm, &it, &md)) return 1;
  if (iters) *iters = it;
  if (max_disp) *max_disp = md;
  return 0;
}

// ---- peer exchange over NVLink (CUDA IPC between the processes of one box)
// handle blob = 64-byte CUDA IPC handle + 16-byte UUID of the device the rank runs on
static int device_uuid(fm_handle* h, char* out16) {
  int dev = 0;
  cudaDeviceProp p;
  FM_CUDA(cudaGetDevice(&dev));
  FM_CUDA(cudaGetDeviceProperties(&p, dev));
  memcpy(out16, &p.uuid, 16);
  return 0;
}
extern "C" int fm_peer_handle(fm_handle* h, void* out80) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t hd;
  FM_CUDA(cudaIpcGetMemHandle(&hd, h->arena));
  memcpy(out80, &hd, sizeof hd);
  return device_uuid(h, (char*)out80 + 64);
}
extern "C" int fm_peer_attach(fm_handle* h, const void* handles) {
  if (h->size < 2 || h->size > 8) { h->fail("fm_peer_attach: size must be 2..8, got %d", h->size); return 1; }
  if (h->peers_attached) { h->fail("fm_peer_attach: already attached"); return 1; }
  FM_CUDA(h->sync());
  char mine[16];
  if (device_uuid(h, mine)) return 1;
  for (int k = 0; k < h->size; k++)
    if (k != h->rank && memcmp(mine, (const char*)handles + 80 * (size_t)k + 64, 16) == 0) {
      // the barrier kernels of two ranks would wait on each other on ONE device, where nothing
      // guarantees that they run at the same time
      h->fail("fm_peer_attach: ranks %d and %d run on the same GPU; the peer exchange needs one GPU per rank "
              "(use fm_lloyd_local + an all-gather instead)", h->rank, k);
      return 1;
    }
  PeerCtx pc = {};
  pc.rank = h->rank; pc.size = h->size;
  {  // how long a rank waits for a peer before it gives up with FM_E_PEER: ~2 s of SM clocks by default;
     // raise it under profilers / kernel replay, or when a rank may be held up on the host
    const char* e = getenv("FM_PEER_TIMEOUT_MS");
    double ms = e ? atof(e) : 2000.0;
    if (!(ms > 0.0)) ms = 2000.0;
    pc.timeout_cycles = (long long)(ms * 2.0e6);  // ~2 GHz
  }
  for (int k = 0; k < h->size; k++) {
    if (k == h->rank) { pc.sync[k] = h->psync; continue; }
    cudaIpcMemHandle_t hd;
    memcpy(&hd, (const char*)handles + 80 * (size_t)k, sizeof hd);
    FM_CUDA(cudaIpcOpenMemHandle(&h->peer_base[k], hd, cudaIpcMemLazyEnablePeerAccess));
    pc.sync[k] = (PeerSync*)((char*)h->peer_base[k] + h->off_sync);
  }
  h->peer_ctx = pc;
  h->peers_attached = true;
  if (h->warm_graph) { cudaGraphExecDestroy(h->warm_graph); h->warm_graph = nullptr; }
  return 0;
}

extern "C" int fm_generator_block(fm_handle* h, void** d_pts, int64_t* n_padded, int64_t* begin, int64_t* count) {
  if (d_pts) *d_pts = h->pts;
  if (n_padded) *n_padded = h->n_pad;
  if (begin) *begin = (int64_t)h->rank * (h->n_pad / h->size);
  if (count) *count = h->n_pad / h->size;
  return cudaStreamSynchronize(h->stream) == cudaSuccess ? 0 : 1;
}

static int xfer_generators(fm_handle* h, double* x, double* y, double* z, bool to_host) {
  int n = h->n;
  double *ix = h->stage, *iy = ix + n, *iz = iy + n;
  int nb = (n + 255) / 256;
  if (to_host) {
    fm_scatter_orig<<<nb, 256, 0, h->stream>>>(n, h->pts, ix, iy, iz);
    FM_CUDA(cudaMemcpyAsync(x, ix, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
    FM_CUDA(cudaMemcpyAsync(y, iy, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
    FM_CUDA(cudaMemcpyAsync(z, iz, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
  } else {
    FM_CUDA(cudaMemcpyAsync(ix, x, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    FM_CUDA(cudaMemcpyAsync(iy, y, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    FM_CUDA(cudaMemcpyAsync(iz, z, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    fm_gather_sorted<<<nb, 256, 0, h->stream>>>(n, h->orig, ix, iy, iz, h->pts, h->inv);
    h->have_cc = false; h->have_vd = false; h->have_edges = false; h->mp.have = false;
    h->dt_current = false;
  }
  h->n_launch++;
  FM_CUDA(h->sync());
  return 0;
}

extern "C" int fm_get_generators(fm_handle* h, double* x, double* y, double* z) {
  return xfer_generators(h, x, y, z, true);
}
extern "C" int fm_set_generators(fm_handle* h, const double* x, const double* y, const double* z) {
  h->star_skip = 0;   // a new set of generators: the stars get their chance again
  return xfer_generators(h, (double*)x, (double*)y, (double*)z, false);
}

// ---- one rank's block of generators (sharded handles): the host side of a rank holds, uploads and
// downloads only the cells it owns
extern "C" int fm_block_ids(fm_handle* h, int32_t* ids) {
  int cnt = h->blk_count;
  if (cnt <= 0) return 0;
  int* d = (int*)h->stage;
  fm_block_ids_kernel<<<(cnt + 255) / 256, 256, 0, h->stream>>>(cnt, h->pts + h->blk_begin, d);
  h->n_launch++;
  FM_CUDA(cudaMemcpyAsync(ids, d, sizeof(int) * cnt, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(h->sync());
  return 0;
}
extern "C" int fm_get_generators_block(fm_handle* h, double* x, double* y, double* z) {
  int cnt = h->blk_count;
  if (cnt <= 0) return 0;
  double *ix = h->stage, *iy = ix + cnt, *iz = iy + cnt;
  fm_block_unpack<<<(cnt + 255) / 256, 256, 0, h->stream>>>(cnt, h->pts + h->blk_begin, ix, iy, iz);
  h->n_launch++;
  FM_CUDA(cudaMemcpyAsync(x, ix, sizeof(double) * cnt, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(cudaMemcpyAsync(y, iy, sizeof(double) * cnt, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(cudaMemcpyAsync(z, iz, sizeof(double) * cnt, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(h->sync());
  return 0;
}
extern "C" int fm_set_generators_block(fm_handle* h, const double* x, const double* y, const double* z) {
  if (h->size > 1 && !h->peers_attached) {
    h->fail("fm_set_generators_block on a sharded handle needs fm_peer_attach (the block is stored into every rank's array)");
    return 1;
  }
  int cnt = h->blk_count;
  double *ix = h->stage, *iy = ix + cnt, *iz = iy + cnt;
  if (cnt > 0) {
    FM_CUDA(cudaMemcpyAsync(ix, x, sizeof(double) * cnt, cudaMemcpyHostToDevice, h->stream));
    FM_CUDA(cudaMemcpyAsync(iy, y, sizeof(double) * cnt, cudaMemcpyHostToDevice, h->stream));
    FM_CUDA(cudaMemcpyAsync(iz, z, sizeof(double) * cnt, cudaMemcpyHostToDevice, h->stream));
    BlockFan f = {};
    f.n_dst = 0;
    f.dst[f.n_dst++] = h->pts + h->blk_begin;
    if (h->peers_attached)
      for (int k = 0; k < h->size; k++)
        if (k != h->rank) f.dst[f.n_dst++] = (P4*)h->peer_base[k] + h->blk_begin;
    h->prof_begin();
    fm_block_pack_fan<<<(cnt + 255) / 256, 256, 0, h->stream>>>(cnt, ix, iy, iz, f);
    h->note_cuda(cudaGetLastError(), "block_fanout");
    h->prof_end("block_fanout", cnt);
  }
  if (h->peers_attached) launch_peer_barrier(h, 3);  // every rank's block has landed everywhere
  h->have_cc = false; h->have_vd = false; h->have_edges = false; h->mp.have = false;
  h->dt_current = false;
  FM_CUDA(h->sync());
  return 0;
}

// ---- mesh-quality statistics (quality_kernels.cuh)
extern "C" int fm_quality(fm_handle* h, const fm_density* rho, fm_quality_stats* out) {
  if (!h->have_dt) { h->fail("fm_quality before fm_triangulate"); return 1; }
  if (ensure_current(h, "fm_quality")) return 1;
  if (ensure_density_table(h, rho)) return 1;
  if (!h->have_cc && circumcentres(h, false)) return 1;
  const int B = 128;
  int n = h->n, T = h->n_tri;
  int nbc = (n + B - 1) / B, nbt = (T + B - 1) / B;
  size_t words = 16 + 6 + 2 + 2 * (size_t)nbt;
  if (!h->qbuf) FM_CUDA(dalloc(&h->qbuf, words));
  double init[24];
  memset(init, 0, sizeof init);
  const double big = 1.0e300;
  init[16] = big; init[18] = big; init[20] = big;   // minima
  FM_CUDA(cudaMemcpyAsync(h->qbuf, init, sizeof init, cudaMemcpyHostToDevice, h->stream));
  QualityCtx q;
  memset(&q, 0, sizeof q);
  q.vd = h->vd_ctx(nullptr);
  q.hist = (unsigned long long*)h->qbuf; q.mm = h->qbuf + 16; q.part = h->qbuf + 24; q.stride = nbt;
  h->prof_begin(); fm_quality_cells<<<nbc, B, 0, h->stream>>>(q); h->prof_end("quality_cells", n);
  h->prof_begin(); fm_quality_triangles<<<nbt, B, 0, h->stream>>>(q); h->prof_end("quality_triangles", T);
  fm_sum_partials<<<1, 256, 0, h->stream>>>(q.part, nbc, 0, h->qbuf + 22);
  fm_sum_partials<<<1, 256, 0, h->stream>>>(q.part + nbt, nbt, 0, h->qbuf + 23);
  h->n_launch += 2;
  // the largest generator-to-centroid chord: the Lloyd kernel over ALL cells, moving nothing
  FM_CUDA(cudaMemsetAsync(h->counters + FM_C_MAXDISP, 0, sizeof(u64), h->stream));
  launch_centroid_kernel(h, rho, false, true);
  double host[24];
  FM_CUDA(cudaMemcpyAsync(host, h->qbuf, sizeof host, cudaMemcpyDeviceToHost, h->stream));
  Readback r;
  if (readback(h, &r, "fm_quality")) return 1;
  FM_CUDA(cudaMemsetAsync(h->counters + FM_C_MAXDISP, 0, sizeof(u64), h->stream));
  memcpy(out->degree_hist, host, 16 * sizeof(int64_t));
  out->area_cell_min = host[16]; out->area_cell_max = host[17];
  out->area_tri_min = host[18]; out->area_tri_max = host[19];
  out->angle_min = host[20]; out->angle_max = host[21];
  out->area_cell_sum = host[22]; out->area_tri_sum = host[23];
  out->max_disp = r.maxdisp;
  return 0;
}

// ---- fm_quality_static: the quality statistics without any allocation (no counterpart in the
// reference; added for callers that cannot call cudaMalloc:
// code under CUDA graph capture, tight serving loops, harnesses that forbid it).
// fm_quality keeps its block partials in a buffer it cudaMallocs on first use (fm_handle::qbuf);
// here the same 24 header words + 2 * ceil(T / 128) partials live in ONE file-scope device array,
// sized for FM_QUALITY_STATIC_MAX_GENERATORS generators (2^22: 131,096 doubles = 1.0 MB, held by
// every GPU that loads the library until the process exits).  All handles and all calls share it
// and it keeps the last call's contents: the header words are re-initialised by every call, every
// partial a call reads is written by that call first, so nothing else has to start at zero; calls
// that overlap on the GPU (two handles = two streams, two host threads) race on it.
// The array's name in host code is a host-side stand-in, not a device address: the address comes
// from cudaGetSymbolAddress, once per device, with that device current.
static const size_t kQualityStaticWords = 24 + 2 * (((size_t)2 * FM_QUALITY_STATIC_MAX_GENERATORS - 4 + 127) / 128);
static __device__ __align__(256) double fm_quality_static_scratch[kQualityStaticWords];

static const int kMaxDevices = 64;
static double* g_quality_static_ptr[kMaxDevices] = {nullptr};
static std::mutex g_quality_static_mu;

// the scratch array's address on the CURRENT device (looked up once per device)
static cudaError_t quality_static_scratch(int dev, double** out) {
  std::lock_guard<std::mutex> lock(g_quality_static_mu);
  if (!g_quality_static_ptr[dev]) {
    double* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, fm_quality_static_scratch);
    if (e != cudaSuccess) return e;
    g_quality_static_ptr[dev] = p;
  }
  *out = g_quality_static_ptr[dev];
  return cudaSuccess;
}

extern "C" int fm_quality_static(fm_handle* h, fm_quality_stats* out) {
  if (!out) { h->fail("fm_quality_static: out is null"); return 1; }
  if (h->n > FM_QUALITY_STATIC_MAX_GENERATORS) {
    h->fail("fm_quality_static: %d generators exceed the static scratch array's limit of %d (FM_QUALITY_STATIC_MAX_GENERATORS); "
            "use fm_quality", h->n, (int)FM_QUALITY_STATIC_MAX_GENERATORS);
    return 1;
  }
  if (!h->have_dt) { h->fail("fm_quality_static before fm_triangulate"); return 1; }
  if (!h->dt_current) {  // a repair may end in a rebuild, whose library calls allocate: not here
    h->fail("fm_quality_static: the generators moved since the last triangulation; call fm_retriangulate first "
            "(this entry point allocates nothing and therefore repairs nothing)");
    return 1;
  }
  int dev = -1;
  FM_CUDA(cudaGetDevice(&dev));
  cudaPointerAttributes pa;
  FM_CUDA(cudaPointerGetAttributes(&pa, h->counters));
  if (dev < 0 || dev >= kMaxDevices || pa.device != dev) {
    h->fail("fm_quality_static: the handle lives on device %d but device %d is current", pa.device, dev);
    return 1;
  }
  double* scratch = nullptr;
  FM_CUDA(quality_static_scratch(dev, &scratch));
  if (!h->have_cc && circumcentres(h, false)) return 1;   // a launch into the arena; no allocation (lonlat = false)
  const int B = 128;
  int n = h->n, T = h->n_tri;
  int nbc = (n + B - 1) / B, nbt = (T + B - 1) / B;
  if (24 + 2 * (size_t)nbt > kQualityStaticWords) { h->fail("fm_quality_static: internal size check failed"); return 1; }
  double init[24];
  memset(init, 0, sizeof init);
  const double big = 1.0e300;
  init[16] = big; init[18] = big; init[20] = big;   // minima
  FM_CUDA(cudaMemcpyAsync(scratch, init, sizeof init, cudaMemcpyHostToDevice, h->stream));
  QualityCtx q;
  memset(&q, 0, sizeof q);
  q.vd = h->vd_ctx(nullptr);
  q.hist = (unsigned long long*)scratch; q.mm = scratch + 16; q.part = scratch + 24; q.stride = nbt;
  fm_quality_cells<<<nbc, B, 0, h->stream>>>(q);
  h->note_cuda(cudaGetLastError(), "quality_cells (static)");
  fm_quality_triangles<<<nbt, B, 0, h->stream>>>(q);
  h->note_cuda(cudaGetLastError(), "quality_triangles (static)");
  fm_sum_partials<<<1, 256, 0, h->stream>>>(q.part, nbc, 0, scratch + 22);
  fm_sum_partials<<<1, 256, 0, h->stream>>>(q.part + nbt, nbt, 0, scratch + 23);
  h->n_launch += 4;
  // the largest generator-to-centroid chord under the uniform density: the Lloyd kernel over all cells, moving nothing
  FM_CUDA(cudaMemsetAsync(h->counters + FM_C_MAXDISP, 0, sizeof(u64), h->stream));
  launch_centroid_kernel(h, nullptr, false, true);
  double host[24];
  FM_CUDA(cudaMemcpyAsync(host, scratch, sizeof host, cudaMemcpyDeviceToHost, h->stream));
  Readback r;
  if (readback(h, &r, "fm_quality_static")) return 1;
  FM_CUDA(cudaMemsetAsync(h->counters + FM_C_MAXDISP, 0, sizeof(u64), h->stream));
  memcpy(out->degree_hist, host, 16 * sizeof(int64_t));
  out->area_cell_min = host[16]; out->area_cell_max = host[17];
  out->area_tri_min = host[18]; out->area_tri_max = host[19];
  out->angle_min = host[20]; out->angle_max = host[21];
  out->area_cell_sum = host[22]; out->area_tri_sum = host[23];
  out->max_disp = r.maxdisp;
  return 0;
}

static Density make_density(const fm_density* rho) {
  Density d;
  d.kind = rho ? rho->kind : 0;
  if (rho) { d.cx = rho->cx; d.cy = rho->cy; d.cz = rho->cz; d.beta = rho->beta; d.alpha = rho->alpha; d.gamma = rho->gamma; }
  else { d.cx = d.cy = 0; d.cz = 1; d.beta = 0; d.alpha = 1; d.gamma = 1; }
  d.k2 = 2.0 / d.alpha;
  d.c2 = -d.beta * d.k2;
  d.tab = nullptr; d.tab_scale = 0.0; d.tab_n = 0;
  return d;
}

// ---- the density function as a table (Density::tab).  rho of kind 1 depends on x only through the
// arc distance d to the centre; written as a function of the chord u = 2 sin(d/2) in [0, 2] it is
// analytic, and on kDensityIntervals equal intervals a degree-7 interpolant at Chebyshev nodes
// reproduces it to round-off for any transition width a mesh can resolve.  The table is CHECKED
// against the formula (25 samples per interval) and only used when its largest relative error is
// below 2e-14; otherwise the kernel evaluates acos / exp directly as before.  FM_DENSITY_TABLE=0
// switches it off.
static const int kDensityIntervals = 512;
// Largest relative error a table interval may have (25 samples against the formula).  The
// interpolant has an ABSOLUTE error of a few 1e-17 (rho <= 1), i.e. a relative 1e-16 / rho where rho
// is small (it bottoms out at gamma); a relative error eps of a weight moves a centroid by at most
// eps times the cell diameter, so 2e-13 keeps the step far inside its 1e-12.  An interval that
// misses it (the last one, next to the antipode of the centre, where the arc is not analytic in the
// chord; transitions much narrower than an interval) is marked and evaluated by the formula.
static const double kDensityTableTol = 2.0e-13;
static double density_of_chord(const Density& d, double u) {
  double arc = 2.0 * asin(u >= 2.0 ? 1.0 : 0.5 * u);
  return 0.5 * (1.0 - d.gamma) * (tanh((d.beta - arc) / d.alpha) + 1.0) + d.gamma;
}
static double density_table(const Density& d, std::vector<double>& tab, int* bad_intervals) {
  const int NI = kDensityIntervals, D = 8;
  int n_bad = 0;
  const double kPiH = 3.14159265358979323846;
  tab.assign((size_t)NI * D, 0.0);
  double T[8][8] = {{0}};   // T[k][m]: coefficient of s^m in the Chebyshev polynomial T_k
  T[0][0] = 1.0; T[1][1] = 1.0;
  for (int k = 2; k < D; k++)
    for (int m = 0; m < D; m++) T[k][m] = (m > 0 ? 2.0 * T[k - 1][m - 1] : 0.0) - T[k - 2][m];
  double worst = 0.0;
  for (int i = 0; i < NI; i++) {
    const double a = 2.0 * i / NI, half = 1.0 / NI, mid = a + half;
    double f[8], c[8];
    for (int j = 0; j < D; j++) f[j] = density_of_chord(d, mid + half * cos(kPiH * (j + 0.5) / D));
    for (int k = 0; k < D; k++) {
      double acc = 0.0;
      for (int j = 0; j < D; j++) acc += f[j] * cos(kPiH * k * (j + 0.5) / D);
      c[k] = acc * (k == 0 ? 1.0 : 2.0) / D;
    }
    double* out = tab.data() + (size_t)i * D;
    for (int m = 0; m < D; m++) {
      double acc = 0.0;
      for (int k = 0; k < D; k++) acc += c[k] * T[k][m];
      out[m] = acc;
    }
    double err_i = 0.0;
    for (int q = 0; q <= 24; q++) {
      const double sx = -1.0 + q / 12.0;
      double p = out[D - 1];
      for (int m = D - 2; m >= 0; m--) p = p * sx + out[m];
      const double want = density_of_chord(d, mid + half * sx);
      const double err = fabs(p / want - 1.0);
      if (!(err <= err_i)) err_i = err;   // (NaN counts as failure)
    }
    if (!(err_i <= kDensityTableTol)) { out[D - 1] = nan(""); n_bad++; }   // this interval is evaluated directly
    else if (err_i > worst) worst = err_i;
  }
  if (bad_intervals) *bad_intervals = n_bad;
  return worst;
}
static bool density_table_wanted() {
  const char* e = getenv("FM_DENSITY_TABLE");
  return !(e && atoi(e) == 0);
}
// Makes sure the handle's table belongs to rho (builds and uploads it when the parameters changed)
static int ensure_density_table(fm_handle* h, const fm_density* rho) {
  if (!rho || rho->kind != 1) return 0;
  if (h->rho_tab && same_density(rho, h->rho_tab_for)) return 0;
  Density d = make_density(rho);
  std::vector<double> tab;
  int bad = 0;
  h->rho_tab_err = density_table(d, tab, &bad);
  h->rho_tab_ok = density_table_wanted() && bad <= kDensityIntervals / 4;   // mostly unfittable: not worth the detour
  if (!h->rho_tab) FM_CUDA(dalloc(&h->rho_tab, (size_t)kDensityIntervals * 8));
  FM_CUDA(cudaMemcpyAsync(h->rho_tab, tab.data(), sizeof(double) * tab.size(), cudaMemcpyHostToDevice, h->stream));
  FM_CUDA(cudaStreamSynchronize(h->stream));
  h->rho_tab_for = *rho;
  if (h->warm_graph) { cudaGraphExecDestroy(h->warm_graph); h->warm_graph = nullptr; }
  return 0;
}
static Density make_density_h(fm_handle* h, const fm_density* rho) {
  Density d = make_density(rho);
  if (d.kind == 1 && h->rho_tab && h->rho_tab_ok && same_density(rho, h->rho_tab_for)) {
    d.tab = h->rho_tab; d.tab_n = kDensityIntervals; d.tab_scale = 0.5 * kDensityIntervals;
  }
  return d;
}

__global__ void fm_density_points(Density d, long long n, const double* x, const double* y, const double* z, double* out) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i < n) out[i] = d.tab ? density_node<true>(d, x[i], y[i], z[i], 1.0) : density_node<false>(d, x[i], y[i], z[i], 1.0);
}
extern "C" int fm_density_eval(const fm_density* rho, int64_t n, const double* x, const double* y, const double* z,
                               double* out) {
  if (n <= 0) return 0;
  double* dev = nullptr;
  size_t bytes = sizeof(double) * (size_t)n;
  if (cudaMalloc((void**)&dev, 4 * bytes) != cudaSuccess) {
    snprintf(g_create_err, sizeof g_create_err, "fm_density_eval: no CUDA device or out of memory");
    return 1;
  }
  bool ok = cudaMemcpy(dev, x, bytes, cudaMemcpyHostToDevice) == cudaSuccess &&
            cudaMemcpy(dev + n, y, bytes, cudaMemcpyHostToDevice) == cudaSuccess &&
            cudaMemcpy(dev + 2 * n, z, bytes, cudaMemcpyHostToDevice) == cudaSuccess;
  double* dtab = nullptr;
  if (ok) {
    Density d = make_density(rho);
    const char* force = getenv("FM_DENSITY_TABLE");
    if (d.kind == 1 && force && atoi(force) == 1) {   // FM_DENSITY_TABLE=1: through the table the Lloyd kernel uses (default: the formula)
      std::vector<double> tab;
      int bad = 0;
      density_table(d, tab, &bad);
      if (bad <= kDensityIntervals / 4 && cudaMalloc((void**)&dtab, sizeof(double) * tab.size()) == cudaSuccess &&
          cudaMemcpy(dtab, tab.data(), sizeof(double) * tab.size(), cudaMemcpyHostToDevice) == cudaSuccess) {
        d.tab = dtab; d.tab_n = kDensityIntervals; d.tab_scale = 0.5 * kDensityIntervals;
      }
    }
    fm_density_points<<<(unsigned)((n + 255) / 256), 256>>>(d, n, dev, dev + n, dev + 2 * n, dev + 3 * n);
    ok = cudaMemcpy(out, dev + 3 * n, bytes, cudaMemcpyDeviceToHost) == cudaSuccess;
  }
  if (dtab) cudaFree(dtab);
  cudaFree(dev);
  if (!ok) snprintf(g_create_err, sizeof g_create_err, "fm_density_eval: %s", cudaGetErrorString(cudaGetLastError()));
  return ok ? 0 : 1;
}

extern "C" int fm_host_alloc(size_t bytes, void** out) {
  return cudaMallocHost(out, bytes) == cudaSuccess ? 0 : 1;
}
extern "C" int fm_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? 0 : 1; }

extern "C" int fm_timer_begin(fm_handle* h) {
  FM_CUDA(h->sync());
  FM_CUDA(cudaEventRecord(h->ev0, h->stream));
  return 0;
}
extern "C" int fm_timer_end(fm_handle* h, double* ms) {
  FM_CUDA(cudaEventRecord(h->ev1, h->stream));
  FM_CUDA(cudaEventSynchronize(h->ev1));
  float f = 0;
  FM_CUDA(cudaEventElapsedTime(&f, h->ev0, h->ev1));
  if (ms) *ms = f;
  return 0;
}

extern "C" int fm_stats(const fm_handle* h, int64_t* s) {
  s[0] = h->tot_rounds; s[1] = h->tot_sub; s[2] = h->tot_flips; s[3] = h->max_active;
  s[4] = h->n_launch; s[5] = h->n_cold; s[6] = h->n_warm; s[7] = h->n_star;
  return 0;
}
extern "C" int fm_star_stats(const fm_handle* h, int64_t* out) {
  out[0] = h->n_star; out[1] = h->n_star_fallback;
  for (int k = 0; k < 8; k++) out[2 + k] = h->star_last[k];
  return 0;
}
extern "C" int fm_stats_reset(fm_handle* h) {
  h->tot_rounds = h->tot_sub = h->tot_flips = 0;
  h->max_active = 0;
  h->n_launch = h->n_cold = h->n_warm = 0;
  h->n_star = h->n_star_fallback = 0;
  h->sync();
  h->prof.clear();
  return 0;
}
extern "C" int fm_profile_enable(fm_handle* h, int32_t on) {
  h->sync();
  h->profile = on != 0;
  return 0;
}
extern "C" int fm_profile_get(fm_handle* h, int32_t cap, char (*names)[32], int64_t* launches,
                              double* ms, int64_t* threads, int32_t* rows) {
  h->sync();
  int k = 0;
  for (auto& r : h->prof) {
    if (k >= cap) break;
    snprintf(names[k], 32, "%s", r.name.c_str());
    launches[k] = r.launches; ms[k] = r.ms; threads[k] = r.threads;
    k++;
  }
  *rows = k;
  return 0;
}
#ifdef FM_COUNT_STEPS
extern "C" int fm_debug_counters(fm_handle* h, int32_t* out) {
  cudaMemcpy(out, h->counters, 16 * sizeof(int), cudaMemcpyDeviceToHost);
  return 0;
}
#endif
// fp64 FMA rate of the current device in TFLOP/s: fm_dfma_chains on every SM, repeated until
// at least min_ms of device time have been measured (CUDA events on the default stream).
extern "C" int fm_fp64_peak(double min_ms, double* tflops) {
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess ||
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 1;
  double* sink = nullptr;
  cudaEvent_t e0, e1;
  if (cudaMalloc((void**)&sink, sizeof(double)) != cudaSuccess) return 1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int blocks = sms * 8, threads = 256, iters = 4096;
  const double flop_per_launch = 2.0 * 64.0 * iters * (double)blocks * threads;   // 64 DFMA per loop trip
  fm_dfma_chains<<<blocks, threads>>>(16, 1.0, sink);   // warm-up
  cudaDeviceSynchronize();
  double best = 0.0, total_ms = 0.0;
  for (int rep = 0; rep < 64 && (rep < 3 || total_ms < min_ms); rep++) {
    cudaEventRecord(e0);
    fm_dfma_chains<<<blocks, threads>>>(iters, 1.0 + rep, sink);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) break;
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    total_ms += ms;
    if (ms > 0 && flop_per_launch / (ms * 1e-3) / 1e12 > best) best = flop_per_launch / (ms * 1e-3) / 1e12;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
  if (tflops) *tflops = best;
  return (best > 0.0 && cudaGetLastError() == cudaSuccess) ? 0 : 1;
}

extern "C" int fm_device_info(char* name, int32_t cap, int32_t* sms) {
  cudaDeviceProp p;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&p, dev) != cudaSuccess) return 1;
  if (name) snprintf(name, cap, "%s", p.name);
  if (sms) *sms = p.multiProcessorCount;
  return 0;
}
