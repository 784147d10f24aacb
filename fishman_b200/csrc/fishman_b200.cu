// fishman_b200.cu -- the C ABI of include/fishman_b200.h on top of the sm_100a kernels.
// Device memory, streams and launches only: no host execution path exists here.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdlib.h>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <string>
#include <vector>

#include "../../include/fishman_b200.h"
#include "dt_driver.h"
#include "vd_kernels.cuh"

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

// ------------------------------------------------------------------ kernels
// Generic launcher: one thread per work item running a body from *_kernels.cuh.
template <class K, class Ctx>
__global__ void __launch_bounds__(256) fm_kernel(const Ctx c, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) K::run(c, i);
}

// Centroid + generator update + max-displacement in one pass: the per-cell chord
// displacement is reduced with warp shuffles, then once per block through shared
// memory, and one atomicMax per block reaches HBM (north-star kernel (d)).
__global__ void __launch_bounds__(256) fm_centroid_update(const VdCtx c, int n) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  double disp = 0.0;
  if (j < n) disp = KCentroid::compute(c, c.cell_begin + j);
  for (int o = 16; o > 0; o >>= 1) disp = fmax(disp, __shfl_xor_sync(0xffffffffu, disp, o));
  __shared__ double wmax[8];
  if ((threadIdx.x & 31) == 0) wmax[threadIdx.x >> 5] = disp;
  __syncthreads();
  if (threadIdx.x < 32) {
    double v = (threadIdx.x < (blockDim.x >> 5)) ? wmax[threadIdx.x] : 0.0;
    for (int o = 4; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    if (threadIdx.x == 0 && v > 0.0) atomicMax(c.maxdisp, (u64)__double_as_longlong(v));
  }
}

__global__ void fm_morton_keys(int n, const double* x, const double* y, const double* z, u64* key,
                               int* idx) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { key[i] = morton_key(x[i], y[i], z[i]); idx[i] = i; }
}

// gather into internal order; inv[orig] = internal
__global__ void fm_gather_sorted(int n, const int* orig, const double* x, const double* y,
                                 const double* z, double* sx, double* sy, double* sz, int* inv) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    int o = orig[i];
    sx[i] = x[o]; sy[i] = y[o]; sz[i] = z[o];
    inv[o] = i;
  }
}

// scatter back to input order
__global__ void fm_scatter_orig(int n, const int* orig, const double* sx, const double* sy,
                                const double* sz, double* x, double* y, double* z) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    int o = orig[i];
    x[o] = sx[i]; y[o] = sy[i]; z[o] = sz[i];
  }
}

// duplicates share a Morton key, so they sit in one run of equal keys
__global__ void fm_find_duplicates(int n, const u64* key, const double* x, const double* y,
                                   const double* z, int* dup) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int j = i + 1; j < n && key[j] == key[i]; j++)
    if (x[j] == x[i] && y[j] == y[i] && z[j] == z[i]) { atomicMin(dup, i); return; }
}

__global__ void fm_lonlat_to_xyz(int n, const double* lon, const double* lat, double* x, double* y,
                                 double* z) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {  // cartesian_transform, src/sphere_geometry_mod.F90:76-85
    double cl = cos(lat[i]);
    x[i] = cl * cos(lon[i]); y[i] = cl * sin(lon[i]); z[i] = sin(lat[i]);
  }
}

// block argmax of p . d ; one packed (ordered value bits | index) atomicMax per block
__global__ void fm_argmax_dot(int n, const double* x, const double* y, const double* z, double dx,
                              double dy, double dz, u64* out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  u64 best = 0;
  if (i < n) {
    float v = (float)(x[i] * dx + y[i] * dy + z[i] * dz);
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

__global__ void fm_export_tris(int nt, const I4* tri, const I4* adj, const int* orig, int* otri,
                               int* oadj) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nt) {
    I4 v = tri[t], a = adj[t];
    otri[3 * t] = orig[v.x] + 1; otri[3 * t + 1] = orig[v.y] + 1; otri[3 * t + 2] = orig[v.z] + 1;
    oadj[3 * t] = (a.x >> 2) + 1; oadj[3 * t + 1] = (a.y >> 2) + 1; oadj[3 * t + 2] = (a.z >> 2) + 1;
  }
}

__global__ void fm_pack_cc(int nt, const double* cx, const double* cy, const double* cz,
                           const double* lon, const double* lat, double* xyz, double* lonlat) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nt) {
    if (xyz) { xyz[3 * t] = cx[t]; xyz[3 * t + 1] = cy[t]; xyz[3 * t + 2] = cz[t]; }
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

// ------------------------------------------------------------------- handle
template <class K> struct KName { static const char* get() { return "kernel"; } };
#define FM_KNAME(K, s) template <> struct KName<K> { static const char* get() { return s; } };
FM_KNAME(KLocateVote, "locate_vote")
FM_KNAME(KSplitFlag, "split_flag")
FM_KNAME(KSplit, "split13")
FM_KNAME(KCheck, "flip_check")
FM_KNAME(KMark, "flip_mark")
FM_KNAME(KFlip, "flip22")
FM_KNAME(KPatch, "flip_patch")
FM_KNAME(KValidate, "validate")
FM_KNAME(KQueueAll, "queue_all")
FM_KNAME(KCircum, "circumcentre")
FM_KNAME(KV2T, "vertex_to_tri")
FM_KNAME(KDegree, "ring_degree")
FM_KNAME(KRingFill, "ring_fill")

struct ProfRow {
  std::string name;
  long long launches = 0, threads = 0;
  double ms = 0.0;
};

struct fm_handle {
  int n = 0, n_pad = 0, n_tri = 0, rank = 0, size = 1;
  int blk_begin = 0, blk_count = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, pev0 = nullptr, pev1 = nullptr;
  // generators (internal order) and the permutation
  double *x = nullptr, *y = nullptr, *z = nullptr;
  int *orig = nullptr, *inv = nullptr;
  // triangulation
  I4 *tri = nullptr, *adj = nullptr;
  int *loc = nullptr, *rankbuf = nullptr, *flip_edge = nullptr, *role = nullptr, *amark = nullptr;
  int *listA = nullptr, *listB = nullptr, *counters = nullptr, *err = nullptr;
  u64 *winner = nullptr, *vote = nullptr, *mailbox = nullptr, *scratch64 = nullptr;
  void* cub_tmp = nullptr;
  size_t cub_bytes = 0;
  int* pinned = nullptr;
  int stamp = 1;
  bool have_dt = false, have_vd = false, have_edges = false;
  // Voronoi
  double *ccx = nullptr, *ccy = nullptr, *ccz = nullptr, *cclon = nullptr, *cclat = nullptr;
  int *v2t = nullptr, *deg = nullptr, *ring_tri = nullptr, *ring_nbr = nullptr;
  int *ecount = nullptr, *de = nullptr, *ve_vvt = nullptr, *ve_vc = nullptr;
  double* stage = nullptr;  // 3n doubles: input-order staging of fm_get/set_generators
  // accounting
  DtStats last = {};
  long long n_launch = 0, n_cold = 0, n_warm = 0, tot_rounds = 0, tot_sub = 0, tot_flips = 0;
  int max_active = 0;
  bool profile = false;
  std::vector<ProfRow> prof;
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
  void prof_begin() { if (profile) cudaEventRecord(pev0, stream); }
  void prof_end(const char* name, int n) {
    n_launch++;
    if (!profile) return;
    cudaEventRecord(pev1, stream);
    cudaEventSynchronize(pev1);
    float ms = 0;
    cudaEventElapsedTime(&ms, pev0, pev1);
    ProfRow& r = row(name);
    r.launches++; r.threads += n; r.ms += ms;
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
    c.n = n; c.x = x; c.y = y; c.z = z; c.orig = orig; c.tri = tri; c.adj = adj; c.loc = loc;
    c.winner = winner; c.rank = rankbuf; c.n_tri = n_tri; c.stamp = stamp; c.vote = vote;
    c.flip_edge = flip_edge; c.role = role; c.mailbox = mailbox; c.amark = amark;
    c.counters = counters; c.err = err;
    return c;
  }
  VdCtx vd_ctx(const Density* rho) {
    VdCtx c;
    memset(&c, 0, sizeof c);
    c.n = n; c.n_tri = n_tri; c.x = x; c.y = y; c.z = z; c.orig = orig; c.tri = tri; c.adj = adj;
    c.ccx = ccx; c.ccy = ccy; c.ccz = ccz; c.cclon = cclon; c.cclat = cclat; c.v2t = v2t;
    c.deg = deg; c.ring_tri = ring_tri; c.ring_nbr = ring_nbr;
    if (rho) c.rho = *rho;
    c.cell_begin = blk_begin; c.cell_end = blk_begin + blk_count;
    c.maxdisp = scratch64; c.err = err;
    return c;
  }
};

struct CudaBackend {
  fm_handle* h;
  template <class K>
  void launch(const DtCtx& c, int n) { h->launch<K, DtCtx>(c, n); }
  void fill64(u64* p, u64 v, size_t n) {
    // only all-ones is ever requested: a byte memset does it
    h->note_cuda(cudaMemsetAsync(p, (int)(v & 0xff), n * sizeof(u64), h->stream), "fill64");
  }
  void fill32(int* p, int v, size_t n) {
    h->note_cuda(cudaMemsetAsync(p, v & 0xff, n * sizeof(int), h->stream), "fill32");
  }
  int exclusive_scan(int* p, int n) {
    h->prof_begin();
    cudaMemsetAsync(p + n, 0, sizeof(int), h->stream);
    size_t bytes = h->cub_bytes;
    h->note_cuda(cub::DeviceScan::ExclusiveSum(h->cub_tmp, bytes, p, p, n + 1, h->stream), "scan");
    h->prof_end("cub_scan", n);
    int total = 0;
    read_ints(&total, p + n, 1);
    return total;
  }
  void read_ints(int* dst, const int* src, int n) {
    h->note_cuda(cudaMemcpyAsync(h->pinned, src, sizeof(int) * n, cudaMemcpyDeviceToHost, h->stream),
                 "read_ints");
    h->note_cuda(cudaStreamSynchronize(h->stream), "sync");
    memcpy(dst, h->pinned, sizeof(int) * n);
  }
  void write_ints(int* dst, const int* src, int n) {
    // sources are tiny host constants; zero is the only value ever written
    bool allzero = true;
    for (int i = 0; i < n; i++) allzero = allzero && (src[i] == 0);
    if (allzero) h->note_cuda(cudaMemsetAsync(dst, 0, sizeof(int) * n, h->stream), "write_ints");
    else {
      h->note_cuda(cudaMemcpyAsync(dst, src, sizeof(int) * n, cudaMemcpyHostToDevice, h->stream), "write_ints");
      h->note_cuda(cudaStreamSynchronize(h->stream), "sync");
    }
  }
  int argmax_dot(const DtCtx& c, const double* d) {
    double nn = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]) + 1e-300;
    cudaMemsetAsync(h->scratch64 + 1, 0, sizeof(u64), h->stream);
    fm_argmax_dot<<<(c.n + 255) / 256, 256, 0, h->stream>>>(c.n, c.x, c.y, c.z, d[0] / nn, d[1] / nn,
                                                            d[2] / nn, h->scratch64 + 1);
    h->n_launch++;
    u64 v = 0;
    h->note_cuda(cudaMemcpyAsync(&v, h->scratch64 + 1, sizeof v, cudaMemcpyDeviceToHost, h->stream), "argmax");
    h->note_cuda(cudaStreamSynchronize(h->stream), "sync");
    return 0x7fffffff - (int)(v & 0xffffffffULL);
  }
  P3 fetch_point(const DtCtx& c, int i) {
    P3 p = {0, 0, 0};
    cudaMemcpyAsync(&p.x, c.x + i, sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    cudaMemcpyAsync(&p.y, c.y + i, sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    cudaMemcpyAsync(&p.z, c.z + i, sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    h->note_cuda(cudaStreamSynchronize(h->stream), "fetch_point");
    return p;
  }
  int fetch_orig(const DtCtx& c, int i) {
    int v = 0;
    cudaMemcpyAsync(&v, c.orig + i, sizeof(int), cudaMemcpyDeviceToHost, h->stream);
    h->note_cuda(cudaStreamSynchronize(h->stream), "fetch_orig");
    return v;
  }
  void write_tris(const DtCtx& c, int t0, int cnt, const I4* tri, const I4* adj) {
    cudaMemcpyAsync(c.tri + t0, tri, sizeof(I4) * cnt, cudaMemcpyHostToDevice, h->stream);
    cudaMemcpyAsync(c.adj + t0, adj, sizeof(I4) * cnt, cudaMemcpyHostToDevice, h->stream);
    h->note_cuda(cudaStreamSynchronize(h->stream), "write_tris");
  }
};

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
  if (size < 1 || rank < 0 || rank >= size) { snprintf(g_create_err, sizeof g_create_err, "bad rank/size"); return 1; }
  fm_handle* h = new fm_handle();
  auto bail = [&](const char* what, cudaError_t err) {
    snprintf(g_create_err, sizeof g_create_err, "fm_create: %s: %s", what,
             err == cudaSuccess ? h->errmsg : cudaGetErrorString(err));
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
  A(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  A(cudaEventCreate(&h->ev0)); A(cudaEventCreate(&h->ev1));
  A(cudaEventCreate(&h->pev0)); A(cudaEventCreate(&h->pev1));
  A(cudaMallocHost((void**)&h->pinned, 64 * sizeof(int)));
  A(dalloc(&h->x, h->n_pad)); A(dalloc(&h->y, h->n_pad)); A(dalloc(&h->z, h->n_pad));
  A(dalloc(&h->orig, n)); A(dalloc(&h->inv, n));
  A(dalloc(&h->tri, T)); A(dalloc(&h->adj, T));
  A(dalloc(&h->loc, n)); A(dalloc(&h->rankbuf, T + 1)); A(dalloc(&h->flip_edge, T));
  A(dalloc(&h->role, T)); A(dalloc(&h->amark, T)); A(dalloc(&h->listA, T)); A(dalloc(&h->listB, T));
  A(dalloc(&h->counters, 8)); A(dalloc(&h->err, 1));
  A(dalloc(&h->winner, T)); A(dalloc(&h->vote, T)); A(dalloc(&h->mailbox, 3 * T)); A(dalloc(&h->scratch64, 4));
  A(cudaMemsetAsync(h->vote, 0xff, sizeof(u64) * T, h->stream));
  A(cudaMemsetAsync(h->role, 0, sizeof(int) * T, h->stream));
  A(cudaMemsetAsync(h->amark, 0, sizeof(int) * T, h->stream));
  A(cudaMemsetAsync(h->mailbox, 0, sizeof(u64) * 3 * T, h->stream));
  A(cudaMemsetAsync(h->err, 0, sizeof(int), h->stream));
  A(cudaMemsetAsync(h->x, 0, sizeof(double) * h->n_pad, h->stream));
  A(cudaMemsetAsync(h->y, 0, sizeof(double) * h->n_pad, h->stream));
  A(cudaMemsetAsync(h->z, 0, sizeof(double) * h->n_pad, h->stream));
  // staging: input order coordinates, keys, sort buffers
  double *ix = nullptr, *iy = nullptr, *iz = nullptr;
  u64 *k0 = nullptr, *k1 = nullptr;
  int *i0 = nullptr, *dup = nullptr;
  A(dalloc(&ix, n)); A(dalloc(&iy, n)); A(dalloc(&iz, n));
  A(dalloc(&k0, n)); A(dalloc(&k1, n)); A(dalloc(&i0, n)); A(dalloc(&dup, 1));
  size_t sort_bytes = 0, scan_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, k0, k1, i0, h->orig, n, 0, 63, h->stream);
  cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, h->rankbuf, h->rankbuf, (int)T + 1, h->stream);
  h->cub_bytes = sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
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
  fm_gather_sorted<<<nb, 256, 0, h->stream>>>(n, h->orig, ix, iy, iz, h->x, h->y, h->z, h->inv);
  int big = 0x7fffffff;
  A(cudaMemcpyAsync(dup, &big, sizeof(int), cudaMemcpyHostToDevice, h->stream));
  fm_find_duplicates<<<nb, 256, 0, h->stream>>>(n, k1, h->x, h->y, h->z, dup);
  int hdup = 0, o1 = 0;
  A(cudaMemcpyAsync(&hdup, dup, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  A(cudaStreamSynchronize(h->stream));
  A(cudaGetLastError());
  h->n_launch += 4;
  if (hdup != big) {
    cudaMemcpy(&o1, h->orig + hdup, sizeof(int), cudaMemcpyDeviceToHost);
    snprintf(h->errmsg, sizeof h->errmsg, "DVT %d coincides with another DVT (duplicate generator)", o1 + 1);
    cudaFree(ix); cudaFree(iy); cudaFree(iz); cudaFree(k0); cudaFree(k1); cudaFree(i0); cudaFree(dup);
    return bail("duplicate check", cudaSuccess);  // DV:286-288
  }
  cudaFree(ix); cudaFree(iy); cudaFree(iz); cudaFree(k0); cudaFree(k1); cudaFree(i0); cudaFree(dup);
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
  void* ptrs[] = {h->x, h->y, h->z, h->orig, h->inv, h->tri, h->adj, h->loc, h->rankbuf, h->flip_edge,
                  h->role, h->amark, h->listA, h->listB, h->counters, h->err, h->winner, h->vote,
                  h->mailbox, h->scratch64, h->cub_tmp, h->ccx, h->ccy, h->ccz, h->cclon, h->cclat,
                  h->v2t, h->deg, h->ring_tri, h->ring_nbr, h->ecount, h->de, h->ve_vvt, h->ve_vc, h->stage};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  if (h->pinned) cudaFreeHost(h->pinned);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->pev0) cudaEventDestroy(h->pev0);
  if (h->pev1) cudaEventDestroy(h->pev1);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

extern "C" const char* fm_last_error(const fm_handle* h) { return h ? h->errmsg : g_create_err; }

static int check_device_err(fm_handle* h, const char* where) {
  if (h->cuda_failed) return 1;
  int bits = 0;
  cudaError_t e = cudaMemcpyAsync(h->pinned, h->err, sizeof(int), cudaMemcpyDeviceToHost, h->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) { h->fail("CUDA error in %s: %s", where, cudaGetErrorString(e)); return 1; }
  bits = h->pinned[0];
  if (bits) {
    cudaMemsetAsync(h->err, 0, sizeof(int), h->stream);
    h->fail("%s: device reported error bits %d (%s%s)", where, bits,
            (bits & FM_E_WALK) ? "point walk did not terminate " : "",
            (bits & FM_E_TOPOLOGY) ? "degenerate triangle / broken ring" : "");
    return 1;
  }
  return 0;
}

static void account(fm_handle* h, const DtStats& st) {
  h->last = st;
  h->tot_rounds += st.rounds; h->tot_sub += st.subrounds; h->tot_flips += st.flips;
  if (st.max_active > h->max_active) h->max_active = st.max_active;
}

// cold build: bootstrap + insertion rounds
static int triangulate_cold(fm_handle* h) {
  CudaBackend be{h};
  DtCtx c = h->dt_ctx();
  int boot[4];
  if (dt_bootstrap(be, c, boot, h->errmsg, sizeof h->errmsg)) return 1;
  fm_init_loc<<<(h->n + 255) / 256, 256, 0, h->stream>>>(h->n, h->loc, boot[0], boot[1], boot[2], boot[3]);
  h->n_launch++;
  DtStats st;
  int rc = dt_run(be, c, h->listA, h->listB, &h->stamp, &st, h->errmsg, sizeof h->errmsg);
  if (h->cuda_failed) return 1;
  if (rc) return 1;
  h->n_tri = c.n_tri;
  if (h->n_tri != 2 * h->n - 4) { h->fail("triangle count %d != 2N-4", h->n_tri); return 1; }
  account(h, st);
  h->n_cold++;
  h->have_dt = true; h->have_vd = false; h->have_edges = false;
  return 0;
}

// warm repair of the existing connectivity; falls back to a cold build
static int triangulate_warm(fm_handle* h) {
  if (!h->have_dt) return triangulate_cold(h);
  CudaBackend be{h};
  DtCtx c = h->dt_ctx();
  DtStats st;
  int rc = dt_repair(be, c, h->listA, h->listB, &h->stamp, &st, h->errmsg, sizeof h->errmsg);
  if (h->cuda_failed) return 1;
  if (rc == 2) return triangulate_cold(h);
  if (rc) return 1;
  account(h, st);
  h->n_warm++;
  h->have_vd = false; h->have_edges = false;
  return 0;
}

extern "C" int fm_triangulate(fm_handle* h) {
  if (triangulate_cold(h)) return 1;
  return check_device_err(h, "fm_triangulate");
}

extern "C" int fm_retriangulate(fm_handle* h, int32_t warm) {
  if (warm ? triangulate_warm(h) : triangulate_cold(h)) return 1;
  return check_device_err(h, "fm_retriangulate");
}

extern "C" int fm_counts(const fm_handle* h, int32_t* n_tri, int32_t* n_edge) {
  if (n_tri) *n_tri = 2 * h->n - 4;
  if (n_edge) *n_edge = 3 * h->n - 6;
  return 0;
}

extern "C" int fm_get_triangles(fm_handle* h, int32_t* tri, int32_t* adj) {
  if (!h->have_dt) { h->fail("fm_get_triangles before fm_triangulate"); return 1; }
  int T = h->n_tri;
  int *dt = nullptr, *da = nullptr;
  FM_CUDA(dalloc(&dt, 3 * (size_t)T));
  FM_CUDA(dalloc(&da, 3 * (size_t)T));
  fm_export_tris<<<(T + 255) / 256, 256, 0, h->stream>>>(T, h->tri, h->adj, h->orig, dt, da);
  h->n_launch++;
  if (tri) FM_CUDA(cudaMemcpyAsync(tri, dt, sizeof(int) * 3 * T, cudaMemcpyDeviceToHost, h->stream));
  if (adj) FM_CUDA(cudaMemcpyAsync(adj, da, sizeof(int) * 3 * T, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(cudaStreamSynchronize(h->stream));
  cudaFree(dt); cudaFree(da);
  return 0;
}

// circumcentres + vertex->triangle map (what the Lloyd step needs every iteration)
static int voronoi_core(fm_handle* h, bool lonlat) {
  size_t T = (size_t)h->n_tri;
  if (!h->ccx) {
    FM_CUDA(dalloc(&h->ccx, T)); FM_CUDA(dalloc(&h->ccy, T)); FM_CUDA(dalloc(&h->ccz, T));
    FM_CUDA(dalloc(&h->v2t, h->n));
  }
  if (lonlat && !h->cclon) { FM_CUDA(dalloc(&h->cclon, T)); FM_CUDA(dalloc(&h->cclat, T)); }
  VdCtx c = h->vd_ctx(nullptr);
  if (!lonlat) { c.cclon = nullptr; c.cclat = nullptr; }
  FM_CUDA(cudaMemsetAsync(h->v2t, 0x7f, sizeof(int) * h->n, h->stream));
  h->launch<KCircum, VdCtx>(c, h->n_tri);
  h->launch<KV2T, VdCtx>(c, h->n_tri);
  return h->cuda_failed ? 1 : 0;
}

extern "C" int fm_voronoi(fm_handle* h, int32_t all) {
  if (!h->have_dt) { h->fail("voronoi_diagram before delaunay_triangulation"); return 1; }
  size_t T = (size_t)h->n_tri;
  int n = h->n;
  if (voronoi_core(h, true)) return 1;
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
  if (check_device_err(h, "fm_voronoi")) return 1;
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
    FM_CUDA(cudaStreamSynchronize(h->stream));
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
  fm_pack_cc<<<(T + 255) / 256, 256, 0, h->stream>>>(T, h->ccx, h->ccy, h->ccz, h->cclon, h->cclat, d3, d2);
  h->n_launch++;
  if (xyz) FM_CUDA(cudaMemcpyAsync(xyz, d3, sizeof(double) * 3 * T, cudaMemcpyDeviceToHost, h->stream));
  if (lonlat) FM_CUDA(cudaMemcpyAsync(lonlat, d2, sizeof(double) * 2 * T, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(cudaStreamSynchronize(h->stream));
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
  FM_CUDA(cudaStreamSynchronize(h->stream));
  return 0;
}

extern "C" int fm_get_edges(fm_handle* h, int32_t* de, int32_t* ve_vvt, int32_t* ve_vc) {
  if (!h->have_edges) { h->fail("fm_get_edges before fm_voronoi(all=1)"); return 1; }
  size_t E2 = 2 * (3 * (size_t)h->n - 6);
  if (de) FM_CUDA(cudaMemcpyAsync(de, h->de, sizeof(int) * E2, cudaMemcpyDeviceToHost, h->stream));
  if (ve_vvt) FM_CUDA(cudaMemcpyAsync(ve_vvt, h->ve_vvt, sizeof(int) * E2, cudaMemcpyDeviceToHost, h->stream));
  if (ve_vc) FM_CUDA(cudaMemcpyAsync(ve_vc, h->ve_vc, sizeof(int) * E2, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(cudaStreamSynchronize(h->stream));
  return 0;
}

// one Lloyd iteration on this rank's block; the displacement stays on the device
static int lloyd_iteration(fm_handle* h, const fm_density* rho, int warm) {
  if (warm ? triangulate_warm(h) : triangulate_cold(h)) return 1;
  if (voronoi_core(h, false)) return 1;
  Density d;
  d.kind = rho ? rho->kind : 0;
  if (rho) { d.cx = rho->cx; d.cy = rho->cy; d.cz = rho->cz; d.beta = rho->beta; d.alpha = rho->alpha; d.gamma = rho->gamma; }
  else { d.cx = d.cy = 0; d.cz = 1; d.beta = 0; d.alpha = 1; d.gamma = 1; }
  VdCtx c = h->vd_ctx(&d);
  FM_CUDA(cudaMemsetAsync(h->scratch64, 0, sizeof(u64), h->stream));
  int cnt = h->blk_count;
  if (cnt > 0) {
    h->prof_begin();
    fm_centroid_update<<<(cnt + 255) / 256, 256, 0, h->stream>>>(c, cnt);
    h->note_cuda(cudaGetLastError(), "centroid_update");
    h->prof_end("centroid_update", cnt);
  }
  h->have_vd = false;  // the generators moved: circumcentres and rings are stale
  h->have_edges = false;
  return h->cuda_failed ? 1 : 0;
}

static int read_maxdisp(fm_handle* h, double* out) {
  u64 bits = 0;
  FM_CUDA(cudaMemcpyAsync(&bits, h->scratch64, sizeof bits, cudaMemcpyDeviceToHost, h->stream));
  FM_CUDA(cudaStreamSynchronize(h->stream));
  memcpy(out, &bits, sizeof bits);
  return 0;
}

extern "C" int fm_lloyd_local(fm_handle* h, const fm_density* rho, int32_t warm, double* max_disp) {
  if (lloyd_iteration(h, rho, warm)) return 1;
  double md = 0.0;
  if (read_maxdisp(h, &md)) return 1;
  if (max_disp) *max_disp = md;
  return check_device_err(h, "fm_lloyd_local");
}

extern "C" int fm_lloyd(fm_handle* h, const fm_density* rho, int32_t max_iter, double tol, int32_t warm,
                        int32_t* iters, double* max_disp) {
  if (h->size != 1) { h->fail("fm_lloyd on a sharded handle: use fm_lloyd_local + all-gather"); return 1; }
  double md = 0.0;
  int it = 0;
  for (; it < max_iter;) {
    if (lloyd_iteration(h, rho, warm)) return 1;
    it++;
    if (read_maxdisp(h, &md)) return 1;  // the convergence test needs the scalar on the host
    if (md <= tol) break;
  }
  if (iters) *iters = it;
  if (max_disp) *max_disp = md;
  return check_device_err(h, "fm_lloyd");
}

extern "C" int fm_generator_block(fm_handle* h, void** dx, void** dy, void** dz, int64_t* n_padded,
                                  int64_t* begin, int64_t* count) {
  if (dx) *dx = h->x;
  if (dy) *dy = h->y;
  if (dz) *dz = h->z;
  if (n_padded) *n_padded = h->n_pad;
  if (begin) *begin = (int64_t)h->rank * (h->n_pad / h->size);
  if (count) *count = h->n_pad / h->size;
  return cudaStreamSynchronize(h->stream) == cudaSuccess ? 0 : 1;
}

static int xfer_generators(fm_handle* h, double* x, double* y, double* z, bool to_host) {
  int n = h->n;
  if (!h->stage) FM_CUDA(dalloc(&h->stage, 3 * (size_t)n));
  double *ix = h->stage, *iy = ix + n, *iz = iy + n;
  int nb = (n + 255) / 256;
  if (to_host) {
    fm_scatter_orig<<<nb, 256, 0, h->stream>>>(n, h->orig, h->x, h->y, h->z, ix, iy, iz);
    FM_CUDA(cudaMemcpyAsync(x, ix, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
    FM_CUDA(cudaMemcpyAsync(y, iy, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
    FM_CUDA(cudaMemcpyAsync(z, iz, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
  } else {
    FM_CUDA(cudaMemcpyAsync(ix, x, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    FM_CUDA(cudaMemcpyAsync(iy, y, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    FM_CUDA(cudaMemcpyAsync(iz, z, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    fm_gather_sorted<<<nb, 256, 0, h->stream>>>(n, h->orig, ix, iy, iz, h->x, h->y, h->z, h->inv);
    h->have_vd = false; h->have_edges = false;
  }
  h->n_launch++;
  FM_CUDA(cudaStreamSynchronize(h->stream));
  return 0;
}

extern "C" int fm_get_generators(fm_handle* h, double* x, double* y, double* z) {
  return xfer_generators(h, x, y, z, true);
}
extern "C" int fm_set_generators(fm_handle* h, const double* x, const double* y, const double* z) {
  return xfer_generators(h, (double*)x, (double*)y, (double*)z, false);
}

extern "C" int fm_host_alloc(size_t bytes, void** out) {
  return cudaMallocHost(out, bytes) == cudaSuccess ? 0 : 1;
}
extern "C" int fm_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? 0 : 1; }

extern "C" int fm_timer_begin(fm_handle* h) {
  FM_CUDA(cudaStreamSynchronize(h->stream));
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
  s[4] = h->n_launch; s[5] = h->n_cold; s[6] = h->n_warm; s[7] = 0;
  return 0;
}
extern "C" int fm_stats_reset(fm_handle* h) {
  h->tot_rounds = h->tot_sub = h->tot_flips = 0;
  h->max_active = 0;
  h->n_launch = h->n_cold = h->n_warm = 0;
  h->prof.clear();
  return 0;
}
extern "C" int fm_profile_enable(fm_handle* h, int32_t on) {
  h->profile = on != 0;
  return 0;
}
extern "C" int fm_profile_get(fm_handle* h, int32_t cap, char (*names)[32], int64_t* launches,
                              double* ms, int64_t* threads, int32_t* rows) {
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
extern "C" int fm_device_info(char* name, int32_t cap, int32_t* sms) {
  cudaDeviceProp p;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&p, dev) != cudaSuccess) return 1;
  if (name) snprintf(name, cap, "%s", p.name);
  if (sms) *sms = p.multiProcessorCount;
  return 0;
}
