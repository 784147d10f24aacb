// quality_kernels.cuh -- mesh-quality statistics on the device (last sentence;
// north_star: "with the final mesh-quality statistics matching"): cell-degree histogram, smallest /
// largest / total Voronoi cell area and Delaunay triangle area by the reference's angle-excess
// formula (calc_area, /root/reference/src/sphere_geometry_mod.F90:198-227, over calc_sphere_angle
// SG:247-285 -- the sums are the reference's run-time checks MM:306 and MM:320), smallest / largest
// interior angle of the Delaunay triangles, and the largest generator-to-centroid chord (the
// Lloyd kernel run dry).  CPU restatement: oracle/fishman_oracle.c fo_quality.
//
// Sums are deterministic (fm_block_sum / fm_sum_partials, mpas_kernels.cuh): a fixed-shape tree
// inside each block, one partial per block, a second fixed-order pass over the partials.
#pragma once
#include "mpas_kernels.cuh"

namespace fm {

__device__ __forceinline__ void fm_atomic_min_pos(double* dst, double v) {  // v >= 0: the bit image is ordered
  atomicMin((unsigned long long*)dst, (unsigned long long)__double_as_longlong(v));
}
__device__ __forceinline__ void fm_atomic_max_pos(double* dst, double v) {
  atomicMax((unsigned long long*)dst, (unsigned long long)__double_as_longlong(v));
}

struct QualityCtx {
  VdCtx vd;
  unsigned long long* hist;  // 16 bins: cells of degree d, d >= 15 in the last one
  double* mm;                // [0] min cell area, [1] max, [2] min triangle area, [3] max, [4] min angle, [5] max angle
  double* part;              // block partials: cells at [0, stride), triangles at [stride, 2 stride)
  int stride;
};

// one thread per cell (internal index): ring walk (extract_incident_DT_and_linked_DVT, DV:704-743),
// area of the polygon of circumcentres
__global__ void __launch_bounds__(128) fm_quality_cells(QualityCtx q) {
  __shared__ double smem[4];
  __shared__ unsigned int sh_hist[16];
  const VdCtx& c = q.vd;
  if (threadIdx.x < 16) sh_hist[threadIdx.x] = 0;
  __syncthreads();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  double area = 0.0;
  if (i < c.n) {
    int k, t0 = ring_start(c, i, k), t = t0, d = 0, tlast = t0;
    do { tlast = t; t = ring_next(c, t, k); d++; } while (t != t0 && d < 4096);
    if (t != t0) atomicOr(c.err, FM_E_TOPOLOGY);
    atomicAdd(&sh_hist[d < 15 ? d : 15], 1u);
    P3 pm = m_p3(c.cc[tlast]), p0 = m_p3(c.cc[t0]);
    t = ring_start(c, i, k);
    for (int j = 0; j < d; j++) {
      t = ring_next(c, t, k);
      P3 pn = m_p3(c.cc[t]);
      area += m_angle(pm, p0, pn);   // calc_area, SG:198-227
      pm = p0; p0 = pn;
    }
    area -= (d - 2) * kPi;
    fm_atomic_min_pos(q.mm + 0, area);
    fm_atomic_max_pos(q.mm + 1, area);
  }
  double s = fm_block_sum(area, smem);
  if (threadIdx.x == 0) q.part[blockIdx.x] = s;
  __syncthreads();
  if (threadIdx.x < 16 && sh_hist[threadIdx.x]) atomicAdd(q.hist + threadIdx.x, (unsigned long long)sh_hist[threadIdx.x]);
}

// one thread per Delaunay triangle: its three interior angles and its area (MM:310-322)
__global__ void __launch_bounds__(128) fm_quality_triangles(QualityCtx q) {
  __shared__ double smem[4];
  const VdCtx& c = q.vd;
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  double area = 0.0;
  if (t < c.n_tri) {
    I4 v = c.row[t].v;
    P3 a = m_p3(c.pts[v.x]), b = m_p3(c.pts[v.y]), cc = m_p3(c.pts[v.z]);
    double A = m_angle(cc, a, b), B = m_angle(a, b, cc), C = m_angle(b, cc, a);
    area = A + B + C - kPi;
    double lo = fmin(A, fmin(B, C)), hi = fmax(A, fmax(B, C));
    fm_atomic_min_pos(q.mm + 2, area);
    fm_atomic_max_pos(q.mm + 3, area);
    fm_atomic_min_pos(q.mm + 4, lo);
    fm_atomic_max_pos(q.mm + 5, hi);
  }
  double s = fm_block_sum(area, smem);
  if (threadIdx.x == 0) q.part[q.stride + blockIdx.x] = s;
}

}  // namespace fm
