// vd_kernels.cuh -- per-thread bodies of the Voronoi / Lloyd kernels.
//
//   KCircum     circumcentres            DV:379-391, calc_circumcircle DV:1323-1348
//   KV2T        one incident triangle per vertex (the reference's DVT%incDT head)
//   KDegree / KRingFill   CSR vertex rings   extract_incident_DT_and_linked_DVT DV:704-743,
//                                             cell rings DV:394-409
//   KCentroid   density-weighted centroid + generator update + displacement
//               (no reference code: SURVEY.md section 7.3 is the specification,
//               restated on the CPU in oracle/fishman_oracle.c fo_centroids)
//   KValidate   orientation check of every triangle after the generators moved
//   KQueueAll   puts every triangle on the flip queue (warm restart)
// "DV" = /root/reference/src/delaunay_voronoi_mod.F90.
#pragma once
#include <string.h>

#include "dt_kernels.cuh"

namespace fm {

#if defined(__CUDA_ARCH__)
#define FM_ATOMIC_MIN32(p, v) atomicMin((int*)(p), (int)(v))
#else
#define FM_ATOMIC_MIN32(p, v) fm_host_min<int>((int*)(p), (int)(v))
#endif

FM_HD inline u64 dbits(double v) {
#if defined(__CUDA_ARCH__)
  return (u64)__double_as_longlong(v);
#else
  u64 b;
  memcpy(&b, &v, sizeof b);
  return b;
#endif
}

struct Density {   // same layout as fm_density in include/fishman_b200.h
  int kind;        // 0 uniform, 1 tanh refinement
  double cx, cy, cz, beta, alpha, gamma;
};

struct VdCtx {
  int n, n_tri;
  double *x, *y, *z;          // generators, internal order (updated in place by KCentroid)
  const int* orig;            // internal -> original id
  const I4* tri;
  const I4* adj;
  double *ccx, *ccy, *ccz;    // circumcentres per triangle
  double *cclon, *cclat;      // optional (may be null)
  int* v2t;                   // per vertex: lowest incident triangle
  int* deg;                   // per ORIGINAL id: ring length, then its exclusive scan
  int *ring_tri, *ring_nbr;   // CSR payload, 1-based output ids
  Density rho;
  int cell_begin, cell_end;   // range of internal cells this rank updates
  u64* maxdisp;               // bits of the largest chord displacement (>= 0, so ordered)
  int* err;
};

static const double kEps = 2.220446049250313e-16;  // eps of src/const_mod.F90:16
static const double kPi2 = 6.283185307179586476925287;

// ------------------------------------------------------------ circumcentres
struct KCircum {
  FM_HD static void run(const VdCtx& c, int t) {
    I4 v = c.tri[t];
    double ax = c.x[v.x], ay = c.y[v.x], az = c.z[v.x];
    double e3x = c.x[v.y] - ax, e3y = c.y[v.y] - ay, e3z = c.z[v.y] - az;
    double e2x = c.x[v.z] - ax, e2y = c.y[v.z] - ay, e2z = c.z[v.z] - az;
    double nx = e3y * e2z - e3z * e2y;  // cross_product, src/math_mod.F90:12-22
    double ny = e3z * e2x - e3x * e2z;
    double nz = e3x * e2y - e3y * e2x;
    double l = sqrt(nx * nx + ny * ny + nz * nz);
    if (l < kEps) { *c.err |= FM_E_TOPOLOGY; l = 1.0; }  // "collinear", DV:1339-1341
    nx /= l; ny /= l; nz /= l;
    c.ccx[t] = nx; c.ccy[t] = ny; c.ccz[t] = nz;
    if (c.cclon) {  // inverse_cartesian_transform, src/sphere_geometry_mod.F90:112-121
      double lon = atan2(ny, nx);
      if (lon < 0.0) lon += kPi2;
      c.cclon[t] = lon;
      c.cclat[t] = asin(nz);
    }
  }
};

struct KV2T {
  FM_HD static void run(const VdCtx& c, int t) {
    I4 v = c.tri[t];
    FM_ATOMIC_MIN32(&c.v2t[v.x], t);
    FM_ATOMIC_MIN32(&c.v2t[v.y], t);
    FM_ATOMIC_MIN32(&c.v2t[v.z], t);
  }
};

// Ring walk around internal vertex i, CCW: from triangle t with i at index k the
// next triangle lies across the edge opposite vertex k+1, and i sits at index
// e+1 of that neighbour (e = the neighbour's edge index stored in adj).
FM_HD inline int ring_start(const VdCtx& c, int i, int& k) {
  int t = c.v2t[i];
  I4 v = c.tri[t];
  k = (v.x == i) ? 0 : ((v.y == i) ? 1 : 2);
  return t;
}
FM_HD inline int ring_next(const VdCtx& c, int t, int& k) {
  int ne = geti(c.adj[t], (k + 1) % 3);
  k = ((ne & 3) + 1) % 3;
  return ne >> 2;
}

struct KDegree {
  FM_HD static void run(const VdCtx& c, int i) {
    int k, t0 = ring_start(c, i, k), t = t0, d = 0;
    do { t = ring_next(c, t, k); d++; } while (t != t0 && d < 4096);
    if (t != t0) *c.err |= FM_E_TOPOLOGY;
    c.deg[c.orig[i]] = d;
  }
};

// Writes the ring of internal vertex i at deg[orig] (now an offset), rotated so
// that it starts at the smallest neighbour id (ring starts are history dependent
// in the reference and not comparable).  ring_nbr[j] lies on the
// edge shared by ring_tri[j] and ring_tri[j+1] (DV:723-738).
struct KRingFill {
  FM_HD static void run(const VdCtx& c, int i) {
    int k, t0 = ring_start(c, i, k), t = t0, d = 0, best = 0x7fffffff, jbest = 0;
    do {
      int nb = c.orig[geti(c.tri[t], (k + 2) % 3)];
      if (nb < best) { best = nb; jbest = d; }
      t = ring_next(c, t, k);
      d++;
    } while (t != t0 && d < 4096);
    int off = c.deg[c.orig[i]];
    t = ring_start(c, i, k);
    for (int j = 0; j < d; j++) {
      int pos = off + (j - jbest + d) % d;
      c.ring_tri[pos] = t + 1;
      c.ring_nbr[pos] = c.orig[geti(c.tri[t], (k + 2) % 3)] + 1;
      t = ring_next(c, t, k);
    }
  }
};

// ------------------------------------------------------------------- Lloyd
FM_HD inline double density_eval(const Density& r, double x, double y, double z) {
  if (r.kind == 0) return 1.0;
  double dot = x * r.cx + y * r.cy + z * r.cz;
  double d = acos(fmax(fmin(dot, 1.0), -1.0));  // calc_arc_length, sphere_geometry_mod.F90:289-296
  return 0.5 * (1.0 - r.gamma) * (tanh((r.beta - d) / r.alpha) + 1.0) + r.gamma;
}

// degree-5 seven-point rule on a triangle (barycentric weights of g, c_j, c_j+1)
#define FM_QA1 0.10128650732345633880
#define FM_QB1 0.79742698535308732240
#define FM_QA2 0.47014206410511508977
#define FM_QB2 0.05971587178976982045
#define FM_QW0 0.225
#define FM_QW1 0.12593918054482715260
#define FM_QW2 0.13239415278850618074

FM_HD inline void quad_node(const Density& r, double w, double l0, double l1, double l2, double g0,
                            double g1, double g2, double a0, double a1, double a2, double b0,
                            double b1, double b2, double& s0, double& s1, double& s2) {
  double q0 = l0 * g0 + l1 * a0 + l2 * b0;
  double q1 = l0 * g1 + l1 * a1 + l2 * b1;
  double q2 = l0 * g2 + l1 * a2 + l2 * b2;
  double rr = sqrt(q0 * q0 + q1 * q1 + q2 * q2);
  q0 /= rr; q1 /= rr; q2 /= rr;
  double ww = w * density_eval(r, q0, q1, q2);
  s0 += ww * q0; s1 += ww * q1; s2 += ww * q2;
}

struct KCentroid {
  // returns the chord displacement of cell i and moves its generator
  FM_HD static double compute(const VdCtx& c, int i) {
    double g0 = c.x[i], g1 = c.y[i], g2 = c.z[i];
    int k, t0 = ring_start(c, i, k), t = t0, d = 0;
    double a0 = c.ccx[t], a1 = c.ccy[t], a2 = c.ccz[t];
    double M0 = 0.0, M1 = 0.0, M2 = 0.0;
    do {
      int tn = ring_next(c, t, k);
      double b0 = c.ccx[tn], b1 = c.ccy[tn], b2 = c.ccz[tn];
      double u0 = a0 - g0, u1 = a1 - g1, u2 = a2 - g2;
      double v0 = b0 - g0, v1 = b1 - g1, v2 = b2 - g2;
      double c0 = u1 * v2 - u2 * v1, c1 = u2 * v0 - u0 * v2, c2 = u0 * v1 - u1 * v0;
      double A = 0.5 * sqrt(c0 * c0 + c1 * c1 + c2 * c2);
      double s0 = 0.0, s1 = 0.0, s2 = 0.0;
      const double third = 1.0 / 3.0;
      quad_node(c.rho, FM_QW0, third, third, third, g0, g1, g2, a0, a1, a2, b0, b1, b2, s0, s1, s2);
      quad_node(c.rho, FM_QW1, FM_QB1, FM_QA1, FM_QA1, g0, g1, g2, a0, a1, a2, b0, b1, b2, s0, s1, s2);
      quad_node(c.rho, FM_QW1, FM_QA1, FM_QB1, FM_QA1, g0, g1, g2, a0, a1, a2, b0, b1, b2, s0, s1, s2);
      quad_node(c.rho, FM_QW1, FM_QA1, FM_QA1, FM_QB1, g0, g1, g2, a0, a1, a2, b0, b1, b2, s0, s1, s2);
      quad_node(c.rho, FM_QW2, FM_QB2, FM_QA2, FM_QA2, g0, g1, g2, a0, a1, a2, b0, b1, b2, s0, s1, s2);
      quad_node(c.rho, FM_QW2, FM_QA2, FM_QB2, FM_QA2, g0, g1, g2, a0, a1, a2, b0, b1, b2, s0, s1, s2);
      quad_node(c.rho, FM_QW2, FM_QA2, FM_QA2, FM_QB2, g0, g1, g2, a0, a1, a2, b0, b1, b2, s0, s1, s2);
      M0 += A * s0; M1 += A * s1; M2 += A * s2;
      a0 = b0; a1 = b1; a2 = b2;
      t = tn;
      d++;
    } while (t != t0 && d < 4096);
    double nrm = sqrt(M0 * M0 + M1 * M1 + M2 * M2);
    if (!(nrm > 0.0) || t != t0) { *c.err |= FM_E_TOPOLOGY; return 0.0; }
    double n0 = M0 / nrm, n1 = M1 / nrm, n2 = M2 / nrm;
    double dx = n0 - g0, dy = n1 - g1, dz = n2 - g2;
    c.x[i] = n0; c.y[i] = n1; c.z[i] = n2;
    return sqrt(dx * dx + dy * dy + dz * dz);
  }
  FM_HD static void run(const VdCtx& c, int j) {  // generic form (host tests)
    double disp = compute(c, c.cell_begin + j);
    FM_ATOMIC_MAX64(c.maxdisp, dbits(disp));
  }
};

}  // namespace fm
