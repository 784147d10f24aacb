// vd_kernels.cuh -- per-thread bodies of the Voronoi / Lloyd kernels.
//
//   KCircum     circumcentres            DV:379-391, calc_circumcircle DV:1323-1348
//   KDegree / KRingFill   CSR vertex rings   extract_incident_DT_and_linked_DVT DV:704-743,
//                                             cell rings DV:394-409
//   KCentroid   density-weighted centroid + generator update + displacement
//               (no reference code: SURVEY.md section 7.3 is the specification,
//               restated on the CPU in oracle/fishman_oracle.c fo_centroids)
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

struct Density {   // the first seven members: same layout as fm_density in include/fishman_b200.h
  int kind;        // 0 uniform, 1 tanh refinement
  double cx, cy, cz, beta, alpha, gamma;
  double k2, c2;   // 2/alpha and -2 beta/alpha, set by the launcher: -2 (beta - d)/alpha = d k2 + c2
  // The density function staged as a table (north-star (c)): rho as a function of the CHORD
  // u = |x - c| in [0, 2] (analytic there, unlike rho(dot) at dot = 1), tab_n intervals with a
  // degree-7 polynomial in s = 2 frac - 1 each (8 doubles per interval, 64-byte aligned), built and
  // checked on the host by the launcher (fishman_b200.cu: density_table); null = evaluate directly.
  const double* tab;
  double tab_scale;   // tab_n / 2: intervals per unit of chord
  int tab_n;
};

struct VdCtx {
  int n, n_tri;
  P4* pts;                    // generators, internal order (x,y,z updated in place by the Lloyd kernel)
  const Row* row;             // triangle slots (dt_kernels.cuh)
  P4* cc;                     // circumcentres per triangle (w unused)
  double *cclon, *cclat;      // optional (may be null)
  int* ringc;                 // ring cache: 8 ints per vertex (see KCentroid), or null
  const int* v2t;             // per vertex: some incident triangle
  int* deg;                   // per ORIGINAL id: ring length, then its exclusive scan
  int *ring_tri, *ring_nbr;   // CSR payload, 1-based output ids
  Density rho;
  int cell_begin, cell_end;   // range of internal cells this rank updates
  u64* maxdisp;               // bits of the largest chord displacement (>= 0, so ordered)
  const int* abort_flag;      // when non-null and != 0 the Lloyd kernel does nothing
  const int* counters;        // CUDA library: the handle's counter block (halt word of a batch), or null
  int dry;                    // != 0: compute the displacements only, move nothing (fm_quality)
  int* err;
  // peer exchange (fm_peer_attach): the generator arrays of the other ranks, mapped into this
  // process over NVLink; the Lloyd kernel stores every updated generator there as well, so the
  // exchange of the blocks overlaps the quadrature instead of following it as a collective
  P4* peer_pts[7];
  int n_peers;
};

static const double kEps = 2.220446049250313e-16;  // eps of src/const_mod.F90:16
static const double kPi2 = 6.283185307179586476925287;

// ------------------------------------------------------------ circumcentres
struct KCircum {
  FM_HD static void run(const VdCtx& c, int t) {
    I4 v = c.row[t].v;
    const P4 &qa = c.pts[v.x], &qb = c.pts[v.y], &qc = c.pts[v.z];
    bool dg;
    P3 n = circumcentre(P3{qa.x, qa.y, qa.z}, P3{qb.x, qb.y, qb.z}, P3{qc.x, qc.y, qc.z}, dg);
    if (dg) *c.err |= FM_E_TOPOLOGY;
    c.cc[t] = P4{n.x, n.y, n.z, 0};
    if (c.cclon) {  // inverse_cartesian_transform, src/sphere_geometry_mod.F90:112-121
      double lon = atan2(n.y, n.x);
      if (lon < 0.0) lon += kPi2;
      c.cclon[t] = lon;
      c.cclat[t] = asin(n.z);
    }
  }
};

// Ring walk around internal vertex i, CCW: from triangle t with i at index k the
// next triangle lies across the edge opposite vertex k+1, and i sits at index
// e+1 of that neighbour (e = the neighbour's edge index stored in adj).
FM_HD inline int ring_start(const VdCtx& c, int i, int& k) {
  int t = c.v2t[i];
  I4 v = c.row[t].v;
  k = (v.x == i) ? 0 : ((v.y == i) ? 1 : 2);
  return t;
}
FM_HD inline int ring_next(const VdCtx& c, int t, int& k) {
  int ne = geti(c.row[t].a, (k + 1) % 3);
  k = ((ne & 3) + 1) % 3;
  return ne >> 2;
}

struct KDegree {
  FM_HD static void run(const VdCtx& c, int i) {
    int k, t0 = ring_start(c, i, k), t = t0, d = 0;
    do { t = ring_next(c, t, k); d++; } while (t != t0 && d < 4096);
    if (t != t0) *c.err |= FM_E_TOPOLOGY;
    c.deg[(int)c.pts[i].id] = d;
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
      int nb = (int)c.pts[geti(c.row[t].v, (k + 2) % 3)].id;
      if (nb < best) { best = nb; jbest = d; }
      t = ring_next(c, t, k);
      d++;
    } while (t != t0 && d < 4096);
    int off = c.deg[(int)c.pts[i].id];
    t = ring_start(c, i, k);
    for (int j = 0; j < d; j++) {
      int pos = off + (j - jbest + d) % d;
      c.ring_tri[pos] = t + 1;
      c.ring_nbr[pos] = (int)c.pts[geti(c.row[t].v, (k + 2) % 3)].id + 1;
      t = ring_next(c, t, k);
    }
  }
};

// ------------------------------------------------------------------- Lloyd
// Specification: SURVEY.md 7.3 (the reference has no Lloyd step).  For cell i with generator
// g and CCW ring of circumcentres c_1..c_d: sub-triangles (g, c_j, c_j+1), planar area A_j,
// degree-5 seven-point rule with nodes q_k projected to the sphere, moment
// M = sum_j A_j sum_k w_k rho(q^_k) q^_k, new generator M/|M|.
FM_HD inline double density_eval(const Density& r, double x, double y, double z) {
  if (r.kind == 0) return 1.0;
  double dot = x * r.cx + y * r.cy + z * r.cz;
  double d = acos(fmax(fmin(dot, 1.0), -1.0));  // calc_arc_length, sphere_geometry_mod.F90:289-296
  return 0.5 * (1.0 - r.gamma) * (tanh((r.beta - d) / r.alpha) + 1.0) + r.gamma;
}

#if defined(__CUDACC__)
// ---- the same density on the device, 42 evaluations per cell.  The library acos / exp / IEEE
// divisions came to ~195 issued instructions per evaluation, a third of them moves of 64-bit
// immediates (two UMOV per polynomial coefficient) and NaN-exact fmin/fmax sequences.  Here: the
// coefficients sit in constant memory (an operand of the DFMA itself), no range branches, no
// library calls.
//   rho = (1 - gamma) / (1 + exp(-2 (beta - d)/alpha)) + gamma     ((tanh u + 1)/2 = 1/(1 + e^-2u))
// fm_acos: |x| <= 1/2: pi/2 - asin x;  |x| > 1/2: 2 asin sqrt((1 - |x|)/2), pi - that for x < 0;
//   asin s = s + s z P(z), z = s^2, P = degree-12 fit of (asin(sqrt z)/sqrt z - 1)/z on [0, 1/4]
//   (Chebyshev nodes, 50-digit arithmetic, |z P - exact| < 6e-18).  |x| a rounding above 1 gives
//   z < 0 -> s = 0 -> 0 or pi: no clamp needed.  Measured on the device against libm over
//   [-1, 1] incl. 1 - 1e-16: absolute error <= 4.5e-16, relative <= 3.6e-16.
// fm_exp: t in [-700, 700]: n = rint(t log2 e), r = t - n ln2 (two FMAs, ln2 split hi/lo),
//   e^r = 1 + r + r^2 Q(r), Q = degree-10 fit on |r| <= ln2/2 (relative error 4e-19), 2^n by an
//   integer add into the exponent field (the result is normal on that range).
static __constant__ double kAsinP[13] = {
    0.16666666666666669, 0.07499999999998433, 0.04464285714635543, 0.030381944138531247,
    0.02237217294214989, 0.017352392720869973, 0.013971212973552933, 0.011479177415184906,
    0.01032281435018578, 0.005457506718640358, 0.01740087944269402, -0.014851887071247204,
    0.028757851367421566};
static __constant__ double kExpQ[11] = {
    0.5, 0.1666666666666667, 0.04166666666666667, 0.008333333333326136, 0.0013888888888883748,
    0.00019841269874820627, 2.4801587325547743e-05, 2.7557255400206422e-06, 2.75572736431103e-07,
    2.5105217004720745e-08, 2.0914686968086876e-09};

// 1/v for v > 0: hardware seed (MUFU.RCP64H, ~2^-20) + one cubic step y (1 + e + e^2), e = 1 - v y
__device__ __forceinline__ double fm_rcp(double v) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(v));
  double e = fma(-v, y, 1.0);
  return fma(y, fma(e, e, e), y);
}
__device__ __forceinline__ double fm_acos(double x) {
  const double a = fabs(x);
  const bool lo = a <= 0.5;
  const double z = lo ? x * x : 0.5 * (1.0 - a);   // 1 - a is exact for a >= 1/2
  double r;                                         // 1/sqrt(z): seed + one cubic step, as fm_rsqrt
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(z));
  const double e = fma(-(z * r), r, 1.0);
  r = fma(r, fma(0.375, e, 0.5) * e, r);
  const double s = lo ? a : (z > 0.0 ? z * r : 0.0);
  double p = kAsinP[12];
#pragma unroll
  for (int k = 11; k >= 0; k--) p = fma(p, z, kAsinP[k]);
  const double t = fma(s, z * p, s);   // asin s
  const double pio2_hi = 1.5707963267948966, pio2_lo = 6.123233995736766e-17;
  const double u = (lo && x < 0.0) ? -t : t;                  // lo: asin x;  otherwise asin s
  const double v = pio2_hi - (u - pio2_lo);                   // lo: acos x;  x < -1/2: (acos x)/2
  return lo ? v : (x > 0.0 ? 2.0 * t : 2.0 * v);
}
__device__ __forceinline__ double fm_exp(double t) {  // t in [-700, 700]
  const double magic = 6755399441055744.0;            // 1.5 * 2^52: the low word of t' + magic is rint(t')
  const double nm = fma(t, 1.4426950408889634, magic);
  const double n = nm - magic;
  double r = fma(n, -0.6931471805599453, t);
  r = fma(n, -2.3190468138462996e-17, r);
  double q = kExpQ[10];
#pragma unroll
  for (int k = 9; k >= 0; k--) q = fma(q, r, kExpQ[k]);
  const double p = fma(r * r, q, r) + 1.0;
  return __hiloint2double(__double2hiint(p) + (__double2loint(nm) << 20), __double2loint(p));
}
#endif
// density at the projected node q/|q|, given q and inv = 1/|q|
#if defined(__CUDACC__)
// the rare interval the table could not fit (marked by a NaN): out of line, so that the table kernel
// does not carry the registers of acos / exp through its quadrature
__device__ __noinline__ double density_node_slow(const Density& r, double dot) {
  const double d = fm_acos(dot);
  double t = fma(d, r.k2, r.c2);
  t = t > 700.0 ? 700.0 : t;
  t = t < -700.0 ? -700.0 : t;
  return fma(1.0 - r.gamma, fm_rcp(1.0 + fm_exp(t)), r.gamma);
}
#endif
// density at the unit vector whose product with the centre is `dot`
template <bool TABLE>
FM_HD inline double density_dot(const Density& r, double dot) {
#if defined(__CUDA_ARCH__)
  if (TABLE) {
    // ~20 fp64 instructions instead of ~89 (acos + exp + reciprocal): chord, interval, Horner
    double w = fma(-2.0, dot, 2.0);                 // |x - c|^2 = 2 - 2 x.c
    w = w > 0.0 ? w : 0.0;
    const double u = w > 0.0 ? w * fm_rsqrt(w) : 0.0;
    const double f = u * r.tab_scale;
    int k = (int)f;
    k = k < r.tab_n ? k : r.tab_n - 1;
    const double sx = fma(2.0, f - (double)k, -1.0);
    struct alignas(32) D4 { double x, y, z, w; };   // two 256-bit loads per evaluation
    const D4* c4 = reinterpret_cast<const D4*>(r.tab) + 2 * (size_t)k;
    const D4 lo = c4[0], hi = c4[1];
    if (hi.w != hi.w) return density_node_slow(r, dot);
    double p = hi.w;   // Horner (Estrin's three-level tree measured the same: 0.585 against 0.578 ms)
    p = fma(p, sx, hi.z); p = fma(p, sx, hi.y); p = fma(p, sx, hi.x);
    p = fma(p, sx, lo.w); p = fma(p, sx, lo.z); p = fma(p, sx, lo.y); p = fma(p, sx, lo.x);
    p = p < r.gamma ? r.gamma : p;   // the interpolant may leave [gamma, 1] by a rounding
    return p > 1.0 ? 1.0 : p;
  }
  const double d = fm_acos(dot);
  double t = fma(d, r.k2, r.c2);   // -2 (beta - d)/alpha
  t = t > 700.0 ? 700.0 : t;
  t = t < -700.0 ? -700.0 : t;
  return fma(1.0 - r.gamma, fm_rcp(1.0 + fm_exp(t)), r.gamma);
#else
  double dd = acos(fmax(fmin(dot, 1.0), -1.0));
  return 0.5 * (1.0 - r.gamma) * (tanh((r.beta - dd) / r.alpha) + 1.0) + r.gamma;
#endif
}
// density at the projected node q/|q|, given q and inv = 1/|q|
template <bool TABLE>
FM_HD inline double density_node(const Density& r, double qx, double qy, double qz, double inv) {
#if defined(__CUDA_ARCH__)
  return density_dot<TABLE>(r, fma(qz, r.cz, fma(qy, r.cy, qx * r.cx)) * inv);
#else
  return density_eval(r, qx * inv, qy * inv, qz * inv);
#endif
}

// degree-5 seven-point rule on a triangle (barycentric weights of g, c_j, c_j+1)
#define FM_QA1 0.10128650732345633880
#define FM_QB1 0.79742698535308732240
#define FM_QA2 0.47014206410511508977
#define FM_QB2 0.05971587178976982045
#define FM_QW0 0.225
#define FM_QW1 0.12593918054482715260
#define FM_QW2 0.13239415278850618074

// 1/sqrt for the weights of the uniform quadrature: the hardware seed + ONE Newton step, 4 fp64
// instructions instead of fm_rsqrt's 5 (8 of the 117 per sub-triangle; centroid kernel 0.168 ->
// 0.162 ms).  Relative error 1.5 e^2 ~ 1e-13 with the seed's e ~ 2^-22: a weight error of that size
// moves a centroid by at most that fraction of the CELL DIAMETER, i.e. < 2e-13 on any mesh of the
// unit sphere and ~1e-16 at 2.6M generators -- inside the 1e-12 of the specification with room to
// spare (parity tests unchanged: 1e-12 against the oracle, 1e-13 against the 50-digit fixtures).
// -DFM_QUAD_CUBIC restores the cubic step.
FM_HD inline double fm_rsqrt_q(double v) {
#if defined(__CUDA_ARCH__) && !defined(FM_QUAD_CUBIC)
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(v));
  double e = __fma_rn(-__dmul_rn(v, y), y, 1.0);
  return __fma_rn(__dmul_rn(y, 0.5), e, y);
#else
  return fm_rsqrt(v);
#endif
}

// Quadrature of one sub-triangle (g, a, b) added to the moment M.  The seven nodes are
// S/3 and (A S + (B - A) v) for v in {g, a, b}, (A, B) in {(A1, B1), (A2, B2)}, S = g + a + b.
// Only q/|q| enters, so each node is taken as q = S + kappa v (kappa = (B - A)/A) and
//   |q|^2 = S.S + 2 kappa S.v + kappa^2 v.v,      sum_k w_k q_k/|q_k| = cS S + cg g + ca a + cb b
// with scalar coefficients: the node vectors themselves are only formed when the density has
// to be evaluated at them.
// General form (a density, or a ring too long for the staged one).  g, a and b are unit vectors to
// a few 1e-16, so g.g = a.a = b.b = 1 inside |q|^2 (see quad_uniform); and the product of a node
// with the density's centre c is  q.c = S.c + kappa v.c  with S.c = g.c + a.c + b.c: the node
// vectors are never formed, three products with c per ring vertex replace one per node.  gc, ac,
// bc: products of g, a, b with the centre (unused when UNIFORM).
template <bool UNIFORM, bool TABLE>
FM_HD inline void quad_subtriangle(const Density& r, const P3& g, const P3& a, const P3& b, double gc,
                                   double ac, double bc, double& M0, double& M1, double& M2) {
  const double k1 = (FM_QB1 - FM_QA1) / FM_QA1, k2 = (FM_QB2 - FM_QA2) / FM_QA2;
  double u0 = a.x - g.x, u1 = a.y - g.y, u2 = a.z - g.z;
  double v0 = b.x - g.x, v1 = b.y - g.y, v2 = b.z - g.z;
  double c0 = u1 * v2 - u2 * v1, c1 = u2 * v0 - u0 * v2, c2 = u0 * v1 - u1 * v0;
  double n2 = c0 * c0 + c1 * c1 + c2 * c2;
  double A = n2 > 0.0 ? 0.5 * n2 * fm_rsqrt(n2) : 0.0;  // planar area; coincident circumcentres give 0
  double ga = g.x * a.x + g.y * a.y + g.z * a.z;
  double gb = g.x * b.x + g.y * b.y + g.z * b.z;
  double ab = a.x * b.x + a.y * b.y + a.z * b.z;
  double Sg = 1.0 + ga + gb, Sa = ga + 1.0 + ab, Sb = gb + ab + 1.0;
  double SS = Sg + Sa + Sb;
  const double S1 = SS + k1 * k1, S2 = SS + k2 * k2;   // |S + kappa v|^2 = S.S + kappa^2 + 2 kappa S.v
  double i0 = fm_rsqrt(SS);
  double i1g = fm_rsqrt(fma(2.0 * k1, Sg, S1));
  double i1a = fm_rsqrt(fma(2.0 * k1, Sa, S1));
  double i1b = fm_rsqrt(fma(2.0 * k1, Sb, S1));
  double i2g = fm_rsqrt(fma(2.0 * k2, Sg, S2));
  double i2a = fm_rsqrt(fma(2.0 * k2, Sa, S2));
  double i2b = fm_rsqrt(fma(2.0 * k2, Sb, S2));
  double w0 = FM_QW0 * i0;
  double w1g = FM_QW1 * i1g, w1a = FM_QW1 * i1a, w1b = FM_QW1 * i1b;
  double w2g = FM_QW2 * i2g, w2a = FM_QW2 * i2a, w2b = FM_QW2 * i2b;
  if (!UNIFORM) {  // density at the projected nodes q^ = (S + kappa v) / |q|
    const double Sc = (gc + ac) + bc;
    w0 *= density_dot<TABLE>(r, Sc * i0);
    w1g *= density_dot<TABLE>(r, fma(k1, gc, Sc) * i1g);
    w1a *= density_dot<TABLE>(r, fma(k1, ac, Sc) * i1a);
    w1b *= density_dot<TABLE>(r, fma(k1, bc, Sc) * i1b);
    w2g *= density_dot<TABLE>(r, fma(k2, gc, Sc) * i2g);
    w2a *= density_dot<TABLE>(r, fma(k2, ac, Sc) * i2a);
    w2b *= density_dot<TABLE>(r, fma(k2, bc, Sc) * i2b);
  }
  double cS = w0 + ((w1g + w1a) + w1b) + ((w2g + w2a) + w2b);
  double tg = A * (cS + fma(k1, w1g, k2 * w2g));
  double ta = A * (cS + fma(k1, w1a, k2 * w2a));
  double tb = A * (cS + fma(k1, w1b, k2 * w2b));
  M0 = fma(tb, b.x, fma(ta, a.x, fma(tg, g.x, M0)));
  M1 = fma(tb, b.y, fma(ta, a.y, fma(tg, g.y, M1)));
  M2 = fma(tb, b.z, fma(ta, a.z, fma(tg, g.z, M2)));
}

// Uniform density: the same rule regrouped around the ring.  Per ring vertex c the product
// g.c and the edge vector c - g are formed once and shared by the two sub-triangles that
// meet there; S never appears: S.v = g.v + a.v + b.v, and
//   A (cS S + cg g + ca a + cb b) = A (cS + cg) g + A (cS + ca) a + A (cS + cb) b,
// so the g part is a scalar accumulated in Mg and added once per cell.  About 12 % fewer fp64
// instructions than quad_subtriangle<true>; the same value up to rounding.
struct RingVertex {
  P3 p, u;    // circumcentre, circumcentre - g
  double gp;  // g.p
};
FM_HD inline RingVertex ring_vertex(const P3& g, const P4& q) {
  RingVertex r;
  r.p = P3{q.x, q.y, q.z};
  r.u = P3{q.x - g.x, q.y - g.y, q.z - g.z};
  r.gp = g.x * q.x + g.y * q.y + g.z * q.z;
  return r;
}
// g, a and b are unit vectors to a few 1e-16 (generators are projected, circumcentres normalised),
// so g.g = a.a = b.b = 1 here: inside |q|^2 ~ 100 the difference is a relative 1e-18 of a weight.
// That takes the three squared norms out of the loop (7 fp64 instructions per sub-triangle).
FM_HD inline void quad_uniform(const RingVertex& a, const RingVertex& b, double& Mg, double& M0,
                               double& M1, double& M2) {
  const double k1 = (FM_QB1 - FM_QA1) / FM_QA1, k2 = (FM_QB2 - FM_QA2) / FM_QA2;
  double c0 = a.u.y * b.u.z - a.u.z * b.u.y, c1 = a.u.z * b.u.x - a.u.x * b.u.z, c2 = a.u.x * b.u.y - a.u.y * b.u.x;
  double n2 = c0 * c0 + c1 * c1 + c2 * c2;
  double A = n2 > 0.0 ? 0.5 * n2 * fm_rsqrt_q(n2) : 0.0;
  double ab = a.p.x * b.p.x + a.p.y * b.p.y + a.p.z * b.p.z;
  double Sg = 1.0 + a.gp + b.gp, Sa = a.gp + 1.0 + ab, Sb = b.gp + ab + 1.0;
  double SS = Sg + Sa + Sb;
  // weights folded into constants: cS = sum_k w_k / |q_k|, and per vertex v the coefficient
  // A (cS + kappa1 w1 / |q_1v| + kappa2 w2 / |q_2v|)
  const double K1 = k1 * FM_QW1, K2 = k2 * FM_QW2;
  const double S1 = SS + k1 * k1, S2 = SS + k2 * k2;   // |S + kappa v|^2 = S.S + kappa^2 + 2 kappa S.v
  double i0 = fm_rsqrt_q(SS);
  double i1g = fm_rsqrt_q(fma(2.0 * k1, Sg, S1));
  double i1a = fm_rsqrt_q(fma(2.0 * k1, Sa, S1));
  double i1b = fm_rsqrt_q(fma(2.0 * k1, Sb, S1));
  double i2g = fm_rsqrt_q(fma(2.0 * k2, Sg, S2));
  double i2a = fm_rsqrt_q(fma(2.0 * k2, Sa, S2));
  double i2b = fm_rsqrt_q(fma(2.0 * k2, Sb, S2));
  double cS = fma(FM_QW2, (i2g + i2a) + i2b, fma(FM_QW1, (i1g + i1a) + i1b, FM_QW0 * i0));
  double tg = A * fma(K1, i1g, fma(K2, i2g, cS));
  double ta = A * fma(K1, i1a, fma(K2, i2a, cS));
  double tb = A * fma(K1, i1b, fma(K2, i2b, cS));
  Mg += tg;
  M0 = fma(tb, b.p.x, fma(ta, a.p.x, M0));
  M1 = fma(tb, b.p.y, fma(ta, a.p.y, M1));
  M2 = fma(tb, b.p.z, fma(ta, a.p.z, M2));
}

static const int kRingCap = 12;  // ring slots staged per cell; longer rings take the re-walk path

struct KCentroid {
  // Returns the chord displacement of cell i and moves its generator.  ring[j * stride] is
  // scratch for kRingCap triangle ids (shared memory on the device).  The accumulation starts
  // at the LOWEST triangle id of the ring, whatever triangle v2t points at, so the result does
  // not depend on which incident triangle happens to be remembered.
  template <bool UNIFORM, bool TABLE = false>
  FM_HD static double compute(const VdCtx& c, int i, int* ring, int stride) {
    P4 gq = c.pts[i];
    P3 g{gq.x, gq.y, gq.z};
    int k = 0, t0 = 0, t = 0, d = 0, jmin = 0, tmin = 0, kmin = 0;
    // Ring cache.  Walking the ring is a chain of d dependent row loads; between two Lloyd
    // iterations only the rings around flipped edges change (a few thousand of millions of
    // cells), so the ring of a cell -- rotated to start at its lowest triangle id, padded with
    // -1 -- is kept in one 32-byte record.  A flip invalidates the records of its four vertices
    // (KFlip), a rebuild all of them; an invalid or long (> 8) ring is walked, and re-cached.
    bool cached = false;
    if (c.ringc) {
      const I4* rec = (const I4*)(c.ringc + 8 * (size_t)i);
      I4 r0 = rec[0], r1 = rec[1];
      if (r0.x >= 0) {
        cached = true;
        ring[0] = r0.x; ring[stride] = r0.y; ring[2 * stride] = r0.z; ring[3 * stride] = r0.w;
        ring[4 * stride] = r1.x; ring[5 * stride] = r1.y; ring[6 * stride] = r1.z; ring[7 * stride] = r1.w;
        d = 3 + (r0.w >= 0) + (r1.x >= 0) + (r1.y >= 0) + (r1.z >= 0) + (r1.w >= 0);
        tmin = r0.x; jmin = 0;
      }
    }
    if (!cached) {
      t0 = ring_start(c, i, k); t = t0; tmin = t0; kmin = k;
      do {
        if (d < kRingCap) ring[d * stride] = t;
        if (t < tmin) { tmin = t; jmin = d; kmin = k; }
        t = ring_next(c, t, k);
        d++;
      } while (t != t0 && d < 4096);
      if (t != t0) { *c.err |= FM_E_TOPOLOGY; return 0.0; }
      if (c.ringc && d <= 8) {
        int q[8];
        for (int s = 0; s < 8; s++) {
          int js = jmin + s;
          if (js >= d) js -= d;
          q[s] = s < d ? ring[js * stride] : -1;
        }
        I4* rec = (I4*)(c.ringc + 8 * (size_t)i);
        rec[0] = I4{q[0], q[1], q[2], q[3]};
        rec[1] = I4{q[4], q[5], q[6], q[7]};
      }
    }
    // (Tried: prefetch.global.L1 / .L2 of ALL ring circumcentres right here, because a fifth of the
    // stall samples wait for the loop's one-trip-ahead loads.  Slower: 0.180 against 0.173 ms -- the
    // extra address arithmetic and cache pressure cost more than the latency hidden;
    // profiles/r02_i_ab_centroid_prefetch.txt.)
    double M0 = 0.0, M1 = 0.0, M2 = 0.0;
    if (UNIFORM && d <= kRingCap) {
      int j = jmin;
      double Mg = 0.0;
      RingVertex va = ring_vertex(g, c.cc[tmin]);
      j = (j + 1 == d) ? 0 : j + 1;
      P4 qb = c.cc[ring[j * stride]];
      for (int s = 0; s < d; s++) {
        j = (j + 1 == d) ? 0 : j + 1;
        P4 qn = c.cc[ring[j * stride]];  // prefetch: consumed in the next iteration
        RingVertex vb = ring_vertex(g, qb);
        quad_uniform(va, vb, Mg, M0, M1, M2);
        va = vb;
        qb = qn;
      }
      M0 = fma(Mg, g.x, M0); M1 = fma(Mg, g.y, M1); M2 = fma(Mg, g.z, M2);
    } else if (d <= kRingCap) {
      int j = jmin;
      P4 qa = c.cc[tmin];
      j = (j + 1 == d) ? 0 : j + 1;
      P4 qb = c.cc[ring[j * stride]];
      const double gc = g.x * c.rho.cx + g.y * c.rho.cy + g.z * c.rho.cz;
      double ac = qa.x * c.rho.cx + qa.y * c.rho.cy + qa.z * c.rho.cz;
      for (int s = 0; s < d; s++) {
        j = (j + 1 == d) ? 0 : j + 1;
        P4 qn = c.cc[ring[j * stride]];  // prefetch: consumed in the next iteration
        double bc = qb.x * c.rho.cx + qb.y * c.rho.cy + qb.z * c.rho.cz;
        quad_subtriangle<UNIFORM, TABLE>(c.rho, g, P3{qa.x, qa.y, qa.z}, P3{qb.x, qb.y, qb.z}, gc, ac, bc, M0, M1, M2);
        qa = qb;
        qb = qn;
        ac = bc;
      }
    } else {  // long ring: walk again from the lowest triangle
      t = tmin; k = kmin;
      P4 qa = c.cc[t];
      const double gc = g.x * c.rho.cx + g.y * c.rho.cy + g.z * c.rho.cz;
      double ac = qa.x * c.rho.cx + qa.y * c.rho.cy + qa.z * c.rho.cz;
      for (int s = 0; s < d; s++) {
        t = ring_next(c, t, k);
        P4 qb = c.cc[t];
        double bc = qb.x * c.rho.cx + qb.y * c.rho.cy + qb.z * c.rho.cz;
        quad_subtriangle<UNIFORM, TABLE>(c.rho, g, P3{qa.x, qa.y, qa.z}, P3{qb.x, qb.y, qb.z}, gc, ac, bc, M0, M1, M2);
        qa = qb;
        ac = bc;
      }
    }
    const double mm = M0 * M0 + M1 * M1 + M2 * M2;
    if (!(mm > 0.0)) { *c.err |= FM_E_TOPOLOGY; return 0.0; }
#if defined(__CUDA_ARCH__) && !defined(FM_CENTROID_DIVIDE)
    const double rn = fm_rsqrt(mm);   // projection to the sphere: three products instead of sqrt + three divisions
    double n0 = M0 * rn, n1 = M1 * rn, n2 = M2 * rn;
#else
    const double nrm = sqrt(mm);
    double n0 = M0 / nrm, n1 = M1 / nrm, n2 = M2 / nrm;
#endif
    double dx = n0 - g.x, dy = n1 - g.y, dz = n2 - g.z;
    if (!c.dry) {
      const P4 moved{n0, n1, n2, gq.id};  // whole 32-byte records: one store, one full sector / NVLink packet
      c.pts[i] = moved;
      for (int q = 0; q < c.n_peers; q++) c.peer_pts[q][i] = moved;
    }
    const double d2 = dx * dx + dy * dy + dz * dz;
#if defined(__CUDA_ARCH__) && !defined(FM_CENTROID_DIVIDE)
    return d2 > 0.0 ? d2 * fm_rsqrt(d2) : 0.0;
#else
    return sqrt(d2);
#endif
  }
  FM_HD static void run(const VdCtx& c, int j) {  // generic form (host tests)
    int ring[kRingCap];
    double disp = (c.rho.kind == 0) ? compute<true>(c, c.cell_begin + j, ring, 1)
                                    : compute<false>(c, c.cell_begin + j, ring, 1);
    FM_ATOMIC_MAX64(c.maxdisp, dbits(disp));
  }
};

}  // namespace fm
