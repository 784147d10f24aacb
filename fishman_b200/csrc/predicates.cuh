// predicates.cuh -- filtered-then-exact geometric predicates for the spherical
// Delaunay path, usable from device code and (for the host-side bootstrap and
// the CPU logic tests) from host code.
//
// The reference evaluates both predicates as plain fp64 3x3 determinants compared
// with eps = epsilon(1d0):
//   orient     /root/reference/src/delaunay_voronoi_mod.F90:1876-1892
//   in_circle  /root/reference/src/delaunay_voronoi_mod.F90:2032-2063
// and abstains (ORIENT_ON / ON_CIRCLE) inside the +-eps band.  Here the sign is
// always decided: stage 1 is the same fp64 determinant with a forward error
// bound, stage 2 is exact expansion arithmetic (Shewchuk-style two-sum /
// two-product; products through FMA), stage 3 breaks exact ties by Simulation of
// Simplicity on the ORIGINAL point ids, so the triangulation built on top is the
// unique Delaunay triangulation of a symbolically perturbed point set.  Outside
// the reference's eps band both give the same sign ("Predicates").
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define FM_HD __host__ __device__
#else
#define FM_HD
#endif

#if defined(__CUDA_ARCH__)
#define FM_ADD(a, b) __dadd_rn((a), (b))
#define FM_SUB(a, b) __dsub_rn((a), (b))
#define FM_MUL(a, b) __dmul_rn((a), (b))
#define FM_FMA(a, b, c) __fma_rn((a), (b), (c))
#else
// host translation units that include this header are built with -ffp-contract=off
#define FM_ADD(a, b) ((a) + (b))
#define FM_SUB(a, b) ((a) - (b))
#define FM_MUL(a, b) ((a) * (b))
#define FM_FMA(a, b, c) fma((a), (b), (c))
#endif

namespace fm {

struct P3 {
  double x, y, z;
};

// ---------------------------------------------------------------- expansions
// An expansion is a sum of non-overlapping doubles stored in increasing
// magnitude; its sign is the sign of its last component.

FM_HD inline void two_sum(double a, double b, double& x, double& y) {
  x = FM_ADD(a, b);
  double bv = FM_SUB(x, a);
  double av = FM_SUB(x, bv);
  double br = FM_SUB(b, bv);
  double ar = FM_SUB(a, av);
  y = FM_ADD(ar, br);
}

FM_HD inline void fast_two_sum(double a, double b, double& x, double& y) {
  x = FM_ADD(a, b);
  double bv = FM_SUB(x, a);
  y = FM_SUB(b, bv);
}

FM_HD inline void two_prod(double a, double b, double& x, double& y) {
  x = FM_MUL(a, b);
  y = FM_FMA(a, b, -x);
}

// h = e + b (in place allowed: h may alias e).  Returns the new length.
FM_HD inline int exp_grow(int elen, const double* e, double b, double* h) {
  double Q = b, Qn, hh;
  int k = 0;
  for (int i = 0; i < elen; i++) {
    two_sum(Q, e[i], Qn, hh);
    Q = Qn;
    if (hh != 0.0) h[k++] = hh;
  }
  if (Q != 0.0 || k == 0) h[k++] = Q;
  return k;
}

// h += f (h has room for hlen + flen)
FM_HD inline int exp_add(int hlen, double* h, int flen, const double* f) {
  for (int j = 0; j < flen; j++) hlen = exp_grow(hlen, h, f[j], h);
  return hlen;
}

// h = e * b, h must not alias e; room for 2*elen
FM_HD inline int exp_scale(int elen, const double* e, double b, double* h) {
  double Q, hh, p1, p0, sum;
  int k = 0;
  two_prod(e[0], b, Q, hh);
  if (hh != 0.0) h[k++] = hh;
  for (int i = 1; i < elen; i++) {
    two_prod(e[i], b, p1, p0);
    two_sum(Q, p0, sum, hh);
    if (hh != 0.0) h[k++] = hh;
    fast_two_sum(p1, sum, Q, hh);
    if (hh != 0.0) h[k++] = hh;
  }
  if (Q != 0.0 || k == 0) h[k++] = Q;
  return k;
}

FM_HD inline int exp_sign(int len, const double* e) {
  double v = e[len - 1];
  return (v > 0.0) - (v < 0.0);
}

// exact a*d - b*c  (<= 4 components)
FM_HD inline int exp_det2(double a, double b, double c, double d, double* h) {
  double p, q, t[2];
  two_prod(a, d, p, q);
  h[0] = q;
  h[1] = p;
  int n = (q == 0.0) ? (h[0] = p, 1) : 2;
  two_prod(b, c, p, q);
  t[0] = -q;
  t[1] = -p;
  return exp_add(n, h, 2, t);
}

// exact 3x3 determinant of doubles, row major (<= 24 components)
FM_HD inline int exp_det3(const double* m, double* h) {
  double d[4], s[8];
  int hl = 1;
  h[0] = 0.0;
  for (int j = 0; j < 3; j++) {
    int c1 = (j == 0) ? 1 : 0, c2 = (j == 2) ? 1 : 2;
    int dl = exp_det2(m[3 + c1], m[3 + c2], m[6 + c1], m[6 + c2], d);
    int sl = exp_scale(dl, d, (j == 1) ? -m[j] : m[j], s);
    hl = exp_add(hl, h, sl, s);
  }
  return hl;
}

// exact sign of a k x k determinant of doubles (k <= 4), row major with stride k
FM_HD inline int exact_det_sign(const double* m, int k) {
  if (k == 0) return 1;
  if (k == 1) return (m[0] > 0.0) - (m[0] < 0.0);
  if (k == 2) {
    double h[4];
    return exp_sign(exp_det2(m[0], m[1], m[2], m[3], h), h);
  }
  if (k == 3) {
    double h[24];
    return exp_sign(exp_det3(m, h), h);
  }
  // k == 4: cofactor expansion along the first row
  double acc[192], sub[9], d3[24], sc[48];
  int al = 1;
  acc[0] = 0.0;
  for (int j = 0; j < 4; j++) {
    if (m[j] == 0.0) continue;
    int cc = 0;
    for (int r = 1; r < 4; r++)
      for (int c = 0; c < 4; c++)
        if (c != j) sub[cc++] = m[4 * r + c];
    int dl = exp_det3(sub, d3);
    int sl = exp_scale(dl, d3, (j & 1) ? -m[j] : m[j], sc);
    al = exp_add(al, acc, sl, sc);
  }
  return exp_sign(al, acc);
}


// ------------------------------------------------------ SoS tie-break tables
// Generated by tools/gen_sos_tables.py (perturbation scheme documented there).
struct SosTerm {
  signed char sign, n;
  unsigned char rows, cols;
};
#define FM_SOS_TABLE(name, len) \
  FM_HD inline SosTerm name(int i) { const SosTerm t[len] = {
#define FM_SOS_TERM(s, n, r, c) {s, n, r, c},
#define FM_SOS_END \
  };               \
  return t[i];     \
  }
#include "sos_tables.inc"
#undef FM_SOS_TABLE
#undef FM_SOS_TERM
#undef FM_SOS_END
static const int kSosDet3Terms = 18, kSosDet4Terms = 18;

// sign(det(M + E)) for eps -> 0+; M (dim x dim, row major) has its rows sorted by
// point id.  dim = 3: rows (x y z); dim = 4: rows (x y z 1).  Never returns 0.
FM_HD inline int sos_sign(const double* M, int dim) {
  int nterms = (dim == 3) ? kSosDet3Terms : kSosDet4Terms;
  for (int i = 0; i < nterms; i++) {
    SosTerm t = (dim == 3) ? sos_det3(i) : sos_det4(i);
    double sub[16];
    int k = 0;
    for (int r = 0; r < dim; r++) {
      if (!((t.rows >> r) & 1)) continue;
      for (int c = 0; c < dim; c++)
        if ((t.cols >> c) & 1) sub[k++] = M[dim * r + c];
    }
    int s = exact_det_sign(sub, t.n);
    if (s != 0) return s * t.sign;
  }
  return 1;  // unreachable: the last term of each table is a constant
}

// Sorts up to 4 (id, point) rows by id; returns the parity (+1/-1) of the sort.
FM_HD inline int sort_rows(int n, int* id, P3* p) {
  int par = 1;
  for (int i = 1; i < n; i++)
    for (int j = i; j > 0 && id[j - 1] > id[j]; j--) {
      int ti = id[j]; id[j] = id[j - 1]; id[j - 1] = ti;
      P3 tp = p[j]; p[j] = p[j - 1]; p[j - 1] = tp;
      par = -par;
    }
  return par;
}

// ------------------------------------------------------------------- orient
// orient(a,b,c) = sign of a . (b x c): +1 when c lies to the LEFT of the great
// circle a -> b seen from outside the sphere (ORIENT_LEFT of DV:1882-1885).

FM_HD inline double orient_det(const P3& a, const P3& b, const P3& c, double& perm) {
  double m1 = b.y * c.z, m2 = b.z * c.y;
  double m3 = b.z * c.x, m4 = b.x * c.z;
  double m5 = b.x * c.y, m6 = b.y * c.x;
  double det = a.x * (m1 - m2) + a.y * (m3 - m4) + a.z * (m5 - m6);
  perm = fabs(a.x) * (fabs(m1) + fabs(m2)) + fabs(a.y) * (fabs(m3) + fabs(m4)) +
         fabs(a.z) * (fabs(m5) + fabs(m6));
  return det;
}

// exact + SoS stage; kept out of line so the filter path stays small
#if defined(__CUDACC__)
__noinline__
#endif
FM_HD inline int orient_slow(P3 a, P3 b, P3 c, int ia, int ib, int ic) {
  int id[3] = {ia, ib, ic};
  P3 p[3] = {a, b, c};
  int par = sort_rows(3, id, p);
  double M[9] = {p[0].x, p[0].y, p[0].z, p[1].x, p[1].y, p[1].z, p[2].x, p[2].y, p[2].z};
  int s = exact_det_sign(M, 3);
  if (s == 0) s = sos_sign(M, 3);
  return s * par;
}

// Forward error of orient_det: each of the three terms carries <= 3 roundings,
// the two additions <= 2 more; 1e-15 > 8 * 2^-53 covers it with margin.
#define FM_ORIENT_ERR 1.0e-15
// Shewchuk's o3derrboundA is (7 + 56 u) u with u = 2^-53 for the in-circle form
// below; 1e-15 ~ 9 u keeps the same margin.
#define FM_INCIRCLE_ERR 1.0e-15

// stage 1 alone: +1 / -1 when the filter certifies the sign, 0 when it cannot (the caller
// hands the item to a kernel that runs the full predicate)
FM_HD inline int orient_fast(const P3& a, const P3& b, const P3& c) {
  double perm, det = orient_det(a, b, c, perm);
  double bound = FM_ORIENT_ERR * perm;
  return det > bound ? 1 : (-det > bound ? -1 : 0);
}

FM_HD inline int orient(const P3& a, const P3& b, const P3& c, int ia, int ib, int ic) {
  double perm, det = orient_det(a, b, c, perm);
  double bound = FM_ORIENT_ERR * perm;
  if (det > bound) return 1;
  if (-det > bound) return -1;
  return orient_slow(a, b, c, ia, ib, ic);
}

// ----------------------------------------------------------------- incircle
// incircle(a,b,c,d) = +1 when d lies strictly outside the plane of the CCW
// triangle (a,b,c), i.e. inside its circumcircle on the sphere (INSIDE_CIRCLE of
// DV:2051-2056: det[c-d; b-d; a-d] > 0).

FM_HD inline double incircle_det(const P3& a, const P3& b, const P3& c, const P3& d,
                                 double& perm) {
  double adx = a.x - d.x, ady = a.y - d.y, adz = a.z - d.z;
  double bdx = b.x - d.x, bdy = b.y - d.y, bdz = b.z - d.z;
  double cdx = c.x - d.x, cdy = c.y - d.y, cdz = c.z - d.z;
  double bdxcdy = bdx * cdy, cdxbdy = cdx * bdy;
  double cdxady = cdx * ady, adxcdy = adx * cdy;
  double adxbdy = adx * bdy, bdxady = bdx * ady;
  double det = adz * (bdxcdy - cdxbdy) + bdz * (cdxady - adxcdy) + cdz * (adxbdy - bdxady);
  perm = (fabs(bdxcdy) + fabs(cdxbdy)) * fabs(adz) + (fabs(cdxady) + fabs(adxcdy)) * fabs(bdz) +
         (fabs(adxbdy) + fabs(bdxady)) * fabs(cdz);
  return -det;  // det[a-d; b-d; c-d] < 0  <=>  d above the plane of CCW (a,b,c)
}

#if defined(__CUDACC__)
__noinline__
#endif
FM_HD inline int incircle_slow(P3 a, P3 b, P3 c, P3 d, int ia, int ib, int ic, int idd) {
  int id[4] = {ia, ib, ic, idd};
  P3 p[4] = {a, b, c, d};
  int par = sort_rows(4, id, p);
  double M[16];
  for (int r = 0; r < 4; r++) {
    M[4 * r] = p[r].x; M[4 * r + 1] = p[r].y; M[4 * r + 2] = p[r].z; M[4 * r + 3] = 1.0;
  }
  int s = exact_det_sign(M, 4);
  if (s == 0) s = sos_sign(M, 4);
  return -s * par;  // det4[(a 1);(b 1);(c 1);(d 1)] = det3[a-d; b-d; c-d]
}

FM_HD inline int incircle_fast(const P3& a, const P3& b, const P3& c, const P3& d) {
  double perm, det = incircle_det(a, b, c, d, perm);
  double bound = FM_INCIRCLE_ERR * perm;
  return det > bound ? 1 : (-det > bound ? -1 : 0);
}

FM_HD inline int incircle(const P3& a, const P3& b, const P3& c, const P3& d, int ia, int ib,
                          int ic, int idd) {
  double perm, det = incircle_det(a, b, c, d, perm);
  double bound = FM_INCIRCLE_ERR * perm;
  if (det > bound) return 1;
  if (-det > bound) return -1;
  return incircle_slow(a, b, c, d, ia, ib, ic, idd);
}

}  // namespace fm
