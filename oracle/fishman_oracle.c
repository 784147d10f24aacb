/*
 * fishman_oracle.c -- CPU restatement of dongli/fishman's spherical Delaunay /
 * Voronoi path, plus the Lloyd reference step this repo defines.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under fishman_b200/ may include, link or
 * call this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, as the checker.
 *
 * PARITY UNPINNED: the reference (Fortran, /root/reference/src) ships no test,
 * fixture or golden vector for this path and cannot be compiled here (no
 * Fortran compiler).  This file follows the reference source line by line
 * (citations "DV:n" = src/delaunay_voronoi_mod.F90:n, "SG:n" =
 * src/sphere_geometry_mod.F90:n) and is pinned instead against Qhull
 * (scipy.spatial.ConvexHull), analytic polyhedra and the reference's own
 * run-time invariants -- see tests/test_oracle.py.
 *
 * Deviations from the reference, all deliberate and documented in DESIGN.md:
 *  D1  the first three vertices are passed in (the reference draws them from a
 *      wall-clock-seeded RNG, DV:595-610, src/random_number_mod.F90:28).
 *  D2  virtual-vertex removal (DV:772-843) tests the ear (b[k-1],b[k],b[k+1])
 *      that the following flip22 actually clips.  As written the reference
 *      tests (b[k],b[k+1],b[k+2]) but flips the triangle pair that clips
 *      (b[k-1],b[k],b[k+1]) (ring pairing DV:723,738 vs idx_map DV:826), and
 *      keeps walking a list item it has just deallocated (DV:1136-1137 ->
 *      linked_list_mod.F90:552).  That is undefined behaviour; the consistent
 *      ear is what the three checks are evidently meant to guard.
 *  D3  rings are re-extracted from the adjacency after the virtual vertices are
 *      gone instead of being patched in place (DV:1129-1139); ring START
 *      positions are history dependent in the reference and are
 *      never compared.
 *  D4  the Lloyd step does not exist in the reference; it is the
 *      specification of SURVEY 7.3, restated in fo_lloyd_step below.
 *  D5  the MPAS adapter (fo_mpas, "MM:n" = src/mpas_mesh_mod.F90:n) sizes its
 *      per-cell rows by the largest degree found instead of the reference's
 *      fixed maxEdges = 6 (MM:19), zero-initialises weightsOnEdge
 *      (the reference accumulates into an unset allocate, MM:386/546), and
 *      fills the padded slots of the per-cell index rows with 0.
 *
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC (see oracle/Makefile).
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define FO_EPS 2.220446049250313e-16 /* eps = epsilon(1d0), src/const_mod.F90:16 */
#define FO_PI 3.14159265358979323846
#define FO_PI2 (2.0 * FO_PI)
#define FO_PI05 (0.5 * FO_PI)

enum { ORIENT_LEFT = 1, ORIENT_RIGHT = 2, ORIENT_ON = 3 };      /* DV:117-119 */
enum { INSIDE_TRIANGLE = 4, OUTSIDE_TRIANGLE = -4 };            /* DV:120-121 */
enum { INSIDE_CIRCLE = 1, OUTSIDE_CIRCLE = 2, ON_CIRCLE = 3 };  /* DV:122-124 */

static const int IM1[3] = {2, 0, 1}, IP1[3] = {1, 2, 0}; /* DV:126, 0-based */

typedef struct fo_stats {
  long long n_orient, n_orient_on;
  long long n_incircle, n_incircle_on; /* ON_CIRCLE results in validate_DT: F3 band hits */
  long long n_flip13, n_flip24, n_flip22, n_relocate;
  long long n_virtual_replaced;
  double min_abs_incircle; /* smallest |det| seen by validate_DT */
  double min_abs_orient;
} fo_stats;

typedef struct fo_density {
  int kind; /* 0 uniform, 1 tanh refinement (SURVEY 7.3 "x4") */
  double cx, cy, cz; /* unit centre */
  double beta, alpha, gamma;
} fo_density;

typedef struct fo_mesh {
  int n, np;
  double *x, *y, *z, *lon, *lat; /* np = n + 3 (virtual at n..n+2) */
  int virt_alive[3];
  /* triangle pool */
  int cap, nslots;
  int *tv, *ta, *tsub;
  signed char *tnsub;
  unsigned char *tstate; /* 0 free, 1 in global_DT_list, 2 removed but still referenced */
  long long *tstamp, stamp;
  int *freelist, nfree;
  int *pending_free, npending, cappending;
  int *bhead;
  /* point-in-triangle relation: node = 2*p + k */
  int *cnt, *nnext, *nprev, *edge_idx, *inc;
  unsigned char *inserted;
  int *obs, nobs, capobs;
  int *tmp, ntmp, captmp;
  fo_stats st;
  /* final Delaunay triangulation, 0-based, in global_DT_list order */
  int nT;
  int *tri, *adj;
  /* Voronoi */
  double *cx, *cy, *cz, *clon, *clat;
  int *ring_off, *ring_tri, *ring_nbr;
  int have_dt, have_vd;
  struct fo_mpas *mp; /* MPAS adapter arrays (fo_mpas) */
  char err[512];
} fo_mesh;
static void mpas_free(fo_mesh *m);

/* ------------------------------------------------------------------ geometry */

/* SG:112-121 */
static void inverse_cartesian_transform(double x, double y, double z, double *lon, double *lat) {
  *lon = atan2(y, x);
  *lat = asin(z);
  if (*lon < 0.0) *lon += FO_PI2;
}

/* SG:76-85 */
static void cartesian_transform(double lon, double lat, double *x, double *y, double *z) {
  double cos_lat = cos(lat);
  *x = cos_lat * cos(lon);
  *y = cos_lat * sin(lon);
  *z = sin(lat);
}

/* SG:156-185 */
static void inverse_rotation_transform(double lon_p, double lat_p, double *lon_o, double *lat_o,
                                       double lon_r, double lat_r) {
  double sin_lon_r = sin(lon_r), cos_lon_r = cos(lon_r);
  double sin_lat_r = sin(lat_r), cos_lat_r = cos(lat_r);
  double sin_lat_p = sin(lat_p), cos_lat_p = cos(lat_p);
  double tmp1 = cos_lat_r * sin_lon_r;
  double tmp2 = sin_lat_r * cos_lat_p + cos_lat_r * cos_lon_r * sin_lat_p;
  double tmp3;
  if (fabs(tmp2) < FO_EPS) tmp2 = 0.0;
  *lon_o = atan2(tmp1, tmp2);
  *lon_o = lon_p + *lon_o;
  if (*lon_o > FO_PI2) *lon_o -= FO_PI2;
  tmp1 = sin_lat_r * sin_lat_p;
  tmp2 = cos_lat_r * cos_lat_p * cos_lon_r;
  tmp3 = tmp1 - tmp2;
  tmp3 = fmin(fmax(tmp3, -1.0), 1.0);
  *lat_o = asin(tmp3);
}

/* DV:1876-1892 orient1 */
static int orient_raw(fo_mesh *m, double x1, double y1, double z1, double x2, double y2, double z2,
                      double x0, double y0, double z0) {
  double det = x0 * (y1 * z2 - y2 * z1) - y0 * (x1 * z2 - x2 * z1) + z0 * (x1 * y2 - x2 * y1);
  double a = fabs(det);
  m->st.n_orient++;
  if (a < m->st.min_abs_orient) m->st.min_abs_orient = a;
  if (det > FO_EPS) return ORIENT_LEFT;
  if (-det > FO_EPS) return ORIENT_RIGHT;
  m->st.n_orient_on++;
  return ORIENT_ON;
}

/* DV:1894-1900 orient2 */
static int orient(fo_mesh *m, int p1, int p2, int p0) {
  return orient_raw(m, m->x[p1], m->y[p1], m->z[p1], m->x[p2], m->y[p2], m->z[p2], m->x[p0],
                    m->y[p0], m->z[p0]);
}

/* DV:2065-2088 in_circle2 (in_circle1, DV:2032-2063, is the same determinant) */
static int in_circle_pts(fo_mesh *m, int v1, int v2, int v3, int p, double *det_out) {
  double x0 = m->x[p], y0 = m->y[p], z0 = m->z[p];
  double dx1 = m->x[v1] - x0, dy1 = m->y[v1] - y0, dz1 = m->z[v1] - z0;
  double dx2 = m->x[v2] - x0, dy2 = m->y[v2] - y0, dz2 = m->z[v2] - z0;
  double dx3 = m->x[v3] - x0, dy3 = m->y[v3] - y0, dz3 = m->z[v3] - z0;
  double det = dx3 * (dy2 * dz1 - dy1 * dz2) - dy3 * (dx2 * dz1 - dx1 * dz2) +
               dz3 * (dx2 * dy1 - dx1 * dy2);
  m->st.n_incircle++;
  if (det_out) *det_out = det;
  if (det > FO_EPS) return INSIDE_CIRCLE;
  if (-det > FO_EPS) return OUTSIDE_CIRCLE;
  return ON_CIRCLE;
}

/* ------------------------------------------------------------ triangle pool */

static int fo_fail(fo_mesh *m, const char *msg) {
  if (!m->err[0]) snprintf(m->err, sizeof m->err, "%s", msg);
  return 1;
}

static void pool_grow(fo_mesh *m) {
  int ncap = m->cap ? m->cap * 2 : 1024;
  m->tv = (int *)realloc(m->tv, sizeof(int) * 3 * ncap);
  m->ta = (int *)realloc(m->ta, sizeof(int) * 3 * ncap);
  m->tsub = (int *)realloc(m->tsub, sizeof(int) * 3 * ncap);
  m->tnsub = (signed char *)realloc(m->tnsub, ncap);
  m->tstate = (unsigned char *)realloc(m->tstate, ncap);
  m->tstamp = (long long *)realloc(m->tstamp, sizeof(long long) * ncap);
  m->bhead = (int *)realloc(m->bhead, sizeof(int) * ncap);
  m->freelist = (int *)realloc(m->freelist, sizeof(int) * ncap);
  m->cap = ncap;
}

/* DV:1708-1724 create_DT: append to global_DT_list (creation order = stamp) */
static int create_DT(fo_mesh *m) {
  int t;
  if (m->nfree > 0) {
    t = m->freelist[--m->nfree];
  } else {
    if (m->nslots == m->cap) pool_grow(m);
    t = m->nslots++;
  }
  m->tv[3 * t] = m->tv[3 * t + 1] = m->tv[3 * t + 2] = -1;
  m->ta[3 * t] = m->ta[3 * t + 1] = m->ta[3 * t + 2] = -1;
  m->tnsub[t] = 0;
  m->tstate[t] = 1;
  m->tstamp[t] = m->stamp++;
  m->bhead[t] = -1;
  return t;
}

static void set_DVT(fo_mesh *m, int t, int a, int b, int c) {
  m->tv[3 * t] = a; m->tv[3 * t + 1] = b; m->tv[3 * t + 2] = c;
}
static void set_adj(fo_mesh *m, int t, int a, int b, int c) {
  m->ta[3 * t] = a; m->ta[3 * t + 1] = b; m->ta[3 * t + 2] = c;
}
static void add_sub(fo_mesh *m, int t, int s) { m->tsub[3 * t + m->tnsub[t]++] = s; }

/* array_replace_ptr (lib/container/src/array_mod.F90:168-187): first match, fatal if absent */
static int adj_replace(fo_mesh *m, int t, int oldt, int newt) {
  int k;
  for (k = 0; k < 3; k++)
    if (m->ta[3 * t + k] == oldt) { m->ta[3 * t + k] = newt; return 0; }
  return fo_fail(m, "array_replace_ptr: array failed to replace pointer!");
}

static void push_int(int **a, int *n, int *cap, int v) {
  if (*n == *cap) {
    *cap = *cap ? *cap * 2 : 64;
    *a = (int *)realloc(*a, sizeof(int) * (*cap));
  }
  (*a)[(*n)++] = v;
}

/* ------------------------------------------------ point-in-triangle relation */

static void bucket_append(fo_mesh *m, int t, int node) {
  /* order inside a bucket never influences the triangulation (every member is
     re-located independently, DV:306-315); prepend for O(1). */
  int h = m->bhead[t];
  m->nnext[node] = h;
  m->nprev[node] = -1;
  if (h >= 0) m->nprev[h] = node;
  m->bhead[t] = node;
}

static void bucket_remove(fo_mesh *m, int t, int node) {
  int p = m->nprev[node], n = m->nnext[node];
  if (p >= 0) m->nnext[p] = n; else m->bhead[t] = n;
  if (n >= 0) m->nprev[n] = p;
}

/* DV:1244-1262 */
static void record_included_DVT(fo_mesh *m, int p, int t1, int t2) {
  bucket_append(m, t1, 2 * p);
  m->cnt[2 * p] = t1;
  if (t2 >= 0) {
    bucket_append(m, t2, 2 * p + 1);
    m->cnt[2 * p + 1] = t2;
  } else {
    m->cnt[2 * p + 1] = -1;
  }
}

/* DV:1264-1275 */
static void delete_included_DVT(fo_mesh *m, int p) {
  if (m->cnt[2 * p] >= 0) bucket_remove(m, m->cnt[2 * p], 2 * p);
  if (m->cnt[2 * p + 1] >= 0) bucket_remove(m, m->cnt[2 * p + 1], 2 * p + 1);
  m->cnt[2 * p] = m->cnt[2 * p + 1] = -1;
}

/* DV:1967-2020 in_triangle_relaxed */
static int in_triangle_relaxed(fo_mesh *m, int t, int p) {
  const double eps_rel = 1.0e-4; /* real(8), parameter :: eps = 1.0e-4 (single literal) */
  const double eps_loc = (double)(float)eps_rel;
  int i, k = 0, on_plane[2] = {0, 0};
  const int *v = &m->tv[3 * t];
  for (i = 0; i < 3; i++) {
    int q = v[i];
    if (q >= m->n && fabs(m->lon[q] - m->lon[p]) < eps_loc && fabs(m->lat[q] - m->lat[p]) < eps_loc)
      return -(i + 1);
  }
  for (i = 0; i < 3; i++) {
    int ret = orient(m, v[i], v[IP1[i]], p);
    if (ret == ORIENT_RIGHT) return OUTSIDE_TRIANGLE;
    if (ret == ORIENT_ON) {
      if (k < 2) on_plane[k] = IM1[i] + 1;
      k++;
    }
  }
  if (k == 0) return INSIDE_TRIANGLE;
  if (k == 1) return on_plane[0];
  if (k == 2) {
    int a = on_plane[0], b = on_plane[1];
    if ((a == 1 && b == 2) || (a == 2 && b == 1)) return -3;
    if ((a == 2 && b == 3) || (a == 3 && b == 2)) return -1;
    return -2;
  }
  return OUTSIDE_TRIANGLE; /* k == 3: res is left undefined in the reference */
}

/* ------------------------------------------------------------- ring extract */

/* DV:704-743.  Fills tri[]/lnk[] (capacity cap) CCW; lnk[k] lies on the edge
   shared by tri[k] and tri[k+1].  Returns the degree, or -1 on failure. */
static int extract_ring(fo_mesh *m, int v, int t0, int *tri, int *lnk, int cap) {
  int t = t0, d = 0, i, link0 = -1;
  for (;;) {
    const int *tvv = &m->tv[3 * t];
    int l;
    for (i = 0; i < 3; i++) if (tvv[i] == v) break;
    if (i == 3) return -1;
    l = tvv[IM1[i]];
    if (link0 < 0) link0 = l;
    else if (l == link0) break;
    if (d == cap) return -1;
    tri[d] = t;
    lnk[d] = l;
    d++;
    t = m->ta[3 * t + IP1[i]];
  }
  return d;
}

/* -------------------------------------------------------------------- flips */

static int validate_DT(fo_mesh *m, int newDT);

/* DV:1098-1149 (the non-cyclic branch, DV:1140-1147) */
static int flip22(fo_mesh *m, int old1, int old2, int *pnew1, int *pnew2, const int idx_map[4]) {
  int D1 = m->tv[3 * old1 + idx_map[0]], D2 = m->tv[3 * old1 + idx_map[1]];
  int D3 = m->tv[3 * old1 + idx_map[2]], D4 = m->tv[3 * old2 + idx_map[3]];
  int a1 = m->ta[3 * old1 + idx_map[0]], a2 = m->ta[3 * old1 + idx_map[1]];
  int a3 = m->ta[3 * old2 + IP1[idx_map[3]]], a4 = m->ta[3 * old2 + IM1[idx_map[3]]];
  int n1 = create_DT(m), n2 = create_DT(m);
  set_DVT(m, n1, D1, D4, D3);
  set_adj(m, n1, n2, a2, a3);
  set_DVT(m, n2, D4, D2, D3);
  set_adj(m, n2, a1, n1, a4);
  if (adj_replace(m, a1, old1, n2)) return 1;
  if (adj_replace(m, a2, old1, n1)) return 1;
  if (adj_replace(m, a3, old2, n1)) return 1;
  if (adj_replace(m, a4, old2, n2)) return 1;
  m->inc[D1] = n1; m->inc[D2] = n2; m->inc[D3] = n2; m->inc[D4] = n1;
  m->st.n_flip22++;
  *pnew1 = n1; *pnew2 = n2;
  return 0;
}

/* DV:845-882 */
static int validate_DT(fo_mesh *m, int newDT) {
  int ops = m->ta[3 * newDT + 2], i, r, n1, n2;
  double det;
  for (i = 0; i < 3; i++) if (m->ta[3 * ops + i] == newDT) break;
  if (i == 3) return fo_fail(m, "validate_DT: opposite triangle does not point back");
  r = in_circle_pts(m, m->tv[3 * newDT], m->tv[3 * newDT + 1], m->tv[3 * newDT + 2],
                    m->tv[3 * ops + i], &det);
  if (fabs(det) < m->st.min_abs_incircle) m->st.min_abs_incircle = fabs(det);
  if (r == OUTSIDE_CIRCLE) return 0;
  if (r == INSIDE_CIRCLE) {
    int map[4] = {0, 1, 2, 0};
    map[3] = i;
    if (flip22(m, newDT, ops, &n1, &n2, map)) return 1;
    if (validate_DT(m, n1)) return 1;
    if (validate_DT(m, n2)) return 1;
    push_int(&m->tmp, &m->ntmp, &m->captmp, newDT);
    add_sub(m, newDT, n1); add_sub(m, newDT, n2);
    push_int(&m->obs, &m->nobs, &m->capobs, ops);
    add_sub(m, ops, n1); add_sub(m, ops, n2);
    return 0;
  }
  m->st.n_incircle_on++; /* ON_CIRCLE: the reference declines to flip, DV:878-880 */
  return 0;
}

/* DV:957-1004 */
static int flip13(fo_mesh *m, int oldDT, int p) {
  int D1 = m->tv[3 * oldDT], D2 = m->tv[3 * oldDT + 1], D3 = m->tv[3 * oldDT + 2];
  int a1 = m->ta[3 * oldDT + 2], a2 = m->ta[3 * oldDT], a3 = m->ta[3 * oldDT + 1];
  int n1 = create_DT(m), n2 = create_DT(m), n3 = create_DT(m);
  set_DVT(m, n1, D1, D2, p); set_adj(m, n1, n2, n3, a1);
  set_DVT(m, n2, D2, D3, p); set_adj(m, n2, n3, n1, a2);
  set_DVT(m, n3, D3, D1, p); set_adj(m, n3, n1, n2, a3);
  m->inc[p] = n1;
  if (adj_replace(m, a1, oldDT, n1)) return 1;
  if (adj_replace(m, a2, oldDT, n2)) return 1;
  if (adj_replace(m, a3, oldDT, n3)) return 1;
  if (m->inc[D1] == oldDT) m->inc[D1] = n1;
  if (m->inc[D2] == oldDT) m->inc[D2] = n2;
  if (m->inc[D3] == oldDT) m->inc[D3] = n3;
  if (validate_DT(m, n1)) return 1;
  if (validate_DT(m, n2)) return 1;
  if (validate_DT(m, n3)) return 1;
  push_int(&m->obs, &m->nobs, &m->capobs, oldDT);
  add_sub(m, oldDT, n1); add_sub(m, oldDT, n2); add_sub(m, oldDT, n3);
  delete_included_DVT(m, p);
  m->st.n_flip13++;
  return 0;
}

/* linked/array insert_ptr_at after a search (DV:1048-1075): overwrite first match */
static void adj_set_first(fo_mesh *m, int t, int oldt, int newt) {
  int k;
  for (k = 0; k < 3; k++)
    if (m->ta[3 * t + k] == oldt) { m->ta[3 * t + k] = newt; return; }
}

/* DV:1006-1096 */
static int flip24(fo_mesh *m, int old1, int old2, int p) {
  int i, e = m->edge_idx[p];
  int D1, D2, D3, D4, a1, a2, a3, a4, n1, n2, n3, n4;
  for (i = 0; i < 3; i++) if (m->ta[3 * old2 + i] == old1) break;
  if (i == 3) return fo_fail(m, "flip24: triangles are not adjacent");
  D1 = m->tv[3 * old1 + e]; D2 = m->tv[3 * old1 + IP1[e]]; D3 = m->tv[3 * old1 + IM1[e]];
  D4 = m->tv[3 * old2 + i];
  a1 = m->ta[3 * old1 + IP1[e]]; a2 = m->ta[3 * old1 + IM1[e]];
  a3 = m->ta[3 * old2 + IP1[i]]; a4 = m->ta[3 * old2 + IM1[i]];
  n1 = create_DT(m); n2 = create_DT(m); n3 = create_DT(m); n4 = create_DT(m);
  set_DVT(m, n1, D3, D1, p); set_adj(m, n1, n2, n4, a1);
  set_DVT(m, n2, D1, D2, p); set_adj(m, n2, n3, n1, a2);
  set_DVT(m, n3, D2, D4, p); set_adj(m, n3, n4, n2, a3);
  set_DVT(m, n4, D4, D3, p); set_adj(m, n4, n1, n3, a4);
  m->inc[p] = n1;
  adj_set_first(m, a1, old1, n1);
  adj_set_first(m, a2, old1, n2);
  adj_set_first(m, a3, old2, n3);
  adj_set_first(m, a4, old2, n4);
  if (m->inc[D1] == old1) m->inc[D1] = n1;
  if (m->inc[D2] == old1 || m->inc[D2] == old2) m->inc[D2] = n2;
  if (m->inc[D3] == old1 || m->inc[D3] == old2) m->inc[D3] = n4;
  if (m->inc[D4] == old2) m->inc[D4] = n3;
  if (validate_DT(m, n1)) return 1;
  if (validate_DT(m, n2)) return 1;
  if (validate_DT(m, n3)) return 1;
  if (validate_DT(m, n4)) return 1;
  push_int(&m->obs, &m->nobs, &m->capobs, old1);
  add_sub(m, old1, n1); add_sub(m, old1, n2);
  push_int(&m->obs, &m->nobs, &m->capobs, old2);
  add_sub(m, old2, n3); add_sub(m, old2, n4);
  delete_included_DVT(m, p);
  m->st.n_flip24++;
  return 0;
}

/* DV:884-948.  Returns 1 found, 0 not found, -1 fatal. */
static int update_point_in_triangle(fo_mesh *m, int p, int t) {
  if (m->tnsub[t] > 0) {
    int i;
    for (i = 0; i < m->tnsub[t]; i++) {
      int r = update_point_in_triangle(m, p, m->tsub[3 * t + i]);
      if (r != 0) return r;
    }
    return 0;
  } else {
    int on1 = 0, on2 = 0, r;
    int D1 = m->tv[3 * t], D2 = m->tv[3 * t + 1], D3 = m->tv[3 * t + 2];
    r = orient(m, D3, D1, p);
    if (r == ORIENT_RIGHT) return 0;
    if (r == ORIENT_ON) on1 = 2;
    r = orient(m, D2, D3, p);
    if (r == ORIENT_RIGHT) return 0;
    if (r == ORIENT_ON) on2 = 1;
    delete_included_DVT(m, p);
    m->st.n_relocate++;
    if (on1 == 0 && on2 == 0) {
      record_included_DVT(m, p, t, -1);
    } else if (on1 == 2 && on2 == 0) {
      m->edge_idx[p] = 1; /* edge index 2, 0-based 1 */
      record_included_DVT(m, p, t, m->ta[3 * t + 1]);
    } else if (on1 == 0 && on2 == 1) {
      m->edge_idx[p] = 0;
      record_included_DVT(m, p, t, m->ta[3 * t]);
    } else {
      fo_fail(m, "DVT coincides with vertex 3 of a DT (duplicate point)");
      return -1;
    }
    return 1;
  }
}

/* DV:1285-1321: remove from global_DT_list.  The slot is recycled only after
   the current insertion is over (children may still be walked, DV:303-317). */
static void retire_list(fo_mesh *m, int *list, int *n) {
  int k;
  for (k = 0; k < *n; k++) {
    int t = list[k];
    if (m->tstate[t] == 1) {
      m->tstate[t] = 2;
      push_int(&m->pending_free, &m->npending, &m->cappending, t);
    }
  }
  *n = 0;
}

static void flush_pending(fo_mesh *m) {
  int k;
  for (k = 0; k < m->npending; k++) {
    int t = m->pending_free[k];
    m->tstate[t] = 0;
    m->tnsub[t] = 0;
    m->freelist[m->nfree++] = t;
  }
  m->npending = 0;
}

/* -------------------------------------------------------- virtual deletion */

/* flip22 on a finished triangulation (cyclic branch, DV:1129-1139; rings are
   rebuilt afterwards, deviation D3). */
static int flip22_plain(fo_mesh *m, int old1, int old2, int *pn1, int *pn2, const int idx_map[4]) {
  int r = flip22(m, old1, old2, pn1, pn2, idx_map);
  if (r) return r;
  m->tstate[old1] = 2; m->tstate[old2] = 2; /* DV:827-828 */
  push_int(&m->pending_free, &m->npending, &m->cappending, old1);
  push_int(&m->pending_free, &m->npending, &m->cappending, old2);
  return 0;
}

/* DV:1151-1184 */
static int flip31(fo_mesh *m, int o1, int o2, int o3, const int idx_map[5]) {
  int D1 = m->tv[3 * o1 + idx_map[0]], D2 = m->tv[3 * o1 + idx_map[1]];
  int D4 = m->tv[3 * o2 + IM1[idx_map[3]]];
  int a1 = m->ta[3 * o1 + idx_map[2]], a2 = m->ta[3 * o2 + idx_map[3]];
  int a3 = m->ta[3 * o3 + idx_map[4]];
  int nt = create_DT(m);
  set_DVT(m, nt, D1, D2, D4);
  set_adj(m, nt, a2, a3, a1);
  if (adj_replace(m, a1, o1, nt)) return 1;
  if (adj_replace(m, a2, o2, nt)) return 1;
  if (adj_replace(m, a3, o3, nt)) return 1;
  m->inc[D1] = nt; m->inc[D2] = nt; m->inc[D4] = nt;
  m->tstate[o1] = m->tstate[o2] = m->tstate[o3] = 2;
  push_int(&m->pending_free, &m->npending, &m->cappending, o1);
  push_int(&m->pending_free, &m->npending, &m->cappending, o2);
  push_int(&m->pending_free, &m->npending, &m->cappending, o3);
  return 0;
}

/* DV:772-843 with deviation D2 (see header). */
static int delete_DVT(fo_mesh *m, int v) {
  enum { CAP = 4096 };
  static int tri[CAP], lnk[CAP];
  int d = extract_ring(m, v, m->inc[v], tri, lnk, CAP);
  int k = 0, guard = 0;
  if (d < 3) return fo_fail(m, "delete_DVT: ring extraction failed");
  while (d > 3) {
    int kp = (k + 1) % d, km = (k + d - 1) % d;
    int DT1 = tri[k], DT2 = tri[kp];
    int b0 = lnk[km], b1 = lnk[k], b2 = lnk[kp];
    int empty = 1, q, i, j, n1, n2, map[4];
    double det;
    if (++guard > 64 * CAP) return fo_fail(m, "delete_DVT: no ear can be clipped");
    /* Check 1 (DV:793): is the potential DT convex? */
    if (orient(m, b0, b1, b2) == ORIENT_RIGHT) { k = kp; continue; }
    /* Check 2 (DV:799): does the potential DT contain the vertex? */
    if (orient(m, b2, b0, v) == ORIENT_LEFT) { k = kp; continue; }
    /* Check 3 (DV:805-822): empty circumcircle w.r.t. the other link vertices */
    for (q = 0; q < d; q++) {
      if (q == km || q == k || q == kp) continue;
      if (in_circle_pts(m, b0, b1, b2, lnk[q], &det) == INSIDE_CIRCLE) { empty = 0; break; }
    }
    if (!empty) { k = kp; continue; }
    for (i = 0; i < 3; i++) if (m->tv[3 * DT1 + i] == v) break;
    for (j = 0; j < 3; j++) if (m->tv[3 * DT2 + j] == v) break;
    map[0] = IM1[i]; map[1] = i; map[2] = IP1[i]; map[3] = IM1[j]; /* DV:826 */
    if (flip22_plain(m, DT1, DT2, &n1, &n2, map)) return 1;
    /* n2 = (b2, v, b0) stays incident to v and b1 leaves the link
       (DV:1132-1137).  The walk resumes AT n2 (the reference keeps incDT_item
       on the item whose value became newDT2): re-extract the ring from n2 so
       that tri[0] = n2 and lnk[0] = b2 lies on the edge n2 | tri[1]. */
    m->inc[v] = n2;
    d = extract_ring(m, v, n2, tri, lnk, CAP);
    if (d < 3) return fo_fail(m, "delete_DVT: ring extraction failed");
    k = 0;
  }
  {
    int i, j, kk, map[5];
    int DT1 = tri[0], DT2 = tri[1], DT3 = tri[2];
    for (i = 0; i < 3; i++) if (m->tv[3 * DT1 + i] == v) break;
    for (j = 0; j < 3; j++) if (m->tv[3 * DT2 + j] == v) break;
    for (kk = 0; kk < 3; kk++) if (m->tv[3 * DT3 + kk] == v) break;
    map[0] = IP1[i]; map[1] = IM1[i]; map[2] = i; map[3] = j; map[4] = kk; /* DV:838 */
    if (flip31(m, DT1, DT2, DT3, map)) return 1;
  }
  return 0;
}

/* ------------------------------------------------------------------ public */

fo_mesh *fo_create(int n, const double *x, const double *y, const double *z) {
  fo_mesh *m = (fo_mesh *)calloc(1, sizeof(fo_mesh));
  int i;
  m->n = n; m->np = n + 3;
  m->x = (double *)malloc(sizeof(double) * m->np);
  m->y = (double *)malloc(sizeof(double) * m->np);
  m->z = (double *)malloc(sizeof(double) * m->np);
  m->lon = (double *)malloc(sizeof(double) * m->np);
  m->lat = (double *)malloc(sizeof(double) * m->np);
  for (i = 0; i < n; i++) {
    /* DV:1369-1373: x,y,z given -> lon,lat derived */
    m->x[i] = x[i]; m->y[i] = y[i]; m->z[i] = z[i];
    inverse_cartesian_transform(x[i], y[i], z[i], &m->lon[i], &m->lat[i]);
  }
  m->cnt = (int *)malloc(sizeof(int) * 2 * m->np);
  m->nnext = (int *)malloc(sizeof(int) * 2 * m->np);
  m->nprev = (int *)malloc(sizeof(int) * 2 * m->np);
  m->edge_idx = (int *)calloc(m->np, sizeof(int));
  m->inc = (int *)malloc(sizeof(int) * m->np);
  m->inserted = (unsigned char *)calloc(m->np, 1);
  return m;
}

/* DV:204-211: lon/lat entry (cartesian_transform, SG:87-98) */
fo_mesh *fo_create_lonlat(int n, const double *lon, const double *lat) {
  double *x = (double *)calloc((size_t)n * 3 + 1, sizeof(double));
  fo_mesh *m;
  int i;
  for (i = 0; i < n; i++) cartesian_transform(lon[i], lat[i], &x[i], &x[n + i], &x[2 * n + i]);
  m = fo_create(n, x, x + n, x + 2 * n);
  for (i = 0; i < n; i++) { m->lon[i] = lon[i]; m->lat[i] = lat[i]; }
  free(x);
  return m;
}

static void reset_triangulation(fo_mesh *m) {
  int i;
  m->nslots = 0; m->nfree = 0; m->npending = 0; m->nobs = 0; m->ntmp = 0; m->stamp = 0;
  memset(&m->st, 0, sizeof m->st);
  m->st.min_abs_incircle = 1e300; m->st.min_abs_orient = 1e300;
  for (i = 0; i < 2 * m->np; i++) m->cnt[i] = -1;
  for (i = 0; i < m->np; i++) { m->inc[i] = -1; m->inserted[i] = 0; m->edge_idx[i] = 0; }
  m->err[0] = 0;
  m->have_dt = m->have_vd = 0;
}

void fo_destroy(fo_mesh *m) {
  if (!m) return;
  free(m->x); free(m->y); free(m->z); free(m->lon); free(m->lat);
  free(m->tv); free(m->ta); free(m->tsub); free(m->tnsub); free(m->tstate); free(m->tstamp);
  free(m->freelist); free(m->pending_free); free(m->bhead);
  free(m->cnt); free(m->nnext); free(m->nprev); free(m->edge_idx); free(m->inc); free(m->inserted);
  free(m->obs); free(m->tmp); free(m->tri); free(m->adj);
  free(m->cx); free(m->cy); free(m->cz); free(m->clon); free(m->clat);
  free(m->ring_off); free(m->ring_tri); free(m->ring_nbr);
  mpas_free(m);
  free(m);
}

const char *fo_last_error(fo_mesh *m) { return m->err; }
void fo_get_stats(fo_mesh *m, fo_stats *out) { *out = m->st; }

/* DV:612-694 */
static int init_DTs(fo_mesh *m, int i1, int i2, int i3) {
  static const int amap[24] = {5,2,4, 6,3,1, 7,4,2, 8,1,3, 1,8,6, 2,5,7, 3,6,8, 4,7,5};
  static const int vleft[24] = {1,2,3, 1,3,5, 1,5,6, 1,6,2, 4,3,2, 4,5,3, 4,6,5, 4,2,6};
  static const int vright[24] = {1,3,2, 1,2,6, 1,6,5, 1,5,3, 4,2,3, 4,6,2, 4,5,6, 4,3,5};
  const int *vmap;
  int dvt[6], dt[8], k, n = m->n;
  int src[3];
  src[0] = i1; src[1] = i2; src[2] = i3;
  for (k = 0; k < 3; k++) {
    double lon, lat;
    int q = n + k;
    inverse_rotation_transform(m->lon[src[k]], m->lat[src[k]], &lon, &lat, 0.0, -FO_PI05);
    m->lon[q] = lon; m->lat[q] = lat;
    cartesian_transform(lon, lat, &m->x[q], &m->y[q], &m->z[q]); /* DV:1365-1368 */
    m->virt_alive[k] = 1;
  }
  dvt[0] = i1; dvt[1] = i2; dvt[2] = i3; dvt[3] = n; dvt[4] = n + 1; dvt[5] = n + 2;
  for (k = 0; k < 8; k++) dt[k] = create_DT(m);
  for (k = 0; k < 8; k++)
    set_adj(m, dt[k], dt[amap[3 * k] - 1], dt[amap[3 * k + 1] - 1], dt[amap[3 * k + 2] - 1]);
  switch (orient(m, i1, i2, i3)) {
    case ORIENT_LEFT: vmap = vleft; break;
    case ORIENT_RIGHT: vmap = vright; break;
    default: return fo_fail(m, "The first three points are collinear.");
  }
  for (k = 0; k < 8; k++)
    set_DVT(m, dt[k], dvt[vmap[3 * k] - 1], dvt[vmap[3 * k + 1] - 1], dvt[vmap[3 * k + 2] - 1]);
  m->inc[i1] = m->inc[i2] = m->inc[i3] = dt[0];
  m->inc[n] = m->inc[n + 1] = m->inc[n + 2] = dt[6];
  return 0;
}

/* DV:226-356.  i1,i2,i3 are 0-based indices of the bootstrap vertices (D1). */
int fo_triangulate(fo_mesh *m, int i1, int i2, int i3) {
  int n = m->n, i, k, t;
  if (n < 3 || i1 == i2 || i1 == i3 || i2 == i3 || i1 < 0 || i2 < 0 || i3 < 0 || i1 >= n ||
      i2 >= n || i3 >= n)
    return fo_fail(m, "bad bootstrap indices");
  reset_triangulation(m);
  if (init_DTs(m, i1, i2, i3)) return 1;
  m->inserted[i1] = m->inserted[i2] = m->inserted[i3] = 1;

  /* DV:247-293 initial point-in-triangle relation */
  for (i = 0; i < n; i++) {
    if (m->inserted[i]) continue;
    for (t = 0; t < 8; t++) { /* global_DT_list holds exactly the 8 bootstrap triangles */
      int ret = in_triangle_relaxed(m, t, i);
      if (ret == INSIDE_TRIANGLE) { record_included_DVT(m, i, t, -1); break; }
      if (ret == OUTSIDE_TRIANGLE) continue;
      if (ret > 0) {
        m->edge_idx[i] = ret - 1;
        record_included_DVT(m, i, t, m->ta[3 * t + ret - 1]);
        break;
      }
      {
        int q = m->tv[3 * t + (-ret - 1)];
        if (q >= n) { /* CASE 3-1, DV:268-282 */
          enum { CAP = 64 };
          int rt[CAP], rl[CAP], d = extract_ring(m, q, m->inc[q], rt, rl, CAP), j, c;
          if (d < 0) return fo_fail(m, "ring extraction failed for a virtual vertex");
          for (j = 0; j < d; j++)
            for (c = 0; c < 3; c++)
              if (m->tv[3 * rt[j] + c] == q) m->tv[3 * rt[j] + c] = i;
          m->inc[i] = rt[0];
          m->inserted[i] = 1;
          m->virt_alive[q - n] = 0;
          m->st.n_virtual_replaced++;
          break;
        }
        snprintf(m->err, sizeof m->err, "DVT %d coincides to DVT %d of DT %d.", i + 1, -ret, t + 1);
        return 1; /* CASE 3-2, DV:286-288 */
      }
    }
  }

  /* DV:296-321 insertion loop, index order */
  for (i = 0; i < n; i++) {
    int no;
    if (m->inserted[i]) continue;
    m->inserted[i] = 1;
    if (m->cnt[2 * i] < 0) {
      snprintf(m->err, sizeof m->err, "DVT %d is contained by no triangle.", i + 1);
      return 1;
    }
    if (m->cnt[2 * i + 1] >= 0) { /* DV:756-762 */
      if (flip24(m, m->cnt[2 * i], m->cnt[2 * i + 1], i)) return 1;
    } else {
      if (flip13(m, m->cnt[2 * i], i)) return 1;
    }
    no = m->nobs;
    for (k = 0; k < no; k++) {
      int ot = m->obs[k], node = m->bhead[ot];
      while (node >= 0) {
        int nx = m->nnext[node], p = node >> 1;
        int r = update_point_in_triangle(m, p, ot);
        if (r < 0) return 1;
        if (r == 0) {
          snprintf(m->err, sizeof m->err, "Failed to update DVT (%d) after delete DT!", p + 1);
          return 1;
        }
        node = nx;
      }
    }
    retire_list(m, m->obs, &m->nobs);
    retire_list(m, m->tmp, &m->ntmp);
    flush_pending(m);
  }

  /* DV:334-338 delete virtual DVTs */
  for (k = 0; k < 3; k++) {
    if (!m->virt_alive[k]) continue;
    if (delete_DVT(m, n + k)) return 1;
    flush_pending(m);
  }

  /* DV:341-352: number the surviving triangles in global_DT_list (creation) order */
  {
    int cnt = 0, s, *order, *newid;
    long long *key;
    for (s = 0; s < m->nslots; s++) if (m->tstate[s] == 1) cnt++;
    order = (int *)malloc(sizeof(int) * (cnt > 0 ? cnt : 1));
    newid = (int *)malloc(sizeof(int) * (m->nslots > 0 ? m->nslots : 1));
    key = m->tstamp;
    cnt = 0;
    for (s = 0; s < m->nslots; s++) if (m->tstate[s] == 1) order[cnt++] = s;
    /* sort by stamp: shell sort is enough, stamps are nearly sorted by slot */
    {
      int gap, a, b;
      for (gap = cnt / 2; gap > 0; gap /= 2)
        for (a = gap; a < cnt; a++) {
          int tmp = order[a];
          for (b = a; b >= gap && key[order[b - gap]] > key[tmp]; b -= gap) order[b] = order[b - gap];
          order[b] = tmp;
        }
    }
    for (s = 0; s < m->nslots; s++) newid[s] = -1;
    for (s = 0; s < cnt; s++) newid[order[s]] = s;
    free(m->tri); free(m->adj);
    m->tri = (int *)malloc(sizeof(int) * 3 * (cnt > 0 ? cnt : 1));
    m->adj = (int *)malloc(sizeof(int) * 3 * (cnt > 0 ? cnt : 1));
    for (s = 0; s < cnt; s++) {
      int o = order[s], c;
      for (c = 0; c < 3; c++) {
        m->tri[3 * s + c] = m->tv[3 * o + c];
        m->adj[3 * s + c] = newid[m->ta[3 * o + c]];
      }
    }
    for (i = 0; i < n; i++) if (m->inc[i] >= 0) m->inc[i] = newid[m->inc[i]];
    m->nT = cnt;
    free(order); free(newid);
  }
  m->have_dt = 1;
  return 0;
}

int fo_num_triangles(fo_mesh *m) { return m->nT; }

/* 1-based output like the Fortran ids (DV:341-352): tri/adj are 3 x nT, column major */
void fo_get_triangles(fo_mesh *m, int *tri, int *adj) {
  int k;
  for (k = 0; k < 3 * m->nT; k++) {
    if (tri) tri[k] = m->tri[k] + 1;
    if (adj) adj[k] = m->adj[k] + 1;
  }
}

/* DV:379-409: circumcentres (calc_circumcircle, DV:1323-1348) and cell rings.
   Rings are rotated to start at the smallest neighbour id (ring starts are not
   comparable across implementations). */
int fo_voronoi(fo_mesh *m) {
  int n = m->n, nT = m->nT, t, i, total;
  if (!m->have_dt) return fo_fail(m, "voronoi_diagram before delaunay_triangulation");
  free(m->cx); free(m->cy); free(m->cz); free(m->clon); free(m->clat);
  m->cx = (double *)malloc(sizeof(double) * nT); m->cy = (double *)malloc(sizeof(double) * nT);
  m->cz = (double *)malloc(sizeof(double) * nT); m->clon = (double *)malloc(sizeof(double) * nT);
  m->clat = (double *)malloc(sizeof(double) * nT);
  for (t = 0; t < nT; t++) {
    int a = m->tri[3 * t], b = m->tri[3 * t + 1], c = m->tri[3 * t + 2];
    double e3[3], e2[3], nn[3], l;
    e3[0] = m->x[b] - m->x[a]; e3[1] = m->y[b] - m->y[a]; e3[2] = m->z[b] - m->z[a];
    e2[0] = m->x[c] - m->x[a]; e2[1] = m->y[c] - m->y[a]; e2[2] = m->z[c] - m->z[a];
    nn[0] = e3[1] * e2[2] - e3[2] * e2[1]; /* cross_product, src/math_mod.F90:12-22 */
    nn[1] = e3[2] * e2[0] - e3[0] * e2[2];
    nn[2] = e3[0] * e2[1] - e3[1] * e2[0];
    l = sqrt(nn[0] * nn[0] + nn[1] * nn[1] + nn[2] * nn[2]);
    if (l < FO_EPS) {
      snprintf(m->err, sizeof m->err, "Vertices %d, %d, %d are collinear!", a + 1, b + 1, c + 1);
      return 1;
    }
    m->cx[t] = nn[0] / l; m->cy[t] = nn[1] / l; m->cz[t] = nn[2] / l;
    inverse_cartesian_transform(m->cx[t], m->cy[t], m->cz[t], &m->clon[t], &m->clat[t]);
  }
  /* rings on the final arrays */
  free(m->ring_off); free(m->ring_tri); free(m->ring_nbr);
  m->ring_off = (int *)malloc(sizeof(int) * (n + 1));
  total = 3 * nT;
  m->ring_tri = (int *)malloc(sizeof(int) * (total > 0 ? total : 1));
  m->ring_nbr = (int *)malloc(sizeof(int) * (total > 0 ? total : 1));
  {
    int *first = (int *)malloc(sizeof(int) * n), pos = 0;
    for (i = 0; i < n; i++) first[i] = -1;
    for (t = nT - 1; t >= 0; t--)
      for (i = 0; i < 3; i++) first[m->tri[3 * t + i]] = t;
    for (i = 0; i < n; i++) {
      int tt = first[i], d = 0, k, kmin = 0, start = pos;
      m->ring_off[i] = pos;
      if (tt < 0) { free(first); return fo_fail(m, "vertex without incident triangle"); }
      for (;;) {
        int c;
        const int *tv = &m->tri[3 * tt];
        for (c = 0; c < 3; c++) if (tv[c] == i) break;
        if (c == 3 || pos >= total) { free(first); return fo_fail(m, "ring walk failed"); }
        m->ring_tri[pos] = tt;
        m->ring_nbr[pos] = tv[IM1[c]];
        pos++; d++;
        tt = m->adj[3 * tt + IP1[c]];
        if (tt == first[i]) break;
      }
      /* canonical rotation: ring_nbr[start] is the smallest neighbour id */
      for (k = 1; k < d; k++) if (m->ring_nbr[start + k] < m->ring_nbr[start + kmin]) kmin = k;
      if (kmin) {
        int *tmp = (int *)malloc(sizeof(int) * 2 * d);
        for (k = 0; k < d; k++) {
          tmp[k] = m->ring_tri[start + (k + kmin) % d];
          tmp[d + k] = m->ring_nbr[start + (k + kmin) % d];
        }
        memcpy(&m->ring_tri[start], tmp, sizeof(int) * d);
        memcpy(&m->ring_nbr[start], tmp + d, sizeof(int) * d);
        free(tmp);
      }
    }
    m->ring_off[n] = pos;
    free(first);
  }
  m->have_vd = 1;
  return 0;
}

void fo_get_circumcentres(fo_mesh *m, double *xyz /*3 x nT*/, double *lonlat /*2 x nT*/) {
  int t;
  for (t = 0; t < m->nT; t++) {
    if (xyz) { xyz[3 * t] = m->cx[t]; xyz[3 * t + 1] = m->cy[t]; xyz[3 * t + 2] = m->cz[t]; }
    if (lonlat) { lonlat[2 * t] = m->clon[t]; lonlat[2 * t + 1] = m->clat[t]; }
  }
}

/* 1-based; ring_tri[k] is the k-th incident triangle (= Voronoi vertex id) and
   ring_nbr[k] the neighbour cell on the edge ring_tri[k] | ring_tri[k+1] (DV:394-409). */
void fo_get_rings(fo_mesh *m, int *off, int *ring_tri, int *ring_nbr) {
  int i, tot = m->ring_off[m->n];
  for (i = 0; i <= m->n; i++) off[i] = m->ring_off[i];
  for (i = 0; i < tot; i++) { ring_tri[i] = m->ring_tri[i] + 1; ring_nbr[i] = m->ring_nbr[i] + 1; }
}

void fo_get_generators(fo_mesh *m, double *x, double *y, double *z) {
  memcpy(x, m->x, sizeof(double) * m->n);
  memcpy(y, m->y, sizeof(double) * m->n);
  memcpy(z, m->z, sizeof(double) * m->n);
}

void fo_set_generators(fo_mesh *m, const double *x, const double *y, const double *z) {
  int i;
  for (i = 0; i < m->n; i++) {
    m->x[i] = x[i]; m->y[i] = y[i]; m->z[i] = z[i];
    inverse_cartesian_transform(x[i], y[i], z[i], &m->lon[i], &m->lat[i]);
  }
  m->have_dt = m->have_vd = 0;
}

/* ------------------------------------------------------- Lloyd */

static double density_eval(const fo_density *rho, double x, double y, double z) {
  if (rho->kind == 0) return 1.0;
  {
    double dot = x * rho->cx + y * rho->cy + z * rho->cz;
    double d = acos(fmax(fmin(dot, 1.0), -1.0)); /* as calc_arc_length_1, SG:289-296 */
    return 0.5 * (1.0 - rho->gamma) * (tanh((rho->beta - d) / rho->alpha) + 1.0) + rho->gamma;
  }
}

/* rho at n unit vectors (for the parity tests of the density function itself) */
void fo_density_eval(const fo_density *rho, long long n, const double *x, const double *y, const double *z,
                     double *out) {
  long long i;
  for (i = 0; i < n; i++) out[i] = density_eval(rho, x[i], y[i], z[i]);
}

/* degree-5, 7-point symmetric rule on the triangle (Radon) */
static const double QA1 = 0.10128650732345633880, QB1 = 0.79742698535308732240;
static const double QA2 = 0.47014206410511508977, QB2 = 0.05971587178976982045;
static const double QW0 = 0.225, QW1 = 0.12593918054482715260, QW2 = 0.13239415278850618074;

/* Centroids of all cells on the current Voronoi diagram; out arrays of n. */
int fo_centroids(fo_mesh *m, const fo_density *rho, double *gx, double *gy, double *gz) {
  int n = m->n, i, j, k;
  static double L[7][3];
  static double W[7];
  if (!m->have_vd) return fo_fail(m, "centroids before voronoi_diagram");
  L[0][0] = L[0][1] = L[0][2] = 1.0 / 3.0; W[0] = QW0;
  L[1][0] = QB1; L[1][1] = QA1; L[1][2] = QA1; W[1] = QW1;
  L[2][0] = QA1; L[2][1] = QB1; L[2][2] = QA1; W[2] = QW1;
  L[3][0] = QA1; L[3][1] = QA1; L[3][2] = QB1; W[3] = QW1;
  L[4][0] = QB2; L[4][1] = QA2; L[4][2] = QA2; W[4] = QW2;
  L[5][0] = QA2; L[5][1] = QB2; L[5][2] = QA2; W[5] = QW2;
  L[6][0] = QA2; L[6][1] = QA2; L[6][2] = QB2; W[6] = QW2;
  for (i = 0; i < n; i++) {
    int o = m->ring_off[i], d = m->ring_off[i + 1] - o;
    double g0 = m->x[i], g1 = m->y[i], g2 = m->z[i];
    double M0 = 0.0, M1 = 0.0, M2 = 0.0, nrm;
    for (j = 0; j < d; j++) {
      int ta = m->ring_tri[o + j], tb = m->ring_tri[o + (j + 1) % d];
      double a0 = m->cx[ta], a1 = m->cy[ta], a2 = m->cz[ta];
      double b0 = m->cx[tb], b1 = m->cy[tb], b2 = m->cz[tb];
      double u0 = a0 - g0, u1 = a1 - g1, u2 = a2 - g2;
      double v0 = b0 - g0, v1 = b1 - g1, v2 = b2 - g2;
      double c0 = u1 * v2 - u2 * v1, c1 = u2 * v0 - u0 * v2, c2 = u0 * v1 - u1 * v0;
      double A = 0.5 * sqrt(c0 * c0 + c1 * c1 + c2 * c2);
      double s0 = 0.0, s1 = 0.0, s2 = 0.0;
      for (k = 0; k < 7; k++) {
        double q0 = L[k][0] * g0 + L[k][1] * a0 + L[k][2] * b0;
        double q1 = L[k][0] * g1 + L[k][1] * a1 + L[k][2] * b1;
        double q2 = L[k][0] * g2 + L[k][1] * a2 + L[k][2] * b2;
        double r = sqrt(q0 * q0 + q1 * q1 + q2 * q2);
        double w;
        q0 /= r; q1 /= r; q2 /= r;
        w = W[k] * density_eval(rho, q0, q1, q2);
        s0 += w * q0; s1 += w * q1; s2 += w * q2;
      }
      M0 += A * s0; M1 += A * s1; M2 += A * s2;
    }
    nrm = sqrt(M0 * M0 + M1 * M1 + M2 * M2);
    if (!(nrm > 0.0)) return fo_fail(m, "degenerate Voronoi cell in centroid");
    gx[i] = M0 / nrm; gy[i] = M1 / nrm; gz[i] = M2 / nrm;
  }
  return 0;
}

/* One Lloyd iteration on the CURRENT generators: DT + VD + centroid + update.
   Returns max chord displacement through *max_disp. */
int fo_lloyd_step(fo_mesh *m, const fo_density *rho, int i1, int i2, int i3, double *max_disp) {
  int n = m->n, i;
  double *g, md = 0.0;
  if (fo_triangulate(m, i1, i2, i3)) return 1;
  if (fo_voronoi(m)) return 1;
  g = (double *)malloc(sizeof(double) * 3 * n);
  if (fo_centroids(m, rho, g, g + n, g + 2 * n)) { free(g); return 1; }
  for (i = 0; i < n; i++) {
    double dx = g[i] - m->x[i], dy = g[n + i] - m->y[i], dz = g[2 * n + i] - m->z[i];
    double dd = sqrt(dx * dx + dy * dy + dz * dz);
    if (dd > md) md = dd;
  }
  for (i = 0; i < n; i++) {
    m->x[i] = g[i]; m->y[i] = g[n + i]; m->z[i] = g[2 * n + i];
    inverse_cartesian_transform(m->x[i], m->y[i], m->z[i], &m->lon[i], &m->lat[i]);
  }
  free(g);
  *max_disp = md;
  return 0;
}

/* ---------------------------------------------- quality stats (SG:198-263) */

static void cross3(const double *a, const double *b, double *r) {
  r[0] = a[1] * b[2] - a[2] * b[1];
  r[1] = a[2] * b[0] - a[0] * b[2];
  r[2] = a[0] * b[1] - a[1] * b[0];
}
static void norm_vector(double *v) { /* SG:229-243 */
  double nn = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  if (nn != 0) { v[0] /= nn; v[1] /= nn; v[2] /= nn; }
}
/* SG:247-263 calc_sphere_angle_1 */
static double calc_sphere_angle(const double *a, const double *b, const double *c) {
  double nab[3], nbc[3], cr[3], dp, res;
  cross3(a, b, nab); norm_vector(nab);
  cross3(b, c, nbc); norm_vector(nbc);
  dp = nab[0] * nbc[0] + nab[1] * nbc[1] + nab[2] * nbc[2];
  res = acos(-fmax(fmin(dp, 1.0), -1.0));
  cross3(nab, nbc, cr);
  if (cr[0] * a[0] + cr[1] * a[1] + cr[2] * a[2] < 0.0) res = FO_PI2 - res;
  return res;
}
/* SG:198-227 calc_area */
double fo_calc_area(int n, const double *x, const double *y, const double *z) {
  int i;
  double res = 0.0;
  for (i = 0; i < n; i++) {
    int im = i ? i - 1 : n - 1, ip = (i + 1 < n) ? i + 1 : 0;
    double a[3], b[3], c[3];
    a[0] = x[im]; a[1] = y[im]; a[2] = z[im];
    b[0] = x[i]; b[1] = y[i]; b[2] = z[i];
    c[0] = x[ip]; c[1] = y[ip]; c[2] = z[ip];
    res += calc_sphere_angle(a, b, c);
  }
  return res - (n - 2) * FO_PI;
}

/* areaCell (MM:296-308) and areaTriangle (MM:310-322); returns the two sums. */
int fo_areas(fo_mesh *m, double *area_cell, double *area_tri, double *sum_cell, double *sum_tri) {
  int i, t, k;
  double sc = 0.0, stt = 0.0;
  if (!m->have_vd) return fo_fail(m, "areas before voronoi_diagram");
  for (i = 0; i < m->n; i++) {
    int o = m->ring_off[i], d = m->ring_off[i + 1] - o;
    double px[64], py[64], pz[64], a;
    if (d > 64) return fo_fail(m, "cell degree > 64");
    for (k = 0; k < d; k++) {
      int tt = m->ring_tri[o + k];
      px[k] = m->cx[tt]; py[k] = m->cy[tt]; pz[k] = m->cz[tt];
    }
    a = fo_calc_area(d, px, py, pz);
    if (area_cell) area_cell[i] = a;
    sc += a;
  }
  for (t = 0; t < m->nT; t++) {
    double px[3], py[3], pz[3], a;
    for (k = 0; k < 3; k++) {
      int v = m->tri[3 * t + k];
      px[k] = m->x[v]; py[k] = m->y[v]; pz[k] = m->z[v];
    }
    a = fo_calc_area(3, px, py, pz);
    if (area_tri) area_tri[t] = a;
    stt += a;
  }
  *sum_cell = sc; *sum_tri = stt;
  return 0;
}

/* Mesh-quality statistics (last sentence; no reference procedure of its own: the
   building blocks are calc_area SG:198-227, calc_sphere_angle SG:247-285, the Sum = 4 pi checks
   MM:306 / MM:320, and the centroid map of fo_centroids).  Layout = fm_quality_stats. */
typedef struct fo_quality_stats {
  long long degree_hist[16];
  double area_cell_min, area_cell_max, area_cell_sum;
  double area_tri_min, area_tri_max, area_tri_sum;
  double angle_min, angle_max;
  double max_disp;
} fo_quality_stats;

int fo_quality(fo_mesh *m, const fo_density *rho, fo_quality_stats *q) {
  int n = m->n, i, t, k;
  double *ac, *at, *g;
  if (!m->have_vd) return fo_fail(m, "quality before voronoi_diagram");
  memset(q, 0, sizeof *q);
  ac = (double *)malloc(sizeof(double) * n);
  at = (double *)malloc(sizeof(double) * m->nT);
  if (fo_areas(m, ac, at, &q->area_cell_sum, &q->area_tri_sum)) { free(ac); free(at); return 1; }
  q->area_cell_min = q->area_tri_min = q->angle_min = 1e300;
  for (i = 0; i < n; i++) {
    int d = m->ring_off[i + 1] - m->ring_off[i];
    q->degree_hist[d < 15 ? d : 15]++;
    if (ac[i] < q->area_cell_min) q->area_cell_min = ac[i];
    if (ac[i] > q->area_cell_max) q->area_cell_max = ac[i];
  }
  for (t = 0; t < m->nT; t++) {
    double p[3][3];
    if (at[t] < q->area_tri_min) q->area_tri_min = at[t];
    if (at[t] > q->area_tri_max) q->area_tri_max = at[t];
    for (k = 0; k < 3; k++) {
      int v = m->tri[3 * t + k];
      p[k][0] = m->x[v]; p[k][1] = m->y[v]; p[k][2] = m->z[v];
    }
    for (k = 0; k < 3; k++) {
      double a = calc_sphere_angle(p[(k + 2) % 3], p[k], p[(k + 1) % 3]);
      if (a < q->angle_min) q->angle_min = a;
      if (a > q->angle_max) q->angle_max = a;
    }
  }
  free(ac); free(at);
  g = (double *)malloc(sizeof(double) * 3 * n);
  if (fo_centroids(m, rho, g, g + n, g + 2 * n)) { free(g); return 1; }
  for (i = 0; i < n; i++) {
    double dx = g[i] - m->x[i], dy = g[n + i] - m->y[i], dz = g[2 * n + i] - m->z[i];
    double dd = sqrt(dx * dx + dy * dy + dz * dz);
    if (dd > q->max_disp) q->max_disp = dd;
  }
  free(g);
  return 0;
}

/* ------------------------------------------- MPAS adapter arithmetic (MM) */
/* Restates mpas_mesh_adapt (MM:181-345) and calc_weightsOnEdge (MM:442-561) on
   flat arrays, with the container graph the reference walks rebuilt first:
   Delaunay / Voronoi edges in creation order (DV:413-460) and the ordered
   VE-in-VC, VE-in-VVT, VC-in-VVT lists (DV:462-582).  0-based inside, 1-based
   through the getters; rows of per-cell arrays have stride maxEdges, rows of
   edgesOnEdge / weightsOnEdge stride maxEdges2 = 2 maxEdges (deviation D5). */
typedef struct fo_mpas {
  int nCells, nEdges, nVertices, maxEdges, maxEdges2;
  int *cellsOnEdge, *verticesOnEdge, *nEdgesOnEdge, *edgesOnEdge;
  double *xEdge, *yEdge, *zEdge, *lonEdge, *latEdge;
  double *dvEdge, *dcEdge, *dv1Edge, *dv2Edge, *angleEdge, *weightsOnEdge;
  int *nEdgesOnCell, *verticesOnCell, *cellsOnCell, *edgesOnCell;
  double *areaCell;
  int *cellsOnVertex, *edgesOnVertex;
  double *areaTriangle, *kiteAreasOnVertex;
} fo_mpas_t;

static void mpas_free(fo_mesh *m) {
  fo_mpas_t *p = m->mp;
  if (!p) return;
  free(p->cellsOnEdge); free(p->verticesOnEdge); free(p->nEdgesOnEdge); free(p->edgesOnEdge);
  free(p->xEdge); free(p->yEdge); free(p->zEdge); free(p->lonEdge); free(p->latEdge);
  free(p->dvEdge); free(p->dcEdge); free(p->dv1Edge); free(p->dv2Edge); free(p->angleEdge);
  free(p->weightsOnEdge);
  free(p->nEdgesOnCell); free(p->verticesOnCell); free(p->cellsOnCell); free(p->edgesOnCell);
  free(p->areaCell); free(p->cellsOnVertex); free(p->edgesOnVertex);
  free(p->areaTriangle); free(p->kiteAreasOnVertex);
  free(p);
  m->mp = NULL;
}

static double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
/* SG:289-296 calc_arc_length */
static double calc_arc_length(const double *a, const double *b) {
  return acos(fmax(fmin(dot3(a, b), 1.0), -1.0));
}

/* SG:316-363 intersect: great circles (a,b) and (c,d); e = the intersection between
   the end points, through lon/lat and back as the reference does.  0 = none. */
static int sg_intersect(const double *a, const double *b, const double *c, const double *d,
                        double *e, double *elon, double *elat) {
  double n1[3], n2[3], v[3], r, lon1, lon2, lat1, lat2, ae[3], be[3], ce[3], de[3];
  cross3(a, b, n1);
  cross3(c, d, n2);
  cross3(n1, n2, v);
  r = sqrt(dot3(v, v));
  if (r > FO_EPS) { v[0] /= r; v[1] /= r; v[2] /= r; } else return 0;
  lat1 = asin(v[2]);
  lat2 = -lat1;
  lon1 = atan2(v[1], v[0]);
  lon2 = lon1 - FO_PI;
  if (lon1 < 0.0) lon1 += FO_PI2;
  if (lon1 > FO_PI2) lon1 -= FO_PI2;
  if (lon2 < 0.0) lon2 += FO_PI2;
  if (lon2 > FO_PI2) lon2 -= FO_PI2;
  *elon = lon1; *elat = lat1;
  cartesian_transform(lon1, lat1, &e[0], &e[1], &e[2]);
  cross3(a, e, ae); cross3(b, e, be); cross3(c, e, ce); cross3(d, e, de);
  if (dot3(ae, be) >= 0 || dot3(ce, de) >= 0) {
    *elon = lon2; *elat = lat2;
    cartesian_transform(lon2, lat2, &e[0], &e[1], &e[2]);
  }
  return 1;
}

#define MP_ALLOC(field, type, count) p->field = (type *)calloc((size_t)(count) > 0 ? (size_t)(count) : 1, sizeof(type))

/* strict != 0: the reference's run-time checks are fatal (MM:306, :320, :339, :469-475) -- they fire
   on any mesh with an obtuse Delaunay triangle; strict == 0: the arithmetic is carried through. */
int fo_mpas(fo_mesh *m, int strict, int *max_edges_out) {
  int n = m->n, nT = m->nT, nE, i, j, k, e, t, maxEdges = 0, ME, ME2;
  fo_mpas_t *p;
  int *ring_edge; /* per ring position: the edge to ring_nbr (VE-in-VC, ordered: DV:462-503) */
  if (!m->have_vd) return fo_fail(m, "mpas_mesh_output before voronoi_diagram");
  mpas_free(m);
  p = m->mp = (fo_mpas_t *)calloc(1, sizeof(fo_mpas_t));
  nE = nT + n - 2; /* euler_formula, SG:57-71 */
  for (i = 0; i < n; i++) {
    int d = m->ring_off[i + 1] - m->ring_off[i];
    if (d > maxEdges) maxEdges = d;
  }
  ME = p->maxEdges = maxEdges; ME2 = p->maxEdges2 = 2 * maxEdges;
  p->nCells = n; p->nEdges = nE; p->nVertices = nT;
  if (max_edges_out) *max_edges_out = ME;
  MP_ALLOC(cellsOnEdge, int, 2 * nE); MP_ALLOC(verticesOnEdge, int, 2 * nE);
  MP_ALLOC(nEdgesOnEdge, int, nE); MP_ALLOC(edgesOnEdge, int, (size_t)ME2 * nE);
  MP_ALLOC(xEdge, double, nE); MP_ALLOC(yEdge, double, nE); MP_ALLOC(zEdge, double, nE);
  MP_ALLOC(lonEdge, double, nE); MP_ALLOC(latEdge, double, nE);
  MP_ALLOC(dvEdge, double, nE); MP_ALLOC(dcEdge, double, nE); MP_ALLOC(dv1Edge, double, nE);
  MP_ALLOC(dv2Edge, double, nE); MP_ALLOC(angleEdge, double, nE);
  MP_ALLOC(weightsOnEdge, double, (size_t)ME2 * nE);
  MP_ALLOC(nEdgesOnCell, int, n); MP_ALLOC(verticesOnCell, int, (size_t)ME * n);
  MP_ALLOC(cellsOnCell, int, (size_t)ME * n); MP_ALLOC(edgesOnCell, int, (size_t)ME * n);
  MP_ALLOC(areaCell, double, n);
  MP_ALLOC(cellsOnVertex, int, 3 * nT); MP_ALLOC(edgesOnVertex, int, 3 * nT);
  MP_ALLOC(areaTriangle, double, nT); MP_ALLOC(kiteAreasOnVertex, double, 3 * nT);
  for (i = 0; i < (int)((size_t)ME * n); i++) p->verticesOnCell[i] = p->cellsOnCell[i] = p->edgesOnCell[i] = -1;
  for (i = 0; i < ME2 * nE; i++) p->edgesOnEdge[i] = -1;

  /* Delaunay + Voronoi edges, DV:413-460: cell i creates the edge to every ring
     neighbour with a larger id, in ring order; VVT1 = ring(j), VVT2 = ring(j+1) */
  ring_edge = (int *)malloc(sizeof(int) * (3 * (size_t)nT + 1));
  e = 0;
  for (i = 0; i < n; i++) {
    int o = m->ring_off[i], d = m->ring_off[i + 1] - o;
    for (j = 0; j < d; j++) {
      int nb = m->ring_nbr[o + j];
      if (nb > i) {
        if (e >= nE) { free(ring_edge); return fo_fail(m, "more edges than euler_formula allows"); }
        p->cellsOnEdge[2 * e] = i; p->cellsOnEdge[2 * e + 1] = nb;
        p->verticesOnEdge[2 * e] = m->ring_tri[o + j];
        p->verticesOnEdge[2 * e + 1] = m->ring_tri[o + (j + 1) % d];
        ring_edge[o + j] = e;
        e++;
      }
    }
  }
  if (e != nE) { free(ring_edge); return fo_fail(m, "edge count differs from euler_formula"); }
  /* the other cell of each edge finds it at the ring position of the creator */
  for (i = 0; i < n; i++) {
    int o = m->ring_off[i], d = m->ring_off[i + 1] - o;
    for (j = 0; j < d; j++) {
      int nb = m->ring_nbr[o + j];
      if (nb < i) {
        int o2 = m->ring_off[nb], d2 = m->ring_off[nb + 1] - o2, q;
        for (q = 0; q < d2; q++) if (m->ring_nbr[o2 + q] == i) break;
        if (q == d2) { free(ring_edge); return fo_fail(m, "Internal error! (edge not found in the neighbour's ring)"); }
        ring_edge[o + j] = ring_edge[o2 + q];
      }
    }
  }
  /* cells, MM:205-226.  VE-in-VC after the fix-up of DV:462-503: edge k runs from
     vertex k to vertex k+1 with the cell on its left, i.e. it is the edge to adjVC(k) */
  for (i = 0; i < n; i++) {
    int o = m->ring_off[i], d = m->ring_off[i + 1] - o;
    p->nEdgesOnCell[i] = d;
    for (j = 0; j < d; j++) {
      p->verticesOnCell[(size_t)ME * i + j] = m->ring_tri[o + j];
      p->cellsOnCell[(size_t)ME * i + j] = m->ring_nbr[o + j];
      p->edgesOnCell[(size_t)ME * i + j] = ring_edge[o + j];
    }
  }
  /* vertices, MM:228-242.  VE-in-VVT / VC-in-VVT after the fix-ups of DV:504-582:
     counter-clockwise, VC(i) on the left of VE(i) seen from the vertex, starting at
     the first edge appended to the vertex = its lowest edge id (DV:454-455) */
  for (t = 0; t < nT; t++) {
    int ed[3], kmin = 0;
    for (k = 0; k < 3; k++) {
      int v = m->tri[3 * t + k], o = m->ring_off[v], d = m->ring_off[v + 1] - o, q;
      for (q = 0; q < d; q++) if (m->ring_tri[o + q] == t) break;
      if (q == d) { free(ring_edge); return fo_fail(m, "Internal error! (vertex not in the ring of its cell)"); }
      ed[k] = ring_edge[o + q]; /* edge (v[k], v[k-1]): ring_nbr = the vertex before v in t */
      if (m->ring_nbr[o + q] != m->tri[3 * t + IM1[k]]) { free(ring_edge); return fo_fail(m, "Internal error! (ring pairing)"); }
    }
    for (k = 1; k < 3; k++) if (ed[k] < ed[kmin]) kmin = k;
    for (k = 0; k < 3; k++) {
      p->edgesOnVertex[3 * t + k] = ed[(kmin + k) % 3];
      p->cellsOnVertex[3 * t + k] = m->tri[3 * t + (kmin + k) % 3];
    }
    /* order consistency check, DV:570-581 */
    for (k = 0; k < 3; k++) {
      int ee = p->edgesOnVertex[3 * t + k], cc = p->cellsOnVertex[3 * t + k];
      if (p->verticesOnEdge[2 * ee] == t) { if (p->cellsOnEdge[2 * ee] != cc) { free(ring_edge); return fo_fail(m, "Internal error! (DV:576)"); } }
      else if (p->cellsOnEdge[2 * ee + 1] != cc) { free(ring_edge); return fo_fail(m, "Internal error! (DV:578)"); }
    }
  }
  free(ring_edge);

  /* edges, MM:244-294 */
  for (e = 0; e < nE; e++) {
    int c1 = p->cellsOnEdge[2 * e], c2 = p->cellsOnEdge[2 * e + 1];
    int v1 = p->verticesOnEdge[2 * e], v2 = p->verticesOnEdge[2 * e + 1];
    double A[3], B[3], Cc[3], D[3], X[3], xp[3], lon, lat, xlon;
    int i1 = -1, i2 = -1, d1 = p->nEdgesOnCell[c1], d2 = p->nEdgesOnCell[c2];
    A[0] = m->x[c1]; A[1] = m->y[c1]; A[2] = m->z[c1];
    B[0] = m->x[c2]; B[1] = m->y[c2]; B[2] = m->z[c2];
    Cc[0] = m->cx[v1]; Cc[1] = m->cy[v1]; Cc[2] = m->cz[v1];
    D[0] = m->cx[v2]; D[1] = m->cy[v2]; D[2] = m->cz[v2];
    if (!sg_intersect(A, B, Cc, D, X, &lon, &lat)) return fo_fail(m, "Unable to find intersection between DE and VE!");
    p->lonEdge[e] = lon; p->latEdge[e] = lat;
    p->xEdge[e] = X[0]; p->yEdge[e] = X[1]; p->zEdge[e] = X[2];
    p->dvEdge[e] = calc_arc_length(Cc, D);
    p->dcEdge[e] = calc_arc_length(A, B);
    p->dv1Edge[e] = calc_arc_length(Cc, X);
    p->dv2Edge[e] = calc_arc_length(D, X);
    xlon = m->lon[c1] + 0.1 * FO_PI;          /* MM:262-266 */
    if (xlon > FO_PI2) xlon -= FO_PI2;
    cartesian_transform(xlon, m->lat[c1], &xp[0], &xp[1], &xp[2]);
    p->angleEdge[e] = calc_sphere_angle(xp, A, B);
    p->nEdgesOnEdge[e] = d1 + d2 - 2;         /* MM:271 */
    j = 0;
    for (k = 0; k < d1; k++) if (p->edgesOnCell[(size_t)ME * c1 + k] == e) i1 = k; /* index_ptr, MM:274 */
    for (k = 0; k < d2; k++) if (p->edgesOnCell[(size_t)ME * c2 + k] == e) i2 = k;
    if (i1 < 0 || i2 < 0) return fo_fail(m, "Internal error! (edge not on its cells)");
    for (k = i1 + 1; k < d1; k++) p->edgesOnEdge[(size_t)ME2 * e + j++] = p->edgesOnCell[(size_t)ME * c1 + k];
    for (k = 0; k < i1; k++) p->edgesOnEdge[(size_t)ME2 * e + j++] = p->edgesOnCell[(size_t)ME * c1 + k];
    for (k = i2 + 1; k < d2; k++) p->edgesOnEdge[(size_t)ME2 * e + j++] = p->edgesOnCell[(size_t)ME * c2 + k];
    for (k = 0; k < i2; k++) p->edgesOnEdge[(size_t)ME2 * e + j++] = p->edgesOnCell[(size_t)ME * c2 + k];
  }

  /* areas, MM:296-341 (each total within 1e-10 of 4 pi) */
  {
    double total = 0.0, px[64], py[64], pz[64];
    if (ME > 64) return fo_fail(m, "cell degree > 64");
    for (i = 0; i < n; i++) {
      int d = p->nEdgesOnCell[i];
      for (k = 0; k < d; k++) {
        int vv = p->verticesOnCell[(size_t)ME * i + k];
        px[k] = m->cx[vv]; py[k] = m->cy[vv]; pz[k] = m->cz[vv];
      }
      p->areaCell[i] = fo_calc_area(d, px, py, pz);
      total += p->areaCell[i];
    }
    if (strict && fabs(total - 4 * FO_PI) > 1.0e-10) return fo_fail(m, "Too large total cell area error!");
    total = 0.0;
    for (t = 0; t < nT; t++) {
      for (k = 0; k < 3; k++) {
        int v = m->tri[3 * t + k];
        px[k] = m->x[v]; py[k] = m->y[v]; pz[k] = m->z[v];
      }
      p->areaTriangle[t] = fo_calc_area(3, px, py, pz);
      total += p->areaTriangle[t];
    }
    if (strict && fabs(total - 4 * FO_PI) > 1.0e-10) return fo_fail(m, "Too large total triangle area error!");
    total = 0.0;
    for (t = 0; t < nT; t++) {
      for (k = 0; k < 3; k++) {
        const int e1 = p->edgesOnVertex[3 * t + k], e2 = p->edgesOnVertex[3 * t + (k + 1) % 3];
        int vc = p->cellsOnVertex[3 * t + k];
        px[0] = p->xEdge[e1]; py[0] = p->yEdge[e1]; pz[0] = p->zEdge[e1];
        px[1] = m->x[vc]; py[1] = m->y[vc]; pz[1] = m->z[vc];
        px[2] = p->xEdge[e2]; py[2] = p->yEdge[e2]; pz[2] = p->zEdge[e2];
        px[3] = m->cx[t]; py[3] = m->cy[t]; pz[3] = m->cz[t];
        p->kiteAreasOnVertex[3 * t + k] = fo_calc_area(4, px, py, pz);
        total += p->kiteAreasOnVertex[3 * t + k];
      }
    }
    if (strict && fabs(total - 4 * FO_PI) > 1.0e-10) return fo_fail(m, "Too large total kite area error!");
  }

  /* calc_weightsOnEdge, MM:442-561 (weights start at zero: deviation D5) */
  {
    double *R = (double *)calloc((size_t)ME * n, sizeof(double));
    int *nrm = (int *)calloc((size_t)ME * n, sizeof(int)), *tng = (int *)calloc(3 * (size_t)nT, sizeof(int));
    int iVertex, iLocalCell, iLocalVertex, iLocalEdge, iCell, iEdge, bad = 0;
    for (iVertex = 0; iVertex < nT && !bad; iVertex++)          /* MM:458-468 */
      for (iLocalCell = 0; iLocalCell < 3; iLocalCell++) {
        iCell = p->cellsOnVertex[3 * iVertex + iLocalCell];
        for (iLocalVertex = 0; iLocalVertex < p->nEdgesOnCell[iCell]; iLocalVertex++)
          if (p->verticesOnCell[(size_t)ME * iCell + iLocalVertex] == iVertex) break;
        if (iLocalVertex >= p->nEdgesOnCell[iCell]) { bad = 1; break; }
        R[(size_t)ME * iCell + iLocalVertex] = p->kiteAreasOnVertex[3 * iVertex + iLocalCell] / p->areaCell[iCell];
      }
    for (iCell = 0; iCell < n && !bad; iCell++) {               /* MM:469-475 */
      double s = 0.0;
      for (k = 0; k < p->nEdgesOnCell[iCell]; k++) s += R[(size_t)ME * iCell + k];
      if (strict && fabs(s - 1.0) > 1.0e-10) bad = 2;
    }
    for (iCell = 0; iCell < n && !bad; iCell++)                 /* MM:477-489 */
      for (iLocalEdge = 0; iLocalEdge < p->nEdgesOnCell[iCell]; iLocalEdge++) {
        iEdge = p->edgesOnCell[(size_t)ME * iCell + iLocalEdge];
        if (iCell == p->cellsOnEdge[2 * iEdge]) nrm[(size_t)ME * iCell + iLocalEdge] = 1;
        else if (iCell == p->cellsOnEdge[2 * iEdge + 1]) nrm[(size_t)ME * iCell + iLocalEdge] = -1;
        else bad = 3;
      }
    for (iVertex = 0; iVertex < nT && !bad; iVertex++)          /* MM:491-503 */
      for (iLocalEdge = 0; iLocalEdge < 3; iLocalEdge++) {
        iEdge = p->edgesOnVertex[3 * iVertex + iLocalEdge];
        if (iVertex == p->verticesOnEdge[2 * iEdge]) tng[3 * iVertex + iLocalEdge] = 1;
        else if (iVertex == p->verticesOnEdge[2 * iEdge + 1]) tng[3 * iVertex + iLocalEdge] = -1;
        else bad = 4;
      }
    for (iEdge = 0; iEdge < nE && !bad; iEdge++) {              /* MM:505-556 */
      int iCell1 = p->cellsOnEdge[2 * iEdge], iCell2 = p->cellsOnEdge[2 * iEdge + 1];
      int n1 = p->nEdgesOnCell[iCell1], n2 = p->nEdgesOnCell[iCell2];
      int iLocalEdge1 = -1, iLocalEdge2 = -1, ii;
      for (k = 0; k < n1; k++) if (p->edgesOnCell[(size_t)ME * iCell1 + k] == iEdge) { iLocalEdge1 = k + 1; break; }
      for (k = 0; k < n2; k++) if (p->edgesOnCell[(size_t)ME * iCell2 + k] == iEdge) { iLocalEdge2 = k + 1; break; }
      if (iLocalEdge1 < 0 || iLocalEdge2 < 0) { bad = 5; break; }
      for (ii = 1; ii <= p->nEdgesOnEdge[iEdge]; ii++) {          /* 1-based local indices, as MM */
        int iEdgeOnEdge = p->edgesOnEdge[(size_t)ME2 * iEdge + ii - 1];
        int nEdgesOnCell, jEnd, iLocalEdgeOnVertex;
        double w = 0.0;
        iVertex = -1;
        if (p->cellsOnEdge[2 * iEdgeOnEdge] == iCell1 || p->cellsOnEdge[2 * iEdgeOnEdge + 1] == iCell1) {
          iCell = iCell1; nEdgesOnCell = n1; jEnd = n1 - ii; iLocalEdge = iLocalEdge1 + ii;
        } else if (p->cellsOnEdge[2 * iEdgeOnEdge] == iCell2 || p->cellsOnEdge[2 * iEdgeOnEdge + 1] == iCell2) {
          iCell = iCell2; nEdgesOnCell = n2; jEnd = n2 - ii + n1 - 1; iLocalEdge = iLocalEdge2 + ii - n1 + 1;
        } else { bad = 6; break; }
        if (iLocalEdge > nEdgesOnCell) iLocalEdge -= nEdgesOnCell;
        iLocalVertex = iLocalEdge;
        for (j = 1; j <= jEnd; j++) {
          iLocalVertex++;
          if (iLocalVertex > nEdgesOnCell) iLocalVertex = 1;
          iVertex = p->verticesOnCell[(size_t)ME * iCell + iLocalVertex - 1];
          w += R[(size_t)ME * iCell + iLocalVertex - 1];
        }
        if (iVertex < 0) { bad = 7; break; }
        for (iLocalEdgeOnVertex = 0; iLocalEdgeOnVertex < 3; iLocalEdgeOnVertex++)
          if (p->edgesOnVertex[3 * iVertex + iLocalEdgeOnVertex] == iEdge) break;
        if (iLocalEdgeOnVertex >= 3) { bad = 8; break; }
        w = (double)nrm[(size_t)ME * iCell + iLocalEdge - 1] / (double)tng[3 * iVertex + iLocalEdgeOnVertex] * (w - 0.5);
        w = w * p->dvEdge[iEdgeOnEdge] / p->dcEdge[iEdge];
        p->weightsOnEdge[(size_t)ME2 * iEdge + ii - 1] = w;
      }
    }
    free(R); free(nrm); free(tng);
    if (bad) { snprintf(m->err, sizeof m->err, "Internal error! (calc_weightsOnEdge check %d)", bad); return 1; }
  }
  return 0;
}

/* getters: ids 1-based (unused slots 0); any pointer may be NULL */
static void out_ids(int *dst, const int *src, size_t cnt) {
  size_t i;
  if (dst) for (i = 0; i < cnt; i++) dst[i] = src[i] + 1;
}
static void out_dbl(double *dst, const double *src, size_t cnt) {
  if (dst) memcpy(dst, src, sizeof(double) * cnt);
}
int fo_mpas_get_edges(fo_mesh *m, int *cellsOnEdge, int *verticesOnEdge, double *xyz, double *lonlat,
                      double *dv, double *dc, double *dv1, double *dv2, double *angle) {
  fo_mpas_t *p = m->mp;
  int e;
  if (!p) return fo_fail(m, "fo_mpas_get_edges before fo_mpas");
  out_ids(cellsOnEdge, p->cellsOnEdge, 2 * (size_t)p->nEdges);
  out_ids(verticesOnEdge, p->verticesOnEdge, 2 * (size_t)p->nEdges);
  for (e = 0; e < p->nEdges; e++) {
    if (xyz) { xyz[3 * e] = p->xEdge[e]; xyz[3 * e + 1] = p->yEdge[e]; xyz[3 * e + 2] = p->zEdge[e]; }
    if (lonlat) { lonlat[2 * e] = p->lonEdge[e]; lonlat[2 * e + 1] = p->latEdge[e]; }
  }
  out_dbl(dv, p->dvEdge, p->nEdges); out_dbl(dc, p->dcEdge, p->nEdges);
  out_dbl(dv1, p->dv1Edge, p->nEdges); out_dbl(dv2, p->dv2Edge, p->nEdges);
  out_dbl(angle, p->angleEdge, p->nEdges);
  return 0;
}
int fo_mpas_get_cells(fo_mesh *m, int *nEdgesOnCell, int *verticesOnCell, int *cellsOnCell, int *edgesOnCell,
                      double *areaCell) {
  fo_mpas_t *p = m->mp;
  size_t cnt;
  if (!p) return fo_fail(m, "fo_mpas_get_cells before fo_mpas");
  cnt = (size_t)p->maxEdges * p->nCells;
  if (nEdgesOnCell) memcpy(nEdgesOnCell, p->nEdgesOnCell, sizeof(int) * p->nCells);
  out_ids(verticesOnCell, p->verticesOnCell, cnt);
  out_ids(cellsOnCell, p->cellsOnCell, cnt);
  out_ids(edgesOnCell, p->edgesOnCell, cnt);
  out_dbl(areaCell, p->areaCell, p->nCells);
  return 0;
}
int fo_mpas_get_vertices(fo_mesh *m, int *cellsOnVertex, int *edgesOnVertex, double *areaTriangle, double *kite) {
  fo_mpas_t *p = m->mp;
  if (!p) return fo_fail(m, "fo_mpas_get_vertices before fo_mpas");
  out_ids(cellsOnVertex, p->cellsOnVertex, 3 * (size_t)p->nVertices);
  out_ids(edgesOnVertex, p->edgesOnVertex, 3 * (size_t)p->nVertices);
  out_dbl(areaTriangle, p->areaTriangle, p->nVertices);
  out_dbl(kite, p->kiteAreasOnVertex, 3 * (size_t)p->nVertices);
  return 0;
}
int fo_mpas_get_edges_on_edge(fo_mesh *m, int *nEdgesOnEdge, int *edgesOnEdge, double *weightsOnEdge) {
  fo_mpas_t *p = m->mp;
  size_t cnt;
  if (!p) return fo_fail(m, "fo_mpas_get_edges_on_edge before fo_mpas");
  cnt = (size_t)p->maxEdges2 * p->nEdges;
  if (nEdgesOnEdge) memcpy(nEdgesOnEdge, p->nEdgesOnEdge, sizeof(int) * p->nEdges);
  out_ids(edgesOnEdge, p->edgesOnEdge, cnt);
  out_dbl(weightsOnEdge, p->weightsOnEdge, cnt);
  return 0;
}
