// star_kernels.cuh -- cold build "by stars": the Delaunay triangulation of well-spaced generators
// WITHOUT insertion rounds and flips.
//
// What it replaces: delaunay_triangulation (/root/reference/src/delaunay_voronoi_mod.F90:255-321)
// inserts the points one by one (find the containing triangle, split 1 -> 3, validate_DT / flip22
// DV:845-882); the insertion-level build of dt_driver.h does the same in parallel rounds and is
// bound by the latency of ~3 flips per point of ~14 scattered accesses each in ~60 sub-rounds of
// four launches.  For a point set whose spacing does not vary wildly (the meshes this code is
// for: Fibonacci / icosahedral / relaxed SCVT generators, graded or not) every
// generator's STAR -- its Delaunay neighbours in counter-clockwise order -- can be found
// independently from the few dozen points around it:
//
//   1. bin the generators into a uniform 3-D grid over [-1,1]^3 (cell = 2 mean spacings; only the cells the
//      sphere passes through are occupied), sort by cell, tables of column starts / column halves;
//   2. one thread per generator g, in sorted order: the candidates are the points of the (2r+1)^3 block of cells
//      around g's cell, which holds EVERY point within r cells of g.  The nearest candidate p1 is a
//      Delaunay neighbour; from the Delaunay edge (g, p_k) the next neighbour p_k+1 is the
//      candidate q on the left of g -> p_k whose circle (g, p_k, q) holds no other candidate on the
//      left (gift wrapping) until the ring closes at p1.  KStarWrapT::wrap does exactly that over all candidates
//      with the filtered (or, EXACT, the full: expansion arithmetic + simulation of simplicity, the flips' own)
//      predicates; KStarWrapT::wrap_near, the form nearly every generator goes through, does it on a short list
//      of the candidates near g kept in shared memory as fp32 offsets: the next neighbour is guessed in fp32 and
//      the guess certified in fp32, then fp64, then -- rarely -- by the exact kernel (see there).  The star is
//      final when every circumcircle found lies inside the searched block: no unseen point can be inside.
//      Otherwise: a longer list, a wider block (r + 1, compact retry lists), the full wrap, the exact kernel;
//   3. triangle (g, p_k, p_k+1) belongs to its smallest vertex: mask + count per generator (written by the wrap),
//      one scan gives the slots -- already in the order sort_triangles would produce -- and the rows
//      (vertices + adjacency) are written by looking the neighbours up in the rings.
//
// The result is checked, not trusted: every triangle of every star must be found in the star of
// its owner, the counts must be T = 2n - 4 triangles and 3T ring entries (=> a closed surface of
// Euler characteristic 2), and the warm sweep (orientation + local Delaunay test of every edge)
// then runs over the rows; anything else -- a star that does not close within kStarMaxR cells or
// kStarCap neighbours (lat-lon grids at the poles, clustered points) -- and the build falls back to
// the insertion-level path; a sampling pass sends sets the stars are slow at (uniform random points) there
// before any work is done.  Both give the Delaunay triangulation of the SoS-perturbed points,
// which is unique: same triangles, bit for bit.
#pragma once
#include "dt_kernels.cuh"

namespace fm {

static const int kStarCap = 24;   // ring slots per generator (fixed stride: 96 bytes = 3 sectors)
static const int kStarMaxR = 4;   // widest block searched: (2 * 4 + 1)^3 cells
static const int kStarMaxG = 1600;  // cells per axis: G^3 < 2^32 (32-bit cell keys)
#ifndef FM_STAR_LIST
#define FM_STAR_LIST 16
#endif
static const int kStarList = FM_STAR_LIST;  // near candidates per generator kept by the filter kernels (wrap_near), regular sets
static const int kStarMaxCandidates = 1024;  // candidates in a block beyond which wrap_near gives up (8x: the full wrap)
static const int kStarListLong = 96;        // ... for the generators whose list overflowed (KStarSlowNear)
static const int kStarListWide = 40;       // ... in the wide mode (irregular sets: uniform random points)
enum { FM_S_OK = 0, FM_S_RADIUS = 1, FM_S_UNSURE = 2, FM_S_FAIL = 3, FM_S_SLOW = 4 };
// slots of StarCtx::sc
enum { FM_SC_LIST0 = 0, FM_SC_LIST1 = 1, FM_SC_EXACT = 2, FM_SC_FAIL = 3, FM_SC_BROKEN = 4, FM_SC_RETRIED = 5, FM_SC_SLOW = 6, FM_SC_MULTI = 7 };

struct StarCtx {
  int n, G;
  double half_g;         // G / 2: cell index = floor((x + 1) * half_g)
  double cell;           // cell size 2 / G
  const P4* pts;         // generators, internal order
  unsigned* key;         // per generator: cell key (ix * G + iy) * G + iz   [KStarKey]
  int* idx;              //                identity, the sort's payload
  const unsigned* skey;  // sorted keys
  const int* sidx;       // sorted position -> internal index
  P4* spts;              // generator records in sorted order (id = the original id, the SoS key)
  int* colstart;         // G * G + 1: first sorted position of column (ix, iy)
  int* colmid;           // G * G: first sorted position of the column's upper half (iz >= G / 2)
  int* ring;             // n * kStarCap: the star of internal index i, CCW, internal indices
  int* deg;              // n: its length
  unsigned* own;         // n: bit k = triangle (i, ring[k], ring[k+1]) has i as its smallest vertex
  u64* off;              // n + 1: exclusive scan of (owned count | degree << 32)
  int* stri;             // n * kStarCap: slot of triangle k of the star
  int *list_in, *list_out, *list_exact;   // sorted positions to retry with a wider block / exactly
  int* list_slow;        // ... to redo with the full wrap (more near candidates than wrap_near's list holds)
  int* sc;               // counters, FM_SC_*
  int r;                 // block radius of this pass (cells)
  int dry;               // sampling pass: statuses are only counted (sc[FM_SC_RETRIED] not OK, sc[FM_SC_MULTI] OK after >= 3 rounds)
  int list_cap;          // capacity of wrap_near's near list (kStarList, or kStarListWide in the wide mode)
  double near2;          // scale of the first near-list radius, squared (1 = ~7.5 expected entries)
  int in_idx, out_idx;   // which of sc[0], sc[1] counts list_in / list_out
  Row* row;
  int* v2t;
};

FM_HD inline int star_cell1(const StarCtx& s, double x) {
  int i = (int)((x + 1.0) * s.half_g);
  return i < 0 ? 0 : (i >= s.G ? s.G - 1 : i);
}
FM_HD inline unsigned star_key(const StarCtx& s, int ix, int iy, int iz) {
  return ((unsigned)ix * (unsigned)s.G + (unsigned)iy) * (unsigned)s.G + (unsigned)iz;
}

struct KStarKey {
  FM_HD static void run(const StarCtx& s, int i) {
    P4 p = s.pts[i];
    s.key[i] = star_key(s, star_cell1(s, p.x), star_cell1(s, p.y), star_cell1(s, p.z));
    s.idx[i] = i;
  }
};
struct KStarGather {
  FM_HD static void run(const StarCtx& s, int j) { s.spts[j] = s.pts[s.sidx[j]]; }
};
// first sorted position whose key is >= k, in [lo, hi)
FM_HD inline int star_lower(const unsigned* key, int lo, int hi, unsigned k) {
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (key[mid] < k) lo = mid + 1; else hi = mid;
  }
  return lo;
}
struct KStarCols {   // over G * G + 1 columns
  FM_HD static void run(const StarCtx& s, int col) {
    s.colstart[col] = col == s.G * s.G ? s.n : star_lower(s.skey, 0, s.n, (unsigned)col * (unsigned)s.G);
    if (col < s.G * s.G) s.colmid[col] = star_lower(s.skey, 0, s.n, (unsigned)col * (unsigned)s.G + (unsigned)(s.G / 2));
  }
};

// ------------------------------------------------------------------ the star of one generator
// Filter constants: both determinants below are sums of three products of a difference (w) with a
// 2x2 minor (m or n) of differences; each minor carries <= 3 roundings relative to the sum of the
// magnitudes of its two products (mabs / nabs), each term of the dot product two more, the two
// additions two more: 7 * 2^-53 < 8e-16 of sum |w_i| * minor_abs_i.  2e-15 keeps a margin of 2.5.
#define FM_STAR_ERR 2.0e-15

#define FM_STAR_FOR(q)                        \
  for (int a_ = 0; a_ < nr; a_++)             \
    for (int q = rng[(2 * a_) * stride], e_ = rng[(2 * a_ + 1) * stride]; q < e_; q++)

// A star is complete: its length, which of its triangles the generator owns (it is their smallest vertex), and the
// (owned count | degree << 32) word the scan turns into slot offsets.  The ring was just written by this thread.
FM_HD inline int star_popc(unsigned v) {
#if defined(__CUDA_ARCH__)
  return __popc(v);
#else
  return __builtin_popcount(v);
#endif
}
FM_HD inline void star_finish(const StarCtx& s, int i, int d) {
  const int* ring = s.ring + (size_t)i * kStarCap;
  unsigned mask = 0;
  int prev = ring[0];
  for (int k = 0; k < d; k++) {
    const int nxt = ring[k + 1 == d ? 0 : k + 1];
    if (i < prev && i < nxt) mask |= 1u << k;
    prev = nxt;
  }
  s.deg[i] = d;
  s.own[i] = mask;
  s.off[i] = (u64)star_popc(mask) | ((u64)(unsigned)d << 32);
}

template <bool EXACT>
struct KStarWrapT {
  // The star of the generator at sorted position j, searched in the block of radius r cells.
  // rng[k * stride]: scratch for 2 (2r+1)^2 ints (shared memory on the device).
  FM_HD static int wrap(const StarCtx& s, int j, int r, int* rng, int stride) {
    const P4 gq = s.spts[j];
    const P3 g{gq.x, gq.y, gq.z};
    const int idg = (int)gq.id;
    const int ix = star_cell1(s, g.x), iy = star_cell1(s, g.y), iz = star_cell1(s, g.z);
    const int zlo = iz - r < 0 ? 0 : iz - r, zhi = iz + r >= s.G ? s.G - 1 : iz + r;
    int nr = 0, total = 0;
    for (int X = ix - r; X <= ix + r; X++) {
      if (X < 0 || X >= s.G) continue;
      for (int Y = iy - r; Y <= iy + r; Y++) {
        if (Y < 0 || Y >= s.G) continue;
        const int col = X * s.G + Y;
        int lo = s.colstart[col], hi = s.colstart[col + 1];
        if (lo == hi) continue;
        const unsigned base = (unsigned)col * (unsigned)s.G;
        if (hi - lo > 4) {
          lo = star_lower(s.skey, lo, hi, base + (unsigned)zlo);
          hi = star_lower(s.skey, lo, hi, base + (unsigned)zhi + 1u);
        } else {   // a short column: walk in from both ends
          while (lo < hi && s.skey[lo] < base + (unsigned)zlo) lo++;
          while (lo < hi && s.skey[hi - 1] > base + (unsigned)zhi) hi--;
        }
        if (lo < hi) { rng[(2 * nr) * stride] = lo; rng[(2 * nr + 1) * stride] = hi; nr++; total += hi - lo; }
      }
    }
    if (total > 8 * kStarMaxCandidates) return FM_S_FAIL;   // thousands of generators in a few cells: a cluster, the insertion build's business
    // the nearest candidate: a Delaunay neighbour (the circle on the diameter g p1 is empty), provided
    // it lies within the distance the block covers completely
    double bd = 1.0e300;
    int p1 = -1;
    FM_STAR_FOR(q) {
      if (q == j) continue;
      const P4 Q = s.spts[q];
      const double dx = Q.x - g.x, dy = Q.y - g.y, dz = Q.z - g.z;
      const double d2 = dx * dx + dy * dy + dz * dz;
      if (d2 < bd) { bd = d2; p1 = q; }
    }
    // The searched block is the box [lo, hi] (cell boundaries; a point strictly inside it, by delta, is
    // binned into one of its cells).  p1 is the true nearest neighbour when it is closer than the box's faces.
    const double delta = 1.0e-9 * s.cell;
    const double lox = (ix - r) * s.cell - 1.0 + delta, hix = (ix + r + 1) * s.cell - 1.0 - delta;
    const double loy = (iy - r) * s.cell - 1.0 + delta, hiy = (iy + r + 1) * s.cell - 1.0 - delta;
    const double loz = (iz - r) * s.cell - 1.0 + delta, hiz = (iz + r + 1) * s.cell - 1.0 - delta;
    {
      const double dmin = fmin(fmin(fmin(g.x - lox, hix - g.x), fmin(g.y - loy, hiy - g.y)), fmin(g.z - loz, hiz - g.z));
      if (p1 < 0 || !(dmin > 0.0) || bd > dmin * dmin) return FM_S_RADIUS;
    }
    const int gi = s.sidx[j];
    int* ring = s.ring + (size_t)gi * kStarCap;
    int pk = p1, d = 0;
    for (;;) {
      if (d == kStarCap) return FM_S_FAIL;
      ring[d++] = s.sidx[pk];
      const P4 Pk = s.spts[pk];
      const P3 pkp{Pk.x, Pk.y, Pk.z};
      const double ux = Pk.x - g.x, uy = Pk.y - g.y, uz = Pk.z - g.z;
      // orient(g, pk, q) = det[g, pk, q] = (q - g) . (g x (pk - g)) = w . m
      const double mx = g.y * uz - g.z * uy, my = g.z * ux - g.x * uz, mz = g.x * uy - g.y * ux;
      const double max_ = fabs(g.y * uz) + fabs(g.z * uy), may = fabs(g.z * ux) + fabs(g.x * uz),
                   maz = fabs(g.x * uy) + fabs(g.y * ux);
      // phase 1: any candidate certainly on the left of g -> pk
      int best = -1;
      P4 Bq = Pk;
      for (int a_ = 0; a_ < nr && best < 0; a_++)
        for (int q = rng[(2 * a_) * stride], e_ = rng[(2 * a_ + 1) * stride]; q < e_; q++) {
          if (q == j || q == pk) continue;
          const P4 Q = s.spts[q];
          bool left;
          if (EXACT) {
            left = orient(g, pkp, P3{Q.x, Q.y, Q.z}, idg, (int)Pk.id, (int)Q.id) > 0;
          } else {
            const double wx = Q.x - g.x, wy = Q.y - g.y, wz = Q.z - g.z;
            const double det = wx * mx + wy * my + wz * mz;
            left = det > FM_STAR_ERR * (fabs(wx) * max_ + fabs(wy) * may + fabs(wz) * maz);
          }
          if (left) { best = q; Bq = Q; break; }
        }
      if (best < 0) return FM_S_RADIUS;
      double vx = Bq.x - g.x, vy = Bq.y - g.y, vz = Bq.z - g.z;
      double nx = uy * vz - uz * vy, ny = uz * vx - ux * vz, nz = ux * vy - uy * vx;
      double nax = fabs(uy * vz) + fabs(uz * vy), nay = fabs(uz * vx) + fabs(ux * vz), naz = fabs(ux * vy) + fabs(uy * vx);
      // phase 2: a candidate inside the circle (g, pk, best) AND on the left replaces best; the part of the circle
      // on the left only shrinks, so what was outside it or on the right stays so
      FM_STAR_FOR(q) {
        if (q == j || q == pk || q == best) continue;
        const P4 Q = s.spts[q];
        bool take;
        if (EXACT) {
          const P3 qp{Q.x, Q.y, Q.z};
          take = incircle(g, pkp, P3{Bq.x, Bq.y, Bq.z}, qp, idg, (int)Pk.id, (int)Bq.id, (int)Q.id) > 0 &&
                 orient(g, pkp, qp, idg, (int)Pk.id, (int)Q.id) > 0;
        } else {
          const double wx = Q.x - g.x, wy = Q.y - g.y, wz = Q.z - g.z;
          const double det = wx * nx + wy * ny + wz * nz;
          const double bound = FM_STAR_ERR * (fabs(wx) * nax + fabs(wy) * nay + fabs(wz) * naz);
          if (det < -bound) continue;            // certainly outside
          if (!(det > bound)) return FM_S_UNSURE;
          const double od = wx * mx + wy * my + wz * mz;
          const double ob = FM_STAR_ERR * (fabs(wx) * max_ + fabs(wy) * may + fabs(wz) * maz);
          if (od < -ob) continue;                // inside, but on the right
          if (!(od > ob)) return FM_S_UNSURE;
          take = true;
        }
        if (take) {
          best = q; Bq = Q;
          vx = Q.x - g.x; vy = Q.y - g.y; vz = Q.z - g.z;
          nx = uy * vz - uz * vy; ny = uz * vx - ux * vz; nz = ux * vy - uy * vx;
          nax = fabs(uy * vz) + fabs(uz * vy); nay = fabs(uz * vx) + fabs(ux * vz); naz = fabs(ux * vy) + fabs(uy * vx);
        }
      }
      // Is the circle of (g, pk, best) inside the searched block?  Every sphere point inside the circle lies in the
      // ball around the circle's centre in its plane, P = n (n.g) / |n|^2, of the circle's radius rho = |g - P|:
      // the star is final when that ball is inside the box for all its triangles (no unseen point can be
      // inside a circle).  A sliver's normal is not trusted.
      {
        const double ng = nx * g.x + ny * g.y + nz * g.z, nn = nx * nx + ny * ny + nz * nz;
        const double uu = ux * ux + uy * uy + uz * uz, vv = vx * vx + vy * vy + vz * vz;
        if (!(ng > 0.0) || nn < 1.0e-12 * uu * vv) return FM_S_RADIUS;
        const double f = ng / nn;
        const double Px = nx * f, Py = ny * f, Pz = nz * f;
        const double ex = g.x - Px, ey = g.y - Py, ez = g.z - Pz;
        const double rho2 = (ex * ex + ey * ey + ez * ez) * (1.0 + 1.0e-6);
        const double a0 = Px - lox, a1 = hix - Px, b0 = Py - loy, b1 = hiy - Py, c0 = Pz - loz, c1 = hiz - Pz;
        const double m = fmin(fmin(fmin(a0, a1), fmin(b0, b1)), fmin(c0, c1));
        if (!(m > 0.0) || m * m < rho2) return FM_S_RADIUS;
      }
      if (best == p1) break;
      pk = best;
    }
    star_finish(s, gi, d);
    return FM_S_OK;
  }
  // The filter kernels' form.  The wrap above visits every candidate of the block in every step -- a 32-byte
  // load and ~25 fp64 instructions per visit, in loops whose trip counts and branches differ from lane to
  // lane.  ncu on it, and on two rearrangements of it, as the main kernel at 2.6M generators: 7 to 16 of 32
  // lanes active, and the fp64 pipe (64 lanes per clock and SM) the floor at ~2000 fp64 instructions per
  // generator.  This form does the geometry of the wrap in FP32 on a short list kept in shared memory:
  //   round   one flat loop over the block's candidates (fp64): w = q - g, distance, nearest neighbour; those
  //           within R of g go on the list as (sorted position, (float) w);
  //   step    from the Delaunay edge (g, pk): the next neighbour is GUESSED as the entry on the left with the
  //           steepest plane through the axis g pk (fp32, selects), and then CERTIFIED: every other entry must
  //           be outside the circle (g, pk, best) or on the right of g -> pk, best itself on the left.  The
  //           certificate is tried in fp32 first (inputs rounded to 2^-24, error bound 1.2e-6 of the
  //           permanent: conclusive unless the entry is within ~1e-3 of the typical distance from the
  //           circle), and the few entries it leaves open are retested in fp64 with the bound of the filters
  //           above; what fp64 leaves open sends the generator to the exact kernel.
  // The star is final when all its circumcircles lie inside the box AND within R of g (radius rho, through g:
  // within 2 rho of g; fp64) -- no candidate off the list and no point outside the block can be inside one.
  // Otherwise R grows towards what the circles found ask for and the round is repeated.
  // lst[(4 k + c) * stride]: s.list_cap entries of (position, wx, wy, wz), the last three float bit patterns.
#define FM_STAR_ERR32 1.2e-6f
#ifndef FM_STAR_PREFETCH
#define FM_STAR_PREFETCH 1
#endif
#ifndef FM_STAR_CONSTB
#define FM_STAR_CONSTB 1
#endif
#ifndef FM_STAR_TRIM
#define FM_STAR_TRIM 0   // A/B at 2.6M: counting the keys outside the wanted cells with 12 independent loads, star_wrap 2.26 -> 2.40 ms
#endif
#ifndef FM_STAR_PIPE
#define FM_STAR_PIPE 1   // A/B at 2.6M: star_wrap 2.40 -> 2.26 ms
#endif
  FM_HD static float as_float(int v) {
#if defined(__CUDA_ARCH__)
    return __int_as_float(v);
#else
    float f; memcpy(&f, &v, 4); return f;
#endif
  }
  FM_HD static int as_int(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_int(f);
#else
    int v; memcpy(&v, &f, 4); return v;
#endif
  }
  FM_HD static int wrap_near(const StarCtx& s, int j, int r, int* rng, int stride, int* lst) {
    const P4 gq = s.spts[j];
    const P3 g{gq.x, gq.y, gq.z};
    const float gfx = (float)g.x, gfy = (float)g.y, gfz = (float)g.z;
    const int ix = star_cell1(s, g.x), iy = star_cell1(s, g.y), iz = star_cell1(s, g.z);
    const int zlo = iz - r < 0 ? 0 : iz - r, zhi = iz + r >= s.G ? s.G - 1 : iz + r;
    int nr = 0, total = 0;
    for (int X = ix - r; X <= ix + r; X++) {
      if (X < 0 || X >= s.G) continue;
      for (int Y = iy - r; Y <= iy + r; Y++) {
        if (Y < 0 || Y >= s.G) continue;
        const int col = X * s.G + Y;
        int lo = s.colstart[col], hi = s.colstart[col + 1];
        if (lo == hi) continue;
        const unsigned base = (unsigned)col * (unsigned)s.G;
        // a column pierces the shell twice: away from the equator of the z axis the wanted cells lie in ONE of its
        // halves, and most of that half is wanted -- take the half from the table, trim its ends
        if (zlo >= s.G / 2) lo = s.colmid[col];
        else if (zhi < s.G / 2) hi = s.colmid[col];
        if (hi - lo > 12) {
          lo = star_lower(s.skey, lo, hi, base + (unsigned)zlo);
          hi = star_lower(s.skey, lo, hi, base + (unsigned)zhi + 1u);
        } else {
#if FM_STAR_TRIM
          // the keys are sorted: how many lie below / above the wanted cells -- twelve independent loads instead of
          // a walk in from both ends, one dependent load per step
          int below = 0, above = 0;
          const unsigned klo = base + (unsigned)zlo, khi = base + (unsigned)zhi;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
          for (int t = 0; t < 12; t++) {
            const bool valid = lo + t < hi;
            const unsigned key = valid ? s.skey[lo + t] : klo;
            below += (valid && key < klo) ? 1 : 0;
            above += (valid && key > khi) ? 1 : 0;
          }
          lo += below; hi -= above;
#else
          while (lo < hi && s.skey[lo] < base + (unsigned)zlo) lo++;
          while (lo < hi && s.skey[hi - 1] > base + (unsigned)zhi) hi--;
#endif
        }
        if (lo < hi) {
          rng[(2 * nr) * stride] = lo; rng[(2 * nr + 1) * stride] = hi; nr++; total += hi - lo;
#if defined(__CUDA_ARCH__) && FM_STAR_PREFETCH
          for (int q = lo & ~3; q < hi; q += 4) asm volatile("prefetch.global.L1 [%0];" ::"l"(s.spts + q));   // 128-byte lines
#endif
        }
      }
    }
    if (total > kStarMaxCandidates) return FM_S_SLOW;   // a clustered cell: not for the list form (and never in a loop of six rounds)
    const double delta = 1.0e-9 * s.cell;
    const double lox = (ix - r) * s.cell - 1.0 + delta, hix = (ix + r + 1) * s.cell - 1.0 - delta;
    const double loy = (iy - r) * s.cell - 1.0 + delta, hiy = (iy + r + 1) * s.cell - 1.0 - delta;
    const double loz = (iz - r) * s.cell - 1.0 + delta, hiz = (iz + r + 1) * s.cell - 1.0 - delta;
    const double dmin = fmin(fmin(fmin(g.x - lox, hix - g.x), fmin(g.y - loy, hiy - g.y)), fmin(g.z - loz, hiz - g.z));
    const int gi = s.sidx[j];
    int* ring = s.ring + (size_t)gi * kStarCap;
    // first radius: the one expected to put ~7.5 candidates on the list at the LOCAL density (the block, of area
    // ~ ((2r+1) cell)^2 on the sphere, holds `total` generators) -- 1.55 mean spacings on a uniform mesh; on a
    // graded mesh it follows the spacing.  near2 scales it (FM_STAR_NEAR).
    double R2 = s.near2 * 2.39 * ((2 * r + 1) * s.cell) * ((2 * r + 1) * s.cell) / (total > 1 ? total : 1);
    for (int round = 0; round < 6; round++) {
      int m = 0, over = 0, k1 = -1;
      double bd = 1.0e300;
      {
        int a_ = 0, q = nr ? rng[0] : 0, e_ = nr ? rng[stride] : 0;
#if FM_STAR_PIPE
        // the next candidate's record is requested before this one is worked on
        int qn = q;
        P4 Qn = total > 0 ? s.spts[qn] : gq;
        for (int k = 0; k < total; k++) {
          const P4 Q = Qn;
          q = qn;
          qn++;
          if (qn == e_ && a_ + 1 < nr) { a_++; qn = rng[(2 * a_) * stride]; e_ = rng[(2 * a_ + 1) * stride]; }
          if (k + 1 < total) Qn = s.spts[qn];
#else
#if defined(__CUDA_ARCH__)
#pragma unroll 4
#endif
        for (int k = 0; k < total; k++) {
          const P4 Q = s.spts[q];
#endif
          const double dx = Q.x - g.x, dy = Q.y - g.y, dz = Q.z - g.z;
          const double d2 = dx * dx + dy * dy + dz * dz;
          const bool other = q != j;
          const bool in = other && d2 <= R2, room = m < s.list_cap;
          const bool nearer = in && room && d2 < bd;   // (the nearest neighbour is wanted only when it is on the list)
          bd = (other && d2 < bd) ? d2 : bd;
          k1 = nearer ? m : k1;
          if (in && room) {
            lst[(4 * m) * stride] = q;
            lst[(4 * m + 1) * stride] = as_int((float)dx);
            lst[(4 * m + 2) * stride] = as_int((float)dy);
            lst[(4 * m + 3) * stride] = as_int((float)dz);
          }
          over |= (in && !room) ? 1 : 0;
          m += (in && room) ? 1 : 0;
#if !FM_STAR_PIPE
          q++;
          if (q == e_ && a_ + 1 < nr) { a_++; q = rng[(2 * a_) * stride]; e_ = rng[(2 * a_ + 1) * stride]; }
#endif
        }
      }
      if (round == 0 && (total < 2 || !(dmin > 0.0) || bd > dmin * dmin)) return FM_S_RADIUS;
      if (over) return FM_S_SLOW;     // more near candidates than the list holds: the full wrap
      const bool complete = (m == total - 1);
      const float wmaxf = sqrtf((float)R2) * 1.000001f;
      bool grow = bd > R2 || k1 < 0, inbox = true;   // (not even the nearest neighbour on the list: grow)
      int unsure = 0;
      double need2 = 0.0;
      int kk = k1, d = 0;
      P4 Pk = gq;
      if (!grow) Pk = s.spts[lst[(4 * k1) * stride]];
      while (!grow) {
        if (d == kStarCap) return FM_S_FAIL;
        const int pk = lst[(4 * kk) * stride];
        ring[d++] = s.sidx[pk];
        const float ux = as_float(lst[(4 * kk + 1) * stride]), uy = as_float(lst[(4 * kk + 2) * stride]),
                    uz = as_float(lst[(4 * kk + 3) * stride]);
        // m = g x u: w . m = det[g, pk, q] > 0 on the left of g -> pk;  e = u x m completes the frame about the axis u
        const float mx = gfy * uz - gfz * uy, my = gfz * ux - gfx * uz, mz = gfx * uy - gfy * ux;
        const float ex = uy * mz - uz * my, ey = uz * mx - ux * mz, ez = ux * my - uy * mx;
        int kb = -1;
        float ba = 1.0f, bb = -3.0e38f;
#if defined(__CUDA_ARCH__)
#pragma unroll 4
#endif
        for (int k = 0; k < m; k++) {
          const float wx = as_float(lst[(4 * k + 1) * stride]), wy = as_float(lst[(4 * k + 2) * stride]),
                      wz = as_float(lst[(4 * k + 3) * stride]);
          const float a = wx * mx + wy * my + wz * mz, b = wx * ex + wy * ey + wz * ez;
          const bool better = k != kk && a > 0.0f && b * ba > bb * a;
          kb = better ? k : kb;
          ba = better ? a : ba;
          bb = better ? b : bb;
        }
        if (kb < 0) { grow = true; need2 = 4.0 * R2; break; }   // nobody on the left within R
        const float vx = as_float(lst[(4 * kb + 1) * stride]), vy = as_float(lst[(4 * kb + 2) * stride]),
                    vz = as_float(lst[(4 * kb + 3) * stride]);
        const float nx = uy * vz - uz * vy, ny = uz * vx - ux * vz, nz = ux * vy - uy * vx;
        const float nax = fabsf(uy * vz) + fabsf(uz * vy), nay = fabsf(uz * vx) + fabsf(ux * vz), naz = fabsf(ux * vy) + fabsf(uy * vx);
        const float max_ = fabsf(gfy * uz) + fabsf(gfz * uy), may = fabsf(gfz * ux) + fabsf(gfx * uz),
                    maz = fabsf(gfx * uy) + fabsf(gfy * ux);
        // the certificate in fp32: entries it leaves open are counted
        int open = 0;
        {
          const float od = vx * mx + vy * my + vz * mz;
          open += !(od > FM_STAR_ERR32 * (fabsf(vx) * max_ + fabsf(vy) * may + fabsf(vz) * maz)) ? 1 : 0;
        }
#if FM_STAR_CONSTB
        // one bound per step instead of one per entry: |w_i| <= |w| <= R for every entry of the list
        const float cbn = FM_STAR_ERR32 * (nax + nay + naz) * wmaxf, cbm = FM_STAR_ERR32 * (max_ + may + maz) * wmaxf;
#endif
#if defined(__CUDA_ARCH__)
#pragma unroll 4
#endif
        for (int k = 0; k < m; k++) {
          const float wx = as_float(lst[(4 * k + 1) * stride]), wy = as_float(lst[(4 * k + 2) * stride]),
                      wz = as_float(lst[(4 * k + 3) * stride]);
          const float det = wx * nx + wy * ny + wz * nz;
          const float od = wx * mx + wy * my + wz * mz;
#if FM_STAR_CONSTB
          const float bound = cbn, ob = cbm;
#else
          const float bound = FM_STAR_ERR32 * (fabsf(wx) * nax + fabsf(wy) * nay + fabsf(wz) * naz);
          const float ob = FM_STAR_ERR32 * (fabsf(wx) * max_ + fabsf(wy) * may + fabsf(wz) * maz);
#endif
          open += (k != kk && k != kb && !(det < -bound) && !(od < -ob)) ? 1 : 0;
        }
        // fp64: the frame of the step (also needed for the circle), and the retest of what fp32 left open.  An entry
        // that is certainly inside the circle and on the left (the fp32 guess was wrong) replaces best: the part of
        // the circle on the left only shrinks, so what was outside it or on the right stays so.
        P4 B0 = s.spts[lst[(4 * kb) * stride]];
        const double Ux = Pk.x - g.x, Uy = Pk.y - g.y, Uz = Pk.z - g.z;
        double Vx = B0.x - g.x, Vy = B0.y - g.y, Vz = B0.z - g.z;
        double Nx = Uy * Vz - Uz * Vy, Ny = Uz * Vx - Ux * Vz, Nz = Ux * Vy - Uy * Vx;
        if (open) {
          const double Mx = g.y * Uz - g.z * Uy, My = g.z * Ux - g.x * Uz, Mz = g.x * Uy - g.y * Ux;
          const double Max = fabs(g.y * Uz) + fabs(g.z * Uy), May = fabs(g.z * Ux) + fabs(g.x * Uz),
                       Maz = fabs(g.x * Uy) + fabs(g.y * Ux);
          double Nax = fabs(Uy * Vz) + fabs(Uz * Vy), Nay = fabs(Uz * Vx) + fabs(Ux * Vz), Naz = fabs(Ux * Vy) + fabs(Uy * Vx);
          {
            const double od = Vx * Mx + Vy * My + Vz * Mz;
            unsure |= !(od > FM_STAR_ERR * (fabs(Vx) * Max + fabs(Vy) * May + fabs(Vz) * Maz)) ? 1 : 0;
          }
          for (int k = 0; k < m; k++) {
            if (k == kk || k == kb) continue;
            const P4 Q = s.spts[lst[(4 * k) * stride]];
            const double wx = Q.x - g.x, wy = Q.y - g.y, wz = Q.z - g.z;
            const double det = wx * Nx + wy * Ny + wz * Nz;
            const double bound = FM_STAR_ERR * (fabs(wx) * Nax + fabs(wy) * Nay + fabs(wz) * Naz);
            if (det < -bound) continue;
            const double od = wx * Mx + wy * My + wz * Mz;
            const double ob = FM_STAR_ERR * (fabs(wx) * Max + fabs(wy) * May + fabs(wz) * Maz);
            if (od < -ob) continue;
            if (!(det > bound) || !(od > ob)) { unsure = 1; continue; }
            kb = k; B0 = Q;
            Vx = wx; Vy = wy; Vz = wz;
            Nx = Uy * Vz - Uz * Vy; Ny = Uz * Vx - Ux * Vz; Nz = Ux * Vy - Uy * Vx;
            Nax = fabs(Uy * Vz) + fabs(Uz * Vy); Nay = fabs(Uz * Vx) + fabs(Ux * Vz); Naz = fabs(Ux * Vy) + fabs(Uy * Vx);
          }
        }
        {
          const double ng = Nx * g.x + Ny * g.y + Nz * g.z, nn = Nx * Nx + Ny * Ny + Nz * Nz;
          const double uu = Ux * Ux + Uy * Uy + Uz * Uz, vv = Vx * Vx + Vy * Vy + Vz * Vz;
          const bool sliver = !(ng > 0.0) || nn < 1.0e-12 * uu * vv;
          const double f = ng / nn;
          const double Px = Nx * f, Py = Ny * f, Pz = Nz * f;
          const double fx = g.x - Px, fy = g.y - Py, fz = g.z - Pz;
          const double rho2 = (fx * fx + fy * fy + fz * fz) * (1.0 + 1.0e-6);
          const double a0 = Px - lox, a1 = hix - Px, b0 = Py - loy, b1 = hiy - Py, c0 = Pz - loz, c1 = hiz - Pz;
          const double mm = fmin(fmin(fmin(a0, a1), fmin(b0, b1)), fmin(c0, c1));
          inbox = inbox && !sliver && mm > 0.0 && mm * mm >= rho2;
          const double want = sliver ? 1.0e300 : 4.0 * rho2;
          need2 = want > need2 ? want : need2;
        }
        if (kb == k1) break;
        kk = kb;
        Pk = B0;
      }
      // an inconclusive proof on a list that is not yet wide enough may just be a bad guess: only the last word counts
      if (!grow && inbox && (complete || need2 <= R2)) {
        if (unsure) return FM_S_UNSURE;
        star_finish(s, gi, d);
        if (s.dry && round >= 2) FM_COUNT_INC(&s.sc[FM_SC_MULTI]);
        return FM_S_OK;
      }
      if (complete) return unsure ? FM_S_UNSURE : FM_S_RADIUS;   // every candidate of the block was on the list
      R2 = need2 < 1.3 * R2 ? 1.3 * R2 : (need2 > 2.0 * R2 ? 2.0 * R2 : need2);   // partial stars overestimate: grow gently
      if (round == 4) R2 = 1.0e300;   // last round: every candidate of the block, if the list holds them
    }
    return FM_S_SLOW;
  }
  FM_HD static void record(const StarCtx& s, int j, int st) {
    if (s.dry) {
      if (st != FM_S_OK) FM_COUNT_INC(&s.sc[FM_SC_RETRIED]);
      return;
    }
    if (st == FM_S_RADIUS) {
      if (s.r < kStarMaxR) s.list_out[FM_COUNT_INC(&s.sc[s.out_idx])] = j;
      else FM_COUNT_INC(&s.sc[FM_SC_FAIL]);
    } else if (st == FM_S_UNSURE) {
      s.list_exact[FM_COUNT_INC(&s.sc[FM_SC_EXACT])] = j;
    } else if (st == FM_S_SLOW) {
      s.list_slow[FM_COUNT_INC(&s.sc[FM_SC_SLOW])] = j;
    } else if (st == FM_S_FAIL) {
      FM_COUNT_INC(&s.sc[FM_SC_FAIL]);
    }
  }
};

// first pass: every generator, block radius 1
template <bool EXACT>
struct KStarAllT {   // rng: 2 (2r+1)^2 + 4 list_cap ints per thread
  FM_HD static void run(const StarCtx& s, int j, int* rng, int stride) {
    const int st = EXACT ? KStarWrapT<true>::wrap(s, j, s.r, rng, stride)
                         : KStarWrapT<false>::wrap_near(s, j, s.r, rng, stride, rng + 2 * (2 * s.r + 1) * (2 * s.r + 1) * stride);
    KStarWrapT<EXACT>::record(s, j, st);
  }
};
// retry passes: the generators on list_in, block radius s.r
struct KStarRetry {
  FM_HD static void run(const StarCtx& s, int i, int* rng, int stride) {
    const int j = s.list_in[i];
    KStarWrapT<false>::record(s, j, KStarWrapT<false>::wrap_near(s, j, s.r, rng, stride,
                                                                 rng + 2 * (2 * s.r + 1) * (2 * s.r + 1) * stride));
  }
};
// the generators whose near list overflowed, first attempt: the same list form with a long list (kStarListLong
// entries; few generators get here, so the shared memory does not matter), widening the block as needed.  What
// overflows even that goes on list_in (counted in sc[FM_SC_LIST0]) for the full wrap.
struct KStarSlowNear {
  FM_HD static void run(const StarCtx& s, int i, int* rng, int stride) {
    const int j = s.list_slow[i];
    int st = FM_S_RADIUS;
    for (int r = 1; r <= kStarMaxR && st == FM_S_RADIUS; r++)
      st = KStarWrapT<false>::wrap_near(s, j, r, rng, stride, rng + 2 * (2 * kStarMaxR + 1) * (2 * kStarMaxR + 1) * stride);
    if (st == FM_S_UNSURE) s.list_exact[FM_COUNT_INC(&s.sc[FM_SC_EXACT])] = j;
    else if (st == FM_S_SLOW) s.list_in[FM_COUNT_INC(&s.sc[FM_SC_LIST0])] = j;
    else if (st != FM_S_OK) FM_COUNT_INC(&s.sc[FM_SC_FAIL]);
  }
};
// ... second attempt: the full wrap with the filters (list_in, sc[FM_SC_LIST0])
struct KStarSlow {
  FM_HD static void run(const StarCtx& s, int i, int* rng, int stride) {
    const int j = s.list_in[i];
    int st = FM_S_RADIUS;
    for (int r = 1; r <= kStarMaxR && st == FM_S_RADIUS; r++) st = KStarWrapT<false>::wrap(s, j, r, rng, stride);
    if (st == FM_S_UNSURE) s.list_exact[FM_COUNT_INC(&s.sc[FM_SC_EXACT])] = j;
    else if (st != FM_S_OK) FM_COUNT_INC(&s.sc[FM_SC_FAIL]);
  }
};
// the generators with an inconclusive filter: full predicates, widening the block as needed
struct KStarExact {
  FM_HD static void run(const StarCtx& s, int i, int* rng, int stride) {
    const int j = s.list_exact[i];
    int st = FM_S_RADIUS;
    for (int r = 1; r <= kStarMaxR && st == FM_S_RADIUS; r++) st = KStarWrapT<true>::wrap(s, j, r, rng, stride);
    if (st != FM_S_OK) FM_COUNT_INC(&s.sc[FM_SC_FAIL]);
  }
};

// ------------------------------------------------------------------ from stars to rows
// position of the triangle (o, x, y) -- x followed by y -- in the star of o, or -1
FM_HD inline int star_find(const StarCtx& s, int o, int x, int y) {
  const int d = s.deg[o];
  const int* ring = s.ring + (size_t)o * kStarCap;
  for (int k = 0; k < d; k++)
    if (ring[k] == x) return ring[k + 1 == d ? 0 : k + 1] == y ? k : -1;
  return -1;
}
FM_HD inline int star_slot_of(const StarCtx& s, int o, int k) {
  return (int)(s.off[o] & 0xffffffffULL) + star_popc(s.own[o] & ((1u << k) - 1u));
}
// slot of every triangle of every star (the triangle is looked up in the star of its owner:
// a triangle that is not there breaks the build)
struct KStarSlots {
  FM_HD static void run(const StarCtx& s, int i) {
    const int d = s.deg[i];
    const int* ring = s.ring + (size_t)i * kStarCap;
    int* out = s.stri + (size_t)i * kStarCap;
    const unsigned mask = s.own[i];
    int bad = 0;
    for (int k = 0; k < d; k++) {
      const int a = ring[k], b = ring[k + 1 == d ? 0 : k + 1];   // CCW (i, a, b)
      int slot = -1;
      if ((mask >> k) & 1u) slot = star_slot_of(s, i, k);
      else {
        const int o = a < b ? a : b;
        const int pos = o == a ? star_find(s, a, b, i) : star_find(s, b, i, a);   // (a, b, i) / (b, i, a)
        if (pos >= 0 && ((s.own[o] >> pos) & 1u)) slot = star_slot_of(s, o, pos);
      }
      if (slot < 0) { bad++; slot = 0; }
      out[k] = slot;
    }
    if (bad) FM_ATOMIC_ADD32(&s.sc[FM_SC_BROKEN], bad);
  }
};
// index, in the stored vertex order of the CCW triangle (p, q, r) -- smallest vertex first -- of w
FM_HD inline int star_edge_index(int p, int q, int r, int w) {
  int v0 = p, v1 = q, v2 = r;
  if (q < p && q < r) { v0 = q; v1 = r; v2 = p; }
  else if (r < p && r < q) { v0 = r; v1 = p; v2 = q; }
  (void)v0;
  return w == v1 ? 1 : (w == v2 ? 2 : 0);
}
// the rows of the triangles a generator owns: T = (i, x, y), a[k] = across the edge opposite vertex k
struct KStarRows {
  FM_HD static void run(const StarCtx& s, int i) {
    const int d = s.deg[i];
    const int* ring = s.ring + (size_t)i * kStarCap;
    const int* tri = s.stri + (size_t)i * kStarCap;
    const unsigned mask = s.own[i];
    s.v2t[i] = tri[0];
    int bad = 0;
    for (int k = 0; k < d; k++) {
      if (!((mask >> k) & 1u)) continue;
      const int kp = k + 1 == d ? 0 : k + 1, kpp = kp + 1 == d ? 0 : kp + 1, km = k == 0 ? d - 1 : k - 1;
      const int x = ring[k], y = ring[kp], z = ring[kpp], w = ring[km];
      // opposite x: the edge (y, i); the neighbour is (i, y, z), star position k+1, its far vertex z
      const int a1 = enc(tri[kp], star_edge_index(i, y, z, z));
      // opposite y: the edge (i, x); the neighbour is (i, w, x), star position k-1, its far vertex w
      const int a2 = enc(tri[km], star_edge_index(i, w, x, w));
      // opposite i: the edge (x, y); the neighbour holds y -> x: in the star of y, x is followed by its far vertex
      int a0 = 0;
      {
        const int dy = s.deg[y];
        const int* ry = s.ring + (size_t)y * kStarCap;
        int pos = -1;
        for (int q = 0; q < dy; q++)
          if (ry[q] == x) { pos = q; break; }
        if (pos < 0 || ry[pos == 0 ? dy - 1 : pos - 1] != i) bad++;   // T itself must be in the star of y: (y, i, x)
        else {
          const int far = ry[pos + 1 == dy ? 0 : pos + 1];
          a0 = enc(s.stri[(size_t)y * kStarCap + pos], star_edge_index(y, x, far, far));
        }
      }
      s.row[tri[k]] = Row{I4{i, x, y, 0}, I4{a0, a1, a2, 0}};
    }
    if (bad) FM_ATOMIC_ADD32(&s.sc[FM_SC_BROKEN], bad);
  }
};

}  // namespace fm
