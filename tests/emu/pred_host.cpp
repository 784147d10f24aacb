// Host build of fishman_b200/csrc/predicates.cuh for the CPU-side predicate
// tests (tests/test_predicates.py).  TEST INFRASTRUCTURE ONLY -- the product
// library never links this file.
#include "../../fishman_b200/csrc/predicates.cuh"
extern "C" {
int fmt_orient(const double* a, const double* b, const double* c, int ia, int ib, int ic) {
  fm::P3 A{a[0], a[1], a[2]}, B{b[0], b[1], b[2]}, C{c[0], c[1], c[2]};
  return fm::orient(A, B, C, ia, ib, ic);
}
int fmt_orient_slow(const double* a, const double* b, const double* c, int ia, int ib, int ic) {
  fm::P3 A{a[0], a[1], a[2]}, B{b[0], b[1], b[2]}, C{c[0], c[1], c[2]};
  return fm::orient_slow(A, B, C, ia, ib, ic);
}
int fmt_incircle(const double* a, const double* b, const double* c, const double* d, int ia, int ib,
                 int ic, int id) {
  fm::P3 A{a[0], a[1], a[2]}, B{b[0], b[1], b[2]}, C{c[0], c[1], c[2]}, D{d[0], d[1], d[2]};
  return fm::incircle(A, B, C, D, ia, ib, ic, id);
}
int fmt_incircle_slow(const double* a, const double* b, const double* c, const double* d, int ia,
                      int ib, int ic, int id) {
  fm::P3 A{a[0], a[1], a[2]}, B{b[0], b[1], b[2]}, C{c[0], c[1], c[2]}, D{d[0], d[1], d[2]};
  return fm::incircle_slow(A, B, C, D, ia, ib, ic, id);
}
int fmt_exact_det_sign(const double* m, int k) { return fm::exact_det_sign(m, k); }
}
