// TEST ONLY. Host instantiation of the product's minimiser source (mundy_b200/csrc/minimize.cuh is __host__ __device__),
// so that the exact text the device kernels are compiled from can be checked bit for bit against the reference's outputs
// on a machine without a GPU (tests/test_primitives_cpu.py). On the host sin / cos are glibc's, as in the reference, so
// equality is exact; on the device they are CUDA's and the GPU tests state a tolerance. Not part of libmundy_b200.so.
#include "minimize.cuh"

using namespace mb200;

static inline V3 h3(const double* p, long i) { return V3{p[3 * i], p[3 * i + 1], p[3 * i + 2]}; }
static inline Q4 h4(const double* p, long i) { return Q4{p[4 * i], p[4 * i + 1], p[4 * i + 2], p[4 * i + 3]}; }
static inline void put(double* p, long i, const V3& v) { p[3 * i] = v.x, p[3 * i + 1] = v.y, p[3 * i + 2] = v.z; }

extern "C" {

void hostchk_point_ellipsoid(long n, const double* x, const double* c, const double* quat, const double* radii,
                             double* dist, double* cp, double* normal) {
  for (long i = 0; i < n; ++i) {
    const prim::Ellipsoid e{h3(c, i), h4(quat, i), h3(radii, i)};
    V3 p, nn;
    dist[i] = prim::dist_point_ellipsoid(h3(x, i), e, p, nn);
    put(cp, i, p), put(normal, i, nn);
  }
}

void hostchk_ellipsoid_ellipsoid(long n, const double* c1, const double* q1, const double* rad1, const double* c2,
                                 const double* q2, const double* rad2, double* dist, double* cp1, double* cp2,
                                 double* n1) {
  for (long i = 0; i < n; ++i) {
    const prim::Ellipsoid e1{h3(c1, i), h4(q1, i), h3(rad1, i)}, e2{h3(c2, i), h4(q2, i), h3(rad2, i)};
    V3 p1, p2, nn;
    dist[i] = prim::dist_ellipsoid_ellipsoid(e1, e2, p1, p2, nn);
    put(cp1, i, p1), put(cp2, i, p2), put(n1, i, nn);
  }
}

void hostchk_circle_circle(long n, const double* c1, const double* q1, const double* r1, const double* c2,
                           const double* q2, const double* r2, double* dist, double* cp1, double* cp2, double* n1) {
  for (long i = 0; i < n; ++i) {
    const prim::Circle k1{h3(c1, i), h4(q1, i), r1[i]}, k2{h3(c2, i), h4(q2, i), r2[i]};
    V3 p1, p2, nn;
    dist[i] = prim::dist_circle_circle(k1, k2, p1, p2, nn);
    put(cp1, i, p1), put(cp2, i, p2), put(n1, i, nn);
  }
}

}  // extern "C"
