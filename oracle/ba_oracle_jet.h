// ba_oracle_jet.h — CPU ORACLE (test infrastructure; see ba_oracle.h): the dual-number type and the
// rotation helper the reference's functors are instantiated on.  Shared by ba_oracle.cpp and by the
// Ceres stand-in headers (oracle/ref_standins/ceres/*.h) that the reference's OWN functor text is
// compiled against for oracle/_ref (oracle/make_ref.py).
#ifndef BA_ORACLE_JET_H_
#define BA_ORACLE_JET_H_
#include <cmath>
#include <limits>

namespace ba_oracle {

// ------------------------------------------------------------------------------------------
// Dual numbers with ceres::Jet<double,N> rules (Ceres 1.12 include/ceres/jet.h, published
// upstream; not in /root/reference).  Value a, infinitesimal part v[N].
// ------------------------------------------------------------------------------------------
template <int N>
struct Jet {
  double a;
  double v[N];
  Jet() : a(0.0) { for (int i = 0; i < N; ++i) v[i] = 0.0; }
  explicit Jet(double s) : a(s) { for (int i = 0; i < N; ++i) v[i] = 0.0; }
  Jet(double s, int k) : a(s) { for (int i = 0; i < N; ++i) v[i] = 0.0; v[k] = 1.0; }
};
template <int N> inline Jet<N> operator+(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h; h.a = f.a + g.a; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] + g.v[i]; return h; }
template <int N> inline Jet<N> operator-(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h; h.a = f.a - g.a; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] - g.v[i]; return h; }
template <int N> inline Jet<N> operator*(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h; h.a = f.a * g.a; for (int i = 0; i < N; ++i) h.v[i] = f.a * g.v[i] + f.v[i] * g.a; return h; }
template <int N> inline Jet<N> operator/(const Jet<N>& f, const Jet<N>& g) {
  // ceres: g_a_inverse = 1/g.a; f_a_by_g_a = f.a * g_a_inverse; (f.v - f_a_by_g_a * g.v) * g_a_inverse
  const double gi = 1.0 / g.a; const double q = f.a * gi;
  Jet<N> h; h.a = q; for (int i = 0; i < N; ++i) h.v[i] = (f.v[i] - q * g.v[i]) * gi; return h; }
template <int N> inline Jet<N> operator/(double s, const Jet<N>& g) {
  const double m = -s / (g.a * g.a);
  Jet<N> h; h.a = s / g.a; for (int i = 0; i < N; ++i) h.v[i] = g.v[i] * m; return h; }
template <int N> inline Jet<N> operator+=(Jet<N>& f, const Jet<N>& g) { f = f + g; return f; }
template <int N> inline bool operator>(const Jet<N>& f, const Jet<N>& g) { return f.a > g.a; }
template <int N> inline Jet<N> sqrt(const Jet<N>& f) {
  const double t = std::sqrt(f.a); const double two_a_inverse = 1.0 / (2.0 * t);
  Jet<N> h; h.a = t; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * two_a_inverse; return h; }
template <int N> inline Jet<N> cos(const Jet<N>& f) {
  const double ms = -std::sin(f.a);
  Jet<N> h; h.a = std::cos(f.a); for (int i = 0; i < N; ++i) h.v[i] = ms * f.v[i]; return h; }
template <int N> inline Jet<N> sin(const Jet<N>& f) {
  const double c = std::cos(f.a);
  Jet<N> h; h.a = std::sin(f.a); for (int i = 0; i < N; ++i) h.v[i] = c * f.v[i]; return h; }
inline double sqrt(double x) { return std::sqrt(x); }
inline double cos(double x) { return std::cos(x); }
inline double sin(double x) { return std::sin(x); }

// ceres::AngleAxisRotatePoint (Ceres 1.12 include/ceres/rotation.h).
template <typename T>
inline void AngleAxisRotatePoint(const T angle_axis[3], const T pt[3], T result[3]) {
  const T theta2 = angle_axis[0] * angle_axis[0] + angle_axis[1] * angle_axis[1] + angle_axis[2] * angle_axis[2];
  if (theta2 > T(std::numeric_limits<double>::epsilon())) {
    const T theta = sqrt(theta2);
    const T costheta = cos(theta);
    const T sintheta = sin(theta);
    const T theta_inverse = 1.0 / theta;
    const T w[3] = {angle_axis[0] * theta_inverse, angle_axis[1] * theta_inverse, angle_axis[2] * theta_inverse};
    const T w_cross_pt[3] = {w[1] * pt[2] - w[2] * pt[1], w[2] * pt[0] - w[0] * pt[2], w[0] * pt[1] - w[1] * pt[0]};
    const T tmp = (w[0] * pt[0] + w[1] * pt[1] + w[2] * pt[2]) * (T(1.0) - costheta);
    result[0] = pt[0] * costheta + w_cross_pt[0] * sintheta + w[0] * tmp;
    result[1] = pt[1] * costheta + w_cross_pt[1] * sintheta + w[1] * tmp;
    result[2] = pt[2] * costheta + w_cross_pt[2] * sintheta + w[2] * tmp;
  } else {
    const T w_cross_pt[3] = {angle_axis[1] * pt[2] - angle_axis[2] * pt[1], angle_axis[2] * pt[0] - angle_axis[0] * pt[2],
                             angle_axis[0] * pt[1] - angle_axis[1] * pt[0]};
    result[0] = pt[0] + w_cross_pt[0];
    result[1] = pt[1] + w_cross_pt[1];
    result[2] = pt[2] + w_cross_pt[2];
  }
}

}  // namespace ba_oracle
#endif
