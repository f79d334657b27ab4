// The one-line switch: this program includes <celerite2/celerite2.h> exactly as a user of the reference does
// (c++/include/celerite2/celerite2.h) and is compiled with -I<repo>/include, so the names resolve to the shim headers
// over libcelerite2_b200.so.  It follows c++/test/test_to_dense.cpp, test_terms.cpp and the failure section of
// test_factor.cpp: kernels built from celerite2::RealTerm / ComplexTerm / SHOTerm and operator+, matrices from
// kernel.get_celerite_matrices, to_dense against the kernel function, factor + solves against dense algebra, and the
// reference's failure semantics (forward.hpp:128: rows after the first non-positive pivot stay untouched).
#include <Eigen/Dense>
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include <celerite2/celerite2.h>

typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> Matrix;
typedef Eigen::VectorXd Vector;

#define REQUIRE(cond)                                                                  \
  do {                                                                                 \
    if (!(cond)) {                                                                     \
      std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);                    \
      std::exit(1);                                                                    \
    }                                                                                  \
  } while (0)

template <typename Kernel>
static double kernel_value(const Kernel &k, double tau) {
  Vector ar, cr, ac, bc, cc, dc;
  std::tie(ar, cr, ac, bc, cc, dc) = k.get_coefficients();
  tau      = std::abs(tau);
  double v = 0.0;
  for (int i = 0; i < ar.rows(); ++i) v += ar(i) * std::exp(-cr(i) * tau);
  for (int i = 0; i < ac.rows(); ++i) v += std::exp(-cc(i) * tau) * (ac(i) * std::cos(dc(i) * tau) + bc(i) * std::sin(dc(i) * tau));
  return v;
}

template <typename Kernel>
static void check_kernel(const Kernel &kernel, int J, const Vector &x, const Vector &diag, const Matrix &Y) {
  using namespace celerite2::core;
  const int N = (int)x.rows();
  typename Kernel::CoeffVector c;
  Vector a;
  typename Kernel::LowRank U, V;
  std::tie(c, a, U, V) = kernel.get_celerite_matrices(x, diag);
  REQUIRE(U.rows() == N && U.cols() == J && V.cols() == J && c.rows() == J && a.rows() == N);
  // to_dense against the kernel function (c++/test/test_to_dense.cpp, 1e-8 there)
  Matrix K;
  to_dense(x, c, a, U, V, K);
  REQUIRE(K.rows() == N && K.cols() == N);
  double worst = 0.0;
  for (int n = 0; n < N; ++n)
    for (int m = 0; m < N; ++m) worst = std::max(worst, std::abs(K(n, m) - (kernel_value(kernel, x(n) - x(m)) + (n == m ? diag(n) : 0.0))));
  REQUIRE(worst < 1e-11);
  // factor and the two solves against dense Cholesky
  Eigen::LLT<Eigen::MatrixXd> llt{Eigen::MatrixXd(K)};
  const Eigen::MatrixXd Lc = llt.matrixL();
  Vector d;
  typename Kernel::LowRank W;
  REQUIRE(factor(x, c, a, U, V, d, W) == 0);
  REQUIRE((d.array() - Lc.diagonal().array().square()).abs().maxCoeff() < 1e-11);
  Matrix Z = Y;
  solve_lower(x, c, U, W, Z, Z);
  Z.array().colwise() /= d.array();
  solve_upper(x, c, U, W, Z, Z);
  REQUIRE((Eigen::MatrixXd(K) * Eigen::MatrixXd(Z) - Eigen::MatrixXd(Y)).cwiseAbs().maxCoeff() < 1e-9);
}

int main() {
  const int N = 50, nrhs = 5;
  Vector x(N), diag(N);
  Matrix Y(N, nrhs);
  for (int i = 0; i < N; ++i) {
    const double delta = double(i) / (N - 1);
    x(i)    = 10 * delta + delta * delta;
    diag(i) = 0.5;
    for (int j = 0; j < nrhs; ++j) Y(i, j) = std::sin(x(i) + double(j) / nrhs);
  }
  // the kernels of c++/test/helpers.hpp:27-62
  celerite2::RealTerm<double> real(1.0, 0.1);
  celerite2::ComplexTerm<double> cplx(0.8, 0.03, 1.0, 0.1);
  celerite2::SHOTerm<double> sho_over(1.2, 0.3, 0.1), sho_under(0.1, 1.3, 5.3);
  check_kernel(real, 1, x, diag, Y);
  check_kernel(cplx, 2, x, diag, Y);
  check_kernel(sho_over, 2, x, diag, Y);
  check_kernel(sho_under, 2, x, diag, Y);
  check_kernel(real + cplx, 3, x, diag, Y);
  check_kernel(real + cplx + sho_over, 5, x, diag, Y);
  check_kernel(real + cplx + sho_over + sho_under, 7, x, diag, Y);
  check_kernel(sho_over + sho_under, 4, x, diag, Y);
  REQUIRE((real + cplx).get_width() == 3);

  // dimension mismatch (terms.hpp:114)
  bool threw = false;
  try {
    Vector short_diag = diag.head(N - 1);
    real.get_celerite_matrices(x, short_diag);
  } catch (const celerite2::dimension_mismatch &) {
    threw = true;
  }
  REQUIRE(threw);

  // failure semantics: first non-positive pivot at row 7; d[0..7] written, W[0..6] written, everything after untouched
  {
    using namespace celerite2::core;
    auto k = sho_over + sho_under;
    decltype(k)::CoeffVector c;
    Vector a;
    decltype(k)::LowRank U, V;
    std::tie(c, a, U, V) = k.get_celerite_matrices(x, diag);
    Vector good_d;
    decltype(k)::LowRank good_W;
    REQUIRE(factor(x, c, a, U, V, good_d, good_W) == 0);
    Vector bad = a;
    bad(7)     = -5.0;
    Vector d   = Vector::Constant(N, 123.0);
    decltype(k)::LowRank W = decltype(k)::LowRank::Constant(N, 4, 321.0);
    Matrix S               = Matrix::Constant(N, 16, 77.0);
    REQUIRE(factor(x, c, bad, U, V, d, W, S) == 7);
    for (int n = 0; n < 7; ++n) REQUIRE(std::abs(d(n) - good_d(n)) < 1e-13 && (W.row(n) - good_W.row(n)).cwiseAbs().maxCoeff() < 1e-13);
    REQUIRE(d(7) <= 0.0);
    for (int n = 8; n < N; ++n) REQUIRE(d(n) == 123.0);
    for (int n = 7; n < N; ++n) REQUIRE((W.row(n).array() == 321.0).all());
    for (int n = 8; n < N; ++n) REQUIRE((S.row(n).array() == 77.0).all());
    REQUIRE((S.row(0).array() == 0.0).all());
    // in place (d aliases a, W aliases V): rows after the failure keep the INPUT values
    Vector a_ip = bad;
    decltype(k)::LowRank V_ip = V;
    REQUIRE(factor(x, c, a_ip, U, V_ip, a_ip, V_ip) == 7);
    for (int n = 8; n < N; ++n) REQUIRE(a_ip(n) == bad(n) && (V_ip.row(n) - V.row(n)).cwiseAbs().maxCoeff() == 0.0);
  }
  std::printf("OK\n");
  return 0;
}
