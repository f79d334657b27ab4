// TEST INFRASTRUCTURE ONLY -- never linked into, imported by or executed from the product path.
//
// Thin extern "C" wrappers that call the *unmodified* reference templates
// `celerite2::core::*` (headers included from /root/reference at build time, nothing copied)
// with plain pointers.  Built by oracle/build_oracle.py into oracle/_ref/ (git-ignored).
//
// Why call the core templates directly instead of the reference's pybind modules: the generated
// bindings type the S workspace as Matrix<double, Dynamic, SIZE*SIZE> which collapses to an N x 1
// column when SIZE == Eigen::Dynamic (python/celerite2/backprop.cpp:60,136), so factor_fwd/factor_rev
// are wrong for J > 4.  c++/test/test_factor_rev.cpp:13-33 shows the correct usage: row-major dynamic
// S and F, which is what these wrappers do.
//
// J dispatch: fixed-size LowRank for the widths listed in C2REF_CASES (what the reference's
// UNWRAP_CASES macros do, python/celerite2/driver.hpp:27-101), Eigen::Dynamic otherwise.

#include <cstdint>
#include <cmath>
#include <Eigen/Core>
#include <celerite2/celerite2.h>

namespace {

typedef Eigen::Matrix<double, Eigen::Dynamic, 1> Vec;
typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> RowMat;

template <int SIZE>
struct order {
  const static int value = (SIZE == 1) ? Eigen::ColMajor : Eigen::RowMajor;
};

template <int SIZE>
using LowRankT = Eigen::Matrix<double, Eigen::Dynamic, SIZE, order<SIZE>::value>;
template <int SIZE>
using CoeffT = Eigen::Matrix<double, SIZE, 1>;

} // namespace

#ifdef C2REF_FAST
#define C2REF_CASES                                                                                                                                  \
  switch (J) {                                                                                                                                       \
    case 1: BODY(1); break;                                                                                                                          \
    case 2: BODY(2); break;                                                                                                                          \
    case 3: BODY(3); break;                                                                                                                          \
    case 4: BODY(4); break;                                                                                                                          \
    case 5: BODY(5); break;                                                                                                                          \
    case 6: BODY(6); break;                                                                                                                          \
    case 7: BODY(7); break;                                                                                                                          \
    case 8: BODY(8); break;                                                                                                                          \
    case 16: BODY(16); break;                                                                                                                        \
    default: BODY(Eigen::Dynamic);                                                                                                                   \
  }
#else
#define C2REF_CASES                                                                                                                                  \
  switch (J) {                                                                                                                                       \
    case 1: BODY(1); break;                                                                                                                          \
    case 2: BODY(2); break;                                                                                                                          \
    case 3: BODY(3); break;                                                                                                                          \
    case 4: BODY(4); break;                                                                                                                          \
    default: BODY(Eigen::Dynamic);                                                                                                                   \
  }
#endif

#define MAP_T Eigen::Map<const Vec> t_(t, N, 1)
#define MAP_C(SIZE) Eigen::Map<const CoeffT<SIZE>> c_(c, J, 1)
#define MAP_LR(SIZE, name) Eigen::Map<const LowRankT<SIZE>> name##_(name, N, J)
#define MAP_LR_OUT(SIZE, name) Eigen::Map<LowRankT<SIZE>> name##_(name, N, J)

extern "C" {

// core::factor, forward.hpp:69-135 (S != NULL) / interface.hpp:37-48 (S == NULL)
int64_t c2ref_factor(int64_t N, int64_t J, const double *t, const double *c, const double *a, const double *U, const double *V, double *d,
                     double *W, double *S) {
  Eigen::Index flag = 0;
#define BODY(SIZE)                                                                                                                                   \
  {                                                                                                                                                  \
    MAP_T;                                                                                                                                           \
    MAP_C(SIZE);                                                                                                                                     \
    Eigen::Map<const Vec> a_(a, N, 1);                                                                                                               \
    MAP_LR(SIZE, U);                                                                                                                                 \
    MAP_LR(SIZE, V);                                                                                                                                 \
    Eigen::Map<Vec> d_(d, N, 1);                                                                                                                     \
    MAP_LR_OUT(SIZE, W);                                                                                                                             \
    if (S) {                                                                                                                                         \
      Eigen::Map<RowMat> S_(S, N, J * J);                                                                                                            \
      flag = celerite2::core::factor(t_, c_, a_, U_, V_, d_, W_, S_);                                                                                \
    } else {                                                                                                                                         \
      flag = celerite2::core::factor(t_, c_, a_, U_, V_, d_, W_);                                                                                    \
    }                                                                                                                                                \
  }
  C2REF_CASES
#undef BODY
  return (int64_t)flag;
}

// core::factor_rev, reverse.hpp:10-85
void c2ref_factor_rev(int64_t N, int64_t J, const double *t, const double *c, const double *a, const double *U, const double *V,
                      const double *d, const double *W, const double *S, const double *bd, const double *bW, double *bt, double *bc,
                      double *ba, double *bU, double *bV) {
#define BODY(SIZE)                                                                                                                                   \
  {                                                                                                                                                  \
    MAP_T;                                                                                                                                           \
    MAP_C(SIZE);                                                                                                                                     \
    Eigen::Map<const Vec> a_(a, N, 1);                                                                                                               \
    MAP_LR(SIZE, U);                                                                                                                                 \
    MAP_LR(SIZE, V);                                                                                                                                 \
    Eigen::Map<const Vec> d_(d, N, 1);                                                                                                               \
    MAP_LR(SIZE, W);                                                                                                                                 \
    Eigen::Map<const RowMat> S_(S, N, J * J);                                                                                                        \
    Eigen::Map<const Vec> bd_(bd, N, 1);                                                                                                             \
    MAP_LR(SIZE, bW);                                                                                                                                \
    Eigen::Map<Vec> bt_(bt, N, 1);                                                                                                                   \
    Eigen::Map<CoeffT<SIZE>> bc_(bc, J, 1);                                                                                                          \
    Eigen::Map<Vec> ba_(ba, N, 1);                                                                                                                   \
    MAP_LR_OUT(SIZE, bU);                                                                                                                            \
    MAP_LR_OUT(SIZE, bV);                                                                                                                            \
    celerite2::core::factor_rev(t_, c_, a_, U_, V_, d_, W_, S_, bd_, bW_, bt_, bc_, ba_, bU_, bV_);                                                  \
  }
  C2REF_CASES
#undef BODY
}

// The four linear sweeps share one wrapper shape.
//   op: 0 solve_lower (forward.hpp:156-170)  1 solve_upper (193-207)  2 matmul_lower (228-239)  3 matmul_upper (260-271)
// F == NULL selects the no-workspace overloads (interface.hpp:70-152).
void c2ref_sweep(int32_t op, int64_t N, int64_t J, int64_t K, const double *t, const double *c, const double *U, const double *W,
                 const double *Y, double *Z, double *F) {
#define CALL(FN)                                                                                                                                     \
  if (F) {                                                                                                                                           \
    Eigen::Map<RowMat> F_(F, N, J * K);                                                                                                              \
    celerite2::core::FN(t_, c_, U_, W_, Y_, Z_, F_);                                                                                                 \
  } else {                                                                                                                                           \
    celerite2::core::FN(t_, c_, U_, W_, Y_, Z_);                                                                                                     \
  }
#define BODY(SIZE)                                                                                                                                   \
  {                                                                                                                                                  \
    MAP_T;                                                                                                                                           \
    MAP_C(SIZE);                                                                                                                                     \
    MAP_LR(SIZE, U);                                                                                                                                 \
    MAP_LR(SIZE, W);                                                                                                                                 \
    if (K == 1) {                                                                                                                                    \
      Eigen::Map<const Vec> Y_(Y, N, 1);                                                                                                             \
      Eigen::Map<Vec> Z_(Z, N, 1);                                                                                                                   \
      switch (op) {                                                                                                                                  \
        case 0: CALL(solve_lower); break;                                                                                                            \
        case 1: CALL(solve_upper); break;                                                                                                            \
        case 2: CALL(matmul_lower); break;                                                                                                           \
        default: CALL(matmul_upper); break;                                                                                                          \
      }                                                                                                                                              \
    } else {                                                                                                                                         \
      Eigen::Map<const RowMat> Y_(Y, N, K);                                                                                                          \
      Eigen::Map<RowMat> Z_(Z, N, K);                                                                                                                \
      switch (op) {                                                                                                                                  \
        case 0: CALL(solve_lower); break;                                                                                                            \
        case 1: CALL(solve_upper); break;                                                                                                            \
        case 2: CALL(matmul_lower); break;                                                                                                           \
        default: CALL(matmul_upper); break;                                                                                                          \
      }                                                                                                                                              \
    }                                                                                                                                                \
  }
  C2REF_CASES
#undef BODY
#undef CALL
}

// Reverse-mode sweeps, reverse.hpp:87-217.  op numbering as in c2ref_sweep.
void c2ref_sweep_rev(int32_t op, int64_t N, int64_t J, int64_t K, const double *t, const double *c, const double *U, const double *W,
                     const double *Y, const double *Z, const double *F, const double *bZ, double *bt, double *bc, double *bU, double *bW,
                     double *bY) {
#define BODY(SIZE)                                                                                                                                   \
  {                                                                                                                                                  \
    MAP_T;                                                                                                                                           \
    MAP_C(SIZE);                                                                                                                                     \
    MAP_LR(SIZE, U);                                                                                                                                 \
    MAP_LR(SIZE, W);                                                                                                                                 \
    Eigen::Map<const RowMat> Y_(Y, N, K);                                                                                                            \
    Eigen::Map<const RowMat> Z_(Z, N, K);                                                                                                            \
    Eigen::Map<const RowMat> F_(F, N, J * K);                                                                                                        \
    Eigen::Map<const RowMat> bZ_(bZ, N, K);                                                                                                          \
    Eigen::Map<Vec> bt_(bt, N, 1);                                                                                                                   \
    Eigen::Map<CoeffT<SIZE>> bc_(bc, J, 1);                                                                                                          \
    MAP_LR_OUT(SIZE, bU);                                                                                                                            \
    MAP_LR_OUT(SIZE, bW);                                                                                                                            \
    Eigen::Map<RowMat> bY_(bY, N, K);                                                                                                                \
    switch (op) {                                                                                                                                    \
      case 0: celerite2::core::solve_lower_rev(t_, c_, U_, W_, Y_, Z_, F_, bZ_, bt_, bc_, bU_, bW_, bY_); break;                                     \
      case 1: celerite2::core::solve_upper_rev(t_, c_, U_, W_, Y_, Z_, F_, bZ_, bt_, bc_, bU_, bW_, bY_); break;                                     \
      case 2: celerite2::core::matmul_lower_rev(t_, c_, U_, W_, Y_, Z_, F_, bZ_, bt_, bc_, bU_, bW_, bY_); break;                                    \
      default: celerite2::core::matmul_upper_rev(t_, c_, U_, W_, Y_, Z_, F_, bZ_, bt_, bc_, bU_, bW_, bY_); break;                                   \
    }                                                                                                                                                \
  }
  C2REF_CASES
#undef BODY
}

// core::general_matmul_lower / _upper, forward.hpp:285-392 (F == NULL: interface.hpp:171-220).  upper != 0 selects _upper.
void c2ref_general_matmul(int32_t upper, int64_t N, int64_t M, int64_t J, int64_t K, const double *t1, const double *t2, const double *c,
                          const double *U, const double *V, const double *Y, double *Z, double *F) {
#define BODY(SIZE)                                                                                                                                   \
  {                                                                                                                                                  \
    Eigen::Map<const Vec> t1_(t1, N, 1);                                                                                                             \
    Eigen::Map<const Vec> t2_(t2, M, 1);                                                                                                             \
    MAP_C(SIZE);                                                                                                                                     \
    Eigen::Map<const LowRankT<SIZE>> U_(U, N, J);                                                                                                    \
    Eigen::Map<const LowRankT<SIZE>> V_(V, M, J);                                                                                                    \
    Eigen::Map<const RowMat> Y_(Y, M, K);                                                                                                            \
    Eigen::Map<RowMat> Z_(Z, N, K);                                                                                                                  \
    if (F) {                                                                                                                                         \
      Eigen::Map<RowMat> F_(F, M, J * K);                                                                                                            \
      if (upper)                                                                                                                                     \
        celerite2::core::general_matmul_upper(t1_, t2_, c_, U_, V_, Y_, Z_, F_);                                                                     \
      else                                                                                                                                           \
        celerite2::core::general_matmul_lower(t1_, t2_, c_, U_, V_, Y_, Z_, F_);                                                                     \
    } else {                                                                                                                                         \
      if (upper)                                                                                                                                     \
        celerite2::core::general_matmul_upper(t1_, t2_, c_, U_, V_, Y_, Z_);                                                                         \
      else                                                                                                                                           \
        celerite2::core::general_matmul_lower(t1_, t2_, c_, U_, V_, Y_, Z_);                                                                         \
    }                                                                                                                                                \
  }
  C2REF_CASES
#undef BODY
}

// core::to_dense, forward.hpp:19-42.  K_out is N x N row-major.
void c2ref_to_dense(int64_t N, int64_t J, const double *t, const double *c, const double *a, const double *U, const double *V, double *K_out) {
  Eigen::Map<const Vec> t_(t, N, 1);
  Eigen::Map<const Vec> c_(c, J, 1);
  Eigen::Map<const Vec> a_(a, N, 1);
  Eigen::Map<const RowMat> U_(U, N, J);
  Eigen::Map<const RowMat> V_(V, N, J);
  Eigen::Map<RowMat> K_(K_out, N, N);
  celerite2::core::to_dense(t_, c_, a_, U_, V_, K_);
}

// Term::get_celerite_matrices, terms.hpp:112-139 (blocked column order [real | cos | sin]).
// c_out (J), a_out (N), U_out/V_out (N, J) row-major.
void c2ref_get_celerite_matrices(int64_t N, int64_t Jr, int64_t Jc, const double *ar, const double *cr, const double *ac, const double *bc,
                                 const double *cc, const double *dc, const double *x, const double *diag, double *c_out, double *a_out,
                                 double *U_out, double *V_out) {
  typedef celerite2::Term<double, Eigen::Dynamic> TermT;
  TermT term;
  Eigen::Map<const Vec> ar_(ar, Jr), cr_(cr, Jr), ac_(ac, Jc), bc_(bc, Jc), cc_(cc, Jc), dc_(dc, Jc);
  term.set_coefficients(ar_, cr_, ac_, bc_, cc_, dc_);
  Eigen::Map<const Vec> x_(x, N), diag_(diag, N);
  auto mats    = term.get_celerite_matrices(x_, diag_);
  int64_t J    = Jr + 2 * Jc;
  const auto &c = std::get<0>(mats);
  const auto &a = std::get<1>(mats);
  const auto &U = std::get<2>(mats);
  const auto &V = std::get<3>(mats);
  for (int64_t j = 0; j < J; ++j) c_out[j] = c(j);
  for (int64_t n = 0; n < N; ++n) {
    a_out[n] = a(n);
    for (int64_t j = 0; j < J; ++j) {
      U_out[n * J + j] = U(n, j);
      V_out[n * J + j] = V(n, j);
    }
  }
}

// 1 when built with the wide fixed-size dispatch (-DC2REF_FAST), else 0
int32_t c2ref_is_fast() {
#ifdef C2REF_FAST
  return 1;
#else
  return 0;
#endif
}

} // extern "C"
