// Drop-in replacement for the reference's <celerite2/core.hpp>: the same celerite2::core::* function templates
// (names, template heads, argument order, in-place resizing of the outputs) with bodies that stage the operands to
// the GPU and call the C ABI of libcelerite2_b200.so.  Needs Eigen (the caller's own copy) and
// include/celerite2_b200.h; no CUDA headers.
//
//   reference                                              here
//   c++/include/celerite2/forward.hpp:69-135   factor            -> c2b_factor
//   forward.hpp:156-207                        solve_lower/upper -> c2b_solve_lower / c2b_solve_upper
//   forward.hpp:228-271                        matmul_lower/upper-> c2b_matmul_lower / c2b_matmul_upper
//   forward.hpp:285-392                        general_matmul_*  -> c2b_general_matmul_lower / _upper
//   interface.hpp:37-220                       the no-workspace overloads
//   reverse.hpp:10-217                         *_rev             -> c2b_*_rev
//   forward.hpp:19-42                          to_dense          -> c2b_to_dense
// The headers under include/celerite2/ (celerite2.h, core.hpp, forward.hpp, interface.hpp, reverse.hpp, terms.hpp) carry
// the reference's file names, so `#include <celerite2/celerite2.h>` with -I<this repo>/include is the whole switch.
//
// Operands may be arbitrary Eigen expressions (they are evaluated into contiguous row-major host buffers first,
// which also makes the reference's aliasing contracts hold trivially); outputs are resized like the reference's
// CAST_* macros do (internal.hpp:11-19).  Errors of the C ABI become std::runtime_error; `factor` returns the
// reference's flag (0, or the first row with a non-positive pivot).  There is no CPU fallback.
#ifndef CELERITE2_B200_CORE_HPP_
#define CELERITE2_B200_CORE_HPP_

#include <Eigen/Core>
#include <stdexcept>
#include <string>
#include <vector>

#include "../celerite2_b200.h"

namespace celerite2 {
namespace b200 {

typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> RowMat;
typedef Eigen::Matrix<double, Eigen::Dynamic, 1> Vec;

inline void check(int rc) {
  if (rc != C2B_OK) throw std::runtime_error(std::string("celerite2_b200: ") + c2b_last_error());
}

// A device buffer of doubles with host <-> device copies.
class DeviceArray {
  public:
  explicit DeviceArray(size_t n) : n_(n), p_((double *)c2b_device_malloc(n * sizeof(double))) {
    if (!p_) throw std::runtime_error(std::string("celerite2_b200: ") + c2b_last_error());
  }
  template <typename Derived>
  explicit DeviceArray(const Eigen::MatrixBase<Derived> &x) : DeviceArray((size_t)x.size()) {
    RowMat h = x;  // evaluates expressions, converts to row-major
    check(c2b_copy_to_device(p_, h.data(), n_ * sizeof(double)));
  }
  DeviceArray(const DeviceArray &) = delete;
  DeviceArray &operator=(const DeviceArray &) = delete;
  ~DeviceArray() { c2b_device_free(p_); }
  double *ptr() const { return p_; }
  // copy back into an Eigen output of the given shape (resized in place, any storage order)
  template <typename Derived>
  void fetch(Eigen::MatrixBase<Derived> const &out_, Eigen::Index rows, Eigen::Index cols) const {
    Eigen::MatrixBase<Derived> &out = const_cast<Eigen::MatrixBase<Derived> &>(out_);
    RowMat h(rows, cols);
    check(c2b_copy_to_host(h.data(), p_, n_ * sizeof(double)));
    out.derived().resize(rows, cols);
    out = h;
  }

  // after a failed factorisation: only the first `nrows` rows are the reference's values; the rest of the caller's
  // array stays as it was (forward.hpp:128 returns before touching later rows)
  template <typename Derived>
  void fetch_rows(Eigen::MatrixBase<Derived> const &out_, Eigen::Index rows, Eigen::Index cols, Eigen::Index nrows) const {
    Eigen::MatrixBase<Derived> &out = const_cast<Eigen::MatrixBase<Derived> &>(out_);
    RowMat h(rows, cols);
    check(c2b_copy_to_host(h.data(), p_, n_ * sizeof(double)));
    out.derived().resize(rows, cols);  // a no-op (contents kept) when the caller's array already has this shape
    if (nrows > rows) nrows = rows;
    if (nrows > 0) out.topRows(nrows) = h.topRows(nrows);
  }

  private:
  size_t n_;
  double *p_;
};

}  // namespace b200

namespace core {

// ---- to_dense (forward.hpp:19-42) ----------------------------------------------------------------------------------
template <typename Input, typename Coeffs, typename Diag, typename LowRank, typename Dense>
void to_dense(const Eigen::MatrixBase<Input> &t, const Eigen::MatrixBase<Coeffs> &c, const Eigen::MatrixBase<Diag> &a,
              const Eigen::MatrixBase<LowRank> &U, const Eigen::MatrixBase<LowRank> &V, Eigen::MatrixBase<Dense> const &K_out) {
  using namespace b200;
  const Eigen::Index N = U.rows(), J = U.cols();
  DeviceArray t_(t), c_(c), a_(a), U_(U), V_(V), K_((size_t)(N * N));
  check(c2b_to_dense(N, (int32_t)J, t_.ptr(), c_.ptr(), a_.ptr(), U_.ptr(), V_.ptr(), K_.ptr(), nullptr));
  check(c2b_stream_synchronize(nullptr));
  K_.fetch(K_out, N, N);
}

// ---- factor --------------------------------------------------------------------------------------------------
template <bool update_workspace = true, typename Input, typename Coeffs, typename Diag, typename LowRank, typename DiagOut,
          typename LowRankOut, typename Work>
Eigen::Index factor(const Eigen::MatrixBase<Input> &t, const Eigen::MatrixBase<Coeffs> &c, const Eigen::MatrixBase<Diag> &a,
                    const Eigen::MatrixBase<LowRank> &U, const Eigen::MatrixBase<LowRank> &V, Eigen::MatrixBase<DiagOut> const &d_out,
                    Eigen::MatrixBase<LowRankOut> const &W_out, Eigen::MatrixBase<Work> const &S_out) {
  using namespace b200;
  const Eigen::Index N = U.rows(), J = U.cols();
  DeviceArray t_(t), c_(c), a_(a), U_(U), V_(V), d_((size_t)N), W_((size_t)(N * J));
  DeviceArray S_(update_workspace ? (size_t)(N * J * J) : 1), status_(1);
  check(c2b_factor(N, (int32_t)J, 1, t_.ptr(), c_.ptr(), a_.ptr(), U_.ptr(), V_.ptr(), d_.ptr(), W_.ptr(),
                   update_workspace ? S_.ptr() : nullptr, (int64_t *)status_.ptr(), 0, nullptr, 0, nullptr));
  check(c2b_stream_synchronize(nullptr));
  int64_t flag = 0;
  check(c2b_copy_to_host(&flag, status_.ptr(), sizeof(flag)));
  if (flag != 0) {
    // the reference's failure semantics (forward.hpp:107-128): d is written up to and including the failing row, W up
    // to the row before it, S up to and including it; later rows of the caller's arrays are not touched.  (A long
    // series runs as several chunks on the device, where rows after the failure hold unspecified values: they never
    // leave the device here.)
    d_.fetch_rows(d_out, N, 1, (Eigen::Index)flag + 1);
    W_.fetch_rows(W_out, N, J, (Eigen::Index)flag);
    if (update_workspace) S_.fetch_rows(S_out, N, J * J, (Eigen::Index)flag + 1);
    return (Eigen::Index)flag;
  }
  d_.fetch(d_out, N, 1);
  W_.fetch(W_out, N, J);
  if (update_workspace) S_.fetch(S_out, N, J * J);
  return 0;
}

template <typename Input, typename Coeffs, typename Diag, typename LowRank, typename DiagOut, typename LowRankOut>
Eigen::Index factor(const Eigen::MatrixBase<Input> &t, const Eigen::MatrixBase<Coeffs> &c, const Eigen::MatrixBase<Diag> &a,
                    const Eigen::MatrixBase<LowRank> &U, const Eigen::MatrixBase<LowRank> &V, Eigen::MatrixBase<DiagOut> const &d_out,
                    Eigen::MatrixBase<LowRankOut> const &W_out) {
  b200::RowMat S;
  return factor<false>(t, c, a, U, V, d_out, W_out, S);
}

// ---- the four sweeps -----------------------------------------------------------------------------------------------
namespace detail {
typedef int (*sweep_fn)(int64_t, int32_t, int32_t, int64_t, const double *, const double *, const double *, const double *,
                        const double *, double *, double *, int32_t, void *, size_t, void *);

template <bool update_workspace, bool accumulate, typename Input, typename Coeffs, typename LowRank, typename RHS, typename RHSOut,
          typename Work>
void sweep(sweep_fn fn, const Eigen::MatrixBase<Input> &t, const Eigen::MatrixBase<Coeffs> &c, const Eigen::MatrixBase<LowRank> &U,
           const Eigen::MatrixBase<LowRank> &W, const Eigen::MatrixBase<RHS> &Y, Eigen::MatrixBase<RHSOut> const &Z_out,
           Eigen::MatrixBase<Work> const &F_out) {
  using namespace b200;
  const Eigen::Index N = U.rows(), J = U.cols(), K = Y.cols();
  DeviceArray t_(t), c_(c), U_(U), W_(W), Y_(Y), F_(update_workspace ? (size_t)(N * J * K) : 1);
  // matmuls accumulate into the caller's Z (forward.hpp:238,270); solves overwrite it
  RowMat Z0 = RowMat::Zero(N, K);
  if (accumulate) Z0 = Z_out;
  DeviceArray Z_(Z0);
  check(fn(N, (int32_t)J, (int32_t)K, 1, t_.ptr(), c_.ptr(), U_.ptr(), W_.ptr(), Y_.ptr(), Z_.ptr(), update_workspace ? F_.ptr() : nullptr,
           0, nullptr, 0, nullptr));
  check(c2b_stream_synchronize(nullptr));
  Z_.fetch(Z_out, N, K);
  if (update_workspace) F_.fetch(F_out, N, J * K);
}
}  // namespace detail

#define C2B_DEFINE_SWEEP(NAME, ACC)                                                                                                       \
  template <bool update_workspace = true, typename Input, typename Coeffs, typename LowRank, typename RightHandSide,                      \
            typename RightHandSideOut, typename Work>                                                                                     \
  void NAME(const Eigen::MatrixBase<Input> &t, const Eigen::MatrixBase<Coeffs> &c, const Eigen::MatrixBase<LowRank> &U,                   \
            const Eigen::MatrixBase<LowRank> &W, const Eigen::MatrixBase<RightHandSide> &Y,                                               \
            Eigen::MatrixBase<RightHandSideOut> const &Z_out, Eigen::MatrixBase<Work> const &F_out) {                                     \
    detail::sweep<update_workspace, ACC>(c2b_##NAME, t, c, U, W, Y, Z_out, F_out);                                                        \
  }                                                                                                                                       \
  template <typename Input, typename Coeffs, typename LowRank, typename RightHandSide, typename RightHandSideOut>                         \
  void NAME(const Eigen::MatrixBase<Input> &t, const Eigen::MatrixBase<Coeffs> &c, const Eigen::MatrixBase<LowRank> &U,                   \
            const Eigen::MatrixBase<LowRank> &W, const Eigen::MatrixBase<RightHandSide> &Y,                                               \
            Eigen::MatrixBase<RightHandSideOut> const &Z_out) {                                                                           \
    b200::RowMat F;                                                                                                                       \
    detail::sweep<false, ACC>(c2b_##NAME, t, c, U, W, Y, Z_out, F);                                                                       \
  }
C2B_DEFINE_SWEEP(solve_lower, false)
C2B_DEFINE_SWEEP(solve_upper, false)
C2B_DEFINE_SWEEP(matmul_lower, true)
C2B_DEFINE_SWEEP(matmul_upper, true)
#undef C2B_DEFINE_SWEEP

// (continued below: general_matmul_*, reverse mode)

// ---- rectangular products ------------------------------------------------------------------------------------------
namespace detail {
typedef int (*gm_fn)(int64_t, int64_t, int32_t, int32_t, const double *, const double *, const double *, const double *, const double *,
                     const double *, double *, double *, void *, size_t, void *);

template <bool do_update, typename Input, typename Coeffs, typename LowRank, typename RHS, typename RHSOut, typename Work>
void general_matmul(gm_fn fn, const Eigen::MatrixBase<Input> &t1, const Eigen::MatrixBase<Input> &t2, const Eigen::MatrixBase<Coeffs> &c,
                    const Eigen::MatrixBase<LowRank> &U, const Eigen::MatrixBase<LowRank> &V, const Eigen::MatrixBase<RHS> &Y,
                    Eigen::MatrixBase<RHSOut> const &Z_out, Eigen::MatrixBase<Work> const &F_out) {
  using namespace b200;
  const Eigen::Index N = t1.rows(), M = t2.rows(), J = c.rows(), K = Y.cols();
  DeviceArray t1_(t1), t2_(t2), c_(c), U_(U), V_(V), Y_(Y);
  RowMat Z0 = Z_out;  // accumulated into (forward.hpp:330,390)
  DeviceArray Z_(Z0);
  RowMat F0 = RowMat::Zero(M, J * K);
  DeviceArray F_(F0);
  check(fn(N, M, (int32_t)J, (int32_t)K, t1_.ptr(), t2_.ptr(), c_.ptr(), U_.ptr(), V_.ptr(), Y_.ptr(), Z_.ptr(), do_update ? F_.ptr() : nullptr,
           nullptr, 0, nullptr));
  check(c2b_stream_synchronize(nullptr));
  Z_.fetch(Z_out, N, K);
  if (do_update) F_.fetch(F_out, M, J * K);
}
}  // namespace detail

#define C2B_DEFINE_GM(NAME)                                                                                                               \
  template <bool do_update = true, typename Input, typename Coeffs, typename LowRank, typename RightHandSide, typename RightHandSideOut,  \
            typename Work>                                                                                                                \
  void NAME(const Eigen::MatrixBase<Input> &t1, const Eigen::MatrixBase<Input> &t2, const Eigen::MatrixBase<Coeffs> &c,                   \
            const Eigen::MatrixBase<LowRank> &U, const Eigen::MatrixBase<LowRank> &V, const Eigen::MatrixBase<RightHandSide> &Y,          \
            Eigen::MatrixBase<RightHandSideOut> const &Z_out, Eigen::MatrixBase<Work> const &F_out) {                                     \
    detail::general_matmul<do_update>(c2b_##NAME, t1, t2, c, U, V, Y, Z_out, F_out);                                                      \
  }                                                                                                                                       \
  template <typename Input, typename Coeffs, typename LowRank, typename RightHandSide, typename RightHandSideOut>                         \
  void NAME(const Eigen::MatrixBase<Input> &t1, const Eigen::MatrixBase<Input> &t2, const Eigen::MatrixBase<Coeffs> &c,                   \
            const Eigen::MatrixBase<LowRank> &U, const Eigen::MatrixBase<LowRank> &V, const Eigen::MatrixBase<RightHandSide> &Y,          \
            Eigen::MatrixBase<RightHandSideOut> const &Z_out) {                                                                           \
    b200::RowMat F;                                                                                                                       \
    detail::general_matmul<false>(c2b_##NAME, t1, t2, c, U, V, Y, Z_out, F);                                                              \
  }
C2B_DEFINE_GM(general_matmul_lower)
C2B_DEFINE_GM(general_matmul_upper)
#undef C2B_DEFINE_GM

// ---- reverse mode --------------------------------------------------------------------------------------------------
template <typename Input, typename Coeffs, typename Diag, typename LowRank, typename Work, typename InputOut, typename CoeffsOut,
          typename DiagOut, typename LowRankOut>
void factor_rev(const Eigen::MatrixBase<Input> &t, const Eigen::MatrixBase<Coeffs> &c, const Eigen::MatrixBase<Diag> &a,
                const Eigen::MatrixBase<LowRank> &U, const Eigen::MatrixBase<LowRank> &V, const Eigen::MatrixBase<Diag> &d,
                const Eigen::MatrixBase<LowRank> &W, const Eigen::MatrixBase<Work> &S, const Eigen::MatrixBase<Diag> &bd,
                const Eigen::MatrixBase<LowRank> &bW, Eigen::MatrixBase<InputOut> const &bt_out, Eigen::MatrixBase<CoeffsOut> const &bc_out,
                Eigen::MatrixBase<DiagOut> const &ba_out, Eigen::MatrixBase<LowRankOut> const &bU_out,
                Eigen::MatrixBase<LowRankOut> const &bV_out) {
  using namespace b200;
  const Eigen::Index N = U.rows(), J = U.cols();
  DeviceArray t_(t), c_(c), a_(a), U_(U), V_(V), d_(d), W_(W), S_(S), bd_(bd), bW_(bW);
  DeviceArray bt_((size_t)N), bc_((size_t)J), ba_((size_t)N), bU_((size_t)(N * J)), bV_((size_t)(N * J));
  check(c2b_factor_rev(N, (int32_t)J, 1, t_.ptr(), c_.ptr(), a_.ptr(), U_.ptr(), V_.ptr(), d_.ptr(), W_.ptr(), S_.ptr(), bd_.ptr(), bW_.ptr(),
                       bt_.ptr(), bc_.ptr(), ba_.ptr(), bU_.ptr(), bV_.ptr(), 0, nullptr, 0, nullptr));
  check(c2b_stream_synchronize(nullptr));
  bt_.fetch(bt_out, N, 1);
  bc_.fetch(bc_out, J, 1);
  ba_.fetch(ba_out, N, 1);
  bU_.fetch(bU_out, N, J);
  bV_.fetch(bV_out, N, J);
}

namespace detail {
typedef int (*sweep_rev_fn)(int64_t, int32_t, int32_t, int64_t, const double *, const double *, const double *, const double *,
                            const double *, const double *, const double *, const double *, double *, double *, double *, double *, double *,
                            int32_t, void *, size_t, void *);
}

#define C2B_DEFINE_SWEEP_REV(NAME)                                                                                                        \
  template <typename Input, typename Coeffs, typename LowRank, typename RightHandSide, typename Work, typename InputOut,                  \
            typename CoeffsOut, typename LowRankOut, typename RightHandSideOut>                                                           \
  void NAME(const Eigen::MatrixBase<Input> &t, const Eigen::MatrixBase<Coeffs> &c, const Eigen::MatrixBase<LowRank> &U,                   \
            const Eigen::MatrixBase<LowRank> &W, const Eigen::MatrixBase<RightHandSide> &Y, const Eigen::MatrixBase<RightHandSide> &Z,    \
            const Eigen::MatrixBase<Work> &F, const Eigen::MatrixBase<RightHandSide> &bZ, Eigen::MatrixBase<InputOut> const &bt_out,      \
            Eigen::MatrixBase<CoeffsOut> const &bc_out, Eigen::MatrixBase<LowRankOut> const &bU_out,                                      \
            Eigen::MatrixBase<LowRankOut> const &bW_out, Eigen::MatrixBase<RightHandSideOut> const &bY_out) {                             \
    using namespace b200;                                                                                                                 \
    const Eigen::Index N = U.rows(), J = U.cols(), K = Y.cols();                                                                          \
    DeviceArray t_(t), c_(c), U_(U), W_(W), Y_(Y), Z_(Z), F_(F), bZ_(bZ);                                                                 \
    DeviceArray bt_((size_t)N), bc_((size_t)J), bU_((size_t)(N * J)), bW_((size_t)(N * J)), bY_((size_t)(N * K));                         \
    check(c2b_##NAME(N, (int32_t)J, (int32_t)K, 1, t_.ptr(), c_.ptr(), U_.ptr(), W_.ptr(), Y_.ptr(), Z_.ptr(), F_.ptr(), bZ_.ptr(),       \
                     bt_.ptr(), bc_.ptr(), bU_.ptr(), bW_.ptr(), bY_.ptr(), 0, nullptr, 0, nullptr));                                     \
    check(c2b_stream_synchronize(nullptr));                                                                                               \
    bt_.fetch(bt_out, N, 1);                                                                                                              \
    bc_.fetch(bc_out, J, 1);                                                                                                              \
    bU_.fetch(bU_out, N, J);                                                                                                              \
    bW_.fetch(bW_out, N, J);                                                                                                              \
    bY_.fetch(bY_out, N, K);                                                                                                              \
  }
C2B_DEFINE_SWEEP_REV(solve_lower_rev)
C2B_DEFINE_SWEEP_REV(solve_upper_rev)
C2B_DEFINE_SWEEP_REV(matmul_lower_rev)
C2B_DEFINE_SWEEP_REV(matmul_upper_rev)
#undef C2B_DEFINE_SWEEP_REV

}  // namespace core
}  // namespace celerite2

#endif  // CELERITE2_B200_CORE_HPP_
