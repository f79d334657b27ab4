// Kernel terms under the reference's name and interface (c++/include/celerite2/terms.hpp): Term<T, J> with
// set_coefficients / get_coefficients / get_celerite_matrices / operator+, RealTerm, ComplexTerm, SHOTerm.  The O(J)
// coefficient algebra stays on the host, as in the reference; get_celerite_matrices (terms.hpp:112-139, the N x J sin/cos
// evaluation) runs on the GPU through c2b_get_celerite_matrices (blocked column order [real.. | cos.. | sin..]).
//
// One deliberate difference: operator+ concatenates bc, cc, dc of the two operands.  The reference's operator+
// (terms.hpp:160-162) copies ac into all three, which only its own self-consistent C++ tests tolerate.
#ifndef CELERITE2_B200_SHIM_TERMS_HPP_
#define CELERITE2_B200_SHIM_TERMS_HPP_

#include <Eigen/Core>
#include <algorithm>
#include <cmath>
#include <exception>
#include <tuple>
#include <type_traits>

#include "../celerite2_b200/core.hpp"

namespace celerite2 {

#ifndef CELERITE_MAX_WIDTH
#define CELERITE_MAX_WIDTH 32
#endif

struct dimension_mismatch : public std::exception {
  const char *what() const throw() { return "dimension mismatch"; }
};

template <int J1, int J2>
struct sum_width {
  constexpr static int value = (J1 == Eigen::Dynamic || J2 == Eigen::Dynamic) ? Eigen::Dynamic : (J1 + J2);
};

template <typename T, int J_ = Eigen::Dynamic>
class Term {
  static_assert(std::is_same<T, double>::value, "the GPU path is FP64 only");

  public:
  constexpr static int Width = ((0 < J_) && (J_ <= CELERITE_MAX_WIDTH)) ? J_ : Eigen::Dynamic;
  static constexpr int Order = (Width != 1) ? Eigen::RowMajor : Eigen::ColMajor;
  typedef T Scalar;
  typedef Eigen::Matrix<Scalar, Eigen::Dynamic, 1> Vector;
  typedef Eigen::Matrix<Scalar, Eigen::Dynamic, Width, Order> LowRank;
  typedef Eigen::Matrix<Scalar, Width, 1> CoeffVector;
  typedef std::tuple<Vector, Vector, Vector, Vector, Vector, Vector> Coeffs;
  typedef std::tuple<CoeffVector, Vector, LowRank, LowRank> Matrices;

  Term() {}
  int get_width() const { return Width; }

  void set_coefficients(const Vector &ar, const Vector &cr, const Vector &ac, const Vector &bc, const Vector &cc, const Vector &dc) {
    if (cr.rows() != ar.rows() || bc.rows() != ac.rows() || cc.rows() != ac.rows() || dc.rows() != ac.rows()) throw dimension_mismatch();
    ar_ = ar;
    cr_ = cr;
    ac_ = ac;
    bc_ = bc;
    cc_ = cc;
    dc_ = dc;
  }
  Coeffs get_coefficients() const { return std::make_tuple(ar_, cr_, ac_, bc_, cc_, dc_); }

  // (c, a, U, V) for the inputs x with the extra diagonal `diag`, evaluated on the GPU
  Matrices get_celerite_matrices(const Vector &x, const Vector &diag) const {
    using namespace b200;
    const Eigen::Index N = x.rows();
    if (diag.rows() != N) throw dimension_mismatch();
    const Eigen::Index nr = ar_.rows(), nc = ac_.rows(), J = nr + 2 * nc;
    if (Width != Eigen::Dynamic && Width != J) throw dimension_mismatch();
    if (N < 1 || J < 1) throw dimension_mismatch();
    DeviceArray ar(ar_), cr(cr_), ac(ac_), bc(bc_), cc(cc_), dc(dc_);  // (an empty vector still gets a valid device pointer)
    DeviceArray x_(x), diag_(diag), c_((size_t)J), a_((size_t)N), U_((size_t)(N * J)), V_((size_t)(N * J));
    check(c2b_get_celerite_matrices(0, N, (int32_t)nr, (int32_t)nc, 1, ar.ptr(), cr.ptr(), ac.ptr(), bc.ptr(), cc.ptr(), dc.ptr(),
                                    x_.ptr(), diag_.ptr(), c_.ptr(), a_.ptr(), U_.ptr(), V_.ptr(), 0, nullptr));
    check(c2b_stream_synchronize(nullptr));
    CoeffVector c(J);
    Vector a(N);
    LowRank U(N, J), V(N, J);
    c_.fetch(c, J, 1);
    a_.fetch(a, N, 1);
    U_.fetch(U, N, J);
    V_.fetch(V, N, J);
    return std::make_tuple(c, a, U, V);
  }

  template <typename Other>
  Term<typename std::common_type<Scalar, typename Other::Scalar>::type, sum_width<Width, Other::Width>::value> operator+(const Other &other) const {
    typedef typename std::common_type<Scalar, typename Other::Scalar>::type NewScalar;
    typedef Eigen::Matrix<NewScalar, Eigen::Dynamic, 1> NewVector;
    const auto o = other.get_coefficients();
    auto cat     = [](const Vector &p, const typename Other::Vector &q) {
      NewVector r(p.rows() + q.rows());
      r << p, q;
      return r;
    };
    Term<NewScalar, sum_width<Width, Other::Width>::value> sum;
    sum.set_coefficients(cat(ar_, std::get<0>(o)), cat(cr_, std::get<1>(o)), cat(ac_, std::get<2>(o)), cat(bc_, std::get<3>(o)),
                         cat(cc_, std::get<4>(o)), cat(dc_, std::get<5>(o)));
    return sum;
  }

  private:
  Vector ar_, cr_, ac_, bc_, cc_, dc_;
};

// k(tau) = a exp(-c tau)                                                        (terms.hpp:181-198)
template <typename T>
class RealTerm : public Term<T, 1> {
  public:
  typedef T Scalar;
  constexpr static int Width = 1;
  using typename Term<Scalar, 1>::Vector;
  using typename Term<Scalar, 1>::LowRank;
  RealTerm(const Scalar &a, const Scalar &c) {
    Vector ar(1), cr(1), e;
    ar << a;
    cr << c;
    this->set_coefficients(ar, cr, e, e, e, e);
  }
};

// k(tau) = exp(-c tau) (a cos(d tau) + b sin(d tau))                           (terms.hpp:209-228)
template <typename T>
class ComplexTerm : public Term<T, 2> {
  public:
  typedef T Scalar;
  constexpr static int Width = 2;
  using typename Term<Scalar, 2>::Vector;
  using typename Term<Scalar, 2>::LowRank;
  ComplexTerm(const Scalar &a, const Scalar &b, const Scalar &c, const Scalar &d) {
    Vector e, ac(1), bc(1), cc(1), dc(1);
    ac << a;
    bc << b;
    cc << c;
    dc << d;
    this->set_coefficients(e, e, ac, bc, cc, dc);
  }
};

// stochastically driven damped harmonic oscillator (terms.hpp:239-277): two real terms when overdamped (Q < 1/2), one
// complex term otherwise; eps regularises the critically damped case
template <typename T>
class SHOTerm : public Term<T, 2> {
  public:
  typedef T Scalar;
  constexpr static int Width = 2;
  using typename Term<Scalar, 2>::Vector;
  using typename Term<Scalar, 2>::LowRank;
  SHOTerm(const Scalar &S0, const Scalar &w0, const Scalar &Q, const Scalar &eps = 1e-5) {
    Vector e;
    if (Q < 0.5) {
      const Scalar f = std::sqrt(std::max(Scalar(1.0 - 4.0 * Q * Q), eps));
      const Scalar amp = 0.5 * S0 * w0 * Q, rate = 0.5 * w0 / Q;
      Vector ar(2), cr(2);
      ar << amp * (1 + 1 / f), amp * (1 - 1 / f);
      cr << rate * (1 - f), rate * (1 + f);
      this->set_coefficients(ar, cr, e, e, e, e);
    } else {
      const Scalar f = std::sqrt(std::max(Scalar(4.0 * Q * Q - 1), eps));
      const Scalar amp = S0 * w0 * Q, rate = 0.5 * w0 / Q;
      Vector ac(1), bc(1), cc(1), dc(1);
      ac << amp;
      bc << amp / f;
      cc << rate;
      dc << rate * f;
      this->set_coefficients(e, e, ac, bc, cc, dc);
    }
  }
};

}  // namespace celerite2

#endif  // CELERITE2_B200_SHIM_TERMS_HPP_
