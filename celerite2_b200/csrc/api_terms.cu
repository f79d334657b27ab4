// C ABI: kernel-term algebra on the device: hyper-parameters -> the six coefficient vectors
// (ar, cr, ac, bc, cc, dc) that the assembly consumes, for a BATCH of parameter sets that already live on the GPU (a
// sampler's proposals, an optimiser's population), and the adjoint maps back to the parameters, so that a whole
// log-likelihood + gradient evaluation with respect to the physical parameters never touches the host.
//
// Formulas restated from python/celerite2/terms.py: RealTerm 485-521, ComplexTerm 524-569, SHOTerm 604-691,
// Matern32Term 694-745, RotationTerm 748-814, TermSum 210-235 (concatenation: the off_* arguments), TermProduct 238-300.
// One thread per parameter set; O(J) work each, exactly what the reference does per call in Python.
#include "host_common.cuh"

namespace c2b {

// where a term writes inside the coefficient arrays of the whole kernel: row b of ar / cr has Jr_tot entries, this term
// owns [r_off, r_off + its real terms); the same for the four complex vectors with Jc_tot / c_off
struct TermSlots {
  int Jr_tot, Jc_tot, r_off, c_off;
  double *ar, *cr, *ac, *bc, *cc, *dc;              // forward: outputs.  adjoint: the cotangents (read)
};

C2B_DEV void put_real(const TermSlots &s, int64_t b, int i, double a, double c) {
  s.ar[b * s.Jr_tot + s.r_off + i] = a;
  s.cr[b * s.Jr_tot + s.r_off + i] = c;
}
C2B_DEV void put_complex(const TermSlots &s, int64_t b, int i, double a, double bb, double c, double d) {
  const int64_t o = b * s.Jc_tot + s.c_off + i;
  s.ac[o] = a;
  s.bc[o] = bb;
  s.cc[o] = c;
  s.dc[o] = d;
}

// ---- SHOTerm (terms.py:658-691) ------------------------------------------------------------------------------------
// underdamped (Q >= 1/2): one complex term; overdamped: two real terms.  A batch must be all of one kind (the shapes of
// the coefficient arrays differ); `under` says which, samples on the wrong side of 1/2 are reported through `bad`.
C2B_DEV void sho_fwd(const TermSlots &s, int64_t b, int slot, bool under, double S0, double w0, double Q, double eps) {
  if (under) {
    const double f = sqrt(fmax(4.0 * Q * Q - 1.0, eps));
    const double a = S0 * w0 * Q, c = 0.5 * w0 / Q;
    put_complex(s, b, slot, a, a / f, c, c * f);
  } else {
    const double f = sqrt(fmax(1.0 - 4.0 * Q * Q, eps));
    const double a = 0.5 * S0 * w0 * Q, c = 0.5 * w0 / Q;
    put_real(s, b, 2 * slot, a * (1.0 + 1.0 / f), c * (1.0 - f));
    put_real(s, b, 2 * slot + 1, a * (1.0 - 1.0 / f), c * (1.0 + f));
  }
}
// cotangents of the term's coefficients -> cotangents of (S0, w0, Q)
C2B_DEV void sho_rev(const TermSlots &s, int64_t b, int slot, bool under, double S0, double w0, double Q, double eps, double *bS0,
                     double *bw0, double *bQ) {
  if (under) {
    const double x  = 4.0 * Q * Q - 1.0;
    const double f  = sqrt(fmax(x, eps));
    const double fQ = x > eps ? 4.0 * Q / f : 0.0;
    const double a = S0 * w0 * Q, c = 0.5 * w0 / Q;
    const int64_t o = b * s.Jc_tot + s.c_off + slot;
    const double ga = s.ac[o], gb = s.bc[o], gc = s.cc[o], gd = s.dc[o];
    // ac = a, bc = a / f, cc = c, dc = c f
    const double A  = ga + gb / f;            // d / da
    const double C  = gc + gd * f;            // d / dc
    const double F  = -gb * a / (f * f) + gd * c;  // d / df
    *bS0 = A * w0 * Q;
    *bw0 = A * S0 * Q + C * 0.5 / Q;
    *bQ  = A * S0 * w0 - C * 0.5 * w0 / (Q * Q) + F * fQ;
  } else {
    const double x  = 1.0 - 4.0 * Q * Q;
    const double f  = sqrt(fmax(x, eps));
    const double fQ = x > eps ? -4.0 * Q / f : 0.0;
    const double a = 0.5 * S0 * w0 * Q, c = 0.5 * w0 / Q;
    const int64_t o = b * s.Jr_tot + s.r_off + 2 * slot;
    const double ga0 = s.ar[o], ga1 = s.ar[o + 1], gc0 = s.cr[o], gc1 = s.cr[o + 1];
    // ar = a (1 +- 1/f), cr = c (1 -+ f)
    const double A = ga0 * (1.0 + 1.0 / f) + ga1 * (1.0 - 1.0 / f);
    const double C = gc0 * (1.0 - f) + gc1 * (1.0 + f);
    const double F = (-ga0 + ga1) * a / (f * f) + (-gc0 + gc1) * c;
    *bS0 = A * 0.5 * w0 * Q;
    *bw0 = A * 0.5 * S0 * Q + C * 0.5 / Q;
    *bQ  = A * 0.5 * S0 * w0 - C * 0.5 * w0 / (Q * Q) + F * fQ;
  }
}

enum TermKind { kReal = 0, kComplex = 1, kSHO = 2, kMatern32 = 3, kRotation = 4 };

// frequency and power of one mode of the rotation term (terms.py:796-806): w = k pi Q / (P sqrt(4 Q^2 - 1)), S = A / (w Q)
struct RotMode {
  double Q, w, S, s;  // s = sqrt(4 Q^2 - 1)
};
C2B_DEV RotMode rot_mode(double k, double Q, double period, double A) {
  RotMode m;
  m.Q = Q;
  m.s = sqrt(4.0 * Q * Q - 1.0);
  m.w = k * 3.14159265358979323846 * Q / (period * m.s);
  m.S = A / (m.w * Q);
  return m;
}

// params: (batch, P) row-major; P = 2 (real: a, c), 4 (complex: a, b, c, d), 3 (SHO: S0, w0, Q), 2 (Matern-3/2: sigma, rho),
// 5 (rotation: sigma, period, Q0, dQ, f)
static __global__ void __launch_bounds__(128) term_fwd_kernel(int kind, int64_t batch, const double *params, double eps, int under,
                                                              TermSlots s, int32_t *bad) {
  pdl_prologue();
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  switch (kind) {
    case kReal: {
      const double *p = params + b * 2;
      put_real(s, b, 0, p[0], p[1]);
    } break;
    case kComplex: {
      const double *p = params + b * 4;
      put_complex(s, b, 0, p[0], p[1], p[2], p[3]);
    } break;
    case kSHO: {
      const double *p = params + b * 3;
      if (bad && ((p[2] < 0.5) == (under != 0))) atomicOr(bad, 1);
      sho_fwd(s, b, 0, under != 0, p[0], p[1], p[2], eps);
    } break;
    case kMatern32: {
      const double *p = params + b * 2;
      const double w0 = sqrt(3.0) / p[1], sig2 = p[0] * p[0];
      put_complex(s, b, 0, sig2, w0 * sig2 / eps, w0, eps);  // w0 S0 = sigma^2, w0^2 S0 / eps (terms.py:733-745)
    } break;
    case kRotation: {
      const double *p = params + b * 5;
      const double sigma = p[0], period = p[1], Q0 = p[2], dQ = p[3], f = p[4];
      const double amp = sigma * sigma / (1.0 + f);
      const RotMode m1 = rot_mode(4.0, 0.5 + Q0 + dQ, period, amp), m2 = rot_mode(8.0, 0.5 + Q0, period, f * amp);
      if (bad && (m1.Q < 0.5 || m2.Q < 0.5)) atomicOr(bad, 1);
      sho_fwd(s, b, 0, true, m1.S, m1.w, m1.Q, 1e-5);  // the two SHO terms use their default eps (terms.py:658)
      sho_fwd(s, b, 1, true, m2.S, m2.w, m2.Q, 1e-5);
    } break;
  }
}

static __global__ void __launch_bounds__(128) term_rev_kernel(int kind, int64_t batch, const double *params, double eps, int under,
                                                              TermSlots s, double *bparams) {
  pdl_prologue();
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  switch (kind) {
    case kReal: {
      bparams[b * 2]     = s.ar[b * s.Jr_tot + s.r_off];
      bparams[b * 2 + 1] = s.cr[b * s.Jr_tot + s.r_off];
    } break;
    case kComplex: {
      const int64_t o    = b * s.Jc_tot + s.c_off;
      bparams[b * 4]     = s.ac[o];
      bparams[b * 4 + 1] = s.bc[o];
      bparams[b * 4 + 2] = s.cc[o];
      bparams[b * 4 + 3] = s.dc[o];
    } break;
    case kSHO: {
      const double *p = params + b * 3;
      sho_rev(s, b, 0, under != 0, p[0], p[1], p[2], eps, bparams + b * 3, bparams + b * 3 + 1, bparams + b * 3 + 2);
    } break;
    case kMatern32: {
      const double *p = params + b * 2;
      const int64_t o = b * s.Jc_tot + s.c_off;
      const double w0 = sqrt(3.0) / p[1], sig2 = p[0] * p[0], w0r = -w0 / p[1];  // d w0 / d rho
      const double ga = s.ac[o], gb = s.bc[o], gc = s.cc[o];
      bparams[b * 2]     = 2.0 * p[0] * (ga + gb * w0 / eps);
      bparams[b * 2 + 1] = (gb * sig2 / eps + gc) * w0r;
    } break;
    case kRotation: {
      const double *p = params + b * 5;
      const double sigma = p[0], period = p[1], Q0 = p[2], dQ = p[3], f = p[4];
      const double amp = sigma * sigma / (1.0 + f);
      const double A[2] = {amp, f * amp}, kk[2] = {4.0, 8.0}, Qm[2] = {0.5 + Q0 + dQ, 0.5 + Q0};
      double bsig = 0.0, bper = 0.0, bQ0 = 0.0, bdQ = 0.0, bf = 0.0;
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const RotMode m = rot_mode(kk[i], Qm[i], period, A[i]);
        double gS, gw, gQ;
        sho_rev(s, b, i, true, m.S, m.w, m.Q, 1e-5, &gS, &gw, &gQ);
        // S = A / (w Q): dS/dA = 1/(w Q), dS/dw = -S/w, dS/dQ = -S/Q;  w(Q, P): dw/dQ = -k pi / (P s^3), dw/dP = -w/P
        const double gA  = gS / (m.w * m.Q);
        const double gwt = gw - gS * m.S / m.w;                 // total cotangent of w
        const double gQt = gQ - gS * m.S / m.Q - gwt * kk[i] * 3.14159265358979323846 / (period * m.s * m.s * m.s);
        bper -= gwt * m.w / period;
        bQ0 += gQt;
        if (i == 0) bdQ += gQt;
        // A_0 = amp, A_1 = f amp; amp = sigma^2 / (1 + f)
        const double gamp = i == 0 ? gA : gA * f;
        bsig += gamp * 2.0 * sigma / (1.0 + f);
        bf += -gamp * amp / (1.0 + f) + (i == 1 ? gA * amp : 0.0);
      }
      bparams[b * 5]     = bsig;
      bparams[b * 5 + 1] = bper;
      bparams[b * 5 + 2] = bQ0;
      bparams[b * 5 + 3] = bdQ;
      bparams[b * 5 + 4] = bf;
    } break;
  }
}

static int term_shape(int kind, int under, int *P, int *nr, int *nc) {
  switch (kind) {
    case kReal: *P = 2; *nr = 1; *nc = 0; return C2B_OK;
    case kComplex: *P = 4; *nr = 0; *nc = 1; return C2B_OK;
    case kSHO: *P = 3; *nr = under ? 0 : 2; *nc = under ? 1 : 0; return C2B_OK;
    case kMatern32: *P = 2; *nr = 0; *nc = 1; return C2B_OK;
    case kRotation: *P = 5; *nr = 0; *nc = 2; return C2B_OK;
  }
  set_error("unknown term kind %d", kind);
  return C2B_ERR_INVALID_ARGUMENT;
}

}  // namespace c2b

namespace c2b {

// ---- TermProduct (terms.py:238-300): coefficients of k1(tau) k2(tau) ---------------------------------------------------
// out real:    Jr1 Jr2            (aj ak, cj + ck), term-1 index outermost
// out complex: Jr1 Jc2 (real of 1 x complex of 2), then Jr2 Jc1 (real of 2 x complex of 1), then 2 Jc1 Jc2 (difference and
//              sum frequencies of every complex pair)
struct TermIn {
  int Jr, Jc;
  const double *ar, *cr, *ac, *bc, *cc, *dc;  // (batch, Jr) / (batch, Jc)
};
struct TermGrad {
  double *ar, *cr, *ac, *bc, *cc, *dc;
};

static __global__ void __launch_bounds__(128) term_product_kernel(int64_t batch, TermIn x, TermIn y, TermSlots s) {
  pdl_prologue();
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  const double *xar = x.ar + b * x.Jr, *xcr = x.cr + b * x.Jr, *xac = x.ac + b * x.Jc, *xbc = x.bc + b * x.Jc, *xcc = x.cc + b * x.Jc,
               *xdc = x.dc + b * x.Jc;
  const double *yar = y.ar + b * y.Jr, *ycr = y.cr + b * y.Jr, *yac = y.ac + b * y.Jc, *ybc = y.bc + b * y.Jc, *ycc = y.cc + b * y.Jc,
               *ydc = y.dc + b * y.Jc;
  int r = 0, c = 0;
  for (int j = 0; j < x.Jr; ++j)
    for (int k = 0; k < y.Jr; ++k) put_real(s, b, r++, xar[j] * yar[k], xcr[j] + ycr[k]);
  for (int j = 0; j < x.Jr; ++j)
    for (int k = 0; k < y.Jc; ++k) put_complex(s, b, c++, xar[j] * yac[k], xar[j] * ybc[k], xcr[j] + ycc[k], ydc[k]);
  for (int j = 0; j < y.Jr; ++j)
    for (int k = 0; k < x.Jc; ++k) put_complex(s, b, c++, yar[j] * xac[k], yar[j] * xbc[k], ycr[j] + xcc[k], xdc[k]);
  for (int j = 0; j < x.Jc; ++j)
    for (int k = 0; k < y.Jc; ++k) {
      const double aj = xac[j], bj = xbc[j], ak = yac[k], bk = ybc[k];
      put_complex(s, b, c++, 0.5 * (aj * ak + bj * bk), 0.5 * (bj * ak - aj * bk), xcc[j] + ycc[k], xdc[j] - ydc[k]);
      put_complex(s, b, c++, 0.5 * (aj * ak - bj * bk), 0.5 * (bj * ak + aj * bk), xcc[j] + ycc[k], xdc[j] + ydc[k]);
    }
}

// adjoint: cotangents of the product's coefficients (in s, read) -> cotangents of both factors (gx, gy: overwritten)
static __global__ void __launch_bounds__(128) term_product_rev_kernel(int64_t batch, TermIn x, TermIn y, TermSlots s, TermGrad gx,
                                                                      TermGrad gy) {
  pdl_prologue();
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  const double *xar = x.ar + b * x.Jr, *xac = x.ac + b * x.Jc, *xbc = x.bc + b * x.Jc;
  const double *yar = y.ar + b * y.Jr, *yac = y.ac + b * y.Jc, *ybc = y.bc + b * y.Jc;
  double *gxar = gx.ar + b * x.Jr, *gxcr = gx.cr + b * x.Jr, *gxac = gx.ac + b * x.Jc, *gxbc = gx.bc + b * x.Jc, *gxcc = gx.cc + b * x.Jc,
         *gxdc = gx.dc + b * x.Jc;
  double *gyar = gy.ar + b * y.Jr, *gycr = gy.cr + b * y.Jr, *gyac = gy.ac + b * y.Jc, *gybc = gy.bc + b * y.Jc, *gycc = gy.cc + b * y.Jc,
         *gydc = gy.dc + b * y.Jc;
  for (int j = 0; j < x.Jr; ++j) gxar[j] = gxcr[j] = 0.0;
  for (int j = 0; j < x.Jc; ++j) gxac[j] = gxbc[j] = gxcc[j] = gxdc[j] = 0.0;
  for (int j = 0; j < y.Jr; ++j) gyar[j] = gycr[j] = 0.0;
  for (int j = 0; j < y.Jc; ++j) gyac[j] = gybc[j] = gycc[j] = gydc[j] = 0.0;
  const double *par = s.ar + b * s.Jr_tot + s.r_off, *pcr = s.cr + b * s.Jr_tot + s.r_off;
  const double *pac = s.ac + b * s.Jc_tot + s.c_off, *pbc = s.bc + b * s.Jc_tot + s.c_off, *pcc = s.cc + b * s.Jc_tot + s.c_off,
               *pdc = s.dc + b * s.Jc_tot + s.c_off;
  int r = 0, c = 0;
  for (int j = 0; j < x.Jr; ++j)
    for (int k = 0; k < y.Jr; ++k, ++r) {
      gxar[j] += par[r] * yar[k];
      gyar[k] += par[r] * xar[j];
      gxcr[j] += pcr[r];
      gycr[k] += pcr[r];
    }
  for (int j = 0; j < x.Jr; ++j)
    for (int k = 0; k < y.Jc; ++k, ++c) {
      gxar[j] += pac[c] * yac[k] + pbc[c] * ybc[k];
      gyac[k] += pac[c] * xar[j];
      gybc[k] += pbc[c] * xar[j];
      gxcr[j] += pcc[c];
      gycc[k] += pcc[c];
      gydc[k] += pdc[c];
    }
  for (int j = 0; j < y.Jr; ++j)
    for (int k = 0; k < x.Jc; ++k, ++c) {
      gyar[j] += pac[c] * xac[k] + pbc[c] * xbc[k];
      gxac[k] += pac[c] * yar[j];
      gxbc[k] += pbc[c] * yar[j];
      gycr[j] += pcc[c];
      gxcc[k] += pcc[c];
      gxdc[k] += pdc[c];
    }
  for (int j = 0; j < x.Jc; ++j)
    for (int k = 0; k < y.Jc; ++k, c += 2) {
      const double aj = xac[j], bj = xbc[j], ak = yac[k], bk = ybc[k];
      const double a0 = 0.5 * pac[c], b0 = 0.5 * pbc[c], a1 = 0.5 * pac[c + 1], b1 = 0.5 * pbc[c + 1];
      // entry c:     ac = (aj ak + bj bk)/2, bc = (bj ak - aj bk)/2;  entry c+1: ac = (aj ak - bj bk)/2, bc = (bj ak + aj bk)/2
      gxac[j] += a0 * ak - b0 * bk + a1 * ak + b1 * bk;
      gxbc[j] += a0 * bk + b0 * ak - a1 * bk + b1 * ak;
      gyac[k] += a0 * aj + b0 * bj + a1 * aj + b1 * bj;
      gybc[k] += a0 * bj - b0 * aj - a1 * bj + b1 * aj;
      gxcc[j] += pcc[c] + pcc[c + 1];
      gycc[k] += pcc[c] + pcc[c + 1];
      gxdc[j] += pdc[c] + pdc[c + 1];
      gydc[k] += -pdc[c] + pdc[c + 1];
    }
}

}  // namespace c2b

extern "C" {

int c2b_term_coefficients(int32_t kind, int64_t batch, const double *params, double eps, int32_t underdamped, int32_t Jr_total,
                          int32_t Jc_total, int32_t r_off, int32_t c_off, double *ar, double *cr, double *ac, double *bc, double *cc,
                          double *dc, int32_t *bad, void *stream) {
  c2b::OpScope _op("c2b_term_coefficients");
  using namespace c2b;
  int P, nr, nc;
  int rc = term_shape(kind, underdamped, &P, &nr, &nc);
  if (rc) return rc;
  C2B_REQUIRE(batch >= 1 && params, "bad arguments");
  C2B_REQUIRE(r_off >= 0 && c_off >= 0 && r_off + nr <= Jr_total && c_off + nc <= Jc_total, "term does not fit its slots");
  C2B_REQUIRE((nr == 0 || (ar && cr)) && (nc == 0 || (ac && bc && cc && dc)), "null coefficient pointer");
  TermSlots s{Jr_total, Jc_total, r_off, c_off, ar, cr, ac, bc, cc, dc};
  C2B_LAUNCH("term_fwd_kernel", (cudaStream_t)stream,
             (launch_k(term_fwd_kernel, (unsigned)((batch + 127) / 128), 128, 0, (cudaStream_t)stream, kind, batch, params, eps, underdamped, s, bad)));
  return C2B_OK;
}

int c2b_term_coefficients_rev(int32_t kind, int64_t batch, const double *params, double eps, int32_t underdamped, int32_t Jr_total,
                              int32_t Jc_total, int32_t r_off, int32_t c_off, const double *b_ar, const double *b_cr, const double *b_ac,
                              const double *b_bc, const double *b_cc, const double *b_dc, double *b_params, void *stream) {
  c2b::OpScope _op("c2b_term_coefficients_rev");
  using namespace c2b;
  int P, nr, nc;
  int rc = term_shape(kind, underdamped, &P, &nr, &nc);
  if (rc) return rc;
  C2B_REQUIRE(batch >= 1 && params && b_params, "bad arguments");
  C2B_REQUIRE(r_off >= 0 && c_off >= 0 && r_off + nr <= Jr_total && c_off + nc <= Jc_total, "term does not fit its slots");
  C2B_REQUIRE((nr == 0 || (b_ar && b_cr)) && (nc == 0 || (b_ac && b_bc && b_cc && b_dc)), "null cotangent pointer");
  TermSlots s{Jr_total, Jc_total, r_off, c_off, (double *)b_ar, (double *)b_cr, (double *)b_ac, (double *)b_bc, (double *)b_cc,
              (double *)b_dc};
  C2B_LAUNCH("term_rev_kernel", (cudaStream_t)stream,
             (launch_k(term_rev_kernel, (unsigned)((batch + 127) / 128), 128, 0, (cudaStream_t)stream, kind, batch, params, eps, underdamped, s, b_params)));
  return C2B_OK;
}

int c2b_term_product(int64_t batch, int32_t Jr1, int32_t Jc1, const double *ar1, const double *cr1, const double *ac1, const double *bc1,
                     const double *cc1, const double *dc1, int32_t Jr2, int32_t Jc2, const double *ar2, const double *cr2,
                     const double *ac2, const double *bc2, const double *cc2, const double *dc2, double *ar, double *cr, double *ac,
                     double *bc, double *cc, double *dc, void *stream) {
  c2b::OpScope _op("c2b_term_product");
  using namespace c2b;
  C2B_REQUIRE(batch >= 1 && Jr1 >= 0 && Jc1 >= 0 && Jr2 >= 0 && Jc2 >= 0, "bad arguments");
  const int Jr = Jr1 * Jr2, Jc = Jr1 * Jc2 + Jr2 * Jc1 + 2 * Jc1 * Jc2;
  C2B_REQUIRE(Jr + 2 * Jc >= 1 && Jr + 2 * Jc <= C2B_MAX_J, "product too wide (J = Jr + 2 Jc must be in 1..32)");
  TermIn x{Jr1, Jc1, ar1, cr1, ac1, bc1, cc1, dc1}, y{Jr2, Jc2, ar2, cr2, ac2, bc2, cc2, dc2};
  TermSlots s{Jr, Jc, 0, 0, ar, cr, ac, bc, cc, dc};
  C2B_LAUNCH("term_product_kernel", (cudaStream_t)stream,
             (launch_k(term_product_kernel, (unsigned)((batch + 127) / 128), 128, 0, (cudaStream_t)stream, batch, x, y, s)));
  return C2B_OK;
}

int c2b_term_product_rev(int64_t batch, int32_t Jr1, int32_t Jc1, const double *ar1, const double *cr1, const double *ac1,
                         const double *bc1, const double *cc1, const double *dc1, int32_t Jr2, int32_t Jc2, const double *ar2,
                         const double *cr2, const double *ac2, const double *bc2, const double *cc2, const double *dc2,
                         const double *b_ar, const double *b_cr, const double *b_ac, const double *b_bc, const double *b_cc,
                         const double *b_dc, double *g1_ar, double *g1_cr, double *g1_ac, double *g1_bc, double *g1_cc, double *g1_dc,
                         double *g2_ar, double *g2_cr, double *g2_ac, double *g2_bc, double *g2_cc, double *g2_dc, void *stream) {
  c2b::OpScope _op("c2b_term_product_rev");
  using namespace c2b;
  C2B_REQUIRE(batch >= 1 && Jr1 >= 0 && Jc1 >= 0 && Jr2 >= 0 && Jc2 >= 0, "bad arguments");
  const int Jr = Jr1 * Jr2, Jc = Jr1 * Jc2 + Jr2 * Jc1 + 2 * Jc1 * Jc2;
  TermIn x{Jr1, Jc1, ar1, cr1, ac1, bc1, cc1, dc1}, y{Jr2, Jc2, ar2, cr2, ac2, bc2, cc2, dc2};
  TermSlots s{Jr, Jc, 0, 0, (double *)b_ar, (double *)b_cr, (double *)b_ac, (double *)b_bc, (double *)b_cc, (double *)b_dc};
  TermGrad gx{g1_ar, g1_cr, g1_ac, g1_bc, g1_cc, g1_dc}, gy{g2_ar, g2_cr, g2_ac, g2_bc, g2_cc, g2_dc};
  C2B_LAUNCH("term_product_rev_kernel", (cudaStream_t)stream,
             (launch_k(term_product_rev_kernel, (unsigned)((batch + 127) / 128), 128, 0, (cudaStream_t)stream, batch, x, y, s, gx, gy)));
  return C2B_OK;
}

}  // extern "C"
