// dk_mvn.cuh — deterministic multivariate normal CDFs for dimension <= 4 on the device.
// Stand-in for mnormt::pmnorm at the call sites /root/reference/R/qEI.R:160,166,179 (Genz bvn / tvn /
// sadmvn with abseps 1e-6 in the reference).  Same algorithms and the same Gauss-Legendre tables as
// oracle/mvnorm.py:
//   Phi_2 : Genz (2004) BVND with the 20-point rule
//   Phi_3 : Plackett's formula along r_1j(t) = sin(t asin r_1j) (Genz TVPACK), inner factor Phi_1
//   Phi_4 : the same reduction with the conditional bivariate CDF as inner factor
// Path integrals: 2 panels x 16-point Gauss-Legendre on t in [0, 1].
#pragma once
#include "dk_gl_tables.cuh"

namespace dk {

#define DK_TWOPI 6.283185307179586476925286766559
#define DK_SQRT_TWOPI 2.506628274631000502415765284811
#define DK_HALFPI 1.570796326794896619231321691640

// 1 / x for a positive normal x: hardware seed (about 20 bits) + two Newton steps, a few ulp
__device__ __forceinline__ double rcp_nr(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = fma(fma(-x, r, 1.0), r, r);
  r = fma(fma(-x, r, 1.0), r, r);
  return r;
}

// Polynomial coefficients live in the constant bank so that the FMAs read them as operands (literal doubles
// would be rebuilt by two moves each inside the quadrature loops).
//   [0..10]  sin:  -1/3!, 1/5!, ... , -1/23!   (odd Taylor series, |x| <= pi/2, truncation below 2e-18)
//   [11..22] exp:  1/2!, 1/3!, ... , 1/13!     (Taylor series on |r| <= ln2 / 2, truncation below 4e-18)
//   [23..25] log2(e), ln2 high part (32 trailing zero bits), ln2 low part
static __device__ __constant__ double MVN_K[26] = {
    -1.6666666666666666e-01, 8.333333333333333e-03,  -1.984126984126984e-04, 2.7557319223985893e-06,
    -2.505210838544172e-08,  1.6059043836821613e-10, -7.647163731819816e-13, 2.8114572543455206e-15,
    -8.22063524662433e-18,   1.9572941063391263e-20, -3.868170170630684e-23,
    5.0e-01,                 1.6666666666666666e-01, 4.1666666666666664e-02, 8.333333333333333e-03,
    1.388888888888889e-03,   1.984126984126984e-04,  2.48015873015873e-05,   2.7557319223985893e-06,
    2.755731922398589e-07,   2.505210838544172e-08,  2.08767569878681e-09,   1.6059043836821613e-10,
    1.4426950408889634,      6.93147180369123816490e-01, 1.90821492927058770002e-10};

// sin(x) for |x| <= pi/2 (every argument on the integration paths below); no range reduction, no slow path
__device__ __forceinline__ double sin_path(double x) {
  const double x2 = x * x;
  double p = MVN_K[10];
#pragma unroll
  for (int i = 9; i >= 0; i--) p = fma(p, x2, MVN_K[i]);
  return fma(x * x2, p, x);
}

// exp(x) for x < 709; returns 0 below -708 (the integrands here are <= 1 and such terms are far below the
// rounding of the sums they enter).  k = rint(x log2 e) by the 2^52 + 2^51 shift, r = x - k ln2, 2^k by exponent bits.
__device__ __forceinline__ double exp_path(double x) {
  const double shift = 6755399441055744.0;
  const double t = fma(x, MVN_K[23], shift);
  const int k = __double2loint(t);
  const double kf = t - shift;
  double r = fma(kf, -MVN_K[24], x);
  r = fma(kf, -MVN_K[25], r);
  double p = MVN_K[22];
#pragma unroll
  for (int i = 21; i >= 11; i--) p = fma(p, r, MVN_K[i]);
  p = fma(p, r, 1.0);
  p = fma(p, r, 1.0);
  const double v = __hiloint2double(__double2hiint(p) + k * 1048576, __double2loint(p));
  return (x < -708.0) ? 0.0 : v;
}

// One out-of-line copy of Phi for the CDF kernels: the inlined erfc is ~150 instructions and the quadrature code
// calls it from a dozen places; the hot path has to stay inside the 32 KB instruction cache.
static __device__ __noinline__ double phi_cdf(double x) { return normcdf(x); }

// |r| >= 0.925 branch of BVND (rare for the conditional correlations met here; kept out of line)
static __device__ __noinline__ double bvnu_far(double h, double k, double r) {
  double hk = h * k, bvn = 0.0;
  if (r < 0.0) { k = -k; hk = -hk; }
  if (fabs(r) < 1.0) {
    const double as = (1.0 - r) * (1.0 + r);
    double a = sqrt(as);
    const double bs = (h - k) * (h - k), c = (4.0 - hk) / 8.0, d = (12.0 - hk) / 16.0;
    double asr = -(bs / as + hk) / 2.0;
    if (asr > -100.0) bvn = a * exp(asr) * (1.0 - c * (bs - as) * (1.0 - d * bs / 5.0) / 3.0 + c * d * as * as / 5.0);
    if (-hk < 100.0) {
      const double b = sqrt(bs);
      bvn -= exp(-hk / 2.0) * DK_SQRT_TWOPI * phi_cdf(-b / a) * b * (1.0 - c * bs * (1.0 - d * bs / 5.0) / 3.0);
    }
    a = a / 2.0;
#pragma unroll 1
    for (int i = 0; i < 10; i++) {
#pragma unroll
      for (int s = 0; s < 2; s++) {
        const double t = a * ((s ? 1.0 : -1.0) * GL20_X[i] + 1.0);
        const double xs = t * t, rs = sqrt(1.0 - xs);
        asr = -(bs / xs + hk) / 2.0;
        if (asr > -100.0) {
          const double sp = 1.0 + c * xs * (1.0 + d * xs);
          const double ep = exp_path(-hk * xs / (2.0 * (1.0 + rs) * (1.0 + rs))) / rs;
          bvn += a * GL20_W[i] * exp_path(asr) * (ep - sp);
        }
      }
    }
    bvn = -bvn / DK_TWOPI;
  }
  if (r > 0.0) {
    bvn += phi_cdf(-fmax(h, k));
  } else {
    bvn = -bvn;
    if (k > h) bvn += (h < 0.0) ? (phi_cdf(k) - phi_cdf(h)) : (phi_cdf(-h) - phi_cdf(-k));
  }
  return bvn;
}

// P(X > dh, Y > dk), correlation r
static __device__ __noinline__ double bvnu_dev(double dh, double dk, double r) {
  if (dh == INFINITY || dk == INFINITY) return 0.0;
  if (dh == -INFINITY) return phi_cdf(-dk);
  if (dk == -INFINITY) return phi_cdf(-dh);
  double h = dh, k = dk, hk = h * k, bvn = 0.0;
  if (fabs(r) < 0.925) {
    if (fabs(r) > 0.0) {
      const double hs = (h * h + k * k) / 2.0, asr = asin(r);
#pragma unroll 1
      for (int i = 0; i < 10; i++) {
#pragma unroll
        for (int s = 0; s < 2; s++) {
          const double sn = sin_path(asr * ((s ? 1.0 : -1.0) * GL20_X[i] + 1.0) / 2.0);
          bvn += GL20_W[i] * exp_path((sn * hk - hs) * rcp_nr(1.0 - sn * sn));  // |sn| < 0.925
        }
      }
      bvn = bvn * asr / (2.0 * DK_TWOPI);
    }
    bvn += phi_cdf(-h) * phi_cdf(-k);
  } else {
    bvn = bvnu_far(h, k, r);
  }
  return fmin(1.0, fmax(0.0, bvn));
}

__device__ __forceinline__ double bvn_cdf_dev(double h, double k, double r) { return bvnu_dev(-h, -k, r); }

// sin(x), cos(x)^2 with Genz's series near |x| = pi/2 (TVPACK SINCS)
__device__ __forceinline__ void sincs_dev(double x, double& sx, double& cs) {
  const double e = DK_HALFPI - fabs(x), ee = e * e;
  if (ee < 5e-5) {
    sx = copysign(1.0 - ee * (1.0 - ee / 12.0) / 2.0, x);
    cs = ee * (1.0 - ee * (1.0 - 2.0 * ee / 15.0) / 3.0);
  } else {
    sx = sin_path(x);
    cs = 1.0 - sx * sx;
  }
}

__device__ __forceinline__ double tvn_term(double h1, double hj, double hk, double rj, double csj, double rk, double rjk) {
  if (csj <= 0.0) return 0.0;
  const double icsj = (csj < 1e-290) ? 1.0 / csj : rcp_nr(csj);
  const double e = exp_path(-0.5 * (h1 * h1 - 2.0 * rj * h1 * hj + hj * hj) * icsj);
  const double mu = ((rk - rj * rjk) * h1 + (rjk - rj * rk) * hj) * icsj;
  const double var = 1.0 - (rk * rk - 2.0 * rj * rk * rjk + rjk * rjk) * icsj;
  if (var <= 0.0) return e * (hk >= mu ? 1.0 : 0.0);
  return e * phi_cdf((hk - mu) * rsqrt(var));
}

static __device__ __noinline__ double tvn_cdf_dev(double h1, double h2, double h3, double r12, double r13, double r23) {
  const double base = phi_cdf(h1) * bvn_cdf_dev(h2, h3, r23);
  const double a2 = asin(r12), a3 = asin(r13);
  if (fabs(a2) + fabs(a3) == 0.0) return base;
  double tot = 0.0;
#pragma unroll 1
  for (int pnl = 0; pnl < 2; pnl++) {
    const double lo = 0.5 * pnl;
#pragma unroll 1
    for (int i = 0; i < 16; i++) {
      const double t = lo + (GL16_X[i] + 1.0) / 4.0;
      double s2, c2, s3, c3;
      sincs_dev(a2 * t, s2, c2);
      sincs_dev(a3 * t, s3, c3);
      double v = 0.0;
      if (a2 != 0.0) v += a2 * tvn_term(h1, h2, h3, s2, c2, s3, r23);
      if (a3 != 0.0) v += a3 * tvn_term(h1, h3, h2, s3, c3, s2, r23);
      tot += GL16_W[i] / 4.0 * v;
    }
  }
  return fmin(1.0, fmax(0.0, base + tot / DK_TWOPI));
}

// One term of the Phi_4 integrand: phi_2(h0, hj; rj) (up to 1 / (2 pi sqrt(csj))) times the bivariate CDF of the
// two remaining variables (limits hk, hl) given the first two.  rk1, rl1: their path correlations with variable
// 0; rkj, rlj, rkl: fixed correlations.  Out of line: one copy serves the three correlations of row 0.
static __device__ __noinline__ double qvn_pair_term(double h0, double hj, double hk, double hl, double rj, double csj,
                                             double rk1, double rl1, double rkj, double rlj, double rkl) {
  const double icsj = (csj < 1e-290) ? 1.0 / csj : rcp_nr(csj);
  const double e = exp_path(-0.5 * (h0 * h0 - 2.0 * rj * h0 * hj + hj * hj) * icsj);
  const double mk = ((rk1 - rj * rkj) * h0 + (rkj - rj * rk1) * hj) * icsj;
  const double ml = ((rl1 - rj * rlj) * h0 + (rlj - rj * rl1) * hj) * icsj;
  const double vk = 1.0 - (rk1 * rk1 - 2.0 * rj * rk1 * rkj + rkj * rkj) * icsj;
  const double vl = 1.0 - (rl1 * rl1 - 2.0 * rj * rl1 * rlj + rlj * rlj) * icsj;
  const double ckl = rkl - (rk1 * rl1 - rj * (rk1 * rlj + rl1 * rkj) + rkj * rlj) * icsj;
  double inner;
  if (vk <= 0.0 || vl <= 0.0) {
    const double pk = (vk <= 0.0) ? (hk >= mk ? 1.0 : 0.0) : phi_cdf((hk - mk) / sqrt(vk));
    const double pl = (vl <= 0.0) ? (hl >= ml ? 1.0 : 0.0) : phi_cdf((hl - ml) / sqrt(vl));
    inner = (vk <= 0.0 && vl <= 0.0) ? fmin(pk, pl) : pk * pl;
  } else {
    const double isk = rsqrt(vk), isl = rsqrt(vl);
    const double rho = fmax(-1.0, fmin(1.0, ckl * isk * isl));
    inner = bvn_cdf_dev((hk - mk) * isk, (hl - ml) * isl, rho);
  }
  return e * inner;
}

// Integrand of the Phi_4 path integral at parameter t: sum over the three correlations of row 0 of
// a_j * phi_2(h_0, h_j; r_j(t)) * Phi_2(the other two | those two), r_j(t) = sin(a_j t).
__device__ __forceinline__ double qvn_node(const double* h, const double* R, const double* a, double t) {
  double r1[3], cs[3];
  for (int j = 0; j < 3; j++) sincs_dev(a[j] * t, r1[j], cs[j]);
  double v = 0.0;
#pragma unroll
  for (int jj = 0; jj < 3; jj++) {
    if (a[jj] == 0.0 || cs[jj] <= 0.0) continue;
    const int j = jj + 1;
    const int k = (j == 1) ? 2 : 1, l = (j == 3) ? 2 : 3;
    v += a[jj] * qvn_pair_term(h[0], h[j], h[k], h[l], r1[jj], cs[jj], r1[k - 1], r1[l - 1], R[k * 4 + j], R[l * 4 + j],
                               R[k * 4 + l]);
  }
  return v;
}

// h[4], R row-major 4 x 4 correlation matrix
static __device__ __noinline__ double qvn_cdf_dev(const double* h, const double* R) {
  const double base = phi_cdf(h[0]) * tvn_cdf_dev(h[1], h[2], h[3], R[1 * 4 + 2], R[1 * 4 + 3], R[2 * 4 + 3]);
  double a[3];
  for (int j = 0; j < 3; j++) a[j] = asin(R[j + 1]);
  if (a[0] == 0.0 && a[1] == 0.0 && a[2] == 0.0) return base;
  double tot = 0.0;
#pragma unroll 1
  for (int i = 0; i < 32; i++) {
    const double t = 0.5 * (i >> 4) + (GL16_X[i & 15] + 1.0) / 4.0;
    tot += GL16_W[i & 15] / 4.0 * qvn_node(h, R, a, t);
  }
  return fmin(1.0, fmax(0.0, base + tot / DK_TWOPI));
}

// ---- cooperative versions: 8 consecutive lanes (an aligned group) share one CDF; the 32 nodes of a path
// integral are dealt round-robin to the lanes and summed by a fixed xor tree.  Every lane of the warp must call
// these (inactive groups pass active = false and only take part in the shuffles). ----
__device__ __forceinline__ double group8_sum(double v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  return v;
}

__device__ __forceinline__ double tvn_path_part(double h1, double h2, double h3, double r12, double r13, double r23, int s) {
  const double a2 = asin(r12), a3 = asin(r13);
  if (fabs(a2) + fabs(a3) == 0.0) return 0.0;
  double tot = 0.0;
#pragma unroll 1
  for (int i = s; i < 32; i += 8) {
    const double t = 0.5 * (i >> 4) + (GL16_X[i & 15] + 1.0) / 4.0;
    double s2, c2, s3, c3;
    sincs_dev(a2 * t, s2, c2);
    sincs_dev(a3 * t, s3, c3);
    double v = 0.0;
    if (a2 != 0.0) v += a2 * tvn_term(h1, h2, h3, s2, c2, s3, r23);
    if (a3 != 0.0) v += a3 * tvn_term(h1, h3, h2, s3, c3, s2, r23);
    tot += GL16_W[i & 15] / 4.0 * v;
  }
  return tot / DK_TWOPI;
}

__device__ __forceinline__ double tvn_coop(bool active, double h1, double h2, double h3, double r12, double r13, double r23,
                                           int s) {
  double part = active ? tvn_path_part(h1, h2, h3, r12, r13, r23, s) : 0.0;
  if (active && s == 0) part += phi_cdf(h1) * bvn_cdf_dev(h2, h3, r23);
  return fmin(1.0, fmax(0.0, group8_sum(part)));
}

__device__ __forceinline__ double qvn_path_part(const double* h, const double* R, int s) {
  double a[3];
  for (int j = 0; j < 3; j++) a[j] = asin(R[j + 1]);
  if (a[0] == 0.0 && a[1] == 0.0 && a[2] == 0.0) return 0.0;
  double tot = 0.0;
#pragma unroll 1
  for (int i = s; i < 32; i += 8) {
    const double t = 0.5 * (i >> 4) + (GL16_X[i & 15] + 1.0) / 4.0;
    tot += GL16_W[i & 15] / 4.0 * qvn_node(h, R, a, t);
  }
  return tot / DK_TWOPI;
}

// pmnorm(x, varcov = S), q = 1..4 uniform over the warp, shared by the 8 lanes of a group
__device__ __forceinline__ double pmnorm_coop(bool active, int q, const double* x, const double* S, int s) {
  double z[4] = {0.0, 0.0, 0.0, 0.0}, R[16];
  for (int i = 0; i < 16; i++) R[i] = 0.0;
  if (active) {
    for (int i = 0; i < q; i++) z[i] = x[i] / sqrt(S[i * 4 + i]);
    for (int i = 0; i < q; i++)
      for (int j = 0; j < q; j++) {
        const double r = S[i * 4 + j] / (sqrt(S[i * 4 + i]) * sqrt(S[j * 4 + j]));
        R[i * 4 + j] = (i == j) ? 1.0 : fmax(-1.0, fmin(1.0, r));
      }
  }
  if (q == 1) return phi_cdf(z[0]);
  if (q == 2) return group8_sum((active && s == 0) ? bvn_cdf_dev(z[0], z[1], R[1]) : 0.0);
  if (q == 3) return tvn_coop(active, z[0], z[1], z[2], R[1], R[2], R[1 * 4 + 2], s);
  const double tv = tvn_coop(active, z[1], z[2], z[3], R[1 * 4 + 2], R[1 * 4 + 3], R[2 * 4 + 3], s);
  double part = active ? qvn_path_part(z, R, s) : 0.0;
  if (active && s == 0) part += phi_cdf(z[0]) * tv;
  return fmin(1.0, fmax(0.0, group8_sum(part)));
}

// pmnorm(x, varcov = S) for q = 1..4; S row-major with leading dimension 4.  Out of line and rolled: the callers
// reach it from a dozen sites and the square roots / divisions expand to ~100 instructions each.
static __device__ __noinline__ double pmnorm_dev(int q, const double* x, const double* S) {
  double sd[4], z[4], R[16];
#pragma unroll 1
  for (int i = 0; i < q; i++) {
    sd[i] = sqrt(S[i * 4 + i]);
    z[i] = x[i] / sd[i];
  }
  if (q == 1) return phi_cdf(z[0]);
#pragma unroll 1
  for (int i = 0; i < q; i++)
#pragma unroll 1
    for (int j = 0; j < q; j++) {
      const double r = S[i * 4 + j] / (sd[i] * sd[j]);
      R[i * 4 + j] = (i == j) ? 1.0 : fmax(-1.0, fmin(1.0, r));
    }
  if (q == 2) return bvn_cdf_dev(z[0], z[1], R[1]);
  if (q == 3) return tvn_cdf_dev(z[0], z[1], z[2], R[1], R[2], R[1 * 4 + 2]);
  return qvn_cdf_dev(z, R);
}

}  // namespace dk
