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

// P(X > dh, Y > dk), correlation r
__device__ __noinline__ double bvnu_dev(double dh, double dk, double r) {
  if (dh == INFINITY || dk == INFINITY) return 0.0;
  if (dh == -INFINITY) return normcdf(-dk);
  if (dk == -INFINITY) return normcdf(-dh);
  double h = dh, k = dk, hk = h * k, bvn = 0.0;
  if (fabs(r) < 0.925) {
    if (fabs(r) > 0.0) {
      const double hs = (h * h + k * k) / 2.0, asr = asin(r);
      for (int i = 0; i < 10; i++) {
#pragma unroll
        for (int s = 0; s < 2; s++) {
          const double sn = sin(asr * ((s ? 1.0 : -1.0) * GL20_X[i] + 1.0) / 2.0);
          bvn += GL20_W[i] * exp((sn * hk - hs) / (1.0 - sn * sn));
        }
      }
      bvn = bvn * asr / (2.0 * DK_TWOPI);
    }
    bvn += normcdf(-h) * normcdf(-k);
  } else {
    if (r < 0.0) { k = -k; hk = -hk; }
    if (fabs(r) < 1.0) {
      const double as = (1.0 - r) * (1.0 + r);
      double a = sqrt(as);
      const double bs = (h - k) * (h - k), c = (4.0 - hk) / 8.0, d = (12.0 - hk) / 16.0;
      double asr = -(bs / as + hk) / 2.0;
      if (asr > -100.0) bvn = a * exp(asr) * (1.0 - c * (bs - as) * (1.0 - d * bs / 5.0) / 3.0 + c * d * as * as / 5.0);
      if (-hk < 100.0) {
        const double b = sqrt(bs);
        bvn -= exp(-hk / 2.0) * DK_SQRT_TWOPI * normcdf(-b / a) * b * (1.0 - c * bs * (1.0 - d * bs / 5.0) / 3.0);
      }
      a = a / 2.0;
      for (int i = 0; i < 10; i++) {
#pragma unroll
        for (int s = 0; s < 2; s++) {
          const double t = a * ((s ? 1.0 : -1.0) * GL20_X[i] + 1.0);
          const double xs = t * t, rs = sqrt(1.0 - xs);
          asr = -(bs / xs + hk) / 2.0;
          if (asr > -100.0) {
            const double sp = 1.0 + c * xs * (1.0 + d * xs);
            const double ep = exp(-hk * xs / (2.0 * (1.0 + rs) * (1.0 + rs))) / rs;
            bvn += a * GL20_W[i] * exp(asr) * (ep - sp);
          }
        }
      }
      bvn = -bvn / DK_TWOPI;
    }
    if (r > 0.0) {
      bvn += normcdf(-fmax(h, k));
    } else {
      bvn = -bvn;
      if (k > h) bvn += (h < 0.0) ? (normcdf(k) - normcdf(h)) : (normcdf(-h) - normcdf(-k));
    }
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
    sx = sin(x);
    cs = 1.0 - sx * sx;
  }
}

__device__ __forceinline__ double tvn_term(double h1, double hj, double hk, double rj, double csj, double rk, double rjk) {
  if (csj <= 0.0) return 0.0;
  const double e = exp(-(h1 * h1 - 2.0 * rj * h1 * hj + hj * hj) / (2.0 * csj));
  const double mu = ((rk - rj * rjk) * h1 + (rjk - rj * rk) * hj) / csj;
  const double var = 1.0 - (rk * rk - 2.0 * rj * rk * rjk + rjk * rjk) / csj;
  if (var <= 0.0) return e * (hk >= mu ? 1.0 : 0.0);
  return e * normcdf((hk - mu) / sqrt(var));
}

__device__ __noinline__ double tvn_cdf_dev(double h1, double h2, double h3, double r12, double r13, double r23) {
  const double base = normcdf(h1) * bvn_cdf_dev(h2, h3, r23);
  const double a2 = asin(r12), a3 = asin(r13);
  if (fabs(a2) + fabs(a3) == 0.0) return base;
  double tot = 0.0;
  for (int pnl = 0; pnl < 2; pnl++) {
    const double lo = 0.5 * pnl;
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

// h[4], R row-major 4 x 4 correlation matrix
__device__ __noinline__ double qvn_cdf_dev(const double* h, const double* R) {
  const double base = normcdf(h[0]) * tvn_cdf_dev(h[1], h[2], h[3], R[1 * 4 + 2], R[1 * 4 + 3], R[2 * 4 + 3]);
  double a[3];
  for (int j = 0; j < 3; j++) a[j] = asin(R[j + 1]);
  if (a[0] == 0.0 && a[1] == 0.0 && a[2] == 0.0) return base;
  double tot = 0.0;
  for (int pnl = 0; pnl < 2; pnl++) {
    const double lo = 0.5 * pnl;
    for (int i = 0; i < 16; i++) {
      const double t = lo + (GL16_X[i] + 1.0) / 4.0;
      double r1[3], cs[3];
      for (int j = 0; j < 3; j++) sincs_dev(a[j] * t, r1[j], cs[j]);
      double v = 0.0;
      for (int jj = 0; jj < 3; jj++) {
        if (a[jj] == 0.0 || cs[jj] <= 0.0) continue;
        const int j = jj + 1;
        const int k = (j == 1) ? 2 : 1, l = (j == 3) ? 2 : 3;
        const double rj = r1[jj], csj = cs[jj], hj = h[j];
        const double e = exp(-(h[0] * h[0] - 2.0 * rj * h[0] * hj + hj * hj) / (2.0 * csj));
        const double rk1 = r1[k - 1], rl1 = r1[l - 1];
        const double rkj = R[k * 4 + j], rlj = R[l * 4 + j], rkl = R[k * 4 + l];
        const double mk = ((rk1 - rj * rkj) * h[0] + (rkj - rj * rk1) * hj) / csj;
        const double ml = ((rl1 - rj * rlj) * h[0] + (rlj - rj * rl1) * hj) / csj;
        const double vk = 1.0 - (rk1 * rk1 - 2.0 * rj * rk1 * rkj + rkj * rkj) / csj;
        const double vl = 1.0 - (rl1 * rl1 - 2.0 * rj * rl1 * rlj + rlj * rlj) / csj;
        const double ckl = rkl - (rk1 * rl1 - rj * (rk1 * rlj + rl1 * rkj) + rkj * rlj) / csj;
        double inner;
        if (vk <= 0.0 || vl <= 0.0) {
          const double pk = (vk <= 0.0) ? (h[k] >= mk ? 1.0 : 0.0) : normcdf((h[k] - mk) / sqrt(vk));
          const double pl = (vl <= 0.0) ? (h[l] >= ml ? 1.0 : 0.0) : normcdf((h[l] - ml) / sqrt(vl));
          inner = (vk <= 0.0 && vl <= 0.0) ? fmin(pk, pl) : pk * pl;
        } else {
          const double rho = fmax(-1.0, fmin(1.0, ckl / sqrt(vk * vl)));
          inner = bvn_cdf_dev((h[k] - mk) / sqrt(vk), (h[l] - ml) / sqrt(vl), rho);
        }
        v += a[jj] * e * inner;
      }
      tot += GL16_W[i] / 4.0 * v;
    }
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
  if (active && s == 0) part += normcdf(h1) * bvn_cdf_dev(h2, h3, r23);
  return fmin(1.0, fmax(0.0, group8_sum(part)));
}

__device__ __forceinline__ double qvn_path_part(const double* h, const double* R, int s) {
  double a[3];
  for (int j = 0; j < 3; j++) a[j] = asin(R[j + 1]);
  if (a[0] == 0.0 && a[1] == 0.0 && a[2] == 0.0) return 0.0;
  double tot = 0.0;
  for (int i = s; i < 32; i += 8) {
    const double t = 0.5 * (i >> 4) + (GL16_X[i & 15] + 1.0) / 4.0;
    double r1[3], cs[3];
    for (int j = 0; j < 3; j++) sincs_dev(a[j] * t, r1[j], cs[j]);
    double v = 0.0;
    for (int jj = 0; jj < 3; jj++) {
      if (a[jj] == 0.0 || cs[jj] <= 0.0) continue;
      const int j = jj + 1;
      const int k = (j == 1) ? 2 : 1, l = (j == 3) ? 2 : 3;
      const double rj = r1[jj], csj = cs[jj], hj = h[j];
      const double e = exp(-(h[0] * h[0] - 2.0 * rj * h[0] * hj + hj * hj) / (2.0 * csj));
      const double rk1 = r1[k - 1], rl1 = r1[l - 1];
      const double rkj = R[k * 4 + j], rlj = R[l * 4 + j], rkl = R[k * 4 + l];
      const double mk = ((rk1 - rj * rkj) * h[0] + (rkj - rj * rk1) * hj) / csj;
      const double ml = ((rl1 - rj * rlj) * h[0] + (rlj - rj * rl1) * hj) / csj;
      const double vk = 1.0 - (rk1 * rk1 - 2.0 * rj * rk1 * rkj + rkj * rkj) / csj;
      const double vl = 1.0 - (rl1 * rl1 - 2.0 * rj * rl1 * rlj + rlj * rlj) / csj;
      const double ckl = rkl - (rk1 * rl1 - rj * (rk1 * rlj + rl1 * rkj) + rkj * rlj) / csj;
      double inner;
      if (vk <= 0.0 || vl <= 0.0) {
        const double pk = (vk <= 0.0) ? (h[k] >= mk ? 1.0 : 0.0) : normcdf((h[k] - mk) / sqrt(vk));
        const double pl = (vl <= 0.0) ? (h[l] >= ml ? 1.0 : 0.0) : normcdf((h[l] - ml) / sqrt(vl));
        inner = (vk <= 0.0 && vl <= 0.0) ? fmin(pk, pl) : pk * pl;
      } else {
        const double rho = fmax(-1.0, fmin(1.0, ckl / sqrt(vk * vl)));
        inner = bvn_cdf_dev((h[k] - mk) / sqrt(vk), (h[l] - ml) / sqrt(vl), rho);
      }
      v += a[jj] * e * inner;
    }
    tot += GL16_W[i & 15] / 4.0 * v;
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
  if (q == 1) return normcdf(z[0]);
  if (q == 2) return group8_sum((active && s == 0) ? bvn_cdf_dev(z[0], z[1], R[1]) : 0.0);
  if (q == 3) return tvn_coop(active, z[0], z[1], z[2], R[1], R[2], R[1 * 4 + 2], s);
  const double tv = tvn_coop(active, z[1], z[2], z[3], R[1 * 4 + 2], R[1 * 4 + 3], R[2 * 4 + 3], s);
  double part = active ? qvn_path_part(z, R, s) : 0.0;
  if (active && s == 0) part += normcdf(z[0]) * tv;
  return fmin(1.0, fmax(0.0, group8_sum(part)));
}

// pmnorm(x, varcov = S) for q = 1..4; S row-major with leading dimension 4.
__device__ __forceinline__ double pmnorm_dev(int q, const double* x, const double* S) {
  double z[4], R[16];
  for (int i = 0; i < q; i++) z[i] = x[i] / sqrt(S[i * 4 + i]);
  if (q == 1) return normcdf(z[0]);
  for (int i = 0; i < q; i++)
    for (int j = 0; j < q; j++) {
      const double r = S[i * 4 + j] / (sqrt(S[i * 4 + i]) * sqrt(S[j * 4 + j]));
      R[i * 4 + j] = (i == j) ? 1.0 : fmax(-1.0, fmin(1.0, r));
    }
  if (q == 2) return bvn_cdf_dev(z[0], z[1], R[1]);
  if (q == 3) return tvn_cdf_dev(z[0], z[1], z[2], R[1], R[2], R[1 * 4 + 2]);
  return qvn_cdf_dev(z, R);
}

}  // namespace dk
