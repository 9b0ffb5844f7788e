// jb_device.cuh — device-side building blocks of the B200 IMC/DDMC hot path:
// per-particle Philox streams, the EOS / opacity / scattering closures, and the
// per-segment physics (IMC step, DDMC albedo + step, isotropic scatter, Planck sampler,
// block-face resampling). Everything here is FP64 and is written as explicit IEEE
// operations (the library is compiled with -fmad=false; fused operations only appear as
// explicit fma() inside jb_math.h) so that a particle history is bit-identical to the
// CPU oracle's on the same stream.
//
// Reference behaviour followed (lanl/jaybenne @ 04947dfa):
//   src/jaybenne/transport_utils.hpp:23-39,111-160,163-277,279-397
//   src/jaybenne/scattering.hpp:21-29, src/jaybenne/planck.hpp:27-50
//   src/jaybenne/sample_ddmc_bface.cpp:24-78
//   singularity-opac photons/{gray,epbremsstrahlung,non_cgs}_photons.hpp,
//   photons/{gray_s,thomson_s,non_cgs_s}_photons.hpp, singularity-eos eos_ideal.hpp:55-58,100-103
#ifndef JAYBENNE_B200_JB_DEVICE_CUH_
#define JAYBENNE_B200_JB_DEVICE_CUH_

#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>

#include "../../include/jaybenne_b200.h"
#include "jb_math.h"

namespace jbd {

// stream selectors folded into the Philox key (particle histories / per-cell rounding)
static constexpr uint32_t STREAM_PARTICLE = 0u;
static constexpr uint32_t STREAM_CELL = 0x5EED0001u;

static constexpr double JB_PI = 3.14159265358979323846;
static constexpr double EPS10 = 10 * DBL_EPSILON;            // parthenon::robust::EPS
static constexpr double EPS_IMC = 1.0e6 * EPS10;             // transport_utils.hpp:23
static constexpr double EPS_DDMC = 1.0e8 * EPS10;            // transport_utils.hpp:24
static constexpr double LAM_EXT = 0.7104;                    // extrapolation distance

// ---------------------------------------------------------------------------------------
// Philox4x32-10 counter generator. One particle owns the stream
//   counter = (draw >> 1, birth index, birth gid, source call), key = (seed_lo, seed_hi ^ stream)
// and double #n is built from words (2*(n&1), 2*(n&1)+1) of block n>>1. The generated
// block is cached so two consecutive draws cost one Philox evaluation.
struct Rng {
  uint32_t id0, id1, id2, n, k0, k1;
  uint32_t w[4];
  uint32_t have;  // block index held in w (0xffffffff = none)

  __device__ __forceinline__ void init(uint32_t a, uint32_t b, uint32_t c, uint32_t draws, uint32_t key0,
                                       uint32_t key1) {
    id0 = a; id1 = b; id2 = c; n = draws; k0 = key0; k1 = key1;
    have = 0xffffffffu;
  }
  __device__ __forceinline__ void refill(uint32_t blk) {
    uint32_t c0 = blk, c1 = id0, c2 = id1, c3 = id2, a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      c0 = hi1 ^ c1 ^ a;
      c1 = lo1;
      c2 = hi0 ^ c3 ^ b;
      c3 = lo0;
      a += 0x9E3779B9u;
      b += 0xBB67AE85u;
    }
    w[0] = c0; w[1] = c1; w[2] = c2; w[3] = c3;
    have = blk;
  }
  __device__ __forceinline__ double drand() {
    const uint32_t blk = n >> 1;
    if (blk != have) refill(blk);
    const bool odd = (n & 1u) != 0;
    const uint32_t lo = odd ? w[2] : w[0], hi = odd ? w[3] : w[1];
    const uint64_t u = ((uint64_t)hi << 32) | lo;
    ++n;
    return (double)(u >> 11) * 1.1102230246251565e-16;
  }
};

// The ten round keys of one Philox key pair, computed once on the host (k[2r] = key0 + r * 0x9E3779B9,
// k[2r+1] = key1 + r * 0xBB67AE85). A kernel that holds them in its parameter block reads each one as a
// constant-bank operand of the round's XOR; bumping the keys in the loop cost two uniform adds per round, a third
// of the generator's instructions (profiles/r02_v11_*.md).
struct PKeys {
  uint32_t k[20];
};
__host__ __device__ inline PKeys make_pkeys(uint32_t k0, uint32_t k1) {
  PKeys K;
  for (int r = 0; r < 10; ++r) {
    K.k[2 * r] = k0 + (uint32_t)r * 0x9E3779B9u;
    K.k[2 * r + 1] = k1 + (uint32_t)r * 0xBB67AE85u;
  }
  return K;
}
__device__ __forceinline__ void philox_pair(uint32_t blk, uint32_t id0, uint32_t id1, uint32_t id2, const PKeys &K,
                                            double &d0, double &d1) {
  uint32_t c0 = blk, c1 = id0, c2 = id1, c3 = id2;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ K.k[2 * r];
    c1 = lo1;
    c2 = hi0 ^ c3 ^ K.k[2 * r + 1];
    c3 = lo0;
  }
  d0 = (double)((((uint64_t)c1 << 32) | c0) >> 11) * 1.1102230246251565e-16;
  d1 = (double)((((uint64_t)c3 << 32) | c2) >> 11) * 1.1102230246251565e-16;
}

// doubles #2*blk and #2*blk+1 of a stream in one straight-line Philox evaluation (what two
// consecutive Rng::drand() calls starting at an even draw counter return)
__device__ __forceinline__ void philox_pair(uint32_t blk, uint32_t id0, uint32_t id1, uint32_t id2, uint32_t k0,
                                            uint32_t k1, double &d0, double &d1) {
  uint32_t c0 = blk, c1 = id0, c2 = id1, c3 = id2, a = k0, b = k1;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ a;
    c1 = lo1;
    c2 = hi0 ^ c3 ^ b;
    c3 = lo0;
    a += 0x9E3779B9u;
    b += 0xBB67AE85u;
  }
  d0 = (double)((((uint64_t)c1 << 32) | c0) >> 11) * 1.1102230246251565e-16;
  d1 = (double)((((uint64_t)c3 << 32) | c2) >> 11) * 1.1102230246251565e-16;
}

// ---------------------------------------------------------------------------------------
// The three closures jaybenne::Initialize is handed (jaybenne.cpp:227-233), flattened to
// plain numbers. The host fills this once (jb_lib.cu: make_closure) with the same libm
// calls the CPU side uses for the derived constants.
struct Closure {
  int opacity_model, scattering_model, gray_mode;
  double kappa, sigma, apm, cv;
  double time_unit, length_unit, temp_unit, rho_unit, freq_unit, inv_emiss_unit;
  double c_code, sb_code;
  double bt_num, bt_den;            // 8 pi^5 kb^4, 15 c^2 h^3
  double sigma_thomson;
  double ep_mmw, ep_prefac;
  double h, kb, c, me;              // CGS constants as singularity-opac derives them
  double two_h_over_c2;             // unused in parity arithmetic; kept for diagnostics
  double pi4;                       // pow(pi, 4) (planck.hpp:30)
};

// IdealGas::TemperatureFromDensityInternalEnergy, eos_ideal.hpp:55-58
__device__ __forceinline__ double eos_temperature(const Closure &cl, double sie) {
  const double q = sie / cl.cv;
  return 0.0 > q ? 0.0 : q;
}

// PlanckDistribution::ThermalDistributionOfTNu, thermal_distributions_photons.hpp:31-37
__device__ __forceinline__ double planck_bnu(const Closure &cl, double temp, double nu) {
  const double x = cl.h * nu / (cl.kb * temp);
  return (2. * cl.h * nu * nu * nu / (cl.c * cl.c)) * 1. / jbm::expm1(x);
}
// singularity robust::ratio, robust_utils.hpp:49-52
__device__ __forceinline__ double robust_ratio(double a, double b) {
  const int sgn = (0.0 <= b) - (b < 0.0);
  return a / (b + sgn * (10 * DBL_MIN));
}
// the same for a denominator known to be >= +0 or NaN (B_nu of a positive frequency): sgn is 1 (or the
// result NaN either way), so the two compares, the int -> double conversion and the multiply drop out
__device__ __forceinline__ double robust_ratio_pos(double a, double b) { return a / (b + 10 * DBL_MIN); }
// EPBremss::EmissivityPerNuOmega, epbremsstrahlung_opacity_photons.hpp:88-96
__device__ __forceinline__ double epbremss_jnu(const Closure &cl, double rho, double temp, double nu) {
  const double thetaE = cl.kb * temp / (cl.me * cl.c * cl.c);
  const double x = cl.h * nu / (cl.kb * temp);
  const double ne = rho / cl.ep_mmw / 2.;
  const double ni = ne;
  return cl.ep_prefac / sqrt(thetaE) * ne * ni * jbm::exp(-x) * 1.2;
}

// Opacity::AbsorptionCoefficient (NonCGSUnits<Gray|EPBremss>)
__device__ __forceinline__ double absorption_coeff(const Closure &cl, double rho, double temp, double nu) {
  const double r = cl.rho_unit * rho, T = cl.temp_unit * temp, f = nu * cl.freq_unit;
  double alpha;
  if (cl.opacity_model == JB_OPAC_GRAY) {
    if (cl.gray_mode == JB_GRAY_DIRECT) {
      alpha = r * cl.kappa;
    } else {
      const double Bnu = planck_bnu(cl, T, f);
      const double jnu = r * cl.kappa * Bnu;
      alpha = robust_ratio(jnu, Bnu);
    }
  } else {
    const double Bnu = planck_bnu(cl, T, f);
    const double jnu = epbremss_jnu(cl, r, T, f);
    alpha = robust_ratio(jnu, Bnu);
  }
  return alpha * cl.length_unit;
}

// The same closure with the factors that depend on the photon alone taken out of the segment
// loop: h*nu and 2 h nu^3 / c^2 are evaluated once per photon by exactly the operations
// planck_bnu() would perform (association included), so the value is bit-identical.
struct PhotonFreq {
  double f, hnu, bpref;  // nu in CGS, h*nu, (2*h*nu*nu*nu/(c*c)) * 1.
};
__device__ __forceinline__ PhotonFreq photon_freq(const Closure &cl, double nu) {
  PhotonFreq q;
  q.f = nu * cl.freq_unit;
  q.hnu = cl.h * q.f;
  q.bpref = (2. * cl.h * q.f * q.f * q.f / (cl.c * cl.c)) * 1.;
  return q;
}
__device__ __forceinline__ double absorption_coeff_hoisted(const Closure &cl, double rho, double temp,
                                                           const PhotonFreq &q) {
  const double r = cl.rho_unit * rho, T = cl.temp_unit * temp;
  double alpha;
  if (cl.opacity_model == JB_OPAC_GRAY && cl.gray_mode == JB_GRAY_DIRECT) {
    alpha = r * cl.kappa;
  } else {
    const double x = q.hnu / (cl.kb * T);
    const double Bnu = q.bpref / jbm::expm1(x);
    const double jnu = cl.opacity_model == JB_OPAC_GRAY ? r * cl.kappa * Bnu : epbremss_jnu(cl, r, T, q.f);
    alpha = robust_ratio(jnu, Bnu);
  }
  return alpha * cl.length_unit;
}

// Scattering::TotalScatteringCoefficient (NonCGSUnitsS<GrayS|ThomsonS>)
__device__ __forceinline__ double scattering_coeff(const Closure &cl, double rho) {
  const double r = cl.rho_unit * rho;
  double alpha;
  if (cl.scattering_model == JB_SCAT_GRAY) {
    alpha = (r / cl.apm) * cl.sigma;
  } else {
    const double ne = (r / cl.apm) / 2.;
    alpha = ne * cl.sigma_thomson;
  }
  return alpha * cl.length_unit;
}

// Opacity::Emissivity (frequency- and angle-integrated)
__device__ __forceinline__ double emissivity(const Closure &cl, double rho, double temp) {
  const double r = rho * cl.rho_unit, T = temp * cl.temp_unit;
  double J;
  if (cl.opacity_model == JB_OPAC_GRAY) {
    const double B = cl.bt_num * jbm::pow4(T) / cl.bt_den;
    J = r * cl.kappa * B;
  } else {
    const double thetaE = cl.kb * T / (cl.me * cl.c * cl.c);
    const double ne = r / cl.ep_mmw / 2.;
    const double ni = ne;
    J = 4. * JB_PI * cl.kb * T / cl.h * cl.ep_prefac / sqrt(thetaE) * ne * ni * 1.2;
  }
  return J * cl.inv_emiss_unit * cl.time_unit;
}

// ---------------------------------------------------------------------------------------
// One photon in registers while its history is advanced.
struct Photon {
  double x, y, z, t, vx, vy, vz;
  int ip, jp, kp;
};
// The cell the photon sits in: bounds + the quantities looked up per segment.
struct CellEnv {
  double xl, yl, zl, xu, yu, zu;
  double ff, aa, ss;
};

// cosine-law direction about axis 1 (transport_utils.hpp:26-39)
__device__ __forceinline__ void face_iso_dir(const double vv, Rng &rng, double &v1, double &v2, double &v3) {
  const double mu = sqrt(rng.drand());
  const double nu = sqrt(1.0 - mu * mu);
  const double phi = 2.0 * JB_PI * rng.drand();
  double sn, cs;
  jbm::sincos(phi, &sn, &cs);
  v1 = vv * mu;
  v2 = vv * nu * cs;
  v3 = vv * nu * sn;
}

// isotropic scatter (scattering.hpp:21-29)
__device__ __forceinline__ void scatter_iso(Rng &rng, const double vv, double &vx, double &vy, double &vz) {
  const double mu = 2.0 * rng.drand() - 1.0;
  const double phi = 2.0 * JB_PI * rng.drand();
  const double stheta = sqrt(1.0 - mu * mu);
  double sn, cs;
  jbm::sincos(phi, &sn, &cs);
  vx = vv * stheta * cs;
  vy = vv * stheta * sn;
  vz = vv * mu;
}

// Everett & Cashwell Planck sampler with the five draws handed in. d_planck_cum[l] = sum_{j<=l} j^-4
// accumulated in the reference's order with the same IEEE operations (host-filled once, see
// launch_source_emit), so `first l with cum[l] >= rhs` is the loop of planck.hpp:33-41 without
// its divides.
static __constant__ double d_planck_cum[100];
__device__ __forceinline__ double planck_energy_draws(const Closure &cl, const double xi0, const double xi1,
                                                      const double xi2, const double xi3, const double xi4,
                                                      const double sb, const double temp) {
  const double rhs = jbm::div_normal(xi0 * cl.pi4, 90.0);  // (numerator 0 or >= 1e-14: the IEEE quotient)
  int l = 1;
  double ll = 1.0;
  while (l < 100) {
    if (d_planck_cum[l] >= rhs) {
      ll = (double)l;
      break;
    }
    l++;
  }
  return -jbm::div_normal(1.0, ll) * jbm::log(xi1 * xi2 * xi3 * xi4) * sb * temp;  // ll = 1 .. 99
}

// Everett & Cashwell Planck sampler (planck.hpp:27-50); running sum of j^-4
__device__ __forceinline__ double planck_energy(const Closure &cl, Rng &rng, const double sb, const double temp) {
  const double xi0 = rng.drand();
  const double rhs = xi0 * cl.pi4 / 90.0;
  int l = 1;
  double ll = 1.0;
  double lhs = 0.0;
  while (l < 100) {
    const double j = (double)l;
    lhs += 1.0 / ((j * j) * (j * j));
    if (lhs >= rhs) {
      ll = j;
      break;
    }
    l++;
  }
  const double xi1 = rng.drand();
  const double xi2 = rng.drand();
  const double xi3 = rng.drand();
  const double xi4 = rng.drand();
  return -(1.0 / ll) * jbm::log(xi1 * xi2 * xi3 * xi4) * sb * temp;
}

// One IMC segment (ptcl_transport_step, transport_utils.hpp:111-160): distances to
// effective absorption / effective scatter / census / cell faces, capped at the smallest
// cell width; advance; nudge across a face the photon has reached.
// di/dj/dk report which face nudge fired (-1 lower, +1 upper, 0 none): after a nudge the
// photon sits eps_imc_offset cell widths beyond that face, every other end point is at least
// that far inside, so the cell index moves by exactly that amount (see k_transport_imc).
template <int NDIM>
__device__ __forceinline__ void imc_step_idx(Rng &rng, const double t_end, const CellEnv &c, const double vv,
                                             const double dx_cell, Photon &p, bool &is_absorbed, bool &is_scattered,
                                             int &di, int &dj, int &dk) {
  constexpr bool multi_d = NDIM >= 2, three_d = NDIM == 3;
  const double rmin = DBL_MIN;
  const double lam_abs = 1.0 / (c.ff * c.aa + rmin);
  const double lam_sc = 1.0 / (c.ss + (1.0 - c.ff) * c.aa + rmin);
  const double dx_abs = -lam_abs * jbm::log(rng.drand());
  const double dx_sc = -lam_sc * jbm::log(rng.drand());
  const double dx_end = vv * (t_end - p.t);
  double dx_push = jbm::min2(dx_cell, dx_end);
  if (p.vx > 0.0) {
    dx_push = jbm::min2(dx_push, vv * (c.xu - p.x) / p.vx);
  } else if (p.vx < 0.0) {
    dx_push = jbm::min2(dx_push, vv * (c.xl - p.x) / p.vx);
  }
  if (multi_d) {
    if (p.vy > 0.0) {
      dx_push = jbm::min2(dx_push, vv * (c.yu - p.y) / p.vy);
    } else if (p.vy < 0.0) {
      dx_push = jbm::min2(dx_push, vv * (c.yl - p.y) / p.vy);
    }
  }
  if (three_d) {
    if (p.vz > 0.0) {
      dx_push = jbm::min2(dx_push, vv * (c.zu - p.z) / p.vz);
    } else if (p.vz < 0.0) {
      dx_push = jbm::min2(dx_push, vv * (c.zl - p.z) / p.vz);
    }
  }
  is_absorbed = ((dx_abs < dx_push) && (dx_abs < dx_sc));
  is_scattered = (!is_absorbed && (dx_sc < dx_push));
  const double dt_push = (is_absorbed ? dx_abs : (is_scattered ? dx_sc : dx_push)) / vv;
  p.t += dt_push;
  p.x += p.vx * dt_push;
  p.y += (multi_d ? 1.0 : 0.0) * p.vy * dt_push;
  p.z += (three_d ? 1.0 : 0.0) * p.vz * dt_push;
  const double fdx = EPS_IMC * (c.xu - c.xl);
  const double fdy = EPS_IMC * (c.yu - c.yl);
  const double fdz = EPS_IMC * (c.zu - c.zl);
  const bool nxl = fabs(p.x - c.xl) < fdx;
  p.x = nxl ? c.xl - fdx : p.x;
  const bool nxu = fabs(p.x - c.xu) < fdx;
  p.x = nxu ? c.xu + fdx : p.x;
  const bool nyl = multi_d && fabs(p.y - c.yl) < fdy;
  p.y = nyl ? c.yl - fdy : p.y;
  const bool nyu = multi_d && fabs(p.y - c.yu) < fdy;
  p.y = nyu ? c.yu + fdy : p.y;
  const bool nzl = three_d && fabs(p.z - c.zl) < fdz;
  p.z = nzl ? c.zl - fdz : p.z;
  const bool nzu = three_d && fabs(p.z - c.zu) < fdz;
  p.z = nzu ? c.zu + fdz : p.z;
  di = nxu ? 1 : (nxl ? -1 : 0);
  dj = nyu ? 1 : (nyl ? -1 : 0);
  dk = nzu ? 1 : (nzl ? -1 : 0);
}
template <int NDIM>
__device__ __forceinline__ void imc_step(Rng &rng, const double t_end, const CellEnv &c, const double vv,
                                         const double dx_cell, Photon &p, bool &is_absorbed, bool &is_scattered) {
  int di, dj, dk;
  imc_step_idx<NDIM>(rng, t_end, c, vv, dx_cell, p, is_absorbed, is_scattered, di, dj, dk);
}

__device__ __forceinline__ bool fuzzy_eq(const double a, const double b, const double c, const double eps) {
  return fabs(a - b) < c * eps;  // jaybenne_utils.hpp:44-49
}

// (v1, v2, v3) of a direction sampled about `axis` (normal, then the two transverse components in
// cyclic order) back to (vx, vy, vz): the argument orders of the reference's six call sites
// (p.vx,p.vy,p.vz | p.vy,p.vz,p.vx | p.vz,p.vx,p.vy, transport_utils.hpp:206-262, 301-386).
__device__ __forceinline__ void unpermute_dir(const int axis, const double v1, const double v2, const double v3,
                                              double &vx, double &vy, double &vz) {
  vx = axis == 0 ? v1 : (axis == 1 ? v3 : v2);
  vy = axis == 0 ? v2 : (axis == 1 ? v1 : v3);
  vz = axis == 0 ? v3 : (axis == 1 ? v2 : v1);
}

// DDMC interface condition for a photon that has just arrived on a cell face from an IMC
// region (ptcl_ddmc_albedo, transport_utils.hpp:279-397): admit with the asymptotic
// transmission probability, otherwise send it back out with a cosine-law direction.
// The reference's six per-face branches are one evaluation here: the face is found by the same
// tests in the same order, its parameters are selected, and the arithmetic is done once — same
// operations on the same operands, so the same bits, without six inlined copies of the sampler
// (profiles/r01_v4_ncu_transport_ddmc_cells64.md: 9.8 of 32 lanes active, instruction-fetch stalls).
template <int NDIM>
__device__ __forceinline__ void ddmc_albedo(Rng &rng, const CellEnv &c, const double vv, Photon &p,
                                            bool &is_rejected) {
  constexpr bool multi_d = NDIM >= 2, three_d = NDIM == 3;
  const double dx = c.xu - c.xl, dy = c.yu - c.yl, dz = c.zu - c.zl;
  const double tol = 2.5 * EPS_IMC;
  const int face = fuzzy_eq(p.x, c.xl, dx, tol) ? 0
                 : fuzzy_eq(p.x, c.xu, dx, tol) ? 1
                 : (fuzzy_eq(p.y, c.yl, dy, tol) && multi_d) ? 2
                 : (fuzzy_eq(p.y, c.yu, dy, tol) && multi_d) ? 3
                 : (fuzzy_eq(p.z, c.zl, dz, tol) && three_d) ? 4
                 : (fuzzy_eq(p.z, c.zu, dz, tol) && three_d) ? 5 : -1;
  if (face >= 0) {
    const int axis = face >> 1;
    const bool upper = (face & 1) != 0;
    const double dcell = axis == 0 ? dx : (axis == 1 ? dy : dz);
    const double vn = axis == 0 ? p.vx : (axis == 1 ? p.vy : p.vz);
    const double sign_in = upper ? -1.0 : 1.0;
    const double Pf = (2.0 / 3.0) / ((c.aa + c.ss) * dcell + 2.0 * LAM_EXT);
    const double P = 2.0 * Pf * (1.0 + sign_in * 1.5 * vn / vv);
    if (rng.drand() > P) {
      double v1, v2, v3;
      face_iso_dir(upper ? vv : -vv, rng, v1, v2, v3);
      unpermute_dir(axis, v1, v2, v3, p.vx, p.vy, p.vz);
      const double lo = axis == 0 ? c.xl : (axis == 1 ? c.yl : c.zl);
      const double hi = axis == 0 ? c.xu : (axis == 1 ? c.yu : c.zu);
      const double out = upper ? hi + EPS_IMC * dcell : lo - EPS_IMC * dcell;
      p.x = axis == 0 ? out : p.x;
      p.y = axis == 1 ? out : p.y;
      p.z = axis == 2 ? out : p.z;
      is_rejected = true;
    }
  }
  if (!is_rejected) {
    p.x = 0.5 * (c.xl + c.xu);
    p.y = 0.5 * (c.yl + c.yu);
    p.z = 0.5 * (c.zl + c.zu);
  }
}

// One DDMC event (ptcl_ddmc_step, transport_utils.hpp:163-277): time to the next
// absorption-or-leak event; pick absorb / one of six faces; a leak puts the photon just
// inside the neighbour at the face centre with an outward cosine-law direction; no event
// before census resamples position and direction inside the cell. The face is chosen by the
// reference's cumulative thresholds (same partial sums, same order); the move is done once.
template <int NDIM>
__device__ __forceinline__ void ddmc_step(Rng &rng, const double t_end, const CellEnv &c, const double vv,
                                          const double P[6] /* xl, yl, zl, xu, yu, zu */, Photon &p,
                                          bool &is_absorbed) {
  constexpr int multi_d = NDIM >= 2 ? 1 : 0, three_d = NDIM == 3 ? 1 : 0;
  const double rmin = DBL_MIN;
  const double eps = EPS_DDMC;
  const double dx = c.xu - c.xl, dy = c.yu - c.yl, dz = c.zu - c.zl;
  const double leakx_l = P[0] / dx, leakx_u = P[3] / dx;
  const double leaky_l = P[1] / dy, leaky_u = P[4] / dy;
  const double leakz_l = P[2] / dz, leakz_u = P[5] / dz;
  const double th1 = leakx_l + leakx_u;
  const double th2 = th1 + leaky_l;
  const double th3 = th2 + leaky_u;
  const double th4 = th3 + leakz_l;
  const double leak_tot = th4 + leakz_u;
  const double cdf_ddmc = c.ff * c.aa + leak_tot + rmin;
  const double dt_ddmc = -jbm::log(rng.drand()) / (vv * cdf_ddmc);
  const double dt_end = t_end - p.t;
  const bool is_ddmc_event = dt_ddmc < dt_end;
  const double dt_push = jbm::min2(dt_ddmc, dt_end);
  p.t += dt_push;
  if (is_ddmc_event) {
    const double xi = cdf_ddmc * rng.drand();
    if (xi < c.ff * c.aa) {
      is_absorbed = true;
    } else if (xi < c.ff * c.aa + leak_tot) {
      const double xim = xi - c.ff * c.aa;
      const int face = xim < leakx_l ? 0 : (xim < th1 ? 1 : (xim < th2 ? 2 : (xim < th3 ? 3 : (xim < th4 ? 4 : (xim <= leak_tot ? 5 : -1)))));
      if (face >= 0) {
        const int axis = face >> 1;
        const bool upper = (face & 1) != 0;
        const int d = upper ? 1 : -1;
        p.ip += axis == 0 ? d : 0;
        p.jp += axis == 1 ? d * multi_d : 0;
        p.kp += axis == 2 ? d * three_d : 0;
        const double fx = upper ? c.xu + eps * dx : c.xl - eps * dx;
        const double fy = upper ? c.yu + eps * dy : c.yl - eps * dy;
        const double fz = upper ? c.zu + eps * dz : c.zl - eps * dz;
        p.x = axis == 0 ? fx : c.xl + 0.5 * dx;
        p.y = axis == 1 ? fy : c.yl + 0.5 * dy;
        p.z = axis == 2 ? fz : c.zl + 0.5 * dz;
        double v1, v2, v3;
        face_iso_dir(upper ? vv : -vv, rng, v1, v2, v3);
        unpermute_dir(axis, v1, v2, v3, p.vx, p.vy, p.vz);
      }
    }
  } else {
    p.z = c.zl + rng.drand() * dz;
    p.x = c.xl + rng.drand() * dx;
    p.y = c.yl + rng.drand() * dy;
    const double mu = 1.0 - 2.0 * rng.drand();
    const double nu = sqrt(1.0 - mu * mu);
    const double phi = 2.0 * JB_PI * rng.drand();
    double sn, cs;
    jbm::sincos(phi, &sn, &cs);
    p.vz = vv * mu;
    p.vx = vv * nu * cs;
    p.vy = vv * nu * sn;
  }
}

}  // namespace jbd

#endif  // JAYBENNE_B200_JB_DEVICE_CUH_
