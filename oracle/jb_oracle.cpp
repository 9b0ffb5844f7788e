// jb_oracle.cpp — TEST INFRASTRUCTURE. CPU oracle for the jaybenne IMC/DDMC hot path.
//
// This file is the checker, never the product: only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may build, load or call it. Nothing
// under jaybenne_b200/ links or imports it.
//
// What it is: a line-by-line CPU restatement of the reference's kernels
//   src/jaybenne/sourcing.cpp:25-205        -> source_photons()
//   src/jaybenne/transport.cpp:28-216       -> transport_photons(), check_completion()
//   src/jaybenne/transport_ddmc.cpp:28-237  -> transport_photons() (use_ddmc branch)
//   src/jaybenne/sample_ddmc_bface.cpp      -> sample_ddmc_block_face()
//   src/jaybenne/jaybenne.cpp:68-151        -> radiation_step()   (task order)
//   src/jaybenne/jaybenne.cpp:285-492       -> update_derived_transport_fields()
//   src/jaybenne/jaybenne.cpp:514-615       -> evaluate_radiation_energy(), update_fluid()
//   src/jaybenne/boundaries.hpp:24-84       -> reflect in apply_swarm_bc()
//   src/mcblock/mcblock.cpp:208-262         -> update_derived_material(), post_initialization()
// plus the Parthenon semantics those kernels lean on (Xtoijk, neighbour lookup, swarm
// Send/Receive, ghost cells). Parthenon's source is NOT in the reference mount, so those
// semantics are restated from its documented behaviour: PARITY UNPINNED at that boundary
// (see DESIGN.md §3). The per-particle physics IS pinned: see the build flavours.
//
// Build flavours (oracle/Makefile):
//   libjb_oracle.so      restated physics + jb_math.h  -> bit-exact twin of the CUDA path
//   libjb_oracle_std.so  restated physics + libm       -> -DJBO_MATH_STD
//   _ref/libjb_oracle_ref.so  the reference's OWN headers (transport_utils.hpp,
//                        scattering.hpp, planck.hpp, singularity-opac/eos closures),
//                        compiled verbatim from /root/reference + libm
//                                                       -> -DJBO_USE_REFERENCE
// tests/test_oracle_pin.py requires std == ref bit-for-bit (the restatement is faithful)
// and portable ~ std to 1e-12 event by event.
//
// Random numbers: JB_RNG_PHILOX = per-particle counter streams (what the GPU uses);
// JB_RNG_XORSHIFT_POOL = the reference's Kokkos::Random_XorShift64_Pool
// (Kokkos_Random.hpp:731-962) borrowed in serial loop order, one state per "thread".

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <vector>
#if defined(_OPENMP)
#include <omp.h>
#endif

#include "../include/jaybenne_b200.h"
#include "../jaybenne_b200/csrc/jb_math.h"

namespace jbo {

// =======================================================================================
// random numbers
// =======================================================================================
static inline void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    const uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static const uint32_t STREAM_PARTICLE = 0u;
static const uint32_t STREAM_CELL = 0x5EED0001u;

// Kokkos::Random_XorShift64 (Kokkos_Random.hpp:731-870)
struct XorShift64 {
  uint64_t state;
  explicit XorShift64(uint64_t s = 1318319) : state(s == 0 ? uint64_t(1318319) : s) {}
  uint32_t urand() {
    state ^= state >> 12;
    state ^= state << 25;
    state ^= state >> 27;
    uint64_t tmp = state * 2685821657736338717ULL;
    tmp = tmp >> 16;
    return static_cast<uint32_t>(tmp & 0xffffffffu);
  }
  uint64_t urand64() {
    state ^= state >> 12;
    state ^= state << 25;
    state ^= state >> 27;
    return (state * 2685821657736338717ULL) - 1;
  }
  int rand() { return static_cast<int>(urand() / 2); }
  double drand() {
    return urand64() / static_cast<double>(std::numeric_limits<uint64_t>::max());
  }
};

// one pool entry per cache line pair, so threads that each own a state do not share lines
// (Kokkos pads its state view the same way)
struct alignas(128) PoolState : XorShift64 {
  explicit PoolState(uint64_t s = 1318319) : XorShift64(s) {}
};

// The generator handed to the physics functions. In PHILOX mode it is the particle's
// stream: double #n = words (2*(n&1), 2*(n&1)+1) of Philox block n>>1 with counter
// (n>>1, idx, gid, call) and key (seed_lo, seed_hi ^ stream); value = (u64 >> 11) * 2^-53.
struct RngGen {
  int mode = JB_RNG_PHILOX;
  uint32_t id[3] = {0, 0, 0};
  uint32_t n = 0;
  uint32_t key[2] = {0, 0};
  XorShift64 *xs = nullptr;
  double drand() {
    if (mode != JB_RNG_PHILOX) return xs->drand();
    const uint32_t ctr[4] = {n >> 1, id[0], id[1], id[2]};
    uint32_t o[4];
    philox4x32_10(ctr, key, o);
    const int h = (n & 1u) ? 2 : 0;
    const uint64_t u = ((uint64_t)o[h + 1] << 32) | o[h];
    ++n;
    return (double)(u >> 11) * 1.1102230246251565e-16;
  }
};

}  // namespace jbo

// =======================================================================================
// physics providers
// =======================================================================================
#if defined(JBO_USE_REFERENCE)
// ---- the reference's own code, compiled verbatim from /root/reference ------------------
#include <parthenon/driver.hpp>  // oracle/ref_shim
#define JAYBENNE_JAYBENNE_HPP_   // planck.hpp's #include "jaybenne.hpp" becomes a no-op
#include "transport_utils.hpp"   // -I/root/reference/src/jaybenne
#include "scattering.hpp"
#include "planck.hpp"
#include "../mcblock/opacity.hpp"
#include "../mcblock/eos.hpp"
#define JBO_FLAVOUR "reference-headers+libm"
#else
using Real = double;
#if defined(JBO_MATH_STD)
#define JBO_FLAVOUR "restated+libm"
#else
#define JBO_FLAVOUR "restated+jb_math"
#endif
#endif

namespace jbo {

// ---- math selection for the restated code ---------------------------------------------
namespace M {
#if defined(JBO_MATH_STD) || defined(JBO_USE_REFERENCE)
inline double log(double x) { return std::log(x); }
inline double sin(double x) { return std::sin(x); }
inline double cos(double x) { return std::cos(x); }
inline double acos(double x) { return std::acos(x); }
inline double expm1(double x) { return std::expm1(x); }
inline double exp(double x) { return std::exp(x); }
inline double pow4(double x) { return std::pow(x, 4.0); }
inline double powm4(double j) { return std::pow(j, -4.0); }
#else
inline double log(double x) { return jbm::log(x); }
inline double sin(double x) { return jbm::sin(x); }
inline double cos(double x) { return jbm::cos(x); }
inline double acos(double x) { return jbm::acos(x); }
inline double expm1(double x) { return jbm::expm1(x); }
inline double exp(double x) { return jbm::exp(x); }
inline double pow4(double x) { return jbm::pow4(x); }
inline double powm4(double j) { return 1.0 / ((j * j) * (j * j)); }
#endif
}  // namespace M

// position displacements into next cell (transport_utils.hpp:23-24)
static constexpr double EPS10 = 10 * std::numeric_limits<double>::epsilon();
static constexpr double eps_imc_offset = 1.0e6 * EPS10;
static constexpr double eps_ddmc_offset = 1.0e8 * EPS10;

// ---- CGS constants as singularity-opac derives them (constants/constants.hpp:41-178) ---
namespace cgs {
static constexpr double length = 1.e2, mass = 1.e3, time = 1., temperature = 1.;
static constexpr double force = mass * length / (time * time);
static constexpr double energy = force * length;
static constexpr double h = 6.62606957e-34 * energy * time;
static constexpr double kb = 1.380648800e-23 * energy / temperature;
static constexpr double c = 2.99792458e8 * length / time;
static constexpr double sb =
    2.0 * M_PI * M_PI * M_PI * M_PI * M_PI * kb * kb * kb * kb / (15.0 * h * h * h * c * c);
static constexpr double me = 9.10938291e-31 * mass;
static constexpr double mp = 1.672621777e-27 * mass;
static constexpr double alpha = 7.2973525698e-3;
static constexpr double hbar = h / (2.0 * M_PI);
static constexpr double qe = 1.602176565e-19 * 2.997924580e9;
}  // namespace cgs

// The three closures jaybenne is handed (jaybenne.cpp:227-233): restated arithmetic of
// NonCGSUnits<Gray|EPBremss>, NonCGSUnitsS<GrayS|ThomsonS>, IdealGas.
struct Closures {
  jb_config cfg;
  // unit factors, non_cgs_photons.hpp:33-46
  double time_unit, mass_unit, length_unit, temp_unit, rho_unit, freq_unit, inv_emiss_unit;
  // runtime constants in code units, constants.hpp:191-207
  double c_code, sb_code;
  // pieces of PlanckDistribution::ThermalDistributionOfT evaluated once with libm
  double bt_num, bt_den;  // (8 pi^5) kb^4  and  (15 c^2) h^3
  double sigma_thomson;

#if defined(JBO_USE_REFERENCE)
  mcblock::Opacity opacity;
  mcblock::Scattering scattering;
  mcblock::EOS eos;
#endif

  void init(const jb_config &c) {
    cfg = c;
    time_unit = c.time_scale;
    mass_unit = c.mass_scale;
    length_unit = c.length_scale;
    temp_unit = c.temperature_scale;
    rho_unit = mass_unit / (length_unit * length_unit * length_unit);
    freq_unit = 1. / time_unit;
    inv_emiss_unit = length_unit * time_unit * time_unit / mass_unit;
    c_code = cgs::c * time_unit / length_unit;
    sb_code = cgs::sb * std::pow(time_unit, 3) / mass_unit * std::pow(temp_unit, 4);
    bt_num = 8. * std::pow(M_PI, 5) * std::pow(cgs::kb, 4);
    bt_den = 15. * std::pow(cgs::c, 2) * std::pow(cgs::h, 3);
    sigma_thomson = 8. * M_PI / 3. * std::pow(cgs::alpha * cgs::hbar / (cgs::me * cgs::c), 2);
#if defined(JBO_USE_REFERENCE)
    namespace sp = singularity::photons;
    if (c.opacity_model == JB_OPAC_GRAY) {
      opacity = sp::NonCGSUnits<sp::Gray>(sp::Gray(c.opacity_kappa), c.time_scale, c.mass_scale,
                                          c.length_scale, c.temperature_scale);
    } else {
      opacity = sp::NonCGSUnits<sp::EPBremss>(sp::EPBremss(), c.time_scale, c.mass_scale,
                                              c.length_scale, c.temperature_scale);
    }
    if (c.scattering_model == JB_SCAT_GRAY) {
      scattering = sp::NonCGSUnitsS<sp::GrayS>(sp::GrayS(c.scattering_sigma, c.apm), c.time_scale,
                                               c.mass_scale, c.length_scale, c.temperature_scale);
    } else {
      scattering = sp::NonCGSUnitsS<sp::ThomsonS>(sp::ThomsonS(c.apm), c.time_scale, c.mass_scale,
                                                  c.length_scale, c.temperature_scale);
    }
    eos = singularity::IdealGas(c.eos_gm1, c.eos_cv);
    const auto units = opacity.GetRuntimePhysicalConstants();
    c_code = units.c;
    sb_code = units.sb;
#endif
  }

  // IdealGas, eos_ideal.hpp:55-58,100-103
  double temperature(double rho, double sie) const {
#if defined(JBO_USE_REFERENCE)
    return eos.TemperatureFromDensityInternalEnergy(rho, sie);
#else
    (void)rho;
    return 0.0 > sie / cfg.eos_cv ? 0.0 : sie / cfg.eos_cv;
#endif
  }
  double specific_heat(double rho, double sie) const {
#if defined(JBO_USE_REFERENCE)
    return eos.SpecificHeatFromDensityInternalEnergy(rho, sie);
#else
    (void)rho;
    (void)sie;
    return cfg.eos_cv;
#endif
  }

  // PlanckDistribution::ThermalDistributionOfTNu, thermal_distributions_photons.hpp:31-37
  static double planck_bnu(double temp, double nu) {
    const double x = cgs::h * nu / (cgs::kb * temp);
    return (2. * cgs::h * nu * nu * nu / (cgs::c * cgs::c)) * 1. / M::expm1(x);
  }
  // robust::ratio, robust_utils.hpp:49-52
  static double ratio(double a, double b) {
    const int sgn = (0.0 <= b) - (b < 0.0);
    return a / (b + sgn * (10 * std::numeric_limits<double>::min()));
  }

  // Opacity::AbsorptionCoefficient: non_cgs_photons.hpp:57-67 around
  // gray_opacity_photons.hpp:50-54,86-90 / epbremsstrahlung_opacity_photons.hpp:53-58,88-96
  double absorption(double rho, double temp, double nu) const {
#if defined(JBO_USE_REFERENCE)
    return opacity.AbsorptionCoefficient(rho, temp, nu);
#else
    const double r = rho_unit * rho, T = temp_unit * temp, f = nu * freq_unit;
    double alpha;
    if (cfg.opacity_model == JB_OPAC_GRAY) {
      if (cfg.gray_mode == JB_GRAY_DIRECT) {
        alpha = r * cfg.opacity_kappa;
      } else {
        const double Bnu = planck_bnu(T, f);
        const double jnu = r * cfg.opacity_kappa * Bnu;
        alpha = ratio(jnu, Bnu);
      }
    } else {
      const double Bnu = planck_bnu(T, f);
      const double jnu = epbremss_jnu(r, T, f);
      alpha = ratio(jnu, Bnu);
    }
    return alpha * length_unit;
#endif
  }

  // EPBremss::EmissivityPerNuOmega, epbremsstrahlung_opacity_photons.hpp:88-96
  static double epbremss_jnu(double rho, double temp, double nu) {
    const double thetaE = cgs::kb * temp / (cgs::me * cgs::c * cgs::c);
    const double x = cgs::h * nu / (cgs::kb * temp);
    const double ne = rho / ep_mmw() / 2.;
    const double ni = ne;
    return ep_prefac_v() / std::sqrt(thetaE) * ne * ni * M::exp(-x) * 1.2;
  }
  static double ep_mmw();
  static double ep_prefac_v();

  // Scattering::TotalScatteringCoefficient: non_cgs_s_photons.hpp:65-75 around
  // gray_s_opacity_photons.hpp:58-62 / thomson_s_opacity_photons.hpp:59-64
  double scattering_coeff(double rho, double temp, double nu) const {
#if defined(JBO_USE_REFERENCE)
    return scattering.TotalScatteringCoefficient(rho, temp, nu);
#else
    (void)temp;
    (void)nu;
    const double r = rho_unit * rho;
    double alpha;
    if (cfg.scattering_model == JB_SCAT_GRAY) {
      alpha = (r / cfg.apm) * cfg.scattering_sigma;
    } else {
      const double ne = (r / cfg.apm) / 2.;
      alpha = ne * sigma_thomson;
    }
    return alpha * length_unit;
#endif
  }

  // Opacity::Emissivity: non_cgs_photons.hpp:159-165 around gray_opacity_photons.hpp:118-122
  // and thermal_distributions_photons.hpp:48-52
  double emissivity(double rho, double temp) const {
#if defined(JBO_USE_REFERENCE)
    return opacity.Emissivity(rho, temp);
#else
    const double r = rho * rho_unit, T = temp * temp_unit;
    double J;
    if (cfg.opacity_model == JB_OPAC_GRAY) {
      const double B = bt_num * M::pow4(T) / bt_den;
      J = r * cfg.opacity_kappa * B;
    } else {
      const double thetaE = cgs::kb * T / (cgs::me * cgs::c * cgs::c);
      const double ne = r / ep_mmw() / 2.;
      const double ni = ne;
      J = 4. * M_PI * cgs::kb * T / cgs::h * ep_prefac_v() / std::sqrt(thetaE) * ne * ni * 1.2;
    }
    return J * inv_emiss_unit * time_unit;
#endif
  }
};

// private members of EPBremsstrahlungOpacity, epbremsstrahlung_opacity_photons.hpp:186-196:
// mmw_ = (mp + me)/2, gff_ = 1.2, prefac_ = 8 qe^6 / (3 (me c^2)^2) sqrt(2 pi / 3)
double Closures::ep_mmw() { return (cgs::mp + cgs::me) / 2.; }
double Closures::ep_prefac_v() {
  static const double v =
      8. * std::pow(cgs::qe, 6) / (3. * std::pow(cgs::me * cgs::c * cgs::c, 2)) * std::sqrt(2. * M_PI / 3.);
  return v;
}

// ---- per-particle step functions ------------------------------------------------------
#if defined(JBO_USE_REFERENCE)

inline void imc_step(RngGen &rng, const double &t_start, const double &dt, const double &ff,
                     const double &aa, const double &ss, const double &vv, const double &vx,
                     const double &vy, const double &vz, const double &dx_push, const bool &multi_d,
                     const bool &three_d, const double &xl, const double &yl, const double &zl,
                     const double &xu, const double &yu, const double &zu, double &t, double &x,
                     double &y, double &z, bool &is_absorbed, bool &is_scattered) {
  jaybenne::tran_step_args tra{rng, t_start, dt, ff, aa, ss, vv, vx, vy, vz, dx_push, multi_d, three_d,
                               xl,  yl,      zl, xu, yu, zu, t,  x,  y,  z,  is_absorbed, is_scattered};
  jaybenne::ptcl_transport_step(tra);
}
struct DdmcArgs {
  RngGen &rng;
  const double &t_start, &dt, &ff, &aa, &ss, &vv;
  const bool &multi_d, &three_d;
  const double &xl, &yl, &zl, &xu, &yu, &zu;
  const double &Px_l, &Py_l, &Pz_l, &Px_u, &Py_u, &Pz_u;
  double &t, &x, &y, &z, &vx, &vy, &vz;
  int &ip, &jp, &kp;
  bool &is_absorbed, &is_scattered;
};
inline jaybenne::ddmc_step_args to_ref(DdmcArgs &a) {
  return jaybenne::ddmc_step_args{a.rng, a.t_start, a.dt, a.ff, a.aa, a.ss, a.vv, a.multi_d, a.three_d,
                                  a.xl,  a.yl,      a.zl, a.xu, a.yu, a.zu, a.Px_l, a.Py_l,  a.Pz_l,
                                  a.Px_u, a.Py_u,   a.Pz_u, a.t, a.x, a.y, a.z,   a.vx,    a.vy,
                                  a.vz,  a.ip,      a.jp, a.kp, a.is_absorbed, a.is_scattered};
}
inline void ddmc_step(DdmcArgs &a) { jaybenne::ptcl_ddmc_step(to_ref(a)); }
inline void ddmc_albedo(DdmcArgs &a, bool &is_rejected) { jaybenne::ptcl_ddmc_albedo(to_ref(a), is_rejected); }
inline void scatter_iso(RngGen &rng, const double &vv, double &vx, double &vy, double &vz) {
  jaybenne::scatter(rng, vv, vx, vy, vz);
}
inline double planck_energy(RngGen &rng, const double &sb, const double &temp) {
  return jaybenne::sample_Planck_energy(rng, sb, temp);
}
inline void face_iso_dir(const double vv, RngGen &rng, double &v1, double &v2, double &v3) {
  jaybenne::sample_face_iso_dir(vv, rng, v1, v2, v3);
}
inline bool fuzzy_eq(const double a, const double b, const double c, const double eps) {
  return ::fuzzy_equal(a, b, c, eps);
}

#else  // ---- restated ------------------------------------------------------------------

// transport_utils.hpp:26-39
inline void face_iso_dir(const double vv, RngGen &rng, double &v1, double &v2, double &v3) {
  const double mu = std::sqrt(rng.drand());
  const double nu = std::sqrt(1.0 - mu * mu);
  const double phi = 2.0 * M_PI * rng.drand();
  v1 = vv * mu;
  v2 = vv * nu * M::cos(phi);
  v3 = vv * nu * M::sin(phi);
}

// transport_utils.hpp:110-160
inline void imc_step(RngGen &rng, const double &t_start, const double &dt, const double &ff,
                     const double &aa, const double &ss, const double &vv, const double &vx,
                     const double &vy, const double &vz, const double &dx_cell, const bool &multi_d,
                     const bool &three_d, const double &xl, const double &yl, const double &zl,
                     const double &xu, const double &yu, const double &zu, double &t, double &x,
                     double &y, double &z, bool &is_absorbed, bool &is_scattered) {
  const double rmin = std::numeric_limits<double>::min();
  const double lam_abs = 1.0 / (ff * aa + rmin);
  const double lam_sc = 1.0 / (ss + (1.0 - ff) * aa + rmin);
  const double dx_abs = -lam_abs * M::log(rng.drand());
  const double dx_sc = -lam_sc * M::log(rng.drand());
  const double dx_end = vv * ((t_start + dt) - t);
  double dx_push = std::min(dx_cell, dx_end);
  if (vx > 0.0) {
    dx_push = std::min(dx_push, vv * (xu - x) / vx);
  } else if (vx < 0.0) {
    dx_push = std::min(dx_push, vv * (xl - x) / vx);
  }
  if (multi_d) {
    if (vy > 0.0) {
      dx_push = std::min(dx_push, vv * (yu - y) / vy);
    } else if (vy < 0.0) {
      dx_push = std::min(dx_push, vv * (yl - y) / vy);
    }
  }
  if (three_d) {
    if (vz > 0.0) {
      dx_push = std::min(dx_push, vv * (zu - z) / vz);
    } else if (vz < 0.0) {
      dx_push = std::min(dx_push, vv * (zl - z) / vz);
    }
  }
  is_absorbed = ((dx_abs < dx_push) && (dx_abs < dx_sc));
  is_scattered = (!is_absorbed && (dx_sc < dx_push));
  const double dt_push = (is_absorbed ? dx_abs : (is_scattered ? dx_sc : dx_push)) / vv;
  t += dt_push;
  x += vx * dt_push;
  y += multi_d * vy * dt_push;
  z += three_d * vz * dt_push;
  const double fdx = eps_imc_offset * (xu - xl);
  const double fdy = eps_imc_offset * (yu - yl);
  const double fdz = eps_imc_offset * (zu - zl);
  x = (std::abs(x - xl) < fdx) ? xl - fdx : x;
  x = (std::abs(x - xu) < fdx) ? xu + fdx : x;
  y = (multi_d && std::abs(y - yl) < fdy) ? yl - fdy : y;
  y = (multi_d && std::abs(y - yu) < fdy) ? yu + fdy : y;
  z = (three_d && std::abs(z - zl) < fdz) ? zl - fdz : z;
  z = (three_d && std::abs(z - zu) < fdz) ? zu + fdz : z;
}

struct DdmcArgs {
  RngGen &rng;
  const double &t_start, &dt, &ff, &aa, &ss, &vv;
  const bool &multi_d, &three_d;
  const double &xl, &yl, &zl, &xu, &yu, &zu;
  const double &Px_l, &Py_l, &Pz_l, &Px_u, &Py_u, &Pz_u;
  double &t, &x, &y, &z, &vx, &vy, &vz;
  int &ip, &jp, &kp;
  bool &is_absorbed, &is_scattered;
};

// transport_utils.hpp:162-277
inline void ddmc_step(DdmcArgs &d) {
  const double rmin = std::numeric_limits<double>::min();
  const double eps = eps_ddmc_offset;
  const double dx = d.xu - d.xl, dy = d.yu - d.yl, dz = d.zu - d.zl;
  const double leakx_l = d.Px_l / dx, leakx_u = d.Px_u / dx;
  const double leaky_l = d.Py_l / dy, leaky_u = d.Py_u / dy;
  const double leakz_l = d.Pz_l / dz, leakz_u = d.Pz_u / dz;
  const double leak_tot = leakx_l + leakx_u + leaky_l + leaky_u + leakz_l + leakz_u;
  const double cdf_ddmc = d.ff * d.aa + leak_tot + rmin;
  const double dt_ddmc = -M::log(d.rng.drand()) / (d.vv * cdf_ddmc);
  const double dt_end = (d.t_start + d.dt) - d.t;
  const bool is_ddmc_event = dt_ddmc < dt_end;
  const double dt_push = std::min(dt_ddmc, dt_end);
  d.t += dt_push;
  if (is_ddmc_event) {
    const double xi = cdf_ddmc * d.rng.drand();
    if (xi < d.ff * d.aa) {
      d.is_absorbed = true;
    } else if (xi < d.ff * d.aa + leak_tot) {
      const double xim = xi - d.ff * d.aa;
      if (xim < leakx_l) {
        d.ip -= 1;
        d.x = d.xl - eps * dx;
        d.y = d.yl + 0.5 * dy;
        d.z = d.zl + 0.5 * dz;
        face_iso_dir(-d.vv, d.rng, d.vx, d.vy, d.vz);
      } else if (xim < leakx_l + leakx_u) {
        d.ip += 1;
        d.x = d.xu + eps * dx;
        d.y = d.yl + 0.5 * dy;
        d.z = d.zl + 0.5 * dz;
        face_iso_dir(d.vv, d.rng, d.vx, d.vy, d.vz);
      } else if (xim < leakx_l + leakx_u + leaky_l) {
        d.jp -= d.multi_d;
        d.y = d.yl - eps * dy;
        d.z = d.zl + 0.5 * dz;
        d.x = d.xl + 0.5 * dx;
        face_iso_dir(-d.vv, d.rng, d.vy, d.vz, d.vx);
      } else if (xim < leakx_l + leakx_u + leaky_l + leaky_u) {
        d.jp += d.multi_d;
        d.y = d.yu + eps * dy;
        d.z = d.zl + 0.5 * dz;
        d.x = d.xl + 0.5 * dx;
        face_iso_dir(d.vv, d.rng, d.vy, d.vz, d.vx);
      } else if (xim < leakx_l + leakx_u + leaky_l + leaky_u + leakz_l) {
        d.kp -= d.three_d;
        d.z = d.zl - eps * dz;
        d.x = d.xl + 0.5 * dx;
        d.y = d.yl + 0.5 * dy;
        face_iso_dir(-d.vv, d.rng, d.vz, d.vx, d.vy);
      } else if (xim <= leak_tot) {
        d.kp += d.three_d;
        d.z = d.zu + eps * dz;
        d.x = d.xl + 0.5 * dx;
        d.y = d.yl + 0.5 * dy;
        face_iso_dir(d.vv, d.rng, d.vz, d.vx, d.vy);
      }
    }
  } else {
    d.z = d.zl + d.rng.drand() * dz;
    d.x = d.xl + d.rng.drand() * dx;
    d.y = d.yl + d.rng.drand() * dy;
    const double mu = 1.0 - 2.0 * d.rng.drand();
    const double nu = std::sqrt(1.0 - mu * mu);
    const double phi = 2.0 * M_PI * d.rng.drand();
    d.vz = d.vv * mu;
    d.vx = d.vv * nu * M::cos(phi);
    d.vy = d.vv * nu * M::sin(phi);
  }
}

inline bool fuzzy_eq(const double a, const double b, const double c, const double eps) {
  return std::abs(a - b) < c * eps;  // jaybenne_utils.hpp:44-49
}

// transport_utils.hpp:279-397. One helper per face keeps the six cases literal.
inline void albedo_face(DdmcArgs &d, bool &is_rejected, const double dcell, const double vn,
                        const double sign_in, const double vsample, double &v1, double &v2,
                        double &v3, double &coord, const double coord_out) {
  constexpr double lam_ext = 0.7104;
  const double Pf = (2.0 / 3.0) / ((d.aa + d.ss) * dcell + 2.0 * lam_ext);
  const double P = 2.0 * Pf * (1.0 + sign_in * 1.5 * vn / d.vv);
  if (d.rng.drand() > P) {
    face_iso_dir(vsample, d.rng, v1, v2, v3);
    coord = coord_out;
    is_rejected = true;
  }
}
inline void ddmc_albedo(DdmcArgs &d, bool &is_rejected) {
  const double dx = d.xu - d.xl, dy = d.yu - d.yl, dz = d.zu - d.zl;
  const double tol = 2.5 * eps_imc_offset;
  if (fuzzy_eq(d.x, d.xl, dx, tol)) {
    albedo_face(d, is_rejected, dx, d.vx, +1.0, -d.vv, d.vx, d.vy, d.vz, d.x, d.xl - eps_imc_offset * dx);
  } else if (fuzzy_eq(d.x, d.xu, dx, tol)) {
    albedo_face(d, is_rejected, dx, d.vx, -1.0, d.vv, d.vx, d.vy, d.vz, d.x, d.xu + eps_imc_offset * dx);
  } else if (fuzzy_eq(d.y, d.yl, dy, tol) && d.multi_d) {
    albedo_face(d, is_rejected, dy, d.vy, +1.0, -d.vv, d.vy, d.vz, d.vx, d.y, d.yl - eps_imc_offset * dy);
  } else if (fuzzy_eq(d.y, d.yu, dy, tol) && d.multi_d) {
    albedo_face(d, is_rejected, dy, d.vy, -1.0, d.vv, d.vy, d.vz, d.vx, d.y, d.yu + eps_imc_offset * dy);
  } else if (fuzzy_eq(d.z, d.zl, dz, tol) && d.three_d) {
    albedo_face(d, is_rejected, dz, d.vz, +1.0, -d.vv, d.vz, d.vx, d.vy, d.z, d.zl - eps_imc_offset * dz);
  } else if (fuzzy_eq(d.z, d.zu, dz, tol) && d.three_d) {
    albedo_face(d, is_rejected, dz, d.vz, -1.0, d.vv, d.vz, d.vx, d.vy, d.z, d.zu + eps_imc_offset * dz);
  }
  if (!is_rejected) {
    d.x = 0.5 * (d.xl + d.xu);
    d.y = 0.5 * (d.yl + d.yu);
    d.z = 0.5 * (d.zl + d.zu);
  }
}

// scattering.hpp:21-29
inline void scatter_iso(RngGen &rng, const double &vv, double &vx, double &vy, double &vz) {
  const double mu = 2.0 * rng.drand() - 1.0;
  const double phi = 2.0 * M_PI * rng.drand();
  const double stheta = std::sqrt(1.0 - mu * mu);
  vx = vv * stheta * M::cos(phi);
  vy = vv * stheta * M::sin(phi);
  vz = vv * mu;
}

// planck.hpp:27-50 (Everett & Cashwell). The reference recomputes sum_{j<=l} j^-4 from
// j = 1 for every l; the running sum below adds the same terms in the same order.
inline double planck_energy(RngGen &rng, const double &sb, const double &temp) {
  static const double pi4 = std::pow(M_PI, 4.0);
  const double xi0 = rng.drand();
  const double rhs = xi0 * pi4 / 90.0;
  int l = 1;
  double ll = 1.0;
  double lhs = 0.0;
  while (l < 100) {
    lhs += M::powm4(static_cast<double>(l));
    if (lhs >= rhs) {
      ll = static_cast<double>(l);
      break;
    }
    l++;
  }
  const double xi1 = rng.drand();
  const double xi2 = rng.drand();
  const double xi3 = rng.drand();
  const double xi4 = rng.drand();
  return -(1.0 / ll) * M::log(xi1 * xi2 * xi3 * xi4) * sb * temp;
}
#endif  // physics providers

// =======================================================================================
// mesh, fields and the photons swarm
// =======================================================================================
struct Particle {
  double x, y, z, t, w, e, vx, vy, vz;
  int i, j, k;  // interior cell index (the reference's ijk minus the ghost offset)
  uint32_t rid[3];
  uint32_t rn;
  bool marked;
};

struct ExRecord {  // what jb_exchange_take / give move between ranks (96 bytes)
  double x, y, z, t, w, e, vx, vy, vz;
  uint32_t rng[4];
  int32_t dest_gid;
  int32_t pad;
};

struct Block {
  jb_block info;
  double dx[3];
  double xmin0[3];  // coordinate of ghost-inclusive index 0 (UniformCartesian::xmin_)
  std::vector<double> rho, sie, u, tally, fleck, src_ew, src_num, edelta, fp[3];
  std::vector<Particle> pool;
};

struct Ctx {
  jb_config cfg;
  Closures cl;
  jb_mesh mesh;
  std::vector<jb_block> all_blocks;
  std::vector<Block> blocks;           // local
  std::vector<int> gid_to_local;       // -1 if remote
  std::vector<int> lattice;            // finest-level block lattice -> gid
  int nlat[3] = {1, 1, 1};
  int ng[3] = {0, 0, 0};               // ghost offset per dim
  int my_rank = 0, n_ranks = 1;
  bool mesh_set = false;
  uint32_t source_call = 0;
  std::vector<PoolState> pool;        // reference-faithful generator pool
  std::vector<std::vector<ExRecord>> outbox;  // per peer rank
  jb_cycle_stats stats;
  std::vector<jb_event> events;
  bool log_events = false;
  int64_t event_cap = 0;
  std::string err;
  int ncell() const { return mesh.nxb[0] * mesh.nxb[1] * mesh.nxb[2]; }
};

static jb_status fail(Ctx *c, jb_status s, const std::string &m) {
  if (c) c->err = m;
  return s;
}

// which pool state the calling thread borrows: state 0 in the serial, reproducible mode; one
// state per OpenMP thread in JB_RNG_XORSHIFT_THREADS (what Random_XorShift64_Pool hands out on
// the Kokkos OpenMP backend, where num_states = concurrency()) — timing builds only
static inline int pool_slot(int mode) {
#if defined(_OPENMP)
  if (mode == JB_RNG_XORSHIFT_THREADS) return omp_get_thread_num();
#endif
  (void)mode;
  return 0;
}

// Kokkos::Random_XorShift64_Pool::init (Kokkos_Random.hpp:891-940)
static void seed_pool(Ctx *c, int num_states) {
  uint64_t seed = (uint64_t)c->cfg.seed;  // Q2: the raw seed, never seed + rank
  if (seed == 0) seed = uint64_t(1318319);
  XorShift64 gen(seed);
  for (int i = 0; i < 17; i++) gen.rand();
  c->pool.clear();
  for (int i = 0; i < num_states; i++) {
    int n1 = gen.rand();
    int n2 = gen.rand();
    int n3 = gen.rand();
    int n4 = gen.rand();
    uint64_t s = (((static_cast<uint64_t>(n1)) & 0xffff) << 00) | (((static_cast<uint64_t>(n2)) & 0xffff) << 16) |
                 (((static_cast<uint64_t>(n3)) & 0xffff) << 32) | (((static_cast<uint64_t>(n4)) & 0xffff) << 48);
    c->pool.push_back(PoolState(s));
  }
}

static RngGen particle_rng(Ctx *c, const Particle &p) {
  RngGen g;
  g.mode = c->cfg.rng_mode;
  if (g.mode != JB_RNG_PHILOX) {
    g.xs = &c->pool[pool_slot(g.mode)];
    return g;
  }
  g.id[0] = p.rid[0];
  g.id[1] = p.rid[1];
  g.id[2] = p.rid[2];
  g.n = p.rn;
  g.key[0] = (uint32_t)((uint64_t)c->cfg.seed);
  g.key[1] = (uint32_t)((uint64_t)c->cfg.seed >> 32) ^ STREAM_PARTICLE;
  return g;
}
static RngGen cell_rng(Ctx *c, uint32_t cell, uint32_t gid, uint32_t call) {
  RngGen g;
  g.mode = c->cfg.rng_mode;
  if (g.mode != JB_RNG_PHILOX) {
    g.xs = &c->pool[pool_slot(g.mode)];
    return g;
  }
  g.id[0] = cell;
  g.id[1] = gid;
  g.id[2] = call;
  g.n = 0;
  g.key[0] = (uint32_t)((uint64_t)c->cfg.seed);
  g.key[1] = (uint32_t)((uint64_t)c->cfg.seed >> 32) ^ STREAM_CELL;
  return g;
}

// UniformCartesian::Xc<dir>(idx) = xmin_ + (idx + 0.5) * dx_, idx ghost-inclusive
static inline double Xc(const Ctx *c, const Block &b, int d, int interior_idx) {
  return b.xmin0[d] + ((interior_idx + c->ng[d]) + 0.5) * b.dx[d];
}
// SwarmDeviceContext::Xtoijk (interior index; inactive dims -> 0)
static inline void Xtoijk(const Ctx *c, const Block &b, double x, double y, double z, int &i, int &j, int &k) {
  i = static_cast<int>(std::floor((x - b.info.xmin[0]) / b.dx[0]));
  j = (c->mesh.ndim > 1) ? static_cast<int>(std::floor((y - b.info.xmin[1]) / b.dx[1])) : 0;
  k = (c->mesh.ndim > 2) ? static_cast<int>(std::floor((z - b.info.xmin[2]) / b.dx[2])) : 0;
}
// GetNeighborBlockIndex(...)'s on_current_mesh_block
static inline bool on_block(const Ctx *c, int i, int j, int k) {
  return i >= 0 && i < c->mesh.nxb[0] && j >= 0 && j < c->mesh.nxb[1] && k >= 0 && k < c->mesh.nxb[2];
}
static inline int cidx(const Ctx *c, int k, int j, int i) {
  return (k * c->mesh.nxb[1] + j) * c->mesh.nxb[0] + i;
}

static void log_event(Ctx *c, const Particle &p, uint32_t draw, int kind, int gid, const double v[3]) {
  if (!c->log_events || (int64_t)c->events.size() >= c->event_cap) return;
  jb_event e;
  memset(&e, 0, sizeof(e));
  e.id_idx = p.rid[0];
  e.id_gid = p.rid[1];
  e.id_call = p.rid[2];
  e.draw = draw;
  e.kind = kind;
  e.block = gid;
  e.i = p.i;
  e.j = p.j;
  e.k = p.k;
  e.t = p.t;
  e.x = p.x;
  e.y = p.y;
  e.z = p.z;
  e.v0 = v[0];
  e.v1 = v[1];
  e.v2 = v[2];
  c->events.push_back(e);
}

// ---- block lookup -----------------------------------------------------------------------
static int lattice_at(const Ctx *c, int lx, int ly, int lz) {
  return c->lattice[((size_t)lz * c->nlat[1] + ly) * c->nlat[0] + lx];
}

// gid of the block containing (x,y,z) (already inside the domain), -1 on failure
static int find_block(const Ctx *c, const double pos[3]) {
  int l[3] = {0, 0, 0};
  for (int d = 0; d < c->mesh.ndim; ++d) {
    const double w = (c->mesh.xmax[d] - c->mesh.xmin[d]) / c->nlat[d];
    int v = static_cast<int>(std::floor((pos[d] - c->mesh.xmin[d]) / w));
    l[d] = std::min(std::max(v, 0), c->nlat[d] - 1);
  }
  for (int it = 0; it < 8; ++it) {
    const int gid = lattice_at(c, l[0], l[1], l[2]);
    const jb_block &b = c->all_blocks[gid];
    bool ok = true;
    for (int d = 0; d < c->mesh.ndim; ++d) {
      const double dxd = (b.xmax[d] - b.xmin[d]) / c->mesh.nxb[d];
      const int ii = static_cast<int>(std::floor((pos[d] - b.xmin[d]) / dxd));
      const int s = 1 << (c->mesh.max_level - b.level);
      if (ii < 0) {
        l[d] = (int)b.lx[d] * s - 1;
        ok = false;
      } else if (ii >= c->mesh.nxb[d]) {
        l[d] = (int)(b.lx[d] + 1) * s;
        ok = false;
      }
      if (l[d] < 0 || l[d] >= c->nlat[d]) return -1;
    }
    if (ok) return gid;
  }
  return -1;
}

// Swarm boundary conditions applied when a particle is handed to Send: periodic wrap,
// PhotonReflectBC (boundaries.hpp:24-84) or outflow removal. Returns false if the
// particle left through an outflow face.
static bool apply_swarm_bc(const Ctx *c, Particle &p) {
  double *pos[3] = {&p.x, &p.y, &p.z};
  double *vel[3] = {&p.vx, &p.vy, &p.vz};
  for (int d = 0; d < c->mesh.ndim; ++d) {
    const double lo = c->mesh.xmin[d], hi = c->mesh.xmax[d];
    if (*pos[d] < lo) {
      const int bc = c->mesh.swarm_bc[2 * d];
      if (bc == JB_BC_PERIODIC) {
        *pos[d] = *pos[d] + (hi - lo);
      } else if (bc == JB_BC_REFLECT) {
        *pos[d] = lo + (lo - *pos[d]);
        *vel[d] = -*vel[d];
      } else {
        return false;
      }
    } else if (*pos[d] >= hi) {
      const int bc = c->mesh.swarm_bc[2 * d + 1];
      if (bc == JB_BC_PERIODIC) {
        *pos[d] = *pos[d] - (hi - lo);
      } else if (bc == JB_BC_REFLECT) {
        if (*pos[d] > hi) {
          *pos[d] = hi - (*pos[d] - hi);
          *vel[d] = -*vel[d];
        }
      } else {
        return false;
      }
    }
  }
  return true;
}

// value of a cell field at level-L global cell G (ghost fill semantics: same level copy,
// coarser neighbour -> covering coarse cell, finer neighbour -> mean of the covering fine
// cells; outflow -> clamp to the edge cell; periodic -> wrap). Needs every block local.
static double sample_cell(const Ctx *c, int field /*0 rho, 1 sie*/, int L, const long Gin[3]) {
  long G[3] = {Gin[0], Gin[1], Gin[2]};
  for (int d = 0; d < 3; ++d) {
    const long n = (long)c->mesh.nrb[d] * c->mesh.nxb[d] << (d < c->mesh.ndim ? L : 0);
    if (G[d] < 0 || G[d] >= n) {
      const bool periodic = c->mesh.mesh_bc[2 * d + (G[d] < 0 ? 0 : 1)] == JB_BC_PERIODIC;
      if (periodic) {
        G[d] = ((G[d] % n) + n) % n;
      } else {
        G[d] = std::min(std::max(G[d], 0L), n - 1);
      }
    }
  }
  int l[3] = {0, 0, 0};
  for (int d = 0; d < c->mesh.ndim; ++d) {
    l[d] = (int)((G[d] << (c->mesh.max_level - L)) / c->mesh.nxb[d]);
  }
  const int gid = lattice_at(c, l[0], l[1], l[2]);
  const jb_block &nb = c->all_blocks[gid];
  const int loc = c->gid_to_local[gid];
  if (loc < 0) return std::numeric_limits<double>::quiet_NaN();
  const Block &B = c->blocks[loc];
  const std::vector<double> &f = field == 0 ? B.rho : B.sie;
  if (nb.level <= L) {
    int idx[3] = {0, 0, 0};
    for (int d = 0; d < c->mesh.ndim; ++d) {
      idx[d] = (int)((G[d] >> (L - nb.level)) - nb.lx[d] * c->mesh.nxb[d]);
    }
    return f[cidx(c, idx[2], idx[1], idx[0])];
  }
  // finer neighbour (one level): mean over the children cells. The children may live in
  // different blocks along the face-parallel directions, so look each one up.
  double sum = 0.0;
  int cnt = 0;
  const int nk = c->mesh.ndim > 2 ? 2 : 1, nj = c->mesh.ndim > 1 ? 2 : 1;
  for (int ck = 0; ck < nk; ++ck)
    for (int cj = 0; cj < nj; ++cj)
      for (int ci = 0; ci < 2; ++ci) {
        long Gc[3] = {2 * G[0] + ci, c->mesh.ndim > 1 ? 2 * G[1] + cj : G[1], c->mesh.ndim > 2 ? 2 * G[2] + ck : G[2]};
        sum += sample_cell(c, field, L + 1, Gc);
        cnt++;
      }
  return sum / cnt;
}

// level of the leaf across face (dir, side) of block b; returns b's own level at a
// non-periodic domain boundary (IsPhysicalBoundary branch, jaybenne.cpp:345-351)
static int neighbor_level(const Ctx *c, const jb_block &b, int dir, int side) {
  const int s = 1 << (c->mesh.max_level - b.level);
  int l[3];
  for (int d = 0; d < 3; ++d) l[d] = d < c->mesh.ndim ? (int)b.lx[d] * s : 0;
  l[dir] = side == 0 ? (int)b.lx[dir] * s - 1 : (int)(b.lx[dir] + 1) * s;
  if (l[dir] < 0 || l[dir] >= c->nlat[dir]) {
    if (c->mesh.mesh_bc[2 * dir + side] != JB_BC_PERIODIC) return b.level;
    l[dir] = (l[dir] + c->nlat[dir]) % c->nlat[dir];
  }
  return c->all_blocks[lattice_at(c, l[0], l[1], l[2])].level;
}

// =======================================================================================
// kernels
// =======================================================================================

// UpdateDerivedTransportFields, jaybenne.cpp:285-492
static jb_status update_derived_transport_fields(Ctx *c, double dt) {
  const int nx1 = c->mesh.nxb[0], nx2 = c->mesh.nxb[1], nx3 = c->mesh.nxb[2];
  for (Block &b : c->blocks) {
    for (int k = 0; k < nx3; ++k)
      for (int j = 0; j < nx2; ++j)
        for (int i = 0; i < nx1; ++i) {
          const int ci = cidx(c, k, j, i);
          const double rho = b.rho[ci], sie = b.sie[ci];
          const double temp = c->cl.temperature(rho, sie);
          const double cv = c->cl.specific_heat(rho, sie);
          const double emis = c->cl.emissivity(rho, temp);
          b.fleck[ci] = 1.0 / (1.0 + (4.0 * emis / (rho * cv * temp)) * dt);
        }
  }
  if (!c->cfg.use_ddmc) return JB_OK;
  for (size_t lb = 0; lb < c->blocks.size(); ++lb)
    if (c->gid_to_local[c->blocks[lb].info.gid] < 0) return fail(c, JB_ERR_STATE, "ddmc needs local blocks");
  if ((int)c->blocks.size() != c->mesh.nblocks)
    return fail(c, JB_ERR_UNSUPPORTED, "oracle DDMC face probabilities need every block on one rank");
  constexpr double lam_ext = 0.7104;
  const double tau_ddmc = c->cfg.tau_ddmc;
  for (Block &b : c->blocks) {
    const int L = b.info.level;
    const double rlev = static_cast<double>(L);
    for (int dir = 0; dir < c->mesh.ndim; ++dir) {
      const double rlev_l = static_cast<double>(neighbor_level(c, b.info, dir, 0));
      const double rlev_u = static_cast<double>(neighbor_level(c, b.info, dir, 1));
      const double dxd = b.dx[dir];
      int n[3] = {nx1, nx2, nx3};
      const int nu = n[dir];  // face index runs 0..n[dir]
      n[dir] += 1;
      b.fp[dir].assign((size_t)n[0] * n[1] * n[2], 0.0);
      for (int k = 0; k < n[2]; ++k)
        for (int j = 0; j < n[1]; ++j)
          for (int i = 0; i < n[0]; ++i) {
            const int f[3] = {i, j, k};
            const double dx_l = f[dir] == 0 ? std::pow(2.0, rlev - rlev_l) * dxd : dxd;
            const double dx_u = f[dir] == nu ? std::pow(2.0, rlev - rlev_u) * dxd : dxd;
            long Gl[3], Gu[3];
            for (int d = 0; d < 3; ++d) {
              const long base = d < c->mesh.ndim ? b.info.lx[d] * c->mesh.nxb[d] : 0;
              Gu[d] = base + f[d];
              Gl[d] = base + f[d] - (d == dir ? 1 : 0);
            }
            const double rho_l = sample_cell(c, 0, L, Gl), sie_l = sample_cell(c, 1, L, Gl);
            const double rho_u = sample_cell(c, 0, L, Gu), sie_u = sample_cell(c, 1, L, Gu);
            const double temp_l = c->cl.temperature(rho_l, sie_l);
            const double temp_u = c->cl.temperature(rho_u, sie_u);
            const double ss_l = c->cl.scattering_coeff(rho_l, temp_l, 1.0);
            const double aa_l = c->cl.absorption(rho_l, temp_l, 1.0);
            const double ss_u = c->cl.scattering_coeff(rho_u, temp_u, 1.0);
            const double aa_u = c->cl.absorption(rho_u, temp_u, 1.0);
            double tau_l = dx_l * (ss_l + aa_l);
            double tau_u = dx_u * (ss_u + aa_u);
            tau_l = tau_l > tau_ddmc ? tau_l : 2.0 * lam_ext;
            tau_u = tau_u > tau_ddmc ? tau_u : 2.0 * lam_ext;
            b.fp[dir][((size_t)k * n[1] + j) * n[0] + i] = 2.0 / (3.0 * (tau_l + tau_u));
          }
    }
  }
  return JB_OK;
}

// SourcePhotons<T, ST>, sourcing.cpp:25-205. `blocks` = the MeshData / MeshBlockData the
// reference call covers; nblocks_npc = vmesh.GetNBlocks() of that call.
static jb_status source_photons(Ctx *c, const std::vector<int> &which, int source_type, double t_start,
                                double dt, int nblocks_npc) {
  if (c->cfg.source_strategy == JB_STRATEGY_ENERGY)
    return fail(c, JB_ERR_UNSUPPORTED, "Energy source strategy not implemented!");
  if (source_type == JB_SOURCE_EMISSION && !c->cfg.do_emission) return JB_OK;
  const double vv = c->cl.c_code, sb = c->cl.sb_code;
  const int nx1 = c->mesh.nxb[0], nx2 = c->mesh.nxb[1], nx3 = c->mesh.nxb[2];
  const int num_cells = nx1 * nx2 * nx3;
  const double npc =
      static_cast<double>(c->cfg.num_particles) / num_cells / (nblocks_npc * c->mesh.nblocks);
  const uint32_t call = c->source_call;
  std::vector<std::vector<int>> prefix(which.size(), std::vector<int>(num_cells, 0));
  const bool serial = c->cfg.rng_mode == JB_RNG_XORSHIFT_POOL;
  (void)serial;
  // pass 1 (SourcePhotons1); timing builds take the blocks in parallel (every cell has its own stream)
#if defined(_OPENMP)
#pragma omp parallel for schedule(dynamic, 1) if (!serial)
#endif
  for (size_t w = 0; w < which.size(); ++w) {
    Block &b = c->blocks[which[w]];
    const double dv = b.dx[0] * b.dx[1] * b.dx[2];
    for (int k = 0; k < nx3; ++k)
      for (int j = 0; j < nx2; ++j)
        for (int i = 0; i < nx1; ++i) {
          const int ci = cidx(c, k, j, i);
          RngGen rng = cell_rng(c, (uint32_t)ci, (uint32_t)b.info.gid, call);
          const double rho = b.rho[ci], sie = b.sie[ci];
          const double temp = c->cl.temperature(rho, sie);
          double erad = 0.0;
          if (source_type == JB_SOURCE_THERMAL) {
            erad = (4.0 * sb / vv) * M::pow4(temp) * dv;
          } else {
            const double emis = c->cl.emissivity(rho, temp);
            erad = b.fleck[ci] * emis * dv * dt;
          }
          double snpc = std::floor(npc);
          snpc += ((npc - snpc) > rng.drand());
          b.src_num[ci] = snpc;
          b.src_ew[ci] = erad / snpc;
        }
    int update = 0;
    for (int idx = 0; idx < num_cells; ++idx) {
      prefix[w][idx] = update;
      update += static_cast<int>(std::round(b.src_num[idx]));
    }
  }
  // pass 2 (SourcePhotons2). The new photons of cell ci occupy pool slots base + prefix[ci]..,
  // i.e. the order a serial push_back would give, so timing builds can run the cells of all blocks in
  // one parallel region (no barrier between blocks); sums are kept per block and added in block order.
  std::vector<size_t> bases(which.size(), 0);
#if defined(_OPENMP)
#pragma omp parallel for schedule(dynamic, 1) if (!serial)
#endif
  for (size_t w = 0; w < which.size(); ++w) {
    Block &b = c->blocks[which[w]];
    bases[w] = b.pool.size();
    const int last = num_cells - 1;
    const size_t nnew = (size_t)prefix[w][last] + (size_t)static_cast<int>(std::round(b.src_num[last]));
    Particle blank;
    memset(&blank, 0, sizeof(blank));
    b.pool.resize(bases[w] + nnew, blank);
  }
  std::vector<long long> sum_n(which.size(), 0);
  std::vector<double> sum_e(which.size(), 0.0);
#if defined(_OPENMP)
#pragma omp parallel if (!serial)
#endif
  for (size_t w = 0; w < which.size(); ++w) {
    Block &b = c->blocks[which[w]];
    const size_t base = bases[w];
    long long r_n = 0;
    double r_e = 0.0;
#if defined(_OPENMP)
#pragma omp for collapse(3) schedule(static) nowait
#endif
    for (int k = 0; k < nx3; ++k)
      for (int j = 0; j < nx2; ++j)
        for (int i = 0; i < nx1; ++i) {
          const int ci = cidx(c, k, j, i);
          const double xi = Xc(c, b, 0, i), yi = Xc(c, b, 1, j), zi = Xc(c, b, 2, k);
          const int pstart = prefix[w][ci];
          const int npart = static_cast<int>(std::round(b.src_num[ci]));
          double &dejbn = b.edelta[ci];
          dejbn = 0.0;
          for (int np = pstart; np < pstart + npart; np++) {
            Particle p;
            memset(&p, 0, sizeof(p));
            p.i = i;
            p.j = j;
            p.k = k;
            p.rid[0] = (uint32_t)np;
            p.rid[1] = (uint32_t)b.info.gid;
            p.rid[2] = call;
            p.rn = 0;
            RngGen rng = particle_rng(c, p);
            p.x = xi + b.dx[0] * (rng.drand() - 0.5);
            p.y = yi + b.dx[1] * (rng.drand() - 0.5);
            p.z = zi + b.dx[2] * (rng.drand() - 0.5);
            const double theta = M::acos(2.0 * rng.drand() - 1.0);
            const double phi = 2.0 * M_PI * rng.drand();
            const double stheta = M::sin(theta);
            p.vx = vv * stheta * M::cos(phi);
            p.vy = vv * stheta * M::sin(phi);
            p.vz = vv * M::cos(theta);
            const double rho = b.rho[ci], sie = b.sie[ci];
            const double temp = c->cl.temperature(rho, sie);
            p.e = planck_energy(rng, sb, temp);
            p.w = b.src_ew[ci];
            if (source_type == JB_SOURCE_EMISSION) {
              dejbn -= p.w;
              p.t = t_start + rng.drand() * dt;
            } else {
              p.t = 0.;
            }
            p.rn = rng.n;
            b.pool[base + (size_t)np] = p;
            r_n++;
            r_e += p.w;
          }
        }
#if defined(_OPENMP)
#pragma omp critical(jbo_source_sums)
#endif
    {
      sum_n[w] += r_n;
      sum_e[w] += r_e;
    }
  }
  for (size_t w = 0; w < which.size(); ++w) {
    c->stats.n_sourced += sum_n[w];
    c->stats.e_sourced += sum_e[w];
  }
  c->source_call++;
  return JB_OK;
}

static void remove_marked(Block &b) {
  b.pool.erase(std::remove_if(b.pool.begin(), b.pool.end(), [](const Particle &p) { return p.marked; }),
               b.pool.end());
}

// TransportPhotons (transport.cpp:28-181) and TransportPhotons_DDMC
// (transport_ddmc.cpp:28-237); the DDMC flavour reduces to the IMC one when
// is_ddmc_step is false, exactly as in the reference.
static jb_status transport_photons(Ctx *c, double t_start, double dt) {
  const double vv = c->cl.c_code;
  const bool multi_d = c->mesh.ndim >= 2, three_d = c->mesh.ndim == 3;
  const bool use_ddmc = c->cfg.use_ddmc != 0;
  const double tau_ddmc = c->cfg.tau_ddmc;
  const int nx1 = c->mesh.nxb[0], nx2 = c->mesh.nxb[1];
  // Timing builds (-fopenmp) spread the photons of ALL blocks over the host threads in one parallel region
  // without a barrier between blocks — the flat (block, particle) par_for of the reference's MeshData pack
  // on the Kokkos OpenMP backend (transport.cpp:64-68). Counters are summed per block (thread partial sums
  // added under a critical section), so a build without OpenMP — the parity builds — runs the very loop
  // and the very summation order it always had; the tally becomes an atomic add.
  const size_t nblk = c->blocks.size();
  struct BlkSums { long long seg = 0, ddmc = 0, abs = 0, scat = 0; double eabs = 0.0; };
  std::vector<BlkSums> sums(nblk);
  const bool serial = c->log_events || c->cfg.rng_mode == JB_RNG_XORSHIFT_POOL;
  (void)serial;
#if defined(_OPENMP)
#pragma omp parallel if (!serial)
#endif
  for (size_t bi = 0; bi < nblk; ++bi) {
    Block &b = c->blocks[bi];
    const double dx_i = b.dx[0], dx_j = b.dx[1], dx_k = b.dx[2];
    const double dx_push = std::min(dx_i, std::min(dx_j, dx_k));
    long long r_seg = 0, r_ddmc = 0, r_abs = 0, r_scat = 0;
    double r_eabs = 0.0;
    const long long npool = (long long)b.pool.size();
#if defined(_OPENMP)
#pragma omp for schedule(dynamic, 1024) nowait
#endif
    for (long long pi = 0; pi < npool; ++pi) {
      Particle &p = b.pool[pi];
      RngGen rng = particle_rng(c, p);
      double &t = p.t, &vx = p.vx, &vy = p.vy, &vz = p.vz;
      const double ww = p.w, ee = p.e;
      double &x = p.x, &y = p.y, &z = p.z;
      int &ip = p.i, &jp = p.j, &kp = p.k;
      Xtoijk(c, b, x, y, z, ip, jp, kp);
      while (t < t_start + dt) {
        const uint32_t draw0 = rng.n;
        r_seg++;
        double xl = Xc(c, b, 0, ip) - 0.5 * dx_i, xu = Xc(c, b, 0, ip) + 0.5 * dx_i;
        double yl = Xc(c, b, 1, jp) - 0.5 * dx_j, yu = Xc(c, b, 1, jp) + 0.5 * dx_j;
        double zl = Xc(c, b, 2, kp) - 0.5 * dx_k, zu = Xc(c, b, 2, kp) + 0.5 * dx_k;
        const int ci = cidx(c, kp, jp, ip);
        const double rho = b.rho[ci], sie = b.sie[ci];
        const double temp = c->cl.temperature(rho, sie);
        const double ff = b.fleck[ci];
        const double ss = c->cl.scattering_coeff(rho, temp, ee);
        const double aa = c->cl.absorption(rho, temp, ee);
        bool is_absorbed = false, is_scattered = false, is_rejected = false;
        const bool is_ddmc_step = use_ddmc && (dx_push * (ss + aa) > tau_ddmc);
        if (is_ddmc_step) {
          r_ddmc++;
          Xtoijk(c, b, x, y, z, ip, jp, kp);
          xl = Xc(c, b, 0, ip) - 0.5 * dx_i, xu = Xc(c, b, 0, ip) + 0.5 * dx_i;
          yl = Xc(c, b, 1, jp) - 0.5 * dx_j, yu = Xc(c, b, 1, jp) + 0.5 * dx_j;
          zl = Xc(c, b, 2, kp) - 0.5 * dx_k, zu = Xc(c, b, 2, kp) + 0.5 * dx_k;
          const double Px_l = b.fp[0][((size_t)kp * nx2 + jp) * (nx1 + 1) + ip];
          const double Px_u = b.fp[0][((size_t)kp * nx2 + jp) * (nx1 + 1) + ip + 1];
          const double Py_l = multi_d ? b.fp[1][((size_t)kp * (nx2 + 1) + jp) * nx1 + ip] : 0.0;
          const double Py_u = multi_d ? b.fp[1][((size_t)kp * (nx2 + 1) + jp + 1) * nx1 + ip] : 0.0;
          const double Pz_l = three_d ? b.fp[2][((size_t)kp * nx2 + jp) * nx1 + ip] : 0.0;
          const double Pz_u = three_d ? b.fp[2][((size_t)(kp + 1) * nx2 + jp) * nx1 + ip] : 0.0;
          DdmcArgs dia{rng, t_start, dt, ff, aa, ss, vv, multi_d, three_d, xl, yl, zl, xu, yu, zu,
                       Px_l, Py_l, Pz_l, Px_u, Py_u, Pz_u, t, x, y, z, vx, vy, vz, ip, jp, kp,
                       is_absorbed, is_scattered};
          ddmc_albedo(dia, is_rejected);
          if (!is_rejected) ddmc_step(dia);
        } else {
          imc_step(rng, t_start, dt, ff, aa, ss, vv, vx, vy, vz, dx_push, multi_d, three_d, xl, yl, zl,
                   xu, yu, zu, t, x, y, z, is_absorbed, is_scattered);
        }
        Xtoijk(c, b, x, y, z, ip, jp, kp);
        int kind = (is_ddmc_step ? JB_EV_DDMC : 0) | (is_rejected ? JB_EV_ALBEDO_REJECT : 0);
        if (!on_block(c, ip, jp, kp)) {
          if (use_ddmc) {
            const bool vmask = !(is_ddmc_step && multi_d && !is_rejected);
            vx *= vmask;
            vy *= vmask;
            vz *= vmask;
          }
          if (c->log_events) {
            const double v[3] = {vx, vy, vz};
            p.rn = rng.n;
            log_event(c, p, draw0, kind | JB_EV_LEFT_BLOCK | (!(t < t_start + dt) ? JB_EV_CENSUS : 0), b.info.gid, v);
          }
          break;
        }
        if (is_absorbed) {
          double &dejbn = b.edelta[cidx(c, kp, jp, ip)];
#if defined(_OPENMP)
#pragma omp atomic
#endif
          dejbn += ww;
          p.marked = true;
          r_abs++;
          r_eabs += ww;
          if (c->log_events) {
            const double v[3] = {vx, vy, vz};
            log_event(c, p, draw0, kind | JB_EV_ABSORB, b.info.gid, v);
          }
          break;
        }
        if (is_scattered) {
          scatter_iso(rng, vv, vx, vy, vz);
          r_scat++;
          kind |= JB_EV_SCATTER;
        }
        if (c->log_events) {
          const double v[3] = {vx, vy, vz};
          log_event(c, p, draw0, kind | (!(t < t_start + dt) ? JB_EV_CENSUS : 0), b.info.gid, v);
        }
      }
      p.rn = rng.n;
    }
#if defined(_OPENMP)
#pragma omp critical(jbo_transport_sums)
#endif
    {
      sums[bi].seg += r_seg; sums[bi].ddmc += r_ddmc; sums[bi].abs += r_abs; sums[bi].scat += r_scat;
      sums[bi].eabs += r_eabs;
    }
  }
  for (size_t bi = 0; bi < nblk; ++bi) {
    c->stats.n_segments += sums[bi].seg;
    c->stats.n_ddmc_segments += sums[bi].ddmc;
    c->stats.n_absorbed += sums[bi].abs;
    c->stats.n_scatters += sums[bi].scat;
    c->stats.e_absorbed += sums[bi].eabs;
  }
#if defined(_OPENMP)
#pragma omp parallel for schedule(dynamic, 1) if (!serial)
#endif
  for (size_t bi = 0; bi < nblk; ++bi) remove_marked(c->blocks[bi]);
  return JB_OK;
}

// MeshResetCommunication + MeshSend + MeshReceive (jaybenne.cpp:26-61): every particle
// that sits off its block gets the swarm boundary conditions applied and is handed to
// the block that now contains it (same rank: appended to that pool; other rank: outbox).
static jb_status swarm_exchange_local(Ctx *c) {
  // the scan of every pool for off-block particles runs over the blocks in parallel (timing builds); the
  // leavers are then delivered in block order, which is the order the serial loop always produced
  const size_t nblk = c->blocks.size();
  std::vector<std::vector<std::pair<int, Particle>>> moved_by(nblk);
  std::vector<long long> esc_n(nblk, 0), cross_n(nblk, 0), lost(nblk, 0);
  std::vector<double> esc_e(nblk, 0.0);
  const bool serial = c->cfg.rng_mode == JB_RNG_XORSHIFT_POOL;
  (void)serial;
#if defined(_OPENMP)
#pragma omp parallel for schedule(dynamic, 1) if (!serial)
#endif
  for (size_t bi = 0; bi < nblk; ++bi) {
    Block &b = c->blocks[bi];
    for (Particle &p : b.pool) {
      if (on_block(c, p.i, p.j, p.k)) continue;
      Particle q = p;
      p.marked = true;
      if (!apply_swarm_bc(c, q)) {
        esc_n[bi]++;
        esc_e[bi] += q.w;
        continue;
      }
      const double pos[3] = {q.x, q.y, q.z};
      const int gid = find_block(c, pos);
      if (gid < 0) {
        lost[bi]++;
        continue;
      }
      cross_n[bi]++;
      moved_by[bi].push_back({gid, q});
    }
    remove_marked(b);
  }
  std::vector<std::pair<int, Particle>> moved;
  for (size_t bi = 0; bi < nblk; ++bi) {
    if (lost[bi]) return fail(c, JB_ERR_STATE, "particle left the domain without a boundary condition");
    c->stats.n_escaped += esc_n[bi];
    c->stats.e_escaped += esc_e[bi];
    c->stats.n_block_crossings += cross_n[bi];
    moved.insert(moved.end(), moved_by[bi].begin(), moved_by[bi].end());
  }
  for (auto &m : moved) {
    const int gid = m.first;
    Particle &q = m.second;
    q.marked = false;
    const int loc = c->gid_to_local[gid];
    if (loc >= 0) {
      Block &nb = c->blocks[loc];
      Xtoijk(c, nb, q.x, q.y, q.z, q.i, q.j, q.k);
      nb.pool.push_back(q);
    } else {
      ExRecord r;
      memset(&r, 0, sizeof(r));
      r.x = q.x; r.y = q.y; r.z = q.z; r.t = q.t; r.w = q.w; r.e = q.e;
      r.vx = q.vx; r.vy = q.vy; r.vz = q.vz;
      r.rng[0] = q.rid[0]; r.rng[1] = q.rid[1]; r.rng[2] = q.rid[2]; r.rng[3] = q.rn;
      r.dest_gid = gid;
      c->outbox[c->all_blocks[gid].rank].push_back(r);
      c->stats.n_sent++;
    }
  }
  return JB_OK;
}

// SampleFace2D / SampleFace3D, sample_ddmc_bface.cpp:24-78
static void sample_face_2d(const int i_l, const double dx, const double P_l, const double P_u, RngGen &rng,
                           int &i, double &x) {
  const int i_u = i_l + 1;
  const double xi = (P_l + P_u) * rng.drand();
  if (xi < P_l) {
    x -= dx * rng.drand();
    i = i_l;
  } else {
    x += dx * rng.drand();
    i = i_u;
  }
}
static void sample_face_3d(const int i1_l, const int i2_l, const double dx1, const double dx2, const double P_ll,
                           const double P_lu, const double P_ul, const double P_uu, RngGen &rng, int &i1,
                           int &i2, double &x1, double &x2) {
  const int i1_u = i1_l + 1, i2_u = i2_l + 1;
  const double xi = (P_ll + P_lu + P_ul + P_uu) * rng.drand();
  if (xi < P_ll) {
    x1 -= dx1 * rng.drand(); i1 = i1_l;
    x2 -= dx2 * rng.drand(); i2 = i2_l;
  } else if (xi < P_ll + P_lu) {
    x1 += dx1 * rng.drand(); i1 = i1_u;
    x2 -= dx2 * rng.drand(); i2 = i2_l;
  } else if (xi < P_ll + P_lu + P_ul) {
    x1 -= dx1 * rng.drand(); i1 = i1_l;
    x2 += dx2 * rng.drand(); i2 = i2_u;
  } else {
    x1 += dx1 * rng.drand(); i1 = i1_u;
    x2 += dx2 * rng.drand(); i2 = i2_u;
  }
}

// SampleDDMCBlockFace, sample_ddmc_bface.cpp:81-427. The reference reads the face probabilities through
// ghost-inclusive indices; on a 2:1 mesh the two fine cells under a coarse cell never straddle a block
// edge, so a read one index outside the block cannot happen: it is counted and fails the task (the
// reference would read an exchanged ghost face there) instead of being clamped silently.
static jb_status sample_ddmc_block_face(Ctx *c) {
  if (!(c->mesh.ndim > 1)) return JB_OK;
  const double vv = c->cl.c_code;
  const bool three_d = c->mesh.ndim > 2;
  constexpr double eps = EPS10;
  const int n1 = c->mesh.nxb[0], n2 = c->mesh.nxb[1], n3 = c->mesh.nxb[2];
  long long ghost_reads = 0;
  auto inblock = [&](int v, int lo, int hi) {
    if (v < lo || v > hi) ++ghost_reads;
    return std::min(std::max(v, lo), hi);
  };
  for (Block &b : c->blocks) {
    auto F1 = [&](int k, int j, int i) {
      return b.fp[0][((size_t)inblock(k, 0, n3 - 1) * n2 + inblock(j, 0, n2 - 1)) * (n1 + 1) + inblock(i, 0, n1)];
    };
    auto F2 = [&](int k, int j, int i) {
      return b.fp[1][((size_t)inblock(k, 0, n3 - 1) * (n2 + 1) + inblock(j, 0, n2)) * n1 + inblock(i, 0, n1 - 1)];
    };
    auto F3 = [&](int k, int j, int i) {
      return b.fp[2][((size_t)inblock(k, 0, n3) * n2 + inblock(j, 0, n2 - 1)) * n1 + inblock(i, 0, n1 - 1)];
    };
    const double bx0 = b.info.xmin[0], bx1 = b.info.xmax[0];
    const double by0 = b.info.xmin[1], by1 = b.info.xmax[1];
    const double bz0 = b.info.xmin[2], bz1 = b.info.xmax[2];
    for (Particle &p : b.pool) {
      double &vx = p.vx, &vy = p.vy, &vz = p.vz;
      if (!(vx * vx + vy * vy + vz * vz < eps * vv * vv)) continue;
      double &x = p.x, &y = p.y, &z = p.z;
      int &ip = p.i, &jp = p.j, &kp = p.k;
      Xtoijk(c, b, x, y, z, ip, jp, kp);
      const double dx_i = b.dx[0], dy_j = b.dx[1], dz_k = b.dx[2];
      const double x_i = Xc(c, b, 0, ip) - 0.5 * dx_i;
      const double y_j = Xc(c, b, 1, jp) - 0.5 * dy_j;
      const double z_k = Xc(c, b, 2, kp) - 0.5 * dz_k;
      RngGen rng = particle_rng(c, p);
      const bool at_block_x_min = fuzzy_eq(x, bx0 + 2.0 * eps_ddmc_offset * dx_i, dx_i, eps);
      const bool at_block_x_max = fuzzy_eq(x, bx1 - 2.0 * eps_ddmc_offset * dx_i, dx_i, eps);
      const bool at_block_y_min = fuzzy_eq(y, by0 + 2.0 * eps_ddmc_offset * dy_j, dy_j, eps);
      const bool at_block_y_max = fuzzy_eq(y, by1 - 2.0 * eps_ddmc_offset * dy_j, dy_j, eps);
      const bool at_block_z_min = three_d && fuzzy_eq(z, bz0 + 2.0 * eps_ddmc_offset * dz_k, dz_k, eps);
      const bool at_block_z_max = three_d && fuzzy_eq(z, bz1 - 2.0 * eps_ddmc_offset * dz_k, dz_k, eps);
      if (at_block_x_min || at_block_x_max) {
        const double dir_sgn = at_block_x_min ? 1.0 : -1.0;
        face_iso_dir(dir_sgn * vv, rng, vx, vy, vz);
        const int ip_b = at_block_x_min ? ip : ip + 1;
        const bool at_y_edge_u = fuzzy_eq(y, y_j, dy_j, eps);
        const bool at_y_edge_l = fuzzy_eq(y, y_j + dy_j, dy_j, eps);
        if (!three_d) {
          if (at_y_edge_u || at_y_edge_l) {
            const int jp_u = at_y_edge_u ? jp : jp + 1;
            const int jp_l = at_y_edge_u ? jp - 1 : jp;
            ++c->stats.n_bface_resampled;
            sample_face_2d(jp_l, dy_j, F1(kp, jp_l, ip_b), F1(kp, jp_u, ip_b), rng, jp, y);
          }
        } else {
          const bool at_z_edge_u = fuzzy_eq(z, z_k, dz_k, eps);
          const bool at_z_edge_l = fuzzy_eq(z, z_k + dz_k, dz_k, eps);
          if ((at_y_edge_u || at_y_edge_l) && (at_z_edge_u || at_z_edge_l)) {
            const int jp_u = at_y_edge_u ? jp : jp + 1, jp_l = at_y_edge_u ? jp - 1 : jp;
            const int kp_u = at_z_edge_u ? kp : kp + 1, kp_l = at_z_edge_u ? kp - 1 : kp;
            const double Px_uu = F1(kp_u, jp_u, ip_b), Px_ul = F1(kp_u, jp_l, ip_b);
            const double Px_lu = F1(kp_l, jp_u, ip_b), Px_ll = F1(kp_l, jp_l, ip_b);
            ++c->stats.n_bface_resampled;
            sample_face_3d(jp_l, kp_l, dy_j, dz_k, Px_ll, Px_lu, Px_ul, Px_uu, rng, jp, kp, y, z);
          }
        }
      } else if (at_block_y_min || at_block_y_max) {
        const double dir_sgn = at_block_y_min ? 1.0 : -1.0;
        face_iso_dir(dir_sgn * vv, rng, vy, vz, vx);
        const int jp_b = at_block_y_min ? jp : jp + 1;
        const bool at_x_edge_u = fuzzy_eq(x, x_i, dx_i, eps);
        const bool at_x_edge_l = fuzzy_eq(x, x_i + dx_i, dx_i, eps);
        if (!three_d) {
          if (at_x_edge_u || at_x_edge_l) {
            const int ip_u = at_x_edge_u ? ip : ip + 1;
            const int ip_l = at_x_edge_u ? ip - 1 : ip;
            ++c->stats.n_bface_resampled;
            sample_face_2d(ip_l, dx_i, F2(kp, jp_b, ip_l), F2(kp, jp_b, ip_u), rng, ip, x);
          }
        } else {
          const bool at_z_edge_u = fuzzy_eq(z, z_k, dz_k, eps);
          const bool at_z_edge_l = fuzzy_eq(z, z_k + dz_k, dz_k, eps);
          if ((at_z_edge_u || at_z_edge_l) && (at_x_edge_u || at_x_edge_l)) {
            const int kp_u = at_z_edge_u ? kp : kp + 1, kp_l = at_z_edge_u ? kp - 1 : kp;
            const int ip_u = at_x_edge_u ? ip : ip + 1, ip_l = at_x_edge_u ? ip - 1 : ip;
            const double Py_uu = F2(kp_u, jp_b, ip_u), Py_ul = F2(kp_u, jp_b, ip_l);
            const double Py_lu = F2(kp_l, jp_b, ip_u), Py_ll = F2(kp_l, jp_b, ip_l);
            ++c->stats.n_bface_resampled;
            sample_face_3d(ip_l, kp_l, dx_i, dz_k, Py_ll, Py_lu, Py_ul, Py_uu, rng, ip, kp, x, z);
          }
        }
      } else if (at_block_z_min || at_block_z_max) {
        const double dir_sgn = at_block_z_min ? 1.0 : -1.0;
        face_iso_dir(dir_sgn * vv, rng, vz, vx, vy);
        const int kp_b = at_block_z_min ? kp : kp + 1;
        const bool at_x_edge_u = fuzzy_eq(x, x_i, dx_i, eps);
        const bool at_x_edge_l = fuzzy_eq(x, x_i + dx_i, dx_i, eps);
        const bool at_y_edge_u = fuzzy_eq(y, y_j, dy_j, eps);
        const bool at_y_edge_l = fuzzy_eq(y, y_j + dy_j, dy_j, eps);
        if ((at_x_edge_u || at_x_edge_l) && (at_y_edge_u || at_y_edge_l)) {
          const int ip_u = at_x_edge_u ? ip : ip + 1, ip_l = at_x_edge_u ? ip - 1 : ip;
          const int jp_u = at_y_edge_u ? jp : jp + 1, jp_l = at_y_edge_u ? jp - 1 : jp;
          const double Pz_uu = F3(kp_b, jp_u, ip_u), Pz_ul = F3(kp_b, jp_u, ip_l);
          const double Pz_lu = F3(kp_b, jp_l, ip_u), Pz_ll = F3(kp_b, jp_l, ip_l);
          ++c->stats.n_bface_resampled;
          sample_face_3d(ip_l, jp_l, dx_i, dy_j, Pz_ll, Pz_lu, Pz_ul, Pz_uu, rng, ip, jp, x, y);
        }
      }
      p.rn = rng.n;
    }
  }
  if (ghost_reads)
    return fail(c, JB_ERR_STATE, "SampleDDMCBlockFace indexed a face probability outside the block (not a 2:1 mesh?)");
  return JB_OK;
}

// CheckCompletion, transport.cpp:187-216 (local count; the global sum is the host's job)
static int64_t check_completion(Ctx *c, double t_end) {
  int64_t num_unfinished = 0;
  for (Block &b : c->blocks)
    for (Particle &p : b.pool)
      if (p.t < t_end) num_unfinished++;
  return num_unfinished;
}

// EvaluateRadiationEnergy, jaybenne.cpp:514-564
static void evaluate_radiation_energy(Ctx *c) {
#if defined(_OPENMP)
#pragma omp parallel for schedule(dynamic, 1)
#endif
  for (size_t bi = 0; bi < c->blocks.size(); ++bi) {
    Block &b = c->blocks[bi];
    std::fill(b.tally.begin(), b.tally.end(), 0.0);
    const double dv = b.dx[0] * b.dx[1] * b.dx[2];
    for (Particle &p : b.pool) b.tally[cidx(c, p.k, p.j, p.i)] += p.w / dv;
  }
}

// UpdateFluid, jaybenne.cpp:583-615
static void update_fluid(Ctx *c) {
  if (!c->cfg.do_feedback) return;
#if defined(_OPENMP)
#pragma omp parallel for schedule(static)
#endif
  for (size_t bi = 0; bi < c->blocks.size(); ++bi) {
    Block &b = c->blocks[bi];
    const double dv = b.dx[0] * b.dx[1] * b.dx[2];
    for (size_t ci = 0; ci < b.u.size(); ++ci) {
      const double delta = b.edelta[ci] / dv;
      b.u[ci] += delta;
    }
  }
}

static void census_sums(Ctx *c, int64_t &n, double &e) {
#if defined(_OPENMP)
  // timing builds: per-block partial sums in parallel, added in block order
  const size_t nblk = c->blocks.size();
  std::vector<int64_t> bn(nblk, 0);
  std::vector<double> be(nblk, 0.0);
#pragma omp parallel for schedule(dynamic, 1)
  for (size_t bi = 0; bi < nblk; ++bi) {
    int64_t nn = 0;
    double ee = 0.0;
    for (Particle &p : c->blocks[bi].pool) {
      nn++;
      ee += p.w;
    }
    bn[bi] = nn;
    be[bi] = ee;
  }
  n = 0;
  e = 0.0;
  for (size_t bi = 0; bi < nblk; ++bi) {
    n += bn[bi];
    e += be[bi];
  }
#else
  n = 0;
  e = 0.0;
  for (Block &b : c->blocks)
    for (Particle &p : b.pool) {
      n++;
      e += p.w;
    }
#endif
}

}  // namespace jbo

// =======================================================================================
// C ABI (jbo_* mirrors jb_* of include/jaybenne_b200.h)
// =======================================================================================
using jbo::Ctx;
struct jb_ctx : public jbo::Ctx {};

extern "C" {

const char *jbo_flavour(void) { return JBO_FLAVOUR; }

// host threads of the timing builds (bench.py's CPU arm sets and reports them; parity builds have one)
int32_t jbo_max_threads(void) {
#if defined(_OPENMP)
  return (int32_t)omp_get_max_threads();
#else
  return 1;
#endif
}
void jbo_set_num_threads(int32_t n) {
#if defined(_OPENMP)
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

void jbo_config_default(jb_config *cfg) {
  memset(cfg, 0, sizeof(*cfg));
  cfg->num_particles = 0;
  cfg->dt = std::numeric_limits<double>::max();
  cfg->min_swarm_occupancy = 0.0;
  cfg->numin = std::numeric_limits<double>::min();
  cfg->numax = std::numeric_limits<double>::max();
  cfg->unique_rank_seeds = 1;
  cfg->seed = 123;
  cfg->max_transport_iterations = 10000;
  cfg->use_ddmc = 0;
  cfg->tau_ddmc = 5.0;
  cfg->source_strategy = JB_STRATEGY_UNIFORM;
  cfg->do_emission = 1;
  cfg->do_feedback = 1;
  cfg->opacity_model = JB_OPAC_GRAY;
  cfg->opacity_kappa = 0.0;
  cfg->scattering_model = JB_SCAT_GRAY;
  cfg->scattering_sigma = 0.0;
  cfg->apm = 1.0;
  cfg->time_scale = cfg->mass_scale = cfg->length_scale = cfg->temperature_scale = 1.0;
  cfg->eos_gm1 = 1.66666666667 - 1.;
  cfg->eos_cv = 1. / (1.66666666667 - 1.);
  cfg->rng_mode = JB_RNG_PHILOX;
  cfg->gray_mode = JB_GRAY_KIRCHHOFF;
  cfg->source_nblocks = 0;
  cfg->sort_interval = 0;
}

jb_status jbo_create(const jb_config *cfg, jb_ctx **out) {
  if (!cfg || !out) return JB_ERR_INVALID;
  if (!(cfg->min_swarm_occupancy >= 0 && cfg->min_swarm_occupancy < 1.0)) return JB_ERR_INVALID;
  jb_ctx *c = new jb_ctx();
  c->cfg = *cfg;
  c->cl.init(*cfg);
  memset(&c->stats, 0, sizeof(c->stats));
  memset(&c->mesh, 0, sizeof(c->mesh));
  {
    int states = 1;
#if defined(_OPENMP)
    if (cfg->rng_mode == JB_RNG_XORSHIFT_THREADS) states = omp_get_max_threads();
#endif
    jbo::seed_pool(c, states);
  }
  *out = c;
  return JB_OK;
}
void jbo_destroy(jb_ctx *c) { delete c; }
const char *jbo_last_error(const jb_ctx *c) { return c ? c->err.c_str() : "null context"; }

jb_status jbo_get_constants(const jb_ctx *c, double *sol, double *sb) {
  if (sol) *sol = c->cl.c_code;
  if (sb) *sb = c->cl.sb_code;
  return JB_OK;
}

jb_status jbo_set_mesh(jb_ctx *c, const jb_mesh *m, int32_t my_rank, int32_t n_ranks) {
  if (!m || m->nblocks <= 0 || m->ndim < 1 || m->ndim > 3) return jbo::fail(c, JB_ERR_INVALID, "bad mesh");
  c->mesh = *m;
  c->all_blocks.assign(m->blocks, m->blocks + m->nblocks);
  c->mesh.blocks = c->all_blocks.data();
  c->my_rank = my_rank;
  c->n_ranks = n_ranks;
  for (int d = 0; d < 3; ++d) {
    c->nlat[d] = d < m->ndim ? (m->nrb[d] << m->max_level) : 1;
    c->ng[d] = (d < m->ndim) ? m->nghost : 0;
  }
  c->lattice.assign((size_t)c->nlat[0] * c->nlat[1] * c->nlat[2], -1);
  c->gid_to_local.assign(m->nblocks, -1);
  c->blocks.clear();
  const int nc = m->nxb[0] * m->nxb[1] * m->nxb[2];
  for (int g = 0; g < m->nblocks; ++g) {
    const jb_block &b = c->all_blocks[g];
    if (b.gid != g) return jbo::fail(c, JB_ERR_INVALID, "blocks must be listed in gid order");
    const int s = 1 << (m->max_level - b.level);
    int lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
    for (int d = 0; d < m->ndim; ++d) {
      lo[d] = (int)b.lx[d] * s;
      hi[d] = lo[d] + s;
    }
    for (int z = lo[2]; z < hi[2]; ++z)
      for (int y = lo[1]; y < hi[1]; ++y)
        for (int x = lo[0]; x < hi[0]; ++x) c->lattice[((size_t)z * c->nlat[1] + y) * c->nlat[0] + x] = g;
    if (b.rank == my_rank) {
      jbo::Block B;
      B.info = b;
      for (int d = 0; d < 3; ++d) {
        B.dx[d] = (b.xmax[d] - b.xmin[d]) / m->nxb[d];
        B.xmin0[d] = b.xmin[d] - c->ng[d] * B.dx[d];
      }
      B.rho.assign(nc, 0.0); B.sie.assign(nc, 0.0); B.u.assign(nc, 0.0); B.tally.assign(nc, 0.0);
      B.fleck.assign(nc, 0.0); B.src_ew.assign(nc, 0.0); B.src_num.assign(nc, 0.0); B.edelta.assign(nc, 0.0);
      c->gid_to_local[g] = (int)c->blocks.size();
      c->blocks.push_back(B);
    }
  }
  for (int v : c->lattice)
    if (v < 0) return jbo::fail(c, JB_ERR_INVALID, "block table does not tile the domain");
  c->outbox.assign(n_ranks, {});
  c->mesh_set = true;
  return JB_OK;
}

jb_status jbo_set_capacity(jb_ctx *, int64_t) { return JB_OK; }

static std::vector<double> *field_vec(jb_ctx *c, jbo::Block &b, int field) {
  switch (field) {
    case JB_FIELD_DENSITY: return &b.rho;
    case JB_FIELD_SIE: return &b.sie;
    case JB_FIELD_INTERNAL_ENERGY: return &b.u;
    case JB_FIELD_ENERGY_TALLY: return &b.tally;
    case JB_FIELD_FLECK_FACTOR: return &b.fleck;
    case JB_FIELD_SOURCE_EW_PER_CELL: return &b.src_ew;
    case JB_FIELD_SOURCE_NUM_PER_CELL: return &b.src_num;
    case JB_FIELD_ENERGY_DELTA: return &b.edelta;
    case JB_FIELD_DDMC_FACE_PROB_X1: return &b.fp[0];
    case JB_FIELD_DDMC_FACE_PROB_X2: return &b.fp[1];
    case JB_FIELD_DDMC_FACE_PROB_X3: return &b.fp[2];
  }
  (void)c;
  return nullptr;
}

int64_t jbo_field_size(const jb_ctx *c, int32_t field) {
  const int64_t n1 = c->mesh.nxb[0], n2 = c->mesh.nxb[1], n3 = c->mesh.nxb[2];
  if (field == JB_FIELD_DDMC_FACE_PROB_X1) return (n1 + 1) * n2 * n3;
  if (field == JB_FIELD_DDMC_FACE_PROB_X2) return n1 * (n2 + 1) * n3;
  if (field == JB_FIELD_DDMC_FACE_PROB_X3) return n1 * n2 * (n3 + 1);
  return n1 * n2 * n3;
}

jb_status jbo_set_field(jb_ctx *c, int32_t gid, int32_t field, const double *src) {
  if (!c->mesh_set) return jbo::fail(c, JB_ERR_STATE, "mesh not set");
  if (gid < 0 || gid >= c->mesh.nblocks || c->gid_to_local[gid] < 0) return jbo::fail(c, JB_ERR_INVALID, "block not local");
  auto *v = field_vec(c, c->blocks[c->gid_to_local[gid]], field);
  if (!v) return jbo::fail(c, JB_ERR_INVALID, "bad field");
  v->assign(src, src + jbo_field_size(c, field));
  return JB_OK;
}
jb_status jbo_get_field(jb_ctx *c, int32_t gid, int32_t field, double *dst) {
  if (!c->mesh_set) return jbo::fail(c, JB_ERR_STATE, "mesh not set");
  if (gid < 0 || gid >= c->mesh.nblocks || c->gid_to_local[gid] < 0) return jbo::fail(c, JB_ERR_INVALID, "block not local");
  auto *v = field_vec(c, c->blocks[c->gid_to_local[gid]], field);
  if (!v) return jbo::fail(c, JB_ERR_INVALID, "bad field");
  const int64_t n = jbo_field_size(c, field);
  for (int64_t i = 0; i < n; ++i) dst[i] = i < (int64_t)v->size() ? (*v)[i] : 0.0;
  return JB_OK;
}

jb_status jbo_update_derived_transport_fields(jb_ctx *c, double dt) {
  return jbo::update_derived_transport_fields(c, dt);
}

jb_status jbo_source_photons(jb_ctx *c, int32_t source_type, double t_start, double dt) {
  std::vector<int> all;
  for (size_t i = 0; i < c->blocks.size(); ++i) all.push_back((int)i);
  const int nb = c->cfg.source_nblocks > 0 ? c->cfg.source_nblocks : (int)c->blocks.size();
  return jbo::source_photons(c, all, source_type, t_start, dt, nb);
}

jb_status jbo_evaluate_radiation_energy(jb_ctx *c) {
  jbo::evaluate_radiation_energy(c);
  return JB_OK;
}

// InitializeRadiation per block (jaybenne.cpp:570-578): MeshBlockData packs => nblocks = 1.
// All blocks of one initialisation share one source-call id (their gid tells them apart).
jb_status jbo_initialize_radiation(jb_ctx *c, int32_t is_thermal) {
  if (!c->mesh_set) return jbo::fail(c, JB_ERR_STATE, "mesh not set");
  if (is_thermal) {
    const uint32_t call = c->source_call;
    for (size_t i = 0; i < c->blocks.size(); ++i) {
      c->source_call = call;
      jb_status s = jbo::source_photons(c, {(int)i}, JB_SOURCE_THERMAL, 0.0, 0.0, 1);
      if (s != JB_OK) return s;
    }
    c->source_call = call + 1;
  }
  jbo::evaluate_radiation_energy(c);
  return JB_OK;
}

jb_status jbo_transport_photons(jb_ctx *c, double t_start, double dt) { return jbo::transport_photons(c, t_start, dt); }
jb_status jbo_swarm_exchange(jb_ctx *c) { return jbo::swarm_exchange_local(c); }
jb_status jbo_sample_ddmc_block_face(jb_ctx *c) { return jbo::sample_ddmc_block_face(c); }
jb_status jbo_check_completion(jb_ctx *c, double t_end, int64_t *n) {
  *n = jbo::check_completion(c, t_end);
  return JB_OK;
}
jb_status jbo_update_fluid(jb_ctx *c) {
  jbo::update_fluid(c);
  return JB_OK;
}
// mcblock::UpdateDerived (mcblock.cpp:208-232): sie = u / rho
jb_status jbo_update_derived_material(jb_ctx *c) {
  for (auto &b : c->blocks)
    for (size_t i = 0; i < b.u.size(); ++i) b.sie[i] = b.u[i] / b.rho[i];
  return JB_OK;
}
// mcblock::PostInitialization (mcblock.cpp:237-262): u = rho * sie
jb_status jbo_post_initialization(jb_ctx *c) {
  for (auto &b : c->blocks)
    for (size_t i = 0; i < b.u.size(); ++i) b.u[i] = b.rho[i] * b.sie[i];
  return JB_OK;
}

// Stats bracket used by jbo_radiation_step and by hosts that drive the tasks themselves.
jb_status jbo_stats_begin(jb_ctx *c) {
  memset(&c->stats, 0, sizeof(c->stats));
  jbo::census_sums(c, c->stats.n_start, c->stats.e_start);
  return JB_OK;
}
jb_status jbo_get_stats(jb_ctx *c, jb_cycle_stats *s) {
  jbo::census_sums(c, c->stats.n_end, c->stats.e_end);
  *s = c->stats;
  return JB_OK;
}

// RadiationStep(...).Execute(), jaybenne.cpp:68-151 — single-rank form. With more than one
// rank the host drives the same tasks and moves the outboxes (tests/test_multirank.py).
jb_status jbo_radiation_step(jb_ctx *c, double t_start, double dt, jb_cycle_stats *stats) {
  if (!c->mesh_set) return jbo::fail(c, JB_ERR_STATE, "mesh not set");
  jbo_stats_begin(c);
  jb_status s = jbo::update_derived_transport_fields(c, dt);
  if (s != JB_OK) return s;
  s = jbo_source_photons(c, JB_SOURCE_EMISSION, t_start, dt);
  if (s != JB_OK) return s;
  bool done = false;
  for (int it = 1; it <= c->cfg.max_transport_iterations; ++it) {
    c->stats.n_iterations = it;
    s = jbo::transport_photons(c, t_start, dt);
    if (s != JB_OK) return s;
    s = jbo::swarm_exchange_local(c);
    if (s != JB_OK) return s;
    if (c->n_ranks > 1) return jbo::fail(c, JB_ERR_UNSUPPORTED, "multi-rank: drive the tasks from the host");
    if (c->cfg.use_ddmc) {
      s = jbo::sample_ddmc_block_face(c);
      if (s != JB_OK) return s;
    }
    if (jbo::check_completion(c, t_start + dt) == 0) {
      done = true;
      break;
    }
  }
  if (!done) return jbo::fail(c, JB_ERR_ITERATIONS, "max_transport_iterations exhausted");
  jbo::evaluate_radiation_energy(c);
  jbo::update_fluid(c);
  if (stats) jbo_get_stats(c, stats);
  return JB_OK;
}

jb_status jbo_num_particles(jb_ctx *c, int64_t *n) {
  int64_t t = 0;
  for (auto &b : c->blocks) t += (int64_t)b.pool.size();
  *n = t;
  return JB_OK;
}

jb_status jbo_get_particles(jb_ctx *c, const jb_particles *d, int64_t capacity, int64_t *n) {
  int64_t m = 0;
  for (auto &b : c->blocks)
    for (auto &p : b.pool) {
      if (m >= capacity) break;
      if (d->x) d->x[m] = p.x;
      if (d->y) d->y[m] = p.y;
      if (d->z) d->z[m] = p.z;
      if (d->time) d->time[m] = p.t;
      if (d->weight) d->weight[m] = p.w;
      if (d->energy) d->energy[m] = p.e;
      if (d->v0) d->v0[m] = p.vx;
      if (d->v1) d->v1[m] = p.vy;
      if (d->v2) d->v2[m] = p.vz;
      if (d->i) d->i[m] = p.i + c->ng[0];
      if (d->j) d->j[m] = p.j + c->ng[1];
      if (d->k) d->k[m] = p.k + c->ng[2];
      if (d->block) d->block[m] = b.info.gid;
      if (d->rng) {
        d->rng[4 * m + 0] = p.rid[0];
        d->rng[4 * m + 1] = p.rid[1];
        d->rng[4 * m + 2] = p.rid[2];
        d->rng[4 * m + 3] = p.rn;
      }
      ++m;
    }
  *n = m;
  return JB_OK;
}

jb_status jbo_reset_streams(jb_ctx *c) {
  c->source_call = 0;
  return JB_OK;
}
jb_status jbo_set_particles(jb_ctx *c, const jb_particles *s, int64_t n) {
  for (auto &b : c->blocks) b.pool.clear();
  for (int64_t m = 0; m < n; ++m) {
    jbo::Particle p;
    memset(&p, 0, sizeof(p));
    p.x = s->x[m]; p.y = s->y ? s->y[m] : 0.0; p.z = s->z ? s->z[m] : 0.0;
    p.t = s->time[m]; p.w = s->weight[m]; p.e = s->energy[m];
    p.vx = s->v0[m]; p.vy = s->v1[m]; p.vz = s->v2[m];
    if (s->rng) {
      p.rid[0] = s->rng[4 * m]; p.rid[1] = s->rng[4 * m + 1]; p.rid[2] = s->rng[4 * m + 2]; p.rn = s->rng[4 * m + 3];
    }
    const int gid = s->block[m];
    if (gid < 0 || gid >= c->mesh.nblocks || c->gid_to_local[gid] < 0) return jbo::fail(c, JB_ERR_INVALID, "block not local");
    jbo::Block &b = c->blocks[c->gid_to_local[gid]];
    jbo::Xtoijk(c, b, p.x, p.y, p.z, p.i, p.j, p.k);
    b.pool.push_back(p);
  }
  return JB_OK;
}

jb_status jbo_event_log_enable(jb_ctx *c, int64_t capacity) {
  c->log_events = capacity > 0;
  c->event_cap = capacity;
  c->events.clear();
  return JB_OK;
}
jb_status jbo_event_log_read(jb_ctx *c, jb_event *dst, int64_t capacity, int64_t *n) {
  const int64_t m = std::min<int64_t>(capacity, (int64_t)c->events.size());
  if (m > 0) memcpy(dst, c->events.data(), m * sizeof(jb_event));
  *n = (int64_t)c->events.size();
  return JB_OK;
}

int64_t jbo_exchange_record_bytes(void) { return (int64_t)sizeof(jbo::ExRecord); }
jb_status jbo_exchange_count(jb_ctx *c, int32_t peer, int64_t *n) {
  *n = (int64_t)c->outbox[peer].size();
  return JB_OK;
}
jb_status jbo_exchange_take(jb_ctx *c, int32_t peer, void *dst, int64_t capacity, int64_t *n) {
  auto &o = c->outbox[peer];
  if ((int64_t)o.size() > capacity) return jbo::fail(c, JB_ERR_CAPACITY, "exchange buffer too small");
  if (!o.empty()) memcpy(dst, o.data(), o.size() * sizeof(jbo::ExRecord));
  *n = (int64_t)o.size();
  o.clear();
  return JB_OK;
}
jb_status jbo_exchange_give(jb_ctx *c, const void *src, int64_t n) {
  const jbo::ExRecord *r = (const jbo::ExRecord *)src;
  for (int64_t m = 0; m < n; ++m) {
    const int gid = r[m].dest_gid;
    if (gid < 0 || gid >= c->mesh.nblocks || c->gid_to_local[gid] < 0) return jbo::fail(c, JB_ERR_COMM, "record for a block that is not local");
    jbo::Block &b = c->blocks[c->gid_to_local[gid]];
    jbo::Particle p;
    memset(&p, 0, sizeof(p));
    p.x = r[m].x; p.y = r[m].y; p.z = r[m].z; p.t = r[m].t; p.w = r[m].w; p.e = r[m].e;
    p.vx = r[m].vx; p.vy = r[m].vy; p.vz = r[m].vz;
    p.rid[0] = r[m].rng[0]; p.rid[1] = r[m].rng[1]; p.rid[2] = r[m].rng[2]; p.rn = r[m].rng[3];
    jbo::Xtoijk(c, b, p.x, p.y, p.z, p.i, p.j, p.k);
    b.pool.push_back(p);
    c->stats.n_received++;
  }
  return JB_OK;
}

// ---- unit-level probes used to pin the restatement against the reference headers -------
double jbo_probe_absorption(jb_ctx *c, double rho, double temp, double nu) { return c->cl.absorption(rho, temp, nu); }
double jbo_probe_scattering(jb_ctx *c, double rho, double temp, double nu) { return c->cl.scattering_coeff(rho, temp, nu); }
double jbo_probe_emissivity(jb_ctx *c, double rho, double temp) { return c->cl.emissivity(rho, temp); }
double jbo_probe_temperature(jb_ctx *c, double rho, double sie) { return c->cl.temperature(rho, sie); }
double jbo_probe_specific_heat(jb_ctx *c, double rho, double sie) { return c->cl.specific_heat(rho, sie); }
double jbo_probe_eps(int which) { return which == 0 ? jbo::eps_imc_offset : jbo::eps_ddmc_offset; }

// stream of doubles of one particle identity (PHILOX) — pins the generator itself
void jbo_probe_philox(uint64_t seed, uint32_t stream, uint32_t idx, uint32_t gid, uint32_t call, uint32_t first,
                      int32_t n, double *out) {
  jbo::RngGen g;
  g.mode = JB_RNG_PHILOX;
  g.id[0] = idx; g.id[1] = gid; g.id[2] = call;
  g.n = first;
  g.key[0] = (uint32_t)seed;
  g.key[1] = (uint32_t)(seed >> 32) ^ stream;
  for (int i = 0; i < n; ++i) out[i] = g.drand();
}
void jbo_probe_philox_raw(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  jbo::philox4x32_10(ctr, key, out);
}
void jbo_probe_xorshift(uint64_t seed, int32_t n, double *out) {
  jb_config cfg;
  jbo_config_default(&cfg);
  cfg.seed = (int64_t)seed;
  jb_ctx c;
  c.cfg = cfg;
  jbo::seed_pool(&c, 1);
  for (int i = 0; i < n; ++i) out[i] = c.pool[0].drand();
}

// One call of each per-particle function with explicit inputs and a PHILOX stream, so the
// restated and reference-header flavours can be compared bit for bit on random inputs.
// io = {t, x, y, z, vx, vy, vz}; flags out = {is_absorbed, is_scattered}; returns draws used.
int32_t jbo_probe_imc_step(uint64_t seed, uint32_t idx, double t_start, double dt, double ff, double aa, double ss,
                           double vv, double dx_push, int32_t ndim, const double lo[3], const double hi[3],
                           double io[7], int32_t flags[2]) {
  jbo::RngGen g;
  g.id[0] = idx;
  g.key[0] = (uint32_t)seed;
  g.key[1] = (uint32_t)(seed >> 32);
  bool a = false, s = false;
  const bool multi_d = ndim >= 2, three_d = ndim == 3;
  jbo::imc_step(g, t_start, dt, ff, aa, ss, vv, io[4], io[5], io[6], dx_push, multi_d, three_d, lo[0], lo[1], lo[2],
                hi[0], hi[1], hi[2], io[0], io[1], io[2], io[3], a, s);
  if (s) jbo::scatter_iso(g, vv, io[4], io[5], io[6]);
  flags[0] = a;
  flags[1] = s;
  return (int32_t)g.n;
}
// DDMC: albedo then (if admitted) step. P = {Px_l,Py_l,Pz_l,Px_u,Py_u,Pz_u};
// ijk in/out; flags = {is_absorbed, is_rejected}
int32_t jbo_probe_ddmc_step(uint64_t seed, uint32_t idx, double t_start, double dt, double ff, double aa, double ss,
                            double vv, int32_t ndim, const double lo[3], const double hi[3], const double P[6],
                            double io[7], int32_t ijk[3], int32_t flags[2]) {
  jbo::RngGen g;
  g.id[0] = idx;
  g.key[0] = (uint32_t)seed;
  g.key[1] = (uint32_t)(seed >> 32);
  bool a = false, s = false, rej = false;
  const bool multi_d = ndim >= 2, three_d = ndim == 3;
  jbo::DdmcArgs d{g, t_start, dt, ff, aa, ss, vv, multi_d, three_d, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2],
                  P[0], P[1], P[2], P[3], P[4], P[5], io[0], io[1], io[2], io[3], io[4], io[5], io[6],
                  ijk[0], ijk[1], ijk[2], a, s};
  jbo::ddmc_albedo(d, rej);
  if (!rej) jbo::ddmc_step(d);
  flags[0] = a;
  flags[1] = rej;
  return (int32_t)g.n;
}
double jbo_probe_planck(uint64_t seed, uint32_t idx, double sb, double temp) {
  jbo::RngGen g;
  g.id[0] = idx;
  g.key[0] = (uint32_t)seed;
  g.key[1] = (uint32_t)(seed >> 32);
  return jbo::planck_energy(g, sb, temp);
}

// jb_math.h probes (tests/test_math.py pins them against libm)
double jbo_math(int32_t fn, double x) {
  switch (fn) {
    case 0: return jbm::log(x);
    case 1: return jbm::sin(x);
    case 2: return jbm::cos(x);
    case 3: return jbm::acos(x);
    case 4: return jbm::expm1(x);
    case 5: return jbm::exp(x);
    case 6: return jbm::pow4(x);
    case 7: return jbm::log_unit(x);
  }
  return 0.0;
}
void jbo_math_array(int32_t fn, const double *x, double *y, int64_t n) {
  for (int64_t i = 0; i < n; ++i) y[i] = jbo_math(fn, x[i]);
}

}  // extern "C"
