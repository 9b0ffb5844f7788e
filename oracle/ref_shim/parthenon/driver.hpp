// TEST INFRASTRUCTURE — minimal stand-in for <parthenon/driver.hpp> so that the
// reference's per-particle headers (src/jaybenne/transport_utils.hpp, scattering.hpp,
// planck.hpp) compile VERBATIM from /root/reference with plain g++ (Parthenon itself is
// absent from the reference mount: external/parthenon is an empty submodule).
// Only the handful of names those three headers touch are provided.
#ifndef JBO_REF_SHIM_PARTHENON_DRIVER_HPP_
#define JBO_REF_SHIM_PARTHENON_DRIVER_HPP_

#include <cmath>
#include <cstdint>
#include <limits>

using Real = double;  // ports-of-call portability.hpp:84-91 (Real = double)

#ifndef KOKKOS_FORCEINLINE_FUNCTION
#define KOKKOS_FORCEINLINE_FUNCTION inline
#endif
#ifndef KOKKOS_INLINE_FUNCTION
#define KOKKOS_INLINE_FUNCTION inline
#endif
#ifndef PARTHENON_DEBUG_REQUIRE
#define PARTHENON_DEBUG_REQUIRE(cond, msg) ((void)0)
#endif

namespace parthenon {
namespace robust {
// UNVERIFIED against Parthenon (source absent); same idiom as
// singularity-opac/base/robust_utils.hpp:29-32.
template <typename T = Real>
constexpr T EPS() {
  return 10 * std::numeric_limits<T>::epsilon();
}
}  // namespace robust
}  // namespace parthenon

// The reference takes its generator by reference everywhere (tran_step_args::rng_gen,
// scatter, sample_Planck_energy), so any type with drand() can stand in for
// Kokkos::Random_XorShift64. The oracle's generator is defined in jb_oracle.cpp.
namespace jbo {
struct RngGen;
}
typedef jbo::RngGen RngGen;

// src/jaybenne/jaybenne_utils.hpp:44-49 (transport_utils.hpp uses it without including it)
inline bool fuzzy_equal(const Real &a, const Real &b, const Real &c, const Real &eps) {
  return std::abs(a - b) < c * eps;
}

#endif  // JBO_REF_SHIM_PARTHENON_DRIVER_HPP_
