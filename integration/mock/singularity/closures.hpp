// DECLARATION-ONLY stand-in for the closure types jaybenne is compiled with (JAYBENNE_EOS_TYPE,
// JAYBENNE_ABSORPTION_OPACITY_TYPE, JAYBENNE_SCATTERING_OPACITY_TYPE, jaybenne_config.hpp.in:19-30): the
// members integration/jaybenne_adapter.cpp calls on singularity-eos / singularity-opac variants
// (photon_variant.hpp, s_photon_variant.hpp, eos_variant.hpp), as mcblock instantiates them
// (src/mcblock/opacity.hpp:23-30, src/mcblock/eos.hpp:21-22). The real headers exist under
// /root/reference/external and are what the oracle's _ref build compiles; this stand-in only keeps the
// syntax check independent of that tree.
#ifndef JB_MOCK_SINGULARITY_CLOSURES_HPP_
#define JB_MOCK_SINGULARITY_CLOSURES_HPP_
namespace singularity {
using Real = double;
struct RuntimePhysicalConstants {
  Real time, mass, length, temp, c, sb;
};
struct IdealGas {};
namespace photons {
struct Gray {};
struct EPBremss {};
struct GrayS {};
struct ThomsonS {};
template <typename T>
struct NonCGSUnits {};
template <typename T>
struct NonCGSUnitsS {};
}  // namespace photons
namespace impl {
template <typename... Ts>
struct OpacVariant {
  template <typename T>
  bool IsType() const noexcept;
  RuntimePhysicalConstants GetRuntimePhysicalConstants() const;
  Real AbsorptionCoefficient(Real rho, Real temp, Real nu, Real *lambda = nullptr) const;
  Real Emissivity(Real rho, Real temp, Real *lambda = nullptr) const;
  OpacVariant GetOnDevice();
};
template <typename... Ts>
struct ScatVariant {
  template <typename T>
  bool IsType() const noexcept;
  Real TotalScatteringCoefficient(Real rho, Real temp, Real nu, Real *lambda = nullptr) const;
  ScatVariant GetOnDevice();
};
template <typename... Ts>
struct EosVariant {
  template <typename T>
  bool IsType() const noexcept;
  Real TemperatureFromDensityInternalEnergy(Real rho, Real sie, Real *lambda = nullptr) const;
  Real SpecificHeatFromDensityTemperature(Real rho, Real temp, Real *lambda = nullptr) const;
  Real GruneisenParamFromDensityTemperature(Real rho, Real temp, Real *lambda = nullptr) const;
  EosVariant GetOnDevice();
};
}  // namespace impl
}  // namespace singularity
namespace mcblock {
using Opacity = singularity::impl::OpacVariant<singularity::photons::NonCGSUnits<singularity::photons::Gray>,
                                               singularity::photons::NonCGSUnits<singularity::photons::EPBremss>>;
using Scattering = singularity::impl::ScatVariant<singularity::photons::NonCGSUnitsS<singularity::photons::GrayS>,
                                                  singularity::photons::NonCGSUnitsS<singularity::photons::ThomsonS>>;
using EOS = singularity::impl::EosVariant<singularity::IdealGas>;
}  // namespace mcblock
#endif
