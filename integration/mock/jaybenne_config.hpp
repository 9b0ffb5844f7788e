// What CMake generates from src/jaybenne/jaybenne_config.hpp.in:19-30 for the mcblock host
// (src/CMakeLists.txt:18-45), with the stand-in closure types of mock/singularity/closures.hpp.
#ifndef JAYBENNE_JAYBENNE_CONFIG_HPP_
#define JAYBENNE_JAYBENNE_CONFIG_HPP_
#include <string>

#include "singularity/closures.hpp"
using mcblock::EOS;
using mcblock::Opacity;
using mcblock::Scattering;
// src/mcblock/mcblock_variables.hpp: the host's variable types (only ::name() is used)
namespace field::material {
struct density { static std::string name(); };
struct sie { static std::string name(); };
struct internal_energy { static std::string name(); };
}  // namespace field::material
#define HOST_DENSITY field::material::density
#define HOST_SPECIFIC_INTERNAL_ENERGY field::material::sie
#define HOST_UPDATE_ENERGY field::material::internal_energy
#endif
