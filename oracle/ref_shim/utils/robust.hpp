// TEST INFRASTRUCTURE — stand-in for Parthenon's <utils/robust.hpp>; see ../parthenon/driver.hpp.
#include "../parthenon/driver.hpp"
