// TEST INFRASTRUCTURE — see driver.hpp in this directory.
#include "driver.hpp"
