// see package.hpp in this directory: declaration-only stand-in
#include "package.hpp"
