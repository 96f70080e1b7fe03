// Same include path as the reference header include/mini/spatial/with_viscosity.hpp; the B200 build of the type lives in
// mini/b200/shim.hpp.
#ifndef MINI_B200_FWD_SPATIAL_WITH_VISCOSITY_HPP_
#define MINI_B200_FWD_SPATIAL_WITH_VISCOSITY_HPP_
#include "mini/b200/shim.hpp"
#endif
