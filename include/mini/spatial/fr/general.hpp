// Same include path as the reference header include/mini/spatial/fr/general.hpp; the B200 build of the type lives in
// mini/b200/shim.hpp.
#ifndef MINI_B200_FWD_SPATIAL_FR_GENERAL_HPP_
#define MINI_B200_FWD_SPATIAL_FR_GENERAL_HPP_
#include "mini/b200/shim.hpp"
#endif
