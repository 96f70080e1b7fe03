// Same include path as the reference header include/mini/limiter/majority.hpp; the B200 build of the type lives in
// mini/b200/shim.hpp.
#ifndef MINI_B200_FWD_LIMITER_MAJORITY_HPP_
#define MINI_B200_FWD_LIMITER_MAJORITY_HPP_
#include "mini/b200/shim.hpp"
#endif
