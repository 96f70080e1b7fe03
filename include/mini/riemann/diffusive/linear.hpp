// Same include path as the reference header include/mini/riemann/diffusive/linear.hpp; the B200 build of the type lives in
// mini/b200/shim.hpp.
#ifndef MINI_B200_FWD_RIEMANN_DIFFUSIVE_LINEAR_HPP_
#define MINI_B200_FWD_RIEMANN_DIFFUSIVE_LINEAR_HPP_
#include "mini/b200/shim.hpp"
#endif
