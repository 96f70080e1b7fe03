// Same include path as the reference header include/mini/constant/index.hpp: the index names drivers use on vectors
// (demo/laminar/blasius.cpp:21-22).
#ifndef MINI_B200_CONSTANT_INDEX_HPP_
#define MINI_B200_CONSTANT_INDEX_HPP_
namespace mini {
namespace constant {
namespace index {
static constexpr int X{0};
static constexpr int Y{1};
static constexpr int Z{2};
}  // namespace index
}  // namespace constant
}  // namespace mini
#endif
