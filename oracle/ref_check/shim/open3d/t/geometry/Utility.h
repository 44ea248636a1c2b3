#pragma once
#include <open3d/core/Tensor.h>
namespace open3d { namespace t { namespace geometry { inline void CheckIntrinsicTensor(const core::Tensor&) {} }}}
