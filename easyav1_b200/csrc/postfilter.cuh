// postfilter.cuh — whole-frame in-loop filter kernels (deblock, CDEF, loop restoration).
#pragma once
#include "dev_common.cuh"
namespace eav1 {
} // namespace eav1
