// stand-in for <cuda_runtime.h> (device_global.hpp:4): the host build under oracle/_ref needs only
// the handle typedefs, which ref_prelude.h provides.
#pragma once
