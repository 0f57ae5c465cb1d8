// Resolves `#include <hip/hip_runtime.h>` (fortran/hip/heat_kernel.cpp:1) to the host shim.
#pragma once
#include "../gpu_host_shim.h"
