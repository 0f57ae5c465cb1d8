// Host stand-in for the CUDA/HIP kernel vocabulary — TEST INFRASTRUCTURE ONLY.
//
// Lets the reference's own GPU kernel sources (fortran/hip/heat_kernel.cpp and the kernel string
// in python/cuda/cuda.py) be compiled UNMODIFIED with g++ and executed on the CPU, one emulated
// thread at a time, so that their arithmetic can pin the oracle (oracle/_ref, see oracle/Makefile).
// Kernels in those sources use no shared memory and no synchronisation that matters, so running
// the threads of a launch sequentially is equivalent.
#pragma once
#include <cstddef>

struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct uint3_shim { unsigned x, y, z; };

inline thread_local uint3_shim threadIdx, blockIdx;
inline thread_local dim3 blockDim, gridDim;

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline

template <typename K, typename... A>
inline void shim_launch(K kernel, dim3 grid, dim3 block, A... args) {
    gridDim = grid;
    blockDim = block;
    for (unsigned bz = 0; bz < grid.z; ++bz)
    for (unsigned by = 0; by < grid.y; ++by)
    for (unsigned bx = 0; bx < grid.x; ++bx) {
        blockIdx = {bx, by, bz};
        for (unsigned tz = 0; tz < block.z; ++tz)
        for (unsigned ty = 0; ty < block.y; ++ty)
        for (unsigned tx = 0; tx < block.x; ++tx) {
            threadIdx = {tx, ty, tz};
            kernel(args...);
        }
    }
}

#define hipLaunchKernelGGL(kernel, grid, block, shmem, stream, ...) \
    shim_launch(kernel, dim3(grid), dim3(block), __VA_ARGS__)
