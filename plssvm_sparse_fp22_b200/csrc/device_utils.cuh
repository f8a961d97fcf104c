// Device helpers: warp/block reductions, deterministic grid-wide sums, kernel-function epilogues.
#pragma once

#include "internal.cuh"

namespace b200 {

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { v += __shfl_xor_sync(0xffffffffu, v, o); }
    return v;
}

// Sum over the block; result valid in thread 0.  `red` holds at least 32 elements of T.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T *red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();  // protect `red` against a previous use
    if (lane == 0) { red[warp] = v; }
    __syncthreads();
    if (warp == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        v = lane < nw ? red[lane] : T(0);
        v = warp_sum(v);
    }
    return v;
}

// Deterministic grid-wide reduction of NV values per block.
//   partials: [NV][MAX_PARTIAL_BLOCKS], ticket: one unsigned int, zero before the launch (re-armed by the last block).
// Returns true in thread 0 of the LAST block to finish, with `out[k]` = sum over blocks (fixed order => bit-reproducible
// for a fixed grid size).  All other threads / blocks return false.
template <typename T, int NV>
__device__ __forceinline__ bool grid_sum(T (&v)[NV], T *partials, unsigned int *ticket, T *red, T (&out)[NV]) {
    __shared__ bool is_last;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const T s = block_sum(v[k], red);
        if (threadIdx.x == 0) { partials[k * MAX_PARTIAL_BLOCKS + blockIdx.x] = s; }
    }
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned int t = atomicAdd(ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) { return false; }
    __threadfence();
    if (threadIdx.x < 32) {
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            T s = 0;
            for (unsigned int b = threadIdx.x; b < gridDim.x; b += 32) { s += __ldcg(&partials[k * MAX_PARTIAL_BLOCKS + b]); }
            out[k] = warp_sum(s);
        }
        if (threadIdx.x == 0) { *ticket = 0u; }
    }
    return threadIdx.x == 0;
}

// x^e for integer e (the reference calls std::pow(x, (real) degree), include/plssvm/kernel_types.hpp:74-77)
template <typename T>
__device__ __forceinline__ T ipow(T x, int e) {
    const bool neg = e < 0;
    unsigned int u = neg ? static_cast<unsigned int>(-e) : static_cast<unsigned int>(e);
    T r = T(1);
    while (u != 0u) {
        if (u & 1u) { r *= x; }
        x *= x;
        u >>= 1;
    }
    return neg ? T(1) / r : r;
}

__device__ __forceinline__ float fast_exp(float x) { return __expf(x); }
__device__ __forceinline__ double fast_exp(double x) { return exp(x); }

// Kernel value from the Gram entry g = <a, b> and the squared norms (include/plssvm/kernel_types.hpp:63-85):
//   linear g ; polynomial (gamma g + coef0)^degree ; rbf exp(-gamma |a-b|^2) with |a-b|^2 = na + nb - 2 g.
struct kparams {
    int degree;
    double gamma, coef0;
};

template <typename T, int KERNEL>
__device__ __forceinline__ T kernel_value(T g, T na, T nb, int degree, T gamma, T coef0) {
    if constexpr (KERNEL == B200SVM_LINEAR) {
        return g;
    } else if constexpr (KERNEL == B200SVM_POLYNOMIAL) {
        return ipow(fma(gamma, g, coef0), degree);
    } else {
        T d2 = fma(T(-2), g, na + nb);
        d2 = d2 > T(0) ? d2 : T(0);
        return fast_exp(-gamma * d2);
    }
}

}  // namespace b200
