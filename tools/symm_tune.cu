// tools/symm_tune.cu -- variants of the multicast allreduce kernel (development tool, not in the library).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -shared -Xcompiler -fPIC -o tools/libsymm_tune.so tools/symm_tune.cu
#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ uint4 ldr_relaxed(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
    return r;
}
__device__ __forceinline__ uint4 ldr_weak(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.weak.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
    return r;
}
__device__ __forceinline__ uint4 ldr_noacc(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
    return r;
}
__device__ __forceinline__ void mst_relaxed(void* mc, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void mst_weak(void* mc, const uint4& v) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// SEM: 0 relaxed.sys, 1 weak, 2 relaxed without fp32 accumulate
template <int T, int U, int SEM>
__global__ void __launch_bounds__(T) k_fused(char* mc, int64_t first, int64_t count) {
    const int64_t stride = int64_t(gridDim.x) * T;
    int64_t i = int64_t(blockIdx.x) * T + threadIdx.x;
    for (; i + (U - 1) * stride < count; i += U * stride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { char* a = mc + ((first + i + u * stride) << 4); v[u] = SEM == 0 ? ldr_relaxed(a) : (SEM == 1 ? ldr_weak(a) : ldr_noacc(a)); }
#pragma unroll
        for (int u = 0; u < U; ++u) { char* a = mc + ((first + i + u * stride) << 4); if (SEM == 1) mst_weak(a, v[u]); else mst_relaxed(a, v[u]); }
    }
    for (; i < count; i += stride) { char* a = mc + ((first + i) << 4); uint4 v = ldr_relaxed(a); mst_relaxed(a, v); }
}
// reduce only: result to my own copy (unicast store)
template <int T, int U>
__global__ void __launch_bounds__(T) k_reduce(char* mc, char* mine, int64_t first, int64_t count) {
    const int64_t stride = int64_t(gridDim.x) * T;
    int64_t i = int64_t(blockIdx.x) * T + threadIdx.x;
    for (; i + (U - 1) * stride < count; i += U * stride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = ldr_relaxed(mc + ((first + i + u * stride) << 4));
#pragma unroll
        for (int u = 0; u < U; ++u) *reinterpret_cast<uint4*>(mine + ((first + i + u * stride) << 4)) = v[u];
    }
    for (; i < count; i += stride) *reinterpret_cast<uint4*>(mine + ((first + i) << 4)) = ldr_relaxed(mc + ((first + i) << 4));
}
// broadcast only: my slice of my copy to every copy
template <int T, int U>
__global__ void __launch_bounds__(T) k_bcast(char* mc, const char* mine, int64_t first, int64_t count) {
    const int64_t stride = int64_t(gridDim.x) * T;
    int64_t i = int64_t(blockIdx.x) * T + threadIdx.x;
    for (; i + (U - 1) * stride < count; i += U * stride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = *reinterpret_cast<const uint4*>(mine + ((first + i + u * stride) << 4));
#pragma unroll
        for (int u = 0; u < U; ++u) mst_relaxed(mc + ((first + i + u * stride) << 4), v[u]);
    }
    for (; i < count; i += stride) mst_relaxed(mc + ((first + i) << 4), *reinterpret_cast<const uint4*>(mine + ((first + i) << 4)));
}

extern "C" int symm_tune_launch(int variant, uint64_t mc_ptr, uint64_t my_ptr, int rank, int world, int64_t numel_f16,
                                int ctas, void* stream) {
    const int64_t count = numel_f16 / 8 / world, first = count * rank;
    char* mc = reinterpret_cast<char*>(mc_ptr);
    char* mine = reinterpret_cast<char*>(my_ptr);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    switch (variant) {
        case 0: k_fused<512, 4, 0><<<ctas, 512, 0, s>>>(mc, first, count); break;
        case 1: k_fused<1024, 4, 0><<<ctas, 1024, 0, s>>>(mc, first, count); break;
        case 2: k_fused<512, 8, 0><<<ctas, 512, 0, s>>>(mc, first, count); break;
        case 3: k_fused<256, 8, 0><<<ctas, 256, 0, s>>>(mc, first, count); break;
        case 4: k_fused<512, 4, 1><<<ctas, 512, 0, s>>>(mc, first, count); break;
        case 5: k_fused<512, 4, 2><<<ctas, 512, 0, s>>>(mc, first, count); break;
        case 6: k_fused<512, 1, 0><<<ctas, 512, 0, s>>>(mc, first, count); break;
        case 7: k_fused<512, 16, 0><<<ctas, 512, 0, s>>>(mc, first, count); break;
        case 10: k_reduce<512, 4><<<ctas, 512, 0, s>>>(mc, mine, first, count); break;
        case 11: k_bcast<512, 4><<<ctas, 512, 0, s>>>(mc, mine, first, count); break;
        case 12: k_reduce<512, 8><<<ctas, 512, 0, s>>>(mc, mine, first, count); k_bcast<512, 8><<<ctas, 512, 0, s>>>(mc, mine, first, count); break;
        default: return -1;
    }
    return int(cudaPeekAtLastError());
}

// explicit 16-byte-vector range [first, first+count): for chunked, two-stream overlap experiments
extern "C" int symm_tune_launch_range(int variant, uint64_t mc_ptr, uint64_t my_ptr, int64_t first, int64_t count,
                                      int ctas, void* stream) {
    char* mc = reinterpret_cast<char*>(mc_ptr);
    char* mine = reinterpret_cast<char*>(my_ptr);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    switch (variant) {
        case 0: k_fused<512, 4, 0><<<ctas, 512, 0, s>>>(mc, first, count); break;
        case 10: k_reduce<512, 4><<<ctas, 512, 0, s>>>(mc, mine, first, count); break;
        case 11: k_bcast<512, 4><<<ctas, 512, 0, s>>>(mc, mine, first, count); break;
        default: return -1;
    }
    return int(cudaPeekAtLastError());
}
