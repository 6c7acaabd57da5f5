// symm_allreduce.cu -- in-place SUM of a flat bucket across the GPUs of one NVLink domain,
// written against peer-mapped ("symmetric") memory instead of calling NCCL.
//
// Reference behaviour being replaced: `dist.all_reduce(coalesced)` inside apply_flat_dist_call
// (apex/parallel/distributed.py:15-17) -- NCCL's generic allreduce of the flat bucket.
//
// Two-shot algorithm, rank r owns slice r of the buffer:
//   multicast (NVSwitch, NVLS):  v = multimem.ld_reduce(slice r)   -- the switch adds the n copies
//                                multimem.st(slice r, v)           -- the switch writes all n copies
//   peer-to-peer (no multicast): v = sum over peers of LDG(peer buffer, slice r)  (fp32 accumulate)
//                                STG(every peer buffer, slice r, v)
// Per GPU and direction about one buffer size crosses NVLink however many ranks there are
// (a ring moves 2(n-1)/n of it through every link hop by hop).  fp16 sums are accumulated in fp32
// and rounded once to fp16, so a sum beyond 65504 still becomes inf (the averaging that follows
// must keep tripping the loss scaler).
//
// The caller (apex_b200/parallel/symm.py) orders the phases across ranks with the symmetric-memory
// barrier of torch.distributed: all ranks' gathers before this kernel, this kernel on all ranks
// before anyone scatters.  The kernel itself never waits on another GPU.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/apex_b200.h"

namespace apex_b200 {

constexpr int kArThreads = 512;
constexpr int kArUnroll = 4;
constexpr int kMaxPeers = 16;

struct PeerPtrs { char* p[kMaxPeers]; };

__device__ __forceinline__ uint4 mc_ld_reduce_f16(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
    return r;
}
__device__ __forceinline__ uint4 mc_ld_reduce_f32(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
    return r;
}
__device__ __forceinline__ void mc_st(void* mc, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_peer(const void* p) {
    uint4 r;
    asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void st_peer(void* p, const uint4& v) {
    asm volatile("st.global.relaxed.sys.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// 16-byte vectors [first, first + count) of the buffer belong to this rank.
template <bool kHalf>
__global__ void __launch_bounds__(kArThreads)
nvls_allreduce_kernel(char* mc, int64_t first, int64_t count) {
    const int64_t stride = int64_t(gridDim.x) * kArThreads;
    int64_t i = int64_t(blockIdx.x) * kArThreads + threadIdx.x;
    for (; i + (kArUnroll - 1) * stride < count; i += kArUnroll * stride) {
        uint4 v[kArUnroll];
#pragma unroll
        for (int u = 0; u < kArUnroll; ++u) {
            char* a = mc + ((first + i + u * stride) << 4);
            v[u] = kHalf ? mc_ld_reduce_f16(a) : mc_ld_reduce_f32(a);
        }
#pragma unroll
        for (int u = 0; u < kArUnroll; ++u) mc_st(mc + ((first + i + u * stride) << 4), v[u]);
    }
    for (; i < count; i += stride) {
        char* a = mc + ((first + i) << 4);
        mc_st(a, kHalf ? mc_ld_reduce_f16(a) : mc_ld_reduce_f32(a));
    }
}

template <bool kHalf>
__device__ __forceinline__ void accumulate(float* acc, const uint4& v) {
    if (kHalf) {
        const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
        for (int k = 0; k < 4; ++k) { const float2 f = __half22float2(h[k]); acc[2 * k] += f.x; acc[2 * k + 1] += f.y; }
    } else {
        const float* f = reinterpret_cast<const float*>(&v);
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[k] += f[k];
    }
}

template <bool kHalf>
__global__ void __launch_bounds__(kArThreads)
p2p_allreduce_kernel(PeerPtrs peers, int world, int rank, int64_t first, int64_t count) {
    const int64_t stride = int64_t(gridDim.x) * kArThreads;
    for (int64_t i = int64_t(blockIdx.x) * kArThreads + threadIdx.x; i < count; i += stride) {
        const int64_t off = (first + i) << 4;
        // Rank k visits the peers in the order k, k+1, ...: at any moment the ranks pull from
        // different GPUs instead of all draining the same one.  Each slice is summed by exactly one
        // rank and then written everywhere, so every rank ends with identical bits.
        uint4 in[kMaxPeers];
#pragma unroll
        for (int j = 0; j < kMaxPeers; ++j)
            if (j < world) in[j] = ld_peer(peers.p[(rank + j) % world] + off);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int j = 0; j < kMaxPeers; ++j)
            if (j < world) accumulate<kHalf>(acc, in[j]);
        uint4 out;
        if (kHalf) {
            __half2* h = reinterpret_cast<__half2*>(&out);
#pragma unroll
            for (int k = 0; k < 4; ++k) h[k] = __floats2half2_rn(acc[2 * k], acc[2 * k + 1]);
        } else {
            float* f = reinterpret_cast<float*>(&out);
#pragma unroll
            for (int k = 0; k < 4; ++k) f[k] = acc[k];
        }
#pragma unroll
        for (int j = 0; j < kMaxPeers; ++j)
            if (j < world) st_peer(peers.p[(rank + j) % world] + off, out);
    }
}

}  // namespace apex_b200

using namespace apex_b200;

extern "C" int apex_b200_symm_allreduce(const uint64_t* peer_ptrs, uint64_t multicast_ptr, int rank, int world,
                                        int64_t numel, int dtype, int ctas, void* stream) {
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || numel < 0) return APEX_B200_EINVAL;
    if (dtype != 0 && dtype != 1) return APEX_B200_EINVAL;
    if (multicast_ptr == 0 && peer_ptrs == nullptr) return APEX_B200_EINVAL;
    const int per_vec = dtype == 0 ? 8 : 4;
    if (numel % (int64_t(per_vec) * world) != 0) return APEX_B200_EINVAL;     // caller pads the buffer
    if (numel == 0 || world == 1) return 0;
    const int64_t count = numel / per_vec / world;          // 16-byte vectors in this rank's slice
    const int64_t first = count * rank;
    if (ctas <= 0) ctas = 64;
    const int64_t need = (count + kArThreads - 1) / kArThreads;
    const int grid = int(need < ctas ? need : ctas);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (multicast_ptr != 0) {
        char* mc = reinterpret_cast<char*>(multicast_ptr);
        if (dtype == 0) nvls_allreduce_kernel<true><<<grid, kArThreads, 0, s>>>(mc, first, count);
        else nvls_allreduce_kernel<false><<<grid, kArThreads, 0, s>>>(mc, first, count);
    } else {
        PeerPtrs pp;
        for (int r = 0; r < kMaxPeers; ++r) pp.p[r] = r < world ? reinterpret_cast<char*>(peer_ptrs[r]) : nullptr;
        if (dtype == 0) p2p_allreduce_kernel<true><<<grid, kArThreads, 0, s>>>(pp, world, rank, first, count);
        else p2p_allreduce_kernel<false><<<grid, kArThreads, 0, s>>>(pp, world, rank, first, count);
    }
    cudaError_t e = cudaPeekAtLastError();
    return e == cudaSuccess ? 0 : int(e);
}
