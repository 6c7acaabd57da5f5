// symm_allreduce.cu -- averaging of a flat gradient bucket across the GPUs of one NVLink domain in ONE
// kernel per rank: cross-rank barrier, two-shot SUM, x (1/world), inf/NaN check, cross-rank barrier.
//
// Reference behaviour being replaced (apex/parallel/distributed.py:11-23, apply_flat_dist_call):
//     dist.all_reduce(coalesced)                 NCCL's generic allreduce          (:15-17)
//     coalesced /= dist.get_world_size()         one more pass over the bucket     (:19-20)
// and, on the B200 path, the overflow scan FP16_Optimizer would otherwise run over the averaged
// gradients (apex/fp16_utils/fp16_optimizer.py:213-221, loss_scaler.py:84-110).
//
// The bucket lives in peer-mapped ("symmetric") memory: every rank can address every rank's copy and,
// on NVSwitch, a multicast alias of all copies.  Rank r owns slice r:
//   multicast (NVSwitch, NVLS):  v = multimem.ld_reduce(slice r)     the switch adds the n copies
//                                v = T(float(v) * (1/world)), check  exactly the `/=` of the reference
//                                multimem.st(slice r, v)             the switch writes all n copies
//   peer-to-peer:                v = sum over peers of LD(peer copy, slice r)   (fp32 accumulate,
//                                    several 16 B loads per peer in flight per thread)
//                                scale + check as above, ST to every peer's copy
// fp16 sums are accumulated in fp32 and rounded once to fp16 BEFORE the multiply, so a sum beyond
// 65504 becomes inf and stays inf through the division, as it does in the reference (the loss scaler
// must keep tripping on it).
//
// Cross-rank ordering is inside the kernel (no separate barrier launches): CTA c of every rank
// exchanges one 32-bit epoch word with CTA c of every other rank through a signal area in the same
// symmetric allocation -- once before touching peer data ("every rank's bucket is final": the peer's
// kernel has started, so everything ordered before it on the peer's stream is done) and once after
// its last store ("my slice is written everywhere").  A kernel therefore exits only after every CTA
// of every rank has finished its stores, and whatever follows it in stream order may read the whole
// averaged bucket.  Epochs only grow (2*seq, 2*seq+1), so nothing is ever reset.  The kernels wait
// on OTHER GPUs only; one process per GPU as with NCCL (never several ranks on one GPU).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/apex_b200.h"

namespace apex_b200 {

constexpr int kArThreads = 512;
constexpr int kMaxPeers = APEX_B200_SYMM_MAX_WORLD;

struct SymmArgs {
    char* peer[kMaxPeers];         // the bucket on every rank
    char* mc;                      // multicast alias of the bucket (nvls kernels)
    uint32_t* sig[kMaxPeers];      // signal area of this channel on every rank: [MAX_CTAS][MAX_WORLD]
    int32_t* flag[kMaxPeers];      // overflow word on every rank, or all null
    int64_t first, count;          // this rank's 16-byte vectors: [first, first + count)
    unsigned long long timeout_ns; // 0 = wait forever
    float mul;
    int apply_mul;
    int check;
    int rank, world;
    uint32_t epoch;                // 2 * seq
};

// ---- cross-rank barrier -----------------------------------------------------------------------------

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// Thread t < world tells peer t "CTA blockIdx.x of rank `rank` reached `epoch`" and waits for the
// same word from peer t.  The release is cumulative: everything the signalling thread has observed
// through the CTA barrier -- every earlier store of the whole CTA, data and flag -- is visible to a
// peer that has acquired the word (SASS: MEMBAR.ALL.SYS in front of the signal store; an extra
// __threadfence_system() here only added a second, sequentially consistent fence per barrier).
__device__ __forceinline__ void symm_barrier(const SymmArgs& a, uint32_t epoch) {
    __syncthreads();
    if (threadIdx.x < a.world) {
        const int peer = threadIdx.x;
        const size_t slot = size_t(blockIdx.x) * kMaxPeers;
        st_release_sys(a.sig[peer] + slot + a.rank, epoch);
        const uint32_t* mine = a.sig[a.rank] + slot + peer;
        unsigned long long t0 = 0;
        uint32_t spins = 0;
        while (int32_t(ld_acquire_sys(mine) - epoch) < 0) {
            if (a.timeout_ns != 0 && (++spins & 0x3ff) == 0) {
                const unsigned long long now = global_ns();
                if (t0 == 0) t0 = now;
                else if (now - t0 > a.timeout_ns) {
                    printf("apex_b200: rank %d CTA %d waited too long for rank %d (epoch %u)\n",
                           a.rank, blockIdx.x, peer, epoch);
                    __trap();
                }
            }
        }
    }
    __syncthreads();
}

__device__ __forceinline__ void raise_everywhere(const SymmArgs& a, bool bad) {
    if (!a.check) return;
    const int block_bad = __syncthreads_or(bad ? 1 : 0);
    if (block_bad && threadIdx.x < a.world) atomicMax_system(a.flag[threadIdx.x], 0x7fffffff);
}

// ---- 16-byte vectors ------------------------------------------------------------------------------------

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

// `coalesced /= world` on a vector that already holds the sum in the bucket dtype: torch's CUDA
// true-divide by a python scalar multiplies by float(1.0 / world) in fp32 and rounds to the tensor
// dtype (the arithmetic of mt_scale_unflatten in mt_convert.cu).  Also the inf/NaN test of the value
// that is written.
template <bool kHalf>
__device__ __forceinline__ uint4 scale_check(uint4 v, float mul, int apply, bool& bad) {
    if (kHalf) {
        __half2* h = reinterpret_cast<__half2*>(&v);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (apply) {
                float2 f = __half22float2(h[k]);
                f.x = __fmul_rn(f.x, mul);
                f.y = __fmul_rn(f.y, mul);
                h[k] = __floats2half2_rn(f.x, f.y);
            }
            const uint32_t w = *reinterpret_cast<const uint32_t*>(&h[k]);
            bad |= ((w & 0x7c00u) == 0x7c00u) | ((w & 0x7c000000u) == 0x7c000000u);
        }
    } else {
        float* f = reinterpret_cast<float*>(&v);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (apply) f[k] = __fmul_rn(f[k], mul);
            bad |= (__float_as_uint(f[k]) & 0x7f800000u) == 0x7f800000u;
        }
    }
    return v;
}

// ---- NVSwitch multicast variant -------------------------------------------------------------------------

constexpr int kMcUnroll = 4;

template <bool kHalf>
__global__ void __launch_bounds__(kArThreads)
nvls_allreduce_kernel(SymmArgs a) {
    symm_barrier(a, a.epoch);
    const int64_t stride = int64_t(gridDim.x) * kArThreads;
    int64_t i = int64_t(blockIdx.x) * kArThreads + threadIdx.x;
    bool bad = false;
    for (; i + (kMcUnroll - 1) * stride < a.count; i += kMcUnroll * stride) {
        uint4 v[kMcUnroll];
#pragma unroll
        for (int u = 0; u < kMcUnroll; ++u) {
            char* p = a.mc + ((a.first + i + u * stride) << 4);
            v[u] = kHalf ? mc_ld_reduce_f16(p) : mc_ld_reduce_f32(p);
        }
#pragma unroll
        for (int u = 0; u < kMcUnroll; ++u)
            mc_st(a.mc + ((a.first + i + u * stride) << 4), scale_check<kHalf>(v[u], a.mul, a.apply_mul, bad));
    }
    for (; i < a.count; i += stride) {
        char* p = a.mc + ((a.first + i) << 4);
        mc_st(p, scale_check<kHalf>(kHalf ? mc_ld_reduce_f16(p) : mc_ld_reduce_f32(p), a.mul, a.apply_mul, bad));
    }
    raise_everywhere(a, bad);
    symm_barrier(a, a.epoch + 1);
}

// ---- peer-to-peer variant -------------------------------------------------------------------------------

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
__device__ __forceinline__ uint4 round_sum(const float* acc) {
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
    return out;
}

// W = compile-time world size (2, 4, 8) with U vectors per peer in flight per thread; W = 0 is the
// generic kernel (any world size up to kMaxPeers, one vector per peer in flight).
template <bool kHalf, int W, int U>
__global__ void __launch_bounds__(kArThreads)
p2p_allreduce_kernel(SymmArgs a) {
    symm_barrier(a, a.epoch);
    constexpr int kW = W == 0 ? kMaxPeers : W;
    const int world = W == 0 ? a.world : W;
    const int64_t stride = int64_t(gridDim.x) * kArThreads;
    int64_t i = int64_t(blockIdx.x) * kArThreads + threadIdx.x;
    bool bad = false;
    // Rank k visits the peers in the order k, k+1, ...: at any moment the ranks pull from different
    // GPUs instead of all draining the same one.  Each slice is summed by exactly one rank and then
    // written everywhere, so every rank ends with identical bits.
    char* src[kW];
#pragma unroll
    for (int j = 0; j < kW; ++j) src[j] = j < world ? a.peer[(a.rank + j) % world] : nullptr;
    for (; i + (U - 1) * stride < a.count; i += U * stride) {
        uint4 in[U][kW];
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int j = 0; j < kW; ++j)
                if (j < world) in[u][j] = ld_peer(src[j] + ((a.first + i + u * stride) << 4));
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
            for (int j = 0; j < kW; ++j)
                if (j < world) accumulate<kHalf>(acc, in[u][j]);
            const uint4 out = scale_check<kHalf>(round_sum<kHalf>(acc), a.mul, a.apply_mul, bad);
#pragma unroll
            for (int j = 0; j < kW; ++j)
                if (j < world) st_peer(src[j] + ((a.first + i + u * stride) << 4), out);
        }
    }
    for (; i < a.count; i += stride) {
        const int64_t off = (a.first + i) << 4;
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int j = 0; j < kW; ++j)
            if (j < world) accumulate<kHalf>(acc, ld_peer(src[j] + off));
        const uint4 out = scale_check<kHalf>(round_sum<kHalf>(acc), a.mul, a.apply_mul, bad);
#pragma unroll
        for (int j = 0; j < kW; ++j)
            if (j < world) st_peer(src[j] + off, out);
    }
    raise_everywhere(a, bad);
    symm_barrier(a, a.epoch + 1);
}

template <bool kHalf>
static void launch_p2p(const SymmArgs& a, int grid, cudaStream_t s) {
    switch (a.world) {
        case 2: p2p_allreduce_kernel<kHalf, 2, 8><<<grid, kArThreads, 0, s>>>(a); break;
        case 4: p2p_allreduce_kernel<kHalf, 4, 4><<<grid, kArThreads, 0, s>>>(a); break;
        case 8: p2p_allreduce_kernel<kHalf, 8, 2><<<grid, kArThreads, 0, s>>>(a); break;
        default: p2p_allreduce_kernel<kHalf, 0, 1><<<grid, kArThreads, 0, s>>>(a); break;
    }
}

}  // namespace apex_b200

using namespace apex_b200;

extern "C" int apex_b200_symm_allreduce(const apex_b200_symm_domain* d, uint64_t offset, int64_t numel,
                                        int dtype, double divisor, int apply_mul, int64_t flag_offset,
                                        int channel, uint32_t seq, int use_multicast, int ctas, void* stream) {
    if (d == nullptr) return APEX_B200_EINVAL;
    const int world = d->world, rank = d->rank;
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || numel < 0) return APEX_B200_EINVAL;
    if (dtype != 0 && dtype != 1) return APEX_B200_EINVAL;
    if (channel < 0 || channel >= APEX_B200_SYMM_CHANNELS || seq == 0 || seq >= 0x7fffffffu) return APEX_B200_EINVAL;
    if (apply_mul && divisor == 0.0) return APEX_B200_EINVAL;
    if ((offset & 15) != 0 || (flag_offset >= 0 && (flag_offset & 3) != 0)) return APEX_B200_EINVAL;
    if (use_multicast && d->multicast_base == 0) return APEX_B200_EINVAL;
    const int per_vec = dtype == 0 ? 8 : 4;
    if (numel % (int64_t(per_vec) * world) != 0) return APEX_B200_EINVAL;     // caller pads the bucket
    if (ctas > APEX_B200_SYMM_MAX_CTAS) return APEX_B200_ERANGE;
    if (numel == 0) return 0;

    SymmArgs a;
    for (int r = 0; r < kMaxPeers; ++r) {
        const bool live = r < world;
        if (live && d->peer_base[r] == 0) return APEX_B200_EINVAL;
        a.peer[r] = live ? reinterpret_cast<char*>(d->peer_base[r] + offset) : nullptr;
        a.sig[r] = live ? reinterpret_cast<uint32_t*>(d->peer_base[r] + d->signal_offset) +
                              size_t(channel) * APEX_B200_SYMM_MAX_CTAS * kMaxPeers
                        : nullptr;
        a.flag[r] = (live && flag_offset >= 0) ? reinterpret_cast<int32_t*>(d->peer_base[r] + uint64_t(flag_offset)) : nullptr;
    }
    a.mc = use_multicast ? reinterpret_cast<char*>(d->multicast_base + offset) : nullptr;
    a.count = numel / per_vec / world;
    a.first = a.count * rank;
    a.timeout_ns = d->timeout_ns;
    a.mul = apply_mul ? static_cast<float>(1.0 / divisor) : 1.0f;
    a.apply_mul = apply_mul != 0;
    a.check = flag_offset >= 0;
    a.rank = rank;
    a.world = world;
    a.epoch = 2u * seq;

    // Every rank must launch the same number of CTAs (CTA c pairs with CTA c of the peers): the grid
    // depends only on arguments that are identical on every rank.
    if (ctas <= 0) ctas = use_multicast ? 32 : 148;
    const int64_t need = (a.count + kArThreads - 1) / kArThreads;
    const int grid = int(need < ctas ? need : ctas);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (use_multicast) {
        if (dtype == 0) nvls_allreduce_kernel<true><<<grid, kArThreads, 0, s>>>(a);
        else nvls_allreduce_kernel<false><<<grid, kArThreads, 0, s>>>(a);
    } else {
        if (dtype == 0) launch_p2p<true>(a, grid, s);
        else launch_p2p<false>(a, grid, s);
    }
    cudaError_t e = cudaPeekAtLastError();
    return e == cudaSuccess ? 0 : int(e);
}
