// mt_overflow.cu -- inf/NaN detection and in-place unscale over a tensor table.
//
// Reference behaviour being replaced (per tensor, with a host sync each):
//   DynamicLossScaler._has_inf_or_nan   apex/fp16_utils/loss_scaler.py:92-110
//   scale_check_overflow                apex/amp/scaler.py:6-18
//   FP16_Optimizer._downscale_master    apex/fp16_utils/fp16_optimizer.py:236-241
// The reference decides by `float(x.float().sum())` being inf/NaN.  For fp16
// inputs that is equivalent to "some element is inf/NaN" (the fp32 sum of
// finite halfs cannot overflow); for fp32 inputs it is equivalent whenever
// sum|x| < FLT_MAX ("overflow decision").  Here: one launch, one
// device word, no host sync.
#include "mt_common.cuh"

namespace apex_b200 {

constexpr int kCheckUnroll = 4;

// acc = x * 0 + acc: stays 0 for finite x, turns NaN for inf/NaN and stays NaN.
__device__ __forceinline__ void poison_h8(const H8& h, __half2& acc) {
    const __half2 z = __float2half2_rn(0.f);
#pragma unroll
    for (int i = 0; i < 4; ++i)
        acc = __hfma2(*reinterpret_cast<const __half2*>(&h.w[i]), z, acc);
}
__device__ __forceinline__ void poison_f8(const F8& f, float& acc) {
#pragma unroll
    for (int i = 0; i < 8; ++i) acc = __fmaf_rn(f.v[i], 0.f, acc);
}

template <typename G> struct Poison;
template <> struct Poison<__half> {
    using Acc = __half2;
    using Vec = H8;
    static __device__ __forceinline__ Acc zero() { return __float2half2_rn(0.f); }
    static __device__ __forceinline__ Vec ld(const __half* p) { return ld_h8(p); }
    static __device__ __forceinline__ void add(const Vec& v, Acc& a) { poison_h8(v, a); }
    static __device__ __forceinline__ void add1(const __half* p, Acc& a) {
        a = __hfma2(__half2half2(*p), __float2half2_rn(0.f), a);
    }
    static __device__ __forceinline__ bool bad(const Acc& a) {
        const float2 f = __half22float2(a);
        return f.x != f.x || f.y != f.y;
    }
};
template <> struct Poison<float> {
    using Acc = float;
    using Vec = F8;
    static __device__ __forceinline__ Acc zero() { return 0.f; }
    static __device__ __forceinline__ Vec ld(const float* p) { return ld_f8(p); }
    static __device__ __forceinline__ void add(const Vec& v, Acc& a) { poison_f8(v, a); }
    static __device__ __forceinline__ void add1(const float* p, Acc& a) { a = __fmaf_rn(*p, 0.f, a); }
    static __device__ __forceinline__ bool bad(const Acc& a) { return a != a; }
};

template <typename G>
__device__ __forceinline__ bool check_chunk(const G* __restrict__ g, const MtChunk& c) {
    using P = Poison<G>;
    typename P::Acc acc = P::zero();
    g += c.offset;
    if (!(c.flags & APEX_B200_TF_UNALIGNED)) {
        const int nvec = c.len >> 3;
        int i = threadIdx.x;
        for (; i + (kCheckUnroll - 1) * kThreads < nvec; i += kCheckUnroll * kThreads) {
            typename P::Vec v[kCheckUnroll];
#pragma unroll
            for (int u = 0; u < kCheckUnroll; ++u) v[u] = P::ld(g + size_t(i + u * kThreads) * kVec);
#pragma unroll
            for (int u = 0; u < kCheckUnroll; ++u) P::add(v[u], acc);
        }
        for (; i < nvec; i += kThreads) {
            typename P::Vec v = P::ld(g + size_t(i) * kVec);
            P::add(v, acc);
        }
        const int t = (nvec << 3) + threadIdx.x;
        if (t < c.len) P::add1(g + t, acc);
    } else {
        for (int i = threadIdx.x; i < c.len; i += kThreads) P::add1(g + i, acc);
    }
    return P::bad(acc);
}

// `post` (optional): what the LAST CTA to finish does once the word is final -- publish it to
// pinned host memory (so that no device-to-host copy sits on the stream between this kernel and
// the update it predicates) and re-arm the word the NEXT check will use (so that no memset does).
struct CheckPost {
    int32_t* ticket;      // device counter, zero between launches
    int32_t* host_mirror; // mapped pinned host word, or nullptr
    int32_t* rearm;       // another flag word to set to zero, or nullptr
};

__global__ void __launch_bounds__(kThreads)
mt_check_overflow_kernel(MtView v, int grad_list, int32_t* __restrict__ flag, CheckPost post) {
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        bool bad;
        if (c.flags & APEX_B200_TF_GRAD_F32)
            bad = check_chunk<float>(mt_ptr<const float>(v, grad_list, c.tensor), c);
        else
            bad = check_chunk<__half>(mt_ptr<const __half>(v, grad_list, c.tensor), c);
        mt_raise_flag(bad, c.tensor, flag);       // (ends with a block-wide barrier + thread 0's atomic)
    }
    if (post.ticket != nullptr && threadIdx.x == 0) {
        __threadfence();                                          // my atomicMax before my ticket
        if (atomicAdd(post.ticket, 1) == int(gridDim.x) - 1) {    // every other CTA is done
            __threadfence();
            const int32_t final_word = atomicMax(flag, 0);
            if (post.host_mirror != nullptr) {
                *reinterpret_cast<volatile int32_t*>(post.host_mirror) = final_word;
                __threadfence_system();
            }
            if (post.rearm != nullptr) *post.rearm = 0;
            *post.ticket = 0;
        }
    }
}

// mode 0: skip everything when the flag is set; mode 1: only tensors before
// the first bad one are scaled (amp's partial mutation, apex/amp/scaler.py:35-40).
__global__ void __launch_bounds__(kThreads)
mt_unscale_kernel(MtView v, int grad_list, float inv_scale,
                  const int32_t* __restrict__ flag, int mode) {
    int first_bad = 0x7fffffff;
    if (flag != nullptr) {
        const int32_t f = *flag;
        if (f != 0) {
            if (mode == 0) return;
            first_bad = APEX_B200_FLAG_FIRST_BAD(f);
        }
    }
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        if (c.tensor >= first_bad) continue;
        if (c.flags & APEX_B200_TF_GRAD_F32) {
            float* g = mt_ptr<float>(v, grad_list, c.tensor) + c.offset;
            if (!(c.flags & APEX_B200_TF_UNALIGNED)) {
                const int nvec = c.len >> 3;
                int i = threadIdx.x;
                for (; i + kThreads < nvec; i += 2 * kThreads) {
                    F8 a = ld_f8(g + size_t(i) * kVec);
                    F8 b = ld_f8(g + size_t(i + kThreads) * kVec);
#pragma unroll
                    for (int k = 0; k < 8; ++k) { a.v[k] = __fmul_rn(a.v[k], inv_scale); b.v[k] = __fmul_rn(b.v[k], inv_scale); }
                    st_f8(g + size_t(i) * kVec, a);
                    st_f8(g + size_t(i + kThreads) * kVec, b);
                }
                for (; i < nvec; i += kThreads) {
                    F8 a = ld_f8(g + size_t(i) * kVec);
#pragma unroll
                    for (int k = 0; k < 8; ++k) a.v[k] = __fmul_rn(a.v[k], inv_scale);
                    st_f8(g + size_t(i) * kVec, a);
                }
                const int t = (nvec << 3) + threadIdx.x;
                if (t < c.len) g[t] = __fmul_rn(g[t], inv_scale);
            } else {
                for (int i = threadIdx.x; i < c.len; i += kThreads) g[i] = __fmul_rn(g[i], inv_scale);
            }
        } else {
            // fp16 in place: torch's mul_ computes in fp32 and rounds to nearest even.
            __half* g = mt_ptr<__half>(v, grad_list, c.tensor) + c.offset;
            if (!(c.flags & APEX_B200_TF_UNALIGNED)) {
                const int nvec = c.len >> 3;
                for (int i = threadIdx.x; i < nvec; i += kThreads) {
                    F8 a = h8_to_f8(ld_h8(g + size_t(i) * kVec));
#pragma unroll
                    for (int k = 0; k < 8; ++k) a.v[k] = __fmul_rn(a.v[k], inv_scale);
                    st_h8(g + size_t(i) * kVec, f8_to_h8(a));
                }
                const int t = (nvec << 3) + threadIdx.x;
                if (t < c.len) g[t] = __float2half_rn(__fmul_rn(__half2float(g[t]), inv_scale));
            } else {
                for (int i = threadIdx.x; i < c.len; i += kThreads)
                    g[i] = __float2half_rn(__fmul_rn(__half2float(g[i]), inv_scale));
            }
        }
    }
}

}  // namespace apex_b200

using namespace apex_b200;

extern "C" int apex_b200_flag_clear(int32_t* d_flag, void* stream) {
    if (d_flag == nullptr) return APEX_B200_EINVAL;
    cudaError_t e = cudaMemsetAsync(d_flag, 0, sizeof(int32_t), static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? 0 : int(e);
}

extern "C" int apex_b200_mt_check_overflow(const apex_b200_mt_table* t, int grad_list,
                                           int32_t* d_flag, void* stream) {
    return apex_b200_mt_check_overflow_publish(t, grad_list, d_flag, nullptr, nullptr, nullptr, stream);
}

extern "C" int apex_b200_mt_check_overflow_publish(const apex_b200_mt_table* t, int grad_list, int32_t* d_flag,
                                                   int32_t* d_ticket, int32_t* h_mirror, int32_t* d_rearm,
                                                   void* stream) {
    if (t == nullptr || d_flag == nullptr || grad_list < 0 || grad_list >= t->n_lists) return APEX_B200_EINVAL;
    if (d_ticket == nullptr && (h_mirror != nullptr || d_rearm != nullptr)) return APEX_B200_EINVAL;
    if (d_rearm == d_flag && d_flag != nullptr && d_rearm != nullptr) return APEX_B200_EINVAL;
    if (t->total_chunks == 0) {
        // nothing to scan, but the caller still expects the mirror and the re-armed word
        if (d_ticket == nullptr) return 0;
        CheckPost post{d_ticket, h_mirror, d_rearm};
        MtView empty{};
        mt_check_overflow_kernel<<<1, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(empty, 0, d_flag, post);
        return mt_cuda_status();
    }
    if (t->d_table == nullptr) return APEX_B200_EINVAL;
    CheckPost post{d_ticket, h_mirror, d_rearm};
    mt_check_overflow_kernel<<<mt_grid(t->total_chunks, 8), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        mt_view(t), grad_list, d_flag, post);
    return mt_cuda_status();
}

extern "C" int apex_b200_mt_unscale(const apex_b200_mt_table* t, int grad_list, double inv_scale,
                                    const int32_t* d_flag, int mode, void* stream) {
    if (t == nullptr || grad_list < 0 || grad_list >= t->n_lists || (mode != 0 && mode != 1)) return APEX_B200_EINVAL;
    if (t->total_chunks == 0) return 0;
    if (t->d_table == nullptr) return APEX_B200_EINVAL;
    mt_unscale_kernel<<<mt_grid(t->total_chunks, 8), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        mt_view(t), grad_list, static_cast<float>(inv_scale), d_flag, mode);
    return mt_cuda_status();
}
