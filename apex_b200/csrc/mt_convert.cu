// mt_convert.cu -- fp16<->fp32 copies, zeroing, and the bucket gather/scatter.
//
// Reference behaviour being replaced (one torch launch per tensor each):
//   model_grads_to_master_grads    apex/fp16_utils/fp16util.py:93-112
//   master_params_to_model_params  apex/fp16_utils/fp16util.py:115-129
//   FP16_Optimizer.zero_grad       apex/fp16_utils/fp16_optimizer.py:187-211
//   apply_flat_dist_call           apex/parallel/distributed.py:11-23
#include "mt_common.cuh"

namespace apex_b200 {

// ---- g16 -> g32 (x inv_scale) ----------------------------------------------

__global__ void __launch_bounds__(kThreads)
mt_grads_to_master_kernel(MtView v, int src_list, int dst_list, float inv_scale,
                          int apply_scale, const int32_t* __restrict__ flag,
                          const float* __restrict__ clip_norm, float clip_max_norm) {
    if (flag != nullptr && *flag != 0) return;
    if (!apply_scale) inv_scale = 1.0f;       // exact widening unless the caller asked for the unscale
    // optional second factor: clip_grad_norm_'s clamp(max_norm / (norm + 1e-6), max=1)
    float coef = 1.0f;
    if (clip_norm != nullptr) {
        coef = __fdiv_rn(clip_max_norm, __fadd_rn(__ldg(clip_norm), 1e-6f));
        coef = coef > 1.0f ? 1.0f : coef;
        apply_scale = 1;                      // (x * 1.0f) * coef from here on
    }
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        const __half* s = mt_ptr<const __half>(v, src_list, c.tensor) + c.offset;
        float* d = mt_ptr<float>(v, dst_list, c.tensor) + c.offset;
        if (!(c.flags & APEX_B200_TF_UNALIGNED)) {
            const int nvec = c.len >> 3;
            int i = threadIdx.x;
            for (; i + 3 * kThreads < nvec; i += 4 * kThreads) {
                H8 h[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) h[u] = ld_h8(s + size_t(i + u * kThreads) * kVec);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    F8 f = h8_to_f8(h[u]);
                    if (apply_scale) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) f.v[k] = __fmul_rn(__fmul_rn(f.v[k], inv_scale), coef);
                    }
                    st_f8(d + size_t(i + u * kThreads) * kVec, f);
                }
            }
            for (; i < nvec; i += kThreads) {
                F8 f = h8_to_f8(ld_h8(s + size_t(i) * kVec));
                if (apply_scale) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) f.v[k] = __fmul_rn(__fmul_rn(f.v[k], inv_scale), coef);
                }
                st_f8(d + size_t(i) * kVec, f);
            }
            const int t = (nvec << 3) + threadIdx.x;
            if (t < c.len) {
                float x = __half2float(s[t]);
                d[t] = apply_scale ? __fmul_rn(__fmul_rn(x, inv_scale), coef) : x;
            }
        } else {
            for (int i = threadIdx.x; i < c.len; i += kThreads) {
                float x = __half2float(s[i]);
                d[i] = apply_scale ? __fmul_rn(__fmul_rn(x, inv_scale), coef) : x;
            }
        }
    }
}

// ---- p32 -> p16 (RNE) -------------------------------------------------------

__global__ void __launch_bounds__(kThreads)
mt_master_to_model_kernel(MtView v, int src_list, int dst_list, const int32_t* __restrict__ flag) {
    if (flag != nullptr && *flag != 0) return;
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        const float* s = mt_ptr<const float>(v, src_list, c.tensor) + c.offset;
        __half* d = mt_ptr<__half>(v, dst_list, c.tensor) + c.offset;
        if (!(c.flags & APEX_B200_TF_UNALIGNED)) {
            const int nvec = c.len >> 3;
            int i = threadIdx.x;
            for (; i + 3 * kThreads < nvec; i += 4 * kThreads) {
                F8 f[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) f[u] = ld_f8(s + size_t(i + u * kThreads) * kVec);
#pragma unroll
                for (int u = 0; u < 4; ++u) st_h8(d + size_t(i + u * kThreads) * kVec, f8_to_h8(f[u]));
            }
            for (; i < nvec; i += kThreads)
                st_h8(d + size_t(i) * kVec, f8_to_h8(ld_f8(s + size_t(i) * kVec)));
            const int t = (nvec << 3) + threadIdx.x;
            if (t < c.len) d[t] = __float2half_rn(s[t]);
        } else {
            for (int i = threadIdx.x; i < c.len; i += kThreads) d[i] = __float2half_rn(s[i]);
        }
    }
}

// ---- byte-wise helpers for zero / flatten / scatter ------------------------
// These kernels treat tensors as bytes; `es` is the element size.  The aligned
// path moves 16 B per thread per step.

__device__ __forceinline__ uint4 ld_b16(const void* p) {
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_b16(void* p, const uint4& r) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w) : "memory");
}

__global__ void __launch_bounds__(kThreads)
mt_zero_kernel(MtView v, int list, int elem_size) {
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        const int es = (c.flags & APEX_B200_TF_GRAD_F32) ? 4 : elem_size;
        char* d = mt_ptr<char>(v, list, c.tensor) + c.offset * es;
        const size_t bytes = size_t(c.len) * es;
        if (!(c.flags & APEX_B200_TF_UNALIGNED)) {
            const size_t nvec = bytes >> 4;
            const uint4 z = make_uint4(0, 0, 0, 0);
            for (size_t i = threadIdx.x; i < nvec; i += kThreads) st_b16(d + (i << 4), z);
            for (size_t t = (nvec << 4) + threadIdx.x; t < bytes; t += kThreads) d[t] = 0;
        } else {
            for (size_t t = threadIdx.x; t < bytes; t += kThreads) d[t] = 0;
        }
    }
}

// Copy `len` elements of size `es` from s to d with 16 B vectors when both ends allow.
__device__ __forceinline__ void copy_span(const char* s, char* d, int len, int es, bool aligned) {
    const size_t bytes = size_t(len) * es;
    if (aligned) {
        const size_t nvec = bytes >> 4;
        size_t i = threadIdx.x;
        for (; i + 3 * kThreads < nvec; i += 4 * kThreads) {
            uint4 r[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) r[u] = ld_b16(s + ((i + u * kThreads) << 4));
#pragma unroll
            for (int u = 0; u < 4; ++u) st_b16(d + ((i + u * kThreads) << 4), r[u]);
        }
        for (; i < nvec; i += kThreads) st_b16(d + (i << 4), ld_b16(s + (i << 4)));
        for (size_t t = (nvec << 4) + threadIdx.x; t < bytes; t += kThreads) d[t] = s[t];
    } else if (es == 2) {
        const uint16_t* s2 = reinterpret_cast<const uint16_t*>(s);
        uint16_t* d2 = reinterpret_cast<uint16_t*>(d);
        for (int i = threadIdx.x; i < len; i += kThreads) d2[i] = s2[i];
    } else if (es == 4) {
        const uint32_t* s4 = reinterpret_cast<const uint32_t*>(s);
        uint32_t* d4 = reinterpret_cast<uint32_t*>(d);
        for (int i = threadIdx.x; i < len; i += kThreads) d4[i] = s4[i];
    } else {
        const uint64_t* s8 = reinterpret_cast<const uint64_t*>(s);
        uint64_t* d8 = reinterpret_cast<uint64_t*>(d);
        for (int i = threadIdx.x; i < len; i += kThreads) d8[i] = s8[i];
    }
}

// tensors -> flat (gather).  The scatter direction has its own kernel because
// it also scales and checks.
__global__ void __launch_bounds__(kThreads)
mt_flatten_kernel(MtView v, int src_list, char* __restrict__ flat, int es) {
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        const char* s = mt_ptr<const char>(v, src_list, c.tensor) + c.offset * es;
        char* d = flat + (__ldg(v.elem_start + c.tensor) + c.offset) * es;
        const bool aligned = !(c.flags & APEX_B200_TF_UNALIGNED) && ((reinterpret_cast<uintptr_t>(d) & 15) == 0);
        copy_span(s, d, c.len, es, aligned);
    }
}

// tensor i of src_list -> tensor i of dst_list, same dtype.
__global__ void __launch_bounds__(kThreads)
mt_copy_kernel(MtView v, int src_list, int dst_list, int es, const int32_t* __restrict__ flag) {
    if (flag != nullptr && *flag != 0) return;
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        const char* s = mt_ptr<const char>(v, src_list, c.tensor) + c.offset * es;
        char* d = mt_ptr<char>(v, dst_list, c.tensor) + c.offset * es;
        copy_span(s, d, c.len, es, !(c.flags & APEX_B200_TF_UNALIGNED));
    }
}

// ---- flat -> tensors with x(1/world), inf/NaN check -------------------------
// torch's CUDA true-divide by a python scalar multiplies by the reciprocal
// (computed in double, narrowed to the op's math type); the same is done here so
// that the averaged gradients equal the reference's on the same reduced sums.

template <typename T> struct ScaleOps;
template <> struct ScaleOps<__half> {
    using M = float;
    static __device__ __forceinline__ __half run(__half x, float m, int apply, bool& bad) {
        float f = __half2float(x);
        if (apply) f = __fmul_rn(f, m);
        const __half r = __float2half_rn(f);
        bad |= (__half_as_ushort(r) & 0x7c00u) == 0x7c00u;
        return r;
    }
};
template <> struct ScaleOps<float> {
    using M = float;
    static __device__ __forceinline__ float run(float x, float m, int apply, bool& bad) {
        if (apply) x = __fmul_rn(x, m);
        bad |= is_nonfinite(x);
        return x;
    }
};
template <> struct ScaleOps<double> {
    using M = double;
    static __device__ __forceinline__ double run(double x, double m, int apply, bool& bad) {
        if (apply) x = __dmul_rn(x, m);
        bad |= !isfinite(x);
        return x;
    }
};

template <typename T, int N>
struct alignas(16) Pack { T v[N]; };

template <typename T>
__device__ __forceinline__ bool scale_span(const T* s, T* d, int len,
                                           typename ScaleOps<T>::M m, int apply, bool vec_ok) {
    constexpr int N = 16 / sizeof(T);
    using P = Pack<T, N>;
    bool bad = false;
    if (vec_ok) {
        const int nvec = len / N;
        int i = threadIdx.x;
        for (; i + 3 * kThreads < nvec; i += 4 * kThreads) {
            uint4 r[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) r[u] = ld_b16(reinterpret_cast<const P*>(s) + i + u * kThreads);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                P p = *reinterpret_cast<P*>(&r[u]);
#pragma unroll
                for (int k = 0; k < N; ++k) p.v[k] = ScaleOps<T>::run(p.v[k], m, apply, bad);
                st_b16(reinterpret_cast<P*>(d) + i + u * kThreads, *reinterpret_cast<uint4*>(&p));
            }
        }
        for (; i < nvec; i += kThreads) {
            uint4 r = ld_b16(reinterpret_cast<const P*>(s) + i);
            P p = *reinterpret_cast<P*>(&r);
#pragma unroll
            for (int k = 0; k < N; ++k) p.v[k] = ScaleOps<T>::run(p.v[k], m, apply, bad);
            st_b16(reinterpret_cast<P*>(d) + i, *reinterpret_cast<uint4*>(&p));
        }
        for (int t = nvec * N + threadIdx.x; t < len; t += kThreads) d[t] = ScaleOps<T>::run(s[t], m, apply, bad);
    } else {
        for (int i = threadIdx.x; i < len; i += kThreads) d[i] = ScaleOps<T>::run(s[i], m, apply, bad);
    }
    return bad;
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
mt_scale_unflatten_kernel(MtView v, const T* __restrict__ flat, int dst_list,
                          typename ScaleOps<T>::M m, int apply, int32_t* __restrict__ flag) {
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        const T* s = flat + __ldg(v.elem_start + c.tensor) + c.offset;
        T* d = mt_ptr<T>(v, dst_list, c.tensor) + c.offset;
        const bool vec_ok = !(c.flags & APEX_B200_TF_UNALIGNED) && ((reinterpret_cast<uintptr_t>(s) & 15) == 0);
        const bool bad = scale_span<T>(s, d, c.len, m, apply, vec_ok);
        if (flag != nullptr) mt_raise_flag(bad, c.tensor, flag);
    }
}

// One contiguous buffer, in place.  Chunks of 16384 elements, grid-stride.
template <typename T>
__global__ void __launch_bounds__(kThreads)
scale_flat_kernel(T* flat, int64_t numel, typename ScaleOps<T>::M m, int32_t* __restrict__ flag) {
    constexpr int kChunk = 16384;
    const int64_t nchunks = (numel + kChunk - 1) / kChunk;
    const bool vec_ok = (reinterpret_cast<uintptr_t>(flat) & 15) == 0;
    bool bad = false;
    for (int64_t chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x) {
        const int64_t off = chunk * kChunk;
        const int len = int(numel - off < kChunk ? numel - off : kChunk);
        bad |= scale_span<T>(flat + off, flat + off, len, m, 1, vec_ok);
    }
    if (flag != nullptr) mt_raise_flag(bad, 0, flag);
}

}  // namespace apex_b200

using namespace apex_b200;

#define APEX_CHECK_TABLE(t)                                              \
    do {                                                                 \
        if ((t) == nullptr) return APEX_B200_EINVAL;                     \
        if ((t)->total_chunks == 0) return 0;                            \
        if ((t)->d_table == nullptr) return APEX_B200_EINVAL;            \
    } while (0)
#define APEX_CHECK_LIST(t, l) \
    do { if ((l) < 0 || (l) >= (t)->n_lists) return APEX_B200_EINVAL; } while (0)

extern "C" int apex_b200_mt_grads_to_master(const apex_b200_mt_table* t, int src_list, int dst_list,
                                            double inv_scale, int apply_scale,
                                            const int32_t* d_flag, const float* d_clip_norm,
                                            double clip_max_norm, void* stream) {
    if (t == nullptr) return APEX_B200_EINVAL;
    APEX_CHECK_LIST(t, src_list);
    APEX_CHECK_LIST(t, dst_list);
    APEX_CHECK_TABLE(t);
    mt_grads_to_master_kernel<<<mt_grid(t->total_chunks, 8), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        mt_view(t), src_list, dst_list, static_cast<float>(inv_scale), apply_scale, d_flag, d_clip_norm,
        static_cast<float>(clip_max_norm));
    return mt_cuda_status();
}

extern "C" int apex_b200_mt_master_to_model(const apex_b200_mt_table* t, int src_list, int dst_list,
                                            const int32_t* d_flag, void* stream) {
    if (t == nullptr) return APEX_B200_EINVAL;
    APEX_CHECK_LIST(t, src_list);
    APEX_CHECK_LIST(t, dst_list);
    APEX_CHECK_TABLE(t);
    mt_master_to_model_kernel<<<mt_grid(t->total_chunks, 8), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        mt_view(t), src_list, dst_list, d_flag);
    return mt_cuda_status();
}

extern "C" int apex_b200_mt_zero(const apex_b200_mt_table* t, int list, int elem_size, void* stream) {
    if (t == nullptr || (elem_size != 2 && elem_size != 4)) return APEX_B200_EINVAL;
    APEX_CHECK_LIST(t, list);
    APEX_CHECK_TABLE(t);
    mt_zero_kernel<<<mt_grid(t->total_chunks, 8), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        mt_view(t), list, elem_size);
    return mt_cuda_status();
}

extern "C" int apex_b200_mt_flatten(const apex_b200_mt_table* t, int src_list, void* flat,
                                    int elem_size, void* stream) {
    if (t == nullptr || (elem_size != 2 && elem_size != 4 && elem_size != 8)) return APEX_B200_EINVAL;
    APEX_CHECK_LIST(t, src_list);
    APEX_CHECK_TABLE(t);
    if (flat == nullptr) return APEX_B200_EINVAL;
    mt_flatten_kernel<<<mt_grid(t->total_chunks, 8), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        mt_view(t), src_list, static_cast<char*>(flat), elem_size);
    return mt_cuda_status();
}

extern "C" int apex_b200_mt_copy(const apex_b200_mt_table* t, int src_list, int dst_list, int elem_size,
                                 const int32_t* d_flag, void* stream) {
    if (t == nullptr || (elem_size != 2 && elem_size != 4 && elem_size != 8)) return APEX_B200_EINVAL;
    APEX_CHECK_LIST(t, src_list);
    APEX_CHECK_LIST(t, dst_list);
    APEX_CHECK_TABLE(t);
    mt_copy_kernel<<<mt_grid(t->total_chunks, 8), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        mt_view(t), src_list, dst_list, elem_size, d_flag);
    return mt_cuda_status();
}

extern "C" int apex_b200_mt_scale_unflatten(const apex_b200_mt_table* t, const void* flat, int dst_list,
                                            int dtype, double divisor, int apply_mul,
                                            int32_t* d_flag, void* stream) {
    if (t == nullptr || dtype < 0 || dtype > 2) return APEX_B200_EINVAL;
    APEX_CHECK_LIST(t, dst_list);
    APEX_CHECK_TABLE(t);
    if (flat == nullptr || (apply_mul && divisor == 0.0)) return APEX_B200_EINVAL;
    const double inv = apply_mul ? 1.0 / divisor : 1.0;
    const int grid = mt_grid(t->total_chunks, 8);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (dtype == 0)
        mt_scale_unflatten_kernel<__half><<<grid, kThreads, 0, s>>>(mt_view(t), static_cast<const __half*>(flat), dst_list, float(inv), apply_mul, d_flag);
    else if (dtype == 1)
        mt_scale_unflatten_kernel<float><<<grid, kThreads, 0, s>>>(mt_view(t), static_cast<const float*>(flat), dst_list, float(inv), apply_mul, d_flag);
    else
        mt_scale_unflatten_kernel<double><<<grid, kThreads, 0, s>>>(mt_view(t), static_cast<const double*>(flat), dst_list, inv, apply_mul, d_flag);
    return mt_cuda_status();
}

extern "C" int apex_b200_scale_flat(void* flat, int64_t numel, int dtype, double divisor,
                                    int32_t* d_flag, void* stream) {
    if (dtype < 0 || dtype > 2 || numel < 0 || divisor == 0.0) return APEX_B200_EINVAL;
    if (numel == 0) return 0;
    if (flat == nullptr) return APEX_B200_EINVAL;
    const double inv = 1.0 / divisor;
    const int64_t nchunks = (numel + 16383) / 16384;
    const int grid = mt_grid(int(nchunks < 0x7fffffff ? nchunks : 0x7fffffff), 8);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (dtype == 0)
        scale_flat_kernel<__half><<<grid, kThreads, 0, s>>>(static_cast<__half*>(flat), numel, float(inv), d_flag);
    else if (dtype == 1)
        scale_flat_kernel<float><<<grid, kThreads, 0, s>>>(static_cast<float*>(flat), numel, float(inv), d_flag);
    else
        scale_flat_kernel<double><<<grid, kThreads, 0, s>>>(static_cast<double*>(flat), numel, inv, d_flag);
    return mt_cuda_status();
}
