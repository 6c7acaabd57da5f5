// mt_larc.cu -- LARC (layer-wise adaptive rate) without a host round trip per tensor.
//
// Reference behaviour being replaced (apex/parallel/LARC.py:68-97), per parameter tensor:
//   param_norm = torch.norm(p); grad_norm = torch.norm(p.grad)            2 reductions
//   if param_norm != 0 and grad_norm != 0:                                2 blocking device->host reads
//       adaptive_lr = trust * param_norm / (grad_norm + param_norm * wd + eps)
//       if clip: adaptive_lr = min(adaptive_lr / lr, 1)
//       p.grad += wd * p ; p.grad *= adaptive_lr                          2-3 elementwise launches
// Here: per-chunk sums of squares of p and g (one launch each), one launch that turns them into the
// two norms and the adaptive rate of every tensor (fixed summation order), one launch that rewrites
// all gradients.  Nothing is read back by the host.
#include "mt_common.cuh"

namespace apex_b200 {

__device__ __forceinline__ float larc_block_sum(float x, float* smem) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) smem[w] = x;
    __syncthreads();
    float t = (threadIdx.x < kThreads / 32) ? smem[threadIdx.x] : 0.f;
    if (w == 0) {
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    }
    return t;
}

template <typename T>
__device__ __forceinline__ float sumsq_span(const T* __restrict__ x, const MtChunk& c) {
    float acc = 0.f;
    x += c.offset;
    if (!(c.flags & APEX_B200_TF_UNALIGNED)) {
        const int nvec = c.len >> 3;
        for (int i = threadIdx.x; i < nvec; i += kThreads) {
            const F8 f = GradIo<T>::ld8(x + size_t(i) * kVec);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc = __fmaf_rn(f.v[k], f.v[k], acc);
        }
        const int t = (nvec << 3) + threadIdx.x;
        if (t < c.len) { const float v = GradIo<T>::ld1(x + t); acc = __fmaf_rn(v, v, acc); }
    } else {
        for (int i = threadIdx.x; i < c.len; i += kThreads) { const float v = GradIo<T>::ld1(x + i); acc = __fmaf_rn(v, v, acc); }
    }
    return acc;
}

// partials[chunk] = sum of squares of list `list` over that chunk (fp16 or fp32 per TF_GRAD_F32)
__global__ void __launch_bounds__(kThreads)
mt_chunk_sumsq_kernel(MtView v, int list, float* __restrict__ partials) {
    __shared__ float smem[kThreads / 32];
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        const float acc = (c.flags & APEX_B200_TF_GRAD_F32)
            ? sumsq_span<float>(mt_ptr<const float>(v, list, c.tensor), c)
            : sumsq_span<__half>(mt_ptr<const __half>(v, list, c.tensor), c);
        const float total = larc_block_sum(acc, smem);
        if (threadIdx.x == 0) partials[chunk] = total;
    }
}

// One warp per tensor: norms from the chunk partials (index order, fp64), then the rate.
// rate[t] is the factor of LARC.py:88-95; 1 and "no weight decay" (flag in apply) when a norm is 0.
__global__ void __launch_bounds__(kThreads)
larc_rate_kernel(MtView v, const float* __restrict__ g_part, const float* __restrict__ p_part,
                 float trust, float eps, float weight_decay, float lr, int clip,
                 float* __restrict__ rate, float* __restrict__ norms) {
    const int t = blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
    if (t >= v.n) return;
    const int lane = threadIdx.x & 31;
    const int c0 = __ldg(v.chunk_start + t), c1 = __ldg(v.chunk_start + t + 1);
    double gs = 0.0, ps = 0.0;
    for (int c = c0 + lane; c < c1; c += 32) { gs += double(g_part[c]); ps += double(p_part[c]); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { gs += __shfl_xor_sync(0xffffffffu, gs, o); ps += __shfl_xor_sync(0xffffffffu, ps, o); }
    if (lane == 0) {
        const float gn = float(sqrt(gs)), pn = float(sqrt(ps));
        float r = -1.0f;                                  // "leave this gradient alone"
        if (pn != 0.f && gn != 0.f) {
            // trust * pn / (gn + pn * wd + eps), every step rounded to fp32 like the 0-dim tensor ops
            r = __fdiv_rn(__fmul_rn(trust, pn), __fadd_rn(__fadd_rn(gn, __fmul_rn(pn, weight_decay)), eps));
            if (clip) { r = __fdiv_rn(r, lr); r = r < 1.0f ? r : 1.0f; }
        }
        rate[t] = r;
        if (norms != nullptr) { norms[2 * t] = pn; norms[2 * t + 1] = gn; }
    }
}

// g = (g + wd * p) * rate[t]   (two roundings for the weight decay: `weight_decay * p.data` is a tensor)
template <typename T>
__device__ __forceinline__ void larc_apply_span(T* g, const T* p, const MtChunk& c, float wd, float r);

template <>
__device__ __forceinline__ void larc_apply_span<float>(float* g, const float* p, const MtChunk& c, float wd, float r) {
    g += c.offset; p += c.offset;
    if (!(c.flags & APEX_B200_TF_UNALIGNED)) {
        const int nvec = c.len >> 3;
        for (int i = threadIdx.x; i < nvec; i += kThreads) {
            F8 a = ld_f8(g + size_t(i) * kVec);
            const F8 b = ld_f8(p + size_t(i) * kVec);
#pragma unroll
            for (int k = 0; k < 8; ++k) a.v[k] = __fmul_rn(__fadd_rn(a.v[k], __fmul_rn(wd, b.v[k])), r);
            st_f8(g + size_t(i) * kVec, a);
        }
        const int t = (nvec << 3) + threadIdx.x;
        if (t < c.len) g[t] = __fmul_rn(__fadd_rn(g[t], __fmul_rn(wd, p[t])), r);
    } else {
        for (int i = threadIdx.x; i < c.len; i += kThreads) g[i] = __fmul_rn(__fadd_rn(g[i], __fmul_rn(wd, p[i])), r);
    }
}
template <>
__device__ __forceinline__ void larc_apply_span<__half>(__half* g, const __half* p, const MtChunk& c, float wd, float r) {
    // fp16 tensors: each torch op rounds its result to fp16
    g += c.offset; p += c.offset;
    for (int i = threadIdx.x; i < c.len; i += kThreads) {
        const __half wp = __float2half_rn(__fmul_rn(wd, __half2float(p[i])));
        const __half s = __float2half_rn(__fadd_rn(__half2float(g[i]), __half2float(wp)));
        g[i] = __float2half_rn(__fmul_rn(__half2float(s), r));
    }
}

__global__ void __launch_bounds__(kThreads)
mt_larc_apply_kernel(MtView v, int g_list, int p_list, const float* __restrict__ rate, float weight_decay) {
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        const float r = __ldg(rate + c.tensor);
        if (r < 0.f) continue;                           // a zero norm: the reference skips the tensor
        if (c.flags & APEX_B200_TF_GRAD_F32)
            larc_apply_span<float>(mt_ptr<float>(v, g_list, c.tensor), mt_ptr<const float>(v, p_list, c.tensor), c, weight_decay, r);
        else
            larc_apply_span<__half>(mt_ptr<__half>(v, g_list, c.tensor), mt_ptr<const __half>(v, p_list, c.tensor), c, weight_decay, r);
    }
}

}  // namespace apex_b200

using namespace apex_b200;

extern "C" int apex_b200_mt_larc(const apex_b200_mt_table* t, int grad_list, int param_list,
                                 double trust_coefficient, double eps, double weight_decay, double lr, int clip,
                                 float* d_scratch, int64_t scratch_len, float* d_norms, void* stream) {
    if (t == nullptr || grad_list < 0 || grad_list >= t->n_lists || param_list < 0 || param_list >= t->n_lists)
        return APEX_B200_EINVAL;
    if (t->n_tensors == 0) return 0;
    if (t->d_table == nullptr || d_scratch == nullptr) return APEX_B200_EINVAL;
    const int64_t need = 2 * int64_t(t->total_chunks) + t->n_tensors;
    if (scratch_len < need) return APEX_B200_ENOSPC;
    float* g_part = d_scratch;
    float* p_part = d_scratch + t->total_chunks;
    float* rate = d_scratch + 2 * int64_t(t->total_chunks);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const MtView v = mt_view(t);
    if (t->total_chunks > 0) {
        mt_chunk_sumsq_kernel<<<t->total_chunks, kThreads, 0, s>>>(v, grad_list, g_part);
        mt_chunk_sumsq_kernel<<<t->total_chunks, kThreads, 0, s>>>(v, param_list, p_part);
    }
    const int warps = kThreads / 32;
    larc_rate_kernel<<<(t->n_tensors + warps - 1) / warps, kThreads, 0, s>>>(
        v, g_part, p_part, float(trust_coefficient), float(eps), float(weight_decay), float(lr), clip, rate, d_norms);
    if (t->total_chunks > 0)
        mt_larc_apply_kernel<<<t->total_chunks, kThreads, 0, s>>>(v, grad_list, param_list, rate, float(weight_decay));
    return mt_cuda_status();
}
