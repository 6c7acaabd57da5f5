// mt_norm.cu -- global L2 norm of a tensor table and clip-by-norm.
//
// Reference behaviour being replaced:
//   FP16_Optimizer.clip_master_grads   apex/fp16_utils/fp16_optimizer.py:243-265
//     -> torch.nn.utils.clip_grad_norm_ (alias at apex/fp16_utils/fp16util.py:139-144)
// Two stages.  Stage 1: each chunk's sum of squares (fp32, fixed tree order)
// goes to partials[chunk] -- caller-provided scratch.  Stage 2: one CTA adds the
// partials in index order in fp64 and writes sqrt(total) as fp32.  The result
// does not depend on the grid size or on scheduling.
#include <atomic>

#include "mt_common.cuh"

namespace apex_b200 {

__device__ __forceinline__ float block_sum(float x, float* smem) {
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
    return t;  // valid in thread 0
}

template <typename G>
__device__ __forceinline__ float sumsq_chunk(const G* __restrict__ g, const MtChunk& c, float inv_scale, int apply) {
    float acc = 0.f;
    if (c.flags & APEX_B200_TF_NO_SCALE) apply = 0;
    g += c.offset;
    if (!(c.flags & APEX_B200_TF_UNALIGNED)) {
        const int nvec = c.len >> 3;
        int i = threadIdx.x;
        for (; i + 3 * kThreads < nvec; i += 4 * kThreads) {
            F8 f[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) f[u] = GradIo<G>::ld8(g + size_t(i + u * kThreads) * kVec);
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float x = apply ? __fmul_rn(f[u].v[k], inv_scale) : f[u].v[k];
                    acc = __fmaf_rn(x, x, acc);
                }
        }
        for (; i < nvec; i += kThreads) {
            const F8 f = GradIo<G>::ld8(g + size_t(i) * kVec);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const float x = apply ? __fmul_rn(f.v[k], inv_scale) : f.v[k];
                acc = __fmaf_rn(x, x, acc);
            }
        }
        const int t = (nvec << 3) + threadIdx.x;
        if (t < c.len) {
            float x = GradIo<G>::ld1(g + t);
            if (apply) x = __fmul_rn(x, inv_scale);
            acc = __fmaf_rn(x, x, acc);
        }
    } else {
        for (int i = threadIdx.x; i < c.len; i += kThreads) {
            float x = GradIo<G>::ld1(g + i);
            if (apply) x = __fmul_rn(x, inv_scale);
            acc = __fmaf_rn(x, x, acc);
        }
    }
    return acc;
}

__global__ void __launch_bounds__(kThreads)
mt_sumsq_kernel(MtView v, int grad_list, float inv_scale, int apply, float* __restrict__ partials) {
    __shared__ float smem[kThreads / 32];
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        float acc;
        if (c.flags & APEX_B200_TF_GRAD_F32)
            acc = sumsq_chunk<float>(mt_ptr<const float>(v, grad_list, c.tensor), c, inv_scale, apply);
        else
            acc = sumsq_chunk<__half>(mt_ptr<const __half>(v, grad_list, c.tensor), c, inv_scale, apply);
        const float total = block_sum(acc, smem);
        if (threadIdx.x == 0) partials[chunk] = total;
    }
}

// <<<1, 256>>>: fixed-order fp64 accumulation of the per-chunk partials.
__global__ void __launch_bounds__(kThreads)
norm_finish_kernel(const float* __restrict__ partials, int n, float* __restrict__ norm) {
    __shared__ double sm[kThreads];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += kThreads) acc += double(partials[i]);
    sm[threadIdx.x] = acc;
    __syncthreads();
    for (int o = kThreads / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) *norm = float(sqrt(sm[0]));
}

// torch.nn.utils.clip_grad_norm_: coef = max_norm / (total_norm + 1e-6),
// clamped to 1, multiplied into every gradient (torch/nn/utils/clip_grad.py).
__global__ void __launch_bounds__(kThreads)
mt_clip_kernel(MtView v, int grad_list, const float* __restrict__ norm, float max_norm,
               const int32_t* __restrict__ flag) {
    if (flag != nullptr && *flag != 0) return;
    const float coef = __fdiv_rn(max_norm, __fadd_rn(*norm, 1e-6f));
    if (coef >= 1.0f) return;     // clamp(max=1): multiplying by 1 changes nothing; a NaN coef still poisons, as in torch
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        float* g = mt_ptr<float>(v, grad_list, c.tensor) + c.offset;
        if (!(c.flags & APEX_B200_TF_UNALIGNED)) {
            const int nvec = c.len >> 3;
            for (int i = threadIdx.x; i < nvec; i += kThreads) {
                F8 a = ld_f8(g + size_t(i) * kVec);
#pragma unroll
                for (int k = 0; k < 8; ++k) a.v[k] = __fmul_rn(a.v[k], coef);
                st_f8(g + size_t(i) * kVec, a);
            }
            const int t = (nvec << 3) + threadIdx.x;
            if (t < c.len) g[t] = __fmul_rn(g[t], coef);
        } else {
            for (int i = threadIdx.x; i < c.len; i += kThreads) g[i] = __fmul_rn(g[i], coef);
        }
    }
}

}  // namespace apex_b200

using namespace apex_b200;

extern "C" int apex_b200_mt_l2norm(const apex_b200_mt_table* t, int grad_list, double inv_scale,
                                   int apply_scale, float* d_partials, int64_t partials_len,
                                   float* d_norm, void* stream) {
    if (t == nullptr || d_norm == nullptr || grad_list < 0 || grad_list >= t->n_lists) return APEX_B200_EINVAL;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (t->total_chunks > 0) {
        if (t->d_table == nullptr || d_partials == nullptr) return APEX_B200_EINVAL;
        if (partials_len < t->total_chunks) return APEX_B200_ENOSPC;
        mt_sumsq_kernel<<<mt_grid(t->total_chunks, 8), kThreads, 0, s>>>(
            mt_view(t), grad_list, static_cast<float>(inv_scale), apply_scale, d_partials);
    }
    norm_finish_kernel<<<1, kThreads, 0, s>>>(d_partials, t->total_chunks, d_norm);
    return mt_cuda_status();
}

extern "C" int apex_b200_mt_clip_by_norm(const apex_b200_mt_table* t, int grad_list, const float* d_norm,
                                         double max_norm, const int32_t* d_flag, void* stream) {
    if (t == nullptr || d_norm == nullptr || grad_list < 0 || grad_list >= t->n_lists) return APEX_B200_EINVAL;
    if (t->total_chunks == 0) return 0;
    if (t->d_table == nullptr) return APEX_B200_EINVAL;
    mt_clip_kernel<<<mt_grid(t->total_chunks, 8), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        mt_view(t), grad_list, d_norm, static_cast<float>(max_norm), d_flag);
    return mt_cuda_status();
}

// ---------------------------------------------------------------------------
// The same norm for callers that must not allocate.
//
// apex_b200_mt_l2norm needs one float of scratch per chunk and expects the
// caller to bring it.  Some callers cannot: a gradient-clipping call made while
// the framework's memory pool is full (a job sized to fill the 180 GB) would
// have to cudaMalloc, which can fail or make the pool flush and synchronise the
// whole device; and code running on the autograd thread while the bucket
// allreduce is still in flight on the reduction stream must not call
// cudaMalloc/cudaFree either, because both wait for every stream of the device.
// For them the partial sums live in a fixed array inside the library image.
//
// Cost and limits: APEX_B200_NOALLOC_MAX_CHUNKS floats (256 KiB) on every GPU
// that loads the library, held until the process exits.  All calls on a device
// share the array and it keeps the previous call's contents; each call writes
// slots [0, total_chunks) before it reads them, so nothing has to start at
// zero -- but two calls running at the same time on one GPU (two streams) would
// overwrite each other's partial sums.  Serialise them.
// ---------------------------------------------------------------------------

static __device__ __align__(256) float g_norm_partials[APEX_B200_NOALLOC_MAX_CHUNKS];

namespace {

constexpr int kMaxDevices = 64;
// Device address of g_norm_partials per device ordinal.  In host code the
// symbol's name is only a handle; the address has to be asked for, once per
// device, with that device current.
std::atomic<float*> g_norm_partials_addr[kMaxDevices];

int norm_partials_for_current_device(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return int(e);
    if (dev < 0 || dev >= kMaxDevices) return APEX_B200_ERANGE;
    float* p = g_norm_partials_addr[dev].load(std::memory_order_acquire);
    if (p == nullptr) {
        void* sym = nullptr;
        e = cudaGetSymbolAddress(&sym, g_norm_partials);
        if (e != cudaSuccess) return int(e);
        if (sym == nullptr) return APEX_B200_EINVAL;
        p = static_cast<float*>(sym);
        g_norm_partials_addr[dev].store(p, std::memory_order_release);   // racing threads store the same value
    }
    *out = p;
    return 0;
}

}  // namespace

extern "C" int apex_b200_mt_l2norm_noalloc(const apex_b200_mt_table* t, int grad_list, double inv_scale,
                                           int apply_scale, float* d_norm, void* stream) {
    if (t == nullptr || d_norm == nullptr || grad_list < 0 || grad_list >= t->n_lists) return APEX_B200_EINVAL;
    if (t->total_chunks < 0) return APEX_B200_EINVAL;
    if (t->total_chunks > APEX_B200_NOALLOC_MAX_CHUNKS) return APEX_B200_ERANGE;   // does not fit the fixed array
    if (t->total_chunks > 0 && t->d_table == nullptr) return APEX_B200_EINVAL;
    float* partials = nullptr;
    const int rc = norm_partials_for_current_device(&partials);    // first call on a device: set-up
    if (rc != 0) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (t->total_chunks > 0)
        mt_sumsq_kernel<<<mt_grid(t->total_chunks, 8), kThreads, 0, s>>>(
            mt_view(t), grad_list, static_cast<float>(inv_scale), apply_scale, partials);
    norm_finish_kernel<<<1, kThreads, 0, s>>>(partials, t->total_chunks, d_norm);
    return mt_cuda_status();
}
