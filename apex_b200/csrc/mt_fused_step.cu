// mt_fused_step.cu -- the fused mixed-precision master update.
//
// One pass per parameter element: read the (scaled) gradient, unscale it in
// fp32, apply the optimizer update to the fp32 master and its state, write the
// master/state back and the fp16 model copy next to it.  Algorithmic traffic:
// SGD-momentum 20 B/param, Adam 28 B/param -- nothing is read or
// written twice.
//
// Reference composition being replaced (each line a separate per-tensor launch
// there):
//   model_grads_to_master_grads     apex/fp16_utils/fp16util.py:93-112
//   _downscale_master               apex/fp16_utils/fp16_optimizer.py:236-241
//   self.optimizer.step()           apex/fp16_utils/fp16_optimizer.py:380
//       torch.optim.SGD  foreach    torch/optim/sgd.py:383-473
//       torch.optim.Adam foreach    torch/optim/adam.py:757-800 (non-capturable)
//   master_params_to_model_params   apex/fp16_utils/fp16util.py:115-129
//   skipped step on overflow        apex/fp16_utils/fp16_optimizer.py:372-375
//
// Arithmetic is written with explicit rounding intrinsics so that compiler
// flags cannot change it.  Where torch's CUDA foreach functors compute
// `a + alpha * b` inside ONE kernel nvcc contracts it to an FMA; each separate
// foreach launch rounds to fp32 in between.  The same roundings happen here.
#include <math.h>

#include "mt_common.cuh"

namespace apex_b200 {

struct SgdHp {
    float neg_lr, momentum, one_minus_dampening, weight_decay, inv_scale;
    int use_wd, use_momentum, nesterov, apply_scale;
    const float* clip_norm;     // device: total gradient norm, or nullptr (no clipping)
    float clip_max_norm;
};

struct AdamHp {
    float beta2, one_minus_beta2, lerp_w, one_minus_lerp_w, eps, weight_decay, decay_mul;
    float step_size, bc2_sqrt, inv_scale;
    int use_wd, decoupled, lerp_small, apply_scale;
    const float* clip_norm;
    float clip_max_norm;
};

// torch.nn.utils.clip_grad_norm_: every gradient is multiplied by
// clamp(max_norm / (total_norm + 1e-6), max=1.0) -- always, also when it is 1 (a NaN norm
// poisons the gradients there, and here).  1.0f when no clipping was asked for.
__device__ __forceinline__ float clip_coef(const float* norm, float max_norm) {
    if (norm == nullptr) return 1.0f;
    const float c = __fdiv_rn(max_norm, __fadd_rn(__ldg(norm), 1e-6f));
    return c > 1.0f ? 1.0f : c;      // NaN stays NaN
}

// torch/optim/sgd.py:425-468 (foreach) == :333-380 (single tensor)
template <bool kMom>
__device__ __forceinline__ void sgd_elem(float g, float& p, float& buf, const SgdHp& h, bool first, float inv, float coef) {
    g = __fmul_rn(g, inv);                                          // grad.mul_(1./loss_scale); inv == 1 is exact
    g = __fmul_rn(g, coef);                                         // clip_grad_norm_'s mul_; coef == 1 is exact
    if (h.use_wd) g = __fmaf_rn(h.weight_decay, p, g);              // _foreach_add(grads, params, alpha=wd)
    if (kMom) {
        if (first) {
            buf = g;                                                // clone(grad)
        } else {
            buf = __fmul_rn(buf, h.momentum);                       // _foreach_mul_(bufs, momentum)
            buf = __fmaf_rn(h.one_minus_dampening, g, buf);         // _foreach_add_(bufs, grads, alpha=1-dampening)
        }
        g = h.nesterov ? __fmaf_rn(h.momentum, buf, g) : buf;       // _foreach_add_(grads, bufs, alpha=momentum)
    }
    p = __fmaf_rn(h.neg_lr, g, p);                                  // _foreach_add_(params, grads, alpha=-lr)
}

// torch/optim/adam.py:686-800 (foreach, capturable=False, amsgrad=False)
__device__ __forceinline__ void adam_elem(float g, float& p, float& m, float& v, const AdamHp& h, float inv, float coef) {
    g = __fmul_rn(g, inv);
    g = __fmul_rn(g, coef);
    if (h.use_wd) {
        if (h.decoupled) p = __fmul_rn(p, h.decay_mul);             // _foreach_mul_(params, 1 - lr*wd)
        else g = __fmaf_rn(h.weight_decay, p, g);                   // _foreach_add(grads, params, alpha=wd)
    }
    // _foreach_lerp_(exp_avgs, grads, 1-beta1): ATen/native/Lerp.h
    const float diff = __fsub_rn(g, m);
    m = h.lerp_small ? __fmaf_rn(h.lerp_w, diff, m)
                     : __fmaf_rn(-diff, h.one_minus_lerp_w, g);
    v = __fmul_rn(v, h.beta2);                                      // _foreach_mul_(exp_avg_sqs, beta2)
    v = __fmaf_rn(h.one_minus_beta2, __fmul_rn(g, g), v);           // _foreach_addcmul_(v, g, g, 1-beta2)
    float denom = __fsqrt_rn(v);                                    // _foreach_sqrt
    denom = __fdiv_rn(denom, h.bc2_sqrt);                           // _foreach_div_(.., bias_correction2_sqrt)
    denom = __fadd_rn(denom, h.eps);                                // _foreach_add_(.., eps)
    p = __fmaf_rn(h.step_size, __fdiv_rn(m, denom), p);             // _foreach_addcdiv_(params, m, denom, step_size)
}

constexpr int kStepUnroll = 1;

// ---- SGD ---------------------------------------------------------------------

template <typename G, bool kModel, bool kMom>
__device__ __forceinline__ void sgd_chunk(const MtView& v, const MtChunk& c, const SgdHp& h) {
    const bool first = c.flags & APEX_B200_TF_FIRST_STEP;
    const float inv = (h.apply_scale && !(c.flags & APEX_B200_TF_NO_SCALE)) ? h.inv_scale : 1.0f;
    const float coef = (c.flags & APEX_B200_TF_NO_SCALE) ? 1.0f : clip_coef(h.clip_norm, h.clip_max_norm);
    const G* g = mt_ptr<const G>(v, 0, c.tensor) + c.offset;
    float* p = mt_ptr<float>(v, 1, c.tensor) + c.offset;
    float* b = kMom ? mt_ptr<float>(v, 2, c.tensor) + c.offset : nullptr;
    __half* q = kModel ? mt_ptr<__half>(v, 3, c.tensor) + c.offset : nullptr;
    const bool read_buf = kMom && !first;

    if (!(c.flags & APEX_B200_TF_UNALIGNED)) {
        const int nvec = c.len >> 3;
        int i = threadIdx.x;
#pragma unroll 1
        for (; i + (kStepUnroll - 1) * kThreads < nvec; i += kStepUnroll * kThreads) {
            F8 gv[kStepUnroll], pv[kStepUnroll], bv[kStepUnroll];
#pragma unroll
            for (int u = 0; u < kStepUnroll; ++u) {
                const size_t e = size_t(i + u * kThreads) * kVec;
                gv[u] = GradIo<G>::ld8(g + e);
                pv[u] = ld_f8(p + e);
                if (read_buf) bv[u] = ld_f8(b + e); else bv[u] = F8{};
            }
#pragma unroll
            for (int u = 0; u < kStepUnroll; ++u) {
                const size_t e = size_t(i + u * kThreads) * kVec;
#pragma unroll
                for (int k = 0; k < 8; ++k) sgd_elem<kMom>(gv[u].v[k], pv[u].v[k], bv[u].v[k], h, first, inv, coef);
                st_f8(p + e, pv[u]);
                if (kMom) st_f8(b + e, bv[u]);
                if (kModel) st_h8(q + e, f8_to_h8(pv[u]));
            }
        }
        if (kStepUnroll > 1) {
#pragma unroll 1
            for (; i < nvec; i += kThreads) {
                const size_t e = size_t(i) * kVec;
                F8 gv = GradIo<G>::ld8(g + e), pv = ld_f8(p + e), bv = F8{};
                if (read_buf) bv = ld_f8(b + e);
#pragma unroll
                for (int k = 0; k < 8; ++k) sgd_elem<kMom>(gv.v[k], pv.v[k], bv.v[k], h, first, inv, coef);
                st_f8(p + e, pv);
                if (kMom) st_f8(b + e, bv);
                if (kModel) st_h8(q + e, f8_to_h8(pv));
            }
        }
        const int t = (nvec << 3) + threadIdx.x;
        if (t < c.len) {
            float pe = p[t], be = read_buf ? b[t] : 0.f;
            sgd_elem<kMom>(GradIo<G>::ld1(g + t), pe, be, h, first, inv, coef);
            p[t] = pe;
            if (kMom) b[t] = be;
            if (kModel) q[t] = __float2half_rn(pe);
        }
    } else {
#pragma unroll 1
        for (int i = threadIdx.x; i < c.len; i += kThreads) {
            float pe = p[i], be = read_buf ? b[i] : 0.f;
            sgd_elem<kMom>(GradIo<G>::ld1(g + i), pe, be, h, first, inv, coef);
            p[i] = pe;
            if (kMom) b[i] = be;
            if (kModel) q[i] = __float2half_rn(pe);
        }
    }
}

template <bool kMom>
__global__ void __launch_bounds__(kThreads, 4)
mt_fused_sgd_kernel(MtView v, SgdHp h, const int32_t* __restrict__ flag) {
    if (flag != nullptr && *flag != 0) return;   // skipped step: nothing is touched
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        const bool f32 = c.flags & APEX_B200_TF_GRAD_F32;
        const bool model = !(c.flags & APEX_B200_TF_NO_MODEL);
        if (!f32 && model) sgd_chunk<__half, true, kMom>(v, c, h);
        else if (f32 && model) sgd_chunk<float, true, kMom>(v, c, h);
        else if (f32) sgd_chunk<float, false, kMom>(v, c, h);
        else sgd_chunk<__half, false, kMom>(v, c, h);
    }
}

// ---- Adam --------------------------------------------------------------------

template <typename G, bool kModel>
__device__ __forceinline__ void adam_chunk(const MtView& v, const MtChunk& c, const AdamHp& h) {
    const float inv = (h.apply_scale && !(c.flags & APEX_B200_TF_NO_SCALE)) ? h.inv_scale : 1.0f;
    const float coef = (c.flags & APEX_B200_TF_NO_SCALE) ? 1.0f : clip_coef(h.clip_norm, h.clip_max_norm);
    const G* g = mt_ptr<const G>(v, 0, c.tensor) + c.offset;
    float* p = mt_ptr<float>(v, 1, c.tensor) + c.offset;
    float* m = mt_ptr<float>(v, 2, c.tensor) + c.offset;
    float* s = mt_ptr<float>(v, 3, c.tensor) + c.offset;
    __half* q = kModel ? mt_ptr<__half>(v, 4, c.tensor) + c.offset : nullptr;

    if (!(c.flags & APEX_B200_TF_UNALIGNED)) {
        const int nvec = c.len >> 3;
        int i = threadIdx.x;
#pragma unroll 1
        for (; i + (kStepUnroll - 1) * kThreads < nvec; i += kStepUnroll * kThreads) {
            F8 gv[kStepUnroll], pv[kStepUnroll], mv[kStepUnroll], sv[kStepUnroll];
#pragma unroll
            for (int u = 0; u < kStepUnroll; ++u) {
                const size_t e = size_t(i + u * kThreads) * kVec;
                gv[u] = GradIo<G>::ld8(g + e);
                pv[u] = ld_f8(p + e);
                mv[u] = ld_f8(m + e);
                sv[u] = ld_f8(s + e);
            }
#pragma unroll
            for (int u = 0; u < kStepUnroll; ++u) {
                const size_t e = size_t(i + u * kThreads) * kVec;
#pragma unroll
                for (int k = 0; k < 8; ++k) adam_elem(gv[u].v[k], pv[u].v[k], mv[u].v[k], sv[u].v[k], h, inv, coef);
                st_f8(p + e, pv[u]);
                st_f8(m + e, mv[u]);
                st_f8(s + e, sv[u]);
                if (kModel) st_h8(q + e, f8_to_h8(pv[u]));
            }
        }
#pragma unroll 1
        for (; i < nvec; i += kThreads) {
            const size_t e = size_t(i) * kVec;
            F8 gv = GradIo<G>::ld8(g + e), pv = ld_f8(p + e), mv = ld_f8(m + e), sv = ld_f8(s + e);
#pragma unroll
            for (int k = 0; k < 8; ++k) adam_elem(gv.v[k], pv.v[k], mv.v[k], sv.v[k], h, inv, coef);
            st_f8(p + e, pv);
            st_f8(m + e, mv);
            st_f8(s + e, sv);
            if (kModel) st_h8(q + e, f8_to_h8(pv));
        }
        const int t = (nvec << 3) + threadIdx.x;
        if (t < c.len) {
            float pe = p[t], me = m[t], se = s[t];
            adam_elem(GradIo<G>::ld1(g + t), pe, me, se, h, inv, coef);
            p[t] = pe; m[t] = me; s[t] = se;
            if (kModel) q[t] = __float2half_rn(pe);
        }
    } else {
#pragma unroll 1
        for (int i = threadIdx.x; i < c.len; i += kThreads) {
            float pe = p[i], me = m[i], se = s[i];
            adam_elem(GradIo<G>::ld1(g + i), pe, me, se, h, inv, coef);
            p[i] = pe; m[i] = me; s[i] = se;
            if (kModel) q[i] = __float2half_rn(pe);
        }
    }
}

__global__ void __launch_bounds__(kThreads, 4)
mt_fused_adam_kernel(MtView v, AdamHp h, const int32_t* __restrict__ flag) {
    if (flag != nullptr && *flag != 0) return;
    for (int chunk = blockIdx.x; chunk < v.total_chunks; chunk += gridDim.x) {
        const MtChunk c = mt_locate(v, chunk);
        const bool f32 = c.flags & APEX_B200_TF_GRAD_F32;
        const bool model = !(c.flags & APEX_B200_TF_NO_MODEL);
        if (!f32 && model) adam_chunk<__half, true>(v, c, h);
        else if (f32 && model) adam_chunk<float, true>(v, c, h);
        else if (f32) adam_chunk<float, false>(v, c, h);
        else adam_chunk<__half, false>(v, c, h);
    }
}

}  // namespace apex_b200

using namespace apex_b200;

extern "C" int apex_b200_mt_fused_sgd(const apex_b200_mt_table* t, const apex_b200_sgd_args* a,
                                      const int32_t* d_flag, void* stream) {
    if (t == nullptr || a == nullptr) return APEX_B200_EINVAL;
    if (t->total_chunks == 0) return 0;
    if (t->d_table == nullptr) return APEX_B200_EINVAL;
    SgdHp h;
    h.use_momentum = a->momentum != 0.0;
    h.use_wd = a->weight_decay != 0.0;
    if (t->n_lists < 4) return APEX_B200_EINVAL;
    // python computes -lr and 1 - dampening in double; torch narrows the scalar
    // to the kernel's fp32 math type.
    h.neg_lr = static_cast<float>(-a->lr);
    h.momentum = static_cast<float>(a->momentum);
    h.one_minus_dampening = static_cast<float>(1.0 - a->dampening);
    h.weight_decay = static_cast<float>(a->weight_decay);
    h.inv_scale = static_cast<float>(a->inv_scale);
    h.nesterov = a->nesterov != 0;
    h.apply_scale = a->apply_scale != 0;
    h.clip_norm = a->clip_norm;
    h.clip_max_norm = static_cast<float>(a->clip_max_norm);
    const int grid = mt_grid(t->total_chunks, 4);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (h.use_momentum) mt_fused_sgd_kernel<true><<<grid, kThreads, 0, s>>>(mt_view(t), h, d_flag);
    else mt_fused_sgd_kernel<false><<<grid, kThreads, 0, s>>>(mt_view(t), h, d_flag);
    return mt_cuda_status();
}

extern "C" int apex_b200_mt_fused_adam(const apex_b200_mt_table* t, const apex_b200_adam_args* a,
                                       const int32_t* d_flag, void* stream) {
    if (t == nullptr || a == nullptr) return APEX_B200_EINVAL;
    if (t->total_chunks == 0) return 0;
    if (t->d_table == nullptr || t->n_lists < 5) return APEX_B200_EINVAL;
    AdamHp h;
    const double w = 1.0 - a->beta1;                 // python: 1 - beta1
    h.lerp_w = static_cast<float>(w);
    h.lerp_small = fabsf(h.lerp_w) < 0.5f;           // ATen is_lerp_weight_small
    h.one_minus_lerp_w = 1.0f - h.lerp_w;            // fp32 subtraction, as in Lerp.h
    h.beta2 = static_cast<float>(a->beta2);
    h.one_minus_beta2 = static_cast<float>(1.0 - a->beta2);
    h.eps = static_cast<float>(a->eps);
    h.weight_decay = static_cast<float>(a->weight_decay);
    h.decay_mul = static_cast<float>(1.0 - a->lr * a->weight_decay);
    h.step_size = static_cast<float>(a->step_size);
    h.bc2_sqrt = static_cast<float>(a->bias_correction2_sqrt);
    h.inv_scale = static_cast<float>(a->inv_scale);
    h.use_wd = a->weight_decay != 0.0;
    h.decoupled = a->decoupled_weight_decay != 0;
    h.apply_scale = a->apply_scale != 0;
    h.clip_norm = a->clip_norm;
    h.clip_max_norm = static_cast<float>(a->clip_max_norm);
    mt_fused_adam_kernel<<<mt_grid(t->total_chunks, 4), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        mt_view(t), h, d_flag);
    return mt_cuda_status();
}
