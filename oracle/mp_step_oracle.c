/* mp_step_oracle.c -- TEST INFRASTRUCTURE (see oracle/__init__.py).
 *
 * Element-by-element C restatement of the arithmetic on the hot path, free of
 * torch, so that the CUDA kernels can be checked bit for bit against something
 * that shares no code with them or with PyTorch.  Scalars arrive as doubles and
 * are narrowed to float exactly where torch narrows them.
 *
 * Reference lines restated:
 *   overflow test     apex/fp16_utils/loss_scaler.py:92-110 (elementwise form; equal
 *                     to the sum test for fp16 inputs)
 *   unscale           apex/fp16_utils/fp16_optimizer.py:236-241  g *= float(1/scale)
 *   fp16->fp32 copy   apex/fp16_utils/fp16util.py:106-112
 *   fp32->fp16 copy   apex/fp16_utils/fp16util.py:127-129 (round to nearest even)
 *   inner step        apex/fp16_utils/fp16_optimizer.py:380 -> torch/optim/sgd.py:425-468,
 *                     torch/optim/adam.py:686-800; one rounding per foreach launch,
 *                     `a + alpha*b` inside one launch is a fused multiply-add
 *   bucket average    apex/parallel/distributed.py:19-20; torch's CUDA divide by a
 *                     python scalar multiplies by float(1.0/world)
 *
 * Build: oracle/build.py (gcc -O2 -ffp-contract=off -shared -fPIC).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

typedef _Float16 half_t;

static inline float h2f(uint16_t bits) { half_t h; memcpy(&h, &bits, 2); return (float)h; }
static inline uint16_t f2h(float f) { half_t h = (half_t)f; uint16_t b; memcpy(&b, &h, 2); return b; }
static inline int nonfinite(float x) { uint32_t u; memcpy(&u, &x, 4); return (u & 0x7f800000u) == 0x7f800000u; }

/* index of the first element that is inf/NaN, or -1 */
long long ora_first_nonfinite_f16(const uint16_t* g, long long n) {
    for (long long i = 0; i < n; ++i) if ((g[i] & 0x7c00u) == 0x7c00u) return i;
    return -1;
}
long long ora_first_nonfinite_f32(const float* g, long long n) {
    for (long long i = 0; i < n; ++i) if (nonfinite(g[i])) return i;
    return -1;
}

void ora_f16_to_f32_scaled(const uint16_t* g, float* out, long long n, double inv_scale, int apply) {
    const float inv = (float)inv_scale;
    for (long long i = 0; i < n; ++i) { float x = h2f(g[i]); out[i] = apply ? x * inv : x; }
}
/* the factor torch.nn.utils.clip_grad_norm_ multiplies every gradient by (fp32 arithmetic) */
float ora_clip_coef(float total_norm, double max_norm) {
    const float c = (float)max_norm / (total_norm + 1e-6f);
    return c > 1.0f ? 1.0f : c;
}
void ora_f16_to_f32_scaled_clipped(const uint16_t* g, float* out, long long n, double inv_scale, float coef) {
    const float inv = (float)inv_scale;
    for (long long i = 0; i < n; ++i) { float x = h2f(g[i]) * inv; out[i] = x * coef; }
}
void ora_scale_f32(float* g, long long n, double inv_scale) {
    const float inv = (float)inv_scale;
    for (long long i = 0; i < n; ++i) g[i] = g[i] * inv;
}
void ora_f32_to_f16(const float* p, uint16_t* out, long long n) {
    for (long long i = 0; i < n; ++i) out[i] = f2h(p[i]);
}

/* grad: fp16 bits (g16 != NULL) or fp32 (g32).  buf may be NULL when momentum == 0.
 * p16 may be NULL.  first != 0: momentum buffer is initialised from the gradient. */
void ora_sgd(const uint16_t* g16, const float* g32, float* p, float* buf, uint16_t* p16, long long n,
             double lr, double momentum, double dampening, double weight_decay, int nesterov,
             int first, double inv_scale, int apply_scale, float clip_coef, int apply_clip) {
    const float neg_lr = (float)(-lr), mom = (float)momentum, omd = (float)(1.0 - dampening);
    const float wd = (float)weight_decay, inv = (float)inv_scale;
    for (long long i = 0; i < n; ++i) {
        float g = g16 ? h2f(g16[i]) : g32[i];
        if (apply_scale) g = g * inv;
        if (apply_clip) g = g * clip_coef;
        if (weight_decay != 0.0) g = fmaf(wd, p[i], g);
        if (momentum != 0.0) {
            float b;
            if (first) b = g;
            else { b = buf[i] * mom; b = fmaf(omd, g, b); }
            buf[i] = b;
            g = nesterov ? fmaf(mom, b, g) : b;
        }
        p[i] = fmaf(neg_lr, g, p[i]);
        if (p16) p16[i] = f2h(p[i]);
    }
}

/* flavour 0: torch's CUDA foreach kernels (what the reference runs on a GPU, and what
 *            apex_b200's kernel computes):  v += (1-b2)*(g*g) as fma(1-b2, g*g, v);
 *            p += step_size*(m/denom) as fma(step_size, m/denom, p).
 * flavour 1: torch's CPU kernels (what generated tests/golden): same formula, other
 *            association -- addcmul is v + ((1-b2)*g)*g and addcdiv is p + (step_size*m)/denom
 *            (ATen/native/cpu/PointwiseOpsKernel.cpp), vectorised code contracted by gcc.
 *            `cpu_fma` selects whether the final add of addcmul is fused. */
void ora_adam_flavour(const uint16_t* g16, const float* g32, float* p, float* m, float* v, uint16_t* p16, long long n,
              double lr, double beta1, double beta2, double eps, double weight_decay,
              double step_size, double bc2_sqrt, int decoupled, double inv_scale, int apply_scale,
              int flavour, int cpu_fma, float clip_coef, int apply_clip) {
    const float w = (float)(1.0 - beta1), b2 = (float)beta2, omb2 = (float)(1.0 - beta2);
    const float one_minus_w = 1.0f - w;
    const float e = (float)eps, wd = (float)weight_decay, decay = (float)(1.0 - lr * weight_decay);
    const float ss = (float)step_size, bc = (float)bc2_sqrt, inv = (float)inv_scale;
    const int small = fabsf(w) < 0.5f;
    for (long long i = 0; i < n; ++i) {
        float g = g16 ? h2f(g16[i]) : g32[i];
        if (apply_scale) g = g * inv;
        if (apply_clip) g = g * clip_coef;
        if (weight_decay != 0.0) {
            if (decoupled) p[i] = p[i] * decay;
            else g = fmaf(wd, p[i], g);
        }
        const float diff = g - m[i];
        m[i] = small ? fmaf(w, diff, m[i]) : fmaf(-diff, one_minus_w, g);
        float vv = v[i] * b2;
        if (flavour == 0) vv = fmaf(omb2, g * g, vv);
        else { const float t = omb2 * g; vv = cpu_fma ? fmaf(t, g, vv) : vv + t * g; }
        v[i] = vv;
        float denom = sqrtf(vv);
        denom = denom / bc;
        denom = denom + e;
        if (flavour == 0) p[i] = fmaf(ss, m[i] / denom, p[i]);
        else p[i] = p[i] + (ss * m[i]) / denom;
        if (p16) p16[i] = f2h(p[i]);
    }
}

void ora_adam(const uint16_t* g16, const float* g32, float* p, float* m, float* v, uint16_t* p16, long long n,
              double lr, double beta1, double beta2, double eps, double weight_decay,
              double step_size, double bc2_sqrt, int decoupled, double inv_scale, int apply_scale) {
    ora_adam_flavour(g16, g32, p, m, v, p16, n, lr, beta1, beta2, eps, weight_decay, step_size, bc2_sqrt,
                     decoupled, inv_scale, apply_scale, 0, 0, 1.0f, 0);
}

/* x *= float(1/divisor) in the bucket dtype */
void ora_average_f16(uint16_t* x, long long n, double divisor) {
    const float inv = (float)(1.0 / divisor);
    for (long long i = 0; i < n; ++i) x[i] = f2h(h2f(x[i]) * inv);
}
void ora_average_f32(float* x, long long n, double divisor) {
    const float inv = (float)(1.0 / divisor);
    for (long long i = 0; i < n; ++i) x[i] = x[i] * inv;
}

/* sum of squares in double, for the gradient-norm kernels */
double ora_sumsq_f16(const uint16_t* g, long long n, double inv_scale, int apply) {
    const float inv = (float)inv_scale; double acc = 0.0;
    for (long long i = 0; i < n; ++i) { float x = h2f(g[i]); if (apply) x = x * inv; acc += (double)x * (double)x; }
    return acc;
}
double ora_sumsq_f32(const float* g, long long n, double inv_scale, int apply) {
    const float inv = (float)inv_scale; double acc = 0.0;
    for (long long i = 0; i < n; ++i) { float x = g[i]; if (apply) x = x * inv; acc += (double)x * (double)x; }
    return acc;
}
