/*
 * apex_b200.h -- C ABI of libapex_b200.so (python name: apex_b200._C)
 *
 * One hot path only: apex's mixed-precision gradient step.
 * Every entry point below is the B200 (sm_100a) replacement for a place where
 * the reference (mbrukman/apex-1, pure python) calls a stock torch op per
 * tensor.  The reference has no FFI of its own; its intended native seam is
 * the commented-out `from apex_C import scale_check_overflow`
 * (apex/amp/scaler.py:3) and the TODO at apex/fp16_utils/fp16_optimizer.py:10.
 * Each declaration cites the reference lines it replaces.
 *
 * Conventions
 *  - plain C types only; `stream` is a cudaStream_t passed as void*.
 *  - every call launches asynchronously on `stream` and returns without
 *    synchronising.  Return value: 0 = ok, >0 = cudaError_t, <0 = APEX_B200_E*.
 *    apex_b200_error_string() turns any of them into text.
 *  - the library owns no device memory (tensors, tables and the overflow flag
 *    are allocated by the caller -- torch in practice) and keeps no mutable
 *    global state, so calls from several host threads / streams are fine.
 *  - there is no CPU implementation behind any of these symbols.
 *
 * The multi-tensor table
 *  A "table" describes up to thousands of tensors so that ONE launch covers
 *  them all.  It is packed on the host by apex_b200_mt_table_pack() into a
 *  caller buffer, copied by the caller to device memory, and then handed to
 *  the kernels through `apex_b200_mt_table`.  Layout (little endian):
 *      int32  header[4]                 n_tensors, n_lists, chunk_elems, total_chunks
 *      int64  numel[n]
 *      int64  elem_start[n+1]           exclusive prefix sum of numel (flat-bucket offsets)
 *      int64  ptr[n_lists][n]           device addresses, list-major
 *      int32  flags[n]                  APEX_B200_TF_* bits per tensor
 *      int32  chunk_start[n+1]          exclusive prefix sum of ceil(numel/chunk_elems)
 *      int32  chunk_tensor[total_chunks] tensor index of every chunk
 *  Kernels launch ONE CTA PER CHUNK; a CTA maps its chunk id to (tensor, offset)
 *  with one load from chunk_tensor.  (Every array starts 8-byte aligned.)
 */
#ifndef APEX_B200_H_
#define APEX_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define APEX_B200_VERSION 100

/* negative return codes (positive ones are cudaError_t) */
#define APEX_B200_EINVAL (-1)   /* bad argument (null pointer, bad list index, ...) */
#define APEX_B200_ERANGE (-2)   /* size beyond what the entry point supports        */
#define APEX_B200_ENOSPC (-3)   /* caller buffer too small                          */

/* per-tensor flag bits in the table */
#define APEX_B200_TF_GRAD_F32   1   /* list 0 (the gradient) holds fp32, not fp16         */
#define APEX_B200_TF_FIRST_STEP 2   /* SGD: momentum buffer does not exist yet -> buf = g */
#define APEX_B200_TF_NO_MODEL   4   /* no fp16 model copy to write (fp32-native param)    */
#define APEX_B200_TF_UNALIGNED  8   /* set by the packer: some pointer is not 32 B aligned */
#define APEX_B200_TF_NO_SCALE   16  /* gradient is already unscaled: ignore inv_scale      */

/* value of an overflow flag word: 0 = clean, otherwise INT32_MAX - (index of the
 * first tensor, in table order, that holds an inf or NaN). */
#define APEX_B200_FLAG_FIRST_BAD(flag) (0x7fffffff - (flag))

typedef struct apex_b200_mt_table {
    const void* d_table;      /* device copy of the packed table */
    int32_t n_tensors;
    int32_t n_lists;
    int32_t chunk_elems;
    int32_t total_chunks;
} apex_b200_mt_table;

/* ---- housekeeping ------------------------------------------------------- */

int apex_b200_version(void);
const char* apex_b200_error_string(int code);
/* SM count and compute capability of the current device. */
int apex_b200_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ---- table packing (host only, no CUDA call) ---------------------------- */

/* Number of chunks a table over tensors of these sizes will have (>= 0), or a
 * negative APEX_B200_E* code. */
int64_t apex_b200_mt_count_chunks(int n_tensors, const int64_t* numels, int chunk_elems);

/* Bytes of the packed table; total_chunks from apex_b200_mt_count_chunks. */
size_t apex_b200_mt_table_bytes(int n_tensors, int n_lists, int64_t total_chunks);

/* ptrs is list-major: ptrs[l * n_tensors + i].  flags may be NULL (all zero).
 * chunk_elems must be a positive multiple of 2048.  Tensors with numel == 0
 * are legal and get no chunk.  Fills *out (d_table left NULL). */
int apex_b200_mt_table_pack(void* host_buf, size_t host_buf_bytes,
                            int n_tensors, int n_lists,
                            const int64_t* numels, const uint64_t* ptrs,
                            const int32_t* flags, int chunk_elems,
                            apex_b200_mt_table* out);

/* ---- overflow flag ------------------------------------------------------ */

/* *d_flag = 0 on `stream` (cudaMemsetAsync). */
int apex_b200_flag_clear(int32_t* d_flag, void* stream);

/* Replaces DynamicLossScaler.has_overflow / _has_inf_or_nan
 * (apex/fp16_utils/loss_scaler.py:84-110) as driven by
 * FP16_Optimizer._check_overflow (apex/fp16_utils/fp16_optimizer.py:213-221)
 * and the check half of scale_check_overflow (apex/amp/scaler.py:6-13):
 * one launch scans list `grad_list` of every tensor (fp16, or fp32 where
 * APEX_B200_TF_GRAD_F32 is set) and atomically raises *d_flag to
 * INT32_MAX - tensor_index for every tensor holding an inf/NaN, i.e. the flag
 * ends as the FIRST bad tensor.  No host sync.  The flag is not cleared here. */
int apex_b200_mt_check_overflow(const apex_b200_mt_table* t, int grad_list,
                                int32_t* d_flag, void* stream);

/* In-place fp32 `g *= inv_scale`.
 * mode 0 replaces FP16_Optimizer._downscale_master
 *   (apex/fp16_utils/fp16_optimizer.py:236-241) under the early return of
 *   update_master_grads (:493-495): nothing is touched when *d_flag != 0.
 * mode 1 replaces the mul_ of scale_check_overflow inside
 *   LossScaler.unscale_and_update (apex/amp/scaler.py:17, :35-40): tensors
 *   before the first bad one are scaled, the bad one and all later ones are
 *   left alone (the reference's partial mutation).
 * d_flag may be NULL (always apply).  Tensors are fp32 where TF_GRAD_F32 is
 * set and fp16 otherwise (fp16: half_rn(float(g) * inv_scale), torch's mul_). */
int apex_b200_mt_unscale(const apex_b200_mt_table* t, int grad_list,
                         double inv_scale, const int32_t* d_flag, int mode,
                         void* stream);

/* ---- fp16 <-> fp32 ------------------------------------------------------ */

/* Replaces model_grads_to_master_grads (apex/fp16_utils/fp16util.py:93-112)
 * fused with _downscale_master (fp16_optimizer.py:236-241):
 * dst32[i] = float(src16[i]) * inv_scale (exact widening when apply_scale == 0).
 * With d_clip_norm != NULL the result is further multiplied by
 * clamp(clip_max_norm / (*d_clip_norm + 1e-6), max=1), the factor of
 * clip_master_grads (fp16_optimizer.py:243-265 -> torch clip_grad_norm_).
 * Skipped entirely when d_flag != NULL and *d_flag != 0. */
int apex_b200_mt_grads_to_master(const apex_b200_mt_table* t, int src_list,
                                 int dst_list, double inv_scale, int apply_scale,
                                 const int32_t* d_flag, const float* d_clip_norm,
                                 double clip_max_norm, void* stream);

/* Replaces master_params_to_model_params (apex/fp16_utils/fp16util.py:115-129):
 * dst16[i] = half_rn(src32[i]).  Skipped when d_flag != NULL and *d_flag != 0. */
int apex_b200_mt_master_to_model(const apex_b200_mt_table* t, int src_list,
                                 int dst_list, const int32_t* d_flag,
                                 void* stream);

/* Same-dtype copy, tensor i of src_list -> tensor i of dst_list: the copy_ of
 * model_grads_to_master_grads / master_params_to_model_params
 * (apex/fp16_utils/fp16util.py:110, :129) when model and master share a dtype
 * (fp32 BatchNorm parameters after network_to_half, fp16util.py:21-41).
 * elem_size in {2, 4, 8}.  Skipped when d_flag != NULL and *d_flag != 0. */
int apex_b200_mt_copy(const apex_b200_mt_table* t, int src_list, int dst_list,
                      int elem_size, const int32_t* d_flag, void* stream);

/* Replaces the per-tensor grad.zero_() loops of FP16_Optimizer.zero_grad
 * (apex/fp16_utils/fp16_optimizer.py:187-211).  elem_size is 2 or 4; tensors
 * flagged APEX_B200_TF_GRAD_F32 are always 4. */
int apex_b200_mt_zero(const apex_b200_mt_table* t, int list, int elem_size,
                      void* stream);

/* ---- fused master update ------------------------------------------------ */

/* Lists: 0 = grad (fp16, or fp32 with TF_GRAD_F32), 1 = fp32 master param,
 * 2 = fp32 momentum buffer, 3 = fp16 model param (ignored with TF_NO_MODEL).
 * Replaces, in one pass that reads every byte once:
 *   model_grads_to_master_grads (apex/fp16_utils/fp16util.py:93-112),
 *   _downscale_master (apex/fp16_utils/fp16_optimizer.py:236-241), the inner
 *   torch.optim.SGD step (fp16_optimizer.py:380 -> torch/optim/sgd.py:383-473)
 *   and master_params_to_model_params (apex/fp16_utils/fp16util.py:115-129).
 * Arithmetic order and FMA placement follow torch's CUDA foreach kernels.
 * The whole launch is a no-op when d_flag != NULL and *d_flag != 0
 * (the skipped step of fp16_optimizer.py:372-375). */
typedef struct apex_b200_sgd_args {
    double lr, momentum, dampening, weight_decay;
    int32_t nesterov;
    int32_t apply_scale;      /* 0: gradient is used as is */
    double inv_scale;         /* 1./loss_scale as a python double */
    const float* clip_norm;   /* device: total norm from apex_b200_mt_l2norm, or NULL.  The unscaled
                                 gradient is multiplied by clamp(clip_max_norm/(norm+1e-6), max=1):
                                 clip_master_grads (fp16_optimizer.py:243-265) folded into the step */
    double clip_max_norm;
} apex_b200_sgd_args;

int apex_b200_mt_fused_sgd(const apex_b200_mt_table* t,
                           const apex_b200_sgd_args* a,
                           const int32_t* d_flag, void* stream);

/* Lists: 0 = grad, 1 = fp32 master, 2 = exp_avg, 3 = exp_avg_sq, 4 = fp16 model.
 * Same fusion, with the same reference lines (apex/fp16_utils/fp16util.py:93-129,
 * apex/fp16_utils/fp16_optimizer.py:236-241, :372-380), for an inner
 * torch.optim.Adam / AdamW (torch/optim/adam.py:416-547, foreach branch :757-800).  step_size = (lr / bias_correction1) * -1 and
 * bias_correction2_sqrt are computed by the caller in python doubles exactly
 * as torch does. */
typedef struct apex_b200_adam_args {
    double lr, beta1, beta2, eps, weight_decay;
    double step_size, bias_correction2_sqrt;
    int32_t decoupled_weight_decay;
    int32_t apply_scale;
    double inv_scale;
    const float* clip_norm;   /* as in apex_b200_sgd_args */
    double clip_max_norm;
} apex_b200_adam_args;

int apex_b200_mt_fused_adam(const apex_b200_mt_table* t,
                            const apex_b200_adam_args* a,
                            const int32_t* d_flag, void* stream);

/* ---- bucket path (apex.parallel) ---------------------------------------- */

/* Replaces _flatten_dense_tensors(bucket) in apply_flat_dist_call
 * (apex/parallel/distributed.py:12): tensor i of list `src_list` is copied to
 * flat + elem_start[i] * elem_size.  elem_size in {2, 4, 8}. */
int apex_b200_mt_flatten(const apex_b200_mt_table* t, int src_list, void* flat,
                         int elem_size, void* stream);

/* Replaces `coalesced /= world_size` plus the _unflatten_dense_tensors/copy_
 * loop of apply_flat_dist_call (apex/parallel/distributed.py:19-23) in one
 * pass, with the inf/NaN check of the reduced values for free:
 * dst[i] = T(opmath(flat[..]) * mul) where mul = (opmath)(1.0 / world_size),
 * which is how torch's CUDA true-divide by a python scalar computes it.
 * apply_mul == 0 gives a plain scatter (the broadcast case, :160).
 * dtype: 0 = fp16, 1 = fp32, 2 = fp64.  d_flag may be NULL (no check). */
int apex_b200_mt_scale_unflatten(const apex_b200_mt_table* t, const void* flat,
                                 int dst_list, int dtype, double divisor,
                                 int apply_mul, int32_t* d_flag, void* stream);

/* In-place `flat /= divisor` on one contiguous buffer: the
 * `coalesced /= dist.get_world_size()` of apex/parallel/distributed.py:20 for a
 * bucket that holds a single tensor (nothing to gather or scatter).  Same
 * arithmetic and check as above. */
int apex_b200_scale_flat(void* flat, int64_t numel, int dtype, double divisor,
                         int32_t* d_flag, void* stream);

/* Replaces `dist.all_reduce(coalesced)` of apply_flat_dist_call
 * (apex/parallel/distributed.py:15-17) for a flat bucket that lives in peer-mapped
 * ("symmetric") memory of one NVLink domain: in-place SUM over `world` GPUs, two-shot,
 * rank r reducing slice r and writing it to every copy.  With `multicast_ptr != 0`
 * the NVSwitch does both halves (multimem.ld_reduce / multimem.st); otherwise
 * `peer_ptrs[world]` (host array of the buffer's address on every rank, own rank
 * included) is used with plain loads/stores over NVLink.  fp16 (dtype 0) is summed in
 * fp32 and rounded once; fp32 is dtype 1.  numel must be a multiple of
 * (16 / element size) * world (pad the buffer).  The call only launches this rank's
 * kernel; the caller must order it after every rank has filled its buffer and must
 * not read the result before every rank's kernel has finished (a cross-rank barrier
 * on the stream before and after).  ctas <= 0 picks a default. */
int apex_b200_symm_allreduce(const uint64_t* peer_ptrs, uint64_t multicast_ptr,
                             int rank, int world, int64_t numel, int dtype,
                             int ctas, void* stream);

/* ---- gradient norm (clip_master_grads) ---------------------------------- */

/* Replaces torch.nn.utils.clip_grad_norm_ as called by
 * FP16_Optimizer.clip_master_grads (apex/fp16_utils/fp16_optimizer.py:243-265)
 * for norm_type 2: a two-stage reduction.  Stage 1 writes one fp32 partial
 * sum of squares per chunk into `d_partials` (>= t->total_chunks floats,
 * caller-allocated scratch); stage 2 adds them in a fixed order in fp64 and
 * writes sqrt(sum) as fp32 to *d_norm.  The value each element contributes is
 * float(grad) * inv_scale when apply_scale != 0 (so the norm of the master
 * gradients can be had straight from the fp16 gradients).  Deterministic. */
int apex_b200_mt_l2norm(const apex_b200_mt_table* t, int grad_list,
                        double inv_scale, int apply_scale, float* d_partials,
                        int64_t partials_len, float* d_norm, void* stream);

/* No counterpart in the reference (it has no native code and lets torch allocate
 * whatever it needs).  Same result as apex_b200_mt_l2norm, for callers that cannot
 * allocate the `d_partials` scratch: a clip_master_grads issued while the
 * framework's memory pool is exhausted (cudaMalloc may fail, or flush the pool and
 * synchronise the device), or anything running while the bucket allreduce is still
 * in flight on another stream (cudaMalloc/cudaFree wait for the whole device).
 * The partial sums go to a fixed array of APEX_B200_NOALLOC_MAX_CHUNKS floats
 * (256 KiB) that every GPU loading the library holds until exit; tables with more
 * chunks are refused with APEX_B200_ERANGE.  This entry point calls no cudaMalloc,
 * cudaFree or other allocating routine.  The first call on each device looks up
 * the array's device address (once, with that device current); make that first
 * call before starting any stream capture.  All calls on a device share the
 * array, which keeps the last call's contents (every slot is written before it is
 * read, nothing needs clearing): calls that overlap on the GPU race -- keep them
 * on one stream or order them with events. */
#define APEX_B200_NOALLOC_MAX_CHUNKS 65536
int apex_b200_mt_l2norm_noalloc(const apex_b200_mt_table* t, int grad_list,
                                double inv_scale, int apply_scale,
                                float* d_norm, void* stream);

/* Second half of clip_master_grads (apex/fp16_utils/fp16_optimizer.py:263):
 * in-place `g *= clip_coef` for fp32 gradients where
 * clip_coef = max_norm / (*d_norm + 1e-6), applied only when clip_coef < 1
 * (torch's clip_grad_norm_ clamps the coefficient to 1).  Reads the norm from
 * device memory: no host sync. */
int apex_b200_mt_clip_by_norm(const apex_b200_mt_table* t, int grad_list,
                              const float* d_norm, double max_norm,
                              const int32_t* d_flag, void* stream);

/* ---- LARC ---------------------------------------------------------------- */

/* Replaces the per-tensor loop of LARC.step (apex/parallel/LARC.py:68-97) for one
 * param group: norms of every parameter and gradient, the adaptive rate
 * trust * |p| / (|g| + |p| * wd + eps) (divided by lr and clamped to 1 when clip != 0),
 * and g = (g + wd * p) * rate, for all tensors at once and without any host read
 * (the reference reads two norms per tensor back to compare them with 0; a tensor
 * with a zero norm is left untouched here too).  Tensors are fp32 where
 * TF_GRAD_F32 is set, else fp16 (gradient and parameter share the dtype).
 * d_scratch: >= 2 * total_chunks + n_tensors floats, caller-allocated.
 * d_norms: optional output, [n_tensors][2] = (|p|, |g|). */
int apex_b200_mt_larc(const apex_b200_mt_table* t, int grad_list, int param_list,
                      double trust_coefficient, double eps, double weight_decay,
                      double lr, int clip, float* d_scratch, int64_t scratch_len,
                      float* d_norms, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* APEX_B200_H_ */
