// mt_common.cuh -- the multi-tensor-apply launcher shared by every kernel.
//
// One launch walks a device-resident table of tensors (see include/apex_b200.h
// for the layout).  Work is cut into chunks of `chunk_elems` elements that
// never straddle a tensor, ONE CTA PER CHUNK: measured on B200, short-lived
// CTAs that each move one 2048-element tile (one 8-element vector per thread)
// reach 6.9 TB/s on the 8-stream Adam pass where persistent grid-stride CTAs
// stall at 6.0 TB/s (tools/tune_step.cu; the resident CTAs drift out of phase
// instead of all reading, then all writing).  A CTA finds its tensor with one
// load from the chunk->tensor map; the table is rebuilt only when a pointer
// changes.
//
// Memory access: every stream is read/written exactly once, so loads bypass L1
// allocation and all fp32 streams move as 256-bit vectors (LDG.E.ENL2.256 /
// STG.E.ENL2.256, new on sm_100), fp16 streams as 128-bit vectors: a thread
// owns 8 consecutive elements, a warp 256 -- each warp-level request is one
// fully used, contiguous 1 KiB (fp32) or 512 B (fp16) span.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/apex_b200.h"

namespace apex_b200 {

constexpr int kThreads = 256;          // threads per CTA for the streaming kernels
constexpr int kVec = 8;                // elements per thread per vector step
constexpr int kTile = kThreads * kVec; // elements per CTA per vector step (2048)

// ---- device view of the packed table --------------------------------------

struct MtView {
    const int64_t* numel;       // [n]
    const int64_t* elem_start;  // [n+1]
    const int64_t* ptr;         // [n_lists][n]
    const int32_t* flags;       // [n]
    const int32_t* chunk_start; // [n+1]
    const int32_t* chunk_tensor; // [total_chunks]
    int n;
    int n_lists;
    int chunk_elems;
    int total_chunks;
};

__host__ __device__ inline size_t mt_align8(size_t x) { return (x + 7) & ~size_t(7); }

__host__ __device__ inline size_t mt_off_numel() { return 16; }
__host__ __device__ inline size_t mt_off_elem_start(int n) { return mt_off_numel() + size_t(n) * 8; }
__host__ __device__ inline size_t mt_off_ptr(int n) { return mt_off_elem_start(n) + size_t(n + 1) * 8; }
__host__ __device__ inline size_t mt_off_flags(int n, int l) { return mt_off_ptr(n) + size_t(n) * l * 8; }
__host__ __device__ inline size_t mt_off_chunk_start(int n, int l) { return mt_align8(mt_off_flags(n, l) + size_t(n) * 4); }
__host__ __device__ inline size_t mt_off_chunk_tensor(int n, int l) { return mt_align8(mt_off_chunk_start(n, l) + size_t(n + 1) * 4); }
__host__ __device__ inline size_t mt_total_bytes(int n, int l, long long chunks) { return mt_align8(mt_off_chunk_tensor(n, l) + size_t(chunks) * 4); }

inline MtView mt_view(const apex_b200_mt_table* t) {
    const char* b = static_cast<const char*>(t->d_table);
    MtView v;
    v.n = t->n_tensors;
    v.n_lists = t->n_lists;
    v.chunk_elems = t->chunk_elems;
    v.total_chunks = t->total_chunks;
    v.numel = reinterpret_cast<const int64_t*>(b + mt_off_numel());
    v.elem_start = reinterpret_cast<const int64_t*>(b + mt_off_elem_start(v.n));
    v.ptr = reinterpret_cast<const int64_t*>(b + mt_off_ptr(v.n));
    v.flags = reinterpret_cast<const int32_t*>(b + mt_off_flags(v.n, v.n_lists));
    v.chunk_start = reinterpret_cast<const int32_t*>(b + mt_off_chunk_start(v.n, v.n_lists));
    v.chunk_tensor = reinterpret_cast<const int32_t*>(b + mt_off_chunk_tensor(v.n, v.n_lists));
    return v;
}

// What a CTA knows about the chunk it is working on.
struct MtChunk {
    int tensor;      // index in the table
    int flags;       // APEX_B200_TF_* of that tensor
    int64_t offset;  // first element of the chunk inside the tensor
    int len;         // elements in this chunk (<= chunk_elems)
    int64_t numel;   // elements in the tensor
};

// Uniform across the CTA: one load for the tensor index, then the per-tensor
// fields in parallel (all L1/L2 resident after the first CTAs).
__device__ __forceinline__ MtChunk mt_locate(const MtView& v, int chunk) {
    MtChunk c;
    c.tensor = __ldg(v.chunk_tensor + chunk);
    c.flags = __ldg(v.flags + c.tensor);
    c.numel = __ldg(v.numel + c.tensor);
    c.offset = int64_t(chunk - __ldg(v.chunk_start + c.tensor)) * v.chunk_elems;
    int64_t left = c.numel - c.offset;
    c.len = left < v.chunk_elems ? int(left) : v.chunk_elems;
    return c;
}

template <typename T>
__device__ __forceinline__ T* mt_ptr(const MtView& v, int list, int tensor) {
    return reinterpret_cast<T*>(__ldg(v.ptr + size_t(list) * v.n + tensor));
}

// ---- 8-element vectors -----------------------------------------------------

struct F8 { float v[8]; };
struct H8 { uint32_t w[4]; };  // 8 halfs packed in 4 words

// 256-bit streaming load of 8 floats (needs 32 B alignment).
__device__ __forceinline__ F8 ld_f8(const float* p) {
    F8 r;
    asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]),
                   "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_f8(float* p, const F8& r) {
    asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]),
                    "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7])
                 : "memory");
}
// 128-bit streaming load of 8 halfs (needs 16 B alignment).
__device__ __forceinline__ H8 ld_h8(const __half* p) {
    H8 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3])
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_h8(__half* p, const H8& r) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(r.w[0]), "r"(r.w[1]), "r"(r.w[2]), "r"(r.w[3])
                 : "memory");
}

__device__ __forceinline__ F8 h8_to_f8(const H8& h) {
    F8 f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        __half2 h2 = *reinterpret_cast<const __half2*>(&h.w[i]);
        float2 f2 = __half22float2(h2);
        f.v[2 * i] = f2.x;
        f.v[2 * i + 1] = f2.y;
    }
    return f;
}
__device__ __forceinline__ H8 f8_to_h8(const F8& f) {
    H8 h;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        __half2 h2 = __floats2half2_rn(f.v[2 * i], f.v[2 * i + 1]);
        h.w[i] = *reinterpret_cast<const uint32_t*>(&h2);
    }
    return h;
}

// Gradient loads: the same code path serves fp16 and fp32 gradients.
template <typename G> struct GradIo;
template <> struct GradIo<__half> {
    static __device__ __forceinline__ F8 ld8(const __half* p) { return h8_to_f8(ld_h8(p)); }
    static __device__ __forceinline__ float ld1(const __half* p) { return __half2float(*p); }
};
template <> struct GradIo<float> {
    static __device__ __forceinline__ F8 ld8(const float* p) { return ld_f8(p); }
    static __device__ __forceinline__ float ld1(const float* p) { return *p; }
};

__device__ __forceinline__ bool is_nonfinite(float x) {
    return (__float_as_uint(x) & 0x7f800000u) == 0x7f800000u;
}

// Warp ballot, then a block-wide OR; thread 0 raises the flag to
// INT32_MAX - tensor so that the word ends as the first bad tensor.
__device__ __forceinline__ void mt_raise_flag(bool bad, int tensor, int32_t* flag) {
    const bool warp_bad = __any_sync(0xffffffffu, bad);
    const int block_bad = __syncthreads_or(warp_bad ? 1 : 0);
    if (block_bad && threadIdx.x == 0) atomicMax(flag, 0x7fffffff - tensor);
}

// Grid sizing: one CTA per chunk (see the header comment).
int mt_sm_count();
inline int mt_grid(int total_chunks, int /*unused*/) { return total_chunks; }

inline int mt_cuda_status() {
    cudaError_t e = cudaPeekAtLastError();
    return e == cudaSuccess ? 0 : int(e);
}

}  // namespace apex_b200
