// tools/tune_step.cu -- standalone sweep of launch/caching variants of the fused Adam pass.
// Development tool (not part of the library): same memory access structure as
// apex_b200/csrc/mt_fused_step.cu on one large aligned tensor, timed with CUDA events.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/tune_step tools/tune_step.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <algorithm>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

struct F8 { float v[8]; };
struct H8 { uint32_t w[4]; };

// LD policy: 0 = L1::no_allocate, 1 = default, 2 = L1::no_allocate + L2::evict_first, 3 = nc + no_allocate
template <int LD> __device__ __forceinline__ F8 ld_f8(const float* p) {
    F8 r;
    if (LD == 0)
        asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7]) : "l"(p));
    else if (LD == 1)
        asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7]) : "l"(p));
    else if (LD == 2)
        asm volatile("ld.global.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7]) : "l"(p));
    else
        asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7]) : "l"(p));
    return r;
}
template <int LD> __device__ __forceinline__ H8 ld_h8(const __half* p) {
    H8 r;
    if (LD == 0 || LD == 3)
        asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]) : "l"(p));
    else if (LD == 1)
        asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]) : "l"(p));
    else
        asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]) : "l"(p));
    return r;
}
// ST policy: 0 = L1::no_allocate, 1 = default, 2 = .cs, 3 = L1::no_allocate.L2::evict_first
template <int ST> __device__ __forceinline__ void st_f8(float* p, const F8& r) {
    if (ST == 0)
        asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7]) : "memory");
    else if (ST == 1)
        asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7]) : "memory");
    else if (ST == 2) {
        asm volatile("st.global.cs.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]) : "memory");
        asm volatile("st.global.cs.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p + 4), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7]) : "memory");
    } else
        asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7]) : "memory");
}
template <int ST> __device__ __forceinline__ void st_h8(__half* p, const H8& r) {
    if (ST == 0)
        asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(r.w[0]), "r"(r.w[1]), "r"(r.w[2]), "r"(r.w[3]) : "memory");
    else if (ST == 1)
        asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(r.w[0]), "r"(r.w[1]), "r"(r.w[2]), "r"(r.w[3]) : "memory");
    else if (ST == 2)
        asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(r.w[0]), "r"(r.w[1]), "r"(r.w[2]), "r"(r.w[3]) : "memory");
    else
        asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(r.w[0]), "r"(r.w[1]), "r"(r.w[2]), "r"(r.w[3]) : "memory");
}
__device__ __forceinline__ F8 h8_to_f8(const H8& h) {
    F8 f;
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 f2 = __half22float2(*reinterpret_cast<const __half2*>(&h.w[i])); f.v[2*i] = f2.x; f.v[2*i+1] = f2.y; }
    return f;
}
__device__ __forceinline__ H8 f8_to_h8(const F8& f) {
    H8 h;
#pragma unroll
    for (int i = 0; i < 4; ++i) { __half2 h2 = __floats2half2_rn(f.v[2*i], f.v[2*i+1]); h.w[i] = *reinterpret_cast<const uint32_t*>(&h2); }
    return h;
}
struct Hp { float beta2, omb2, w, eps, step_size, bc2, inv; };
__device__ __forceinline__ void adam_elem(float g, float& p, float& m, float& v, const Hp& h) {
    g = __fmul_rn(g, h.inv);
    m = __fmaf_rn(h.w, __fsub_rn(g, m), m);
    v = __fmul_rn(v, h.beta2);
    v = __fmaf_rn(h.omb2, __fmul_rn(g, g), v);
    float d = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), h.bc2), h.eps);
    p = __fmaf_rn(h.step_size, __fdiv_rn(m, d), p);
}

// Plain variant: persistent CTAs over chunks, U vectors loaded up front per thread.
template <int T, int MINB, int U, int LD, int ST>
__global__ void __launch_bounds__(T, MINB)
adam_plain(const __half* __restrict__ g, float* p, float* m, float* v, __half* q, long long n, int chunk, Hp h) {
    const long long nchunks = (n + chunk - 1) / chunk;
    for (long long c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const long long off = c * chunk;
        const int len = (int)((n - off) < chunk ? (n - off) : chunk);
        const int nvec = len >> 3;
        int i = threadIdx.x;
#pragma unroll 1
        for (; i + (U - 1) * T < nvec; i += U * T) {
            F8 gv[U], pv[U], mv[U], sv[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const long long e = off + (long long)(i + u * T) * 8;
                gv[u] = h8_to_f8(ld_h8<LD>(g + e)); pv[u] = ld_f8<LD>(p + e); mv[u] = ld_f8<LD>(m + e); sv[u] = ld_f8<LD>(v + e);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const long long e = off + (long long)(i + u * T) * 8;
#pragma unroll
                for (int k = 0; k < 8; ++k) adam_elem(gv[u].v[k], pv[u].v[k], mv[u].v[k], sv[u].v[k], h);
                st_f8<ST>(p + e, pv[u]); st_f8<ST>(m + e, mv[u]); st_f8<ST>(v + e, sv[u]); st_h8<ST>(q + e, f8_to_h8(pv[u]));
            }
        }
    }
}

// Software-pipelined variant: loads of iteration i+1 are issued before the math of iteration i.
template <int T, int MINB, int LD, int ST>
__global__ void __launch_bounds__(T, MINB)
adam_pipe(const __half* __restrict__ g, float* p, float* m, float* v, __half* q, long long n, int chunk, Hp h) {
    const long long nvec_total = n >> 3;
    const long long stride = (long long)gridDim.x * T;
    long long i = (long long)blockIdx.x * T + threadIdx.x;
    if (i >= nvec_total) return;
    H8 gn = ld_h8<LD>(g + i * 8); F8 pn = ld_f8<LD>(p + i * 8), mn = ld_f8<LD>(m + i * 8), sn = ld_f8<LD>(v + i * 8);
#pragma unroll 1
    for (; i < nvec_total; i += stride) {
        F8 gv = h8_to_f8(gn), pv = pn, mv = mn, sv = sn;
        const long long nx = i + stride;
        if (nx < nvec_total) { gn = ld_h8<LD>(g + nx * 8); pn = ld_f8<LD>(p + nx * 8); mn = ld_f8<LD>(m + nx * 8); sn = ld_f8<LD>(v + nx * 8); }
#pragma unroll
        for (int k = 0; k < 8; ++k) adam_elem(gv.v[k], pv.v[k], mv.v[k], sv.v[k], h);
        st_f8<ST>(p + i * 8, pv); st_f8<ST>(m + i * 8, mv); st_f8<ST>(v + i * 8, sv); st_h8<ST>(q + i * 8, f8_to_h8(pv));
    }
}

struct Bufs { __half* g; float* p; float* m; float* v; __half* q; long long n; };

template <typename F>
float time_it(F launch, int reps = 5) {
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    launch(); launch(); CK(cudaDeviceSynchronize());
    std::vector<float> t;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(a)); launch(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        float ms; CK(cudaEventElapsedTime(&ms, a, b)); t.push_back(ms);
    }
    CK(cudaGetLastError());
    std::sort(t.begin(), t.end());
    return t[t.size() / 2];
}

static int g_sms = 148;
static Hp g_hp = {0.999f, 0.001f, 0.1f, 1e-8f, -1e-3f, 0.0316f, 1.f / 65536};

template <int T, int MINB, int U, int LD, int ST>
void run_plain(const Bufs& b, int per_sm, int chunk, const char* tag) {
    int grid = g_sms * per_sm;
    float ms = time_it([&] { adam_plain<T, MINB, U, LD, ST><<<grid, T>>>(b.g, b.p, b.m, b.v, b.q, b.n, chunk, g_hp); });
    printf("plain T=%d minb=%d U=%d LD=%d ST=%d per_sm=%d chunk=%d %s: %.3f ms  %.1f GB/s\n", T, MINB, U, LD, ST, per_sm, chunk, tag, ms, 28.0 * b.n / ms / 1e6);
}
template <int T, int MINB, int LD, int ST>
void run_pipe(const Bufs& b, int per_sm) {
    int grid = g_sms * per_sm;
    float ms = time_it([&] { adam_pipe<T, MINB, LD, ST><<<grid, T>>>(b.g, b.p, b.m, b.v, b.q, b.n, 0, g_hp); });
    printf("pipe  T=%d minb=%d LD=%d ST=%d per_sm=%d: %.3f ms  %.1f GB/s\n", T, MINB, LD, ST, per_sm, ms, 28.0 * b.n / ms / 1e6);
}

int main(int argc, char** argv) {
    long long n = 1LL << 29;   // 512 Mi elements: 15 GB of traffic per pass
    if (argc > 1) n = atoll(argv[1]);
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0)); g_sms = prop.multiProcessorCount;
    printf("%s, %d SMs, n=%lld\n", prop.name, g_sms, n);
    Bufs b; b.n = n;
    CK(cudaMalloc(&b.g, n * 2)); CK(cudaMalloc(&b.p, n * 4)); CK(cudaMalloc(&b.m, n * 4)); CK(cudaMalloc(&b.v, n * 4)); CK(cudaMalloc(&b.q, n * 2));
    CK(cudaMemset(b.g, 0x11, n * 2)); CK(cudaMemset(b.p, 0, n * 4)); CK(cudaMemset(b.m, 0, n * 4)); CK(cudaMemset(b.v, 0, n * 4));
    // memcpy reference (the MEASURED_PEAKS number is a torch copy_: read+write bytes)
    { float ms = time_it([&] { CK(cudaMemcpyAsync(b.m, b.v, n * 4, cudaMemcpyDeviceToDevice)); }); printf("cudaMemcpy D2D %lld B: %.3f ms  %.1f GB/s (r+w)\n", n * 4, ms, 2.0 * n * 4 / ms / 1e6); CK(cudaMemset(b.m, 0, n * 4)); }

    int sweep = argc > 2 ? atoi(argv[2]) : 2;
    if (sweep == 2) {
        run_plain<256, 4, 1, 0, 0>(b, 4, 16384, "(library)");
        run_plain<256, 2, 1, 0, 0>(b, 2, 16384, "");
        run_plain<256, 2, 1, 0, 0>(b, 1, 16384, "");
        run_plain<256, 1, 1, 0, 0>(b, 1, 16384, "");
        run_plain<128, 2, 1, 0, 0>(b, 2, 16384, "");
        run_plain<128, 3, 1, 0, 0>(b, 3, 16384, "");
        run_plain<128, 4, 1, 0, 0>(b, 4, 16384, "");
        run_plain<384, 1, 1, 0, 0>(b, 1, 16384, "");
        run_plain<512, 1, 1, 0, 0>(b, 1, 16384, "");
        run_plain<768, 1, 1, 0, 0>(b, 1, 16384, "");
        run_plain<256, 2, 2, 0, 0>(b, 2, 16384, "");
        run_plain<256, 1, 2, 0, 0>(b, 1, 16384, "");
        run_plain<256, 1, 4, 0, 0>(b, 1, 16384, "");
        run_plain<128, 2, 2, 0, 0>(b, 2, 16384, "");
        run_plain<128, 1, 4, 0, 0>(b, 1, 16384, "");
        run_plain<128, 2, 4, 0, 0>(b, 2, 16384, "");
        run_plain<64, 4, 2, 0, 0>(b, 4, 16384, "");
        run_plain<64, 8, 1, 0, 0>(b, 8, 16384, "");
        run_plain<256, 2, 1, 0, 0>(b, 2, 8192, "");
        run_plain<256, 2, 1, 0, 0>(b, 2, 32768, "");
        run_plain<256, 2, 1, 0, 0>(b, 2, 65536, "");
        run_plain<256, 2, 1, 0, 0>(b, 2, 2048, "");
        run_plain<256, 2, 1, 0, 0>(b, 2, 4096, "");
        run_plain<256, 2, 1, 2, 3>(b, 2, 16384, "evict_first");
        run_plain<256, 2, 1, 1, 1>(b, 2, 16384, "default");
        // non persistent: one CTA per chunk, hardware scheduler, 2 resident per SM
        { long long nch = (n + 16383) / 16384; float ms = time_it([&] { adam_plain<256, 2, 1, 0, 0><<<(unsigned)nch, 256>>>(b.g, b.p, b.m, b.v, b.q, b.n, 16384, g_hp); }); printf("nonpersistent T=256 minb=2 chunk=16384: %.3f ms %.1f GB/s\n", ms, 28.0 * n / ms / 1e6); }
        { long long nch = (n + 4095) / 4096; float ms = time_it([&] { adam_plain<256, 2, 1, 0, 0><<<(unsigned)nch, 256>>>(b.g, b.p, b.m, b.v, b.q, b.n, 4096, g_hp); }); printf("nonpersistent T=256 minb=2 chunk=4096: %.3f ms %.1f GB/s\n", ms, 28.0 * n / ms / 1e6); }
        run_pipe<256, 1, 0, 0>(b, 1);
        run_pipe<128, 2, 0, 0>(b, 2);
        run_pipe<128, 3, 0, 0>(b, 3);
        run_pipe<384, 1, 0, 0>(b, 1);
    }
    if (sweep == 3) {
#define NP(T, MINB, U, CH) { long long nch = (n + (CH) - 1) / (CH); float ms = time_it([&] { adam_plain<T, MINB, U, 0, 0><<<(unsigned)nch, T>>>(b.g, b.p, b.m, b.v, b.q, b.n, CH, g_hp); }); printf("nonpersistent T=%d minb=%d U=%d chunk=%d: %.3f ms %.1f GB/s\n", T, MINB, U, CH, ms, 28.0 * n / ms / 1e6); }
        NP(256, 2, 1, 2048) NP(256, 2, 1, 4096) NP(256, 2, 1, 8192) NP(256, 2, 1, 16384)
        NP(256, 4, 1, 2048) NP(256, 4, 1, 4096) NP(256, 4, 1, 8192)
        NP(256, 1, 1, 2048) NP(256, 1, 1, 4096)
        NP(128, 4, 1, 1024) NP(128, 4, 1, 2048) NP(128, 4, 1, 4096) NP(128, 8, 1, 2048) NP(128, 8, 1, 4096)
        NP(512, 1, 1, 4096) NP(512, 1, 1, 8192) NP(512, 2, 1, 4096) NP(512, 2, 1, 8192)
        NP(1024, 1, 1, 8192) NP(1024, 1, 1, 16384)
        NP(256, 2, 2, 4096) NP(256, 2, 2, 8192) NP(128, 4, 2, 2048) NP(128, 4, 2, 4096) NP(128, 2, 4, 4096) NP(256, 1, 4, 8192)
        NP(64, 8, 1, 512) NP(64, 8, 1, 2048) NP(64, 8, 2, 1024) NP(64, 4, 4, 2048)
    }
    return 0;
}
