// c_api.cu -- housekeeping entry points and the host-side table packer.
#include <string.h>

#include "mt_common.cuh"

namespace apex_b200 {

int mt_sm_count() {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) return 148;
    return sms;
}

}  // namespace apex_b200

using namespace apex_b200;

extern "C" int apex_b200_version(void) { return APEX_B200_VERSION; }

extern "C" const char* apex_b200_error_string(int code) {
    switch (code) {
        case 0: return "ok";
        case APEX_B200_EINVAL: return "apex_b200: invalid argument";
        case APEX_B200_ERANGE: return "apex_b200: size beyond the supported maximum";
        case APEX_B200_ENOSPC: return "apex_b200: caller buffer too small";
        default: break;
    }
    if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
    return "apex_b200: unknown error";
}

extern "C" int apex_b200_device_info(int* sm_count, int* cc_major, int* cc_minor) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return int(e);
    int v = 0;
    if (sm_count) {
        e = cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return int(e);
        *sm_count = v;
    }
    if (cc_major) {
        e = cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMajor, dev);
        if (e != cudaSuccess) return int(e);
        *cc_major = v;
    }
    if (cc_minor) {
        e = cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMinor, dev);
        if (e != cudaSuccess) return int(e);
        *cc_minor = v;
    }
    return 0;
}

extern "C" int64_t apex_b200_mt_count_chunks(int n_tensors, const int64_t* numels, int chunk_elems) {
    if (n_tensors < 0 || chunk_elems <= 0 || (n_tensors > 0 && numels == nullptr)) return APEX_B200_EINVAL;
    int64_t chunks = 0;
    for (int i = 0; i < n_tensors; ++i) {
        if (numels[i] < 0) return APEX_B200_EINVAL;
        chunks += (numels[i] + chunk_elems - 1) / chunk_elems;
    }
    return chunks;
}

extern "C" size_t apex_b200_mt_table_bytes(int n_tensors, int n_lists, int64_t total_chunks) {
    if (n_tensors < 0 || n_lists < 1 || total_chunks < 0) return 0;
    return mt_total_bytes(n_tensors, n_lists, total_chunks);
}

extern "C" int apex_b200_mt_table_pack(void* host_buf, size_t host_buf_bytes, int n_tensors, int n_lists,
                                       const int64_t* numels, const uint64_t* ptrs, const int32_t* flags,
                                       int chunk_elems, apex_b200_mt_table* out) {
    if (host_buf == nullptr || out == nullptr || n_tensors < 0 || n_lists < 1) return APEX_B200_EINVAL;
    if (n_tensors > 0 && (numels == nullptr || ptrs == nullptr)) return APEX_B200_EINVAL;
    if (chunk_elems <= 0 || chunk_elems % kTile != 0) return APEX_B200_EINVAL;
    const int64_t want_chunks = apex_b200_mt_count_chunks(n_tensors, numels, chunk_elems);
    if (want_chunks < 0) return APEX_B200_EINVAL;
    if (want_chunks > 0x7fffffff) return APEX_B200_ERANGE;
    const size_t need = mt_total_bytes(n_tensors, n_lists, want_chunks);
    if (host_buf_bytes < need) return APEX_B200_ENOSPC;

    char* b = static_cast<char*>(host_buf);
    memset(b, 0, need);
    int32_t* header = reinterpret_cast<int32_t*>(b);
    int64_t* t_numel = reinterpret_cast<int64_t*>(b + mt_off_numel());
    int64_t* t_estart = reinterpret_cast<int64_t*>(b + mt_off_elem_start(n_tensors));
    int64_t* t_ptr = reinterpret_cast<int64_t*>(b + mt_off_ptr(n_tensors));
    int32_t* t_flags = reinterpret_cast<int32_t*>(b + mt_off_flags(n_tensors, n_lists));
    int32_t* t_cstart = reinterpret_cast<int32_t*>(b + mt_off_chunk_start(n_tensors, n_lists));
    int32_t* t_ctensor = reinterpret_cast<int32_t*>(b + mt_off_chunk_tensor(n_tensors, n_lists));

    int64_t elems = 0, chunks = 0;
    for (int i = 0; i < n_tensors; ++i) {
        if (numels[i] < 0) return APEX_B200_EINVAL;
        t_numel[i] = numels[i];
        t_estart[i] = elems;
        t_cstart[i] = int32_t(chunks);
        elems += numels[i];
        const int64_t mine = (numels[i] + chunk_elems - 1) / chunk_elems;
        for (int64_t c = 0; c < mine; ++c) t_ctensor[chunks + c] = i;
        chunks += mine;
        int32_t f = flags ? (flags[i] & ~APEX_B200_TF_UNALIGNED) : 0;
        for (int l = 0; l < n_lists; ++l) {
            const uint64_t p = ptrs[size_t(l) * n_tensors + i];
            t_ptr[size_t(l) * n_tensors + i] = int64_t(p);
            if (p & 31) f |= APEX_B200_TF_UNALIGNED;
        }
        t_flags[i] = f;
    }
    t_estart[n_tensors] = elems;
    t_cstart[n_tensors] = int32_t(chunks);
    header[0] = n_tensors;
    header[1] = n_lists;
    header[2] = chunk_elems;
    header[3] = int32_t(chunks);

    out->d_table = nullptr;
    out->n_tensors = n_tensors;
    out->n_lists = n_lists;
    out->chunk_elems = chunk_elems;
    out->total_chunks = int32_t(chunks);
    return 0;
}
