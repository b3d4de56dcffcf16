// sfc_group_* — one host process driving several B200s (the shape a Go VSP has): a handle per GPU, the
// table image replicated with ncclBroadcast, the counters summed with ncclAllReduce.  Written on the
// public C ABI only (sfc_table_export_device / _import_device, sfc_counters_device_ptr); NCCL is
// opened with dlopen at group creation so that the library carries no load-time dependency on it (a
// Python process that has torch loaded keeps using torch's copy).  Neither collective is on the
// per-batch path: packets are sharded, every GPU classifies its own range.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <errno.h>
#include <nccl.h>

#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/b200sfc.h"

namespace {
struct Nccl {
    void* lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool open() {
        if (lib) return true;
        lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) return false;
#define SFC_NCCL_SYM(field, name) *(void**)(&field) = dlsym(lib, name)
        SFC_NCCL_SYM(CommInitAll, "ncclCommInitAll");
        SFC_NCCL_SYM(CommDestroy, "ncclCommDestroy");
        SFC_NCCL_SYM(GroupStart, "ncclGroupStart");
        SFC_NCCL_SYM(GroupEnd, "ncclGroupEnd");
        SFC_NCCL_SYM(Broadcast, "ncclBroadcast");
        SFC_NCCL_SYM(AllReduce, "ncclAllReduce");
        SFC_NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef SFC_NCCL_SYM
        return CommInitAll && CommDestroy && GroupStart && GroupEnd && Broadcast && AllReduce && GetErrorString;
    }
};
Nccl g_nccl;
}  // namespace

struct sfc_group {
    std::vector<int> dev;
    std::vector<sfc_handle*> h;
    std::vector<ncclComm_t> comm;
    std::vector<cudaStream_t> st;
    std::string err;
    float last_broadcast_ms = 0.f, last_reduce_ms = 0.f;
};

namespace {
int gfail(sfc_group* g, int code, const char* fmt, ...) {
    char buf[384];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (g) g->err = buf;
    return code;
}
#define G_CUDA(g, call)                                                                          \
    do {                                                                                         \
        cudaError_t e_ = (call);                                                                 \
        if (e_ != cudaSuccess) return gfail(g, -EIO, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)
#define G_NCCL(g, call)                                                                               \
    do {                                                                                              \
        ncclResult_t r_ = (call);                                                                     \
        if (r_ != ncclSuccess) return gfail(g, -EIO, "%s: %s", #call, g_nccl.GetErrorString(r_)); \
    } while (0)

struct DevScope {  // leave the caller's current device as it was
    int prev = 0;
    DevScope() { cudaGetDevice(&prev); }
    ~DevScope() { cudaSetDevice(prev); }
};
}  // namespace

extern "C" {

int sfc_group_create(int n_gpus, const int* device_ordinals, sfc_group** out) {
    if (!out) return -EINVAL;
    *out = nullptr;
    if (n_gpus < 1 || n_gpus > 64) return -EINVAL;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return -ENODEV;  // no CPU stand-in
    sfc_group* g = new sfc_group();
    DevScope scope;
    for (int i = 0; i < n_gpus; ++i) {
        int d = device_ordinals ? device_ordinals[i] : i;
        for (int p : g->dev)
            if (p == d) {
                sfc_group_destroy(g);
                return -EINVAL;  // one handle per GPU
            }
        sfc_handle* h = nullptr;
        int rc = sfc_create(d, &h);
        if (rc) {
            sfc_group_destroy(g);
            return rc;
        }
        g->dev.push_back(d);
        g->h.push_back(h);
        cudaStream_t s = nullptr;
        if (cudaSetDevice(d) != cudaSuccess || cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) {
            sfc_group_destroy(g);
            return -EIO;
        }
        g->st.push_back(s);
    }
    if (n_gpus > 1) {
        if (!g_nccl.open()) {
            sfc_group_destroy(g);
            return -ENOSYS;  // libnccl.so.2 not found
        }
        g->comm.assign((size_t)n_gpus, nullptr);
        if (g_nccl.CommInitAll(g->comm.data(), n_gpus, g->dev.data()) != ncclSuccess) {
            g->comm.clear();
            sfc_group_destroy(g);
            return -EIO;
        }
    }
    *out = g;
    return 0;
}

int sfc_group_destroy(sfc_group* g) {
    if (!g) return -EINVAL;
    DevScope scope;
    for (size_t i = 0; i < g->comm.size(); ++i)
        if (g->comm[i]) g_nccl.CommDestroy(g->comm[i]);
    for (size_t i = 0; i < g->st.size(); ++i) {
        cudaSetDevice(g->dev[i]);
        if (g->st[i]) cudaStreamDestroy(g->st[i]);
    }
    for (sfc_handle* h : g->h)
        if (h) sfc_destroy(h);
    delete g;
    return 0;
}

int sfc_group_size(const sfc_group* g) { return g ? (int)g->h.size() : -EINVAL; }

sfc_handle* sfc_group_handle(sfc_group* g, int rank) {
    return (g && rank >= 0 && rank < (int)g->h.size()) ? g->h[(size_t)rank] : nullptr;
}

const char* sfc_group_last_error(const sfc_group* g) { return g ? g->err.c_str() : "null group"; }

float sfc_group_last_ms(const sfc_group* g, int which) {
    return !g ? -1.f : (which == 0 ? g->last_broadcast_ms : g->last_reduce_ms);
}

int sfc_group_broadcast_table(sfc_group* g) {
    if (!g) return -EINVAL;
    const int n = (int)g->h.size();
    const size_t bytes = sfc_table_size(g->h[0]);
    if (!bytes) return gfail(g, -ENOENT, "sfc_group_broadcast_table: rank 0 has no committed table");
    if (n == 1) return 0;
    DevScope scope;
    std::vector<void*> buf((size_t)n, nullptr);
    auto release = [&]() {
        for (int r = 0; r < n; ++r)
            if (buf[(size_t)r]) {
                cudaSetDevice(g->dev[(size_t)r]);
                cudaFree(buf[(size_t)r]);
            }
    };
    for (int r = 0; r < n; ++r) {
        cudaError_t e = cudaSetDevice(g->dev[(size_t)r]);
        if (e == cudaSuccess) e = cudaMalloc(&buf[(size_t)r], bytes);
        if (e != cudaSuccess) {
            release();
            return gfail(g, -ENOMEM, "staging %zu bytes on GPU %d: %s", bytes, g->dev[(size_t)r], cudaGetErrorString(e));
        }
    }
    int rc = sfc_table_export_device(g->h[0], buf[0], bytes, g->st[0]);
    if (rc) {
        release();
        return gfail(g, rc, "export on rank 0: %s", sfc_last_error(g->h[0]));
    }
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaSetDevice(g->dev[0]);
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, g->st[0]);
    ncclResult_t nr = g_nccl.GroupStart();
    for (int r = 0; r < n && nr == ncclSuccess; ++r)
        nr = g_nccl.Broadcast(buf[(size_t)r], buf[(size_t)r], bytes, ncclUint8, 0, g->comm[(size_t)r], g->st[(size_t)r]);
    ncclResult_t ne = g_nccl.GroupEnd();
    if (nr == ncclSuccess) nr = ne;
    cudaEventRecord(e1, g->st[0]);
    for (int r = 0; r < n && nr == ncclSuccess && rc == 0; ++r) {
        cudaSetDevice(g->dev[(size_t)r]);
        if (cudaStreamSynchronize(g->st[(size_t)r]) != cudaSuccess) rc = gfail(g, -EIO, "broadcast failed on GPU %d", g->dev[(size_t)r]);
    }
    if (nr == ncclSuccess && rc == 0) {
        cudaSetDevice(g->dev[0]);
        cudaEventElapsedTime(&g->last_broadcast_ms, e0, e1);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    for (int r = 1; r < n && nr == ncclSuccess && rc == 0; ++r) {
        rc = sfc_table_import_device(g->h[(size_t)r], buf[(size_t)r], bytes, g->st[(size_t)r]);
        if (rc) gfail(g, rc, "import on rank %d: %s", r, sfc_last_error(g->h[(size_t)r]));
    }
    release();
    if (nr != ncclSuccess) return gfail(g, -EIO, "ncclBroadcast: %s", g_nccl.GetErrorString(nr));
    return rc;
}

int sfc_group_reduce_counters(sfc_group* g) {
    if (!g) return -EINVAL;
    const int n = (int)g->h.size();
    if (n == 1) return 0;
    DevScope scope;
    for (int r = 0; r < n; ++r) {  // counters of launches still in flight must have landed
        G_CUDA(g, cudaSetDevice(g->dev[(size_t)r]));
        G_CUDA(g, cudaDeviceSynchronize());
    }
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaSetDevice(g->dev[0]);
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, g->st[0]);
    ncclResult_t nr = g_nccl.GroupStart();
    for (int r = 0; r < n && nr == ncclSuccess; ++r) {
        void* p = sfc_counters_device_ptr(g->h[(size_t)r]);
        nr = g_nccl.AllReduce(p, p, SFC_CNT_TOTAL, ncclUint64, ncclSum, g->comm[(size_t)r], g->st[(size_t)r]);
    }
    ncclResult_t ne = g_nccl.GroupEnd();
    if (nr == ncclSuccess) nr = ne;
    cudaEventRecord(e1, g->st[0]);
    int rc = 0;
    for (int r = 0; r < n; ++r) {
        cudaSetDevice(g->dev[(size_t)r]);
        if (cudaStreamSynchronize(g->st[(size_t)r]) != cudaSuccess && rc == 0) rc = gfail(g, -EIO, "all-reduce failed on GPU %d", g->dev[(size_t)r]);
    }
    if (rc == 0 && nr == ncclSuccess) {
        cudaSetDevice(g->dev[0]);
        cudaEventElapsedTime(&g->last_reduce_ms, e0, e1);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (nr != ncclSuccess) return gfail(g, -EIO, "ncclAllReduce: %s", g_nccl.GetErrorString(nr));
    return rc;
}

}  // extern "C"
