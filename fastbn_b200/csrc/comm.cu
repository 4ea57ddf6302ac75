// fbn_comm / fbn_reduce (include/fastbn_b200.h): the ONE exchange step of the multi-GPU path — the sum over ranks of
// {n_correct, n_total, sum_mse, sum_hd} — as one ncclAllReduce of four doubles over NVLink.
// The reference has no counterpart (its USE_MPI flag is a dead stub, /root/reference/CMakeLists.txt:10-16).
//
// NCCL is bound at run time (dlopen of libnccl.so.2, preferring a copy the process has already loaded, e.g. the one bundled
// with PyTorch) so that a host process never ends up with two different NCCL builds behind one soname; only the few entry
// points used here are declared, with the ABI of nccl.h 2.x (ncclUniqueId = 128 bytes, ncclFloat64 = 8, ncclSum = 0).
#include <dlfcn.h>

#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/fastbn_b200.h"

int fbn_set_error(int code, const std::string &msg);  // capi.cu

namespace {

typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { kNcclSuccess = 0, kNcclFloat64 = 8, kNcclSum = 0 };

struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    int (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int *) = nullptr;
    std::string error;
};

NcclApi &nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {  // a copy that is already mapped into the process wins
            api.handle = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle)
            for (const char *n : names) {
                api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
                if (api.handle) break;
            }
        if (!api.handle) {
            api.error = std::string("cannot load libnccl.so.2: ") + dlerror();
            return;
        }
#define FBN_SYM(field, name)                                                        \
        *(void **)(&api.field) = dlsym(api.handle, name);                           \
        if (!api.field) { api.error = std::string("libnccl: missing symbol ") + name; return; }
        FBN_SYM(GetUniqueId, "ncclGetUniqueId")
        FBN_SYM(CommInitRank, "ncclCommInitRank")
        FBN_SYM(CommInitAll, "ncclCommInitAll")
        FBN_SYM(CommDestroy, "ncclCommDestroy")
        FBN_SYM(AllReduce, "ncclAllReduce")
        FBN_SYM(GetErrorString, "ncclGetErrorString")
        FBN_SYM(GetVersion, "ncclGetVersion")
#undef FBN_SYM
    });
    return api;
}

}  // namespace

struct fbn_comm {
    ncclComm_t comm = nullptr;
    int device = 0, rank = 0, world = 1;
    double *d_buf = nullptr;   // 4 doubles on the device
    double *h_buf = nullptr;   // pinned
    cudaStream_t stream = nullptr;
};

#define CKC(call)                                                                                      \
    do {                                                                                               \
        cudaError_t e__ = (call);                                                                      \
        if (e__ != cudaSuccess) return fbn_set_error(FBN_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__)); \
    } while (0)
#define CKN(call)                                                                                      \
    do {                                                                                               \
        int r__ = (call);                                                                              \
        if (r__ != kNcclSuccess) return fbn_set_error(FBN_ERR_COMM, std::string(#call) + ": " + nccl().GetErrorString(r__)); \
    } while (0)

static int comm_buffers(fbn_comm *c) {
    CKC(cudaSetDevice(c->device));
    CKC(cudaMalloc((void **)&c->d_buf, 4 * sizeof(double)));
    CKC(cudaMallocHost((void **)&c->h_buf, 4 * sizeof(double)));
    CKC(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    return FBN_OK;
}

extern "C" int fbn_comm_unique_id(void *out128) {
    if (!out128) return fbn_set_error(FBN_ERR_INVALID, "null argument");
    NcclApi &n = nccl();
    if (!n.error.empty()) return fbn_set_error(FBN_ERR_COMM, n.error);
    ncclUniqueId id;
    CKN(n.GetUniqueId(&id));
    memcpy(out128, &id, sizeof(id));
    return FBN_OK;
}

extern "C" int fbn_comm_create(int32_t device, int32_t rank, int32_t world, const void *id128, fbn_comm **out) {
    if (!out || !id128) return fbn_set_error(FBN_ERR_INVALID, "null argument");
    *out = nullptr;
    if (world < 1 || rank < 0 || rank >= world) return fbn_set_error(FBN_ERR_INVALID, "rank / world out of range");
    if (fbn_device_count() <= 0) return fbn_set_error(FBN_ERR_NO_DEVICE, "no CUDA device: fastbn_b200 has no CPU fallback");
    NcclApi &n = nccl();
    if (!n.error.empty()) return fbn_set_error(FBN_ERR_COMM, n.error);
    fbn_comm *c = new fbn_comm();
    c->device = device; c->rank = rank; c->world = world;
    int rc = comm_buffers(c);
    if (rc != FBN_OK) { fbn_comm_destroy(c); return rc; }
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    int r = n.CommInitRank(&c->comm, world, id, rank);
    if (r != kNcclSuccess) {
        fbn_comm_destroy(c);
        return fbn_set_error(FBN_ERR_COMM, std::string("ncclCommInitRank: ") + n.GetErrorString(r));
    }
    *out = c;
    return FBN_OK;
}

extern "C" int fbn_comm_create_all(int32_t n_devices, const int32_t *devices, fbn_comm **out) {
    if (!out || !devices || n_devices < 1) return fbn_set_error(FBN_ERR_INVALID, "bad arguments");
    for (int i = 0; i < n_devices; ++i) out[i] = nullptr;
    if (fbn_device_count() <= 0) return fbn_set_error(FBN_ERR_NO_DEVICE, "no CUDA device: fastbn_b200 has no CPU fallback");
    NcclApi &n = nccl();
    if (!n.error.empty()) return fbn_set_error(FBN_ERR_COMM, n.error);
    std::vector<ncclComm_t> comms((size_t)n_devices);
    std::vector<int> devs(devices, devices + n_devices);
    CKN(n.CommInitAll(comms.data(), n_devices, devs.data()));
    for (int i = 0; i < n_devices; ++i) {
        fbn_comm *c = new fbn_comm();
        c->comm = comms[(size_t)i];
        c->device = devices[i]; c->rank = i; c->world = n_devices;
        out[i] = c;
        int rc = comm_buffers(c);
        if (rc != FBN_OK) {
            for (int k = 0; k < n_devices; ++k) {
                if (out[k]) fbn_comm_destroy(out[k]);
                else n.CommDestroy(comms[(size_t)k]);
                out[k] = nullptr;
            }
            return rc;
        }
    }
    return FBN_OK;
}

extern "C" int fbn_reduce_device(fbn_comm *c, double *dev4, void *stream) {
    if (!c || !dev4) return fbn_set_error(FBN_ERR_INVALID, "null argument");
    CKC(cudaSetDevice(c->device));
    CKN(nccl().AllReduce(dev4, dev4, 4, kNcclFloat64, kNcclSum, c->comm, (cudaStream_t)stream));
    return FBN_OK;
}

extern "C" int fbn_reduce(fbn_comm *c, double inout4[4]) {
    if (!c || !inout4) return fbn_set_error(FBN_ERR_INVALID, "null argument");
    CKC(cudaSetDevice(c->device));
    memcpy(c->h_buf, inout4, 4 * sizeof(double));
    CKC(cudaMemcpyAsync(c->d_buf, c->h_buf, 4 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    CKN(nccl().AllReduce(c->d_buf, c->d_buf, 4, kNcclFloat64, kNcclSum, c->comm, c->stream));
    CKC(cudaMemcpyAsync(c->h_buf, c->d_buf, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CKC(cudaStreamSynchronize(c->stream));
    memcpy(inout4, c->h_buf, 4 * sizeof(double));
    return FBN_OK;
}

extern "C" int fbn_comm_info(const fbn_comm *c, int32_t *rank, int32_t *world, int32_t *nccl_version) {
    if (!c) return fbn_set_error(FBN_ERR_INVALID, "null argument");
    if (rank) *rank = c->rank;
    if (world) *world = c->world;
    if (nccl_version) {
        int v = 0;
        nccl().GetVersion(&v);
        *nccl_version = v;
    }
    return FBN_OK;
}

extern "C" void fbn_comm_destroy(fbn_comm *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) { cudaStreamSynchronize(c->stream); cudaStreamDestroy(c->stream); }
    if (c->comm) nccl().CommDestroy(c->comm);
    cudaFree(c->d_buf);
    cudaFreeHost(c->h_buf);
    cudaGetLastError();
    delete c;
}
