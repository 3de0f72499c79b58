// nccl_shim.cu -- NCCL through dlopen: the library has no link-time dependency on libnccl, so the
// single-GPU path loads anywhere; the multi-GPU entry points (sfm_*_mgpu) resolve the symbols of the
// libnccl.so.2 already mapped into the process (torch's, when the host is Python) or found by the
// dynamic loader, and return SFM_E_NCCL when there is none.
//
// The reference has no communication layer at all; this backs SURVEY.md 8(e):
// one process per GPU, an 8-byte max-allreduce for RANSAC, and S / g_red sum-allreduces per LM
// trial for point-sharded bundle adjustment.
#include "sfm_common.cuh"
#include <dlfcn.h>
#include <mutex>
#include <stdio.h>
#include <string.h>

namespace sfm {

// minimal mirror of nccl.h (stable ABI since NCCL 2.x)
typedef struct { char internal[128]; } nccl_unique_id;
typedef void *nccl_comm_t;
enum { NCCL_SUM = 0, NCCL_MAX = 2 };
enum { NCCL_UINT64 = 5, NCCL_FLOAT64 = 8 };

struct NcclApi {
    int (*GetUniqueId)(nccl_unique_id *);
    int (*CommInitRank)(nccl_comm_t *, int, nccl_unique_id, int);
    int (*CommDestroy)(nccl_comm_t);
    int (*AllReduce)(const void *, void *, size_t, int, int, nccl_comm_t, cudaStream_t);
    const char *(*GetErrorString)(int);
    bool ok;
};

static NcclApi g_nccl = {};
static std::once_flag g_nccl_once;
static char g_nccl_err[256] = "";

static void nccl_load() {
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *n : names) {
        h = dlopen(n, RTLD_NOW | RTLD_NOLOAD);          // prefer the copy already in the process
        if (h) break;
    }
    if (!h) {
        const char *env = getenv("SFM_NCCL_LIB");
        if (env) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
    }
    for (const char *n : names) {
        if (h) break;
        h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    }
    if (!h) {
        snprintf(g_nccl_err, sizeof(g_nccl_err), "libnccl.so.2 not found (%s)", dlerror());
        return;
    }
    g_nccl.GetUniqueId = (int (*)(nccl_unique_id *))dlsym(h, "ncclGetUniqueId");
    g_nccl.CommInitRank = (int (*)(nccl_comm_t *, int, nccl_unique_id, int))dlsym(h, "ncclCommInitRank");
    g_nccl.CommDestroy = (int (*)(nccl_comm_t))dlsym(h, "ncclCommDestroy");
    g_nccl.AllReduce = (int (*)(const void *, void *, size_t, int, int, nccl_comm_t, cudaStream_t))dlsym(h, "ncclAllReduce");
    g_nccl.GetErrorString = (const char *(*)(int))dlsym(h, "ncclGetErrorString");
    g_nccl.ok = g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.CommDestroy && g_nccl.AllReduce;
    if (!g_nccl.ok) snprintf(g_nccl_err, sizeof(g_nccl_err), "libnccl.so.2 lacks a required symbol");
}

static bool nccl_ready() {
    std::call_once(g_nccl_once, nccl_load);
    return g_nccl.ok;
}

static int nccl_check(int rc, const char *what) {
    if (rc == 0) return SFM_OK;
    snprintf(g_nccl_err, sizeof(g_nccl_err), "%s: %s", what,
             g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "NCCL error");
    return SFM_E_NCCL;
}

// in-place allreduce helpers used by the *_mgpu entry points (comm == nullptr: single rank, no-op)
int comm_allreduce_sum_f64(void *comm, double *buf, size_t count, cudaStream_t st) {
    if (!comm || count == 0) return SFM_OK;
    if (!nccl_ready()) return SFM_E_NCCL;
    count_launch();
    return nccl_check(g_nccl.AllReduce(buf, buf, count, NCCL_FLOAT64, NCCL_SUM, comm, st), "ncclAllReduce(sum,f64)");
}

int comm_allreduce_max_u64(void *comm, unsigned long long *buf, size_t count, cudaStream_t st) {
    if (!comm || count == 0) return SFM_OK;
    if (!nccl_ready()) return SFM_E_NCCL;
    count_launch();
    return nccl_check(g_nccl.AllReduce(buf, buf, count, NCCL_UINT64, NCCL_MAX, comm, st), "ncclAllReduce(max,u64)");
}

}  // namespace sfm

using namespace sfm;

extern "C" const char *sfm_nccl_last_error(void) { return g_nccl_err; }

extern "C" int sfm_nccl_unique_id(void *h_id128) {
    if (!h_id128) return SFM_E_INVALID;
    if (!nccl_ready()) return SFM_E_NCCL;
    nccl_unique_id id;
    int rc = nccl_check(g_nccl.GetUniqueId(&id), "ncclGetUniqueId");
    if (rc == SFM_OK) memcpy(h_id128, &id, sizeof(id));
    return rc;
}

extern "C" int sfm_nccl_comm_create(const void *h_id128, int32_t rank, int32_t world, void **comm_out) {
    if (!h_id128 || !comm_out || world < 1 || rank < 0 || rank >= world) return SFM_E_INVALID;
    if (!nccl_ready()) return SFM_E_NCCL;
    nccl_unique_id id;
    memcpy(&id, h_id128, sizeof(id));
    nccl_comm_t c = nullptr;
    int rc = nccl_check(g_nccl.CommInitRank(&c, world, id, rank), "ncclCommInitRank");
    *comm_out = rc == SFM_OK ? c : nullptr;
    return rc;
}

extern "C" int sfm_nccl_comm_destroy(void *comm) {
    if (!comm) return SFM_OK;
    if (!nccl_ready()) return SFM_E_NCCL;
    return nccl_check(g_nccl.CommDestroy(comm), "ncclCommDestroy");
}

// exposed for tests / the RANSAC tail: in-place allreduces on the library's communicator
extern "C" int sfm_allreduce_sum_f64(void *comm, double *d_buf, int64_t count, void *stream) {
    if (!d_buf || count < 0) return SFM_E_INVALID;
    return comm_allreduce_sum_f64(comm, d_buf, (size_t)count, as_stream(stream));
}

extern "C" int sfm_allreduce_max_u64(void *comm, uint64_t *d_buf, int64_t count, void *stream) {
    if (!d_buf || count < 0) return SFM_E_INVALID;
    return comm_allreduce_max_u64(comm, reinterpret_cast<unsigned long long *>(d_buf), (size_t)count, as_stream(stream));
}
