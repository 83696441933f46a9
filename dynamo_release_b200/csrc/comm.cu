// Multi-GPU plumbing: one NCCL communicator owned by the library.  libnccl.so.2 is resolved at run time with
// dlopen (the copy PyTorch already loaded is reused; the library itself has no link-time NCCL dependency,
// so single-GPU use never touches NCCL).  Cells are sharded over ranks; the only traffic of the EM loop is
// the allreduce of the Gram/RHS block plus a handful of scalars ("multi-GPU").
#include <dlfcn.h>
#include <mutex>
#include "common.cuh"
#include "comm.h"

namespace dvf {

namespace {
typedef int (*fn_get_uid)(void*);
typedef int (*fn_init_rank)(void**, int, dvf_nccl_uid, int);
typedef int (*fn_destroy)(void*);
typedef int (*fn_allreduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef const char* (*fn_errstr)(int);

struct nccl_api {
    void* handle = nullptr;
    fn_get_uid get_uid = nullptr;
    fn_init_rank init_rank = nullptr;
    fn_destroy destroy = nullptr;
    fn_allreduce allreduce = nullptr;
    fn_errstr errstr = nullptr;
};
nccl_api g_nccl;
std::once_flag g_once;

void load_nccl() {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        g_nccl.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.handle) break;
    }
    if (!g_nccl.handle) return;
    g_nccl.get_uid = (fn_get_uid)dlsym(g_nccl.handle, "ncclGetUniqueId");
    g_nccl.init_rank = (fn_init_rank)dlsym(g_nccl.handle, "ncclCommInitRank");
    g_nccl.destroy = (fn_destroy)dlsym(g_nccl.handle, "ncclCommDestroy");
    g_nccl.allreduce = (fn_allreduce)dlsym(g_nccl.handle, "ncclAllReduce");
    g_nccl.errstr = (fn_errstr)dlsym(g_nccl.handle, "ncclGetErrorString");
}

int need_nccl() {
    std::call_once(g_once, load_nccl);
    if (!g_nccl.handle || !g_nccl.get_uid || !g_nccl.init_rank || !g_nccl.allreduce) {
        set_error("libnccl.so.2 could not be loaded (%s)", dlerror() ? dlerror() : "symbols missing");
        return DVF_ERR_COMM;
    }
    return DVF_OK;
}

int nccl_fail(int rc, const char* what) {
    set_error("NCCL error %d (%s) in %s", rc, g_nccl.errstr ? g_nccl.errstr(rc) : "?", what);
    return DVF_ERR_COMM;
}
}  // namespace

int comm_allreduce(void* comm, double* buf, int64_t count, bool is_min, cudaStream_t st) {
    if (!comm || count <= 0) return DVF_OK;
    comm_t* c = (comm_t*)comm;
    if (c->world <= 1) return DVF_OK;
    int rc = g_nccl.allreduce(buf, buf, (size_t)count, /*ncclFloat64*/ 8, is_min ? /*ncclMin*/ 3 : /*ncclSum*/ 0, c->nccl, st);
    if (rc != 0) return nccl_fail(rc, "ncclAllReduce");
    return DVF_OK;
}

}  // namespace dvf

using namespace dvf;

extern "C" int dvf_comm_unique_id(void* out128) {
    DVF_ARG(out128, "NULL pointer");
    if (int rc = need_nccl()) return rc;
    int rc = g_nccl.get_uid(out128);
    if (rc != 0) return nccl_fail(rc, "ncclGetUniqueId");
    return DVF_OK;
}

extern "C" int dvf_comm_init(const void* id128, int rank, int world, void** comm) {
    DVF_ARG(id128 && comm && world >= 1 && rank >= 0 && rank < world, "bad argument");
    if (int rc = need_nccl()) return rc;
    comm_t* c = new comm_t();
    c->rank = rank;
    c->world = world;
    dvf_nccl_uid uid;
    memcpy(&uid, id128, sizeof(uid));
    int rc = g_nccl.init_rank(&c->nccl, world, uid, rank);
    if (rc != 0) { delete c; return nccl_fail(rc, "ncclCommInitRank"); }
    *comm = c;
    return DVF_OK;
}

extern "C" int dvf_comm_destroy(void* comm) {
    if (!comm) return DVF_OK;
    comm_t* c = (comm_t*)comm;
    if (c->nccl && g_nccl.destroy) g_nccl.destroy(c->nccl);
    delete c;
    return DVF_OK;
}

extern "C" int dvf_comm_allreduce_sum(void* comm, double* buf, int64_t count, void* stream) {
    DVF_ARG(comm && buf, "NULL pointer");
    return comm_allreduce(comm, buf, count, false, (cudaStream_t)stream);
}
