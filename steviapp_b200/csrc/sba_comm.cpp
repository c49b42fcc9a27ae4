// sba_comm.cpp -- NCCL plumbing, loaded lazily with dlopen so that the single-GPU
// library has no NCCL link dependency and shares whatever libnccl.so.2 the host
// process (e.g. torch) already mapped.  One process per GPU; the ncclUniqueId is
// made by rank 0 and distributed by the launcher (torch.distributed in bench.py).
#include <dlfcn.h>

#include <cstring>

#include "sba_solver.h"

namespace sba {

namespace {
struct NcclId { char internal[128]; };
typedef int (*fn_get_id)(NcclId*);
typedef int (*fn_init_rank)(void**, int, NcclId, int);
typedef int (*fn_allreduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_allgather)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef int (*fn_reducescatter)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_destroy)(void*);
typedef const char* (*fn_errstr)(int);

struct Api {
    void* lib = nullptr;
    fn_get_id get_id = nullptr;
    fn_init_rank init_rank = nullptr;
    fn_allreduce allreduce = nullptr;
    fn_allgather allgather = nullptr;
    fn_reducescatter reducescatter = nullptr;
    fn_destroy destroy = nullptr;
    fn_errstr errstr = nullptr;
} api;

int load(std::string& err) {
    if (api.lib) return 0;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.lib) break;
    }
    if (!api.lib) { err = std::string("NCCL not found: ") + dlerror(); return -1; }
    api.get_id = (fn_get_id)dlsym(api.lib, "ncclGetUniqueId");
    api.init_rank = (fn_init_rank)dlsym(api.lib, "ncclCommInitRank");
    api.allreduce = (fn_allreduce)dlsym(api.lib, "ncclAllReduce");
    api.allgather = (fn_allgather)dlsym(api.lib, "ncclAllGather");
    api.reducescatter = (fn_reducescatter)dlsym(api.lib, "ncclReduceScatter");
    api.destroy = (fn_destroy)dlsym(api.lib, "ncclCommDestroy");
    api.errstr = (fn_errstr)dlsym(api.lib, "ncclGetErrorString");
    if (!api.get_id || !api.init_rank || !api.allreduce || !api.destroy) { err = "NCCL symbols missing"; return -1; }
    return 0;
}
std::string nccl_err(int r) { return api.errstr ? api.errstr(r) : "nccl error"; }
constexpr int NCCL_FLOAT64 = 8, NCCL_SUM = 0, NCCL_MAX = 2;
}  // namespace

int Comm::unique_id(uint8_t* id128, std::string& err) {
    if (load(err)) return -1;
    NcclId id;
    const int r = api.get_id(&id);
    if (r) { err = "ncclGetUniqueId: " + nccl_err(r); return -1; }
    std::memcpy(id128, id.internal, 128);
    return 0;
}

int Comm::init(int rank_, int nranks_, const uint8_t* id128, std::string& err) {
    if (load(err)) return -1;
    NcclId id;
    std::memcpy(id.internal, id128, 128);
    const int r = api.init_rank(&comm, nranks_, id, rank_);
    if (r) { err = "ncclCommInitRank: " + nccl_err(r); return -1; }
    rank = rank_; nranks = nranks_; active = true;
    return 0;
}

int Comm::allreduce_sum(double* buf, size_t n, cudaStream_t st, std::string& err) {
    const int r = api.allreduce(buf, buf, n, NCCL_FLOAT64, NCCL_SUM, comm, st);
    if (r) { err = "ncclAllReduce(sum): " + nccl_err(r); return -1; }
    return 0;
}
int Comm::allreduce_max(double* buf, size_t n, cudaStream_t st, std::string& err) {
    const int r = api.allreduce(buf, buf, n, NCCL_FLOAT64, NCCL_MAX, comm, st);
    if (r) { err = "ncclAllReduce(max): " + nccl_err(r); return -1; }
    return 0;
}
int Comm::allgather_bytes(const void* send, void* recv, size_t bytes_per_rank, cudaStream_t st, std::string& err) {
    if (!api.allgather) { err = "ncclAllGather missing"; return -1; }
    const int r = api.allgather(send, recv, bytes_per_rank, /*ncclChar*/ 0, comm, st);
    if (r) { err = "ncclAllGather: " + nccl_err(r); return -1; }
    return 0;
}
// recv (count doubles) = sum over ranks of the rank-th slice of send (nranks * count doubles)
int Comm::reduce_scatter_sum(const double* send, double* recv, size_t count, cudaStream_t st, std::string& err) {
    if (!api.reducescatter) { err = "ncclReduceScatter missing"; return -1; }
    const int r = api.reducescatter(send, recv, count, /*ncclFloat64*/ 8, /*ncclSum*/ 0, comm, st);
    if (r) { err = "ncclReduceScatter: " + nccl_err(r); return -1; }
    return 0;
}
void Comm::destroy() {
    if (active && comm && api.destroy) api.destroy(comm);
    comm = nullptr; active = false;
}

}  // namespace sba
