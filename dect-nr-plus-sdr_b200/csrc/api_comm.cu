// api_comm.cu -- the one collective of the multi-GPU path: an all-reduce over the experiment's counter vector
// (SURVEY.md 8e).  The reference keeps its counters in one object of one process (result_t::n_pcc / n_pcc_and_plcf /
// n_pdc, lib/src/upper/loopback/result.cpp:78-95; worker statistics, lib/src/phy/pool/worker_tx_rx.cpp:96-198); with one
// context per GPU the same sums come from ncclAllReduce over NVLink / NVSwitch.  NCCL is bound at run time with dlopen:
// the library has no link-time dependency on it and a process that never reduces never loads it.
#include <dlfcn.h>
#include <nccl.h>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>
#include "ctx.hpp"

#define DNRP_REQUIRE(cond, msg)                 \
    do {                                        \
        if (!(cond)) {                          \
            ::dnrp::set_last_error(msg);        \
            return DNRP_ERR_INVALID;            \
        }                                       \
    } while (0)

namespace {

struct nccl_api_t {
    void* so = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

// resolved once per process; nullptr (with the last error set) when libnccl.so.2 is not there
const nccl_api_t* nccl() {
    static nccl_api_t api;
    static std::once_flag once;
    static bool ok = false;
    std::call_once(once, [] {
        // a copy already mapped into the process (e.g. the one bundled with PyTorch) answers by its soname -- and the
        // other way round: a process that loads PyTorch LATER gets this copy for PyTorch's own NCCL calls, so such a
        // process imports torch first (the Python binding does) or names the library to use in DNRP_NCCL_LIB
        const char* path = getenv("DNRP_NCCL_LIB");
        if (path && path[0]) api.so = dlopen(path, RTLD_NOW | RTLD_LOCAL);
        if (!api.so) api.so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
        if (!api.so) api.so = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
        if (!api.so) return;
#define DNRP_SYM(name) api.name = reinterpret_cast<decltype(api.name)>(dlsym(api.so, "nccl" #name))
        DNRP_SYM(GetUniqueId); DNRP_SYM(CommInitRank); DNRP_SYM(CommInitAll); DNRP_SYM(CommDestroy);
        DNRP_SYM(AllReduce); DNRP_SYM(GroupStart); DNRP_SYM(GroupEnd); DNRP_SYM(GetErrorString);
#undef DNRP_SYM
        ok = api.GetUniqueId && api.CommInitRank && api.CommInitAll && api.CommDestroy && api.AllReduce &&
             api.GroupStart && api.GroupEnd && api.GetErrorString;
    });
    if (!ok) {
        dnrp::set_last_error("NCCL (libnccl.so.2) could not be loaded: the counter reduction needs it");
        return nullptr;
    }
    return &api;
}

#define DNRP_NCCL(api, expr)                                                                                \
    do {                                                                                                    \
        ncclResult_t r__ = (expr);                                                                          \
        if (r__ != ncclSuccess) {                                                                           \
            ::dnrp::set_last_error(std::string(#expr) + ": " + (api)->GetErrorString(r__));                 \
            return DNRP_ERR_CUDA;                                                                           \
        }                                                                                                   \
    } while (0)

struct comm_state_t {
    ncclComm_t comm = nullptr;
    int64_t* d_buf = nullptr;     // [DNRP_COMM_MAX_COUNTERS] on the context's device
    int n_ranks = 0, rank = 0;
};
std::mutex g_mutex;
std::unordered_map<const dnrp_ctx_t*, comm_state_t> g_comms;

comm_state_t* find_comm(const dnrp_ctx_t* c) {
    std::lock_guard<std::mutex> lk(g_mutex);
    auto it = g_comms.find(c);
    return it == g_comms.end() ? nullptr : &it->second;
}

int attach(dnrp_ctx_t* c, ncclComm_t comm, int n_ranks, int rank) {
    comm_state_t s;
    s.comm = comm; s.n_ranks = n_ranks; s.rank = rank;
    DNRP_CUDA(cudaSetDevice(c->device));
    DNRP_CUDA(cudaMalloc((void**)&s.d_buf, sizeof(int64_t) * DNRP_COMM_MAX_COUNTERS));
    std::lock_guard<std::mutex> lk(g_mutex);
    g_comms[c] = s;
    return DNRP_OK;
}

ncclRedOp_t red_op(int op) { return op == 1 ? ncclMin : (op == 2 ? ncclMax : ncclSum); }

}  // namespace

extern "C" {

int dnrp_comm_unique_id(uint8_t id[DNRP_COMM_ID_BYTES]) {
    DNRP_REQUIRE(id, "null argument");
    static_assert(sizeof(ncclUniqueId) == DNRP_COMM_ID_BYTES, "ncclUniqueId size");
    const nccl_api_t* api = nccl();
    if (!api) return DNRP_ERR_STATE;
    ncclUniqueId u;
    DNRP_NCCL(api, api->GetUniqueId(&u));
    std::memcpy(id, &u, sizeof(u));
    return DNRP_OK;
}

int dnrp_comm_init_rank(dnrp_ctx_t* c, int n_ranks, int rank, const uint8_t id[DNRP_COMM_ID_BYTES]) {
    DNRP_REQUIRE(c && id, "null argument");
    DNRP_REQUIRE(n_ranks >= 1 && rank >= 0 && rank < n_ranks, "rank out of range");
    DNRP_REQUIRE(!find_comm(c), "the context already has a communicator (dnrp_comm_destroy first)");
    const nccl_api_t* api = nccl();
    if (!api) return DNRP_ERR_STATE;
    ncclUniqueId u;
    std::memcpy(&u, id, sizeof(u));
    DNRP_CUDA(cudaSetDevice(c->device));
    ncclComm_t comm = nullptr;
    DNRP_NCCL(api, api->CommInitRank(&comm, n_ranks, u, rank));
    return attach(c, comm, n_ranks, rank);
}

int dnrp_comm_init_all(dnrp_ctx_t* const* ctxs, int n_ctx) {
    DNRP_REQUIRE(ctxs && n_ctx >= 1, "null argument");
    std::vector<int> devs(n_ctx);
    for (int i = 0; i < n_ctx; ++i) {
        DNRP_REQUIRE(ctxs[i], "null context");
        DNRP_REQUIRE(!find_comm(ctxs[i]), "a context already has a communicator (dnrp_comm_destroy first)");
        devs[i] = ctxs[i]->device;
        for (int j = 0; j < i; ++j) DNRP_REQUIRE(devs[j] != devs[i], "two contexts on one device: NCCL wants one rank per GPU");
    }
    const nccl_api_t* api = nccl();
    if (!api) return DNRP_ERR_STATE;
    std::vector<ncclComm_t> comms(n_ctx, nullptr);
    DNRP_NCCL(api, api->CommInitAll(comms.data(), n_ctx, devs.data()));
    for (int i = 0; i < n_ctx; ++i) {
        const int rc = attach(ctxs[i], comms[i], n_ctx, i);
        if (rc != DNRP_OK) return rc;
    }
    return DNRP_OK;
}

int dnrp_counters_allreduce_all(dnrp_ctx_t* const* ctxs, int n_ctx, int64_t* counters, uint32_t n, int op) {
    DNRP_REQUIRE(ctxs && counters && n_ctx >= 1, "null argument");
    DNRP_REQUIRE(n >= 1 && n <= DNRP_COMM_MAX_COUNTERS, "counter count out of range");
    DNRP_REQUIRE(op >= 0 && op <= 2, "op: 0 sum, 1 min, 2 max");
    std::vector<comm_state_t*> st(n_ctx);
    for (int i = 0; i < n_ctx; ++i) {
        DNRP_REQUIRE(ctxs[i], "null context");
        st[i] = find_comm(ctxs[i]);
        if (!st[i]) {
            dnrp::set_last_error("no communicator on this context (dnrp_comm_init_rank / dnrp_comm_init_all first)");
            return DNRP_ERR_STATE;
        }
    }
    const nccl_api_t* api = nccl();   // after the state checks: a call that is refused does not load NCCL
    if (!api) return DNRP_ERR_STATE;
    for (int i = 0; i < n_ctx; ++i) {
        DNRP_CUDA(cudaSetDevice(ctxs[i]->device));
        DNRP_CUDA(cudaMemcpyAsync(st[i]->d_buf, counters + (size_t)i * n, sizeof(int64_t) * n, cudaMemcpyHostToDevice,
                                  ctxs[i]->stream));
    }
    // several ranks driven by one thread: their calls must form one group, or the first would wait for the others
    DNRP_NCCL(api, api->GroupStart());
    for (int i = 0; i < n_ctx; ++i) {
        ncclResult_t r = api->AllReduce(st[i]->d_buf, st[i]->d_buf, n, ncclInt64, red_op(op), st[i]->comm, ctxs[i]->stream);
        if (r != ncclSuccess) {
            api->GroupEnd();
            dnrp::set_last_error(std::string("ncclAllReduce: ") + api->GetErrorString(r));
            return DNRP_ERR_CUDA;
        }
        ++ctxs[i]->launches;
    }
    DNRP_NCCL(api, api->GroupEnd());
    for (int i = 0; i < n_ctx; ++i) {
        DNRP_CUDA(cudaSetDevice(ctxs[i]->device));
        DNRP_CUDA(cudaMemcpyAsync(counters + (size_t)i * n, st[i]->d_buf, sizeof(int64_t) * n, cudaMemcpyDeviceToHost,
                                  ctxs[i]->stream));
    }
    for (int i = 0; i < n_ctx; ++i) {
        DNRP_CUDA(cudaSetDevice(ctxs[i]->device));
        DNRP_CUDA(cudaStreamSynchronize(ctxs[i]->stream));
    }
    return DNRP_OK;
}

int dnrp_counters_allreduce(dnrp_ctx_t* c, int64_t* counters, uint32_t n, int op) {
    dnrp_ctx_t* one[1] = {c};
    return dnrp_counters_allreduce_all(one, 1, counters, n, op);
}

int dnrp_comm_destroy(dnrp_ctx_t* c) {
    DNRP_REQUIRE(c, "null argument");
    comm_state_t s;
    {
        std::lock_guard<std::mutex> lk(g_mutex);
        auto it = g_comms.find(c);
        if (it == g_comms.end()) return DNRP_OK;
        s = it->second;
        g_comms.erase(it);
    }
    cudaSetDevice(c->device);
    if (s.d_buf) cudaFree(s.d_buf);
    const nccl_api_t* api = nccl();
    if (api && s.comm) api->CommDestroy(s.comm);
    return DNRP_OK;
}

}  // extern "C"
