// Multi-GPU: heading-sharded volume, slabs exchanged with ncclAllGather over NVLink/NVSwitch
// (see include/pp_b200.h). NCCL is bound at run time (dlopen of libnccl.so.2) so
// that a process which already carries a NCCL -- e.g. one that imported torch -- shares that copy,
// and single-GPU users need no NCCL at all.
#include "pp_internal.h"

#include <dlfcn.h>
#include <nccl.h>
#include <cstring>

namespace
{
struct NcclApi
{
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi *nccl_api(std::string &why)
{
    static NcclApi api;
    static bool tried = false;
    if (tried)
        return api.lib ? &api : nullptr;
    tried = true;
    void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib)
    {
        why = dlerror();
        return nullptr;
    }
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(dlsym(lib, "ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(dlsym(lib, "ncclCommInitRank"));
    api.AllGather = reinterpret_cast<decltype(api.AllGather)>(dlsym(lib, "ncclAllGather"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
    if (!api.GetUniqueId || !api.CommInitRank || !api.AllGather || !api.CommDestroy || !api.GetErrorString)
    {
        why = "libnccl.so.2 lacks a required symbol";
        return nullptr;
    }
    api.lib = lib;
    return &api;
}
}  // namespace

extern "C" {

pp_status pp_comm_unique_id(void *id_out)
{
    if (!id_out)
        return PP_ERR_INVALID;
    std::string why;
    NcclApi *n = nccl_api(why);
    if (!n)
        return pp_fail(nullptr, PP_ERR_COMM, "NCCL unavailable: %s", why.c_str());
    static_assert(PP_COMM_ID_BYTES == NCCL_UNIQUE_ID_BYTES, "id size");
    ncclUniqueId id;
    ncclResult_t r = n->GetUniqueId(&id);
    if (r != ncclSuccess)
        return pp_fail(nullptr, PP_ERR_COMM, "ncclGetUniqueId: %s", n->GetErrorString(r));
    std::memcpy(id_out, &id, sizeof(id));
    return PP_OK;
}

pp_status pp_comm_init(pp_handle h, const void *id, int rank, int world)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, id && world >= 1 && rank >= 0 && rank < world, PP_ERR_INVALID, "pp_comm_init: rank %d of %d", rank, world);
    std::string why;
    NcclApi *n = nccl_api(why);
    PP_REQUIRE(h, n, PP_ERR_COMM, "NCCL unavailable: %s", why.c_str());
    PP_CUDA(h, cudaSetDevice(h->device));
    ncclUniqueId uid;
    std::memcpy(&uid, id, sizeof(uid));
    ncclComm_t comm;
    ncclResult_t r = n->CommInitRank(&comm, world, uid, rank);
    PP_REQUIRE(h, r == ncclSuccess, PP_ERR_COMM, "ncclCommInitRank: %s", n->GetErrorString(r));
    h->comm = comm;
    h->rank = rank;
    h->world = world;
    return PP_OK;
}

pp_status pp_se2_volume_allgather(pp_handle h)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->comm, PP_ERR_STATE, "pp_se2_volume_allgather: call pp_comm_init first");
    PP_REQUIRE(h, h->d_vol && h->n_theta > 0, PP_ERR_STATE, "pp_se2_volume_allgather: no volume");
    PP_REQUIRE(h, h->n_theta % h->world == 0, PP_ERR_INVALID, "n_theta %u is not a multiple of the world size %d", h->n_theta, h->world);
    const uint32_t per = h->n_theta / h->world;
    PP_REQUIRE(h, h->t_begin == per * h->rank && h->t_end == per * (h->rank + 1), PP_ERR_STATE,
               "rank %d must have built bins [%u,%u), has [%u,%u)", h->rank, per * h->rank, per * (h->rank + 1), h->t_begin, h->t_end);
    std::string why;
    NcclApi *n = nccl_api(why);
    PP_CUDA(h, cudaSetDevice(h->device));
    const size_t slab = static_cast<size_t>(h->W) * h->H * per;
    ncclResult_t r = n->AllGather(h->d_vol + slab * h->rank, h->d_vol, slab, ncclUint8, static_cast<ncclComm_t>(h->comm), h->stream);
    PP_REQUIRE(h, r == ncclSuccess, PP_ERR_COMM, "ncclAllGather: %s", n->GetErrorString(r));
    return PP_OK;
}

}  // extern "C"


int pp_volume_targets(pp_handle h, uint8_t **out)
{
    int n = 0;
    out[n++] = h->d_vol;
    if (h->push)
        for (int r = 0; r < h->n_peer; ++r)
            if (h->peer_vol[r] && h->peer_vol[r] != h->d_vol)
                out[n++] = h->peer_vol[r];
    return n;
}

extern "C" {

pp_status pp_se2_volume_reserve(pp_handle h, uint32_t n_theta)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_cost, PP_ERR_STATE, "pp_se2_volume_reserve: upload a costmap first (its size fixes the volume size)");
    PP_REQUIRE(h, n_theta > 0 && n_theta <= 4096, PP_ERR_RANGE, "n_theta %u outside 1..4096", n_theta);
    PP_CUDA(h, cudaSetDevice(h->device));
    PP_REQUIRE(h, h->n_peer == 0 || static_cast<size_t>(h->W) * h->H * n_theta <= h->vol_cap, PP_ERR_STATE,
               "pp_se2_volume_reserve: the volume is already mapped by peers and cannot grow");
    PP_CUDA(h, pp_reserve(h->d_vol, h->vol_cap, static_cast<size_t>(h->W) * h->H * n_theta));
    return PP_OK;
}

pp_status pp_se2_volume_ipc_export(pp_handle h, void *handle_out)
{
    if (!h || !handle_out)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_vol, PP_ERR_STATE, "pp_se2_volume_ipc_export: call pp_se2_volume_reserve first");
    static_assert(sizeof(cudaIpcMemHandle_t) == PP_IPC_HANDLE_BYTES, "IPC handle size");
    PP_CUDA(h, cudaSetDevice(h->device));
    cudaIpcMemHandle_t mh;
    PP_CUDA(h, cudaIpcGetMemHandle(&mh, h->d_vol));
    std::memcpy(handle_out, &mh, sizeof(mh));
    return PP_OK;
}

pp_status pp_se2_volume_ipc_open(pp_handle h, const void *handles, int rank, int world)
{
    if (!h || !handles)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, world >= 1 && world <= 8 && rank >= 0 && rank < world, PP_ERR_RANGE, "pp_se2_volume_ipc_open: rank %d of %d (at most 8)", rank, world);
    PP_REQUIRE(h, h->d_vol, PP_ERR_STATE, "pp_se2_volume_ipc_open: call pp_se2_volume_reserve first");
    PP_CUDA(h, cudaSetDevice(h->device));
    for (int r = 0; r < world; ++r)
    {
        if (r == rank)
        {
            h->peer_vol[r] = h->d_vol;
            continue;
        }
        cudaIpcMemHandle_t mh;
        std::memcpy(&mh, static_cast<const char *>(handles) + static_cast<size_t>(r) * PP_IPC_HANDLE_BYTES, sizeof(mh));
        void *p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, mh, cudaIpcMemLazyEnablePeerAccess);
        PP_REQUIRE(h, e == cudaSuccess, PP_ERR_COMM, "cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
        h->peer_vol[r] = static_cast<uint8_t *>(p);
    }
    h->n_peer = world;
    h->rank = rank;
    h->world = world;
    return PP_OK;
}

pp_status pp_se2_volume_set_push(pp_handle h, int enable)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, !enable || h->n_peer > 0, PP_ERR_STATE, "pp_se2_volume_set_push: no peer volumes mapped (pp_se2_volume_ipc_open)");
    h->push = enable != 0;
    return PP_OK;
}

}  // extern "C"

void pp_comm_destroy(pp_handle h)
{
    for (int r = 0; r < h->n_peer; ++r)
        if (h->peer_vol[r] && h->peer_vol[r] != h->d_vol)
            cudaIpcCloseMemHandle(h->peer_vol[r]);
    h->n_peer = 0;
    h->push = false;
    if (!h->comm)
        return;
    std::string why;
    if (NcclApi *n = nccl_api(why))
        n->CommDestroy(static_cast<ncclComm_t>(h->comm));
    h->comm = nullptr;
}
