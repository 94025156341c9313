// Multi-GPU: heading-sharded volume, slabs exchanged with ncclAllGather over NVLink/NVSwitch
// (see include/pp_b200.h). NCCL is bound at run time (dlopen of libnccl.so.2) so
// that a process which already carries a NCCL -- e.g. one that imported torch -- shares that copy,
// and single-GPU users need no NCCL at all.
#include "pp_internal.h"

#include <dlfcn.h>
#include <nccl.h>
#include <cstring>
#include <vector>

namespace
{
struct NcclApi
{
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
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
    api.Broadcast = reinterpret_cast<decltype(api.Broadcast)>(dlsym(lib, "ncclBroadcast"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(dlsym(lib, "ncclAllReduce"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(dlsym(lib, "ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(dlsym(lib, "ncclGroupEnd"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
    if (!api.GetUniqueId || !api.CommInitRank || !api.AllGather || !api.Broadcast || !api.AllReduce || !api.GroupStart || !api.GroupEnd ||
        !api.CommDestroy || !api.GetErrorString)
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

// Heading-sharded build with the exchange hidden behind it (see include/pp_b200.h).
pp_status pp_se2_volume_build_sharded(pp_handle h, uint32_t n_theta, float ox, float oy, float ot, uint32_t n_chunks, uint32_t flags)
{
    if (!h)
        return PP_ERR_INVALID;
    const bool use_push = flags & PP_SHARD_PUSH, exchange_only = flags & PP_SHARD_EXCHANGE_ONLY;
    PP_REQUIRE(h, h->comm, PP_ERR_STATE, "pp_se2_volume_build_sharded: call pp_comm_init first");
    PP_REQUIRE(h, !use_push || h->n_peer == h->world, PP_ERR_STATE, "pp_se2_volume_build_sharded: PP_SHARD_PUSH needs the peers' volumes mapped (pp_se2_volume_ipc_open)");
    PP_REQUIRE(h, !h->push, PP_ERR_STATE, "pp_se2_volume_build_sharded: switch the fused store exchange off first (pp_se2_volume_set_push(h, 0))");
    PP_REQUIRE(h, n_chunks >= 1 && n_chunks <= 16, PP_ERR_RANGE, "pp_se2_volume_build_sharded: n_chunks %u outside 1..16", n_chunks);
    PP_REQUIRE(h, n_theta > 0 && n_theta <= 4096, PP_ERR_RANGE, "n_theta %u outside 1..4096", n_theta);
    PP_REQUIRE(h, h->d_cost, PP_ERR_STATE, "pp_se2_volume_build_sharded: upload a costmap first");
    std::string why;
    NcclApi *n = nccl_api(why);
    PP_CUDA(h, cudaSetDevice(h->device));
    if (!h->comm_stream)
    {
        PP_CUDA(h, cudaStreamCreateWithFlags(&h->comm_stream, cudaStreamNonBlocking));
        PP_CUDA(h, cudaEventCreateWithFlags(&h->ev_comm, cudaEventDisableTiming));
        for (int c = 0; c < 16; ++c)
            PP_CUDA(h, cudaEventCreateWithFlags(&h->ev_chunk[c], cudaEventDisableTiming));
        PP_CUDA(h, cudaMalloc(reinterpret_cast<void **>(&h->d_token), sizeof(int)));
        PP_CUDA(h, cudaMemset(h->d_token, 0, sizeof(int)));
    }
    if (use_push)
        for (int p = 0; p < h->world; ++p)
            if (p != h->rank && !h->push_stream[p])
            {
                PP_CUDA(h, cudaStreamCreateWithFlags(&h->push_stream[p], cudaStreamNonBlocking));
                PP_CUDA(h, cudaEventCreateWithFlags(&h->ev_push[p], cudaEventDisableTiming));
            }
    const size_t plane = static_cast<size_t>(h->W) * h->H;
    PP_REQUIRE(h, h->n_peer == 0 || plane * n_theta <= h->vol_cap, PP_ERR_STATE,
               "pp_se2_volume_build_sharded: the volume is mapped by peers and too small for %u bins", n_theta);
    PP_CUDA(h, pp_reserve(h->d_vol, h->vol_cap, plane * n_theta));
    // units: pairs of bins (u, u + n/2) stay on one rank, so that a centrally symmetric footprint is
    // evaluated once per pair (an odd bin count has single bins as units)
    const bool pairs = n_theta % 2 == 0;
    const uint32_t U = pairs ? n_theta / 2 : n_theta;
    auto unit_range = [&](int rank, uint32_t chunk, uint32_t &a, uint32_t &b) {
        const uint32_t r0 = static_cast<uint32_t>(static_cast<uint64_t>(U) * rank / h->world);
        const uint32_t r1 = static_cast<uint32_t>(static_cast<uint64_t>(U) * (rank + 1) / h->world);
        // the first part is half as long as the others: nothing travels while it is being built
        const uint64_t total = 2ull * n_chunks - 1;
        auto cut = [&](uint32_t c) { return c == 0 ? 0ull : 2ull * c - 1; };
        a = r0 + static_cast<uint32_t>((static_cast<uint64_t>(r1 - r0) * cut(chunk) + total / 2) / total);
        b = chunk + 1 == n_chunks ? r1 : r0 + static_cast<uint32_t>((static_cast<uint64_t>(r1 - r0) * cut(chunk + 1) + total / 2) / total);
    };
    // one table count for all chunks of this call (chunks of different sizes must not flip the plan)
    {
        uint32_t a0, b0, a1, b1;
        unit_range(h->rank, 0, a0, b0);
        unit_range(h->rank, n_chunks - 1, a1, b1);
        const uint32_t per_chunk = (b1 - a0 + n_chunks - 1) / n_chunks * (pairs ? 2 : 1);
        h->few_bins_override = per_chunk <= 12 ? 1 : 0;
    }
    // everything queued so far (the costmap upload) precedes the exchange streams' first use of the volume
    PP_CUDA(h, cudaEventRecord(h->ev_comm, h->stream));
    PP_CUDA(h, cudaStreamWaitEvent(h->comm_stream, h->ev_comm, 0));
    pp_status st = PP_OK;
    for (uint32_t c = 0; c < n_chunks && st == PP_OK; ++c)
    {
        uint32_t a, b;
        unit_range(h->rank, c, a, b);
        if (!exchange_only && b > a)
        {
            std::vector<uint32_t> bins;
            for (uint32_t u = a; u < b; ++u)
            {
                bins.push_back(u);
                if (pairs)
                    bins.push_back(u + U);
            }
            st = pp_se2_volume_build_bins(h, n_theta, ox, oy, ot, bins.data(), static_cast<uint32_t>(bins.size()));
            if (st != PP_OK)
                break;
        }
        PP_CUDA(h, cudaEventRecord(h->ev_chunk[c], h->stream));
        if (use_push)
        {
            // copy engines: this rank's part of the chunk goes to every peer's mapped volume
            for (int p = 0; p < h->world; ++p)
            {
                if (p == h->rank || b == a)
                    continue;
                PP_CUDA(h, cudaStreamWaitEvent(h->push_stream[p], h->ev_chunk[c], 0));
                // both halves of the volume in one strided copy where the pitch allows it
                const size_t off = plane * a, pitch = plane * U;
                if (pairs && pitch <= static_cast<size_t>(0x7FFFFFFF))
                    PP_CUDA(h, cudaMemcpy2DAsync(h->peer_vol[p] + off, pitch, h->d_vol + off, pitch, plane * (b - a), 2, cudaMemcpyDeviceToDevice,
                                                 h->push_stream[p]));
                else
                    for (int half = 0; half < (pairs ? 2 : 1); ++half)
                        PP_CUDA(h, cudaMemcpyAsync(h->peer_vol[p] + off + half * pitch, h->d_vol + off + half * pitch, plane * (b - a),
                                                   cudaMemcpyDeviceToDevice, h->push_stream[p]));
            }
        }
        else
        {
            // NCCL: every rank's part of the chunk is broadcast from its owner, all in one group
            PP_CUDA(h, cudaStreamWaitEvent(h->comm_stream, h->ev_chunk[c], 0));
            ncclResult_t r = n->GroupStart();
            for (int root = 0; root < h->world && r == ncclSuccess; ++root)
            {
                uint32_t ra, rb;
                unit_range(root, c, ra, rb);
                if (rb == ra)
                    continue;
                for (int half = 0; half < (pairs ? 2 : 1) && r == ncclSuccess; ++half)
                {
                    uint8_t *buf = h->d_vol + plane * (ra + static_cast<size_t>(half) * U);
                    r = n->Broadcast(buf, buf, plane * (rb - ra), ncclUint8, root, static_cast<ncclComm_t>(h->comm), h->comm_stream);
                }
            }
            const ncclResult_t r2 = n->GroupEnd();
            PP_REQUIRE(h, r == ncclSuccess && r2 == ncclSuccess, PP_ERR_COMM, "ncclBroadcast (chunk %u): %s", c,
                       n->GetErrorString(r != ncclSuccess ? r : r2));
        }
    }
    h->few_bins_override = -1;
    if (st != PP_OK)
        return st;
    if (use_push)
    {
        // the peers' parts are here once THEY have finished pushing: a one-word all-reduce queued behind this
        // rank's copies completes on a rank only after every rank has reached it
        for (int p = 0; p < h->world; ++p)
            if (p != h->rank)
            {
                PP_CUDA(h, cudaEventRecord(h->ev_push[p], h->push_stream[p]));
                PP_CUDA(h, cudaStreamWaitEvent(h->comm_stream, h->ev_push[p], 0));
            }
        ncclResult_t r = n->AllReduce(h->d_token, h->d_token, 1, ncclInt32, ncclSum, static_cast<ncclComm_t>(h->comm), h->comm_stream);
        PP_REQUIRE(h, r == ncclSuccess, PP_ERR_COMM, "ncclAllReduce: %s", n->GetErrorString(r));
    }
    // the caller's stream continues when the exchange is complete
    PP_CUDA(h, cudaEventRecord(h->ev_comm, h->comm_stream));
    PP_CUDA(h, cudaStreamWaitEvent(h->stream, h->ev_comm, 0));
    // the handle now holds every bin
    h->t_begin = 0;
    h->t_end = n_theta;
    h->bins.resize(n_theta);
    for (uint32_t t = 0; t < n_theta; ++t)
        h->bins[t] = t;
    h->vol_valid = true;
    h->dirty = false;
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
    if (h->comm_stream)
    {
        cudaStreamSynchronize(h->comm_stream);
        cudaStreamDestroy(h->comm_stream);
        cudaEventDestroy(h->ev_comm);
        for (int c = 0; c < 16; ++c)
            cudaEventDestroy(h->ev_chunk[c]);
        cudaFree(h->d_token);
        h->comm_stream = nullptr;
    }
    for (int p = 0; p < 8; ++p)
        if (h->push_stream[p])
        {
            cudaStreamSynchronize(h->push_stream[p]);
            cudaStreamDestroy(h->push_stream[p]);
            cudaEventDestroy(h->ev_push[p]);
            h->push_stream[p] = nullptr;
        }
    if (!h->comm)
        return;
    std::string why;
    if (NcclApi *n = nccl_api(why))
        n->CommDestroy(static_cast<ncclComm_t>(h->comm));
    h->comm = nullptr;
}
