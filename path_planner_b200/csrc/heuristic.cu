// K2: obstacle-aware 2D heuristic field (see include/pp_b200.h).
//
// The reference grows the field lazily with a heap (astar_hybrid.cpp:97-241). Its recurrence is
//   d[goal] = 1e-5f                                                       (:117, sign-encoded)
//   d[v]    = min over 8 neighbours u of  fl( d[u] + travel(u->v) )        (:151-158)
//   travel  = (straight ? 1.0f : sqrtf(2.0f)) * (1.0f + (cost_penalty * cost[v] / 252.0f))
// where v must be an interior cell (:141) with cost[v] < 253 (:143-145). Float addition is
// monotone, so this system has one least fixed point, independent of relaxation order; here it is
// reached by tiled chaotic relaxation: each active 32x32 tile is staged in shared memory with a
// one-cell halo, relaxed in place until nothing in it changes (per-tile convergence flag via
// __syncthreads_or), written back, and the tiles across any border that changed are flagged for
// the next round. min_u fl(d[u] + w) == fl(min_u d[u] + w), so the 4 straight and the 4 diagonal
// neighbours are min-reduced first and each class costs one rounded addition. The travel weights
// come from a 2x256 table evaluated on the host with the reference's exact float expression.
#include "pp_internal.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <limits>
#include <vector>

namespace
{
constexpr int HT = 32;            // tile edge
#ifndef PP_HT_BR
#define PP_HT_BR 4
#endif
constexpr int BR = PP_HT_BR;                 // rows of a warp's band = cells per thread
constexpr int HT_THREADS = 32 * (HT / BR);   // one warp per band, lane = column
constexpr int HT_PITCH = HT + 2 + 1;
#ifndef PP_HT_CHECK
#define PP_HT_CHECK 32
#endif
#ifndef PP_HT_GRID
#define PP_HT_GRID 4
#endif
constexpr int ROUNDS_PER_CHECK = PP_HT_CHECK;
#ifndef PP_HT_LOCAL
#define PP_HT_LOCAL 4
#endif
constexpr int HT_LOCAL = PP_HT_LOCAL;  // warp-local relaxations of a band between two block-wide votes
constexpr uint32_t ROUND_GRID = 148 * PP_HT_GRID;  // CTAs per relaxation round: they walk the round's work list

constexpr uint32_t INACTIVE = 0xFFFFFFFFu;  // tile key: bit pattern of the smallest pending value, or this
constexpr int N_COUNTERS = 16;              // [0..3] tiles flagged per round, [4] visits, [5] sweeps, [6] rounds that had work, [8..11] min key per round

#ifdef PP_HEUR_TIMELINE
// dev build: per tile visit {round, tile, kernel entry, after the dependency wait, work list read, tile staged,
// converged, published (ns, %globaltimer), sweeps}
__device__ unsigned long long heur_tl[9 * (1 << 17)];
__device__ unsigned int heur_tl_n;
__device__ __forceinline__ unsigned long long gtime()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#endif

__global__ void heur_init_kernel(float *field, size_t n, uint32_t *flags, size_t n_flags, uint32_t goal_index,
                                 uint32_t goal_tile)
{
    const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    const uint32_t goal_key = __float_as_uint(0.00001f);
    for (size_t k = i; k < n; k += stride)
        field[k] = (k == goal_index) ? 0.00001f : __int_as_float(0x7f800000);
    // layout: [0,nT) keys gen 0 | [nT,2nT) keys gen 1 | [2nT,3nT) list gen 0 | [3nT,4nT) list gen 1 | counters
    const size_t counters_at = n_flags - N_COUNTERS;
    const size_t nT = counters_at / 4;
    for (size_t k = i; k < n_flags; k += stride)
    {
        uint32_t v = INACTIVE;                       // tile keys of both generations
        if (k == goal_tile)
            v = goal_key;                            // the goal's tile is the only active one in generation 0
        else if (k >= 2 * nT && k < counters_at)
            v = (k == 2 * nT) ? goal_tile : 0u;      // work lists; generation 0 holds the goal's tile
        else if (k >= counters_at)
        {
            const size_t c = k - counters_at;
            v = (c >= 8 && c < 12) ? (c == 8 ? goal_key : INACTIVE) : (c == 0 ? 1u : 0u);
        }
        flags[k] = v;
    }
}

// flags layout: [0, nT) tile keys of generation 0, [nT, 2nT) generation 1, then N_COUNTERS words.
// Scheduling: a tile's key is the smallest value that reached its border since it last ran. A
// round only runs the tiles whose key is within `delta` of the smallest key of the round (a
// delta-stepping bucket over tiles); the others carry their key over to the next round. This keeps
// tiles far behind the wavefront from being re-relaxed every time a slightly better value trickles
// in, without changing the fixed point that is reached.
__global__ void __launch_bounds__(HT_THREADS) heur_round_kernel(float *__restrict__ field, const uint8_t *__restrict__ cost,
                                                                const float *__restrict__ travel, uint32_t W, uint32_t H,
                                                                uint32_t tiles_x, uint32_t tiles_y, uint32_t *flags,
                                                                uint32_t round, float delta)
{
    // Programmatic dependent launch: the CTAs of round r+1 may be scheduled while round r drains (most
    // CTAs of a round find no work and leave at once); they park here until round r has completed and its
    // writes are visible. Without the launch attribute both instructions do nothing.
#ifdef PP_HEUR_TIMELINE
    const unsigned long long tl_entry = gtime();
#endif
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
#ifdef PP_HEUR_TIMELINE
    const unsigned long long tl_wait = gtime();
#endif
    const uint32_t nT = tiles_x * tiles_y;
    uint32_t *cur = flags + (round & 1) * nT;
    uint32_t *nxt = flags + ((round + 1) & 1) * nT;
    const uint32_t *list_cur = flags + (2 + (round & 1)) * nT;
    uint32_t *list_nxt = flags + (2 + ((round + 1) & 1)) * nT;
    uint32_t *counters = flags + 4 * nT;
    if (blockIdx.x == 0 && threadIdx.x == 0)
    {
        counters[(round + 2) & 3] = 0;               // will count the tiles listed during round+1
        counters[8 + ((round + 2) & 3)] = INACTIVE;  // and hold their smallest key
    }
    // The first list entry of this thread is requested together with the entry count it depends on (entries past
    // the count are stale tiles of earlier rounds: loaded, never used): one round trip less at the head of a round.
    const uint32_t first_entry = blockIdx.x + threadIdx.x * gridDim.x;
    const uint32_t first_tile = first_entry < nT ? list_cur[first_entry] : 0u;
    const uint32_t n_active = counters[round & 3];
    if (blockIdx.x == 0 && threadIdx.x == 0 && n_active > 0)
        counters[6] = round + 1;  // rounds that had work (the host launches them in batches and overshoots)
    const float threshold = __uint_as_float(counters[8 + (round & 3)]) + delta;
    __shared__ float F[HT + 2][HT_PITCH];
    __shared__ float WS[HT][HT];
    __shared__ float WD[HT][HT];
    __shared__ uint32_t edge_min[8];  // smallest new value on each border / corner (bit pattern), INACTIVE if unchanged
    __shared__ uint32_t s_moved[3];   // bands that changed in a vote of the in-tile relaxation, three votes in rotation
    __shared__ float s_travel[512];
    for (int i = threadIdx.x; i < 512; i += HT_THREADS)
        s_travel[i] = travel[i];
    // Pass 1, one list entry per thread: entries outside this round's bucket carry their key over to
    // the next generation (their chains of dependent atomics run side by side instead of one after the
    // other), entries inside it are collected for pass 2.
    __shared__ uint32_t s_mine[HT_THREADS];
    __shared__ uint32_t s_n_mine;
    for (uint32_t base = blockIdx.x; base < n_active; base += gridDim.x * HT_THREADS)
    {
    __syncthreads();  // the previous batch's tiles are done with s_mine / the tile arrays
    if (threadIdx.x == 0)
        s_n_mine = 0;
    __syncthreads();
    const uint32_t entry = base + threadIdx.x * gridDim.x;
    if (entry < n_active)
    {
        const uint32_t tile = base == blockIdx.x ? first_tile : list_cur[entry];
        const uint32_t key = cur[tile];
        if (__uint_as_float(key) > threshold)
        {
            cur[tile] = INACTIVE;
            if (atomicMin(&nxt[tile], key) == INACTIVE)
                list_nxt[atomicAdd(&counters[(round + 1) & 3], 1u)] = tile;
            atomicMin(&counters[8 + ((round + 1) & 3)], key);
        }
        else
            s_mine[atomicAdd(&s_n_mine, 1u)] = tile;
    }
    __syncthreads();
    const uint32_t n_mine = s_n_mine;
#ifdef PP_HEUR_TIMELINE
    const unsigned long long tl_list = gtime();
#endif
    // Pass 2: relax the collected tiles, one after the other
    for (uint32_t m = 0; m < n_mine; ++m)
    {
    const uint32_t tile = s_mine[m];
    __syncthreads();  // every thread is done with the previous tile's shared memory

    const uint32_t tx = tile % tiles_x, ty = tile / tiles_x;
    const int x0 = static_cast<int>(tx * HT), y0 = static_cast<int>(ty * HT);
    const int tid = threadIdx.x;
    const float INF = __int_as_float(0x7f800000);
    if (tid < 8)
        edge_min[tid] = INACTIVE;
    if (tid < 3)
        s_moved[tid] = 0;
    // A tile visit is a chain of latencies, not of bandwidth (a round with a single tile takes as long as
    // one with hundreds): every global load of the visit -- 34 x 34 field values, 32 x 32 costs -- is
    // issued before the first one is used, and the travel weights come from the shared-memory copy.
    constexpr int F_LOADS = ((HT + 2) * (HT + 2) + HT_THREADS - 1) / HT_THREADS;  // 5
    constexpr int C_LOADS = HT * HT / HT_THREADS;                                 // 4
    float fv[F_LOADS];
    uint32_t cv[C_LOADS];
#pragma unroll
    for (int j = 0; j < F_LOADS; ++j)
    {
        const int i = tid + j * HT_THREADS;
        const int r = i / (HT + 2), c = i - r * (HT + 2);
        const int gy = y0 + r - 1, gx = x0 + c - 1;
        fv[j] = INF;
        if (i < (HT + 2) * (HT + 2) && gy >= 0 && gy < static_cast<int>(H) && gx >= 0 && gx < static_cast<int>(W))
            fv[j] = field[static_cast<size_t>(gy) * W + gx];
    }
#pragma unroll
    for (int j = 0; j < C_LOADS; ++j)
    {
        const int i = tid + j * HT_THREADS;
        const int r = i / HT, c = i - r * HT;
        const int gy = y0 + r, gx = x0 + c;
        // border cells are never entered (astar_hybrid.cpp:141), nor cells with cost >= 253 (:143): weight INF
        cv[j] = 256u;
        if (gy > 0 && gy < static_cast<int>(H) - 1 && gx > 0 && gx < static_cast<int>(W) - 1)
            cv[j] = cost[static_cast<size_t>(gy) * W + gx];
    }
#pragma unroll
    for (int j = 0; j < F_LOADS; ++j)
    {
        const int i = tid + j * HT_THREADS;
        const int r = i / (HT + 2), c = i - r * (HT + 2);
        if (i < (HT + 2) * (HT + 2))
            F[r][c] = fv[j];
    }
#pragma unroll
    for (int j = 0; j < C_LOADS; ++j)
    {
        const int i = tid + j * HT_THREADS;
        const int r = i / HT, c = i - r * HT;
        WS[r][c] = cv[j] < 256u ? s_travel[cv[j]] : INF;
        WD[r][c] = cv[j] < 256u ? s_travel[256 + cv[j]] : INF;
    }
    __syncthreads();
#ifdef PP_HEUR_TIMELINE
    const unsigned long long tl_staged = gtime();
#endif

    // A warp owns a band of 4 consecutive tile rows, lane = column, and keeps its 4 x 32 values and their
    // travel weights in registers. Left / right neighbours come from the neighbouring lanes by shuffle
    // (the tile's halo columns, constant during a visit, for lanes 0 and 31), the rows above and below
    // the band from shared memory, re-read after every block-wide vote. Between two votes a warp relaxes
    // its band HT_LOCAL times on its own (Jacobi inside the band); stale values across bands are
    // harmless: values only decrease, towards the same fixed point.
    const int lane = tid & 31, band = (tid >> 5) * BR, c = lane + 1;
    float v[BR], ws[BR], wd[BR], edge_l[BR], edge_r[BR], init[BR];
#pragma unroll
    for (int k = 0; k < BR; ++k)
    {
        const int r = band + k + 1;
        v[k] = init[k] = F[r][c];
        ws[k] = WS[r - 1][c - 1];
        wd[k] = WD[r - 1][c - 1];
        edge_l[k] = F[r][0];
        edge_r[k] = F[r][HT + 1];
    }
    uint32_t iters = 0;

    // A band whose own values and whose neighbour bands did not change in the last vote is a fixed point of its
    // local relaxation with unchanged inputs: its warp sits the next vote out (same result, fewer instructions on
    // the schedulers the still-moving bands share). s_moved[n % 3] collects the bands that changed in vote n.
    const uint32_t my_bit = 1u << (tid >> 5);
    const uint32_t near_bits = my_bit | my_bit << 1 | my_bit >> 1;
    uint32_t moved = 0xFFFFFFFFu;  // first vote: every band
    uint32_t slot = 0;
    do
    {
        int local = 0;
        if (moved & near_bits)
        {
            float up[3], dn[3];  // the rows just outside the band: columns c-1, c, c+1
#pragma unroll
            for (int d = 0; d < 3; ++d)
            {
                up[d] = F[band][c - 1 + d];
                dn[d] = F[band + BR + 1][c - 1 + d];
            }
#pragma unroll 1
            for (int it = 0; it < HT_LOCAL; ++it)
            {
                float l[BR], r[BR];
#pragma unroll
                for (int k = 0; k < BR; ++k)
                {
                    l[k] = __shfl_up_sync(0xffffffffu, v[k], 1);
                    r[k] = __shfl_down_sync(0xffffffffu, v[k], 1);
                    if (lane == 0)
                        l[k] = edge_l[k];
                    if (lane == 31)
                        r[k] = edge_r[k];
                }
                float nv[BR];
#pragma unroll
                for (int k = 0; k < BR; ++k)
                {
                    const float a0 = k == 0 ? up[0] : l[k - 1], a1 = k == 0 ? up[1] : v[k - 1], a2 = k == 0 ? up[2] : r[k - 1];
                    const float b0 = k == BR - 1 ? dn[0] : l[k + 1], b1 = k == BR - 1 ? dn[1] : v[k + 1], b2 = k == BR - 1 ? dn[2] : r[k + 1];
                    const float ms = fminf(fminf(a1, b1), fminf(l[k], r[k]));
                    const float md = fminf(fminf(a0, b2), fminf(a2, b0));
                    nv[k] = fminf(__fadd_rn(ms, ws[k]), __fadd_rn(md, wd[k]));
                }
                int step = 0;
#pragma unroll
                for (int k = 0; k < BR; ++k)
                    if (nv[k] < v[k])
                    {
                        v[k] = nv[k];
                        step = 1;
                    }
                local |= step;
                if (!__any_sync(0xffffffffu, step))
                    break;  // the band is a fixed point of its inputs: the remaining steps would change nothing
            }
            if (local)
            {
#pragma unroll
                for (int k = 0; k < BR; ++k)
                    F[band + k + 1][c] = v[k];
            }
            if (__any_sync(0xffffffffu, local) && lane == 0)
                atomicOr(&s_moved[slot], my_bit);
        }
        const uint32_t next_slot = slot == 2 ? 0 : slot + 1;
        if (tid == 0)
            s_moved[next_slot] = 0;  // last read after the vote before the previous one
        __syncthreads();
        moved = s_moved[slot];
        slot = next_slot;
        ++iters;
    } while (moved);
#ifdef PP_HEUR_TIMELINE
    const unsigned long long tl_conv = gtime();
#endif

    // write back what changed and note which borders moved
    if (tid == 0)
        cur[tile] = INACTIVE;  // off the current generation; neighbours may list it again for the next round
    if (tid == 0)
    {
        atomicAdd(&counters[4], 1u);      // tile visits
        atomicAdd(&counters[5], iters);   // in-tile relaxation sweeps
    }
#pragma unroll
    for (int k = 0; k < BR; ++k)
    {
        const int r = band + k, c = lane;
        const float val = v[k];
        if (val < init[k])
        {
            const int gy = y0 + r, gx = x0 + c;
            if (gy < static_cast<int>(H) && gx < static_cast<int>(W))
                field[static_cast<size_t>(gy) * W + gx] = val;
            const uint32_t vb = __float_as_uint(val);
            if (r == 0) atomicMin(&edge_min[0], vb);                       // N
            if (r == HT - 1) atomicMin(&edge_min[1], vb);                  // S
            if (c == 0) atomicMin(&edge_min[2], vb);                       // W
            if (c == HT - 1) atomicMin(&edge_min[3], vb);                  // E
            if (r == 0 && c == 0) atomicMin(&edge_min[4], vb);             // NW
            if (r == 0 && c == HT - 1) atomicMin(&edge_min[5], vb);        // NE
            if (r == HT - 1 && c == 0) atomicMin(&edge_min[6], vb);        // SW
            if (r == HT - 1 && c == HT - 1) atomicMin(&edge_min[7], vb);   // SE
        }
    }
    __syncthreads();
    if (tid < 8 && edge_min[tid] != INACTIVE)
    {
        const int dxs[8] = {0, 0, -1, 1, -1, 1, -1, 1};
        const int dys[8] = {-1, 1, 0, 0, -1, -1, 1, 1};
        const int ntx = static_cast<int>(tx) + dxs[tid], nty = static_cast<int>(ty) + dys[tid];
        if (ntx >= 0 && ntx < static_cast<int>(tiles_x) && nty >= 0 && nty < static_cast<int>(tiles_y))
        {
            const uint32_t nt = nty * tiles_x + ntx;
            const uint32_t k = edge_min[tid];
            if (atomicMin(&nxt[nt], k) == INACTIVE)
                list_nxt[atomicAdd(&counters[(round + 1) & 3], 1u)] = nt;
            atomicMin(&counters[8 + ((round + 1) & 3)], k);
        }
    }
#ifdef PP_HEUR_TIMELINE
    __syncthreads();
    if (tid == 0)
    {
        const unsigned int slot = atomicAdd(&heur_tl_n, 1u);
        if (slot < (1u << 17))
        {
            unsigned long long *o = heur_tl + 9ull * slot;
            o[0] = round, o[1] = tile, o[2] = tl_entry, o[3] = tl_wait, o[4] = tl_list, o[5] = tl_staged, o[6] = tl_conv, o[7] = gtime(), o[8] = iters;
        }
    }
#endif
    }  // pass 2
    }  // batches of the work list
}

__global__ void heur_finalize_kernel(float *field, size_t n)
{
    const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t k = i; k < n; k += stride)
        if (field[k] == __int_as_float(0x7f800000))
            field[k] = 0.0f;  // unreachable: the reference leaves 0 (unseen) there
}
// bucket width of the tile schedule, in cost units (a straight step over a free cell costs 1)
float heur_delta()
{
    static const float d = [] {
        const char *e = getenv("PP_HEUR_DELTA");  // tuning knob for experiments
        const float v = e ? static_cast<float>(atof(e)) : 96.0f;
        return (v >= 0.0f && v <= 1e30f) ? v : 96.0f;  // a negative or NaN width would leave every bucket empty
    }();
    return d;
}

// travel-cost table (2 x 256: straight, diagonal) with the reference's float expression and evaluation
// order (astar_hybrid.cpp:151); entering a cell with cost >= 253 is impossible (:143-145)
void fill_travel_table(float cost_penalty, float (&travel)[512])
{
    const float sqrt_2 = sqrtf(2.0f);
    for (int c = 0; c < 256; ++c)
    {
        const float cost = static_cast<float>(c);
        const float scale = 1.0f + (cost_penalty * cost / 252.0f);
        travel[c] = c >= PP_COST_INSCRIBED ? std::numeric_limits<float>::infinity() : 1.0f * scale;
        travel[256 + c] = c >= PP_COST_INSCRIBED ? std::numeric_limits<float>::infinity() : sqrt_2 * scale;
    }
}

// Guard against a schedule that stops making progress. A round that has active tiles lowers at least one
// cell, and the values along any shortest path are fixed one 32-cell tile (or at worst a few cells of a
// serpentine corridor) per round, so W*H/8 rounds are out of reach of every map; typical: a few hundred.
uint64_t heur_max_rounds(size_t n_cells) { return n_cells / 8 + 1024; }
}  // namespace

extern "C" {

pp_status pp_heuristic_build(pp_handle h, uint32_t goal_x, uint32_t goal_y, float cost_penalty, float *field_out)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_cost, PP_ERR_STATE, "pp_heuristic_build: upload a costmap first");
    PP_REQUIRE(h, goal_x < h->W && goal_y < h->H, PP_ERR_INVALID, "goal (%u,%u) outside %ux%u", goal_x, goal_y, h->W, h->H);
    PP_CUDA(h, cudaSetDevice(h->device));
    const uint32_t W = h->W, H = h->H;
    const size_t n = static_cast<size_t>(W) * H;
    const uint32_t tiles_x = (W + HT - 1) / HT, tiles_y = (H + HT - 1) / HT, nT = tiles_x * tiles_y;
    const size_t n_flags = 4 * static_cast<size_t>(nT) + N_COUNTERS;
    PP_CUDA(h, pp_reserve(h->d_field, h->field_cap, n));
    PP_CUDA(h, pp_reserve(h->d_tile_flags, h->tile_flags_cap, n_flags));
    size_t travel_cap = h->d_travel ? 512 : 0;
    PP_CUDA(h, pp_reserve(h->d_travel, travel_cap, static_cast<size_t>(512)));

    float travel[512];
    fill_travel_table(cost_penalty, travel);
    PP_CUDA(h, cudaMemcpyAsync(h->d_travel, travel, sizeof(travel), cudaMemcpyHostToDevice, h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));  // `travel` lives on this stack frame

    const uint32_t goal_index = goal_y * W + goal_x;
    const uint32_t goal_tile = (goal_y / HT) * tiles_x + goal_x / HT;
    heur_init_kernel<<<148 * 8, 256, 0, h->stream>>>(h->d_field, n, h->d_tile_flags, n_flags, goal_index, goal_tile);
    h->launches++;
    PP_CUDA(h, cudaGetLastError());

    h->heur_rounds = 0;
    h->heur_tile_visits = 0;
    const uint64_t max_rounds = heur_max_rounds(n);
    uint32_t round = 0;
    bool done = false;
    while (!done)
    {
        for (int k = 0; k < ROUNDS_PER_CHECK; ++k, ++round)
        {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(std::min<uint32_t>(nT, ROUND_GRID));
            cfg.blockDim = dim3(HT_THREADS);
            cfg.stream = h->stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;  // see the top of heur_round_kernel
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            PP_CUDA(h, cudaLaunchKernelEx(&cfg, heur_round_kernel, h->d_field, static_cast<const uint8_t *>(h->d_cost),
                                          static_cast<const float *>(h->d_travel), W, H, tiles_x, tiles_y, h->d_tile_flags, round,
                                          heur_delta()));
            h->launches++;
        }
        PP_CUDA(h, cudaGetLastError());
        // counters[(round) & 3] holds the number of tiles flagged for the round that would run next
        PP_CUDA(h, cudaMemcpyAsync(h->h_counter, h->d_tile_flags + 4 * static_cast<size_t>(nT), N_COUNTERS * sizeof(uint32_t),
                                   cudaMemcpyDeviceToHost, h->stream));
        PP_CUDA(h, cudaStreamSynchronize(h->stream));
        done = h->h_counter[round & 3] == 0;
        PP_REQUIRE(h, done || round < max_rounds, PP_ERR_STATE, "pp_heuristic_build: relaxation did not converge");
    }
#ifdef PP_HEUR_TIMELINE
    if (const char *path = getenv("PP_HEUR_TIMELINE_FILE"))
    {
        unsigned int n_tl = 0;
        cudaMemcpyFromSymbol(&n_tl, heur_tl_n, sizeof(n_tl));
        n_tl = std::min(n_tl, 1u << 17);
        std::vector<unsigned long long> tl(9ull * n_tl);
        cudaMemcpyFromSymbol(tl.data(), heur_tl, tl.size() * sizeof(unsigned long long));
        if (FILE *f = fopen(path, "w"))
        {
            for (unsigned int i = 0; i < n_tl; ++i)
            {
                for (int k = 0; k < 9; ++k)
                    fprintf(f, "%llu ", tl[9ull * i + k]);
                fprintf(f, "\n");
            }
            fclose(f);
        }
        n_tl = 0;
        cudaMemcpyToSymbol(heur_tl_n, &n_tl, sizeof(n_tl));
    }
#endif
    h->heur_rounds = h->h_counter[6];  // rounds that had work
    h->heur_tile_visits = h->h_counter[4];
    h->heur_sweeps = h->h_counter[5];
    heur_finalize_kernel<<<148 * 8, 256, 0, h->stream>>>(h->d_field, n);
    h->launches++;
    PP_CUDA(h, cudaGetLastError());
    if (field_out)
    {
        PP_CUDA(h, cudaMemcpyAsync(field_out, h->d_field, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
        PP_CUDA(h, cudaStreamSynchronize(h->stream));
    }
    return PP_OK;
}

pp_status pp_heuristic_ptr(pp_handle h, const float **device_ptr_out)
{
    if (!h || !device_ptr_out)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_field, PP_ERR_STATE, "pp_heuristic_ptr: no field built");
    *device_ptr_out = h->d_field;
    return PP_OK;
}

pp_status pp_heuristic_stats(pp_handle h, uint32_t *rounds, uint64_t *tile_visits)
{
    if (!h)
        return PP_ERR_INVALID;
    if (rounds)
        *rounds = h->heur_rounds;
    if (tile_visits)
        *tile_visits = h->heur_tile_visits;
    return PP_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// pp_heuristic_build_noalloc: the same field, for callers that must not allocate.
//
// Who needs it: a planner thread that runs inside a hard real-time control loop (cudaMalloc and
// cudaFree may take the device-wide allocator lock, synchronise the device and take an
// unbounded time), a plugin loaded into a host process that has already handed all device memory
// to its own pool, and code that runs after start-up in a process whose policy forbids new device
// allocations. Such a caller owns its buffers; what it cannot supply is this kernel's scratch
// memory: the two generations of per-tile activity flags with their round counters, and the
// travel-cost table. Here that scratch lives in a fixed array in device global memory that is part
// of the library image, so every GPU that loads the library carries these ~260 KB until the
// process exits.
//
// Limit: at most 16384 tiles of 32x32 cells, i.e. ceil(W/32)*ceil(H/32) <= 16384 (4096 x 4096 fits);
// anything larger is refused with PP_ERR_RANGE.
//
// Set-up: the array's name means something only to device code; in host code it is a stand-in,
// not an address a kernel can use. So the first call on each device asks the runtime for the real
// device address (cudaGetSymbolAddress) with that device current and remembers it. That first call
// therefore does slightly more work than later ones and should be made before any stream capture
// begins on that device.
//
// Sharing: there is ONE array per device. Each call clears the flags and counters and rewrites the
// travel table before use, so nothing leaks from one call into the next, but two calls whose GPU
// work overlaps (different streams, or different threads) would trample each other's flags:
// serialise them.
namespace
{
constexpr uint32_t NOALLOC_MAX_TILES = (4096 / HT) * (4096 / HT);
constexpr size_t NOALLOC_FLAG_WORDS = 4 * static_cast<size_t>(NOALLOC_MAX_TILES) + N_COUNTERS;
constexpr size_t NOALLOC_WORDS = NOALLOC_FLAG_WORDS + 512;  // flags + counters, then the 2x256 travel table
constexpr int NOALLOC_MAX_DEVICES = 64;

static __device__ __align__(256) uint32_t g_noalloc_scratch[NOALLOC_WORDS];

std::atomic<uint32_t *> g_noalloc_addr[NOALLOC_MAX_DEVICES] = {};  // device address of the array, per device ordinal
}  // namespace

extern "C" pp_status pp_heuristic_build_noalloc(const uint8_t *d_cost, uint32_t W, uint32_t H, uint32_t goal_x,
                                                uint32_t goal_y, float cost_penalty, float *d_field_out,
                                                void *cuda_stream)
{
    if (!d_cost || !d_field_out || W == 0 || H == 0)
        return pp_fail(nullptr, PP_ERR_INVALID, "pp_heuristic_build_noalloc: null buffer or empty map");
    if (goal_x >= W || goal_y >= H)
        return pp_fail(nullptr, PP_ERR_INVALID, "pp_heuristic_build_noalloc: goal (%u,%u) outside %ux%u", goal_x, goal_y, W, H);
    const uint32_t tiles_x = (W + HT - 1) / HT, tiles_y = (H + HT - 1) / HT;
    const uint64_t nT64 = static_cast<uint64_t>(tiles_x) * tiles_y;
    if (nT64 > NOALLOC_MAX_TILES)
        return pp_fail(nullptr, PP_ERR_RANGE,
                       "pp_heuristic_build_noalloc: a %ux%u map needs %llu tiles of %dx%d cells; the built-in scratch "
                       "holds %u (e.g. 4096x4096). Use pp_heuristic_build for larger maps.",
                       W, H, static_cast<unsigned long long>(nT64), HT, HT, NOALLOC_MAX_TILES);
    const uint32_t nT = static_cast<uint32_t>(nT64);
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);

    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return pp_fail(nullptr, PP_ERR_CUDA, "pp_heuristic_build_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES)
        return pp_fail(nullptr, PP_ERR_RANGE, "pp_heuristic_build_noalloc: device ordinal %d not supported", dev);
    if (!g_noalloc_addr[dev].load(std::memory_order_acquire))
    {
        // once per device, with that device current: turn the host-side stand-in into the address
        // the kernels can dereference
        uint32_t *p = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return pp_fail(nullptr, PP_ERR_CUDA, "pp_heuristic_build_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        g_noalloc_addr[dev].store(p, std::memory_order_release);  // two first calls racing store the same address
    }
    uint32_t *flags = g_noalloc_addr[dev].load(std::memory_order_acquire);
    float *travel_dev = reinterpret_cast<float *>(flags + NOALLOC_FLAG_WORDS);

    float travel[512];
    fill_travel_table(cost_penalty, travel);
    e = cudaMemcpyAsync(travel_dev, travel, sizeof(travel), cudaMemcpyHostToDevice, stream);
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(stream);  // `travel` is on this stack frame
    if (e != cudaSuccess)
        return pp_fail(nullptr, PP_ERR_CUDA, "pp_heuristic_build_noalloc: travel table upload: %s", cudaGetErrorString(e));

    const size_t n = static_cast<size_t>(W) * H;
    const size_t n_flags = 4 * static_cast<size_t>(nT) + N_COUNTERS;
    const uint32_t goal_index = goal_y * W + goal_x;
    const uint32_t goal_tile = (goal_y / HT) * tiles_x + goal_x / HT;
    // the init kernel clears every flag and counter this call will touch, and seeds the goal
    heur_init_kernel<<<148 * 8, 256, 0, stream>>>(d_field_out, n, flags, n_flags, goal_index, goal_tile);
    uint32_t round = 0;
    uint32_t counters[N_COUNTERS];
    const uint64_t max_rounds = heur_max_rounds(n);
    for (;;)
    {
        for (int k = 0; k < ROUNDS_PER_CHECK; ++k, ++round)
            heur_round_kernel<<<std::min<uint32_t>(nT, ROUND_GRID), HT_THREADS, 0, stream>>>(d_field_out, d_cost, travel_dev, W, H, tiles_x, tiles_y, flags, round, heur_delta());
        e = cudaGetLastError();
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(counters, flags + 4 * static_cast<size_t>(nT), sizeof(counters), cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess)
            e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess)
            return pp_fail(nullptr, PP_ERR_CUDA, "pp_heuristic_build_noalloc: relaxation round %u: %s", round, cudaGetErrorString(e));
        if (counters[round & 3] == 0)
            break;
        if (round >= max_rounds)
            return pp_fail(nullptr, PP_ERR_STATE, "pp_heuristic_build_noalloc: relaxation did not converge");
    }
    heur_finalize_kernel<<<148 * 8, 256, 0, stream>>>(d_field_out, n);
    e = cudaGetLastError();
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess)
        return pp_fail(nullptr, PP_ERR_CUDA, "pp_heuristic_build_noalloc: %s", cudaGetErrorString(e));
    return PP_OK;
}
