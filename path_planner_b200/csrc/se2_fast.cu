// K1 fast path: tiled SE2 volume with shared-memory "brick" tables.
//
// For one heading bin the reference's footprint test is, away from the few places where float
// rounding makes the integer sample offsets position-dependent, a grey-scale dilation of the
// costmap by a fixed set S_t of integer offsets (collision_checker.cpp:155-166).
// max over uint8 is exact, idempotent and order-free, so S_t may be covered by translated copies
// of small "bricks" B (S_t = U_i (B_i + o_i), overlaps allowed, no point outside S_t), and
//   V[t][y][x] = max_i T_{B_i}[y + o_i.y][x + o_i.x],   T_B = dilation of the costmap tile by B.
// The brick tables are heading-independent: a CTA stages one 128x16 tile (+ halo) of the costmap in
// shared memory, derives the brick tables from it by doubling (each table is the max of two
// shifted earlier tables; the recipes are run-time data, so the library can be chosen per
// footprint), and then serves ALL heading bins of the launch from them, turning the P = 48 gathers
// per cell of the reference into ~15 aligned 32-bit shared-memory loads per four cells.
//
// Table layout ("column-transposed"): the four bytes of table word w of a row are the cells
// x0+w, x0+w+32, x0+w+64, x0+w+96 (x0 = tile origin - halo). A thread that owns output cells
// q, q+32, q+64, q+96 of a row therefore finds the sample at ANY x offset dx in the single aligned
// word q+halo+dx: one copy per table instead of four byte-shifted ones (2.6x less shared memory
// per table, no copy pass in the table build), a warp reads 32 consecutive words (no bank conflict
// for any pitch), and a warp's 32 output words are exactly one 128-byte row of the volume. The price
// is a 4x4 byte transpose across groups of four lanes (2 SHFL + 2 PRMT per word) before the store.
// Bins whose offset sets coincide (t and t + n_theta/2 for a centrally symmetric footprint --
// verified by the planner on the exact offsets, not assumed) are evaluated once per tile and
// stored to all of them.
//
// sm_100a has no native per-byte max (__vmaxu4 is emulated with ~7 instructions). It does have the
// 3-input VIMNMX3.U16x2. A u16 max is decided by the high byte, so on a raw word (b0,b1,b2,b3) it
// yields the correct maxima of b1 and b3, and on the word shifted left by 8 bits those of b0 and
// b2; the low bytes are don't-cares. Two accumulators per word, two samples per instruction;
// the shift is an IMAD on the otherwise idle FMA pipe.
//
// Exactness: the host planner evaluates the reference expression for every column, row, point
// and bin with the same float operations (and the host libm's cosf/sinf), finds the ranges of
// columns / rows over which all offsets are constant, and keeps the two longest per axis and bin.
// A tile inside one range is evaluated with that range's offset set; a tile that the boundary
// between the two ranges crosses ("mixed") is evaluated once per range and every evaluation stores
// only its own columns / rows; all other (tile, bin) pairs go to the exact generic kernel through a
// work list. Covers are minimum exact set covers found by greedy + branch and bound.
//
// Stores: a warp's 16 x 128-byte output block goes to a staging buffer and from there to the volume
// with one bulk-tensor (TMA) store per heading bin that shares the evaluation; the tensor map covers
// this GPU's own volume only. Volumes of peers (fused exchange) are written with plain stores.
#include "se2_plan.h"

#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched through cudaGetDriverEntryPoint)

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <set>
#include <utility>
#include <vector>

namespace
{
#ifndef PP_FT_THREADS
#define PP_FT_THREADS 384
#endif
#ifndef PP_CTAS_PER_SM
#define PP_CTAS_PER_SM 2
#endif
constexpr int FT_W = 128, FT_THREADS = PP_FT_THREADS;  // the tile height (16 or 32 rows) is a template parameter
constexpr int GEN_W = 64, GEN_H = 16;  // work-list tile of the exact generic kernel (se2_volume.cu)
constexpr int CTAS_PER_SM = PP_CTAS_PER_SM;
constexpr int MAX_TABLES = 24;
constexpr int MAX_T_PER_LAUNCH = 160;
constexpr int MAX_LIST_WORDS = 6400;
constexpr int MAX_R = 15;
constexpr int MIN_HALO = 7;
constexpr size_t SMEM_LIMIT = (227 * 1024 - 2048) / CTAS_PER_SM;  // per CTA; 1 KB per CTA is reserved by the system
#ifndef PP_QPT
#define PP_QPT 16
#endif
constexpr size_t STAGE_BYTES = static_cast<size_t>(PP_FT_THREADS / 32) * PP_QPT * 128 + 128;  // per-warp QPT x 128-byte store blocks (+ alignment)

// A brick library: table 0 is the costmap itself (the one-point brick); table k > 0 is
// T_k = max(T_a, T_b shifted by (du, dv)), a and b earlier tables. Listed in dependency-level order.
struct Recipe
{
    int a, b, du, dv;
};

struct FastArgs
{
    // the own volume as a 3-D uint8 tensor (x, y, bin) for bulk-tensor stores of 128 x 8 row blocks
    alignas(64) CUtensorMap vol_map;
    int use_tma;  // 0: plain 32-bit stores (peer volumes mapped, or rows not 16-byte aligned)
    const uint8_t *cost;
    uint8_t *vol;             // own volume
    long long peer_delta[7];  // byte distance from the own volume to each mapped peer volume
    int n_peers;              // fused exchange: every output word is also stored to these
    const uint8_t *xclass;  // [n_theta][tiles_x]: 0/1 = column range class, 255 = not uniform
    const uint8_t *yclass;  // [n_theta][tiles_y]
    uint32_t W, H, tiles_x, tiles_y;
    const uint32_t *tile_order;  // CTA b works on tile tile_order[b] = tile_x | tile_y << 16, the tiles with the most evaluations first
    int32_t n_tables;  // the halo (and with it the row pitch) is a template parameter of the kernel
    // table k >= 1: word i = bytemax(T[rec_a[k] + i], T_b[rec_d[k] + i]) with T_b = T + rec_b[k]; rec_d is the
    // shift dv * pitch + du inside table b (reads that leave table b give 0: they only feed entries whose
    // brick leaves the staged region, which no sample list references)
    int32_t rec_a[MAX_TABLES], rec_b[MAX_TABLES], rec_d[MAX_TABLES];
    uint32_t rec_sync;  // bit k: table k starts a new dependency level (block barrier before it)
    uint32_t n_t;                        // bins of this launch ...
    uint16_t bin[MAX_T_PER_LAUNCH];      // ... and which heading bin each one is
    // mixed tiles: first column (row) of the right-hand (upper) range of a bin and the class of the range before it
    uint16_t xsplit[MAX_T_PER_LAUNCH], ysplit[MAX_T_PER_LAUNCH];
    uint8_t split_left[MAX_T_PER_LAUNCH];  // bit 0: class of the columns before xsplit, bit 1: of the rows before ysplit
    uint16_t list_off[MAX_T_PER_LAUNCH][4];
    uint16_t list_len[MAX_T_PER_LAUNCH][4];
    // (bin, class) pairs of this launch with the SAME offset set form a chain in bin order (for a
    // centrally symmetric footprint bins t and t + n_theta/2 do): link = bin | class << 8, NO_LINK at the ends
    uint16_t chain_prev[MAX_T_PER_LAUNCH][4];
    uint16_t chain_next[MAX_T_PER_LAUNCH][4];
    alignas(16) uint32_t lists[MAX_LIST_WORDS];  // byte offsets into shared memory, relative to the thread's base
#ifdef PP_SE2_TIMELINE
    unsigned long long *timeline;  // dev build: per CTA {start, staged, built, end (ns, %globaltimer), SM id}
#endif
};
static_assert(MAX_T_PER_LAUNCH <= 255, "chain links keep the bin in 8 bits");
static_assert(MAX_TABLES <= 32, "rec_sync is a 32-bit mask");
static_assert(sizeof(FastArgs) < 32764, "kernel parameter block too large");

constexpr uint32_t NO_LINK = 0xFFFFu;

__device__ __forceinline__ uint32_t bytemax4(uint32_t a, uint32_t b)
{
    const uint32_t odd = __vmaxu2(a, b);                 // high byte of each half = max(b1), max(b3)
    const uint32_t even = __vmaxu2(a << 8, b << 8);      // high byte of each half = max(b0), max(b2)
    return __byte_perm(even, odd, 0x7351);
}

constexpr int QPT = PP_QPT;                  // rows per thread and work unit
static_assert(QPT == 16, "tile heights 16 and 32 are whole work units (QPT rows x 128 columns of one bin: one warp, lane = word column)");

template <int PW>
__device__ __forceinline__ void load_pair(const uint8_t *smb, uint32_t d1, uint32_t d2, uint32_t (&u)[QPT], uint32_t (&v)[QPT])
{
    // the row pitch is a template parameter, so the QPT row offsets are immediates: LDS [R + imm]
    const uint32_t *p1 = reinterpret_cast<const uint32_t *>(smb + d1);
    const uint32_t *p2 = reinterpret_cast<const uint32_t *>(smb + d2);
#pragma unroll
    for (int j = 0; j < QPT; ++j)
    {
        u[j] = p1[j * PW];
        v[j] = p2[j * PW];
    }
}

__device__ __forceinline__ void accumulate(uint32_t (&odd)[QPT], uint32_t (&even)[QPT], const uint32_t (&u)[QPT], const uint32_t (&v)[QPT])
{
#pragma unroll
    for (int j = 0; j < QPT; ++j)
    {
        odd[j] = __vimax3_u16x2(odd[j], u[j], v[j]);
        even[j] = __vimax3_u16x2(even[j], u[j] * 256u, v[j] * 256u);
    }
}

__device__ __forceinline__ void start(uint32_t (&odd)[QPT], uint32_t (&even)[QPT], const uint32_t (&u)[QPT], const uint32_t (&v)[QPT])
{
#pragma unroll
    for (int j = 0; j < QPT; ++j)
    {
        odd[j] = __vmaxu2(u[j], v[j]);
        even[j] = __vmaxu2(u[j] * 256u, v[j] * 256u);
    }
}

// A range boundary of a bin crosses the tile ("mixed" tile): an evaluation with the offset set of one class pair
// is valid for the columns / rows of those classes only. Rare (the boundaries sit in the first tiles of a few
// bins), so this is an out-of-line call with plain stores: whole words where all four columns are in, bytes else.
__device__ __noinline__ void store_mixed(const FastArgs &a, uint32_t cls, uint32_t t, uint32_t lc, int gx, int gy, uint32_t r0, uint32_t r1,
                                         uint32_t r2, uint32_t r3, uint32_t r4, uint32_t r5, uint32_t r6, uint32_t r7)
{
    const uint32_t res[8] = {r0, r1, r2, r3, r4, r5, r6, r7};
    const int W = static_cast<int>(a.W), H = static_cast<int>(a.H);
    const uint32_t mcx = cls % 3u, mcy = cls / 3u;
    const int xsplit = a.xsplit[t], ysplit = a.ysplit[t];
    const uint32_t xleft = a.split_left[t] & 1u, yleft = (a.split_left[t] >> 1) & 1u;
    uint32_t xm = 0;  // bit b: column gx + b is inside the map and in this evaluation's column class
#pragma unroll
    for (int b = 0; b < 4; ++b)
        if (gx + b < W && (mcx != 2 || ((gx + b < xsplit) ? xleft : 1u - xleft) == (lc >> 1)))
            xm |= 1u << b;
    if (xm == 0)
        return;
    uint8_t *base = a.vol + static_cast<size_t>(a.W) * a.H * a.bin[t] + gx;
#pragma unroll
    for (int j = 0; j < 8; ++j)
    {
        const int y = gy + j;
        if (y >= H || (mcy == 2 && ((y < ysplit) ? yleft : 1u - yleft) != (lc & 1u)))
            continue;
        uint8_t *dst = base + static_cast<size_t>(y) * a.W;
        for (int p = -1; p < a.n_peers; ++p)  // own volume, then the peers' over NVLink
        {
            uint8_t *d = p < 0 ? dst : dst + a.peer_delta[p];
            if (xm == 0xFu)
                *reinterpret_cast<uint32_t *>(d) = res[j];
            else
                for (int b = 0; b < 4; ++b)
                    if (xm >> b & 1u)
                        d[b] = static_cast<uint8_t>(res[j] >> (8 * b));
        }
    }
}

// Output phase. A work unit is QPT rows of one evaluating bin: one warp, lane q owns the cells q, q+32, q+64,
// q+96 of each row. Warps fetch units from a shared counter, so bins with long and short sample lists
// balance out. The loads of the next sample pair are issued before the max operations of the current one
// (software pipelining by hand). The result is transposed across groups of four lanes, so that lane 4g+i
// holds the four consecutive cells 4g+32i .. 4g+32i+3 and a warp's store is one whole 128-byte row, and is
// stored to every bin of the evaluating bin's chain that is in the matching class on this tile.
// A tile that the boundary between a bin's two column (row) ranges crosses is "mixed": the unit is evaluated
// once per range and the results are blended per column (row).
constexpr uint32_t LEAD_MIXED = 0x80;  // s_lead_lc flag: the evaluation covers only its classes' part of a mixed tile
constexpr uint32_t CLS_MIXED0 = 4;  // s_cls: 0..3 uniform class cx*2+cy; 4 + cx + 3*cy with cx, cy in {0,1,2 = mixed}; 255 generic
template <int PW, int TH>
__device__ __forceinline__ void run_units(const FastArgs &a, const uint32_t *sm, const uint8_t *s_cls, const uint8_t *s_lead, const uint8_t *s_lead_lc,
                                          uint32_t n_lead, uint32_t *s_next, int lane, int tx0, int ty0, uint32_t *stage)
{
    const int W = static_cast<int>(a.W), H = static_cast<int>(a.H);
    // after the transpose this lane holds word (lane >> 2) + 8 * (lane & 3) of the 32-word row
    const int out_word = (lane >> 2) + 8 * (lane & 3);
    const int gx = tx0 + 4 * out_word;
    const bool in_x = gx < W;
    const uint32_t plane = a.W * a.H;  // cells of one bin (maps are at most 32768 x 32768)
    constexpr int UNITS_PER_BIN = TH / QPT;
    const uint32_t n_units = UNITS_PER_BIN * n_lead;
    const uint32_t cell0 = static_cast<uint32_t>(ty0) * a.W + static_cast<uint32_t>(gx);  // < 2^30
    const bool tile_inside = tx0 + FT_W <= W && ty0 + TH <= H;
    // lane roles of the two transpose steps (selectors of the byte permutes; two rows travel in one shuffle)
    const bool odd_lane = lane & 1, upper_lane = lane & 2;
    const uint32_t give1 = odd_lane ? 0x6420u : 0x7531u, keep1a = odd_lane ? 0x3514u : 0x5240u, keep1b = odd_lane ? 0x3716u : 0x7260u;
    const uint32_t give2 = upper_lane ? 0x5410u : 0x7632u, keep2a = upper_lane ? 0x3254u : 0x5410u, keep2b = upper_lane ? 0x3276u : 0x7610u;
    while (true)
    {
        uint32_t unit = 0;
        if (lane == 0)
            asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(unit) : "r"(static_cast<uint32_t>(__cvta_generic_to_shared(s_next))) : "memory");
        unit = __shfl_sync(0xffffffffu, unit, 0);
        if (unit >= n_units)
            break;
        // s_lead lists the (bin, class pair) evaluations of this tile (see the kernel prologue)
        const uint32_t t = s_lead[unit / UNITS_PER_BIN];
        const uint32_t lcm = s_lead_lc[unit / UNITS_PER_BIN];
        const uint32_t lc = lcm & 3u;  // column class * 2 + row class
        const int row0 = QPT * static_cast<int>(unit % UNITS_PER_BIN);
        const uint8_t *smb = reinterpret_cast<const uint8_t *>(sm + row0 * PW + lane);
        const int gy = ty0 + row0;
        const uint32_t *list = a.lists + a.list_off[t][lc];
        const uint32_t n_samples = a.list_len[t][lc];  // >= 1
        const uint32_t len = n_samples & ~1u;           // samples consumed in pairs
        uint32_t odd[QPT], even[QPT];
        uint32_t ua[QPT], va[QPT], ub[QPT], vb[QPT];
        if (len)
        {
            load_pair<PW>(smb, list[0], list[1], ua, va);
            uint32_t i = 2;
            if (i < len)
            {
                load_pair<PW>(smb, list[2], list[3], ub, vb);
                start(odd, even, ua, va);  // the accumulators start from the first pair
                i = 4;
#pragma unroll 1
                for (; i + 2 < len; i += 4)
                {
                    load_pair<PW>(smb, list[i], list[i + 1], ua, va);
                    accumulate(odd, even, ub, vb);
                    load_pair<PW>(smb, list[i + 2], list[i + 3], ub, vb);
                    accumulate(odd, even, ua, va);
                }
                if (i < len)
                {
                    load_pair<PW>(smb, list[i], list[i + 1], ua, va);
                    accumulate(odd, even, ub, vb);
                    accumulate(odd, even, ua, va);
                }
                else
                {
                    accumulate(odd, even, ub, vb);
                }
            }
            else
            {
                start(odd, even, ua, va);
            }
        }
        else
        {
#pragma unroll
            for (int j = 0; j < QPT; ++j)
                odd[j] = even[j] = 0;
        }
        if (n_samples & 1u)  // the last sample of an odd list: one load, used for both operands
        {
            const uint32_t *p1 = reinterpret_cast<const uint32_t *>(smb + list[n_samples - 1]);
#pragma unroll
            for (int j = 0; j < QPT; ++j)
                ua[j] = p1[j * PW];
            accumulate(odd, even, ua, ua);
        }
        // 4x4 byte transpose across lanes 4g .. 4g+3. The valid bytes are the high bytes of the 16-bit halves
        // (odd = (., b1, ., b3), even = (., b0, ., b2), b_i = the lane's cell q + 32 i); they are merged first.
        // Step 1 (lane ^ 1): even lanes collect (b0, b2) of both lanes, odd lanes (b1, b3).
        // Step 2 (lane ^ 2): lower lanes collect bytes 0..1 of both, upper lanes bytes 2..3.
        // A lane gives away half of its bytes in each step, so the halves of two rows share one shuffle
        // (shuffles go through the shared-memory pipe, the one this kernel is bound by).
        uint32_t res[QPT];
#pragma unroll
        for (int j = 0; j < QPT; j += 2)
        {
            const uint32_t m0 = __byte_perm(even[j], odd[j], 0x7351), m1 = __byte_perm(even[j + 1], odd[j + 1], 0x7351);
            const uint32_t got = __shfl_xor_sync(0xffffffffu, __byte_perm(m0, m1, give1), 1);
            const uint32_t p0 = __byte_perm(m0, got, keep1a), p1 = __byte_perm(m1, got, keep1b);
            const uint32_t got2 = __shfl_xor_sync(0xffffffffu, __byte_perm(p0, p1, give2), 2);
            res[j] = __byte_perm(p0, got2, keep2a);
            res[j + 1] = __byte_perm(p1, got2, keep2b);
        }
        if (lcm & LEAD_MIXED)
        {
#pragma unroll
            for (int g = 0; g < QPT; g += 8)
                store_mixed(a, s_cls[t] - CLS_MIXED0, t, lc, gx, gy + g, res[g], res[g + 1], res[g + 2], res[g + 3], res[g + 4], res[g + 5],
                            res[g + 6], res[g + 7]);
            continue;
        }
        uint32_t link = a.chain_next[t][lc];
        if (a.use_tma)
        {
            // The warp's 8 x 128-byte block goes to its staging buffer and from there to every bin of the
            // chain with one bulk-tensor store each (the TMA unit clips at the map's edges). The buffer is
            // reused once the previous unit's stores have read it.
            if (lane == 0)
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
#pragma unroll
            for (int j = 0; j < QPT; ++j)
                stage[j * 32 + out_word] = res[j];
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0)
            {
                const uint32_t src = static_cast<uint32_t>(__cvta_generic_to_shared(stage));
                uint32_t tb_ = t;
                while (true)
                {
                    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(&a.vol_map), "r"(tx0),
                                 "r"(gy), "r"(static_cast<int>(a.bin[tb_])), "r"(src)
                                 : "memory");
                    uint32_t t2 = NO_LINK;
                    while (link != NO_LINK && t2 == NO_LINK)
                    {
                        const uint32_t lt = link & 0xFFu, lc = link >> 8;
                        link = a.chain_next[lt][lc];
                        if (s_cls[lt] == lc)
                            t2 = lt;
                    }
                    if (t2 == NO_LINK)
                        break;
                    tb_ = t2;
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            continue;
        }
        const uint32_t lane_off = cell0 + static_cast<uint32_t>(row0) * a.W;  // this lane's word of the unit's first row, < 2^31
        const int rows_ok = H - gy;  // rows of the unit inside the map
        uint8_t *out = a.vol + (static_cast<uint64_t>(plane) * a.bin[t] + lane_off);
        while (true)
        {
            if (tile_inside && a.n_peers == 0)
            {
                uint8_t *dst = out;  // one 64-bit add per row instead of a product and an add
#pragma unroll
                for (int j = 0; j < QPT; ++j, dst += a.W)
                    *reinterpret_cast<uint32_t *>(dst) = res[j];
            }
            else if (in_x)
            {
#pragma unroll
                for (int j = 0; j < QPT; ++j)
                    if (j < rows_ok)
                    {
                        uint8_t *dst = out + static_cast<uint32_t>(j) * a.W;
                        *reinterpret_cast<uint32_t *>(dst) = res[j];
                        for (int p = 0; p < a.n_peers; ++p)  // peer volumes over NVLink
                            *reinterpret_cast<uint32_t *>(dst + a.peer_delta[p]) = res[j];
                    }
            }
            // next bin of the chain that is in the matching class on this tile
            uint32_t t2 = NO_LINK;
            while (link != NO_LINK && t2 == NO_LINK)
            {
                const uint32_t lt = link & 0xFFu, lc = link >> 8;
                link = a.chain_next[lt][lc];
                if (s_cls[lt] == lc)
                    t2 = lt;
            }
            if (t2 == NO_LINK)
                break;
            out = a.vol + (static_cast<uint64_t>(plane) * a.bin[t2] + lane_off);
        }
    }
    // the staging buffer must outlive the bulk stores that read it
    if (a.use_tma && lane == 0)
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// Shared-memory geometry, compile-time (the halo is a template parameter) so that the row offsets of the
// output loop are immediates.
template <int HALO_, int TH>
struct Geo
{
    static constexpr int HALO = HALO_;
    static constexpr int PW = 32 + 2 * HALO;  // words per table row
    static constexpr int ROWS = TH + 2 * HALO;
    static constexpr int CW = ROWS * PW;      // words per table
};
constexpr int halo_for(int reach) { return reach < MIN_HALO ? MIN_HALO : reach; }

template <int HALO, int TH>
__global__ void __launch_bounds__(FT_THREADS, CTAS_PER_SM) se2_fast_kernel(const __grid_constant__ FastArgs a)
{
    extern __shared__ __align__(128) uint32_t tb[];
    __shared__ uint8_t s_cls[MAX_T_PER_LAUNCH];
    __shared__ uint8_t s_lead[4 * MAX_T_PER_LAUNCH];     // evaluations of this tile: bin ...
    __shared__ uint8_t s_lead_lc[4 * MAX_T_PER_LAUNCH];  // ... and its class pair (| LEAD_MIXED)
    __shared__ uint32_t s_next, s_nlead;
    using G = Geo<HALO, TH>;
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    constexpr int NWARPS = FT_THREADS / 32;
    constexpr int ROW_ITERS = (G::ROWS + NWARPS - 1) / NWARPS;
    constexpr int PW = G::PW, CW = G::CW;
#ifdef PP_SE2_TIMELINE
    unsigned long long tl0 = 0, tl1 = 0, tl2 = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl0));
#endif
    const uint32_t tile_xy = __ldg(a.tile_order + blockIdx.x);
    const uint32_t tile_x = tile_xy & 0xFFFFu, tile_y = tile_xy >> 16;
    const int tx0 = static_cast<int>(tile_x) * FT_W, ty0 = static_cast<int>(tile_y) * TH;
    const int W = static_cast<int>(a.W), H = static_cast<int>(a.H);

    // bin classes of this tile (one bin per thread), requested before anything else
    uint32_t my_cx = 255u, my_cy = 255u;
    if (static_cast<uint32_t>(tid) < a.n_t)
    {
        my_cx = __ldg(a.xclass + static_cast<size_t>(a.bin[tid]) * a.tiles_x + tile_x);
        my_cy = __ldg(a.yclass + static_cast<size_t>(a.bin[tid]) * a.tiles_y + tile_y);
    }
    // Costmap tile with halo, one warp per row. Lane l fetches the four bytes of table word l (cells
    // x0+l+32i: each of the four loads of a warp reads 32 consecutive bytes) and, for l < 2*HALO, the one
    // byte that word 32+l has beyond word l shifted down by a byte. Every global load of the tile is issued
    // before the first use; zeros outside the map (the identity of max: the reference ignores such samples).
    const int x0 = tx0 - HALO;
    uint32_t cell[ROW_ITERS][5];
    if (x0 >= 0 && x0 + FT_W + 2 * HALO <= W && ty0 >= HALO && ty0 + TH + HALO <= H)
    {
        // the tile and its halo lie inside the map: no bounds tests
        const uint8_t *p0 = a.cost + static_cast<size_t>(ty0 - HALO + warp) * a.W + (x0 + lane);
#pragma unroll
        for (int j = 0; j < ROW_ITERS; ++j)
        {
            const uint8_t *rowp = p0 + static_cast<size_t>(j * NWARPS) * a.W;
            const bool row_ok = warp + j * NWARPS < G::ROWS;
#pragma unroll
            for (int i = 0; i < 5; ++i)
                cell[j][i] = (row_ok && (i < 4 || lane < 2 * HALO)) ? static_cast<uint32_t>(__ldg(rowp + 32 * i)) : 0u;
        }
    }
    else
    {
#pragma unroll
        for (int j = 0; j < ROW_ITERS; ++j)
        {
            const int r = warp + j * NWARPS;
            const int gy = ty0 - HALO + r;
            const bool row_ok = r < G::ROWS && gy >= 0 && gy < H;
            const uint8_t *rowp = a.cost + static_cast<size_t>(row_ok ? gy : 0) * a.W;
#pragma unroll
            for (int i = 0; i < 5; ++i)
            {
                const int gxi = x0 + lane + 32 * i;
                const bool ok = row_ok && gxi >= 0 && gxi < W && (i < 4 || lane < 2 * HALO);
                cell[j][i] = ok ? static_cast<uint32_t>(__ldg(rowp + gxi)) : 0u;
            }
        }
    }
    static_assert(MAX_T_PER_LAUNCH <= FT_THREADS, "one bin class per thread");
    // 0..3: uniform class pair; CLS_MIXED0 + cx + 3 cy: a range boundary crosses the tile; 255: generic work list
    const uint32_t my_cls = (my_cx > 2 || my_cy > 2) ? 255u : (my_cx == 2 || my_cy == 2) ? CLS_MIXED0 + my_cx + 3 * my_cy : my_cx * 2 + my_cy;
    if (static_cast<uint32_t>(tid) < a.n_t)
        s_cls[tid] = static_cast<uint8_t>(my_cls);
    if (tid == 0)
        s_next = s_nlead = 0;
#pragma unroll
    for (int j = 0; j < ROW_ITERS; ++j)
    {
        const int r = warp + j * NWARPS;
        if (r < G::ROWS)
        {
            const uint32_t w0 = cell[j][0] | cell[j][1] << 8 | cell[j][2] << 16 | cell[j][3] << 24;
            tb[r * PW + lane] = w0;
            if (lane < 2 * HALO)
                tb[r * PW + 32 + lane] = w0 >> 8 | cell[j][4] << 24;
        }
    }
    const int any_fast = __syncthreads_or(my_cls != 255u);
#ifdef PP_SE2_TIMELINE
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl1));
#endif
    if (!any_fast)
        return;  // every bin of this tile belongs to the generic work list

    // Bins whose offset sets coincide on this tile have identical outputs: the first bin of such a
    // chain (host-built, in bin order) evaluates and stores to all of them. One bin per thread.
    if (my_cls != 255u)
    {
        bool evaluated_earlier = false;  // a mixed tile of a bin is always evaluated for that bin alone
        for (uint32_t link = my_cls < CLS_MIXED0 ? a.chain_prev[tid][my_cls] : NO_LINK; link != NO_LINK && !evaluated_earlier;
             link = a.chain_prev[link & 0xFFu][link >> 8])
            evaluated_earlier = s_cls[link & 0xFFu] == (link >> 8);
        if (my_cls < CLS_MIXED0)
        {
            if (!evaluated_earlier)
            {
                const uint32_t k = atomicAdd(&s_nlead, 1u);
                s_lead[k] = static_cast<uint8_t>(tid);
                s_lead_lc[k] = static_cast<uint8_t>(my_cls);
            }
        }
        else
        {
            // one evaluation per class pair present on the tile
            const uint32_t cx_lo = my_cx == 2 ? 0 : my_cx, cx_hi = my_cx == 2 ? 1 : my_cx;
            const uint32_t cy_lo = my_cy == 2 ? 0 : my_cy, cy_hi = my_cy == 2 ? 1 : my_cy;
            uint32_t k = atomicAdd(&s_nlead, (cx_hi - cx_lo + 1) * (cy_hi - cy_lo + 1));
            for (uint32_t cxv = cx_lo; cxv <= cx_hi; ++cxv)
                for (uint32_t cyv = cy_lo; cyv <= cy_hi; ++cyv, ++k)
                {
                    s_lead[k] = static_cast<uint8_t>(tid);
                    s_lead_lc[k] = static_cast<uint8_t>(cxv * 2 + cyv) | LEAD_MIXED;
                }
        }
    }

    // brick tables: T_k = max(T_a, T_b shifted); tables of one dependency level need no barrier between them
    constexpr int ITERS = (CW + FT_THREADS - 1) / FT_THREADS;
    for (int k = 1; k < a.n_tables; ++k)
    {
        if (a.rec_sync >> k & 1u)
            __syncthreads();
        const uint32_t *pa = tb + a.rec_a[k], *pb = tb + a.rec_b[k];
        const int d = a.rec_d[k];
        uint32_t *pd = tb + k * CW;
        uint32_t va[ITERS], vb[ITERS];
#pragma unroll
        for (int it = 0; it < ITERS; ++it)
        {
            const int i = tid + it * FT_THREADS;
            va[it] = vb[it] = 0;
            if (i < CW)
            {
                va[it] = pa[i];
                if (static_cast<uint32_t>(i + d) < static_cast<uint32_t>(CW))
                    vb[it] = pb[i + d];
            }
        }
#pragma unroll
        for (int it = 0; it < ITERS; ++it)
        {
            const int i = tid + it * FT_THREADS;
            if (i < CW)
                pd[i] = bytemax4(va[it], vb[it]);
        }
    }
    __syncthreads();  // also orders s_lead / s_nlead when no table beyond the costmap is built
#ifdef PP_SE2_TIMELINE
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl2));
#endif

    // output phase: warps fetch (evaluating bin, QPT rows) units until none are left
    // per-warp staging blocks of the bulk stores sit behind the tables, 128-byte aligned
    uint32_t *stage = tb + ((a.n_tables * CW + 31) & ~31) + warp * (QPT * 32);
    run_units<PW, TH>(a, tb, s_cls, s_lead, s_lead_lc, s_nlead, &s_next, lane, tx0, ty0, stage);
#ifdef PP_SE2_TIMELINE
    __syncthreads();
    if (tid == 0 && a.timeline)
    {
        unsigned long long tl3;
        uint32_t smid;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl3));
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        unsigned long long *o = a.timeline + 5ull * blockIdx.x;
        o[0] = tl0, o[1] = tl1, o[2] = tl2, o[3] = tl3, o[4] = smid;
    }
#endif
}

// one instantiation per halo and tile height
template <class F>
auto dispatch_fast(int halo, int th, F f)
{
#define PP_CASE(HL) \
    if (halo == HL) return th == 32 ? f(se2_fast_kernel<HL, 32>) : f(se2_fast_kernel<HL, 16>);
    PP_CASE(7) PP_CASE(8) PP_CASE(9) PP_CASE(10) PP_CASE(11) PP_CASE(12) PP_CASE(13) PP_CASE(14)
#undef PP_CASE
    return th == 32 ? f(se2_fast_kernel<15, 32>) : f(se2_fast_kernel<15, 16>);
}
cudaError_t set_fast_smem(int halo, int th, size_t bytes)
{
    return dispatch_fast(halo, th, [&](auto kernel) { return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)); });
}
void launch_fast(int halo, int th, dim3 grid, size_t smem, cudaStream_t stream, const FastArgs &args)
{
    dispatch_fast(halo, th, [&](auto kernel) { kernel<<<grid, FT_THREADS, smem, stream>>>(args); return cudaSuccess; });
}

// ---------------------------------------------------------------------------------------------
// host planner
// ---------------------------------------------------------------------------------------------
using Library = std::vector<Recipe>;
// Fixed libraries, listed in dependency-level order; the planner picks the one with the smallest covers
// for the footprint at hand and truncates it to the tables that fit.
//   STAIRS  2-runs along x, y and the two diagonals, four 4-cell "stair" bricks (two 2-runs offset by
//           (2,+-1) or (+-1,2)) that follow rasterised oblique edges, then runs of 4 along x and y
//           and of 8 along x (the stairs come first so that a truncated library keeps them), then the two
//           diagonals of 4 and the run of 8 along y (found by tools/lib_search.cpp as the next best bricks)
//   RUNS3   runs of 2/3/6 along x and y, diagonals of 2/3/4
//   BLOCKS  rectangles 2x2 .. 8x8 for solid footprints
//   RUNS    runs of 2/4/8 along x and y, diagonals of 2/4
const std::vector<Library> &fixed_libraries()
{
    static const std::vector<Library> libs = {
        {{0, 0, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}, {0, 0, 1, 1}, {0, 0, 1, -1}, {1, 1, 2, 1}, {1, 1, 2, -1}, {2, 2, 1, 2}, {2, 2, -1, 2}, {1, 1, 2, 0}, {2, 2, 0, 2}, {9, 9, 4, 0}, {4, 4, 2, -2}, {3, 3, 2, 2}, {10, 10, 0, 4}},
        {{0, 0, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}, {0, 0, 1, 1}, {0, 0, 1, -1}, {1, 0, 2, 0}, {2, 0, 0, 2}, {3, 0, 2, 2}, {4, 0, 2, -2}, {3, 3, 2, 2}, {5, 5, 3, 0}, {6, 6, 0, 3}},
        {{0, 0, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}, {1, 1, 0, 1}, {1, 1, 2, 0}, {2, 2, 0, 2}, {3, 3, 2, 0}, {4, 4, 4, 0}, {5, 5, 0, 4}, {6, 6, 0, 2}, {9, 9, 4, 0}, {10, 10, 0, 4}},
        {{0, 0, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}, {0, 0, 1, 1}, {0, 0, 1, -1}, {1, 1, 2, 0}, {2, 2, 0, 2}, {3, 3, 2, 2}, {4, 4, 2, -2}, {5, 5, 4, 0}, {6, 6, 0, 4}},
    };
    return libs;
}
// dependency level of every table of a library (table 0: level 0)
std::vector<int> library_levels(const Library &lib)
{
    std::vector<int> lvl(lib.size(), 0);
    for (size_t k = 1; k < lib.size(); ++k)
        lvl[k] = 1 + std::max(lvl[lib[k].a], lvl[lib[k].b]);
    return lvl;
}

using Offsets = std::vector<std::pair<int, int>>;  // sorted, unique (dx, dy)

struct Placement
{
    int brick, ox, oy;
};

std::vector<Offsets> brick_points(const Library &lib, int n)
{
    std::vector<Offsets> pts(n);
    pts[0] = {{0, 0}};
    for (int k = 1; k < n; ++k)
    {
        Offsets p = pts[lib[k].a];
        for (auto &q : pts[lib[k].b])
            p.emplace_back(q.first + lib[k].du, q.second + lib[k].dv);
        std::sort(p.begin(), p.end());
        p.erase(std::unique(p.begin(), p.end()), p.end());
        pts[k] = p;
    }
    return pts;
}

// Minimum exact cover of S by brick placements that lie entirely inside S (overlaps allowed).
// Cells of S are numbered by rank, so a placement is a bit mask of at most MASK_WORDS*64 bits.
// Greedy gives the first solution; a depth-first branch and bound (branch on the uncovered cell
// with the fewest covering placements, bound by ceil(uncovered / largest brick)) then improves it
// within a node budget. Deterministic.
constexpr int MASK_WORDS = 3;
struct Mask
{
    uint64_t w[MASK_WORDS];
    bool empty() const { return (w[0] | w[1] | w[2]) == 0; }
    int count() const { return __builtin_popcountll(w[0]) + __builtin_popcountll(w[1]) + __builtin_popcountll(w[2]); }
    Mask andnot(const Mask &o) const { return {{w[0] & ~o.w[0], w[1] & ~o.w[1], w[2] & ~o.w[2]}}; }
    int overlap(const Mask &o) const
    {
        return __builtin_popcountll(w[0] & o.w[0]) + __builtin_popcountll(w[1] & o.w[1]) + __builtin_popcountll(w[2] & o.w[2]);
    }
    bool subset_of(const Mask &o) const { return ((w[0] & ~o.w[0]) | (w[1] & ~o.w[1]) | (w[2] & ~o.w[2])) == 0; }
    bool operator==(const Mask &o) const { return w[0] == o.w[0] && w[1] == o.w[1] && w[2] == o.w[2]; }
    bool test(int b) const { return (w[b >> 6] >> (b & 63)) & 1ull; }
};

struct CoverSearch
{
    struct Cand
    {
        Placement pl;
        Mask mask;
        int npts;
    };
    std::vector<Cand> cands;
    std::vector<std::vector<int>> covering;  // per cell: candidates that contain it
    int max_pts = 1;
    std::vector<int> best, stack;
    long nodes = 0, node_budget = 0;

    void dfs(const Mask &unc)
    {
        if (unc.empty())
        {
            if (stack.size() < best.size())
                best = stack;
            return;
        }
        if (++nodes > node_budget)
            return;
        const int left = unc.count();
        if (static_cast<int>(stack.size()) + (left + max_pts - 1) / max_pts >= static_cast<int>(best.size()))
            return;
        // the uncovered cell with the fewest useful placements
        int cell = -1, cell_n = 1 << 30;
        for (int i = 0; i < MASK_WORDS; ++i)
        {
            uint64_t m = unc.w[i];
            while (m)
            {
                const int b = i * 64 + __builtin_ctzll(m);
                m &= m - 1;
                const int n = static_cast<int>(covering[b].size());
                if (n < cell_n)
                {
                    cell_n = n;
                    cell = b;
                }
            }
        }
        // its placements by decreasing gain
        std::vector<std::pair<int, int>> order;
        for (int c : covering[cell])
            order.emplace_back(-cands[c].mask.overlap(unc), c);
        std::sort(order.begin(), order.end());
        for (auto &o : order)
        {
            stack.push_back(o.second);
            dfs(unc.andnot(cands[o.second].mask));
            stack.pop_back();
            if (nodes > node_budget)
                return;
        }
    }
};

std::vector<Placement> cover_set(const Offsets &S, const std::vector<Offsets> &bricks, int R, long node_budget)
{
    const int n = static_cast<int>(S.size());
    const int side = 2 * R + 1;
    std::vector<int> rank(static_cast<size_t>(side) * side, -1);
    auto idx = [&](int x, int y) { return (y + R) * side + (x + R); };
    for (int i = 0; i < n; ++i)
        rank[idx(S[i].first, S[i].second)] = i;
    CoverSearch cs;
    if (n > MASK_WORDS * 64)
    {
        // too many cells for the mask: one point per sample (still exact, just not compressed)
        std::vector<Placement> out;
        for (auto &p : S)
            out.push_back({0, p.first, p.second});
        return out;
    }
    // every valid placement has its first (smallest) brick point on a cell of S, so anchoring that
    // point on each cell of S enumerates each placement exactly once
    for (int k = static_cast<int>(bricks.size()) - 1; k >= 0; --k)
        for (auto &anchor : S)
        {
            const int ox = anchor.first - bricks[k][0].first, oy = anchor.second - bricks[k][0].second;
            CoverSearch::Cand c;
            c.mask = {{0, 0, 0}};
            bool ok = true;
            for (auto &q : bricks[k])
            {
                const int x = ox + q.first, y = oy + q.second;
                if (x < -R || x > R || y < -R || y > R || rank[idx(x, y)] < 0)
                {
                    ok = false;
                    break;
                }
                const int b = rank[idx(x, y)];
                c.mask.w[b >> 6] |= 1ull << (b & 63);
            }
            if (!ok)
                continue;
            c.pl = {k, ox, oy};
            c.npts = static_cast<int>(bricks[k].size());
            cs.cands.push_back(c);
        }
    // drop placements whose cells are a subset of another placement's
    {
        std::vector<CoverSearch::Cand> kept;
        for (size_t i = 0; i < cs.cands.size(); ++i)
        {
            bool dominated = false;
            for (size_t j = 0; j < cs.cands.size() && !dominated; ++j)
                if (i != j && cs.cands[i].mask.subset_of(cs.cands[j].mask) && (!(cs.cands[i].mask == cs.cands[j].mask) || j < i))
                    dominated = true;
            if (!dominated)
                kept.push_back(cs.cands[i]);
        }
        cs.cands.swap(kept);
    }
    cs.covering.assign(n, {});
    for (size_t c = 0; c < cs.cands.size(); ++c)
    {
        cs.max_pts = std::max(cs.max_pts, cs.cands[c].mask.count());
        for (int b = 0; b < n; ++b)
            if (cs.cands[c].mask.test(b))
                cs.covering[b].push_back(static_cast<int>(c));
    }
    // greedy start
    Mask unc = {{0, 0, 0}};
    for (int b = 0; b < n; ++b)
        unc.w[b >> 6] |= 1ull << (b & 63);
    const Mask all = unc;
    while (!unc.empty())
    {
        int bestc = -1, gain = 0;
        for (size_t c = 0; c < cs.cands.size(); ++c)
        {
            const int g = cs.cands[c].mask.overlap(unc);
            if (g > gain)
            {
                gain = g;
                bestc = static_cast<int>(c);
            }
        }
        cs.best.push_back(bestc);
        unc = unc.andnot(cs.cands[bestc].mask);
    }
    cs.node_budget = node_budget;
    cs.dfs(all);
    std::vector<Placement> out;
    for (int c : cs.best)
        out.push_back(cs.cands[c].pl);
    return out;
}

// One axis of one bin: integer offsets floor(f(pos)) - pos for every position and point, the
// maximal ranges of positions over which all of them are constant, and the two longest ranges.
struct AxisClasses
{
    int n_classes = 0;
    int lo[2], hi[2];               // position ranges [lo, hi)
    std::vector<int> offs[2];       // per point
};

}  // namespace

struct Se2Plan
{
    // key
    uint32_t W = 0, H = 0, n_theta = 0;
    float ox = 0, oy = 0, ot = 0;
    std::vector<int32_t> fp;
    bool few_bins = false;  // built for launches of at most FEW_BINS bins: fewer brick tables pay off
    // how hard the cover search tried (0 quick, 1 thorough) and how often the plan has been used
    int effort = 0;
    uint32_t uses = 0;
    uint64_t serial = 0;  // distinguishes plans: the launcher keeps the assembled bin lists of the last (plan, bin set)
    // geometry
    bool usable = false;
    bool tma = false;  // smem_bytes includes the staging blocks of the bulk-tensor stores
    bool tall_ok = false;  // the build covers enough tiles for 32-row tiles (part of the plan's key)
    int R = 0, PW = 0, TH = 16, rows = 0, table_words = 0, n_tables = 0;  // halo, row pitch (words), tile height, table rows
    int max_abs_offset = 0;  // largest |sample offset| of any bin, point, column or row (exact, from the reference expression)
    size_t smem_bytes = 0;
    Library lib;
    int lib_id = 0;
    uint32_t tiles_x = 0, tiles_y = 0;
    uint8_t *d_xclass = nullptr, *d_yclass = nullptr;
    std::vector<uint8_t> h_xclass, h_yclass;
    std::vector<uint16_t> xsplit, ysplit;  // per bin: where the second range starts (mixed tiles)
    std::vector<uint8_t> split_left;      // per bin: bit 0 / 1 = class of the columns / rows before the split
    // per bin: up to 4 sample lists (class = cx*2 + cy)
    std::vector<uint32_t> lists;
    std::vector<uint32_t> list_off;  // [n_theta*4]
    std::vector<uint16_t> list_len;  // [n_theta*4]
    std::vector<uint32_t> list_id;   // [n_theta*4] number of the distinct offset set; NO_ID = class absent
    static constexpr uint32_t NO_ID = 0xFFFFFFFFu;
    // generic work list, sorted by bin
    std::vector<uint32_t> h_work;
    std::vector<uint32_t> work_off;  // [n_theta+1] entry index
    uint32_t *d_work = nullptr;
    // the part of the work list inside the last banded / dirty region (cached by region and bins)
    std::vector<uint32_t> region_work;
    uint32_t *d_region_work = nullptr;
    size_t region_cap = 0;
    uint64_t region_key[6] = {~0ull, 0, 0, 0, 0, 0};
    // dispatch order of the tiles of the last region (cached the same way)
    // [0]: the whole map (kept for the life of the plan), [1]: the last other region (a row band, a dirty rectangle)
    uint32_t *d_tile_order[2] = {nullptr, nullptr};
    size_t order_cap[2] = {0, 0};
    uint64_t order_key[2][5] = {{~0ull, 0, 0, 0, 0}, {~0ull, 0, 0, 0, 0}};
    ~Se2Plan()
    {
        cudaFree(d_tile_order[0]);
        cudaFree(d_tile_order[1]);
        cudaFree(d_region_work);
        cudaFree(d_xclass);
        cudaFree(d_yclass);
        cudaFree(d_work);
    }
};

void pp_se2_plan_free(pp_handle h)
{
    if (h->plan)
    {
        cudaSetDevice(h->device);
        cudaStreamSynchronize(h->stream);
        delete h->plan;
    }
    h->plan = nullptr;
}

namespace
{
// the reference expressions, one float operation at a time (no contraction on the host either)
inline int ref_x_index(float x, float px, float py, float c, float s)
{
    const float a = px * c;
    const float b = py * s;
    const float t = x + a;
    const float xn = t - b;
    return static_cast<int>(floorf(xn));
}
inline int ref_y_index(float y, float px, float py, float c, float s)
{
    const float a = px * s;
    const float b = py * c;
    const float t = y + a;
    const float yn = t + b;
    return static_cast<int>(floorf(yn));
}

AxisClasses classify_axis(bool is_x, uint32_t n, float sample_off, const std::vector<int32_t> &fp, float c, float s,
                          std::vector<int8_t> &scratch, int &max_abs_offset)
{
    const size_t P = fp.size() / 2;
    scratch.resize(P * n);
    for (size_t p = 0; p < P; ++p)
    {
        const float px = static_cast<float>(fp[2 * p]), py = static_cast<float>(fp[2 * p + 1]);
        int8_t *dst = scratch.data() + p * n;
        for (uint32_t i = 0; i < n; ++i)
        {
            const float pos = static_cast<float>(i) + sample_off;
            const int idx = is_x ? ref_x_index(pos, px, py, c, s) : ref_y_index(pos, px, py, c, s);
            int d = idx - static_cast<int>(i);
            max_abs_offset = std::max(max_abs_offset, std::abs(d));
            d = std::max(-127, std::min(127, d));  // storage only: a plan with such offsets is never used (R > MAX_R below)
            dst[i] = static_cast<int8_t>(d);
        }
    }
    // boundaries of constant ranges
    std::vector<uint32_t> starts{0};
    for (uint32_t i = 1; i < n; ++i)
    {
        bool same = true;
        for (size_t p = 0; p < P && same; ++p)
            same = scratch[p * n + i] == scratch[p * n + i - 1];
        if (!same)
            starts.push_back(i);
    }
    starts.push_back(n);
    // two longest ranges
    std::vector<std::pair<uint32_t, uint32_t>> ranges;
    for (size_t k = 0; k + 1 < starts.size(); ++k)
        ranges.emplace_back(starts[k + 1] - starts[k], starts[k]);
    std::sort(ranges.begin(), ranges.end(), [](auto &l, auto &r) { return l.first > r.first; });
    AxisClasses out;
    out.n_classes = static_cast<int>(std::min<size_t>(2, ranges.size()));
    for (int k = 0; k < out.n_classes; ++k)
    {
        out.lo[k] = static_cast<int>(ranges[k].second);
        out.hi[k] = static_cast<int>(ranges[k].second + ranges[k].first);
        out.offs[k].resize(P);
        for (size_t p = 0; p < P; ++p)
            out.offs[k][p] = scratch[p * n + out.lo[k]];
    }
    return out;
}

constexpr uint32_t FEW_BINS = 12;
// 32-row tiles only where a build has enough of them (about three and a half waves of CTA slots; measured on 512-,
// 1024- and 2048-row bands of a 4096-wide map: equal, 3 % and 5 % faster than 16 rows): the whole map, or a wide band
bool wants_tall_tiles(pp_handle h)
{
    const uint32_t rows = h->band_y1 > h->band_y0 ? std::min(h->band_y1, h->H) - std::min(h->band_y0, h->H) : h->H;
    return static_cast<uint64_t>((h->W + FT_W - 1) / FT_W) * ((rows + 31) / 32) >= 1000;
}
bool wants_few_bins(pp_handle h) { return h->few_bins_override >= 0 ? h->few_bins_override != 0 : h->bins.size() <= FEW_BINS; }
bool plan_matches(const Se2Plan *p, pp_handle h)
{
    return p && p->few_bins == wants_few_bins(h) && p->tall_ok == wants_tall_tiles(h) && p->W == h->W && p->H == h->H && p->n_theta == h->n_theta && p->ox == h->ox && p->oy == h->oy &&
           p->ot == h->ot && p->fp == h->fp;
}

pp_status build_plan(pp_handle h, int effort)
{
    Se2Plan *pl = new Se2Plan();
    static std::atomic<uint64_t> plan_serial{0};
    pl->serial = ++plan_serial;
    h->plan = pl;
    pl->effort = effort;
    // branch-and-bound nodes per offset set: a one-off build (pp_plan) should not wait for the last
    // few percent; a plan that keeps being reused is re-planned thoroughly (see pp_se2_launch_fast)
    const long node_budget = effort > 0 ? 400000 : 3000;
    pl->W = h->W;
    pl->H = h->H;
    pl->n_theta = h->n_theta;
    pl->ox = h->ox;
    pl->oy = h->oy;
    pl->ot = h->ot;
    pl->fp = h->fp;
    pl->usable = false;
    const uint32_t nt = h->n_theta;
    pl->tiles_x = (h->W + FT_W - 1) / FT_W;

    // 1. exact offsets and uniform ranges for every bin
    std::vector<AxisClasses> xc(nt), yc(nt);
    std::vector<int8_t> scratch;
    int R = 0;
    for (uint32_t t = 0; t < nt; ++t)
    {
        const float c = h->host_trig[t].x, s = h->host_trig[t].y;
        xc[t] = classify_axis(true, h->W, h->ox, h->fp, c, s, scratch, pl->max_abs_offset);
        yc[t] = classify_axis(false, h->H, h->oy, h->fp, c, s, scratch, pl->max_abs_offset);
        for (int k = 0; k < xc[t].n_classes; ++k)
            for (int d : xc[t].offs[k])
                R = std::max(R, std::abs(d));
        for (int k = 0; k < yc[t].n_classes; ++k)
            for (int d : yc[t].offs[k])
                R = std::max(R, std::abs(d));
    }
    if (R > MAX_R)
        return PP_OK;
    if (h->W % 4 != 0)
        return PP_OK;  // rows would not be 4-byte aligned: leave everything to the generic kernel
    // 2. geometry: the staged halo is a template parameter of the kernel, the row pitch follows from it
    const int halo = halo_for(R);
    pl->R = halo;
    pl->PW = 32 + 2 * halo;
    // Tile height: with 32 rows the halo rows are shared by twice the output rows (table build and staging per cell
    // -26 %) but a table is 47 % bigger; taken when at least 11 tables still fit twice per SM (measured at halo 9:
    // 0.497 against 0.518 ms with 16 rows; the box of config 2, halo 14, is better off with 16 rows and more tables).
    // A build that covers few tiles (a row band of an 8-GPU job) keeps 16 rows: twice as many tiles fill the last wave
    // of CTA slots better (measured at 8 GPUs: 0.089 against 0.096 ms per band).
    pl->tall_ok = wants_tall_tiles(h);
    pl->TH = pl->tall_ok && (SMEM_LIMIT - 1024) / (static_cast<size_t>(32 + 2 * halo) * pl->PW * 4) >= 11 ? 32 : 16;
    if (const char *e = getenv("PP_SE2_TILE_ROWS"))  // tuning knob for experiments
        pl->TH = atoi(e) == 32 ? 32 : 16;
    pl->tiles_y = (h->H + pl->TH - 1) / pl->TH;
    pl->rows = pl->TH + 2 * halo;
    pl->table_words = pl->rows * pl->PW;
    const size_t table_bytes = static_cast<size_t>(pl->table_words) * 4;
    // stores through the TMA unit need per-warp staging blocks behind the tables (PP_SE2_NO_TMA: knob for A/B runs;
    // plain whole-row stores measure within +-3 % of the bulk stores, which way depends on the box)
    pl->tma = getenv("PP_SE2_NO_TMA") == nullptr && h->W % 16 == 0;
    // a large halo leaves room for few tables: there the staging blocks cost more (as samples) than they save
    if ((SMEM_LIMIT - 1024 - STAGE_BYTES) / table_bytes < 12)
        pl->tma = false;
    const size_t stage_bytes = pl->tma ? STAGE_BYTES : 0;
    int slots = static_cast<int>(std::min<size_t>(MAX_TABLES, (SMEM_LIMIT - 1024 - stage_bytes) / table_bytes));
    // Every table costs a fixed amount per tile and saves samples per bin: with few bins per launch
    // (a heading-sharded rank of an 8-GPU job builds 9) a short library is the optimum.
    pl->few_bins = wants_few_bins(h);
    if (pl->few_bins)
    {
        const char *few = getenv("PP_SE2_FEW_TABLES");  // tuning knob for experiments
        slots = std::min(slots, few ? std::max(1, atoi(few)) : 5);
    }
    if (const char *cap = getenv("PP_SE2_MAX_TABLES"))  // tuning knob for experiments
        slots = std::max(1, std::min(slots, atoi(cap)));
    if (slots < 1)
        return PP_OK;

    // 3. distinct offset sets per (bin, class)
    auto make_set = [&](uint32_t t, int cx, int cy) {
        Offsets S;
        const size_t P = h->fp.size() / 2;
        for (size_t p = 0; p < P; ++p)
            S.emplace_back(xc[t].offs[cx][p], yc[t].offs[cy][p]);
        std::sort(S.begin(), S.end());
        S.erase(std::unique(S.begin(), S.end()), S.end());
        return S;
    };
    // 4. choose the library on a sample of bins
    int best_lib = 0;
    size_t best_total = ~static_cast<size_t>(0);
    const uint32_t step = std::max(1u, nt / 18);
    const std::vector<Library> &libs = fixed_libraries();
    for (int cand = 0; cand < static_cast<int>(libs.size()); ++cand)
    {
        const Library &lib = libs[cand];
        const int n = std::min<int>(slots, static_cast<int>(lib.size()));
        auto bricks = brick_points(lib, n);
        size_t total = 0;
        for (uint32_t t = 0; t < nt; t += step)
            total += cover_set(make_set(t, 0, 0), bricks, R, 2000).size();
        if (total < best_total)
        {
            best_total = total;
            best_lib = cand;
        }
    }
    {
        const Library &lib = libs[best_lib];
        pl->lib_id = best_lib;
        pl->n_tables = std::min<int>(slots, static_cast<int>(lib.size()));
        pl->lib.assign(lib.begin(), lib.begin() + pl->n_tables);
    }
    pl->smem_bytes = table_bytes * pl->n_tables + stage_bytes;
    auto bricks = brick_points(pl->lib, pl->n_tables);

    // 5. covers -> sample lists (byte offsets relative to the thread's base address)
    pl->list_off.assign(static_cast<size_t>(nt) * 4, 0);
    pl->list_len.assign(static_cast<size_t>(nt) * 4, 0);
    pl->list_id.assign(static_cast<size_t>(nt) * 4, Se2Plan::NO_ID);
    std::map<Offsets, std::vector<uint32_t>> memo;
    std::map<const std::vector<uint32_t> *, uint32_t> memo_id;  // map nodes do not move
    for (uint32_t t = 0; t < nt; ++t)
        for (int cx = 0; cx < xc[t].n_classes; ++cx)
            for (int cy = 0; cy < yc[t].n_classes; ++cy)
            {
                Offsets S = make_set(t, cx, cy);
                auto it = memo.find(S);
                if (it == memo.end())
                {
                    std::vector<uint32_t> words;
                    for (const Placement &p : cover_set(S, bricks, R, node_budget))
                    {
                        const int ax = halo + p.ox, ay = halo + p.oy;  // >= 0: the offset set lies inside the halo
                        const uint32_t word = static_cast<uint32_t>(p.brick * pl->table_words + ay * pl->PW + ax);
                        words.push_back(word * 4);
                    }
                    it = memo.emplace(S, words).first;
                }
                pl->list_id[t * 4 + cx * 2 + cy] = memo_id.emplace(&it->second, static_cast<uint32_t>(memo_id.size())).first->second;
                pl->list_off[t * 4 + cx * 2 + cy] = static_cast<uint32_t>(pl->lists.size());
                pl->list_len[t * 4 + cx * 2 + cy] = static_cast<uint16_t>(it->second.size());
                pl->lists.insert(pl->lists.end(), it->second.begin(), it->second.end());
            }

    // Every shared-memory address of the output phase is (thread base) + (list word) + (row of the unit) * pitch,
    // with host-made list words: checked here once for all launches (the kernel has no bounds tests in that loop).
    for (uint32_t w : pl->lists)
    {
        const uint32_t first = w / 4, last = first + static_cast<uint32_t>(pl->TH - 1) * pl->PW + 31;
        PP_REQUIRE(h, w % 4 == 0 && last < static_cast<uint32_t>(pl->n_tables) * pl->table_words, PP_ERR_STATE,
                   "internal: a sample list addresses word %u of %u table words", last, pl->n_tables * pl->table_words);
    }

    // 6. tile classes and the generic work list
    pl->h_xclass.assign(static_cast<size_t>(nt) * pl->tiles_x, 255);
    pl->h_yclass.assign(static_cast<size_t>(nt) * pl->tiles_y, 255);
    pl->work_off.assign(nt + 1, 0);
    pl->xsplit.assign(nt, 0);
    pl->ysplit.assign(nt, 0);
    pl->split_left.assign(nt, 0);
    // class of a tile [lo, hi) on one axis: inside one range -> that class; inside the union of the two ranges
    // when they touch -> 2 (mixed: evaluated once per range and blended); anything else -> 255 (generic kernel)
    auto tile_class = [](const AxisClasses &ac, int lo, int hi, int split, int left) -> uint8_t {
        for (int k = 0; k < ac.n_classes; ++k)
            if (lo >= ac.lo[k] && hi <= ac.hi[k])
                return static_cast<uint8_t>(k);
        if (split >= 0 && lo >= ac.lo[left] && hi <= ac.hi[1 - left])
            return 2;
        return 255;
    };
    auto split_of = [](const AxisClasses &ac, int &left) {
        if (ac.n_classes == 2 && ac.hi[0] == ac.lo[1])
        {
            left = 0;
            return ac.hi[0];
        }
        if (ac.n_classes == 2 && ac.hi[1] == ac.lo[0])
        {
            left = 1;
            return ac.hi[1];
        }
        left = 0;
        return -1;
    };
    const bool allow_mixed = getenv("PP_SE2_NO_MIXED") == nullptr;  // tuning knob for experiments
    for (uint32_t t = 0; t < nt; ++t)
    {
        int xleft = 0, yleft = 0;
        int xs = split_of(xc[t], xleft), ys = split_of(yc[t], yleft);
        if (!allow_mixed || xs > 0xFFFF)
            xs = -1;
        if (!allow_mixed || ys > 0xFFFF)
            ys = -1;
        pl->xsplit[t] = static_cast<uint16_t>(std::max(xs, 0));
        pl->ysplit[t] = static_cast<uint16_t>(std::max(ys, 0));
        pl->split_left[t] = static_cast<uint8_t>(xleft | yleft << 1);
        for (uint32_t tx = 0; tx < pl->tiles_x; ++tx)
        {
            const int lo = static_cast<int>(tx * FT_W), hi = static_cast<int>(std::min(h->W, (tx + 1) * FT_W));
            pl->h_xclass[static_cast<size_t>(t) * pl->tiles_x + tx] = tile_class(xc[t], lo, hi, xs, xleft);
        }
        for (uint32_t ty = 0; ty < pl->tiles_y; ++ty)
        {
            const int lo = static_cast<int>(ty * pl->TH), hi = static_cast<int>(std::min<uint32_t>(h->H, (ty + 1) * pl->TH));
            pl->h_yclass[static_cast<size_t>(t) * pl->tiles_y + ty] = tile_class(yc[t], lo, hi, ys, yleft);
        }
        pl->work_off[t] = static_cast<uint32_t>(pl->h_work.size() / 2);
        for (uint32_t ty = 0; ty < pl->tiles_y; ++ty)
            for (uint32_t tx = 0; tx < pl->tiles_x; ++tx)
                if (pl->h_xclass[static_cast<size_t>(t) * pl->tiles_x + tx] == 255 ||
                    pl->h_yclass[static_cast<size_t>(t) * pl->tiles_y + ty] == 255)
                    for (uint32_t sy = 0; sy < static_cast<uint32_t>(pl->TH / GEN_H); ++sy)  // in tiles of the generic kernel
                        for (uint32_t half = 0; half < FT_W / GEN_W; ++half)
                        {
                            const uint32_t gtx = tx * (FT_W / GEN_W) + half, gty = ty * (pl->TH / GEN_H) + sy;
                            if (gtx * GEN_W >= h->W || gty * GEN_H >= h->H)
                                continue;
                            pl->h_work.push_back(gtx | (gty << 16));
                            pl->h_work.push_back(t);
                        }
    }
    pl->work_off[nt] = static_cast<uint32_t>(pl->h_work.size() / 2);

    PP_CUDA(h, cudaMalloc(reinterpret_cast<void **>(&pl->d_xclass), pl->h_xclass.size()));
    PP_CUDA(h, cudaMalloc(reinterpret_cast<void **>(&pl->d_yclass), pl->h_yclass.size()));
    PP_CUDA(h, cudaMemcpyAsync(pl->d_xclass, pl->h_xclass.data(), pl->h_xclass.size(), cudaMemcpyHostToDevice, h->stream));
    PP_CUDA(h, cudaMemcpyAsync(pl->d_yclass, pl->h_yclass.data(), pl->h_yclass.size(), cudaMemcpyHostToDevice, h->stream));
    if (!pl->h_work.empty())
    {
        PP_CUDA(h, cudaMalloc(reinterpret_cast<void **>(&pl->d_work), pl->h_work.size() * sizeof(uint32_t)));
        PP_CUDA(h, cudaMemcpyAsync(pl->d_work, pl->h_work.data(), pl->h_work.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    }
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    PP_CUDA(h, set_fast_smem(pl->R, pl->TH, pl->smem_bytes));
    pl->usable = true;
    return PP_OK;
}
}  // namespace

namespace
{
// The volume as a (x, y, bin) uint8 tensor with 128 x 8 x 1 boxes. The encoder lives in the driver library;
// it is looked up once through the runtime, so the product does not link against libcuda.
using EncodeTiled = CUresult (*)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
bool encode_volume_map(CUtensorMap *map, uint8_t *vol, uint32_t W, uint32_t H, uint32_t n_theta)
{
    static EncodeTiled encode = [] {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            fn = nullptr;
        return reinterpret_cast<EncodeTiled>(fn);
    }();
    if (!encode || W % 16 != 0 || reinterpret_cast<uintptr_t>(vol) % 16 != 0)
        return false;
    const cuuint64_t dims[3] = {W, H, n_theta};
    const cuuint64_t strides[2] = {W, static_cast<cuuint64_t>(W) * H};
    const cuuint32_t box[3] = {FT_W, QPT, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, vol, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
}  // namespace

pp_status pp_se2_prepare_plan(pp_handle h)
{
    if (!plan_matches(h->plan, h))
    {
        pp_se2_plan_free(h);
        pp_status s = build_plan(h, 0);
        if (s != PP_OK)
            return s;
    }
    else if (h->plan->effort == 0 && ++h->plan->uses >= 2)
    {
        // third build with the same map size, footprint, bins and offsets: worth a thorough plan
        pp_se2_plan_free(h);
        pp_status s = build_plan(h, 1);
        if (s != PP_OK)
            return s;
    }
    // Both kernels stage reach + 2 cells around a tile; that this covers every sample follows from
    // |p| <= reach and a sample offset in [-1, 1]. The planner has evaluated every offset exactly, so a
    // configuration that broke the bound is refused here instead of being sampled wrongly.
    PP_REQUIRE(h, h->plan->max_abs_offset <= h->reach + 2, PP_ERR_RANGE,
               "a footprint sample lies %d cells from its pose, beyond the staged reach of %d + 2", h->plan->max_abs_offset, h->reach);
    return PP_OK;
}

pp_status pp_se2_launch_fast(pp_handle h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1)
{
    Se2Plan *pl = h->plan;
    if (!pl->usable)
        return PP_INTERNAL_DECLINED;

    const uint32_t tx_lo = x0 / FT_W, tx_hi = (x1 + FT_W - 1) / FT_W;
    const uint32_t ty_lo = y0 / pl->TH, ty_hi = (y1 + pl->TH - 1) / pl->TH;
    static thread_local FastArgs args;  // ~28 KB: keep it off the stack
    // Assembling the per-bin part of `args` (sample lists, chains) costs tens of microseconds of host time; a rebuild
    // with the same plan and the same bins, all in one launch, reuses what is already in `args`.
    static thread_local uint64_t listed_serial = 0;
    static thread_local std::vector<uint32_t> listed_bins;
    args.cost = h->d_cost;
    args.vol = h->d_vol;
    {
        uint8_t *targets[8];
        const int n = pp_volume_targets(h, targets);
        args.n_peers = n - 1;
        for (int p = 1; p < n; ++p)
            args.peer_delta[p - 1] = static_cast<long long>(reinterpret_cast<uintptr_t>(targets[p]) - reinterpret_cast<uintptr_t>(targets[0]));
    }
    // bulk-tensor stores unless peers' volumes are written too (those go out as plain stores over NVLink)
    {
        static thread_local const void *map_vol = nullptr;  // the tensor map in `args` describes this volume
        static thread_local uint32_t map_dims[3] = {0, 0, 0};
        const bool want = pl->tma && args.n_peers == 0;
        if (want && (map_vol != h->d_vol || map_dims[0] != h->W || map_dims[1] != h->H || map_dims[2] != h->n_theta))
        {
            const bool ok = encode_volume_map(&args.vol_map, h->d_vol, h->W, h->H, h->n_theta);
            map_vol = ok ? h->d_vol : nullptr;
            map_dims[0] = h->W, map_dims[1] = h->H, map_dims[2] = h->n_theta;
        }
        args.use_tma = want && map_vol == h->d_vol;
    }
    if (getenv("PP_SE2_DEBUG"))
        fprintf(stderr, "[pp] se2 fast: halo %d tables %d smem %zu tma %d grid %ux%u\n", pl->R, pl->n_tables, pl->smem_bytes, args.use_tma,
                (x1 + FT_W - 1) / FT_W - x0 / FT_W, ty_hi - ty_lo);
    args.xclass = pl->d_xclass;
    args.yclass = pl->d_yclass;
    args.W = h->W;
    args.H = h->H;
    args.tiles_x = pl->tiles_x;
    args.tiles_y = pl->tiles_y;
    args.n_tables = pl->n_tables;
    {
        const std::vector<int> lvl = library_levels(pl->lib);
        args.rec_sync = 0;
        for (int k = 1; k < pl->n_tables; ++k)
        {
            const Recipe &rc = pl->lib[k];
            args.rec_a[k] = rc.a * pl->table_words;
            args.rec_b[k] = rc.b * pl->table_words;
            args.rec_d[k] = rc.dv * pl->PW + rc.du;
            if (lvl[k] != lvl[k - 1] && lvl[k] > 1)  // level 1 reads only table 0, which the staging barrier covers
                args.rec_sync |= 1u << k;
        }
    }
    const uint32_t n_tiles = (tx_hi - tx_lo) * (ty_hi - ty_lo);  // one CTA per tile
    const int oc = (tx_lo == 0 && ty_lo == 0 && tx_hi >= pl->tiles_x && ty_hi >= pl->tiles_y) ? 0 : 1;  // which cached order
    // Dispatch order: CTAs start in index order, so the tiles with the most work go first (a tile that a range
    // boundary crosses evaluates such a bin once per class pair and stores through masks -- with the boundaries
    // near x = 0 that is the left-hand tile column, twice as long as the others) and none of them ends the launch.
    {
        const uint64_t key[5] = {tx_lo, tx_hi, ty_lo, ty_hi, 1};  // weighed over all bins: one order per region
        if (std::memcmp(key, pl->order_key[oc], sizeof(key)) != 0)
        {
            std::vector<uint32_t> weight(n_tiles, 0), order(n_tiles);
            for (uint32_t t = 0; t < pl->n_theta; ++t)
                for (uint32_t ty = ty_lo; ty < ty_hi; ++ty)
                {
                    const uint32_t cy = pl->h_yclass[static_cast<size_t>(t) * pl->tiles_y + ty];
                    for (uint32_t tx = tx_lo; tx < tx_hi; ++tx)
                    {
                        const uint32_t cx = pl->h_xclass[static_cast<size_t>(t) * pl->tiles_x + tx];
                        if (cx > 2 || cy > 2)
                            continue;  // generic work list
                        const uint32_t pairs = (cx == 2 ? 2u : 1u) * (cy == 2 ? 2u : 1u);
                        weight[(ty - ty_lo) * (tx_hi - tx_lo) + (tx - tx_lo)] += (cx == 2 || cy == 2) ? 2u * pairs : 1u;
                    }
                }
            // tiles whose halo leaves the map stage through the bounds-tested path: ahead of the interior ones as well
            for (uint32_t ty = ty_lo; ty < ty_hi; ++ty)
                for (uint32_t tx = tx_lo; tx < tx_hi; ++tx)
                {
                    const long long cx0 = static_cast<long long>(tx) * FT_W - pl->R, cy0 = static_cast<long long>(ty) * pl->TH - pl->R;
                    const bool inside = cx0 >= 0 && cy0 >= 0 && cx0 + FT_W + 2 * pl->R <= static_cast<long long>(pl->W) &&
                                        cy0 + pl->TH + 2 * pl->R <= static_cast<long long>(pl->H);
                    uint32_t &wv = weight[(ty - ty_lo) * (tx_hi - tx_lo) + (tx - tx_lo)];
                    wv = 2u * wv + (inside ? 0u : 1u);
                }
            for (uint32_t i = 0; i < n_tiles; ++i)
                order[i] = i;
            std::stable_sort(order.begin(), order.end(), [&](uint32_t p, uint32_t q) { return weight[p] > weight[q]; });
            for (uint32_t i = 0; i < n_tiles; ++i)
                order[i] = (tx_lo + order[i] % (tx_hi - tx_lo)) | (ty_lo + order[i] / (tx_hi - tx_lo)) << 16;
            PP_CUDA(h, pp_reserve(pl->d_tile_order[oc], pl->order_cap[oc], order.size()));
            // (pageable source: the call returns once the data has been staged, `order` may go out of scope)
            PP_CUDA(h, cudaMemcpyAsync(pl->d_tile_order[oc], order.data(), order.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
            std::memcpy(pl->order_key[oc], key, sizeof(key));
            if (std::getenv("PP_SE2_DEBUG"))
            {
                std::map<uint32_t, uint32_t> hist;
                for (uint32_t wv : weight)
                    hist[wv]++;
                for (auto &kv : hist)
                    std::fprintf(stderr, "[pp_se2] tile weight %u: %u tiles\n", kv.first, kv.second);
            }
        }
    }
    args.tile_order = pl->d_tile_order[oc];
    const dim3 grid(n_tiles, 1, 1);
    double sample_sum = 0.0;
    uint64_t fast_pairs = 0, generic_pairs = 0;

    const std::vector<uint32_t> &bins = h->bins;  // ascending
    size_t bi = 0;
    while (bi < bins.size())
    {
        uint32_t n = 0;
        const bool reuse = bi == 0 && listed_serial == pl->serial && listed_bins == bins;
        if (reuse)
            n = static_cast<uint32_t>(bins.size());
        else
        {
            listed_serial = 0;
            // as many bins as fit in the parameter block
            uint32_t words = 0;
            std::map<uint32_t, uint32_t> stored;           // offset-set id -> where its list sits in args.lists
            std::map<uint32_t, uint32_t> last_with;        // offset-set id -> last (bin, class) link that has it
            while (bi + n < bins.size() && n < MAX_T_PER_LAUNCH)
            {
                const uint32_t t = bins[bi + n];
                uint32_t need = 0;
                for (int c = 0; c < 4; ++c)
                    if (!stored.count(pl->list_id[t * 4 + c]))
                        need += (pl->list_len[t * 4 + c] + 3u) & ~3u;
                if (words + need > MAX_LIST_WORDS)
                    break;
                args.xsplit[n] = pl->xsplit[t];
                args.ysplit[n] = pl->ysplit[t];
                args.split_left[n] = pl->split_left[t];
                for (int c = 0; c < 4; ++c)
                {
                    const uint32_t len = pl->list_len[t * 4 + c], id = pl->list_id[t * 4 + c];
                    args.list_len[n][c] = static_cast<uint16_t>(len);
                    args.chain_prev[n][c] = args.chain_next[n][c] = static_cast<uint16_t>(NO_LINK);
                    if (id == Se2Plan::NO_ID)
                    {
                        args.list_off[n][c] = 0;
                        continue;
                    }
                    auto st = stored.find(id);
                    if (st == stored.end())
                    {
                        st = stored.emplace(id, words).first;
                        std::memcpy(args.lists + words, pl->lists.data() + pl->list_off[t * 4 + c], len * sizeof(uint32_t));
                        words += (len + 3u) & ~3u;  // keep every list 16-byte aligned
                    }
                    args.list_off[n][c] = static_cast<uint16_t>(st->second);
                    const uint32_t me = n | static_cast<uint32_t>(c) << 8;
                    auto lw = last_with.find(id);
                    if (lw != last_with.end())
                    {
                        args.chain_prev[n][c] = static_cast<uint16_t>(lw->second);
                        args.chain_next[lw->second & 0xFFu][lw->second >> 8] = static_cast<uint16_t>(me);
                        lw->second = me;
                    }
                    else
                        last_with.emplace(id, me);
                }
                args.bin[n] = static_cast<uint16_t>(t);
                ++n;
            }
            if (bi == 0 && n == bins.size())
            {
                listed_serial = pl->serial;
                listed_bins = bins;
            }
        }
        PP_REQUIRE(h, n > 0, PP_ERR_RANGE, "a single heading bin needs more than %d sample words", MAX_LIST_WORDS);
        args.n_t = n;
        // bookkeeping + the exact path for the (tile, bin) pairs outside uniform ranges: queued first, on the
        // side stream, so that its CTAs share the SMs with the tiled kernel instead of running after it
        bool forked = false;
        uint32_t n_work = 0;
        for (uint32_t k = 0; k < n; ++k)
        {
            const uint32_t t = bins[bi + k];
            const uint32_t nw = pl->work_off[t + 1] - pl->work_off[t];
            n_work += nw;
            generic_pairs += nw;
            fast_pairs += static_cast<uint64_t>(pl->tiles_x) * pl->tiles_y - nw / ((FT_W / GEN_W) * (pl->TH / GEN_H));
            if (args.chain_prev[k][0] == NO_LINK)  // bins that share an earlier bin's evaluation load nothing
                sample_sum += pl->list_len[t * 4];
        }
        if (n_work > 0)
        {
            const bool whole_map = tx_lo == 0 && ty_lo == 0 && tx_hi >= pl->tiles_x && ty_hi >= pl->tiles_y;
            const uint32_t t_lo = bins[bi], t_hi = bins[bi + n - 1];
            if (whole_map && t_hi - t_lo + 1 == n)  // consecutive bins: a slice of the bin-sorted work list
            {
                pp_status s = pp_se2_generic_worklist(h, pl->d_work + 2 * static_cast<size_t>(pl->work_off[t_lo]), n_work);
                if (s != PP_OK)
                    return s;
                forked = true;
            }
            else
            {
                // a row band, a dirty region or a scattered set of bins: only the listed (tile, bin) pairs inside
                uint64_t key[6] = {tx_lo, tx_hi, ty_lo, ty_hi, 0, n};
                for (uint32_t k = 0; k < n; ++k)
                    key[4] = key[4] * 1000003ull + bins[bi + k] + 1;
                if (std::memcmp(key, pl->region_key, sizeof(key)) != 0)
                {
                    pl->region_work.clear();
                    for (uint32_t k = 0; k < n; ++k)
                        for (uint32_t e = pl->work_off[bins[bi + k]]; e < pl->work_off[bins[bi + k] + 1]; ++e)
                        {
                            const uint32_t tile = pl->h_work[2 * e], wx = tile & 0xFFFFu, wy = tile >> 16;
                            if (wx / (FT_W / GEN_W) >= tx_lo && wx / (FT_W / GEN_W) < tx_hi && wy / (pl->TH / GEN_H) >= ty_lo && wy / (pl->TH / GEN_H) < ty_hi)
                            {
                                pl->region_work.push_back(tile);
                                pl->region_work.push_back(pl->h_work[2 * e + 1]);
                            }
                        }
                    if (!pl->region_work.empty())
                    {
                        PP_CUDA(h, pp_reserve(pl->d_region_work, pl->region_cap, pl->region_work.size()));
                        PP_CUDA(h, cudaMemcpyAsync(pl->d_region_work, pl->region_work.data(), pl->region_work.size() * sizeof(uint32_t),
                                                   cudaMemcpyHostToDevice, h->stream));
                    }
                    std::memcpy(pl->region_key, key, sizeof(key));
                }
                if (!pl->region_work.empty())
                {
                    pp_status s = pp_se2_generic_worklist(h, pl->d_region_work, static_cast<uint32_t>(pl->region_work.size() / 2));
                    if (s != PP_OK)
                        return s;
                    forked = true;
                }
            }
        }
#ifdef PP_SE2_TIMELINE
        static unsigned long long *d_tl = nullptr;
        static size_t tl_cap = 0;
        const char *tl_path = std::getenv("PP_SE2_TIMELINE_FILE");
        args.timeline = nullptr;
        if (tl_path)
        {
            if (tl_cap < grid.x)
            {
                cudaFree(d_tl);
                cudaMalloc(&d_tl, 5ull * grid.x * sizeof(unsigned long long));
                tl_cap = grid.x;
            }
            cudaMemsetAsync(d_tl, 0, 5ull * grid.x * sizeof(unsigned long long), h->stream);
            args.timeline = d_tl;
        }
#endif
        launch_fast(pl->R, pl->TH, grid, pl->smem_bytes, h->stream, args);
        h->launches++;
        PP_CUDA(h, cudaGetLastError());
#ifdef PP_SE2_TIMELINE
        if (tl_path)
        {
            std::vector<unsigned long long> tl(5ull * grid.x);
            cudaStreamSynchronize(h->stream);
            cudaMemcpy(tl.data(), d_tl, tl.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
            if (FILE *f = std::fopen(tl_path, "w"))  // the last launch wins
            {
                for (uint32_t b = 0; b < grid.x; ++b)
                    std::fprintf(f, "%u %llu %llu %llu %llu %llu\n", b, tl[5 * b], tl[5 * b + 1], tl[5 * b + 2], tl[5 * b + 3], tl[5 * b + 4]);
                std::fclose(f);
            }
        }
#endif
        if (forked)
        {
            pp_status s = pp_se2_join_aux(h);
            if (s != PP_OK)
                return s;
        }
        bi += n;
    }
    h->stat_fast_tiles = fast_pairs;
    h->stat_samples_per_cell = sample_sum / std::max<size_t>(1, h->bins.size());
    return PP_OK;
}
