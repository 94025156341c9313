// K3: TEB obstacle association (see include/pp_b200.h).
//
// K3a  obstacle_extract: teb_follower.cpp:140-169. The reference scans x-outer / y-inner and
//      appends, so the output order is column-major. Each thread walks one column top to bottom
//      (adjacent lanes read adjacent bytes of a row), first counting, then -- after an exclusive
//      scan over the column counts -- emitting at its column's base offset, which reproduces the
//      reference order exactly. World coordinates and the two filters are evaluated in double with
//      separately rounded operations (no FMA), like the reference build.
// K3b  radius_topk: teb_optimizer.cpp:294-311 over nanoflann radiusSearch semantics. One CTA per
//      query pose: brute-force squared distances (d2 = (0 + dx*dx) + dy*dy, nanoflann.hpp:627-638),
//      strict d2 < radius_arg (:404-408), hits compacted into shared memory with warp ballots,
//      then the K smallest are extracted in ascending (d2, index) order by repeated block-wide
//      arg-min. At N ~ 1e4 the whole cloud is 160 KB, so a spatial index would cost more to build
//      than the scan it saves; the kernel is launch-latency bound.
#include "pp_internal.h"

#include <cstdlib>
#include <cstring>

#include <cmath>
#include <algorithm>

namespace
{
struct ExtractArgs
{
    const uint8_t *cost;
    uint32_t W, H;
    double res, origin_x, origin_y, robot_x, robot_y, dir_x, dir_y, max_distance, behind_range;
    // candidates whose distance is within rounding of behind_range (see candidate()): collected in the first
    // pass (count word + (i, j) pairs), decided by the host with its own hypot, handed back as overrides
    uint32_t *uncertain;         // [0] count, then (i, j) pairs; null in the passes that use overrides
    const uint32_t *overrides;   // (i, j, keep) triples
    uint32_t n_overrides;
};
constexpr uint32_t MAX_UNCERTAIN = 256;

__device__ __forceinline__ bool candidate(const ExtractArgs &a, uint32_t i, uint32_t j, double &wx, double &wy)
{
    if (a.cost[static_cast<size_t>(j) * a.W + i] < PP_COST_INSCRIBED)
        return false;
    // costmap_2d::Costmap2D::mapToWorld: origin + (m + 0.5) * resolution
    wx = __dadd_rn(a.origin_x, __dmul_rn(static_cast<double>(i) + 0.5, a.res));
    wy = __dadd_rn(a.origin_y, __dmul_rn(static_cast<double>(j) + 0.5, a.res));
    const double dx = __dsub_rn(wx, a.robot_x), dy = __dsub_rn(wy, a.robot_y);
    if (fabs(dx) > a.max_distance || fabs(dy) > a.max_distance)
        return false;
    if (!(__dadd_rn(__dmul_rn(dx, a.dir_x), __dmul_rn(dy, a.dir_y)) < 0.0))
        return true;  // not behind the robot: the distance does not matter (teb_follower.cpp:161)
    // The reference compares glibc's hypot(dx, dy) with obstacle_behind_ignore_range. Any hypot with an error
    // below one ulp gives the same answer as the exact comparison of the squares unless the distance is within a
    // few ulps of the range: decide those on the host, with the host's own hypot.
    const double s2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), r2 = __dmul_rn(a.behind_range, a.behind_range);
    if (a.behind_range >= 0.0 && fabs(__dsub_rn(s2, r2)) <= __dmul_rn(r2, 0x1p-46))
    {
        for (uint32_t k = 0; k < a.n_overrides; ++k)
            if (a.overrides[3 * k] == i && a.overrides[3 * k + 1] == j)
                return a.overrides[3 * k + 2] != 0;
        if (a.uncertain)
        {
            const uint32_t at = atomicAdd(a.uncertain, 1u);
            if (at < MAX_UNCERTAIN)
            {
                a.uncertain[1 + 2 * at] = i;
                a.uncertain[2 + 2 * at] = j;
            }
        }
        return false;  // provisional: the pass is repeated with the host's decisions
    }
    return !(a.behind_range < 0.0 || s2 > r2);
}

__global__ void extract_count_kernel(ExtractArgs a, uint32_t *col_counts)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.W)
        return;
    uint32_t n = 0;
    double wx, wy;
    for (uint32_t j = 0; j < a.H; ++j)
        n += candidate(a, i, j, wx, wy) ? 1u : 0u;
    col_counts[i] = n;
}

// exclusive scan of col_counts[0..W) in place; total written to col_counts[W]
__global__ void __launch_bounds__(1024) extract_scan_kernel(uint32_t *col_counts, uint32_t W)
{
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0)
        carry = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t base = 0; base < W; base += 1024)
    {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < W ? col_counts[i] : 0;
        uint32_t incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1)
        {
            const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= static_cast<uint32_t>(d))
                incl += o;
        }
        if (lane == 31)
            warp_sums[warp] = incl;
        __syncthreads();
        if (warp == 0)
        {
            uint32_t w = warp_sums[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1)
            {
                const uint32_t o = __shfl_up_sync(0xffffffffu, w, d);
                if (lane >= static_cast<uint32_t>(d))
                    w += o;
            }
            warp_sums[lane] = w;
        }
        __syncthreads();
        const uint32_t before = carry + (warp ? warp_sums[warp - 1] : 0) + incl - v;
        if (i < W)
            col_counts[i] = before;
        __syncthreads();
        if (threadIdx.x == 1023)
            carry = before + v;
        __syncthreads();
    }
    if (threadIdx.x == 0)
        col_counts[W] = carry;
}

__global__ void extract_emit_kernel(ExtractArgs a, const uint32_t *col_offsets, double *pts, uint32_t cap)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.W)
        return;
    uint32_t at = col_offsets[i];
    double wx, wy;
    for (uint32_t j = 0; j < a.H; ++j)
        if (candidate(a, i, j, wx, wy))
        {
            if (at < cap)
            {
                pts[2 * static_cast<size_t>(at)] = wx;
                pts[2 * static_cast<size_t>(at) + 1] = wy;
            }
            ++at;
        }
}

constexpr int RK_THREADS = 256;
constexpr int RK_CAP = 3072;  // hits buffered per query before a reduction to the K best
constexpr int RK_MAXK = 64;
constexpr int RK_BATCH = 4;   // points per thread between two block barriers of the scan

struct Key
{
    unsigned long long d2;  // bit pattern of a non-negative double: integer order == numeric order
    uint32_t idx;
};
__device__ __forceinline__ bool key_less(const Key &a, const Key &b) { return a.d2 < b.d2 || (a.d2 == b.d2 && a.idx < b.idx); }

// col_prefix (optional): the points are those of pp_obstacles_extract, ordered by costmap column, and
// col_prefix[c] is the index of the first point of column c (col_prefix[n_cols] = N). A query then scans only
// the columns whose centres can lie within sqrt(radius_arg) of it in x (one column of slack either side, so
// the exact d2 test below still decides every hit: results are unchanged).
__global__ void __launch_bounds__(RK_THREADS) radius_topk_kernel(const double *__restrict__ pts, uint32_t N,
                                                                 const double *__restrict__ queries, double radius_arg,
                                                                 uint32_t K, uint32_t *__restrict__ idx_out,
                                                                 double *__restrict__ d2_out, uint32_t *__restrict__ cnt_out,
                                                                 const uint32_t *__restrict__ col_prefix, uint32_t n_cols,
                                                                 double col_origin, double col_res)
{
    __shared__ unsigned long long s_d2[RK_CAP + RK_MAXK];
    __shared__ uint32_t s_idx[RK_CAP + RK_MAXK];
    __shared__ uint32_t s_count;
    __shared__ unsigned long long s_best_d2[RK_MAXK];
    __shared__ uint32_t s_best_idx[RK_MAXK];
    __shared__ uint32_t s_nbest;

    const uint32_t q = blockIdx.x;
    const double qx = queries[2 * q], qy = queries[2 * q + 1];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0)
    {
        s_count = 0;
        s_nbest = 0;
    }
    __syncthreads();
    uint32_t n_lo = 0, n_hi = N;
    if (col_prefix)
    {
        const double r = sqrt(fmax(radius_arg, 0.0));
        // column c has its centre at col_origin + (c + 0.5) * col_res
        const double c_lo = floor((qx - r - col_origin) / col_res - 0.5) - 1.0, c_hi = ceil((qx + r - col_origin) / col_res - 0.5) + 1.0;
        const uint32_t lo = c_lo <= 0.0 ? 0u : (c_lo >= static_cast<double>(n_cols) ? n_cols : static_cast<uint32_t>(c_lo));
        const uint32_t hi = c_hi < 0.0 ? 0u : (c_hi + 1.0 >= static_cast<double>(n_cols) ? n_cols : static_cast<uint32_t>(c_hi) + 1u);
        n_lo = col_prefix[lo];
        n_hi = hi > lo ? col_prefix[hi] : n_lo;
    }

    const unsigned long long EMPTY = 0xFFFFFFFFFFFFFFFFull;
    // reduce the buffered hits (plus the incumbents) to the K smallest, ascending
    auto select_best = [&]() {
        __syncthreads();
        uint32_t n = s_count;
        // append incumbents so they compete with the new hits
        for (uint32_t i = tid; i < s_nbest; i += RK_THREADS)
        {
            s_d2[n + i] = s_best_d2[i];
            s_idx[n + i] = s_best_idx[i];
        }
        __syncthreads();
        n += s_nbest;
        const uint32_t want = min(K, n);
        // K rounds of arg-min by warp 0 alone: shuffles instead of block barriers (a few hundred candidates)
        if (warp == 0)
        {
            for (uint32_t r = 0; r < want; ++r)
            {
                Key best = {EMPTY, 0xFFFFFFFFu};
                uint32_t slot = 0xFFFFFFFFu;
                for (uint32_t i = lane; i < n; i += 32)
                {
                    const Key k = {s_d2[i], s_idx[i]};
                    if (key_less(k, best))
                    {
                        best = k;
                        slot = i;
                    }
                }
#pragma unroll
                for (int d = 16; d > 0; d >>= 1)
                {
                    Key o;
                    o.d2 = __shfl_xor_sync(0xffffffffu, best.d2, d);
                    o.idx = __shfl_xor_sync(0xffffffffu, best.idx, d);
                    const uint32_t oslot = __shfl_xor_sync(0xffffffffu, slot, d);
                    if (key_less(o, best))
                    {
                        best = o;
                        slot = oslot;
                    }
                }
                if (lane == 0)
                {
                    s_best_d2[r] = best.d2;
                    s_best_idx[r] = best.idx;
                    s_d2[slot] = EMPTY;  // taken
                    s_idx[slot] = 0xFFFFFFFFu;
                }
                __syncwarp();
            }
        }
        if (tid == 0)
        {
            s_nbest = want;
            s_count = 0;
        }
        __syncthreads();
    };

    // `reduce` is uniform by construction: thread 0 publishes it through the barrier that ends an iteration. It reads
    // the counter before that barrier, so hits of the same iteration that other warps have not added yet may be
    // missing: the threshold leaves room for them (one batch) and for the next iteration (another batch).
    static_assert(RK_CAP >= 3 * RK_BATCH * RK_THREADS, "buffer holds the carried hits and two batches");
    int reduce = 0;
    for (uint32_t base = n_lo; base < n_hi; base += RK_BATCH * RK_THREADS)
    {
        if (reduce)
            select_best();
        // RK_BATCH points per thread between two barriers, all loads issued first
        double px[RK_BATCH], py[RK_BATCH];
#pragma unroll
        for (int b = 0; b < RK_BATCH; ++b)
        {
            const uint32_t i = base + b * RK_THREADS + tid;
            px[b] = i < n_hi ? pts[2 * static_cast<size_t>(i)] : 0.0;
            py[b] = i < n_hi ? pts[2 * static_cast<size_t>(i) + 1] : 0.0;
        }
#pragma unroll
        for (int b = 0; b < RK_BATCH; ++b)
        {
            const uint32_t i = base + b * RK_THREADS + tid;
            // L2_Simple_Adaptor::evalMetric (nanoflann.hpp:627-638): (0 + dx*dx) + dy*dy in double, no contraction
            double d2 = 0.0;
            const double dx = __dsub_rn(qx, px[b]);
            d2 = __dadd_rn(d2, __dmul_rn(dx, dx));
            const double dy = __dsub_rn(qy, py[b]);
            d2 = __dadd_rn(d2, __dmul_rn(dy, dy));
            const bool hit = i < n_hi && d2 < radius_arg;
            const unsigned ballot = __ballot_sync(0xffffffffu, hit);
            uint32_t warp_base = 0;
            if (lane == 0 && ballot)
                warp_base = atomicAdd(&s_count, __popc(ballot));
            warp_base = __shfl_sync(0xffffffffu, warp_base, 0);
            if (hit)
            {
                const uint32_t slot = warp_base + __popc(ballot & ((1u << lane) - 1u));
                s_d2[slot] = static_cast<unsigned long long>(__double_as_longlong(d2));
                s_idx[slot] = i;
            }
        }
        reduce = __syncthreads_or(tid == 0 && s_count + 2 * RK_BATCH * RK_THREADS > RK_CAP);
    }
    select_best();
    const uint32_t n = s_nbest;
    if (tid == 0)
        cnt_out[q] = n;
    for (uint32_t r = tid; r < K; r += RK_THREADS)
    {
        idx_out[static_cast<size_t>(q) * K + r] = r < n ? s_best_idx[r] : 0u;
        d2_out[static_cast<size_t>(q) * K + r] = r < n ? __longlong_as_double(static_cast<long long>(s_best_d2[r])) : 0.0;
    }
}
}  // namespace

extern "C" {

pp_status pp_obstacles_extract(pp_handle h, double resolution, double origin_x, double origin_y, double robot_x,
                               double robot_y, double robot_theta, double max_follow_range, double behind_ignore_range,
                               double *pts_out, uint32_t max_points, uint32_t *n_out)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_cost, PP_ERR_STATE, "pp_obstacles_extract: upload a costmap first");
    PP_REQUIRE(h, n_out, PP_ERR_INVALID, "pp_obstacles_extract: n_out is null");
    PP_CUDA(h, cudaSetDevice(h->device));
    ExtractArgs a;
    a.cost = h->d_cost;
    a.W = h->W;
    a.H = h->H;
    a.res = resolution;
    a.origin_x = origin_x;
    a.origin_y = origin_y;
    a.robot_x = robot_x;
    a.robot_y = robot_y;
    a.dir_x = cos(robot_theta);  // Point2E::direction(), teb_path_follower/commons.h
    a.dir_y = sin(robot_theta);
    a.max_distance = max_follow_range * 1.2;  // teb_follower.cpp:146
    a.behind_range = behind_ignore_range;
    const size_t unc_at = static_cast<size_t>(h->W) + 2, ovr_at = unc_at + 1 + 2 * MAX_UNCERTAIN;
    PP_CUDA(h, pp_reserve(h->d_col_counts, h->col_cap, ovr_at + 3 * MAX_UNCERTAIN));
    const uint32_t blocks = (h->W + 127) / 128;
    a.uncertain = h->d_col_counts + unc_at;
    a.overrides = nullptr;
    a.n_overrides = 0;
    PP_CUDA(h, cudaMemsetAsync(a.uncertain, 0, sizeof(uint32_t), h->stream));
    extract_count_kernel<<<blocks, 128, 0, h->stream>>>(a, h->d_col_counts);
    extract_scan_kernel<<<1, 1024, 0, h->stream>>>(h->d_col_counts, h->W);
    h->launches += 2;
    PP_CUDA(h, cudaGetLastError());
    PP_CUDA(h, cudaMemcpyAsync(h->h_counter, h->d_col_counts + h->W, sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaMemcpyAsync(h->h_counter + 1, a.uncertain, sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    if (const uint32_t n_unc = h->h_counter[1])
    {
        // candidates within rounding of the range: the reference's own expression, with the host libm's hypot
        PP_REQUIRE(h, n_unc <= MAX_UNCERTAIN, PP_ERR_RANGE, "pp_obstacles_extract: %u candidates lie within rounding of behind_ignore_range", n_unc);
        std::vector<uint32_t> list(2 * n_unc), ovr(3 * n_unc);
        PP_CUDA(h, cudaMemcpyAsync(list.data(), a.uncertain + 1, list.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
        PP_CUDA(h, cudaStreamSynchronize(h->stream));
        for (uint32_t k = 0; k < n_unc; ++k)
        {
            const uint32_t i = list[2 * k], j = list[2 * k + 1];
            const double wx = origin_x + (i + 0.5) * resolution, wy = origin_y + (j + 0.5) * resolution;  // Costmap2D::mapToWorld
            const double dx = wx - robot_x, dy = wy - robot_y;
            const double distance = hypot(dx, dy);  // teb_follower.cpp:158; behind the robot is already known
            ovr[3 * k] = i;
            ovr[3 * k + 1] = j;
            ovr[3 * k + 2] = distance > behind_ignore_range ? 0u : 1u;
        }
        PP_CUDA(h, cudaMemcpyAsync(h->d_col_counts + ovr_at, ovr.data(), ovr.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
        a.uncertain = nullptr;
        a.overrides = h->d_col_counts + ovr_at;
        a.n_overrides = n_unc;
        extract_count_kernel<<<blocks, 128, 0, h->stream>>>(a, h->d_col_counts);
        extract_scan_kernel<<<1, 1024, 0, h->stream>>>(h->d_col_counts, h->W);
        h->launches += 2;
        PP_CUDA(h, cudaGetLastError());
        PP_CUDA(h, cudaMemcpyAsync(h->h_counter, h->d_col_counts + h->W, sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
        PP_CUDA(h, cudaStreamSynchronize(h->stream));  // (ovr is read by the copy above: it completes here)
        if (getenv("PP_TEB_DEBUG"))
            fprintf(stderr, "[pp] extract: %u candidates within rounding of the range decided on the host\n", n_unc);
    }
    else
        a.uncertain = nullptr;
    const uint32_t total = h->h_counter[0];
    PP_CUDA(h, pp_reserve(h->d_pts, h->pts_cap, 2 * static_cast<size_t>(std::max(total, 1u))));
    extract_emit_kernel<<<blocks, 128, 0, h->stream>>>(a, h->d_col_counts, h->d_pts, total);
    h->launches++;
    PP_CUDA(h, cudaGetLastError());
    h->n_pts = total;
    // the points are ordered by column: pp_radius_topk prunes with the column prefix
    h->pts_cols = h->W;
    h->pts_col_origin = origin_x;
    h->pts_col_res = resolution;
    *n_out = total;
    if (pts_out && max_points)
        PP_CUDA(h, cudaMemcpyAsync(pts_out, h->d_pts, sizeof(double) * 2 * std::min(total, max_points), cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    return PP_OK;
}

pp_status pp_radius_topk(pp_handle h, const double *pts_xy, uint32_t N, const double *queries_xy, uint32_t Q,
                         double radius_arg, uint32_t K, uint32_t *idx_out, double *d2_out, uint32_t *cnt_out)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, K >= 1 && K <= RK_MAXK, PP_ERR_RANGE, "pp_radius_topk: K=%u outside 1..%d", K, RK_MAXK);
    if (Q == 0)
        return PP_OK;
    PP_REQUIRE(h, queries_xy && idx_out && d2_out && cnt_out, PP_ERR_INVALID, "pp_radius_topk: null buffer");
    PP_CUDA(h, cudaSetDevice(h->device));
    const double *d_pts = h->d_pts;
    if (pts_xy)
    {
        PP_CUDA(h, pp_reserve(h->d_pts, h->pts_cap, 2 * static_cast<size_t>(std::max(N, 1u))));
        PP_CUDA(h, cudaMemcpyAsync(h->d_pts, pts_xy, sizeof(double) * 2 * N, cudaMemcpyHostToDevice, h->stream));
        h->n_pts = N;
        h->pts_cols = 0;  // caller-supplied points have no column order
        d_pts = h->d_pts;
    }
    else
    {
        N = h->n_pts;
        PP_REQUIRE(h, d_pts || N == 0, PP_ERR_STATE, "pp_radius_topk: no extracted points on the device");
    }
    PP_CUDA(h, pp_reserve(h->d_queries, h->q_cap, 2 * static_cast<size_t>(Q)));
    const size_t out_n = static_cast<size_t>(Q) * K;
    if (out_n > h->topk_cap)
    {
        PP_CUDA(h, cudaStreamSynchronize(h->stream));
        cudaFree(h->d_topk_idx);
        cudaFree(h->d_topk_d2);
        h->d_topk_idx = nullptr;
        h->d_topk_d2 = nullptr;
        h->topk_cap = 0;
        PP_CUDA(h, cudaMalloc(reinterpret_cast<void **>(&h->d_topk_idx), out_n * sizeof(uint32_t)));
        PP_CUDA(h, cudaMalloc(reinterpret_cast<void **>(&h->d_topk_d2), out_n * sizeof(double)));
        h->topk_cap = out_n;
    }
    PP_CUDA(h, pp_reserve(h->d_topk_cnt, h->topk_q_cap, static_cast<size_t>(Q)));
    // Results land in a pinned buffer of their own and are copied out on the host: three device-to-host
    // copies into pageable memory are staged and synchronised by the driver one by one, which costs more
    // than the kernel. (The few KB of queries go down inline with the copy command.)
    const size_t d2_bytes = out_n * sizeof(double), idx_bytes = out_n * sizeof(uint32_t), cnt_bytes = Q * sizeof(uint32_t);
    if (d2_bytes + idx_bytes + cnt_bytes > h->teb_host_cap)
    {
        if (h->h_teb)
            PP_CUDA(h, cudaFreeHost(h->h_teb));
        h->h_teb = nullptr;
        h->teb_host_cap = 0;
        PP_CUDA(h, cudaHostAlloc(reinterpret_cast<void **>(&h->h_teb), d2_bytes + idx_bytes + cnt_bytes, cudaHostAllocDefault));
        h->teb_host_cap = d2_bytes + idx_bytes + cnt_bytes;
    }
    PP_CUDA(h, cudaMemcpyAsync(h->d_queries, queries_xy, sizeof(double) * 2 * Q, cudaMemcpyHostToDevice, h->stream));
    const bool prune = h->pts_cols > 0 && h->pts_col_res > 0.0 && getenv("PP_TEB_NO_PRUNE") == nullptr;
    radius_topk_kernel<<<Q, RK_THREADS, 0, h->stream>>>(d_pts, N, h->d_queries, radius_arg, K, h->d_topk_idx, h->d_topk_d2,
                                                        h->d_topk_cnt, prune ? h->d_col_counts : nullptr, h->pts_cols,
                                                        h->pts_col_origin, h->pts_col_res);
    h->launches++;
    PP_CUDA(h, cudaGetLastError());
    h->topk_Q = Q;
    h->topk_K = K;
    uint8_t *s_d2 = h->h_teb, *s_idx = s_d2 + d2_bytes, *s_cnt = s_idx + idx_bytes;
    PP_CUDA(h, cudaMemcpyAsync(s_d2, h->d_topk_d2, d2_bytes, cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaMemcpyAsync(s_idx, h->d_topk_idx, idx_bytes, cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaMemcpyAsync(s_cnt, h->d_topk_cnt, cnt_bytes, cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    std::memcpy(d2_out, s_d2, d2_bytes);
    std::memcpy(idx_out, s_idx, idx_bytes);
    std::memcpy(cnt_out, s_cnt, cnt_bytes);
    return PP_OK;
}

}  // extern "C"
