// N1: pp_plan -- GPU tables + host Hybrid-A* (see include/pp_b200.h, host/hybrid_astar.h).
#include "pp_internal.h"
#include "../host/hybrid_astar.h"

#include <algorithm>
#include <chrono>
#include <vector>

namespace
{
template <class T>
cudaError_t reserve_pinned(T *&ptr, size_t &cap, size_t need)
{
    if (need <= cap && ptr)
        return cudaSuccess;
    if (ptr)
        cudaFreeHost(ptr);
    ptr = nullptr;
    cap = 0;
    cudaError_t e = cudaHostAlloc(reinterpret_cast<void **>(&ptr), need * sizeof(T), cudaHostAllocDefault);
    if (e == cudaSuccess)
        cap = need;
    return e;
}
std::vector<ppb::Query> *query_log(pp_handle h) { return static_cast<std::vector<ppb::Query> *>(h->plan_log); }
// on-demand mirror of the SE2 volume for the host search (see pp_plan)
struct BandPager
{
    static constexpr unsigned int SHIFT = 6, ROWS = 1u << SHIFT;
    pp_handle h;
    uint32_t n_theta;
    std::vector<uint8_t> ready;       // what the search tests: 1 = the band is in the mirror
    cudaError_t error;
    std::vector<uint8_t> issued;      // a copy of the band has been queued on the side stream
    std::vector<cudaEvent_t> done;    // recorded behind each queued copy (created on first use)
    int toward = 0;                   // +1 / -1: the goal lies in higher / lower rows than the start

    void issue(unsigned int band)
    {
        if (issued[band] || error != cudaSuccess)
            return;
        const size_t plane = static_cast<size_t>(h->W) * h->H;
        const uint32_t y0 = band * ROWS, y1 = std::min<uint32_t>(h->H, y0 + ROWS);
        const size_t offset = static_cast<size_t>(y0) * h->W, bytes = static_cast<size_t>(y1 - y0) * h->W;
        // one strided copy: a "row" of it is this band inside one heading bin
        cudaError_t e = cudaMemcpy2DAsync(h->h_vol_mirror + offset, plane, h->d_vol + offset, plane, bytes, n_theta,
                                          cudaMemcpyDeviceToHost, h->aux);
        if (e == cudaSuccess && !done[band])
            e = cudaEventCreateWithFlags(&done[band], cudaEventDisableTiming);
        if (e == cudaSuccess)
            e = cudaEventRecord(done[band], h->aux);
        if (e != cudaSuccess)
            error = e;
        issued[band] = 1;
    }
    // called by the search on its first lookup in a band: wait for that band only, and queue the next
    // band towards the goal so that it travels while the search works on this one
    static void fetch(void *ctx, unsigned int band)
    {
        BandPager *p = static_cast<BandPager *>(ctx);
        p->issue(band);
        const long next = static_cast<long>(band) + p->toward;
        if (p->toward != 0 && next >= 0 && next < static_cast<long>(p->ready.size()))
            p->issue(static_cast<unsigned int>(next));
        if (p->error == cudaSuccess)
        {
            const cudaError_t e = cudaEventSynchronize(p->done[band]);
            if (e != cudaSuccess)
                p->error = e;
        }
        p->ready[band] = 1;  // also after an error: the search must not spin on it; pp_plan reports the error
    }
    ~BandPager()
    {
        if (!done.empty())
            cudaStreamSynchronize(h->aux);  // nothing may still be landing in the mirror when pp_plan returns
        for (cudaEvent_t e : done)
            if (e)
                cudaEventDestroy(e);
    }
};
}  // namespace

extern "C" pp_status pp_plan(pp_handle h, const pp_plan_params *params, double resolution, double origin_x, double origin_y,
                             const float *start_xyt, const float *goal_xyt, float *path_xyt_out, uint32_t max_points,
                             uint32_t *n_points_out, int32_t *iterations_out, int32_t *found_out, double *tables_ms_out,
                             double *search_ms_out)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, params && start_xyt && goal_xyt && n_points_out && iterations_out && found_out, PP_ERR_INVALID, "pp_plan: null argument");
    PP_REQUIRE(h, h->d_cost, PP_ERR_STATE, "pp_plan: upload a costmap first");
    PP_REQUIRE(h, !h->fp.empty(), PP_ERR_STATE, "pp_plan: set a footprint first");
    PP_REQUIRE(h, params->motion_type == 1 || params->motion_type == 2, PP_ERR_INVALID, "pp_plan: motion_type must be 1 (Dubins) or 2 (Reeds-Shepp)");
    PP_REQUIRE(h, params->n_theta > 0 && resolution > 0.0, PP_ERR_INVALID, "pp_plan: n_theta and resolution must be positive");
    // search nodes are numbered in 32 bits, like the reference's (node_hybrid.h: unsigned int index)
    PP_REQUIRE(h, static_cast<uint64_t>(h->W) * h->H * params->n_theta < 0xFFFFFFFFull, PP_ERR_RANGE,
               "pp_plan: %ux%ux%u search nodes do not fit the 32-bit node numbering", h->W, h->H, params->n_theta);
    *n_points_out = 0;
    *iterations_out = 0;
    *found_out = 0;

    ppb::PlanParams pp;
    pp.motion_type = params->motion_type;
    pp.n_theta = params->n_theta;
    pp.minimum_turning_radius = params->minimum_turning_radius;
    pp.cost_penalty = params->cost_penalty;
    pp.change_penalty = params->change_penalty;
    pp.non_straight_penalty = params->non_straight_penalty;
    pp.reverse_penalty = params->reverse_penalty;
    pp.analytic_expansion_ratio = params->analytic_expansion_ratio;
    pp.analytic_expansion_tolerance = params->analytic_expansion_tolerance;
    pp.tolerance = params->tolerance;
    pp.max_iterations = params->max_iterations;
    pp.max_on_approach_iterations = params->max_on_approach_iterations;
    ppb::HybridAStar search(pp, h->W, h->H, resolution, origin_x, origin_y);
    const bool exact_analytic = params->flags & PP_PLAN_EXACT_ANALYTIC, validate = params->flags & PP_PLAN_VALIDATE_PATH;
    h->plan_stats = pp_plan_stats_t{};
    if (h->plan_log_on && !h->plan_log)
        h->plan_log = new std::vector<ppb::Query>();
    if (h->plan_log)
        query_log(h)->clear();
    if (h->plan_log_on)
        search.setQueryLog(query_log(h));
    const ppb::Pose start{start_xyt[0], start_xyt[1], start_xyt[2]}, goal{goal_xyt[0], goal_xyt[1], goal_xyt[2]};

    const auto tic = std::chrono::steady_clock::now();
    unsigned int gx = 0, gy = 0;
    if (!search.goalCell(goal, gx, gy))
        return PP_OK;  // goal outside the map: no path, like worldToMap failing at astar_hybrid.cpp:407
    const float frac = search.headingFraction(start.theta);
    pp_status st = pp_se2_volume_build(h, params->n_theta, 0.5f, 0.5f, frac, 0, params->n_theta);
    if (st != PP_OK)
        return st;
    const size_t cells = static_cast<size_t>(h->W) * h->H;
    PP_CUDA(h, reserve_pinned(h->h_vol_mirror, h->vol_mirror_cap, cells * params->n_theta));
    PP_CUDA(h, reserve_pinned(h->h_field_mirror, h->field_mirror_cap, cells));
    // The volume travels to the host on the side stream while the heuristic field is built on the main
    // one. A small volume goes as a whole; a large one is paged in during the search, by bands of 64 map
    // rows (all heading bins of a band in one strided copy), because a search only touches the rows of
    // its corridor and the host link, not the GPU, is what a 300 MB .. 1.2 GB mirror waits for.
    PP_CUDA(h, cudaEventRecord(h->ev_fork, h->stream));
    PP_CUDA(h, cudaStreamWaitEvent(h->aux, h->ev_fork, 0));
    const bool paged = cells * params->n_theta > (static_cast<size_t>(64) << 20);
    BandPager pager{h, params->n_theta, {}, cudaSuccess, {}, {}, 0};
    if (paged)
    {
        const size_t n_bands = (h->H + BandPager::ROWS - 1) / BandPager::ROWS;
        pager.ready.assign(n_bands, 0);
        pager.issued.assign(n_bands, 0);
        pager.done.assign(n_bands, nullptr);
        pager.toward = goal.y > start.y ? 1 : (goal.y < start.y ? -1 : 0);
    }
    else
        PP_CUDA(h, cudaMemcpyAsync(h->h_vol_mirror, h->d_vol, cells * params->n_theta, cudaMemcpyDeviceToHost, h->aux));
    if (exact_analytic)
    {
        // the continuous-pose re-checks of the search read the costmap on the host
        PP_CUDA(h, reserve_pinned(h->h_cost_mirror, h->cost_mirror_cap, cells));
        PP_CUDA(h, cudaMemcpyAsync(h->h_cost_mirror, h->d_cost, cells, cudaMemcpyDeviceToHost, h->aux));
    }
    st = pp_heuristic_build(h, gx, gy, params->cost_penalty, nullptr);
    if (st != PP_OK)
    {
        cudaStreamSynchronize(h->aux);
        return st;
    }
    PP_CUDA(h, cudaMemcpyAsync(h->h_field_mirror, h->d_field, cells * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->aux));
    const auto mid = std::chrono::steady_clock::now();

    ppb::Tables tb;
    tb.volume = h->h_vol_mirror;
    tb.field = h->h_field_mirror;
    tb.W = h->W;
    tb.H = h->H;
    tb.n_theta = params->n_theta;
    tb.heading_fraction = frac;
    if (paged)
    {
        tb.band_ready = pager.ready.data();
        tb.band_shift = BandPager::SHIFT;
        tb.fetch_band = &BandPager::fetch;
        tb.ctx = &pager;
    }
    if (exact_analytic)
    {
        tb.exact_analytic = true;
        tb.cost = h->h_cost_mirror;
        tb.fp = h->fp.data();
        tb.n_fp = static_cast<unsigned int>(h->fp.size() / 2);
        tb.confirm_ctx = h;
        tb.confirm_batch = [](void *ctx, const float *poses, unsigned int n, uint8_t *flags) {
            std::vector<float> costs(n);
            return pp_footprint_cost_batch(static_cast<pp_handle>(ctx), poses, n, costs.data(), flags) == PP_OK;
        };
    }
    std::vector<ppb::Pose> path;
    int iterations = 0;
    bool found = false;
    const char *fake_hit = getenv("PP_DEBUG_PLAN_FAKE_HIT");  // test hook: pretend that pose k of the first path collides
    for (uint32_t attempt = 0;; ++attempt)
    {
        found = search.plan(tb, start, goal, path, iterations);
        const ppb::PlanStats &ps = search.stats();
        h->plan_stats.lookups += ps.lookups;
        h->plan_stats.analytic_attempts += ps.analytic_attempts;
        h->plan_stats.analytic_confirm_batches += ps.analytic_confirm_batches;
        h->plan_stats.exact_host_checks += ps.exact_host_checks;
        h->plan_stats.false_collisions_overruled += ps.false_collisions_overruled;
        h->plan_stats.false_free_caught += ps.false_free_caught;
        if (!found || !validate)
            break;
        // every pose of the path at its continuous position, exactly, in one batched launch
        const std::vector<ppb::Pose> &pc = search.pathCells();
        std::vector<float> poses(3 * pc.size()), costs(pc.size());
        std::vector<uint8_t> flags(pc.size());
        for (size_t i = 0; i < pc.size(); ++i)
        {
            poses[3 * i] = pc[i].x;
            poses[3 * i + 1] = pc[i].y;
            poses[3 * i + 2] = pc[i].theta;
        }
        pp_status vs = pp_footprint_cost_batch(h, poses.data(), static_cast<uint32_t>(pc.size()), costs.data(), flags.data());
        if (vs != PP_OK)
            return vs;
        if (fake_hit && attempt == 0 && static_cast<size_t>(atoi(fake_hit)) < flags.size())
            flags[atoi(fake_hit)] = 1;
        h->plan_stats.path_poses_checked = static_cast<uint32_t>(pc.size());
        h->plan_stats.path_poses_in_collision = 0;
        uint32_t newly_blocked = 0;
        const std::vector<unsigned int> &nodes = search.pathNodes();
        for (size_t i = 0; i < flags.size(); ++i)
            if (flags[i])
            {
                ++h->plan_stats.path_poses_in_collision;
                if (i > 0 && i + 1 < flags.size() && i < nodes.size())  // the goal and start poses themselves cannot be avoided
                {
                    search.block(nodes[i]);
                    ++newly_blocked;
                }
            }
        // a pose that the volume let through but the continuous check rejects: search again without those nodes
        if (h->plan_stats.path_poses_in_collision == 0 || newly_blocked == 0 || attempt == 3)
            break;
        ++h->plan_stats.replans;
    }
    if (paged)
        cudaStreamSynchronize(h->aux);  // a prefetched band may still be on its way: part of the plan's time
    const auto toc = std::chrono::steady_clock::now();
    PP_REQUIRE(h, pager.error == cudaSuccess, PP_ERR_CUDA, "pp_plan: paging the volume mirror failed: %s",
               cudaGetErrorString(pager.error));

    *found_out = found ? 1 : 0;
    *iterations_out = iterations;
    *n_points_out = static_cast<uint32_t>(path.size());
    if (path_xyt_out)
        for (size_t i = 0; i < path.size() && i < max_points; ++i)
        {
            path_xyt_out[3 * i] = path[i].x;
            path_xyt_out[3 * i + 1] = path[i].y;
            path_xyt_out[3 * i + 2] = path[i].theta;
        }
    if (tables_ms_out)
        *tables_ms_out = std::chrono::duration<double, std::milli>(mid - tic).count();
    if (search_ms_out)
        *search_ms_out = std::chrono::duration<double, std::milli>(toc - mid).count();
    return PP_OK;
}

extern "C" pp_status pp_plan_stats(pp_handle h, pp_plan_stats_t *out)
{
    if (!h || !out)
        return PP_ERR_INVALID;
    *out = h->plan_stats;
    return PP_OK;
}

extern "C" pp_status pp_plan_set_query_log(pp_handle h, int enable)
{
    if (!h)
        return PP_ERR_INVALID;
    h->plan_log_on = enable != 0;
    if (!enable)
        pp_plan_log_free(h);
    return PP_OK;
}

extern "C" pp_status pp_plan_query_log(pp_handle h, float *poses_xyt_out, uint8_t *kind_out, uint8_t *flag_out, uint8_t *cost_out,
                                       uint32_t max_queries, uint32_t *n_out)
{
    if (!h || !n_out)
        return PP_ERR_INVALID;
    static const std::vector<ppb::Query> none;
    const std::vector<ppb::Query> &log = h->plan_log ? *query_log(h) : none;
    *n_out = static_cast<uint32_t>(log.size());
    const size_t n = std::min<size_t>(log.size(), max_queries);
    for (size_t i = 0; i < n; ++i)
    {
        const ppb::Query &q = log[i];
        if (poses_xyt_out)
        {
            poses_xyt_out[3 * i] = q.x;
            poses_xyt_out[3 * i + 1] = q.y;
            poses_xyt_out[3 * i + 2] = q.theta;
        }
        if (kind_out)
            kind_out[i] = q.kind;
        if (flag_out)
            flag_out[i] = q.flag;
        if (cost_out)
            cost_out[i] = q.cost;
    }
    return PP_OK;
}

void pp_plan_log_free(pp_handle h)
{
    delete query_log(h);
    h->plan_log = nullptr;
}
