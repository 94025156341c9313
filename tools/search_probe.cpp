// dev probe (not part of the product): times HybridAStar::plan on the host against tables supplied by the
// caller (tools/search_probe.py builds them with the CPU oracle), so the search can be profiled without a GPU.
#include "../path_planner_b200/host/hybrid_astar.h"

#include <chrono>
#include <cstring>

extern "C" float probe_heading_fraction(int motion_type, unsigned n_theta, unsigned W, unsigned H, double res, float start_theta)
{
    ppb::PlanParams p;
    p.motion_type = motion_type;
    p.n_theta = n_theta;
    ppb::HybridAStar s(p, W, H, res, 0.0, 0.0);
    return s.headingFraction(start_theta);
}

extern "C" int probe_plan(const uint8_t *vol, const float *field, unsigned W, unsigned H, unsigned n_theta, float frac, double res,
                          const float *start, const float *goal, int motion_type, unsigned max_iter, float *path, int max_pts,
                          int *n_out, int *iters, double *ms, int reps)
{
    ppb::PlanParams p;
    p.motion_type = motion_type;
    p.n_theta = n_theta;
    p.max_iterations = max_iter;
    ppb::Tables tb;
    tb.volume = vol, tb.field = field, tb.W = W, tb.H = H, tb.n_theta = n_theta, tb.heading_fraction = frac;
    std::vector<ppb::Pose> out;
    bool ok = false;
    double best = 1e30;
    for (int r = 0; r < reps; ++r)
    {
        ppb::HybridAStar s(p, W, H, res, 0.0, 0.0);
        const auto t0 = std::chrono::steady_clock::now();
        ok = s.plan(tb, {start[0], start[1], start[2]}, {goal[0], goal[1], goal[2]}, out, *iters);
        const double t = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        best = t < best ? t : best;
    }
    *ms = best;
    *n_out = static_cast<int>(out.size());
    for (int i = 0; i < *n_out && i < max_pts; ++i)
        std::memcpy(path + 3 * i, &out[i], 12);
    return ok;
}
