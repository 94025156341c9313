// TEST INFRASTRUCTURE ONLY (oracle). End-to-end "identical path" check.
//
// The reference's own search (astar_hybrid.cpp, motion_hybrid.cpp, node_hybrid.cpp, commons.cpp)
// is copied at build time into the git-ignored oracle/_ref/patched/ with ONE added line after
// astar_hybrid.cpp:438-439 (a call to pp_b200_after_cache_heuristic) and compiled against the
// drop-in CollisionChecker of path_planner_b200/host/ instead of the reference's. Two modes:
//   mode 0 (A)  collision lookups into an ORACLE-built volume handed in by the caller, and the
//               reference's own lazily grown heuristic table (no injection)
//   mode 1 (B)  collision lookups into the GPU-built volume, and the GPU heuristic field
//               injected over the reference's table right after cacheObstacleHeuristic()
//   + 4         the samples of analytic expansions (checkFootPrintCollision, astar_hybrid.cpp:325) are decided
//               at their continuous poses (CollisionChecker::setExactAnalytic): what pp_plan's
//               PP_PLAN_EXACT_ANALYTIC mode must reproduce
//   + 8         every footprint question and the volume's answer are logged (e2e_query_log)
// Pass = same Coordinates element-wise and same iteration count. Both modes quantise poses to
// the cell centre / nearest bin, which is what separates them from the unmodified reference
// (ref_plan in ref_harness.cpp, reported for context only).
#define private public
#include "astar_hybrid.h"
#undef private

#include <chrono>
#include <cstring>

namespace
{
int g_mode = 0;
double g_inject_ms = 0.0;
std::vector<CollisionChecker::Query> g_log;
}

void pp_b200_after_cache_heuristic(AstarHybrid *self)
{
    if (g_mode != 1)
        return;
    auto tic = std::chrono::steady_clock::now();
    const unsigned int gx = static_cast<unsigned int>(self->_goal->coordinate().x);
    const unsigned int gy = static_cast<unsigned int>(self->_goal->coordinate().y);
    std::vector<float> field;
    if (self->_collision_checker->buildHeuristicField(gx, gy, self->_search_info->cost_penalty, field))
        self->_obstacle_heuristic_lookup_table.swap(field);  // all entries >= 0: every later query is the hit at :182-183
    g_inject_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tic).count();
}

extern "C"
{
// fractional part of the search's heading coordinate, formed like AstarHybrid::setStart (:45-55)
float e2e_heading_fraction(float start_theta, unsigned int n_theta)
{
    const float size_theta_float = static_cast<float>(n_theta);
    const float bin = 2.0f * static_cast<float>(M_PI) / size_theta_float;
    float m = start_theta / bin + size_theta_float;
    while (m >= size_theta_float)
        m -= size_theta_float;
    while (m < 0.0)
        m += size_theta_float;
    return m - floorf(m);
}

struct E2EParams
{
    int motion_type;  // 1 Dubins, 2 ReedsShepp
    unsigned int n_theta;
    float minimum_turning_radius, cost_penalty, change_penalty, non_straight_penalty, reverse_penalty;
    float analytic_expansion_ratio, analytic_expansion_tolerance, tolerance;
    unsigned int max_iterations, max_on_approach_iterations;
};

int e2e_plan(int mode, const uint8_t *cost, uint32_t W, uint32_t H, double res, const float *boundary, int nb, int hole,
             const uint8_t *oracle_volume, const E2EParams *p, const float *start_xyt, const float *goal_xyt,
             float *path_out, int max_pts, int *n_pts, int *iterations, double *plan_ms, double *prepare_ms,
             unsigned long long *lookups, unsigned long long *exact_calls)
{
    const bool exact_analytic = mode & 4, log_queries = mode & 8;
    mode &= 3;
    g_mode = mode;
    g_inject_ms = 0.0;
    g_log.clear();
    costmap_2d::Costmap2D map(W, H, res, 0.0, 0.0);
    std::memcpy(map.getCharMap(), cost, static_cast<size_t>(W) * H);
    CollisionChecker checker;
    checker.setCostMap(&map);
    checker.setHole(hole != 0);
    std::vector<float> b(boundary, boundary + nb);
    checker.initBoundary(b);

    SearchInfo info;
    info.size_theta = p->n_theta;
    info.minimum_turning_radius = p->minimum_turning_radius;
    info.cost_penalty = p->cost_penalty;
    info.change_penalty = p->change_penalty;
    info.non_straight_penalty = p->non_straight_penalty;
    info.reverse_penalty = p->reverse_penalty;
    info.analytic_expansion_ratio = p->analytic_expansion_ratio;
    info.analytic_expansion_tolerance = p->analytic_expansion_tolerance;
    info.tolerance = p->tolerance;
    info.max_iterations = p->max_iterations;
    info.max_on_approach_iterations = p->max_on_approach_iterations;
    info.motion_type = p->motion_type == 2 ? MotionType::ReedsShepp : MotionType::Dubins;
    info.max_millsec = 1000000000;  // no wall-clock cut-off: the comparison must be deterministic

    const float frac = e2e_heading_fraction(start_xyt[2], p->n_theta);
    auto tic = std::chrono::steady_clock::now();
    if (mode == 0)
        checker.attachVolume(oracle_volume, p->n_theta, frac);
    else if (!checker.prepare(p->n_theta, frac))
    {
        std::fprintf(stderr, "e2e_plan: prepare failed: %s\n", checker.lastError());
        return -1;
    }
    *prepare_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tic).count();

    checker.setExactAnalytic(exact_analytic);
    if (log_queries)
        checker.setQueryLog(&g_log);
    AstarHybrid astar(&info, &checker);
    Coordinate start(start_xyt[0], start_xyt[1], start_xyt[2]), goal(goal_xyt[0], goal_xyt[1], goal_xyt[2]);
    Coordinates path;
    int it = 0;
    tic = std::chrono::steady_clock::now();
    const bool ok = astar.createPath(start, goal, path, it);
    *plan_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tic).count();
    *prepare_ms += g_inject_ms;
    *plan_ms -= g_inject_ms;
    *iterations = it;
    *lookups = checker.lookups();
    *exact_calls = checker.exactCalls();
    const int n = static_cast<int>(path.size());
    *n_pts = n;
    for (int i = 0; i < n && i < max_pts; ++i)
    {
        path_out[3 * i] = path[i].x;
        path_out[3 * i + 1] = path[i].y;
        path_out[3 * i + 2] = path[i].theta;
    }
    return ok ? 1 : 0;
}

// the footprint questions of the last e2e_plan run with mode + 8: pose (cells, cells, radians), kind
// (0 primitive / end pose, 1 analytic-expansion sample), and the volume's flag and cost
int e2e_query_log(float *xyt, uint8_t *kind, uint8_t *flag, uint8_t *cost, int max_n)
{
    const int n = static_cast<int>(g_log.size());
    for (int i = 0; i < n && i < max_n; ++i)
    {
        xyt[3 * i] = g_log[i].x;
        xyt[3 * i + 1] = g_log[i].y;
        xyt[3 * i + 2] = g_log[i].theta;
        kind[i] = g_log[i].kind;
        flag[i] = g_log[i].flag;
        cost[i] = g_log[i].cost;
    }
    return n;
}
}
