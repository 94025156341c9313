// TEST INFRASTRUCTURE ONLY (oracle). Never linked into, loaded by, or called from the product
// library; only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / reference arm use it.
//
// Thin extern "C" drivers around the reference's OWN sources, which oracle/Makefile compiles
// unmodified from /root/reference into oracle/_ref/libpp_ref.so:
//   src/path_planner/src/planner/{collision_checker,astar_hybrid,motion_hybrid,node_hybrid,commons}.cpp
//   src/path_follower/src/third_party/nanoflann.hpp (header only)
// against the stand-in headers in oracle/shim/. Private members are reached with
// -Dprivate=public on this translation unit only.
//
// Definitions implemented here:
//   volume   V[t][y][x] = (uint8) cost of CollisionChecker::checkAndGetFootprintCost
//            (collision_checker.cpp:144-172) at x=fl(ix+ox), y=fl(iy+oy), theta=fl(fl(t+ot)*bin)
//            with bin = 2.0f*(float)M_PI/(float)n_theta (motion_hybrid.cpp:10)
//   field    raw table of AstarHybrid::cacheObstacleHeuristic / getObstacleHeuristic
//            (astar_hybrid.cpp:97-241): >0 closed, <0 open, 0 unseen
//   radius   KDTree2D::radiusSearch + keep first K, as teb_optimizer.cpp:294-311
#define private public
#include "collision_checker.h"
#include "astar_hybrid.h"
#undef private
#include "nanoflann.hpp"

#include <cstdint>
#include <cstring>
#include <chrono>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace
{
struct Scene
{
    costmap_2d::Costmap2D map;
    CollisionChecker checker;
    Scene(const uint8_t *cost, uint32_t W, uint32_t H, double res, double origin_x, double origin_y,
          const float *boundary, int nb, int hole)
        : map(W, H, res, origin_x, origin_y)
    {
        std::memcpy(map.getCharMap(), cost, static_cast<size_t>(W) * H);
        checker.setCostMap(&map);
        checker.setHole(hole != 0);
        if (nb > 0)
        {
            std::vector<float> b(boundary, boundary + nb);
            checker.initBoundary(b);
        }
    }
};

// restated from teb_path_follower/commons.h:182-199 (that header needs g2o + Eigen, absent here)
struct Pt2 { double x, y; };
struct Cloud2D
{
    std::vector<Pt2> pts;
    inline size_t kdtree_get_point_count() const { return pts.size(); }
    inline double kdtree_get_pt(const size_t idx, const size_t dim) const { return dim == 0 ? pts[idx].x : pts[idx].y; }
    template <class BBOX>
    bool kdtree_get_bbox(BBOX &) const { return false; }
};
using Tree2D = nanoflann::KDTreeSingleIndexAdaptor<nanoflann::L2_Simple_Adaptor<double, Cloud2D>, Cloud2D, 2>;
}

extern "C"
{
int ref_trig_is_float()
{
#ifdef PP_ORACLE_DOUBLE_TRIG
    return 0;
#else
    return 1;
#endif
}

// Footprint offsets exactly as CollisionChecker::initBoundary produces them (collision_checker.cpp:3-81).
int ref_footprint_points(double resolution, const float *boundary, int nb, int hole, int32_t *out_xy, int max_pts)
{
    uint8_t dummy = 0;
    Scene s(&dummy, 1, 1, resolution, 0.0, 0.0, boundary, nb, hole);
    int n = static_cast<int>(s.checker._points.size());
    for (int i = 0; i < n && i < max_pts; ++i)
    {
        out_xy[2 * i] = static_cast<int32_t>(s.checker._points[i].x);
        out_xy[2 * i + 1] = static_cast<int32_t>(s.checker._points[i].y);
    }
    return n;
}

int ref_se2_volume(const uint8_t *cost, uint32_t W, uint32_t H, double res, const float *boundary, int nb, int hole,
                   uint32_t n_theta, float ox, float oy, float ot, uint32_t t0, uint32_t t1, uint32_t y0, uint32_t y1,
                   uint8_t *out, int nthreads)
{
    Scene s(cost, W, H, res, 0.0, 0.0, boundary, nb, hole);
    const float bin = 2.0f * static_cast<float>(M_PI) / static_cast<float>(n_theta);
    const size_t rows = y1 - y0;
#ifdef _OPENMP
    if (nthreads > 0)
        omp_set_num_threads(nthreads);
#endif
    const long long total = static_cast<long long>(t1 - t0) * static_cast<long long>(rows);
#pragma omp parallel for schedule(dynamic, 4)
    for (long long job = 0; job < total; ++job)
    {
        // CollisionChecker is read-only after initBoundary, so threads may share it.
        uint32_t t = t0 + static_cast<uint32_t>(job / static_cast<long long>(rows));
        uint32_t iy = y0 + static_cast<uint32_t>(job % static_cast<long long>(rows));
        float tf = static_cast<float>(t) + ot;
        float theta = tf * bin;
        float y = static_cast<float>(iy) + oy;
        uint8_t *dst = out + (static_cast<size_t>(t - t0) * rows + (iy - y0)) * W;
        for (uint32_t ix = 0; ix < W; ++ix)
        {
            float x = static_cast<float>(ix) + ox;
            float c;
            s.checker.checkAndGetFootprintCost(x, y, theta, c);
            dst[ix] = static_cast<uint8_t>(c);
        }
    }
    return 0;
}

// Arbitrary continuous poses (cell units, radians): cost + flag of checkAndGetFootprintCost and
// the early-exit boolean of checkFootPrintCollision (collision_checker.cpp:126-142).
int ref_footprint_cost_batch(const uint8_t *cost, uint32_t W, uint32_t H, double res, const float *boundary, int nb,
                             int hole, const float *poses_xyt, uint32_t n, float *cost_out, uint8_t *flag_out,
                             uint8_t *collision_out)
{
    Scene s(cost, W, H, res, 0.0, 0.0, boundary, nb, hole);
    for (uint32_t i = 0; i < n; ++i)
    {
        float c = 0.0f;
        bool f = s.checker.checkAndGetFootprintCost(poses_xyt[3 * i], poses_xyt[3 * i + 1], poses_xyt[3 * i + 2], c);
        if (cost_out)
            cost_out[i] = c;
        if (flag_out)
            flag_out[i] = f ? 1 : 0;
        if (collision_out)
            collision_out[i] = s.checker.checkFootPrintCollision(poses_xyt[3 * i], poses_xyt[3 * i + 1], poses_xyt[3 * i + 2]) ? 1 : 0;
    }
    return 0;
}

// Heuristic table. start/goal are continuous cell coordinates as setStart/setGoal receive them
// (astar_hybrid.cpp:45-66). After cacheObstacleHeuristic(), each (qx,qy) in queries is passed
// through getObstacleHeuristic (lazy resume, astar_hybrid.cpp:166-241) and its return value stored.
// A start on the map border (never entered, :141) forces expansion of the whole reachable set.
// Returns 1 if cacheObstacleHeuristic() reported the start closed, else 0.
int ref_heuristic(const uint8_t *cost, uint32_t W, uint32_t H, double res, float start_x, float start_y,
                  float goal_x, float goal_y, float cost_penalty, uint32_t n_theta, const float *queries_xy,
                  uint32_t n_queries, float *query_out, float *table_out, double *seconds_out)
{
    float circle[1] = {static_cast<float>(res)};
    Scene s(cost, W, H, res, 0.0, 0.0, circle, 1, 0);
    SearchInfo info;
    info.size_theta = n_theta;
    info.cost_penalty = cost_penalty;
    info.motion_type = MotionType::Dubins;
    AstarHybrid astar(&info, &s.checker);
    astar.clearGraph();
    astar.setStart(start_x, start_y, 0.0f);
    astar.setGoal(goal_x, goal_y, 0.0f);
    auto tic = std::chrono::steady_clock::now();
    bool ok = astar.cacheObstacleHeuristic();
    auto toc = std::chrono::steady_clock::now();
    if (seconds_out)
        *seconds_out = std::chrono::duration<double>(toc - tic).count();
    for (uint32_t i = 0; i < n_queries; ++i)
    {
        Coordinate q(queries_xy[2 * i], queries_xy[2 * i + 1], 0.0f);
        float h = astar.getObstacleHeuristic(q, astar._goal->coordinate());
        if (query_out)
            query_out[i] = h;
    }
    if (table_out)
        std::memcpy(table_out, astar._obstacle_heuristic_lookup_table.data(), sizeof(float) * static_cast<size_t>(W) * H);
    return ok ? 1 : 0;
}

// teb_optimizer.cpp:294-311 with the vendored nanoflann: radiusSearch(query, radius_arg, results)
// (sorted ascending), then keep the first min(K, n). radius_arg is passed through untouched: the
// L2 adaptor compares SQUARED distances against it (nanoflann.hpp:1747-1758, :404-408).
int ref_radius_topk(const double *pts_xy, uint32_t N, const double *q_xy, uint32_t Q, double radius_arg, uint32_t K,
                    uint32_t *idx_out, double *d2_out, uint32_t *cnt_out, uint32_t *found_out, double *build_seconds,
                    double *query_seconds)
{
    Cloud2D cloud;
    Tree2D tree(2, cloud);   // built over the still-empty cloud, as _kdtree(2, _obstacles) at teb_optimizer.cpp:4
    cloud.pts.resize(N);
    for (uint32_t i = 0; i < N; ++i)
    {
        cloud.pts[i].x = pts_xy[2 * i];
        cloud.pts[i].y = pts_xy[2 * i + 1];
    }
    auto tic = std::chrono::steady_clock::now();
    tree.buildIndex();       // teb_optimizer.h:29
    auto mid = std::chrono::steady_clock::now();
    std::vector<nanoflann::ResultItem<uint32_t, double>> results;
    for (uint32_t q = 0; q < Q; ++q)
    {
        results.clear();
        double query[2] = {q_xy[2 * q], q_xy[2 * q + 1]};
        tree.radiusSearch(query, radius_arg, results);
        uint32_t n = std::min<uint32_t>(K, static_cast<uint32_t>(results.size()));
        if (found_out)
            found_out[q] = static_cast<uint32_t>(results.size());
        cnt_out[q] = n;
        for (uint32_t j = 0; j < n; ++j)
        {
            idx_out[static_cast<size_t>(q) * K + j] = results[j].first;
            d2_out[static_cast<size_t>(q) * K + j] = results[j].second;
        }
    }
    auto toc = std::chrono::steady_clock::now();
    if (build_seconds)
        *build_seconds = std::chrono::duration<double>(mid - tic).count();
    if (query_seconds)
        *query_seconds = std::chrono::duration<double>(toc - mid).count();
    return 0;
}

// The unmodified reference planner (continuous-pose footprint checks, its own lazy heuristic),
// AstarHybrid::createPath astar_hybrid.cpp:392-433. Reported for context next to the table-driven runs.
struct RefPlanParams
{
    int motion_type;  // 1 Dubins, 2 ReedsShepp
    unsigned int n_theta;
    float minimum_turning_radius, cost_penalty, change_penalty, non_straight_penalty, reverse_penalty;
    float analytic_expansion_ratio, analytic_expansion_tolerance, tolerance;
    unsigned int max_iterations, max_on_approach_iterations;
};

int ref_plan(const uint8_t *cost, uint32_t W, uint32_t H, double res, const float *boundary, int nb, int hole,
             const RefPlanParams *p, const float *start_xyt, const float *goal_xyt, float *path_out, int max_pts, int *n_pts,
             int *iterations, double *plan_ms)
{
    Scene s(cost, W, H, res, 0.0, 0.0, boundary, nb, hole);
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
    info.max_millsec = 1000000000;
    AstarHybrid astar(&info, &s.checker);
    Coordinate start(start_xyt[0], start_xyt[1], start_xyt[2]), goal(goal_xyt[0], goal_xyt[1], goal_xyt[2]);
    Coordinates path;
    int it = 0;
    auto tic = std::chrono::steady_clock::now();
    const bool ok = astar.createPath(start, goal, path, it);
    *plan_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tic).count();
    *iterations = it;
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
}
