// Host Hybrid-A* driver over the GPU-built tables.
//
// Same search as the reference's AstarHybrid::createPath (astar_hybrid.cpp:392-515, with
// getNeighbors :264-291, getTravelCost :368-390, getHeuristicCost :242-247,
// tryAnalyticExpansion / getAnalyticExpansion :293-366, MotionHybrid motion_hybrid.cpp:10-92),
// written from scratch against two O(1) tables instead of per-pose work:
//   * footprint checks  -> SE2 volume lookup at the pose's cell and nearest heading bin
//                          (what CollisionChecker::checkAndGetFootprintCost costs P gathers for)
//   * obstacle heuristic -> max(field[cell], diagonal distance): the table-hit branch of
//                          getObstacleHeuristic (:177-183); the lazy A* (:97-165, :184-239) is gone
// Evaluation order, float widths and the priority-queue discipline follow the reference so that,
// on the same tables, the same path and iteration count come out
// (tests/test_gpu_e2e_plan.py compares against the reference's own search).
// The wall-clock cut-off of the reference (max_millsec, :469) is not reproduced: it makes the
// result depend on machine load.
#pragma once
#include "curves.h"

#include <cstdint>
#include <deque>
#include <memory>
#include <queue>
#include <vector>

namespace ppb
{
struct PlanParams  // SearchInfo, planner/commons.h:26-48; defaults of config/params.yaml:32-45
{
    int motion_type = 2;  // 1 Dubins, 2 Reeds-Shepp
    unsigned int n_theta = 72;
    float minimum_turning_radius = 0.84f, cost_penalty = 2.0f, change_penalty = 0.01f, non_straight_penalty = 1.2f,
          reverse_penalty = 3.0f;
    float analytic_expansion_ratio = 1.0f, analytic_expansion_tolerance = 3.0f, tolerance = 0.6f;
    unsigned int max_iterations = 100000, max_on_approach_iterations = 1000;
};

struct Pose
{
    float x, y, theta;
};

struct Tables
{
    const uint8_t *volume = nullptr;  // [n_theta][H][W], built with sample offsets (0.5, 0.5, heading_fraction)
    const float *field = nullptr;     // [H][W] obstacle heuristic for the goal cell, 0 = unreachable
    unsigned int W = 0, H = 0, n_theta = 0;
    float heading_fraction = 0.0f;
    // Optional paging of the volume mirror by bands of (1 << band_shift) map rows: where band_ready is
    // set, a lookup in a band whose flag is 0 first calls fetch_band(ctx, band), which must bring the
    // rows of that band (all heading bins) into `volume` and set the flag. A search touches a corridor.
    uint8_t *band_ready = nullptr;
    unsigned int band_shift = 0;
    void (*fetch_band)(void *ctx, unsigned int band) = nullptr;
    void *ctx = nullptr;
    // Exact analytic expansion (optional). The reference checks every sample of an analytic expansion at the
    // CONTINUOUS pose (astar_hybrid.cpp:315-329 -> collision_checker.cpp:126-142); the volume answers for the
    // cell centre and the nearest heading bin. With exact_analytic set, the volume is only a filter: samples it
    // flags are re-checked exactly on the host against `cost` (the expansion fails if one of them really
    // collides), and an expansion that survives is confirmed by ONE batched exact evaluation of all its samples
    // (confirm_batch: pp_footprint_cost_batch on the GPU). The outcome of every expansion then equals the
    // reference's.
    bool exact_analytic = false;
    const uint8_t *cost = nullptr;   // host copy of the costmap [H][W]
    const int32_t *fp = nullptr;     // footprint offsets x0,y0,x1,y1,... (cells)
    unsigned int n_fp = 0;
    bool (*confirm_batch)(void *ctx, const float *poses_xyt, unsigned int n, uint8_t *flags_out) = nullptr;
    void *confirm_ctx = nullptr;
};

struct Query  // one footprint question of a search, for the quantisation statistics (pp_plan_query_log)
{
    float x, y, theta;     // cells, cells, radians: what the reference would pass to its CollisionChecker
    uint8_t kind;          // 0 = checkAndGetFootprintCost (primitive / end pose), 1 = checkFootPrintCollision (analytic sample)
    uint8_t flag, cost;    // what the volume lookup answered
};

struct PlanStats
{
    unsigned long long lookups = 0;            // volume lookups
    unsigned int analytic_attempts = 0;        // analytic expansions tried
    unsigned int analytic_confirm_batches = 0; // batched exact confirmations (exact mode)
    unsigned int exact_host_checks = 0;        // samples re-checked exactly on the host (exact mode)
    unsigned int false_collisions_overruled = 0;  // ... of which the exact check said free
    unsigned int false_free_caught = 0;        // samples the volume called free that the exact batch rejected
};

class HybridAStar
{
public:
    HybridAStar(const PlanParams &params, unsigned int W, unsigned int H, double resolution, double origin_x, double origin_y);

    // fractional part of the heading coordinate every node of a search from `start_theta` has
    float headingFraction(float start_theta) const;
    // goal cell of a world pose (for building the heuristic field); false if outside the map
    bool goalCell(const Pose &goal_world, unsigned int &gx, unsigned int &gy) const;

    // world poses in (metres, metres, radians); path runs goal -> start like the reference's
    // (path_planner_ros.cpp:140 reverses it afterwards)
    bool plan(const Tables &tables, const Pose &start_world, const Pose &goal_world, std::vector<Pose> &path, int &iterations);
    // after plan(): the poses of the path as the search held them (cells, cells, radians), goal -> start
    const std::vector<Pose> &pathCells() const { return _path_cells; }
    const std::vector<unsigned int> &pathNodes() const { return _path_nodes; }  // node numbers of those poses
    // search nodes the next plan() must not enter (poses of an earlier path that failed the exact check)
    void block(unsigned int node_index) { _blocked.push_back(node_index); }
    void clearBlocked() { _blocked.clear(); }
    const PlanStats &stats() const { return _stats; }
    void setQueryLog(std::vector<Query> *log) { _log = log; }

private:
    struct Node
    {
        Node *parent = nullptr;
        bool visited = false;
        float cell_cost;
        float g;
        unsigned int index;
        unsigned int motion;
        Pose at{0.0f, 0.0f, 0.0f};
        explicit Node(unsigned int i);
    };
    using QItem = std::pair<float, Node *>;
    struct Later
    {
        bool operator()(const QItem &a, const QItem &b) const { return a.first > b.first; }
    };

    Node *node(unsigned int index);
    unsigned int indexOf(unsigned int x, unsigned int y, unsigned int t) const { return t + x * _n_theta + y * _n_theta * _W; }
    void wrapHeading(float &h) const;
    bool footprint(const Tables &tb, float x, float y, float theta_rad, float &cost, uint8_t kind = 0) const;
    bool exactCollisionHost(const Tables &tb, float x, float y, float theta_rad) const;
    float heuristic(const Tables &tb, const Pose &at) const;
    float travelCost(const Node *from, const Node *to) const;
    void expand(const Tables &tb, Node *n, std::vector<Node *> &out);
    Node *analyticExpansion(const Tables &tb, Node *n);
    bool backtrace(Node *n, std::vector<Pose> &path);
    bool blocked(unsigned int index) const;

    PlanParams _p;
    unsigned int _W, _H, _n_theta;
    double _res, _ox, _oy;
    float _n_theta_f, _bin, _turn_cells, _tol_cells, _ae_ratio_cells, _ae_tol_cells, _neutral;
    std::vector<Pose> _motions;
    std::unique_ptr<CurveSpace> _curves;
    // The graph (the reference's std::unordered_map<unsigned, NodeHybrid>, astar_hybrid.h): nodes live in a
    // deque (stable addresses: the queue and the parent links hold pointers) and are found through an
    // open-addressing table of node numbers, which costs one probe where the node-based map costs an
    // allocation per insert and a pointer chase per lookup.
    std::deque<Node> _nodes;
    std::vector<unsigned int> _slot_key;  // node index + 1, 0 = empty
    std::vector<unsigned int> _slot_node; // position in _nodes
    unsigned int _slot_mask = 0;
    void resetGraph(unsigned int expected_nodes);
    void growGraph();
    Node *_start = nullptr, *_goal = nullptr;
    SE2 _goal_se2{0.0, 0.0, 0.0};
    struct HeadingEntry
    {
        uint32_t bits = 0xFFFFFFFFu;  // bit pattern of the heading coordinate this entry was computed for (NaN: none yet)
        double yaw = 0.0;
        CurveSpace::HeadingTrig trig{1.0, 0.0, 0.0, 1.0};
    };
    mutable std::vector<HeadingEntry> _heading_cache;
    std::vector<Pose> _path_cells;
    std::vector<unsigned int> _path_nodes, _blocked;
    mutable PlanStats _stats;
    std::vector<Query> *_log = nullptr;
};
}  // namespace ppb
