// Drop-in replacement for the reference's CollisionChecker
// (src/path_planner/src/planner/collision_checker.h:6-37): same class name, same public member
// functions with the same signatures, so that the reference's astar_hybrid.cpp / smoother.cpp
// compile against it unchanged when this directory precedes the reference's on the include path
// (see INTEGRATION.md). All work goes through the C ABI of libpp_b200.so; there is no CPU path.
//
// Behaviour
//   * initBoundary / setHole / setCostMap mirror the reference (collision_checker.cpp:3-81). As in
//     the reference, successive initBoundary calls APPEND points.
//   * prepare(n_theta, heading_fraction) uploads the grid (Costmap2D::getCharMap()), builds the SE2
//     volume on the GPU for sample offsets (0.5, 0.5, heading_fraction) and mirrors it in host
//     memory. After that, checkAndGetFootprintCost / checkFootPrintCollision at a pose inside the
//     map answer from the volume at the pose's cell and nearest heading bin: O(1), no P-point loop.
//     This QUANTISES the pose to the cell centre; the reference evaluates the continuous pose.
//   * exact mode (setExact(true), or before prepare()): every call is answered by
//     pp_footprint_cost_batch for the continuous pose -- bit-identical to the reference, one kernel
//     launch per call. setExactAnalytic(true) does that for checkFootPrintCollision only, i.e. for the
//     samples of analytic expansions, whose headings are not bin-aligned.
//   * point queries and getCost read the host grid exactly like the reference (:174-214).
#pragma once
#include "commons.h"
#include <costmap_2d/costmap_2d.h>
#include <costmap_2d/cost_values.h>

#include <cstdint>
#include <vector>

struct pp_ctx;

class CollisionChecker
{
public:
    CollisionChecker();
    ~CollisionChecker();
    CollisionChecker(const CollisionChecker &) = delete;
    CollisionChecker &operator=(const CollisionChecker &) = delete;

    void setCostMap(costmap_2d::Costmap2D *costmap);
    void initBoundary(float radius);
    void initBoundary(float width, float height);
    void initBoundary(float x_min, float y_min, float x_max, float y_max);
    void setHole(bool hole) { _hole = hole; }
    void initBoundary(std::vector<float> &boundary);
    void printBoundary();

    bool checkFootPrintCollision(const float &x, const float &y, const float &theta);
    bool checkAndGetFootprintCost(const float &x, const float &y, const float &theta, float &cost);
    bool isInCostMap(const float &x, const float &y);
    bool checkPointCollision(const float &x, const float &y);
    bool checkAndGetPointCost(const float &x, const float &y, float &cost);
    float getCost(const unsigned int &x, const unsigned int &y);
    costmap_2d::Costmap2D *getCostMap() { return _costmap; }

    // ---- additions (no counterpart in the reference) ----
    // Upload the current grid and build + mirror the SE2 volume. heading_fraction is the fractional
    // part of the search's heading coordinate (all nodes of one search share it, motion_hybrid.cpp:70-72).
    bool prepare(unsigned int n_theta, float heading_fraction = 0.0f);
    // Use an externally built volume ([n_theta][H][W], stays owned by the caller) instead of the GPU's.
    void attachVolume(const uint8_t *volume, unsigned int n_theta, float heading_fraction);
    void setExact(bool exact) { _exact = exact; }
    // checkFootPrintCollision -- the analytic-expansion samples of astar_hybrid.cpp:325, whose poses are not
    // bin-aligned -- always at the continuous pose (pp_footprint_cost_batch); checkAndGetFootprintCost -- the
    // motion primitives of :283 -- from the volume
    void setExactAnalytic(bool exact) { _exact_analytic = exact; }
    // record every footprint question with the volume's answer (quantisation statistics)
    struct Query
    {
        float x, y, theta;
        uint8_t kind, flag, cost;  // kind 0: checkAndGetFootprintCost, 1: checkFootPrintCollision
    };
    void setQueryLog(std::vector<Query> *log) { _log = log; }
    // Whole obstacle-heuristic field for a goal cell (replaces AstarHybrid::cacheObstacleHeuristic's table).
    bool buildHeuristicField(unsigned int goal_x, unsigned int goal_y, float cost_penalty, std::vector<float> &field);
    const char *lastError() const;
    pp_ctx *handle() { return _h; }
    unsigned long long lookups() const { return _lookups; }
    unsigned long long exactCalls() const { return _exact_calls; }

private:
    bool lookup(float x, float y, float theta, float &cost, bool &flag);
    bool exactEval(float x, float y, float theta, float &cost, bool &flag);
    bool ensureUploaded();

    std::vector<int32_t> _points;  // x0,y0,x1,y1,... integer cell offsets
    costmap_2d::Costmap2D *_costmap;
    bool _hole;
    pp_ctx *_h;
    bool _uploaded, _footprint_dirty, _exact, _exact_analytic = false;
    std::vector<Query> *_log = nullptr;
    std::vector<uint8_t> _volume_mirror;
    const uint8_t *_volume;
    unsigned int _n_theta;
    float _heading_fraction, _bin_size;
    unsigned long long _lookups, _exact_calls;
};
