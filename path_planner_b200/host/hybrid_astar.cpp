// See hybrid_astar.h. Reference lines cited are in src/path_planner/src/planner/.
#include "hybrid_astar.h"

#include <algorithm>
#include <climits>
#include <cstring>
#include <cmath>
#include <limits>

namespace ppb
{
namespace
{
// commons.cpp:6-12 and :18-24
float wrap_angle_f(float a)
{
    const float r = std::fmod(a + static_cast<float>(M_PI), 2.0f * static_cast<float>(M_PI));
    return r <= 0.0f ? r + static_cast<float>(M_PI) : r - static_cast<float>(M_PI);
}
double wrap_angle_d(double a)
{
    // fmod(s, 2 pi) is s itself for |s| < 2 pi and one exact subtraction (Sterbenz) for 2 pi <= |s| <= 4 pi;
    // fmod is exact, so skipping the library call there gives the same bits
    const double two_pi = 2.0 * M_PI, s = a + M_PI, m = std::fabs(s);
    const double r = m < two_pi ? s : (m <= 2.0 * two_pi ? (s < 0.0 ? s + two_pi : s - two_pi) : std::fmod(s, two_pi));
    return r <= 0.0 ? r + M_PI : r - M_PI;
}
}  // namespace

HybridAStar::Node::Node(unsigned int i)
    : cell_cost(std::numeric_limits<float>::quiet_NaN()), g(std::numeric_limits<float>::max()), index(i),
      motion(std::numeric_limits<unsigned int>::max())
{
}

HybridAStar::HybridAStar(const PlanParams &params, unsigned int W, unsigned int H, double resolution, double origin_x, double origin_y)
    : _p(params), _W(W), _H(H), _n_theta(params.n_theta), _res(resolution), _ox(origin_x), _oy(origin_y)
{
    // astar_hybrid.cpp:8-13: metric parameters in cells, rounded up
    const float res_f = static_cast<float>(resolution);
    _turn_cells = std::ceil(_p.minimum_turning_radius / res_f);
    _tol_cells = std::ceil(_p.tolerance / res_f);
    _ae_ratio_cells = std::ceil(_p.analytic_expansion_ratio / res_f);
    _ae_tol_cells = std::ceil(_p.analytic_expansion_tolerance / res_f);
    _neutral = sqrtf(2.0f);  // NodeHybrid::NeutralCost, node_hybrid.cpp:11
    // motion_hybrid.cpp:5-10
    _n_theta_f = static_cast<float>(_n_theta);
    _bin = 2.0f * static_cast<float>(M_PI) / _n_theta_f;
    // motion_hybrid.cpp:20-36 / :39-55: one arc that changes heading by a whole number of bins and
    // moves at least sqrt(2) cells; the trigonometry there is done in double and narrowed
    float angle = static_cast<float>(2.0 * asin(sqrt(2.0) / static_cast<double>(2 * _turn_cells)));
    float increments = 1.0f;
    if (angle > _bin)
        increments = std::ceil(angle / _bin);
    angle = _bin * increments;
    const float dx = static_cast<float>(static_cast<double>(_turn_cells) * sin(static_cast<double>(angle)));
    const float dy = static_cast<float>(static_cast<double>(_turn_cells) * (1.0 - cos(static_cast<double>(angle))));
    _motions = {{hypotf(dx, dy), 0.0f, 0.0f}, {dx, dy, increments}, {dx, -dy, -increments}};
    if (_p.motion_type == 2)
    {
        _motions.push_back({-hypotf(dx, dy), 0.0f, 0.0f});
        _motions.push_back({-dx, dy, -increments});
        _motions.push_back({-dx, -dy, increments});
    }
    _curves.reset(new CurveSpace(_p.motion_type == 2 ? CurveSpace::REEDS_SHEPP : CurveSpace::DUBINS, _turn_cells));
}

void HybridAStar::wrapHeading(float &h) const
{  // motion_hybrid.cpp:56-62
    while (h >= _n_theta_f)
        h -= _n_theta_f;
    while (h < 0.0)
        h += _n_theta_f;
}

float HybridAStar::headingFraction(float start_theta) const
{  // the heading coordinate setStart stores (astar_hybrid.cpp:49-53); motions add whole bins to it
    float m = start_theta / _bin + _n_theta_f;
    wrapHeading(m);
    return m - std::floor(m);
}

bool HybridAStar::goalCell(const Pose &goal_world, unsigned int &gx, unsigned int &gy) const
{
    if (goal_world.x < _ox || goal_world.y < _oy)
        return false;
    const float x = (goal_world.x - static_cast<float>(_ox)) / static_cast<float>(_res);
    const float y = (goal_world.y - static_cast<float>(_oy)) / static_cast<float>(_res);
    gx = static_cast<unsigned int>(x);  // what cacheObstacleHeuristic truncates (:109-110)
    gy = static_cast<unsigned int>(y);
    return gx < _W && gy < _H;
}

HybridAStar::Node *HybridAStar::node(unsigned int index)
{
    const unsigned int key = index + 1;
    unsigned int slot = (index * 2654435761u) & _slot_mask;  // Fibonacci hashing; linear probing
    while (_slot_key[slot] != 0)
    {
        if (_slot_key[slot] == key)
            return &_nodes[_slot_node[slot]];
        slot = (slot + 1) & _slot_mask;
    }
    if ((_nodes.size() + 1) * 2 > _slot_key.size())
    {
        growGraph();
        return node(index);
    }
    _slot_key[slot] = key;
    _slot_node[slot] = static_cast<unsigned int>(_nodes.size());
    _nodes.emplace_back(index);
    return &_nodes.back();
}

void HybridAStar::resetGraph(unsigned int expected_nodes)
{
    unsigned int slots = 1024;
    while (slots < 2 * expected_nodes)
        slots *= 2;
    _nodes.clear();
    _slot_key.assign(slots, 0u);
    _slot_node.assign(slots, 0u);
    _slot_mask = slots - 1;
}

void HybridAStar::growGraph()
{
    const unsigned int slots = static_cast<unsigned int>(_slot_key.size()) * 2;
    _slot_key.assign(slots, 0u);
    _slot_node.assign(slots, 0u);
    _slot_mask = slots - 1;
    for (unsigned int i = 0; i < _nodes.size(); ++i)
    {
        unsigned int slot = (_nodes[i].index * 2654435761u) & _slot_mask;
        while (_slot_key[slot] != 0)
            slot = (slot + 1) & _slot_mask;
        _slot_key[slot] = _nodes[i].index + 1;
        _slot_node[slot] = i;
    }
}

// Volume lookup in place of collision_checker.cpp:126-172: cell of the pose, nearest heading bin.
// Returns the collision flag; poses outside the map count as collisions (the search never asks).
bool HybridAStar::footprint(const Tables &tb, float x, float y, float theta_rad, float &cost, uint8_t kind) const
{
    const float fx = std::floor(x), fy = std::floor(y);
    if (!(fx >= 0.0f && fy >= 0.0f && fx < static_cast<float>(_W) && fy < static_cast<float>(_H)))
    {
        cost = 254.0f;
        return true;
    }
    ++_stats.lookups;
    if (tb.band_ready)
    {
        const unsigned int band = static_cast<unsigned int>(fy) >> tb.band_shift;
        if (!tb.band_ready[band])
            tb.fetch_band(tb.ctx, band);
    }
    long t = std::lround(theta_rad / _bin - tb.heading_fraction);
    t %= static_cast<long>(_n_theta);
    if (t < 0)
        t += _n_theta;
    const uint8_t v = tb.volume[(static_cast<size_t>(t) * _H + static_cast<size_t>(fy)) * _W + static_cast<size_t>(fx)];
    cost = static_cast<float>(v);
    if (_log)
        _log->push_back({x, y, theta_rad, kind, static_cast<uint8_t>(v >= 253), v});
    return v >= 253;
}

// collision_checker.cpp:126-142 at a continuous pose, on the host copy of the costmap: the same float
// expression, one rounded operation at a time (this file is compiled without contraction), cosf / sinf of
// the host libm like the reference build; a sample outside the map is skipped (:134). The float -> unsigned
// conversion of a negative floor() wraps to a huge index there and fails the bounds test: a sign test here.
bool HybridAStar::exactCollisionHost(const Tables &tb, float x, float y, float theta_rad) const
{
    const float c = cosf(theta_rad), s = sinf(theta_rad);
    for (unsigned int p = 0; p < tb.n_fp; ++p)
    {
        const float px = static_cast<float>(tb.fp[2 * p]), py = static_cast<float>(tb.fp[2 * p + 1]);
        const float a = px * c, b = py * s, u = x + a, xn = u - b;
        const float a2 = px * s, b2 = py * c, u2 = y + a2, yn = u2 + b2;
        const float fx = std::floor(xn), fy = std::floor(yn);
        if (fx >= 0.0f && fy >= 0.0f && fx < static_cast<float>(_W) && fy < static_cast<float>(_H) &&
            tb.cost[static_cast<size_t>(fy) * _W + static_cast<size_t>(fx)] >= 253)
            return true;
    }
    return false;
}

// astar_hybrid.cpp:242-247 with :82-92 (motion heuristic) and the table hit of :166-183
float HybridAStar::heuristic(const Tables &tb, const Pose &at) const
{
    const Pose &goal = _goal->at;
    // every node heading of a search is start fraction + whole bins: a few dozen distinct floats, so the
    // wrapped yaw and the trigonometry of (node heading, goal heading) are remembered per heading value
    uint32_t bits;
    std::memcpy(&bits, &at.theta, sizeof bits);
    HeadingEntry &e = _heading_cache[static_cast<unsigned int>(at.theta) % _heading_cache.size()];
    if (e.bits != bits)
    {
        e.bits = bits;
        e.yaw = wrap_angle_d(static_cast<double>(at.theta * _bin));
        e.trig = CurveSpace::headingTrig(e.yaw, _goal_se2.yaw);
    }
    const SE2 from{static_cast<double>(at.x), static_cast<double>(at.y), e.yaw};
    const float motion_h = static_cast<float>(_curves->distance(from, _goal_se2, e.trig));

    const unsigned int sx = static_cast<unsigned int>(at.x), sy = static_cast<unsigned int>(at.y);
    const float field = tb.field[static_cast<size_t>(sy) * _W + sx];
    const int ax = std::abs(static_cast<int>(at.x) - static_cast<int>(goal.x));
    const int ay = std::abs(static_cast<int>(at.y) - static_cast<int>(goal.y));
    const float diagonal = std::min(ax, ay) * sqrtf(2.0f) + std::abs(ax - ay);
    const float obstacle_h = std::max(field, diagonal);
    return _neutral * std::max(motion_h, obstacle_h);
}

// astar_hybrid.cpp:368-390
float HybridAStar::travelCost(const Node *from, const Node *to) const
{
    const float normalized = static_cast<float>(to->cell_cost / 252.0);
    if (from->motion == std::numeric_limits<unsigned int>::max())
        return _neutral;
    const float raw = _neutral + _p.cost_penalty * normalized;
    float cost;
    if (to->motion == 0 || to->motion == 3)
        cost = raw;
    else if (from->motion == to->motion)
        cost = raw * _p.non_straight_penalty;
    else
        cost = raw * (_p.change_penalty + _p.non_straight_penalty);
    if (to->motion > 2)
        cost *= _p.reverse_penalty;
    return cost;
}

// astar_hybrid.cpp:264-291 with MotionHybrid::getProjection (motion_hybrid.cpp:63-74)
void HybridAStar::expand(const Tables &tb, Node *n, std::vector<Node *> &out)
{
    out.clear();
    const float heading = n->at.theta;
    const float c = static_cast<float>(cos(static_cast<double>(heading * _bin)));
    const float s = static_cast<float>(sin(static_cast<double>(heading * _bin)));
    for (unsigned int i = 0; i < _motions.size(); ++i)
    {
        const Pose &m = _motions[i];
        const float ddx = m.x * c - m.y * s;
        const float ddy = m.x * s + m.y * c;
        float nh = heading + m.theta;
        wrapHeading(nh);
        const Pose next{ddx + n->at.x, ddy + n->at.y, nh};
        const unsigned int mx = static_cast<unsigned int>(next.x), my = static_cast<unsigned int>(next.y);
        const unsigned int mt = static_cast<unsigned int>(next.theta);
        if (mx <= 0 || mx >= _W - 1 || my <= 0 || my >= _H - 1)
            continue;
        Node *nb = node(indexOf(mx, my, mt));
        if (nb->visited || blocked(nb->index))
            continue;
        float cost;
        if (footprint(tb, next.x, next.y, next.theta * _bin, cost))
            continue;
        nb->cell_cost = cost;
        nb->at = next;
        nb->motion = i;
        out.push_back(nb);
    }
}

// astar_hybrid.cpp:293-350
HybridAStar::Node *HybridAStar::analyticExpansion(const Tables &tb, Node *n)
{
    const SE2 from{static_cast<double>(n->at.x), static_cast<double>(n->at.y), wrap_angle_d(static_cast<double>(n->at.theta * _bin))};
    const SE2 to{static_cast<double>(_goal->at.x), static_cast<double>(_goal->at.y),
                 wrap_angle_d(static_cast<double>(_goal->at.theta * _bin))};
    const float distance = static_cast<float>(_curves->distance(from, to));
    const float intervals = std::floor(distance / sqrtf(2.0f));
    struct Step
    {
        unsigned int index;
        Pose at;
    };
    ++_stats.analytic_attempts;
    const bool exact = tb.exact_analytic && tb.cost && tb.fp && tb.confirm_batch;
    std::vector<Step> steps;
    std::vector<float> samples;  // exact mode: x, y, yaw of every sample that still needs an exact answer
    for (float i = 1; i < intervals; i += 1.0f)
    {
        const SE2 q = _curves->interpolate(from, to, i / intervals);
        float heading = static_cast<float>(q.yaw) / _bin + _n_theta_f;
        wrapHeading(heading);
        const unsigned int xi = static_cast<unsigned int>(q.x), yi = static_cast<unsigned int>(q.y);
        if (xi <= 0 || xi >= _W - 1 || yi <= 0 || yi >= _H - 1)
            return nullptr;
        const float sx = static_cast<float>(q.x), sy = static_cast<float>(q.y), syaw = static_cast<float>(q.yaw);
        float ignored;
        if (footprint(tb, sx, sy, syaw, ignored, 1))
        {
            if (!exact)
                return nullptr;
            // the volume speaks for the cell centre and the nearest bin: ask the continuous pose
            ++_stats.exact_host_checks;
            if (exactCollisionHost(tb, sx, sy, syaw))
                return nullptr;
            ++_stats.false_collisions_overruled;
        }
        else if (exact)
        {
            samples.push_back(sx);
            samples.push_back(sy);
            samples.push_back(syaw);
        }
        if (blocked(indexOf(xi, yi, static_cast<unsigned int>(heading))))
            return nullptr;  // a pose of an earlier path that failed the exact check
        steps.push_back({indexOf(xi, yi, static_cast<unsigned int>(heading)), {sx, sy, heading}});
    }
    if (exact && !samples.empty())
    {
        // every sample passed the filter: one batched exact evaluation decides
        std::vector<uint8_t> flags(samples.size() / 3);
        ++_stats.analytic_confirm_batches;
        if (!tb.confirm_batch(tb.confirm_ctx, samples.data(), static_cast<unsigned int>(flags.size()), flags.data()))
            return nullptr;  // a failed query fails closed
        unsigned int hits = 0;
        for (uint8_t f : flags)
            hits += f != 0;
        _stats.false_free_caught += hits;
        if (hits)
            return nullptr;
    }
    if (steps.empty())
        return nullptr;
    Node *prev = n;
    for (const Step &st : steps)
    {
        Node *nb = node(st.index);
        if (nb->visited)
            continue;
        nb->parent = prev;
        nb->visited = true;
        nb->at = st.at;
        prev = nb;
    }
    if (prev != _goal)
    {
        _goal->parent = prev;
        _goal->visited = true;
    }
    return _goal;
}

// astar_hybrid.cpp:248-263
bool HybridAStar::backtrace(Node *n, std::vector<Pose> &path)
{
    if (!n->parent)
        return false;
    Node *cur = n;
    _path_nodes.clear();
    while (cur->parent)
    {
        path.push_back(cur->at);
        _path_nodes.push_back(cur->index);
        cur = cur->parent;
    }
    path.push_back(cur->at);
    _path_nodes.push_back(cur->index);
    return path.size() > 1;
}

bool HybridAStar::blocked(unsigned int index) const
{
    return !_blocked.empty() && std::find(_blocked.begin(), _blocked.end(), index) != _blocked.end();
}

bool HybridAStar::plan(const Tables &tb, const Pose &start_world, const Pose &goal_world, std::vector<Pose> &path, int &iterations)
{
    path.clear();
    _path_cells.clear();
    _path_nodes.clear();
    _stats = PlanStats();
    iterations = 0;
    // astar_hybrid.cpp:392-433: world -> cells (costmap_2d::worldToMap rejects poses outside the map)
    auto inside = [&](const Pose &p) {
        if (p.x < _ox || p.y < _oy)
            return false;
        const unsigned int mx = static_cast<unsigned int>(static_cast<int>((p.x - _ox) / _res));
        const unsigned int my = static_cast<unsigned int>(static_cast<int>((p.y - _oy) / _res));
        return mx < _W && my < _H;
    };
    if (!inside(start_world) || !inside(goal_world))
        return false;
    const float ox = static_cast<float>(_ox), oy = static_cast<float>(_oy), res = static_cast<float>(_res);
    resetGraph(32768);
    auto place = [&](const Pose &w) {  // setStart / setGoal, :45-66
        const float x = (w.x - ox) / res, y = (w.y - oy) / res;
        float h = w.theta / _bin + _n_theta_f;
        wrapHeading(h);
        Node *nd = node(indexOf(static_cast<unsigned int>(std::floor(x)), static_cast<unsigned int>(std::floor(y)),
                                static_cast<unsigned int>(std::floor(h))));
        nd->at = {x, y, h};
        return nd;
    };
    _start = place(start_world);
    _goal = place(goal_world);
    _heading_cache.assign(_n_theta + 1, HeadingEntry());  // keyed to this goal's heading
    _goal_se2 = {static_cast<double>(_goal->at.x), static_cast<double>(_goal->at.y),
                 wrap_angle_d(static_cast<double>(_goal->at.theta * _bin))};  // the same for every heuristic call

    // cacheObstacleHeuristic() succeeds iff the start cell is reached from the goal (:161-164)
    const unsigned int scx = static_cast<unsigned int>(_start->at.x), scy = static_cast<unsigned int>(_start->at.y);
    if (!(tb.field[static_cast<size_t>(scy) * _W + scx] > 0.0f))
        return false;

    // createPath, :434-515
    std::priority_queue<QItem, std::vector<QItem>, Later> queue;
    int approach_iterations = 0, analytic_iterations = 0, closest = INT_MAX;
    const float tolerance = _tol_cells * _neutral;
    std::pair<float, unsigned int> best{std::numeric_limits<float>::max(), 0};
    {  // isInputValid, :32-44
        float c;
        if (footprint(tb, _start->at.x, _start->at.y, _start->at.theta * _bin, c))
            return false;
        _start->cell_cost = c;
        if (footprint(tb, _goal->at.x, _goal->at.y, _goal->at.theta * _bin, c))
            return false;
        _goal->cell_cost = c;
    }
    queue.emplace(0.0f, _start);
    _start->g = 0.0f;
    std::vector<Node *> neighbours;
    bool found = false;
    while (iterations < static_cast<int>(_p.max_iterations) && !queue.empty())
    {
        Node *cur = queue.top().second;
        queue.pop();
        if (cur->visited)
            continue;
        ++iterations;
        cur->visited = true;
        if (best.first < _ae_tol_cells)
        {  // tryAnalyticExpansion, :352-366
            closest = std::min(static_cast<int>(heuristic(tb, cur->at) / _neutral), closest);
            const int desired = std::max(static_cast<int>(closest / _ae_ratio_cells), static_cast<int>(std::ceil(_ae_ratio_cells)));
            analytic_iterations = std::min(analytic_iterations, desired);
            if (analytic_iterations <= 0)
            {
                analytic_iterations = desired;
                if (Node *hit = analyticExpansion(tb, cur))
                    cur = hit;
            }
            else
                --analytic_iterations;
        }
        if (cur == _goal)
        {
            found = backtrace(cur, path);
            break;
        }
        if (best.first < tolerance)
        {
            ++approach_iterations;
            if (approach_iterations > static_cast<int>(_p.max_on_approach_iterations) ||
                iterations + 1 > static_cast<int>(_p.max_iterations))
            {
                _goal->parent = node(best.second);  // exists: it was queued
                found = backtrace(_goal, path);
                break;
            }
        }
        expand(tb, cur, neighbours);
        for (Node *nb : neighbours)
        {
            const float g = cur->g + travelCost(cur, nb);
            if (g < nb->g)
            {
                nb->g = g;
                nb->parent = cur;
                const float h = heuristic(tb, nb->at);
                if (h < best.first)
                    best = {h, nb->index};
                queue.emplace(g + h, nb);
            }
        }
    }
    if (!found)
    {
        path.clear();
        return false;
    }
    for (const Pose &p : path)
        _path_cells.push_back({p.x, p.y, p.theta * _bin});
    // fromCellToWorld, :516-528, then the two end poses are replaced by the requested ones (:428-429)
    for (Pose &p : path)
    {
        const double wx = _ox + (static_cast<unsigned int>(p.x) + 0.5) * _res;
        const double wy = _oy + (static_cast<unsigned int>(p.y) + 0.5) * _res;
        p = {static_cast<float>(wx), static_cast<float>(wy), wrap_angle_f(p.theta * _bin)};
    }
    path.front() = goal_world;
    path.back() = start_world;
    return true;
}
}  // namespace ppb
