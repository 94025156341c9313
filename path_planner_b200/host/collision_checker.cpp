// See collision_checker.h. Reference lines cited are in
// src/path_planner/src/planner/collision_checker.cpp.
#include "collision_checker.h"

#include "pp_b200.h"

#include <cmath>
#include <cstdio>
#include <iostream>

CollisionChecker::CollisionChecker()
    : _costmap(nullptr), _hole(true), _h(nullptr), _uploaded(false), _footprint_dirty(true), _exact(false), _volume(nullptr),
      _n_theta(0), _heading_fraction(0.0f), _bin_size(0.0f), _lookups(0), _exact_calls(0)
{
    if (pp_create(&_h, 0) != PP_OK)
        _h = nullptr;  // every later GPU-backed call reports failure; there is no CPU fallback
}

CollisionChecker::~CollisionChecker() { pp_destroy(_h); }

const char *CollisionChecker::lastError() const { return _h ? pp_last_error(_h) : "pp_create failed: no CUDA device"; }

void CollisionChecker::setCostMap(costmap_2d::Costmap2D *costmap)
{
    _costmap = costmap;
    _uploaded = false;
    _volume = nullptr;
}

// :3-27
void CollisionChecker::initBoundary(float radius)
{
    const int r = static_cast<int>(std::ceil(radius / _costmap->getResolution()));
    const int inner = std::max(0, r - 1);
    for (int i = -r; i <= r; i++)
        for (int j = -r; j <= r; j++)
        {
            const int d = i * i + j * j;
            if (d <= r * r && (!_hole || d >= inner * inner))
            {
                _points.push_back(i);
                _points.push_back(j);
            }
        }
    _footprint_dirty = true;
}
// :28-48
void CollisionChecker::initBoundary(float width, float height)
{
    const int hw = static_cast<int>(std::ceil(0.5 * width / _costmap->getResolution()));
    const int hh = static_cast<int>(std::ceil(0.5 * height / _costmap->getResolution()));
    for (int i = -hw; i <= hw; i++)
        for (int j = -hh; j <= hh; j++)
            if (!_hole || i == -hw || i == hw || j == -hh || j == hh)
            {
                _points.push_back(i);
                _points.push_back(j);
            }
    _footprint_dirty = true;
}
// :49-71
void CollisionChecker::initBoundary(float x_min, float y_min, float x_max, float y_max)
{
    const int x0 = static_cast<int>(std::floor(x_min / _costmap->getResolution()));
    const int y0 = static_cast<int>(std::floor(y_min / _costmap->getResolution()));
    const int x1 = static_cast<int>(std::ceil(x_max / _costmap->getResolution()));
    const int y1 = static_cast<int>(std::ceil(y_max / _costmap->getResolution()));
    for (int i = x0; i <= x1; i++)
        for (int j = y0; j <= y1; j++)
            if (!_hole || i == x0 || i == x1 || j == y0 || j == y1)
            {
                _points.push_back(i);
                _points.push_back(j);
            }
    _footprint_dirty = true;
}
// :72-81
void CollisionChecker::initBoundary(std::vector<float> &boundary)
{
    if (boundary.size() == 1)
        initBoundary(boundary[0]);
    if (boundary.size() == 2)
        initBoundary(boundary[0], boundary[1]);
    if (boundary.size() == 4)
        initBoundary(boundary[0], boundary[1], boundary[2], boundary[3]);
}

// :83-119
void CollisionChecker::printBoundary()
{
    std::cout << "=======================" << std::endl;
    if (_points.empty())
        return;
    int x_min = _points[0], x_max = _points[0], y_min = _points[1], y_max = _points[1];
    for (size_t p = 0; p < _points.size(); p += 2)
    {
        x_min = std::min(x_min, _points[p]);
        x_max = std::max(x_max, _points[p]);
        y_min = std::min(y_min, _points[p + 1]);
        y_max = std::max(y_max, _points[p + 1]);
    }
    std::vector<std::string> rows(y_max - y_min + 1, std::string(2 * (x_max - x_min + 1), ' '));
    for (auto &r : rows)
        for (size_t c = 0; c < r.size(); c += 2)
            r[c] = '-';
    for (size_t p = 0; p < _points.size(); p += 2)
        rows[_points[p + 1] - y_min][2 * (_points[p] - x_min)] = '*';
    for (auto &r : rows)
        std::cout << r << std::endl;
    std::cout << "=======================" << std::endl;
}

bool CollisionChecker::ensureUploaded()
{
    if (!_h || !_costmap)
        return false;
    if (!_uploaded)
    {
        // the reference reads the live grid without a lock; take costmap->getMutex() around this
        // call in a ROS node (costmap_2d updates the grid on its own thread)
        if (pp_costmap_upload(_h, _costmap->getCharMap(), _costmap->getSizeInCellsX(), _costmap->getSizeInCellsY(), 0) != PP_OK)
            return false;
        _uploaded = true;
        _footprint_dirty = true;
    }
    if (_footprint_dirty)
    {
        if (_points.empty() || pp_footprint_set(_h, _points.data(), static_cast<uint32_t>(_points.size() / 2)) != PP_OK)
            return false;
        _footprint_dirty = false;
    }
    return true;
}

bool CollisionChecker::prepare(unsigned int n_theta, float heading_fraction)
{
    _uploaded = false;  // the grid may have changed since the last plan
    if (!ensureUploaded())
        return false;
    if (pp_se2_volume_build(_h, n_theta, 0.5f, 0.5f, heading_fraction, 0, n_theta) != PP_OK)
        return false;
    const size_t n = static_cast<size_t>(_costmap->getSizeInCellsX()) * _costmap->getSizeInCellsY() * n_theta;
    _volume_mirror.resize(n);
    if (pp_se2_volume_download(_h, 0, n_theta, _volume_mirror.data(), 0) != PP_OK)
        return false;
    attachVolume(_volume_mirror.data(), n_theta, heading_fraction);
    return true;
}

void CollisionChecker::attachVolume(const uint8_t *volume, unsigned int n_theta, float heading_fraction)
{
    _volume = volume;
    _n_theta = n_theta;
    _heading_fraction = heading_fraction;
    _bin_size = 2.0f * static_cast<float>(M_PI) / static_cast<float>(n_theta);  // motion_hybrid.cpp:10
}

bool CollisionChecker::buildHeuristicField(unsigned int goal_x, unsigned int goal_y, float cost_penalty, std::vector<float> &field)
{
    if (!ensureUploaded())
        return false;
    field.resize(static_cast<size_t>(_costmap->getSizeInCellsX()) * _costmap->getSizeInCellsY());
    return pp_heuristic_build(_h, goal_x, goal_y, cost_penalty, field.data()) == PP_OK;
}

bool CollisionChecker::lookup(float x, float y, float theta, float &cost, bool &flag)
{
    if (!_volume || _exact)
        return false;
    const float fx = std::floor(x), fy = std::floor(y);
    const unsigned int W = _costmap->getSizeInCellsX(), H = _costmap->getSizeInCellsY();
    if (!(fx >= 0.0f && fy >= 0.0f && fx < static_cast<float>(W) && fy < static_cast<float>(H)))
        return false;
    // nearest heading bin: callers pass (k + fraction) * bin_size (astar_hybrid.cpp:37,40,283)
    const float bins = theta / _bin_size - _heading_fraction;
    long t = std::lround(bins);
    t %= static_cast<long>(_n_theta);
    if (t < 0)
        t += _n_theta;
    const uint8_t v = _volume[(static_cast<size_t>(t) * H + static_cast<size_t>(fy)) * W + static_cast<size_t>(fx)];
    cost = static_cast<float>(v);
    flag = v >= costmap_2d::INSCRIBED_INFLATED_OBSTACLE;
    ++_lookups;
    return true;
}

bool CollisionChecker::exactEval(float x, float y, float theta, float &cost, bool &flag)
{
    if (!ensureUploaded())
        return false;
    const float pose[3] = {x, y, theta};
    uint8_t f = 0;
    if (pp_footprint_cost_batch(_h, pose, 1, &cost, &f) != PP_OK)
        return false;
    flag = f != 0;
    ++_exact_calls;
    return true;
}

// :126-142
bool CollisionChecker::checkFootPrintCollision(const float &x, const float &y, const float &theta)
{
    float cost = 0.0f;
    bool flag = false;
    const bool looked = lookup(x, y, theta, cost, flag);
    if (looked && _log)
        _log->push_back({x, y, theta, 1, static_cast<uint8_t>(flag), static_cast<uint8_t>(cost)});
    if ((!looked || _exact_analytic) && !exactEval(x, y, theta, cost, flag))
    {
        std::fprintf(stderr, "CollisionChecker: GPU query failed: %s\n", lastError());
        return true;  // fail closed: report a collision rather than let the search walk through
    }
    return flag;
}

// :144-172
bool CollisionChecker::checkAndGetFootprintCost(const float &x, const float &y, const float &theta, float &cost)
{
    bool flag = false;
    const bool looked = lookup(x, y, theta, cost, flag);
    if (looked && _log)
        _log->push_back({x, y, theta, 0, static_cast<uint8_t>(flag), static_cast<uint8_t>(cost)});
    if (!looked && !exactEval(x, y, theta, cost, flag))
    {
        std::fprintf(stderr, "CollisionChecker: GPU query failed: %s\n", lastError());
        cost = 254.0f;
        return true;
    }
    return flag;
}

// :174-214: point queries read the host grid exactly like the reference
bool CollisionChecker::isInCostMap(const float &x, const float &y)
{
    const unsigned int x_id = static_cast<unsigned int>(std::floor(x)), y_id = static_cast<unsigned int>(std::floor(y));
    return x_id < _costmap->getSizeInCellsX() && y_id < _costmap->getSizeInCellsY();
}

bool CollisionChecker::checkPointCollision(const float &x, const float &y)
{
    const unsigned int x_id = static_cast<unsigned int>(std::floor(x)), y_id = static_cast<unsigned int>(std::floor(y));
    if (x_id >= _costmap->getSizeInCellsX() || y_id >= _costmap->getSizeInCellsY())
        return false;
    return _costmap->getCost(x_id, y_id) >= costmap_2d::INSCRIBED_INFLATED_OBSTACLE;
}

bool CollisionChecker::checkAndGetPointCost(const float &x, const float &y, float &cost)
{
    const unsigned int x_id = static_cast<unsigned int>(std::floor(x)), y_id = static_cast<unsigned int>(std::floor(y));
    if (x_id >= _costmap->getSizeInCellsX() || y_id >= _costmap->getSizeInCellsY())
        return false;
    const unsigned char c = _costmap->getCost(x_id, y_id);
    cost = static_cast<float>(c);
    return c >= costmap_2d::INSCRIBED_INFLATED_OBSTACLE;
}

float CollisionChecker::getCost(const unsigned int &x, const unsigned int &y)
{
    if (x < _costmap->getSizeInCellsX() && y < _costmap->getSizeInCellsY())
        return static_cast<float>(_costmap->getCost(x, y));
    return 253.0f;
}
