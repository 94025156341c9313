// TEST INFRASTRUCTURE ONLY (oracle). Minimal stand-in for ROS costmap_2d::Costmap2D so that the
// reference's own planner sources compile unmodified from /root/reference. Only the accessors
// the reference calls on the hot path are restated:
//   getCost / getIndex        : row-major my*size_x+mx
//   mapToWorld                : origin + (m + 0.5) * resolution
//   worldToMap                : truncating, false when outside
// The real header transitively brings in <math.h>, <iostream>, <cassert>, <algorithm>, <limits>
// via ROS message headers; <math.h> is decisive because it makes unqualified cos/sin/floor on
// float operands resolve to the float overloads. Define
// PP_ORACLE_DOUBLE_TRIG to drop it and get the double-precision resolution instead.
#pragma once
#ifndef PP_ORACLE_DOUBLE_TRIG
#include <math.h>
#endif
#include <cmath>
#include <cassert>
#include <iostream>
#include <algorithm>
#include <limits>
#include <vector>
#include <cstring>
#include <cstdint>

namespace costmap_2d
{
class Costmap2D
{
public:
    Costmap2D(unsigned int cells_size_x, unsigned int cells_size_y, double resolution,
              double origin_x, double origin_y, unsigned char default_value = 0)
        : size_x_(cells_size_x), size_y_(cells_size_y), resolution_(resolution),
          origin_x_(origin_x), origin_y_(origin_y),
          costmap_(static_cast<size_t>(cells_size_x) * cells_size_y, default_value) {}

    unsigned char getCost(unsigned int mx, unsigned int my) const { return costmap_[getIndex(mx, my)]; }
    void setCost(unsigned int mx, unsigned int my, unsigned char cost) { costmap_[getIndex(mx, my)] = cost; }
    inline unsigned int getIndex(unsigned int mx, unsigned int my) const { return my * size_x_ + mx; }
    void mapToWorld(unsigned int mx, unsigned int my, double &wx, double &wy) const
    {
        wx = origin_x_ + (mx + 0.5) * resolution_;
        wy = origin_y_ + (my + 0.5) * resolution_;
    }
    bool worldToMap(double wx, double wy, unsigned int &mx, unsigned int &my) const
    {
        if (wx < origin_x_ || wy < origin_y_)
            return false;
        mx = (int)((wx - origin_x_) / resolution_);
        my = (int)((wy - origin_y_) / resolution_);
        return mx < size_x_ && my < size_y_;
    }
    unsigned char *getCharMap() { return costmap_.data(); }
    const unsigned char *getCharMap() const { return costmap_.data(); }
    unsigned int getSizeInCellsX() const { return size_x_; }
    unsigned int getSizeInCellsY() const { return size_y_; }
    double getResolution() const { return resolution_; }
    double getOriginX() const { return origin_x_; }
    double getOriginY() const { return origin_y_; }

private:
    unsigned int size_x_, size_y_;
    double resolution_, origin_x_, origin_y_;
    std::vector<unsigned char> costmap_;
};
}
