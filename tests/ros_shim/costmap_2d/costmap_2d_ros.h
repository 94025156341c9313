// TEST INFRASTRUCTURE ONLY. Stand-in for <costmap_2d/costmap_2d_ros.h>: a grid with the accessors a
// global planner plugin uses, and the wrapper that hands it out.
#pragma once
#include <mutex>
#include <string>
#include <vector>

namespace costmap_2d
{
class Costmap2D
{
public:
    typedef std::recursive_mutex mutex_t;
    Costmap2D(unsigned int cells_size_x, unsigned int cells_size_y, double resolution, double origin_x, double origin_y)
        : size_x_(cells_size_x), size_y_(cells_size_y), resolution_(resolution), origin_x_(origin_x), origin_y_(origin_y),
          costmap_(static_cast<size_t>(cells_size_x) * cells_size_y, 0)
    {
    }
    unsigned char *getCharMap() { return costmap_.data(); }
    unsigned int getSizeInCellsX() const { return size_x_; }
    unsigned int getSizeInCellsY() const { return size_y_; }
    double getResolution() const { return resolution_; }
    double getOriginX() const { return origin_x_; }
    double getOriginY() const { return origin_y_; }
    mutex_t *getMutex() { return &access_; }

private:
    unsigned int size_x_, size_y_;
    double resolution_, origin_x_, origin_y_;
    std::vector<unsigned char> costmap_;
    mutex_t access_;
};

class Costmap2DROS
{
public:
    Costmap2DROS(Costmap2D *costmap, const std::string &global_frame) : costmap_(costmap), global_frame_(global_frame) {}
    Costmap2D *getCostmap() { return costmap_; }
    std::string getGlobalFrameID() { return global_frame_; }

private:
    Costmap2D *costmap_;
    std::string global_frame_;
};
}  // namespace costmap_2d
