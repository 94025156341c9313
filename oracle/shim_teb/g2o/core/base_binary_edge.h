// TEST INFRASTRUCTURE ONLY. base_edges.hpp names g2o::BaseBinaryEdge; nothing on the obstacle-edge
// path instantiates it.
#pragma once
#include <iostream>
namespace g2o
{
template <int D, typename E, typename V1, typename V2>
class BaseBinaryEdge
{
public:
    virtual ~BaseBinaryEdge() {}
    virtual bool read(std::istream &is) = 0;
    virtual bool write(std::ostream &os) const = 0;
};
}  // namespace g2o
