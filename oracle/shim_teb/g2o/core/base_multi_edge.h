// TEST INFRASTRUCTURE ONLY. base_edges.hpp names g2o::BaseMultiEdge; nothing on the obstacle-edge
// path instantiates it.
#pragma once
#include <iostream>
namespace g2o
{
template <int D, typename E>
class BaseMultiEdge
{
public:
    virtual ~BaseMultiEdge() {}
    virtual bool read(std::istream &is) = 0;
    virtual bool write(std::ostream &os) const = 0;
};
}  // namespace g2o
