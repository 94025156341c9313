// TEST INFRASTRUCTURE ONLY. Stand-in for g2o::BaseVertex (g2o is not installed here): the members
// the reference's Vertex2E / VertexTimeDiff (teb_path_follower/commons.h:124-180) use, plus the
// push/pop/oplus the numeric Jacobian needs.
#pragma once
#include <iostream>
#include <vector>

namespace g2o
{
template <int D, typename T>
class BaseVertex
{
public:
    static const int Dimension = D;
    virtual ~BaseVertex() {}
    const T &estimate() const { return _estimate; }
    void setEstimate(const T &e) { _estimate = e; }
    void setFixed(bool f) { _fixed = f; }
    bool fixed() const { return _fixed; }
    void push() { _backup.push_back(_estimate); }
    void pop()
    {
        _estimate = _backup.back();
        _backup.pop_back();
    }
    void oplus(const double *v) { oplusImpl(v); }
    virtual void setToOriginImpl() = 0;
    virtual void oplusImpl(const double *update) = 0;
    virtual bool read(std::istream &is) = 0;
    virtual bool write(std::ostream &os) const = 0;

protected:
    T _estimate;
    bool _fixed = false;
    std::vector<T> _backup;
};
}  // namespace g2o
