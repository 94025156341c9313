// TEST INFRASTRUCTURE ONLY (oracle). See StateSpace.h.
#pragma once
#include "StateSpace.h"
namespace ompl
{
namespace base
{
template <class T = StateSpace>
class ScopedState
{
public:
    explicit ScopedState(const StateSpacePtr &space) : _space(space) {}
    double &operator[](unsigned int i) { return _s.v[i]; }
    State *operator()() { return &_s; }
    std::vector<double> reals() const { return {_s.v[0], _s.v[1], _s.v[2]}; }

private:
    StateSpacePtr _space;
    State _s;
};
}
}
