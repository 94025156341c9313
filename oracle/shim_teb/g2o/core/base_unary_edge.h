// TEST INFRASTRUCTURE ONLY. Stand-in for g2o::BaseUnaryEdge (g2o is not installed here; the
// reference installs an un-pinned g2o, README.md). What the reference's edges use of it, plus
// linearizeOplus() -- the central-difference Jacobian g2o computes for an edge that does not
// override it (EdgeObstacle does not), restated from g2o/core/base_unary_edge.hpp: step 1e-9 per
// vertex dimension through push / oplus / computeError / pop, column = (e(+d) - e(-d)) / (2d).
#pragma once
#include <Eigen/Core>
#include <algorithm>
#include <iostream>
#include <vector>

namespace g2o
{
template <int D, typename E, typename VertexXi>
class BaseUnaryEdge
{
public:
    virtual ~BaseUnaryEdge() {}
    void setMeasurement(const E &m) { _measurement = m; }
    const E &measurement() const { return _measurement; }
    void setVertex(size_t, VertexXi *v) { _vertices[0] = v; }
    template <class I>
    void setInformation(const I &)
    {
    }
    virtual void computeError() = 0;
    virtual bool read(std::istream &is) = 0;
    virtual bool write(std::ostream &os) const = 0;
    const Eigen::FixedVector<D> &error() const { return _error; }
    // row-major D x VertexXi::Dimension
    const double *jacobianOplusXi() const { return _jacobian; }

    virtual void linearizeOplus()
    {
        VertexXi *vi = static_cast<VertexXi *>(_vertices[0]);
        if (vi->fixed())
            return;
        const double delta = 1e-9;
        const double scalar = 1 / (2 * delta);
        Eigen::FixedVector<D> error1;
        Eigen::FixedVector<D> errorBeforeNumeric = _error;
        double add_vi[VertexXi::Dimension];
        std::fill(add_vi, add_vi + VertexXi::Dimension, 0.0);
        for (int d = 0; d < VertexXi::Dimension; ++d)
        {
            vi->push();
            add_vi[d] = delta;
            vi->oplus(add_vi);
            computeError();
            error1 = _error;
            vi->pop();
            vi->push();
            add_vi[d] = -delta;
            vi->oplus(add_vi);
            computeError();
            vi->pop();
            add_vi[d] = 0.0;
            for (int r = 0; r < D; ++r)
                _jacobian[r * VertexXi::Dimension + d] = scalar * (error1[r] - _error[r]);
        }
        _error = errorBeforeNumeric;
    }

protected:
    E _measurement;
    Eigen::FixedVector<D> _error;
    VertexXi *_vertices[1] = {nullptr};
    double _jacobian[D * VertexXi::Dimension] = {};
};
}  // namespace g2o
