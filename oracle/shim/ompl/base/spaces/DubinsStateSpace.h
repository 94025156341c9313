// TEST INFRASTRUCTURE ONLY (oracle). From-scratch restatement of shortest Dubins paths
// (Dubins 1957; closed forms as tabulated by Shkel & Lumelsky 2001) behind the OMPL class name
// the reference constructs at motion_hybrid.cpp:35. Six words LSL RSR RSL LSR RLR LRL are
// evaluated in normalised coordinates (turning radius 1) and the shortest kept.
// PARITY UNPINNED: OMPL itself is absent (see StateSpace.h).
#pragma once
#include "../StateSpace.h"

namespace ompl
{
namespace base
{
class DubinsStateSpace : public StateSpace
{
public:
    enum Seg { L = 0, S = 1, R = 2 };
    struct Path
    {
        Seg type[3] = {L, S, L};
        double len[3] = {0.0, std::numeric_limits<double>::max(), 0.0};
        double length() const { return len[0] + len[1] + len[2]; }
    };

    explicit DubinsStateSpace(double turningRadius = 1.0) : rho_(turningRadius) {}

    double distance(const State *a, const State *b) const override { return rho_ * solve(a, b).length(); }

    void interpolate(const State *from, const State *to, double t, State *out) const override
    {
        if (t >= 1.0) { *out = *to; return; }
        if (t <= 0.0) { *out = *from; return; }
        Path p = solve(from, to);
        double seg = t * p.length();
        double x = 0.0, y = 0.0, yaw = from->getYaw();
        for (int i = 0; i < 3 && seg > 0.0; ++i)
        {
            double v = std::min(seg, p.len[i]);
            double phi = yaw;
            seg -= v;
            if (p.type[i] == L)
            {
                x = x + sin(phi + v) - sin(phi);
                y = y - cos(phi + v) + cos(phi);
                yaw = phi + v;
            }
            else if (p.type[i] == R)
            {
                x = x - sin(phi - v) + sin(phi);
                y = y + cos(phi - v) - cos(phi);
                yaw = phi - v;
            }
            else
            {
                x = x + v * cos(phi);
                y = y + v * sin(phi);
            }
        }
        out->v[0] = x * rho_ + from->getX();
        out->v[1] = y * rho_ + from->getY();
        out->v[2] = pp_wrap_yaw(yaw);
    }

    Path solve(const State *a, const State *b) const
    {
        double dx = b->getX() - a->getX(), dy = b->getY() - a->getY();
        double d = sqrt(dx * dx + dy * dy) / rho_;
        double th = atan2(dy, dx);
        return solveNormalised(d, wrap2pi(a->getYaw() - th), wrap2pi(b->getYaw() - th));
    }

private:
    static constexpr double kTwoPi = 2.0 * 3.14159265358979323846;
    static constexpr double kEps = 1e-6;
    static constexpr double kZero = -1e-7;

    static double wrap2pi(double x)
    {
        if (x < 0 && x > kZero)
            return 0;
        double xm = x - kTwoPi * floor(x / kTwoPi);
        if (kTwoPi - xm < 0.5 * kEps)
            xm = 0.0;
        return xm;
    }
    static Path make(Seg a, Seg b, Seg c, double t, double p, double q)
    {
        Path r;
        r.type[0] = a; r.type[1] = b; r.type[2] = c;
        r.len[0] = t; r.len[1] = p; r.len[2] = q;
        return r;
    }
    static void keepShorter(Path &best, double &bestLen, const Path &cand)
    {
        double l = cand.length();
        if (l < bestLen) { bestLen = l; best = cand; }
    }
    static Path solveNormalised(double d, double alpha, double beta)
    {
        if (d < kEps && fabs(alpha - beta) < kEps)
            return make(L, S, L, 0.0, d, 0.0);
        const double ca = cos(alpha), sa = sin(alpha), cb = cos(beta), sb = sin(beta);
        const double cab = ca * cb + sa * sb;
        Path best, cand;
        double tmp;
        // LSL
        tmp = 2.0 + d * d - 2.0 * (cab - d * (sa - sb));
        if (tmp >= kZero)
        {
            double theta = atan2(cb - ca, d + sa - sb);
            best = make(L, S, L, wrap2pi(-alpha + theta), sqrt(std::max(tmp, 0.0)), wrap2pi(beta - theta));
        }
        double bestLen = best.length();
        // RSR
        tmp = 2.0 + d * d - 2.0 * (cab - d * (sb - sa));
        if (tmp >= kZero)
        {
            double theta = atan2(ca - cb, d - sa + sb);
            keepShorter(best, bestLen, make(R, S, R, wrap2pi(alpha - theta), sqrt(std::max(tmp, 0.0)), wrap2pi(-beta + theta)));
        }
        // RSL
        tmp = d * d - 2.0 + 2.0 * (cab - d * (sa + sb));
        if (tmp >= kZero)
        {
            double p = sqrt(std::max(tmp, 0.0));
            double theta = atan2(ca + cb, d - sa - sb) - atan2(2.0, p);
            keepShorter(best, bestLen, make(R, S, L, wrap2pi(alpha - theta), p, wrap2pi(beta - theta)));
        }
        // LSR
        tmp = -2.0 + d * d + 2.0 * (cab + d * (sa + sb));
        if (tmp >= kZero)
        {
            double p = sqrt(std::max(tmp, 0.0));
            double theta = atan2(-ca - cb, d + sa + sb) - atan2(-2.0, p);
            keepShorter(best, bestLen, make(L, S, R, wrap2pi(-alpha + theta), p, wrap2pi(-beta + theta)));
        }
        // RLR
        tmp = 0.125 * (6.0 - d * d + 2.0 * (cab + d * (sa - sb)));
        if (fabs(tmp) < 1.0)
        {
            double p = kTwoPi - acos(tmp);
            double theta = atan2(ca - cb, d - sa + sb);
            double t = wrap2pi(alpha - theta + 0.5 * p);
            keepShorter(best, bestLen, make(R, L, R, t, p, wrap2pi(alpha - beta - t + p)));
        }
        // LRL
        tmp = 0.125 * (6.0 - d * d + 2.0 * (cab - d * (sa - sb)));
        if (fabs(tmp) < 1.0)
        {
            double p = kTwoPi - acos(tmp);
            double theta = atan2(-ca + cb, d + sa - sb);
            double t = wrap2pi(-alpha + theta + 0.5 * p);
            keepShorter(best, bestLen, make(L, R, L, t, p, wrap2pi(beta - alpha - t + p)));
        }
        return best;
    }

    double rho_;
};
}
}
