// TEST INFRASTRUCTURE ONLY (oracle). From-scratch restatement of shortest Reeds-Shepp paths
// (Reeds & Shepp 1990, formulas 8.1-8.11, with the two well-known typo corrections) behind the
// OMPL class name the reference constructs at motion_hybrid.cpp:54. Each base formula is tried
// under the four symmetries (identity, time-flip, reflection, both) and, for the asymmetric
// families, on the reversed problem; candidates are visited family by family in the order
// CSC, CCC, CCCC, CCSC, CCSCC and a candidate replaces the incumbent only if strictly shorter.
// PARITY UNPINNED: OMPL itself is absent (see StateSpace.h).
#pragma once
#include "../StateSpace.h"

namespace ompl
{
namespace base
{
class ReedsSheppStateSpace : public StateSpace
{
public:
    enum Seg { NOP = 0, L = 1, S = 2, R = 3 };
    struct Path
    {
        Seg type[5] = {L, R, L, NOP, NOP};
        double len[5] = {std::numeric_limits<double>::max(), 0.0, 0.0, 0.0, 0.0};
        double total = std::numeric_limits<double>::max();
        double length() const { return total; }
    };

    explicit ReedsSheppStateSpace(double turningRadius = 1.0) : rho_(turningRadius) {}

    double distance(const State *a, const State *b) const override { return rho_ * solve(a, b).length(); }

    void interpolate(const State *from, const State *to, double t, State *out) const override
    {
        if (t >= 1.0) { *out = *to; return; }
        if (t <= 0.0) { *out = *from; return; }
        Path p = solve(from, to);
        double seg = t * p.length();
        double x = 0.0, y = 0.0, yaw = from->getYaw();
        for (int i = 0; i < 5 && seg > 0.0; ++i)
        {
            double v;
            if (p.len[i] < 0)
            {
                v = std::max(-seg, p.len[i]);
                seg += v;
            }
            else
            {
                v = std::min(seg, p.len[i]);
                seg -= v;
            }
            double phi = yaw;
            switch (p.type[i])
            {
            case L:
                x = x + sin(phi + v) - sin(phi);
                y = y - cos(phi + v) + cos(phi);
                yaw = phi + v;
                break;
            case R:
                x = x - sin(phi - v) + sin(phi);
                y = y + cos(phi - v) - cos(phi);
                yaw = phi - v;
                break;
            case S:
                x = x + v * cos(phi);
                y = y + v * sin(phi);
                break;
            case NOP:
                break;
            }
        }
        out->v[0] = x * rho_ + from->getX();
        out->v[1] = y * rho_ + from->getY();
        out->v[2] = pp_wrap_yaw(yaw);
    }

    Path solve(const State *a, const State *b) const
    {
        double dx = b->getX() - a->getX(), dy = b->getY() - a->getY();
        double c = cos(a->getYaw()), s = sin(a->getYaw());
        double x = c * dx + s * dy, y = -s * dx + c * dy, phi = b->getYaw() - a->getYaw();
        return solveNormalised(x / rho_, y / rho_, phi);
    }

private:
    static constexpr double kPi = 3.14159265358979323846;
    static constexpr double kTwoPi = 2.0 * kPi;
    static constexpr double kZero = 10.0 * std::numeric_limits<double>::epsilon();

    static double wrapPi(double x)
    {
        double v = fmod(x, kTwoPi);
        if (v < -kPi)
            v += kTwoPi;
        else if (v > kPi)
            v -= kTwoPi;
        return v;
    }
    static void polar(double x, double y, double &r, double &theta)
    {
        r = sqrt(x * x + y * y);
        theta = atan2(y, x);
    }
    static void tauOmega(double u, double v, double xi, double eta, double phi, double &tau, double &omega)
    {
        double delta = wrapPi(u - v), A = sin(u) - sin(delta), B = cos(u) - cos(delta) - 1.0;
        double t1 = atan2(eta * A - xi * B, xi * A + eta * B), t2 = 2.0 * (cos(delta) - cos(v) - cos(u)) + 3.0;
        tau = (t2 < 0) ? wrapPi(t1 + kPi) : wrapPi(t1);
        omega = wrapPi(tau - u + v - phi);
    }

    // ---- base formulas (all return segment parameters of the "+" variant) ----
    static bool f81_LpSpLp(double x, double y, double phi, double &t, double &u, double &v)
    {
        polar(x - sin(phi), y - 1.0 + cos(phi), u, t);
        if (t >= -kZero)
        {
            v = wrapPi(phi - t);
            if (v >= -kZero)
                return true;
        }
        return false;
    }
    static bool f82_LpSpRp(double x, double y, double phi, double &t, double &u, double &v)
    {
        double t1, u1;
        polar(x + sin(phi), y - 1.0 - cos(phi), u1, t1);
        u1 = u1 * u1;
        if (u1 >= 4.0)
        {
            u = sqrt(u1 - 4.0);
            double theta = atan2(2.0, u);
            t = wrapPi(t1 + theta);
            v = wrapPi(t - phi);
            return t >= -kZero && v >= -kZero;
        }
        return false;
    }
    static bool f83_LpRmL(double x, double y, double phi, double &t, double &u, double &v)
    {
        double xi = x - sin(phi), eta = y - 1.0 + cos(phi), u1, theta;
        polar(xi, eta, u1, theta);
        if (u1 <= 4.0)
        {
            u = -2.0 * asin(0.25 * u1);
            t = wrapPi(theta + 0.5 * u + kPi);
            v = wrapPi(phi - t + u);
            return t >= -kZero && u <= kZero;
        }
        return false;
    }
    static bool f87_LpRupLumRm(double x, double y, double phi, double &t, double &u, double &v)
    {
        double xi = x + sin(phi), eta = y - 1.0 - cos(phi), rho = 0.25 * (2.0 + sqrt(xi * xi + eta * eta));
        if (rho <= 1.0)
        {
            u = acos(rho);
            tauOmega(u, -u, xi, eta, phi, t, v);
            return t >= -kZero && v <= kZero;
        }
        return false;
    }
    static bool f88_LpRumLumRp(double x, double y, double phi, double &t, double &u, double &v)
    {
        double xi = x + sin(phi), eta = y - 1.0 - cos(phi), rho = (20.0 - xi * xi - eta * eta) / 16.0;
        if (rho >= 0 && rho <= 1)
        {
            u = -acos(rho);
            if (u >= -0.5 * kPi)
            {
                tauOmega(u, u, xi, eta, phi, t, v);
                return t >= -kZero && v >= -kZero;
            }
        }
        return false;
    }
    static bool f89_LpRmSmLm(double x, double y, double phi, double &t, double &u, double &v)
    {
        double xi = x - sin(phi), eta = y - 1.0 + cos(phi), rho, theta;
        polar(xi, eta, rho, theta);
        if (rho >= 2.0)
        {
            double r = sqrt(rho * rho - 4.0);
            u = 2.0 - r;
            t = wrapPi(theta + atan2(r, -2.0));
            v = wrapPi(phi - 0.5 * kPi - t);
            return t >= -kZero && u <= kZero && v <= kZero;
        }
        return false;
    }
    static bool f810_LpRmSmRm(double x, double y, double phi, double &t, double &u, double &v)
    {
        double xi = x + sin(phi), eta = y - 1.0 - cos(phi), rho, theta;
        polar(-eta, xi, rho, theta);
        if (rho >= 2.0)
        {
            t = theta;
            u = 2.0 - rho;
            v = wrapPi(t + 0.5 * kPi - phi);
            return t >= -kZero && u <= kZero && v <= kZero;
        }
        return false;
    }
    static bool f811_LpRmSLmRp(double x, double y, double phi, double &t, double &u, double &v)
    {
        double xi = x + sin(phi), eta = y - 1.0 - cos(phi), rho, theta;
        polar(xi, eta, rho, theta);
        if (rho >= 2.0)
        {
            u = 4.0 - sqrt(rho * rho - 4.0);
            if (u <= kZero)
            {
                t = wrapPi(atan2((4.0 - u) * xi - 2.0 * eta, -2.0 * xi + (u - 4.0) * eta));
                v = wrapPi(t - phi);
                return t >= -kZero && v >= -kZero;
            }
        }
        return false;
    }

    typedef bool (*Formula)(double, double, double, double &, double &, double &);

    // A word pair: segment types for the plain and the reflected variant.
    struct Word { Seg plain[5]; Seg mirrored[5]; };

    static void assign(Path &p, const Seg *ty, double a, double b, double c, double d, double e)
    {
        for (int i = 0; i < 5; ++i)
            p.type[i] = ty[i];
        p.len[0] = a; p.len[1] = b; p.len[2] = c; p.len[3] = d; p.len[4] = e;
        p.total = fabs(a) + fabs(b) + fabs(c) + fabs(d) + fabs(e);
    }

    // Visit identity / time-flip / reflect / both for one formula on (x,y,phi). `emit` lays the
    // (t,u,v) triple (already sign-adjusted by sgn) into the five segment lengths.
    template <class Emit>
    static void fourWays(Formula f, double x, double y, double phi, const Word &w, double &Lmin,
                         double uWeight, Path &best, Emit emit)
    {
        const double sx[4] = {1, -1, 1, -1}, sy[4] = {1, 1, -1, -1}, sp[4] = {1, -1, -1, 1};
        const double sg[4] = {1, -1, 1, -1};
        const bool mir[4] = {false, false, true, true};
        for (int k = 0; k < 4; ++k)
        {
            double t, u, v, Lc;
            if (f(sx[k] * x, sy[k] * y, sp[k] * phi, t, u, v) &&
                Lmin > (Lc = fabs(t) + uWeight * fabs(u) + fabs(v)))
            {
                emit(best, mir[k] ? w.mirrored : w.plain, sg[k], t, u, v);
                Lmin = Lc;
            }
        }
    }

    static Path solveNormalised(double x, double y, double phi)
    {
        const double hp = 0.5 * kPi;
        Path best;
        double Lmin;

        static const Word W_LSL = {{L, S, L, NOP, NOP}, {R, S, R, NOP, NOP}};
        static const Word W_LSR = {{L, S, R, NOP, NOP}, {R, S, L, NOP, NOP}};
        static const Word W_LRL = {{L, R, L, NOP, NOP}, {R, L, R, NOP, NOP}};
        static const Word W_LRLR = {{L, R, L, R, NOP}, {R, L, R, L, NOP}};
        static const Word W_LRSL = {{L, R, S, L, NOP}, {R, L, S, R, NOP}};
        static const Word W_LSRL = {{L, S, R, L, NOP}, {R, S, L, R, NOP}};
        static const Word W_LRSR = {{L, R, S, R, NOP}, {R, L, S, L, NOP}};
        static const Word W_RSRL = {{R, S, R, L, NOP}, {L, S, L, R, NOP}};
        static const Word W_LRSLR = {{L, R, S, L, R}, {R, L, S, R, L}};

        auto tuv = [](Path &p, const Seg *ty, double s, double t, double u, double v) { assign(p, ty, s * t, s * u, s * v, 0, 0); };
        auto vut = [](Path &p, const Seg *ty, double s, double t, double u, double v) { assign(p, ty, s * v, s * u, s * t, 0, 0); };

        // CSC
        Lmin = best.length();
        fourWays(f81_LpSpLp, x, y, phi, W_LSL, Lmin, 1.0, best, tuv);
        fourWays(f82_LpSpRp, x, y, phi, W_LSR, Lmin, 1.0, best, tuv);

        // CCC (forward problem, then the reversed one)
        Lmin = best.length();
        fourWays(f83_LpRmL, x, y, phi, W_LRL, Lmin, 1.0, best, tuv);
        const double xb = x * cos(phi) + y * sin(phi), yb = x * sin(phi) - y * cos(phi);
        fourWays(f83_LpRmL, xb, yb, phi, W_LRL, Lmin, 1.0, best, vut);

        // CCCC
        Lmin = best.length();
        fourWays(f87_LpRupLumRm, x, y, phi, W_LRLR, Lmin, 2.0, best,
                 [](Path &p, const Seg *ty, double s, double t, double u, double v) { assign(p, ty, s * t, s * u, -s * u, s * v, 0); });
        fourWays(f88_LpRumLumRp, x, y, phi, W_LRLR, Lmin, 2.0, best,
                 [](Path &p, const Seg *ty, double s, double t, double u, double v) { assign(p, ty, s * t, s * u, s * u, s * v, 0); });

        // CCSC
        Lmin = best.length() - hp;
        auto t_q_u_v = [hp](Path &p, const Seg *ty, double s, double t, double u, double v) { assign(p, ty, s * t, -s * hp, s * u, s * v, 0); };
        auto v_u_q_t = [hp](Path &p, const Seg *ty, double s, double t, double u, double v) { assign(p, ty, s * v, s * u, -s * hp, s * t, 0); };
        fourWays(f89_LpRmSmLm, x, y, phi, W_LRSL, Lmin, 1.0, best, t_q_u_v);
        fourWays(f810_LpRmSmRm, x, y, phi, W_LRSR, Lmin, 1.0, best, t_q_u_v);
        fourWays(f89_LpRmSmLm, xb, yb, phi, W_LSRL, Lmin, 1.0, best, v_u_q_t);
        fourWays(f810_LpRmSmRm, xb, yb, phi, W_RSRL, Lmin, 1.0, best, v_u_q_t);

        // CCSCC
        Lmin = best.length() - kPi;
        fourWays(f811_LpRmSLmRp, x, y, phi, W_LRSLR, Lmin, 1.0, best,
                 [hp](Path &p, const Seg *ty, double s, double t, double u, double v) { assign(p, ty, s * t, -s * hp, s * u, -s * hp, s * v); });
        return best;
    }

    double rho_;
};
}
}
