// Shortest Dubins and Reeds-Shepp curves for the host Hybrid-A* driver (hybrid_astar.h).
// The reference gets these from OMPL (motion_hybrid.cpp:35,54; astar_hybrid.cpp:84-91,297-317):
// DubinsStateSpace / ReedsSheppStateSpace distance() and interpolate(). OMPL is an un-vendored,
// un-pinned dependency of the reference and is not available here, so the published algorithms
// are implemented from scratch: Dubins 1957 with the closed forms tabulated by Shkel & Lumelsky
// (six words LSL RSR RSL LSR RLR LRL), Reeds & Shepp 1990 formulas 8.1-8.11 under the four
// symmetries (identity, time-flip, reflection, both), families visited in the order CSC, CCC,
// CCCC, CCSC, CCSCC, a candidate replacing the incumbent only if strictly shorter. States are
// (x, y, yaw) in doubles; segment lengths are in units of the turning radius.
#pragma once
#include <algorithm>
#include <cmath>
#include <limits>

namespace ppb
{
struct SE2
{
    double x, y, yaw;
};

class CurveSpace
{
public:
    enum Kind { DUBINS = 1, REEDS_SHEPP = 2 };
    CurveSpace(Kind kind, double turning_radius) : kind_(kind), rho_(turning_radius) {}

    double distance(const SE2 &a, const SE2 &b) const { return rho_ * solve(a, b).total; }

    // The four trigonometric values a Reeds-Shepp query starts with, for callers that ask about the same
    // pair of headings again and again (a search has one goal heading and a few dozen node headings):
    // values of exactly the calls distance() makes, so the result is the same bits.
    struct HeadingTrig
    {
        double cos_a, sin_a;      // of a.yaw
        double sin_phi, cos_phi;  // of b.yaw - a.yaw
    };
    static HeadingTrig headingTrig(double a_yaw, double b_yaw)
    {
        const double phi = b_yaw - a_yaw;
        return {cos(a_yaw), sin(a_yaw), sin(phi), cos(phi)};
    }
    double distance(const SE2 &a, const SE2 &b, const HeadingTrig &h) const
    {
        if (kind_ == DUBINS)
            return distance(a, b);
        const double dx = b.x - a.x, dy = b.y - a.y;
        return rho_ * reeds_shepp((h.cos_a * dx + h.sin_a * dy) / rho_, (-h.sin_a * dx + h.cos_a * dy) / rho_, b.yaw - a.yaw,
                                  h.sin_phi, h.cos_phi).total;
    }

    SE2 interpolate(const SE2 &from, const SE2 &to, double t) const
    {
        if (t >= 1.0)
            return to;
        if (t <= 0.0)
            return from;
        const Word w = solve(from, to);
        double seg = t * w.total;
        double x = 0.0, y = 0.0, yaw = from.yaw;
        for (int i = 0; i < 5 && seg > 0.0; ++i)
        {
            double v;
            if (w.len[i] < 0)
            {
                v = std::max(-seg, w.len[i]);
                seg += v;
            }
            else
            {
                v = std::min(seg, w.len[i]);
                seg -= v;
            }
            const double phi = yaw;
            switch (w.seg[i])
            {
            case 'L':
                x = x + sin(phi + v) - sin(phi);
                y = y - cos(phi + v) + cos(phi);
                yaw = phi + v;
                break;
            case 'R':
                x = x - sin(phi - v) + sin(phi);
                y = y + cos(phi - v) - cos(phi);
                yaw = phi - v;
                break;
            case 'S':
                x = x + v * cos(phi);
                y = y + v * sin(phi);
                break;
            default:
                break;
            }
        }
        return {x * rho_ + from.x, y * rho_ + from.y, wrap_pi_half_open(yaw)};
    }

private:
    struct Word
    {
        char seg[5] = {'L', 'S', 'L', 0, 0};
        double len[5] = {0, 0, 0, 0, 0};
        double total = std::numeric_limits<double>::max();
    };
    static constexpr double kPi = 3.14159265358979323846;
    static constexpr double kTwoPi = 2.0 * kPi;

    static double wrap_pi_half_open(double yaw)  // into [-pi, pi), as an SO(2) component is kept
    {
        double v = std::fmod(yaw, kTwoPi);
        if (v < -kPi)
            v += kTwoPi;
        else if (v >= kPi)
            v -= kTwoPi;
        return v;
    }

    Word solve(const SE2 &a, const SE2 &b) const
    {
        const double dx = b.x - a.x, dy = b.y - a.y;
        if (kind_ == DUBINS)
        {
            const double d = sqrt(dx * dx + dy * dy) / rho_;
            const double th = atan2(dy, dx);
            return dubins(d, dub_mod(a.yaw - th), dub_mod(b.yaw - th));
        }
        const double c = cos(a.yaw), s = sin(a.yaw);
        return reeds_shepp((c * dx + s * dy) / rho_, (-s * dx + c * dy) / rho_, b.yaw - a.yaw);
    }

    // ---------------- Dubins ----------------
    static constexpr double kDubEps = 1e-6, kDubZero = -1e-7;
    static double dub_mod(double x)
    {
        if (x < 0 && x > kDubZero)
            return 0;
        double xm = x - kTwoPi * floor(x / kTwoPi);
        if (kTwoPi - xm < 0.5 * kDubEps)
            xm = 0.0;
        return xm;
    }
    static Word word3(char a, char b, char c, double t, double p, double q)
    {
        Word w;
        w.seg[0] = a; w.seg[1] = b; w.seg[2] = c;
        w.len[0] = t; w.len[1] = p; w.len[2] = q;
        w.total = t + p + q;
        return w;
    }
    static void keep_shorter(Word &best, const Word &cand)
    {
        if (cand.total < best.total)
            best = cand;
    }
    static Word dubins(double d, double alpha, double beta)
    {
        if (d < kDubEps && fabs(alpha - beta) < kDubEps)
            return word3('L', 'S', 'L', 0.0, d, 0.0);
        const double ca = cos(alpha), sa = sin(alpha), cb = cos(beta), sb = sin(beta);
        const double cab = ca * cb + sa * sb;
        Word best;  // total = max: "no path yet"
        double tmp = 2.0 + d * d - 2.0 * (cab - d * (sa - sb));
        if (tmp >= kDubZero)
        {
            const double th = atan2(cb - ca, d + sa - sb);
            best = word3('L', 'S', 'L', dub_mod(-alpha + th), sqrt(std::max(tmp, 0.0)), dub_mod(beta - th));
        }
        tmp = 2.0 + d * d - 2.0 * (cab - d * (sb - sa));
        if (tmp >= kDubZero)
        {
            const double th = atan2(ca - cb, d - sa + sb);
            keep_shorter(best, word3('R', 'S', 'R', dub_mod(alpha - th), sqrt(std::max(tmp, 0.0)), dub_mod(-beta + th)));
        }
        tmp = d * d - 2.0 + 2.0 * (cab - d * (sa + sb));
        if (tmp >= kDubZero)
        {
            const double p = sqrt(std::max(tmp, 0.0));
            const double th = atan2(ca + cb, d - sa - sb) - atan2(2.0, p);
            keep_shorter(best, word3('R', 'S', 'L', dub_mod(alpha - th), p, dub_mod(beta - th)));
        }
        tmp = -2.0 + d * d + 2.0 * (cab + d * (sa + sb));
        if (tmp >= kDubZero)
        {
            const double p = sqrt(std::max(tmp, 0.0));
            const double th = atan2(-ca - cb, d + sa + sb) - atan2(-2.0, p);
            keep_shorter(best, word3('L', 'S', 'R', dub_mod(-alpha + th), p, dub_mod(-beta + th)));
        }
        tmp = 0.125 * (6.0 - d * d + 2.0 * (cab + d * (sa - sb)));
        if (fabs(tmp) < 1.0)
        {
            const double p = kTwoPi - acos(tmp);
            const double th = atan2(ca - cb, d - sa + sb);
            const double t = dub_mod(alpha - th + 0.5 * p);
            keep_shorter(best, word3('R', 'L', 'R', t, p, dub_mod(alpha - beta - t + p)));
        }
        tmp = 0.125 * (6.0 - d * d + 2.0 * (cab - d * (sa - sb)));
        if (fabs(tmp) < 1.0)
        {
            const double p = kTwoPi - acos(tmp);
            const double th = atan2(-ca + cb, d + sa - sb);
            const double t = dub_mod(-alpha + th + 0.5 * p);
            keep_shorter(best, word3('L', 'R', 'L', t, p, dub_mod(beta - alpha - t + p)));
        }
        if (best.total == std::numeric_limits<double>::max())
            best = word3('L', 'S', 'L', 0.0, std::numeric_limits<double>::max(), 0.0);
        return best;
    }

    // ---------------- Reeds-Shepp ----------------
    static constexpr double kRsZero = 10.0 * std::numeric_limits<double>::epsilon();
    // fmod(x, 2 pi) without the library call where the result is known exactly: |x| < 2 pi leaves x,
    // 2 pi <= |x| <= 4 pi is one exact subtraction (Sterbenz); fmod itself is exact, so this is bit-identical
    static double fmod_two_pi(double x)
    {
        const double a = fabs(x);
        if (a < kTwoPi)
            return x;
        if (a <= 2.0 * kTwoPi)
            return x < 0 ? x + kTwoPi : x - kTwoPi;
        return fmod(x, kTwoPi);
    }
    static double rs_mod(double x)
    {
        double v = fmod_two_pi(x);
        if (v < -kPi)
            v += kTwoPi;
        else if (v > kPi)
            v -= kTwoPi;
        return v;
    }
    static void tau_omega(double u, double v, double xi, double eta, double phi, double &tau, double &omega)
    {
        const double delta = rs_mod(u - v), A = sin(u) - sin(delta), B = cos(u) - cos(delta) - 1.0;
        const double t1 = atan2(eta * A - xi * B, xi * A + eta * B), t2 = 2.0 * (cos(delta) - cos(v) - cos(u)) + 3.0;
        tau = (t2 < 0) ? rs_mod(t1 + kPi) : rs_mod(t1);
        omega = rs_mod(tau - u + v - phi);
    }
    // Sign pre-tests. A formula is rejected when its first arc t is negative beyond the tolerance; whether
    // it is can often be seen from the sign of a sine that costs two multiplications instead of two
    // atan2. Only clear cases are decided here (margins far above the rounding error of the test itself
    // and far above kRsZero); everything else takes the full evaluation, so the outcome is unchanged.
    static bool clearly_negative_atan2(double y, double x)  // atan2(y, x) < -kRsZero for certain
    {
        return y < 0.0 && !(x > 0.0 && -y <= 1e-14 * x);
    }
    static bool clearly_negative_sine(double a, double b)  // a + b < 0 for certain, a and b products of the inputs
    {
        return a + b < -1e-12 * (fabs(a) + fabs(b));
    }

    // Every formula receives sin(phi) and cos(phi) (computed once per query; sin is odd and cos even
    // bit for bit in the host libm, checked in tests/test_curves.py) and the incumbent bound Lmin: a
    // candidate replaces the incumbent only if |t| + w|u| + |v| < Lmin, so once a partial sum reaches
    // Lmin the remaining inverse trigonometry is skipped -- the outcome is the same.
    static bool f81(double x, double y, double phi, double sp, double cp, double Lmin, double &t, double &u, double &v)  // L+S+L+
    {
        const double xi = x - sp, eta = y - 1.0 + cp;
        u = sqrt(xi * xi + eta * eta);
        if (!(u < Lmin) || clearly_negative_atan2(eta, xi))
            return false;
        t = atan2(eta, xi);
        if (t >= -kRsZero)
        {
            v = rs_mod(phi - t);
            if (v >= -kRsZero)
                return true;
        }
        return false;
    }
    static bool f82(double x, double y, double phi, double sp, double cp, double Lmin, double &t, double &u, double &v)  // L+S+R+
    {
        const double xi = x + sp, eta = y - 1.0 - cp;
        double u1 = sqrt(xi * xi + eta * eta);
        u1 = u1 * u1;
        if (u1 >= 4.0)
        {
            u = sqrt(u1 - 4.0);
            if (!(u < Lmin) || clearly_negative_sine(eta * u, xi * 2.0))  // sin(t1 + theta) up to a positive factor
                return false;
            const double t1 = atan2(eta, xi);
            const double theta = atan2(2.0, u);
            t = rs_mod(t1 + theta);
            v = rs_mod(t - phi);
            return t >= -kRsZero && v >= -kRsZero;
        }
        return false;
    }
    static bool f83(double x, double y, double phi, double sp, double cp, double Lmin, double &t, double &u, double &v)  // L+R-L
    {
        const double xi = x - sp, eta = y - 1.0 + cp;
        const double u1 = sqrt(xi * xi + eta * eta);
        if (u1 <= 4.0)
        {
            const double theta = atan2(eta, xi);
            u = -2.0 * asin(0.25 * u1);
            t = rs_mod(theta + 0.5 * u + kPi);
            v = rs_mod(phi - t + u);
            return t >= -kRsZero && u <= kRsZero;
        }
        (void)Lmin;
        return false;
    }
    static bool f87(double x, double y, double phi, double sp, double cp, double Lmin, double &t, double &u, double &v)  // L+R+uL-uR-
    {
        const double xi = x + sp, eta = y - 1.0 - cp, rho = 0.25 * (2.0 + sqrt(xi * xi + eta * eta));
        if (rho <= 1.0)
        {
            u = acos(rho);
            tau_omega(u, -u, xi, eta, phi, t, v);
            return t >= -kRsZero && v <= kRsZero;
        }
        (void)Lmin;
        return false;
    }
    static bool f88(double x, double y, double phi, double sp, double cp, double Lmin, double &t, double &u, double &v)  // L+R-uL-uR+
    {
        const double xi = x + sp, eta = y - 1.0 - cp, rho = (20.0 - xi * xi - eta * eta) / 16.0;
        if (rho >= 0 && rho <= 1)
        {
            u = -acos(rho);
            if (u >= -0.5 * kPi)
            {
                tau_omega(u, u, xi, eta, phi, t, v);
                return t >= -kRsZero && v >= -kRsZero;
            }
        }
        (void)Lmin;
        return false;
    }
    static bool f89(double x, double y, double phi, double sp, double cp, double Lmin, double &t, double &u, double &v)  // L+R-S-L-
    {
        const double xi = x - sp, eta = y - 1.0 + cp;
        const double rho = sqrt(xi * xi + eta * eta);
        if (rho >= 2.0)
        {
            const double r = sqrt(rho * rho - 4.0);
            u = 2.0 - r;
            if (!(u <= kRsZero) || !(fabs(u) < Lmin) || clearly_negative_sine(eta * -2.0, xi * r))
                return false;
            const double theta = atan2(eta, xi);
            t = rs_mod(theta + atan2(r, -2.0));
            v = rs_mod(phi - 0.5 * kPi - t);
            return t >= -kRsZero && u <= kRsZero && v <= kRsZero;
        }
        return false;
    }
    static bool f810(double x, double y, double phi, double sp, double cp, double Lmin, double &t, double &u, double &v)  // L+R-S-R-
    {
        const double xi = x + sp, eta = y - 1.0 - cp;
        const double rho = sqrt(eta * eta + xi * xi);
        if (rho >= 2.0)
        {
            u = 2.0 - rho;
            if (!(u <= kRsZero) || !(fabs(u) < Lmin) || clearly_negative_atan2(xi, -eta))
                return false;
            t = atan2(xi, -eta);
            v = rs_mod(t + 0.5 * kPi - phi);
            return t >= -kRsZero && u <= kRsZero && v <= kRsZero;
        }
        return false;
    }
    static bool f811(double x, double y, double phi, double sp, double cp, double Lmin, double &t, double &u, double &v)  // L+R-S-L-R+
    {
        const double xi = x + sp, eta = y - 1.0 - cp;
        const double rho = sqrt(xi * xi + eta * eta);
        if (rho >= 2.0)
        {
            u = 4.0 - sqrt(rho * rho - 4.0);
            if (u <= kRsZero)
            {
                if (!(fabs(u) < Lmin))
                    return false;
                const double ty = (4.0 - u) * xi - 2.0 * eta, tx = -2.0 * xi + (u - 4.0) * eta;
                if (clearly_negative_atan2(ty, tx))
                    return false;
                t = rs_mod(atan2(ty, tx));
                v = rs_mod(t - phi);
                return t >= -kRsZero && v >= -kRsZero;
            }
        }
        return false;
    }

    typedef bool (*Formula)(double, double, double, double, double, double, double &, double &, double &);
    // layout of (t, u, v) over the five segments: 0 tuv, 1 vut, 2 t,u,-u,v, 3 t,u,u,v,
    // 4 t,-pi/2,u,v, 5 v,u,-pi/2,t, 6 t,-pi/2,u,-pi/2,v
    static void lay(Word &w, const char *segs, int layout, double s, double t, double u, double v)
    {
        const double hp = 0.5 * kPi;
        double l[5] = {0, 0, 0, 0, 0};
        switch (layout)
        {
        case 0: l[0] = s * t; l[1] = s * u; l[2] = s * v; break;
        case 1: l[0] = s * v; l[1] = s * u; l[2] = s * t; break;
        case 2: l[0] = s * t; l[1] = s * u; l[2] = -s * u; l[3] = s * v; break;
        case 3: l[0] = s * t; l[1] = s * u; l[2] = s * u; l[3] = s * v; break;
        case 4: l[0] = s * t; l[1] = -s * hp; l[2] = s * u; l[3] = s * v; break;
        case 5: l[0] = s * v; l[1] = s * u; l[2] = -s * hp; l[3] = s * t; break;
        default: l[0] = s * t; l[1] = -s * hp; l[2] = s * u; l[3] = -s * hp; l[4] = s * v; break;
        }
        w.total = 0.0;
        for (int i = 0; i < 5; ++i)
        {
            w.seg[i] = segs[i] == '-' ? 0 : segs[i];
            w.len[i] = l[i];
            w.total += fabs(l[i]);
        }
    }
    // one formula under the four symmetries; plain / mirrored give the segment letters
    static void four_ways(Formula f, double x, double y, double phi, double sphi, double cphi, const char *plain,
                          const char *mirrored, int layout, double u_weight, double &Lmin, Word &best)
    {
        const double sx[4] = {1, -1, 1, -1}, sy[4] = {1, 1, -1, -1}, sp[4] = {1, -1, -1, 1}, sg[4] = {1, -1, 1, -1};
        for (int k = 0; k < 4; ++k)
        {
            double t, u, v, L;
            // bound on |u| alone: w|u| >= |u| for w >= 1
            if (f(sx[k] * x, sy[k] * y, sp[k] * phi, sp[k] * sphi, cphi, Lmin, t, u, v) &&
                Lmin > (L = fabs(t) + u_weight * fabs(u) + fabs(v)))
            {
                lay(best, k >= 2 ? mirrored : plain, layout, sg[k], t, u, v);
                Lmin = L;
            }
        }
    }
    static Word reeds_shepp(double x, double y, double phi) { return reeds_shepp(x, y, phi, sin(phi), cos(phi)); }
    static Word reeds_shepp(double x, double y, double phi, double sphi, double cphi)
    {
        Word best;
        best.seg[0] = 'L'; best.seg[1] = 'R'; best.seg[2] = 'L';
        double Lmin = best.total;
        four_ways(f81, x, y, phi, sphi, cphi, "LSL--", "RSR--", 0, 1.0, Lmin, best);
        four_ways(f82, x, y, phi, sphi, cphi, "LSR--", "RSL--", 0, 1.0, Lmin, best);
        Lmin = best.total;
        four_ways(f83, x, y, phi, sphi, cphi, "LRL--", "RLR--", 0, 1.0, Lmin, best);
        const double xb = x * cphi + y * sphi, yb = x * sphi - y * cphi;
        four_ways(f83, xb, yb, phi, sphi, cphi, "LRL--", "RLR--", 1, 1.0, Lmin, best);
        Lmin = best.total;
        four_ways(f87, x, y, phi, sphi, cphi, "LRLR-", "RLRL-", 2, 2.0, Lmin, best);
        four_ways(f88, x, y, phi, sphi, cphi, "LRLR-", "RLRL-", 3, 2.0, Lmin, best);
        Lmin = best.total - 0.5 * kPi;
        four_ways(f89, x, y, phi, sphi, cphi, "LRSL-", "RLSR-", 4, 1.0, Lmin, best);
        four_ways(f810, x, y, phi, sphi, cphi, "LRSR-", "RLSL-", 4, 1.0, Lmin, best);
        four_ways(f89, xb, yb, phi, sphi, cphi, "LSRL-", "RSLR-", 5, 1.0, Lmin, best);
        four_ways(f810, xb, yb, phi, sphi, cphi, "RSRL-", "LSLR-", 5, 1.0, Lmin, best);
        Lmin = best.total - kPi;
        four_ways(f811, x, y, phi, sphi, cphi, "LRSLR", "RLSRL", 6, 1.0, Lmin, best);
        return best;
    }

    Kind kind_;
    double rho_;
};
}  // namespace ppb
