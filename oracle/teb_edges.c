/* TEST INFRASTRUCTURE ONLY (oracle). CPU restatement of the obstacle-edge evaluation that follows
 * the TEB association. Pinned against the reference's own code
 * compiled into oracle/_ref/libpp_teb_ref.so (tests/test_oracle_teb_edges.py, fixture
 * tests/golden/teb_edges_v1.npz). The numeric Jacobian is g2o's (un-vendored, un-pinned; restated
 * from g2o/core/base_unary_edge.hpp) over the reference's own Vertex2E::oplusImpl.
 * All paths are relative to /root/reference/src/path_follower/src/teb_path_follower.
 */
#include "port.h"
#include <float.h>
#include <math.h>
#include <string.h>

typedef struct
{
    double x, y;
} P2;

/* commons.cpp:3-6 */
static double p2_distance(P2 a, P2 b) { return hypot(a.x - b.x, a.y - b.y); }

/* commons.cpp:42-56 */
static P2 nearest_point_on_line(P2 point, P2 s, P2 e)
{
    double dx = e.x - s.x;
    double dy = e.y - s.y;
    double sq_norm = dx * dx + dy * dy;
    if (sq_norm < 1e-6)
        return s;
    double u = ((point.x - s.x) * dx + (point.y - s.y) * dy) / sq_norm;
    if (u <= 0)
        return s;
    if (u >= 1)
        return e;
    P2 r = {s.x + u * dx, s.y + u * dy};
    return r;
}

/* commons.cpp:36-40 */
static double point_to_line_distance(P2 point, P2 s, P2 e) { return p2_distance(point, nearest_point_on_line(point, s, e)); }

/* RobotModelHelper::createRobotModel, robot_model.cpp:68-95: 1 number = circle radius, 2 = width and
 * height of a rectangle, otherwise polygon vertices x0,y0,x1,y1,... Returns the vertex count (0 = circle),
 * -1 on an odd count or fewer than 3 vertices (the reference throws). */
int port_robot_model(const double *numbers, int n, double *verts_xy, int max_verts, double *inscribed, double *circum)
{
    if (n == 1)
    {
        *inscribed = *circum = numbers[0]; /* robot_model.h:24-25 */
        return 0;
    }
    int nv;
    if (n == 2)
    {
        double hw = numbers[0] / 2.0, hh = numbers[1] / 2.0;
        double r[8] = {hw, hh, hw, -hh, -hw, -hh, -hw, hh};
        if (max_verts < 4)
            return -1;
        memcpy(verts_xy, r, sizeof r);
        nv = 4;
    }
    else
    {
        if (n % 2 != 0 || n / 2 < 3 || n / 2 > max_verts)
            return -1;
        memcpy(verts_xy, numbers, sizeof(double) * n);
        nv = n / 2;
    }
    /* PolygonRobotModel::computeRadius, robot_model.cpp:15-37 */
    const P2 *v = (const P2 *)verts_xy;
    P2 center = {0.0, 0.0};
    double distance = 0.0;
    double min_dist = DBL_MAX;
    double max_dist = DBL_MIN; /* numeric_limits<double>::min(): the smallest positive normal */
    for (int i = 0; i < nv; ++i)
    {
        distance = p2_distance(center, v[i]);
        min_dist = fmin(min_dist, distance);
        max_dist = fmax(max_dist, distance);
    }
    for (int i = 0; i < nv - 1; ++i)
    {
        distance = point_to_line_distance(center, v[i], v[i + 1]);
        min_dist = fmin(min_dist, distance);
        max_dist = fmax(max_dist, distance);
    }
    distance = point_to_line_distance(center, v[nv - 1], v[0]);
    *inscribed = fmin(min_dist, distance);
    *circum = fmax(max_dist, distance);
    return nv;
}

typedef struct
{
    int nv; /* 0 = circle */
    const P2 *v;
    double inscribed, circum;
} Robot;

/* CircleRobotModel::distanceToPoint robot_model.cpp:3-6, PolygonRobotModel::distanceToPoint :39-66 */
static double distance_to_point(const Robot *r, double px, double py, double ptheta, P2 point)
{
    if (r->nv == 0)
        return hypot(point.x - px, point.y - py) - r->circum;
    P2 pose = {px, py};
    double distance = p2_distance(pose, point);
    if (distance <= r->inscribed)
        return distance - r->circum;
    double dist = DBL_MAX;
    for (int i = 0; i < r->nv; ++i)
    {
        const P2 a = r->v[i], b = r->v[(i + 1) % r->nv];
        P2 s = {cos(ptheta) * a.x - sin(ptheta) * a.y + px, sin(ptheta) * a.x + cos(ptheta) * a.y + py};
        P2 e = {cos(ptheta) * b.x - sin(ptheta) * b.y + px, sin(ptheta) * b.x + cos(ptheta) * b.y + py};
        distance = point_to_line_distance(point, s, e);
        dist = fmin(dist, distance);
    }
    return dist;
}

/* commons.h:53-62 */
static double penalty_below(double var, double a, double epsilon)
{
    if (var > a + epsilon)
        return 0.0;
    return a + epsilon - var;
}

/* commons.h:11-19 */
static double normalize_theta(double angle)
{
    const double PI = M_PI;
    double result = fmod(angle + PI, 2 * PI);
    if (result <= 0)
        return result + PI;
    return result - PI;
}

/* EdgeObstacle::computeError, g2o_edges/edge_obstacle.hpp:19-26 */
static void compute_error(const Robot *r, const double *pose, P2 obstacle, const double *info4, double *err)
{
    double dist = distance_to_point(r, pose[0], pose[1], pose[2], obstacle);
    err[0] = penalty_below(dist, info4[0], info4[1]);
    err[1] = penalty_below(dist, info4[2], info4[3]);
}

/* Same contract as teb_ref_obstacle_edges (oracle/teb_ref_harness.cpp). */
int port_obstacle_edges(const double *poses_xyt, uint32_t Q, const double *pts_xy, const uint32_t *idx,
                        const uint32_t *cnt, uint32_t K, const double *robot_numbers, int n_robot, const double *info4,
                        double *err_out, double *jac_out)
{
    double verts[2 * 64];
    Robot robot;
    robot.nv = port_robot_model(robot_numbers, n_robot, verts, 64, &robot.inscribed, &robot.circum);
    if (robot.nv < 0)
        return -1;
    robot.v = (const P2 *)verts;
    memset(err_out, 0, sizeof(double) * Q * K * 2);
    if (jac_out)
        memset(jac_out, 0, sizeof(double) * Q * K * 6);
    for (uint32_t q = 0; q < Q; ++q)
        for (uint32_t k = 0; k < cnt[q] && k < K; ++k)
        {
            const uint32_t i = idx[q * K + k];
            const P2 obstacle = {pts_xy[2 * i], pts_xy[2 * i + 1]};
            const double *pose = poses_xyt + 3 * q;
            compute_error(&robot, pose, obstacle, info4, err_out + (q * K + k) * 2);
            if (!jac_out)
                continue;
            /* g2o BaseUnaryEdge::linearizeOplus over Vertex2E::oplusImpl (commons.cpp:75-80) */
            const double delta = 1e-9;
            const double scalar = 1 / (2 * delta);
            for (int d = 0; d < 3; ++d)
            {
                double e1[2], e2[2], p[3];
                for (int sgn = 0; sgn < 2; ++sgn)
                {
                    double add[3] = {0.0, 0.0, 0.0};
                    add[d] = sgn == 0 ? delta : -delta;
                    p[0] = pose[0] + add[0];
                    p[1] = pose[1] + add[1];
                    p[2] = normalize_theta(pose[2] + add[2]);
                    compute_error(&robot, p, obstacle, info4, sgn == 0 ? e1 : e2);
                }
                for (int r = 0; r < 2; ++r)
                    jac_out[(q * K + k) * 6 + r * 3 + d] = scalar * (e1[r] - e2[r]);
            }
        }
    return 0;
}
