// N3: obstacle-edge errors and Jacobians for the TEB optimiser, batched.
// The step after the association of K3: TebOptimizer::addObstacleEdges (teb_optimizer.cpp:283-313)
// creates one EdgeObstacle per (pose, kept obstacle); g2o then calls its computeError()
// (g2o_edges/edge_obstacle.hpp:19-26: PolygonRobotModel::distanceToPoint, robot_model.cpp:39-66,
// two penaltyBelow terms, commons.h:53-62) and, because the edge has no analytic Jacobian, g2o's
// central-difference linearizeOplus() (step 1e-9 through Vertex2E::oplusImpl, commons.cpp:75-80) on
// every Levenberg-Marquardt iteration. Here all edges of an iteration are evaluated in one launch
// from the association that pp_radius_topk left on the device: 8 lanes per edge, lane e evaluates
// the error at the base pose (e = 0) or at one of the six perturbed poses, the Jacobian columns are
// formed by shuffles. Double precision, no FMA contraction (-fmad=false), like the reference build.
// A few thousand point-segment distances: the launch is latency-bound (SURVEY.md section 8d says the
// same of K3); what it saves is 7 x edges virtual calls per iteration on the host.
#include "pp_internal.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>

namespace
{
constexpr int EDGE_LANES = 8, EDGE_THREADS = 256;

struct EdgeArgs
{
    const double *poses;  // Q x 3
    const double *pts;    // N x 2
    const uint32_t *idx;  // Q x K
    const uint32_t *cnt;  // Q
    double *out;          // Q x K x 8: error[2], jacobian[2][3]
    uint32_t Q, K;
    int want_jacobian;
    int nv;  // 0 = circle
    double inscribed, circum;
    double collision_dist, collision_margin, inflation_dist, inflation_margin;
    double verts[2 * PP_ROBOT_MAX_VERTS];
};

struct P2
{
    double x, y;
};

// commons.cpp:3-6
__device__ __forceinline__ double p2_distance(P2 a, P2 b) { return hypot(a.x - b.x, a.y - b.y); }

// commons.cpp:36-56
__device__ double point_to_line_distance(P2 point, P2 s, P2 e)
{
    const double dx = e.x - s.x;
    const double dy = e.y - s.y;
    const double sq_norm = dx * dx + dy * dy;
    P2 nearest = s;
    if (!(sq_norm < 1e-6))
    {
        const double u = ((point.x - s.x) * dx + (point.y - s.y) * dy) / sq_norm;
        if (u >= 1)
            nearest = e;
        else if (!(u <= 0))
            nearest = P2{s.x + u * dx, s.y + u * dy};
    }
    return p2_distance(point, nearest);
}

// robot_model.cpp:3-6 (circle) and :39-66 (polygon)
__device__ double distance_to_point(const EdgeArgs &a, double px, double py, double ptheta, P2 point)
{
    if (a.nv == 0)
        return hypot(point.x - px, point.y - py) - a.circum;
    double distance = p2_distance(P2{px, py}, point);
    if (distance <= a.inscribed)
        return distance - a.circum;
    double sn, cs;
    sincos(ptheta, &sn, &cs);
    double dist = DBL_MAX;
    P2 s{cs * a.verts[0] - sn * a.verts[1] + px, sn * a.verts[0] + cs * a.verts[1] + py};
    const P2 first = s;
    for (int i = 1; i <= a.nv; ++i)
    {
        P2 e = first;
        if (i < a.nv)
            e = P2{cs * a.verts[2 * i] - sn * a.verts[2 * i + 1] + px, sn * a.verts[2 * i] + cs * a.verts[2 * i + 1] + py};
        distance = point_to_line_distance(point, s, e);
        dist = fmin(dist, distance);
        s = e;
    }
    return dist;
}

// commons.h:53-62
__device__ __forceinline__ double penalty_below(double var, double a, double epsilon)
{
    return var > a + epsilon ? 0.0 : a + epsilon - var;
}

// commons.h:11-19
__device__ __forceinline__ double normalize_theta(double angle)
{
    const double PI = 3.14159265358979323846;
    const double result = fmod(angle + PI, 2 * PI);
    return result <= 0 ? result + PI : result - PI;
}

__global__ void __launch_bounds__(EDGE_THREADS) obstacle_edges_kernel(const __grid_constant__ EdgeArgs a)
{
    const uint32_t slot = (blockIdx.x * EDGE_THREADS + threadIdx.x) / EDGE_LANES;
    const int e = threadIdx.x % EDGE_LANES;  // 0 base, 1 + 2d plus step, 2 + 2d minus step in dimension d
    const bool live = slot < a.Q * a.K;
    const uint32_t q = live ? slot / a.K : 0, k = live ? slot - q * a.K : 0;
    const bool used = live && k < a.cnt[q];
    double err0 = 0.0, err1 = 0.0;
    if (used && e < 7 && (e == 0 || a.want_jacobian))
    {
        const uint32_t i = a.idx[slot];
        const P2 obstacle{a.pts[2 * i], a.pts[2 * i + 1]};
        double px = a.poses[3 * q], py = a.poses[3 * q + 1], pt = a.poses[3 * q + 2];
        if (e > 0)
        {
            // g2o BaseUnaryEdge::linearizeOplus: oplus with +-1e-9 in one dimension, zero in the others --
            // Vertex2E::oplusImpl normalises the heading on every update, also when its increment is zero
            const int d = (e - 1) >> 1;
            const double step = (e & 1) ? 1e-9 : -1e-9;
            px = px + (d == 0 ? step : 0.0);
            py = py + (d == 1 ? step : 0.0);
            pt = normalize_theta(pt + (d == 2 ? step : 0.0));
        }
        const double dist = distance_to_point(a, px, py, pt, obstacle);
        err0 = penalty_below(dist, a.collision_dist, a.collision_margin);
        err1 = penalty_below(dist, a.inflation_dist, a.inflation_margin);
    }
    // lanes of one edge are 8 consecutive lanes of a warp
    const unsigned base = (threadIdx.x % 32) & ~(EDGE_LANES - 1);
    const double scalar = 1 / (2 * 1e-9);
    double jac[6];
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
        const double p0 = __shfl_sync(0xffffffffu, err0, base + 1 + 2 * d), m0 = __shfl_sync(0xffffffffu, err0, base + 2 + 2 * d);
        const double p1 = __shfl_sync(0xffffffffu, err1, base + 1 + 2 * d), m1 = __shfl_sync(0xffffffffu, err1, base + 2 + 2 * d);
        jac[d] = scalar * (p0 - m0);
        jac[3 + d] = scalar * (p1 - m1);
    }
    if (live && e == 0)
    {
        double *o = a.out + static_cast<size_t>(slot) * 8;
        o[0] = err0;
        o[1] = err1;
#pragma unroll
        for (int j = 0; j < 6; ++j)
            o[2 + j] = used && a.want_jacobian ? jac[j] : 0.0;
    }
}

// host side of robot_model.cpp:15-37 and commons.cpp:36-56
double host_point_to_line(double px, double py, double sx, double sy, double ex, double ey)
{
    const double dx = ex - sx, dy = ey - sy;
    const double sq_norm = dx * dx + dy * dy;
    double nx = sx, ny = sy;
    if (!(sq_norm < 1e-6))
    {
        const double u = ((px - sx) * dx + (py - sy) * dy) / sq_norm;
        if (u >= 1)
            nx = ex, ny = ey;
        else if (!(u <= 0))
            nx = sx + u * dx, ny = sy + u * dy;
    }
    return std::hypot(px - nx, py - ny);
}
}  // namespace

pp_status pp_robot_model_set(pp_handle h, const double *numbers, uint32_t n, double *inscribed_out, double *circum_out)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, numbers && n >= 1, PP_ERR_INVALID, "pp_robot_model_set: no numbers");
    if (n == 1)
    {
        h->robot_nv = 0;
        h->robot_inscribed = h->robot_circum = numbers[0];
    }
    else
    {
        uint32_t nv;
        if (n == 2)
        {
            const double hw = numbers[0] / 2.0, hh = numbers[1] / 2.0;
            const double r[8] = {hw, hh, hw, -hh, -hw, -hh, -hw, hh};
            std::memcpy(h->robot_verts, r, sizeof r);
            nv = 4;
        }
        else
        {
            PP_REQUIRE(h, n % 2 == 0, PP_ERR_INVALID, "pp_robot_model_set: a polygon needs an even count of numbers (got %u)", n);
            PP_REQUIRE(h, n / 2 >= 3, PP_ERR_INVALID, "pp_robot_model_set: a polygon needs at least 3 vertices");
            PP_REQUIRE(h, n / 2 <= PP_ROBOT_MAX_VERTS, PP_ERR_RANGE, "pp_robot_model_set: %u vertices exceed the limit of %d",
                       n / 2, PP_ROBOT_MAX_VERTS);
            std::memcpy(h->robot_verts, numbers, sizeof(double) * n);
            nv = n / 2;
        }
        const double *v = h->robot_verts;
        double min_dist = DBL_MAX, max_dist = DBL_MIN, distance = 0.0;
        for (uint32_t i = 0; i < nv; ++i)
        {
            distance = std::hypot(0.0 - v[2 * i], 0.0 - v[2 * i + 1]);
            min_dist = std::min(min_dist, distance);
            max_dist = std::max(max_dist, distance);
        }
        for (uint32_t i = 0; i + 1 < nv; ++i)
        {
            distance = host_point_to_line(0.0, 0.0, v[2 * i], v[2 * i + 1], v[2 * i + 2], v[2 * i + 3]);
            min_dist = std::min(min_dist, distance);
            max_dist = std::max(max_dist, distance);
        }
        distance = host_point_to_line(0.0, 0.0, v[2 * nv - 2], v[2 * nv - 1], v[0], v[1]);
        h->robot_inscribed = std::min(min_dist, distance);
        h->robot_circum = std::max(max_dist, distance);
        h->robot_nv = static_cast<int>(nv);
    }
    if (inscribed_out)
        *inscribed_out = h->robot_inscribed;
    if (circum_out)
        *circum_out = h->robot_circum;
    return PP_OK;
}

pp_status pp_obstacle_edges(pp_handle h, const double *poses_xyt, uint32_t Q, uint32_t K,
                            const pp_obstacle_edge_params *params, double *error_out, double *jacobian_out)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->robot_nv >= 0, PP_ERR_STATE, "pp_obstacle_edges: no robot model (pp_robot_model_set)");
    PP_REQUIRE(h, h->topk_Q > 0 && h->d_topk_idx, PP_ERR_STATE, "pp_obstacle_edges: no association on the device (pp_radius_topk)");
    PP_REQUIRE(h, Q == h->topk_Q && K == h->topk_K, PP_ERR_INVALID,
               "pp_obstacle_edges: Q=%u K=%u do not match the last pp_radius_topk (Q=%u K=%u)", Q, K, h->topk_Q, h->topk_K);
    PP_REQUIRE(h, poses_xyt && params && error_out, PP_ERR_INVALID, "pp_obstacle_edges: null buffer");
    PP_CUDA(h, cudaSetDevice(h->device));
    const size_t slots = static_cast<size_t>(Q) * K;
    PP_CUDA(h, pp_reserve(h->d_poses, h->poses_cap, 3 * static_cast<size_t>(Q)));
    PP_CUDA(h, pp_reserve(h->d_edge_out, h->edge_cap, 8 * slots));
    pp_status s = pp_ensure_stage(h, 8 * slots * sizeof(double));
    if (s != PP_OK)
        return s;
    PP_CUDA(h, cudaMemcpyAsync(h->d_poses, poses_xyt, sizeof(double) * 3 * Q, cudaMemcpyHostToDevice, h->stream));
    EdgeArgs a{};
    a.poses = h->d_poses, a.pts = h->d_pts, a.idx = h->d_topk_idx, a.cnt = h->d_topk_cnt, a.out = h->d_edge_out;
    a.Q = Q, a.K = K, a.want_jacobian = jacobian_out != nullptr;
    a.nv = h->robot_nv, a.inscribed = h->robot_inscribed, a.circum = h->robot_circum;
    a.collision_dist = params->obstacle_collision_dist, a.collision_margin = params->margin_obstacle_collision;
    a.inflation_dist = params->obstacle_inflation_dist, a.inflation_margin = params->margin_obstacle_inflation;
    std::memcpy(a.verts, h->robot_verts, sizeof(a.verts));
    const unsigned blocks = static_cast<unsigned>((slots * EDGE_LANES + EDGE_THREADS - 1) / EDGE_THREADS);
    obstacle_edges_kernel<<<blocks, EDGE_THREADS, 0, h->stream>>>(a);
    PP_CUDA(h, cudaGetLastError());
    h->launches++;
    double *stage = reinterpret_cast<double *>(h->h_stage);
    PP_CUDA(h, cudaMemcpyAsync(stage, h->d_edge_out, sizeof(double) * 8 * slots, cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    for (size_t i = 0; i < slots; ++i)
    {
        error_out[2 * i] = stage[8 * i];
        error_out[2 * i + 1] = stage[8 * i + 1];
        if (jacobian_out)
            std::memcpy(jacobian_out + 6 * i, stage + 8 * i + 2, sizeof(double) * 6);
    }
    return PP_OK;
}
