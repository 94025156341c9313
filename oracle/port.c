/* TEST INFRASTRUCTURE ONLY (oracle). See port.h for the pinning status.
 *
 * Plain-C restatement of the reference algorithms on the hot path. Every function cites the
 * reference lines (relative to /root/reference) it follows. Built without FMA contraction
 * (-ffp-contract=off), like the reference (C++17, extensions off, no -march).
 */
#include "port.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---------------------------------------------------------------------------------------------
 * Footprint offsets: src/path_planner/src/planner/collision_checker.cpp:3-81.
 * boundary has 1 (circle radius), 2 (width, height) or 4 (x_min,y_min,x_max,y_max) numbers; hole
 * keeps only the rim. Offsets are emitted x-outer / y-inner like the reference loops.
 * ------------------------------------------------------------------------------------------- */
int port_footprint_points(double resolution, const float *boundary, int nb, int hole, int32_t *out_xy, int max_pts)
{
    int n = 0;
#define EMIT(i, j)                    \
    do                                \
    {                                 \
        if (n < max_pts)              \
        {                             \
            out_xy[2 * n] = (i);      \
            out_xy[2 * n + 1] = (j);  \
        }                             \
        n++;                          \
    } while (0)
    if (nb == 1)
    { /* :3-27 */
        int r = (int)ceil(boundary[0] / resolution);
        int r_sq = r * r;
        int rf = r - 1 > 0 ? r - 1 : 0;
        int rf_sq = rf * rf;
        for (int i = -r; i <= r; i++)
            for (int j = -r; j <= r; j++)
            {
                int d = i * i + j * j;
                if (hole ? (d <= r_sq && d >= rf_sq) : (d <= r_sq))
                    EMIT(i, j);
            }
    }
    else if (nb == 2)
    { /* :28-48 */
        int hw = (int)ceil(0.5 * boundary[0] / resolution);
        int hh = (int)ceil(0.5 * boundary[1] / resolution);
        for (int i = -hw; i <= hw; i++)
            for (int j = -hh; j <= hh; j++)
                if (!hole || i == -hw || i == hw || j == -hh || j == hh)
                    EMIT(i, j);
    }
    else if (nb == 4)
    { /* :49-71 */
        int x0 = (int)floor(boundary[0] / resolution);
        int y0 = (int)floor(boundary[1] / resolution);
        int x1 = (int)ceil(boundary[2] / resolution);
        int y1 = (int)ceil(boundary[3] / resolution);
        for (int i = x0; i <= x1; i++)
            for (int j = y0; j <= y1; j++)
                if (!hole || i == x0 || i == x1 || j == y0 || j == y1)
                    EMIT(i, j);
    }
#undef EMIT
    return n;
}

/* ---------------------------------------------------------------------------------------------
 * One footprint evaluation: collision_checker.cpp:144-172 (cost + flag) and :126-142 (early-exit
 * boolean, returned through *collision when non-NULL).
 * Float path: unqualified cos/sin/floor resolve to the float overloads once <math.h> is visible
 * (SURVEY.md §7 H1). Double path: only <cmath> visible -> double trig, double expression,
 * rounded to float on assignment to x_new / y_new.
 * The unsigned conversion of a negative float is what x86-64 gcc emits (64-bit truncation, low
 * 32 bits kept), which makes negative coordinates fail the "< size" test.
 * ------------------------------------------------------------------------------------------- */
static int footprint_eval(const uint8_t *cost, uint32_t W, uint32_t H, const int32_t *pts, int P, float x, float y,
                          float theta, int trig_double, float *cost_out, int *collision)
{
    int flag = 0;
    float temp = 0.0f;
    int coll = 0, coll_done = 0;
    float cf = 0.f, sf = 0.f;
    double cd = 0., sd = 0.;
    if (trig_double)
    {
        cd = cos((double)theta);
        sd = sin((double)theta);
    }
    else
    {
        cf = cosf(theta);
        sf = sinf(theta);
    }
    for (int p = 0; p < P; ++p)
    {
        float px = (float)pts[2 * p], py = (float)pts[2 * p + 1];
        float x_new, y_new;
        if (trig_double)
        {
            x_new = (float)((double)x + (double)px * cd - (double)py * sd);
            y_new = (float)((double)y + (double)px * sd + (double)py * cd);
        }
        else
        {
            x_new = x + px * cf - py * sf;
            y_new = y + px * sf + py * cf;
        }
        unsigned int xi = (unsigned int)(long long)floorf(x_new);
        unsigned int yi = (unsigned int)(long long)floorf(y_new);
        if (xi < W && yi < H)
        {
            uint8_t c = cost[(size_t)yi * W + xi];
            if (c >= 253 && !coll_done)
            {
                coll = 1;
                coll_done = 1;
            }
            if (c == 0)
                continue;
            if (c >= 253)
                flag = 1;
            temp = ((float)c > temp) ? (float)c : temp; /* std::max(static_cast<float>(c), temp) */
        }
    }
    *cost_out = temp;
    if (collision)
        *collision = coll;
    return flag;
}

/* Dense volume: SURVEY.md §8 a-1 definition, pose formed as at astar_hybrid.cpp:283 with
 * bin_size from motion_hybrid.cpp:10. */
int port_se2_volume(const uint8_t *cost, uint32_t W, uint32_t H, const int32_t *pts_xy, int P, uint32_t n_theta,
                    float ox, float oy, float ot, uint32_t t0, uint32_t t1, uint32_t y0, uint32_t y1, uint8_t *out,
                    int nthreads, int trig_double)
{
    const float bin = 2.0f * (float)M_PI / (float)n_theta;
    const long long rows = (long long)y1 - y0;
    const long long total = (long long)(t1 - t0) * rows;
#ifdef _OPENMP
    if (nthreads > 0)
        omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 4)
    for (long long job = 0; job < total; ++job)
    {
        uint32_t t = t0 + (uint32_t)(job / rows);
        uint32_t iy = y0 + (uint32_t)(job % rows);
        float tf = (float)t + ot;
        float theta = tf * bin;
        float y = (float)iy + oy;
        uint8_t *dst = out + ((size_t)(t - t0) * (size_t)rows + (iy - y0)) * W;
        for (uint32_t ix = 0; ix < W; ++ix)
        {
            float x = (float)ix + ox;
            float c;
            footprint_eval(cost, W, H, pts_xy, P, x, y, theta, trig_double, &c, NULL);
            dst[ix] = (uint8_t)c;
        }
    }
    return 0;
}

int port_footprint_cost_batch(const uint8_t *cost, uint32_t W, uint32_t H, const int32_t *pts_xy, int P,
                              const float *poses_xyt, uint32_t n, float *cost_out, uint8_t *flag_out,
                              uint8_t *collision_out, int trig_double)
{
    for (uint32_t i = 0; i < n; ++i)
    {
        float c;
        int coll;
        int f = footprint_eval(cost, W, H, pts_xy, P, poses_xyt[3 * i], poses_xyt[3 * i + 1], poses_xyt[3 * i + 2],
                               trig_double, &c, &coll);
        if (cost_out)
            cost_out[i] = c;
        if (flag_out)
            flag_out[i] = (uint8_t)f;
        if (collision_out)
            collision_out[i] = (uint8_t)coll;
    }
    return 0;
}

/* ---------------------------------------------------------------------------------------------
 * Binary min-heap on (key, index) with the element movement of libstdc++'s std::push_heap /
 * std::pop_heap / std::make_heap under ObstacleHeuristicComparator (astar_hybrid.h:23-29:
 * comp(a,b) = a.first > b.first), so that ties pop in the reference's order.
 * ------------------------------------------------------------------------------------------- */
typedef struct { float key; uint32_t idx; } HeapItem;
typedef struct { HeapItem *a; size_t n, cap; } Heap;

static void heap_reserve(Heap *h, size_t cap)
{
    if (cap > h->cap)
    {
        h->a = (HeapItem *)realloc(h->a, cap * sizeof(HeapItem));
        h->cap = cap;
    }
}
static void sift_up(HeapItem *a, size_t hole, size_t top, HeapItem v)
{
    while (hole > top)
    {
        size_t parent = (hole - 1) / 2;
        if (!(a[parent].key > v.key))
            break;
        a[hole] = a[parent];
        hole = parent;
    }
    a[hole] = v;
}
static void adjust(HeapItem *a, size_t hole, size_t len, HeapItem v)
{
    const size_t top = hole;
    size_t child = hole;
    while (len > 0 && child < (len - 1) / 2)
    {
        child = 2 * (child + 1);
        if (a[child].key > a[child - 1].key)
            child--;
        a[hole] = a[child];
        hole = child;
    }
    if ((len & 1) == 0 && len >= 2 && child == (len - 2) / 2)
    {
        child = 2 * (child + 1);
        a[hole] = a[child - 1];
        hole = child - 1;
    }
    sift_up(a, hole, top, v);
}
static void heap_push(Heap *h, float key, uint32_t idx)
{
    if (h->n == h->cap)
        heap_reserve(h, h->cap ? 2 * h->cap : 1024);
    HeapItem v = {key, idx};
    h->n++;
    sift_up(h->a, h->n - 1, 0, v);
}
static HeapItem heap_pop(Heap *h)
{
    HeapItem front = h->a[0];
    if (h->n > 1)
    {
        HeapItem v = h->a[h->n - 1];
        h->a[h->n - 1] = front;
        adjust(h->a, 0, h->n - 1, v);
    }
    h->n--;
    return front;
}
static void heap_make(Heap *h)
{
    size_t len = h->n;
    if (len < 2)
        return;
    size_t parent = (len - 2) / 2;
    for (;;)
    {
        HeapItem v = h->a[parent];
        adjust(h->a, parent, len, v);
        if (parent == 0)
            return;
        parent--;
    }
}

static float dist2d(uint32_t x1, uint32_t y1, uint32_t x2, uint32_t y2)
{ /* astar_hybrid.cpp:93-96 */
    return hypotf((float)x1 - (float)x2, (float)y1 - (float)y2);
}

/* CollisionChecker::getCost, collision_checker.cpp:203-214 */
static float get_cost(const uint8_t *cost, uint32_t W, uint32_t H, uint32_t x, uint32_t y)
{
    if (x < W && y < H)
        return (float)cost[(size_t)y * W + x];
    return 253.0f;
}

/* The expansion loop shared by cacheObstacleHeuristic (astar_hybrid.cpp:123-163) and the resume in
 * getObstacleHeuristic (:199-239). Returns the last popped index (or UINT32_MAX). */
static uint32_t expand(const uint8_t *cost, uint32_t W, uint32_t H, float cost_penalty, float *table, Heap *q,
                       uint32_t start_x, uint32_t start_y)
{
    static const int nb[8][2] = {{0, 1}, {1, 0}, {0, -1}, {-1, 0}, {1, 1}, {-1, -1}, {1, -1}, {-1, 1}}; /* :119 */
    const uint32_t start_index = start_y * W + start_x;
    const float sqrt_2 = sqrtf(2.0f);
    uint32_t idx = 0xFFFFFFFFu;
    while (q->n)
    {
        idx = heap_pop(q).idx;
        float c_cost = table[idx];
        if (c_cost > 0.0f)
            continue; /* already closed */
        c_cost = -c_cost;
        table[idx] = c_cost;
        uint32_t my_idx = idx / W, mx_idx = idx - my_idx * W;
        for (unsigned i = 0; i < 8; ++i)
        {
            uint32_t mx = mx_idx + (uint32_t)nb[i][0], my = my_idx + (uint32_t)nb[i][1];
            /* :141 -- unsigned compare: -1 wraps and is rejected; border cells are never entered */
            if (mx <= 0 || mx >= (uint32_t)((int)W - 1) || my <= 0 || my >= (uint32_t)((int)H - 1))
                continue;
            float c = get_cost(cost, W, H, mx, my);
            if (c >= 253.0)
                continue;
            uint32_t new_idx = my * W + mx;
            float existing = table[new_idx];
            if (existing <= 0.0f)
            {
                float travel = ((i <= 3) ? 1.0f : sqrt_2) * (1.0f + (cost_penalty * c / 252.0f)); /* :151 */
                float new_cost = c_cost + travel;
                if (existing == 0.0f || -existing > new_cost)
                {
                    table[new_idx] = -new_cost;
                    heap_push(q, new_cost + dist2d(mx, my, start_x, start_y), new_idx);
                }
            }
        }
        if (idx == start_index)
            break;
    }
    return idx;
}

/* cacheObstacleHeuristic (astar_hybrid.cpp:97-165) followed by getObstacleHeuristic
 * (:166-241) for each query; same contract as ref_heuristic in ref_harness.cpp. */
int port_heuristic(const uint8_t *cost, uint32_t W, uint32_t H, float start_xf, float start_yf, float goal_xf,
                   float goal_yf, float cost_penalty, const float *queries_xy, uint32_t n_queries, float *query_out,
                   float *table_out)
{
    const size_t size = (size_t)W * H;
    float *table = table_out;
    memset(table, 0, size * sizeof(float));
    Heap q = {0, 0, 0};
    heap_reserve(&q, 1024);
    uint32_t goal_x = (uint32_t)goal_xf, goal_y = (uint32_t)goal_yf;
    uint32_t start_x = (uint32_t)start_xf, start_y = (uint32_t)start_yf;
    const uint32_t goal_index = goal_y * W + goal_x, start_index = start_y * W + start_x;
    heap_push(&q, dist2d(start_x, start_y, goal_x, goal_y), goal_index);
    table[goal_index] = -0.00001f; /* :117 */
    uint32_t last = expand(cost, W, H, cost_penalty, table, &q, start_x, start_y);
    int ok = (last == start_index);

    const float sqrt_2 = sqrtf(2.0f);
    for (uint32_t k = 0; k < n_queries; ++k)
    {
        float qx = queries_xy[2 * k], qy = queries_xy[2 * k + 1];
        uint32_t sx = (uint32_t)qx, sy = (uint32_t)qy;
        uint32_t s_index = sy * W + sx;
        int dxi = abs((int)qx - (int)goal_xf), dyi = abs((int)qy - (int)goal_yf);
        float diagonal = (float)(dxi < dyi ? dxi : dyi) * sqrt_2 + (float)abs(dxi - dyi); /* :180 */
        float h;
        if (table[s_index] > 0.0f)
        {
            h = table[s_index] > diagonal ? table[s_index] : diagonal;
        }
        else
        {
            for (size_t j = 0; j < q.n; ++j)
            { /* re-key the open set to the new query cell, :184-191 */
                uint32_t c_idx = q.a[j].idx, c_y = c_idx / W, c_x = c_idx - c_y * W;
                q.a[j].key = -table[c_idx] + dist2d(c_x, c_y, sx, sy);
            }
            heap_make(&q);
            expand(cost, W, H, cost_penalty, table, &q, sx, sy);
            h = table[s_index] > diagonal ? table[s_index] : diagonal; /* std::max(a,b): a<b ? b : a */
        }
        if (query_out)
            query_out[k] = h;
    }
    free(q.a);
    return ok;
}

/* Least fixed point of d[v] = min_u fl(d[u] + travel(u->v)) with d[goal] = 1e-5f and the
 * reference's travel-cost expression (astar_hybrid.cpp:151): plain Dijkstra keyed by g only.
 * Float addition is monotone and travel > 0, so closing order is non-decreasing and closed values
 * are final. This is the value the GPU relaxation converges to; the reference's A* ordering can
 * close a cell 1 ulp above it. Unreached cells stay 0. */
int port_heuristic_fixed_point(const uint8_t *cost, uint32_t W, uint32_t H, uint32_t goal_x, uint32_t goal_y,
                               float cost_penalty, float *field_out)
{
    static const int nb[8][2] = {{0, 1}, {1, 0}, {0, -1}, {-1, 0}, {1, 1}, {-1, -1}, {1, -1}, {-1, 1}};
    const size_t size = (size_t)W * H;
    float *table = field_out;
    memset(table, 0, size * sizeof(float));
    Heap q = {0, 0, 0};
    heap_reserve(&q, 1024);
    const uint32_t goal_index = goal_y * W + goal_x;
    table[goal_index] = -0.00001f;
    heap_push(&q, 0.00001f, goal_index);
    const float sqrt_2 = sqrtf(2.0f);
    while (q.n)
    {
        HeapItem it = heap_pop(&q);
        uint32_t idx = it.idx;
        float c_cost = table[idx];
        if (c_cost > 0.0f)
            continue;
        if (-c_cost != it.key)
            continue; /* stale entry */
        c_cost = -c_cost;
        table[idx] = c_cost;
        uint32_t my_idx = idx / W, mx_idx = idx - my_idx * W;
        for (unsigned i = 0; i < 8; ++i)
        {
            uint32_t mx = mx_idx + (uint32_t)nb[i][0], my = my_idx + (uint32_t)nb[i][1];
            if (mx <= 0 || mx >= (uint32_t)((int)W - 1) || my <= 0 || my >= (uint32_t)((int)H - 1))
                continue;
            float c = get_cost(cost, W, H, mx, my);
            if (c >= 253.0)
                continue;
            uint32_t new_idx = my * W + mx;
            float existing = table[new_idx];
            if (existing <= 0.0f)
            {
                float travel = ((i <= 3) ? 1.0f : sqrt_2) * (1.0f + (cost_penalty * c / 252.0f));
                float new_cost = c_cost + travel;
                if (existing == 0.0f || -existing > new_cost)
                {
                    table[new_idx] = -new_cost;
                    heap_push(&q, new_cost, new_idx);
                }
            }
        }
    }
    free(q.a);
    return 0;
}

/* ---------------------------------------------------------------------------------------------
 * TEB candidate extraction: src/path_follower/src/teb_path_follower/teb_follower.cpp:140-169.
 * x outer / y inner scan, cells with cost >= 253 -> cell centre in world (costmap_2d mapToWorld),
 * box filter at 1.2*max_follow_range, behind-the-robot filter beyond obstacle_behind_ignore_range.
 * abs() on doubles resolves to std::abs(double) there (<math.h> is visible through the costmap
 * headers). Returns the number of candidates (may exceed max_pts; only max_pts are written).
 * ------------------------------------------------------------------------------------------- */
int port_extract_obstacles(const uint8_t *cost, uint32_t W, uint32_t H, double res, double origin_x, double origin_y,
                           double robot_x, double robot_y, double robot_theta, double max_follow_range,
                           double behind_ignore_range, double *pts_out, uint32_t max_pts)
{
    const double dir_x = cos(robot_theta), dir_y = sin(robot_theta); /* commons.h: Point2E::direction */
    const double max_distance = max_follow_range * 1.2;
    uint32_t n = 0;
    for (uint32_t i = 0; i < W; i++)
        for (uint32_t j = 0; j < H; j++)
        {
            uint8_t c = cost[(size_t)j * W + i];
            if (c < 253)
                continue;
            double wx = origin_x + (i + 0.5) * res, wy = origin_y + (j + 0.5) * res;
            double dx = wx - robot_x, dy = wy - robot_y;
            double distance = hypot(dx, dy);
            if (fabs(dx) > max_distance || fabs(dy) > max_distance)
                continue;
            if (distance > behind_ignore_range && (dx * dir_x + dy * dir_y < 0))
                continue;
            if (n < max_pts)
            {
                pts_out[2 * n] = wx;
                pts_out[2 * n + 1] = wy;
            }
            n++;
        }
    return (int)n;
}

/* ---------------------------------------------------------------------------------------------
 * TEB obstacle association: teb_optimizer.cpp:294-311 over nanoflann v1.6.2 semantics:
 * result set = { i : d2_i < radius_arg } (strict; RadiusResultSet::addPoint nanoflann.hpp:404-408;
 * the L2 adaptor compares squared distances, :627-638: d2 = (0 + dx*dx) + dy*dy, dx = q.x - p.x),
 * sorted ascending by d2 (:427-431), first min(K, n) kept. nanoflann's std::sort leaves the order
 * of equal distances unspecified; here ties are broken by index, and tests compare tie-tolerantly.
 * ------------------------------------------------------------------------------------------- */
typedef struct { double d2; uint32_t idx; } Hit;
static int hit_cmp(const void *a, const void *b)
{
    const Hit *x = (const Hit *)a, *y = (const Hit *)b;
    if (x->d2 < y->d2) return -1;
    if (x->d2 > y->d2) return 1;
    return (x->idx > y->idx) - (x->idx < y->idx);
}
int port_radius_topk(const double *pts_xy, uint32_t N, const double *q_xy, uint32_t Q, double radius_arg, uint32_t K,
                     uint32_t *idx_out, double *d2_out, uint32_t *cnt_out, uint32_t *found_out)
{
    Hit *hits = (Hit *)malloc(sizeof(Hit) * (N ? N : 1));
    for (uint32_t q = 0; q < Q; ++q)
    {
        uint32_t n = 0;
        for (uint32_t i = 0; i < N; ++i)
        {
            double d2 = 0.0;
            double dx = q_xy[2 * q] - pts_xy[2 * i];
            d2 += dx * dx;
            double dy = q_xy[2 * q + 1] - pts_xy[2 * i + 1];
            d2 += dy * dy;
            if (d2 < radius_arg)
            {
                hits[n].d2 = d2;
                hits[n].idx = i;
                n++;
            }
        }
        qsort(hits, n, sizeof(Hit), hit_cmp);
        if (found_out)
            found_out[q] = n;
        uint32_t keep = n < K ? n : K;
        cnt_out[q] = keep;
        for (uint32_t j = 0; j < keep; ++j)
        {
            idx_out[(size_t)q * K + j] = hits[j].idx;
            d2_out[(size_t)q * K + j] = hits[j].d2;
        }
    }
    free(hits);
    return 0;
}
