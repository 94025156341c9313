/* TEST INFRASTRUCTURE ONLY (oracle). CPU restatement of the costmap ingestion that runs BEFORE the
 * hot path: map image -> occupancy grid -> static layer -> inflation.
 *
 * Pinning status: PARITY UNPINNED for the two layer steps. The image thresholding follows the
 * reference's own loop (src/map_manager/src/map_manager_ros.cpp:93-108). The static and inflation
 * layers live in costmap_2d (ros-planning/navigation, the noetic 1.17.x series that README.md:15
 * installs; neither vendored under /root/reference nor version-pinned), so they are restated
 * from that package's published algorithm, named function by function below, and the reference
 * holds no test or fixture for them. What the reference does pin is their configuration
 * (src/path_planner/config/params.yaml:15-26).
 */
#include "port.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

enum
{
    FREE_SPACE = 0,
    INSCRIBED = 253,
    LETHAL = 254,
    NO_INFORMATION = 255
};

/* map_manager_ros.cpp:93-108. `image` is the 8-bit grayscale raster, row 0 at the top; the grid is
 * written bottom row first (MAP_IDX(width, j, height - i - 1), map_manager_ros.h:15). Thresholds
 * are compared with the 0..255 pixel value as they are read from the YAML (maps/sim.yaml has
 * free_thresh 100, occupied_thresh 200). */
int port_occupancy_from_image(const uint8_t *image, uint32_t W, uint32_t H, double occ_th, double free_th, int negate,
                              int8_t *occ_out)
{
    for (uint32_t i = 0; i < H; ++i)
        for (uint32_t j = 0; j < W; ++j)
        {
            int value = image[(size_t)i * W + j];
            if (negate)
                value = 255 - value;
            if (value > occ_th)
                value = 100;
            else if (value < free_th)
                value = 0;
            else
                value = (int)((value - free_th) / (occ_th - free_th) * 100);
            occ_out[(size_t)W * (H - i - 1) + j] = (int8_t)value;
        }
    return 0;
}

/* costmap_2d StaticLayer::interpretValue (plugins/static_layer.cpp). `value` is the occupancy
 * byte reinterpreted as unsigned char, unknown_cost_value likewise (-1 -> 255). */
static uint8_t interpret_value(uint8_t value, const port_layer_params *p)
{
    const uint8_t unknown = (uint8_t)p->unknown_cost_value;
    if (p->track_unknown_space && value == unknown)
        return NO_INFORMATION;
    else if (!p->track_unknown_space && value == unknown)
        return FREE_SPACE;
    else if (value >= (uint8_t)p->lethal_threshold)
        return LETHAL;
    else if (p->trinary_costmap)
        return FREE_SPACE;
    double scale = (double)value / (uint8_t)p->lethal_threshold;
    return (uint8_t)(scale * LETHAL);
}

/* StaticLayer::incomingMap + updateCosts with the layer first in the plugin list (params.yaml:24-26):
 * the master grid is overwritten with the interpreted values (updateWithTrueOverwrite). */
int port_static_layer(const int8_t *occ, uint32_t W, uint32_t H, const port_layer_params *p, uint8_t *cost_out)
{
    for (size_t i = 0; i < (size_t)W * H; ++i)
        cost_out[i] = interpret_value((uint8_t)occ[i], p);
    return 0;
}

/* Costmap2D::cellDistance (costmap_2d.cpp): max(0, ceil(world / resolution)). */
uint32_t port_cell_inflation_radius(const port_layer_params *p)
{
    double cells = ceil(p->inflation_radius / p->resolution);
    return (uint32_t)(cells > 0.0 ? cells : 0.0);
}

/* InflationLayer::computeCost (inflation_layer.h), distance in cells. */
static uint8_t compute_cost(double distance, const port_layer_params *p)
{
    uint8_t cost = 0;
    if (distance == 0)
        cost = LETHAL;
    else if (distance * p->resolution <= p->inscribed_radius)
        cost = INSCRIBED;
    else
    {
        double euclidean_distance = distance * p->resolution;
        double factor = exp(-1.0 * p->cost_scaling_factor * (euclidean_distance - p->inscribed_radius));
        cost = (uint8_t)((INSCRIBED - 1) * factor);
    }
    return cost;
}

/* InflationLayer::computeCaches: cached_distances_[i][j] = hypot(i, j), cached_costs_[i][j] =
 * computeCost(that), 0 <= i, j <= R + 1. table is (R+2) x (R+2), row i, column j. */
int port_inflation_cost_table(const port_layer_params *p, uint8_t *table)
{
    const uint32_t R = port_cell_inflation_radius(p);
    for (uint32_t i = 0; i <= R + 1; ++i)
        for (uint32_t j = 0; j <= R + 1; ++j)
            table[i * (R + 2) + j] = compute_cost(hypot((double)i, (double)j), p);
    return (int)R;
}

/* InflationLayer::updateCosts, the master-grid write. */
static void write_cost(uint8_t *master, size_t index, uint8_t cost, const port_layer_params *p)
{
    uint8_t old_cost = master[index];
    if (old_cost == NO_INFORMATION && (p->inflate_unknown ? (cost > FREE_SPACE) : (cost >= INSCRIBED)))
        master[index] = cost;
    else
        master[index] = old_cost > cost ? old_cost : cost;
}

/* ---------------------------------------------------------------------------------------------
 * costmap_2d's own propagation: InflationLayer::updateCosts + enqueue (plugins/inflation_layer.cpp)
 * over the whole grid. Cells are queued in bins keyed by their distance to the obstacle cell that
 * queued them (std::map<double, std::vector<CellData>>); bins are visited in ascending order, a
 * cell takes the cost of the first entry that reaches it and hands its source to its four
 * neighbours. An entry queued into a bin that was already visited is never processed.
 * ------------------------------------------------------------------------------------------- */
typedef struct
{
    uint32_t index, x, y, sx, sy;
} CellData;

typedef struct
{
    CellData *v;
    size_t n, cap;
} Bin;

static void bin_push(Bin *b, CellData c)
{
    if (b->n == b->cap)
    {
        b->cap = b->cap ? b->cap * 2 : 1024;
        b->v = (CellData *)realloc(b->v, b->cap * sizeof(CellData));
    }
    b->v[b->n++] = c;
}

static int cmp_double(const void *a, const void *b)
{
    double x = *(const double *)a, y = *(const double *)b;
    return (x > y) - (x < y);
}

int port_inflate_propagate(uint8_t *master, uint32_t W, uint32_t H, const port_layer_params *p)
{
    const uint32_t R = port_cell_inflation_radius(p);
    const uint32_t T = R + 2;
    uint8_t *costs = (uint8_t *)malloc(T * T);
    double *dists = (double *)malloc(sizeof(double) * T * T);
    port_inflation_cost_table(p, costs);
    for (uint32_t i = 0; i < T; ++i)
        for (uint32_t j = 0; j < T; ++j)
            dists[i * T + j] = hypot((double)i, (double)j);
    /* the distinct keys the map can hold, ascending; level_of[i][j] = rank of hypot(i, j) */
    double *keys = (double *)malloc(sizeof(double) * T * T);
    memcpy(keys, dists, sizeof(double) * T * T);
    qsort(keys, T * T, sizeof(double), cmp_double);
    uint32_t n_keys = 0;
    for (uint32_t k = 0; k < T * T; ++k)
        if (n_keys == 0 || keys[k] != keys[n_keys - 1])
            keys[n_keys++] = keys[k];
    uint32_t *level_of = (uint32_t *)malloc(sizeof(uint32_t) * T * T);
    for (uint32_t k = 0; k < T * T; ++k)
    {
        uint32_t lo = 0;
        while (keys[lo] != dists[k])
            ++lo;
        level_of[k] = lo;
    }
    Bin *bins = (Bin *)calloc(n_keys, sizeof(Bin));
    uint8_t *seen = (uint8_t *)calloc((size_t)W * H, 1);

    for (uint32_t j = 0; j < H; ++j)
        for (uint32_t i = 0; i < W; ++i)
        {
            size_t index = (size_t)j * W + i;
            if (master[index] == LETHAL)
            {
                CellData c = {(uint32_t)index, i, j, i, j};
                bin_push(&bins[0], c);
            }
        }

#define ENQUEUE(IDX, MX, MY, SX, SY)                                              \
    do                                                                            \
    {                                                                             \
        if (!seen[IDX])                                                           \
        {                                                                         \
            uint32_t ddx = (MX) > (SX) ? (MX) - (SX) : (SX) - (MX);               \
            uint32_t ddy = (MY) > (SY) ? (MY) - (SY) : (SY) - (MY);               \
            if (ddx < T && ddy < T && !(dists[ddx * T + ddy] > (double)R))        \
            {                                                                     \
                CellData e = {(uint32_t)(IDX), (MX), (MY), (SX), (SY)};           \
                bin_push(&bins[level_of[ddx * T + ddy]], e);                      \
            }                                                                     \
        }                                                                         \
    } while (0)

    for (uint32_t b = 0; b < n_keys; ++b)
        for (size_t i = 0; i < bins[b].n; ++i)
        {
            const CellData cell = bins[b].v[i];
            if (seen[cell.index])
                continue;
            seen[cell.index] = 1;
            uint32_t ddx = cell.x > cell.sx ? cell.x - cell.sx : cell.sx - cell.x;
            uint32_t ddy = cell.y > cell.sy ? cell.y - cell.sy : cell.sy - cell.y;
            write_cost(master, cell.index, costs[ddx * T + ddy], p);
            if (cell.x > 0)
                ENQUEUE(cell.index - 1, cell.x - 1, cell.y, cell.sx, cell.sy);
            if (cell.y > 0)
                ENQUEUE(cell.index - W, cell.x, cell.y - 1, cell.sx, cell.sy);
            if (cell.x < W - 1)
                ENQUEUE(cell.index + 1, cell.x + 1, cell.y, cell.sx, cell.sy);
            if (cell.y < H - 1)
                ENQUEUE(cell.index + W, cell.x, cell.y + 1, cell.sx, cell.sy);
        }
#undef ENQUEUE
    for (uint32_t b = 0; b < n_keys; ++b)
        free(bins[b].v);
    free(bins);
    free(seen);
    free(level_of);
    free(keys);
    free(dists);
    free(costs);
    return 0;
}

/* ---------------------------------------------------------------------------------------------
 * The definition the GPU path is held to: every cell within R cells (Euclidean, d <= R) of a
 * lethal cell takes cached_costs_ of its NEAREST lethal cell -- the value costmap_2d's propagation
 * yields wherever its four-neighbour hand-over reaches a cell from its nearest obstacle, which is
 * almost everywhere; where three obstacle cells shadow one another (e.g. lethal cells at (2,2),
 * (3,0), (0,3) around (0,0)) the propagation assigns a farther source and a lower cost. The
 * nearest-obstacle cost is therefore >= the propagated cost in every cell. Brute force over the
 * (2R+1)^2 window; for tests only.
 * ------------------------------------------------------------------------------------------- */
int port_inflate_nearest(uint8_t *master, uint32_t W, uint32_t H, const port_layer_params *p)
{
    const int R = (int)port_cell_inflation_radius(p);
    const int T = R + 2;
    uint8_t *costs = (uint8_t *)malloc((size_t)T * T);
    port_inflation_cost_table(p, costs);
    uint8_t *lethal = (uint8_t *)malloc((size_t)W * H);
    for (size_t i = 0; i < (size_t)W * H; ++i)
        lethal[i] = master[i] == LETHAL;
    for (int y = 0; y < (int)H; ++y)
        for (int x = 0; x < (int)W; ++x)
        {
            int best = R * R + 1, bi = 0, bj = 0;
            for (int dy = -R; dy <= R; ++dy)
            {
                int yy = y + dy;
                if (yy < 0 || yy >= (int)H)
                    continue;
                for (int dx = -R; dx <= R; ++dx)
                {
                    int xx = x + dx;
                    if (xx < 0 || xx >= (int)W || !lethal[(size_t)yy * W + xx])
                        continue;
                    int d2 = dx * dx + dy * dy;
                    if (d2 < best)
                    {
                        best = d2;
                        bi = abs(dx);
                        bj = abs(dy);
                    }
                }
            }
            if (best <= R * R)
                write_cost(master, (size_t)y * W + x, costs[bi * T + bj], p);
        }
    free(lethal);
    free(costs);
    return 0;
}
