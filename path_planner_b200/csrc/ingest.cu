// N2: costmap ingestion on the device: map image -> occupancy grid ->
// static layer + inflation -> the handle's costmap, with dirty-rectangle propagation into the SE2
// volume refresh. Replaces MapManagerROS::loadMap's threshold loop
// (src/map_manager/src/map_manager_ros.cpp:93-108) and the two costmap_2d layers configured at
// src/path_planner/config/params.yaml:15-26.
//
// One kernel does both layers: a CTA stages the interpreted (static-layer) cost of a 128x32 tile
// plus an R-cell halo in shared memory, computes for every column the distance to the nearest
// lethal cell above/below (two sequential sweeps per column), then every cell takes the minimum of
// dx^2 + g(x+dx)^2 over |dx| <= R -- the exact squared distance to the nearest lethal cell within R
// -- and looks its cost up in a table indexed by squared distance that the HOST filled with its
// own libm exp()/hypot(), so no device transcendental touches the result. Tiles without a lethal
// cell in reach (most of a map) skip both passes. The work is one byte in and one byte out per cell
// (33.5 MB at 4096^2, ~5 us of HBM time); the kernel is bound by the 2R+1 shared-memory bytes it
// reads per cell and is not worth more engineering than this.
#include "pp_internal.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace
{
constexpr int IT_W = 128, IT_H = 32, IT_THREADS = 256, MAX_R = 32;
constexpr int PITCH = IT_W + 2 * MAX_R;
constexpr uint8_t FREE_SPACE = 0, INSCRIBED = 253, LETHAL = 254, NO_INFORMATION = 255;

struct IngestArgs
{
    const int8_t *occ;
    uint8_t *cost;
    const uint8_t *infl;  // cost by squared distance, 0..R*R
    uint32_t W, H;          // the costmap that is written
    uint32_t occ_W, occ_H;  // the occupancy grid it is cut from: costmap cell (x, y) is occupancy cell (x + ox, y + oy)
    int ox, oy;             // (0, 0 and equal sizes for the whole map; a rolling window otherwise)
    uint32_t rx0, ry0, rx1, ry1;  // cells to write, half-open
    uint32_t tile_x0, tile_y0;    // tile index of block (0, 0)
    int R;
    uint8_t unknown, lethal_threshold;
    int track_unknown, trinary, inflate_unknown;
};

// costmap_2d StaticLayer::interpretValue
__device__ __forceinline__ uint8_t interpret(uint8_t v, const IngestArgs &a)
{
    if (v == a.unknown)
        return a.track_unknown ? NO_INFORMATION : FREE_SPACE;
    if (v >= a.lethal_threshold)
        return LETHAL;
    if (a.trinary)
        return FREE_SPACE;
    const double scale = __ddiv_rn(static_cast<double>(v), static_cast<double>(a.lethal_threshold));
    return static_cast<uint8_t>(__dmul_rn(scale, 254.0));
}

__global__ void __launch_bounds__(IT_THREADS) ingest_kernel(const IngestArgs a)
{
    __shared__ uint8_t s_stat[(IT_H + 2 * MAX_R) * PITCH];  // static-layer cost, tile + halo
    __shared__ uint8_t s_g[IT_H * PITCH];                    // per column: rows to the nearest lethal cell, 255 = none within R
    __shared__ uint8_t s_tab[MAX_R * MAX_R + 1];
    const int R = a.R;
    const int x_base = static_cast<int>((a.tile_x0 + blockIdx.x) * IT_W) - R;
    const int y_base = static_cast<int>((a.tile_y0 + blockIdx.y) * IT_H) - R;
    const int cols = IT_W + 2 * R, rows = IT_H + 2 * R;

    for (int i = threadIdx.x; i <= R * R; i += IT_THREADS)
        s_tab[i] = a.infl[i];
    int any_lethal = 0;
    constexpr int BATCH = 8;  // global loads in flight per thread: the staging loop is latency-bound otherwise
    for (int i0 = threadIdx.x; i0 < cols * rows; i0 += BATCH * IT_THREADS)
    {
        int8_t raw[BATCH];
        bool inside[BATCH];
#pragma unroll
        for (int b = 0; b < BATCH; ++b)
        {
            const int i = i0 + b * IT_THREADS;
            const int r = i / cols, c = i - r * cols;
            const int x = x_base + c, y = y_base + r;
            inside[b] = i < cols * rows && x >= 0 && y >= 0 && x < static_cast<int>(a.W) && y < static_cast<int>(a.H);
            // a window cell beyond the occupancy grid keeps the master grid's default value, which is what
            // the static layer makes of an unknown cell
            const int xs = x + a.ox, ys = y + a.oy;
            const bool in_occ = xs >= 0 && ys >= 0 && xs < static_cast<int>(a.occ_W) && ys < static_cast<int>(a.occ_H);
            raw[b] = inside[b] && in_occ ? __ldg(a.occ + static_cast<size_t>(ys) * a.occ_W + xs) : static_cast<int8_t>(a.unknown);
        }
#pragma unroll
        for (int b = 0; b < BATCH; ++b)
        {
            const int i = i0 + b * IT_THREADS;
            if (i >= cols * rows)
                break;
            const int r = i / cols, c = i - r * cols;
            // cells outside the costmap hold no obstacle (a window is inflated from its own cells only)
            const uint8_t v = inside[b] ? interpret(static_cast<uint8_t>(raw[b]), a) : FREE_SPACE;
            s_stat[r * PITCH + c] = v;
            any_lethal |= (v == LETHAL);
        }
    }
    any_lethal = __syncthreads_or(any_lethal);

    if (any_lethal)
    {
        for (int c = threadIdx.x; c < cols; c += IT_THREADS)
        {
            int d = 255;
            for (int r = 0; r < rows; ++r)
            {
                d = s_stat[r * PITCH + c] == LETHAL ? 0 : min(d + 1, 255);
                if (r >= R && r < R + IT_H)
                    s_g[(r - R) * PITCH + c] = static_cast<uint8_t>(d);
            }
            d = 255;
            for (int r = rows - 1; r >= R; --r)
            {
                d = s_stat[r * PITCH + c] == LETHAL ? 0 : min(d + 1, 255);
                if (r < R + IT_H)
                {
                    const int g = min(static_cast<int>(s_g[(r - R) * PITCH + c]), d);
                    s_g[(r - R) * PITCH + c] = static_cast<uint8_t>(g > R ? 255 : g);
                }
            }
        }
        __syncthreads();
    }

    const int R2 = R * R;
    for (int i = threadIdx.x; i < IT_W * IT_H; i += IT_THREADS)
    {
        const int r = i / IT_W, c = i - r * IT_W;
        const uint32_t x = static_cast<uint32_t>(x_base + R + c), y = static_cast<uint32_t>(y_base + R + r);
        if (x < a.rx0 || x >= a.rx1 || y < a.ry0 || y >= a.ry1)
            continue;
        uint8_t out = s_stat[(r + R) * PITCH + c + R];
        if (any_lethal)
        {
            int best = R2 + 1;
            const uint8_t *g = &s_g[r * PITCH + c + R];
            for (int dx = -R; dx <= R; ++dx)
            {
                const int gv = g[dx];
                best = min(best, gv * gv + dx * dx);
            }
            if (best <= R2)
            {
                // costmap_2d InflationLayer::updateCosts, the master-grid write
                const uint8_t cost = s_tab[best];
                if (out == NO_INFORMATION && (a.inflate_unknown ? cost > FREE_SPACE : cost >= INSCRIBED))
                    out = cost;
                else
                    out = max(out, cost);
            }
        }
        a.cost[static_cast<size_t>(y) * a.W + x] = out;
    }
}

// map_manager_ros.cpp:93-108, one thread per pixel
__global__ void threshold_kernel(const uint8_t *image, int8_t *occ, uint32_t W, uint32_t H, double occ_th, double free_th,
                                 int negate)
{
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y * blockDim.y + threadIdx.y;
    if (j >= W || i >= H)
        return;
    int value = image[static_cast<size_t>(i) * W + j];
    if (negate)
        value = 255 - value;
    if (value > occ_th)
        value = 100;
    else if (value < free_th)
        value = 0;
    else
        value = static_cast<int>(__dmul_rn(__ddiv_rn(__dsub_rn(value, free_th), __dsub_rn(occ_th, free_th)), 100.0));
    occ[static_cast<size_t>(W) * (H - i - 1) + j] = static_cast<int8_t>(value);
}

// costmap_2d InflationLayer::computeCost, distance in cells
uint8_t compute_cost(double distance, const pp_layer_params &p)
{
    if (distance == 0)
        return LETHAL;
    if (distance * p.resolution <= p.inscribed_radius)
        return INSCRIBED;
    const double euclidean = distance * p.resolution;
    const double factor = std::exp(-1.0 * p.cost_scaling_factor * (euclidean - p.inscribed_radius));
    return static_cast<uint8_t>((INSCRIBED - 1) * factor);
}

pp_status set_occupancy_size(pp_handle h, uint32_t W, uint32_t H, const char *who)
{
    PP_REQUIRE(h, W > 0 && H > 0, PP_ERR_INVALID, "%s: empty grid", who);
    PP_REQUIRE(h, W <= 32768 && H <= 32768, PP_ERR_RANGE, "%s: %ux%u exceeds 32768x32768", who, W, H);
    PP_CUDA(h, cudaSetDevice(h->device));
    PP_CUDA(h, pp_reserve(h->d_occ, h->occ_cap, static_cast<size_t>(W) * H));
    h->occ_W = W;
    h->occ_H = H;
    h->occ_all = true;
    h->occ_dirty = false;
    return PP_OK;
}
}  // namespace

pp_status pp_occupancy_from_image(pp_handle h, const uint8_t *image, uint32_t W, uint32_t H, double occupied_thresh,
                                  double free_thresh, int negate)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, image, PP_ERR_INVALID, "pp_occupancy_from_image: null image");
    pp_status s = set_occupancy_size(h, W, H, "pp_occupancy_from_image");
    if (s != PP_OK)
        return s;
    const size_t bytes = static_cast<size_t>(W) * H;
    // pinned staging copy, async H2D, then one thread per pixel
    s = pp_ensure_stage(h, bytes);
    if (s != PP_OK)
        return s;
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    std::memcpy(h->h_stage, image, bytes);
    PP_CUDA(h, pp_reserve(h->d_img, h->img_cap, bytes));
    PP_CUDA(h, cudaMemcpyAsync(h->d_img, h->h_stage, bytes, cudaMemcpyHostToDevice, h->stream));
    const dim3 block(64, 4), grid((W + 63) / 64, (H + 3) / 4);
    threshold_kernel<<<grid, block, 0, h->stream>>>(h->d_img, h->d_occ, W, H, occupied_thresh, free_thresh, negate);
    PP_CUDA(h, cudaGetLastError());
    h->launches++;
    return PP_OK;
}

pp_status pp_occupancy_upload(pp_handle h, const int8_t *occ, uint32_t W, uint32_t H)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, occ, PP_ERR_INVALID, "pp_occupancy_upload: null grid");
    pp_status s = set_occupancy_size(h, W, H, "pp_occupancy_upload");
    if (s != PP_OK)
        return s;
    const size_t bytes = static_cast<size_t>(W) * H;
    s = pp_ensure_stage(h, bytes);
    if (s != PP_OK)
        return s;
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    std::memcpy(h->h_stage, occ, bytes);
    PP_CUDA(h, cudaMemcpyAsync(h->d_occ, h->h_stage, bytes, cudaMemcpyHostToDevice, h->stream));
    return PP_OK;
}

pp_status pp_occupancy_update_rect(pp_handle h, uint32_t x0, uint32_t y0, uint32_t w, uint32_t hgt, const int8_t *src,
                                   size_t src_pitch)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_occ, PP_ERR_STATE, "pp_occupancy_update_rect: no occupancy grid yet");
    PP_REQUIRE(h, src && w > 0 && hgt > 0 && src_pitch >= w, PP_ERR_INVALID, "pp_occupancy_update_rect: bad source");
    PP_REQUIRE(h, static_cast<uint64_t>(x0) + w <= h->occ_W && static_cast<uint64_t>(y0) + hgt <= h->occ_H, PP_ERR_INVALID,
               "pp_occupancy_update_rect: rect (%u,%u)+(%u,%u) outside %ux%u", x0, y0, w, hgt, h->occ_W, h->occ_H);
    PP_CUDA(h, cudaSetDevice(h->device));
    const size_t bytes = static_cast<size_t>(w) * hgt;
    pp_status s = pp_ensure_stage(h, bytes);
    if (s != PP_OK)
        return s;
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    for (uint32_t r = 0; r < hgt; ++r)
        std::memcpy(h->h_stage + static_cast<size_t>(r) * w, src + static_cast<size_t>(r) * src_pitch, w);
    PP_CUDA(h, cudaMemcpy2DAsync(h->d_occ + static_cast<size_t>(y0) * h->occ_W + x0, h->occ_W, h->h_stage, w, w, hgt,
                                 cudaMemcpyHostToDevice, h->stream));
    if (!h->occ_dirty)
    {
        h->occ_dirty = true;
        h->occ_x0 = x0;
        h->occ_y0 = y0;
        h->occ_x1 = x0 + w;
        h->occ_y1 = y0 + hgt;
    }
    else
    {
        h->occ_x0 = std::min(h->occ_x0, x0);
        h->occ_y0 = std::min(h->occ_y0, y0);
        h->occ_x1 = std::max(h->occ_x1, x0 + w);
        h->occ_y1 = std::max(h->occ_y1, y0 + hgt);
    }
    return PP_OK;
}

pp_status pp_occupancy_download(pp_handle h, int8_t *occ_out)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_occ && occ_out, PP_ERR_STATE, "pp_occupancy_download: no occupancy grid or null output");
    PP_CUDA(h, cudaSetDevice(h->device));
    PP_CUDA(h, cudaMemcpyAsync(occ_out, h->d_occ, static_cast<size_t>(h->occ_W) * h->occ_H, cudaMemcpyDeviceToHost,
                               h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    return PP_OK;
}

namespace
{
// Checks the layer parameters, fills the handle's cost-by-squared-distance table when they changed; returns the
// clamped parameters, the inflation radius in cells and whether the table is new.
pp_status prepare_layers(pp_handle h, const pp_layer_params *params, const char *who, pp_layer_params &p, int &R, bool &params_changed)
{
    PP_REQUIRE(h, params, PP_ERR_INVALID, "%s: null params", who);
    p = *params;
    PP_REQUIRE(h, p.resolution > 0.0 && p.inflation_radius >= 0.0 && p.inscribed_radius >= 0.0, PP_ERR_INVALID,
               "%s: resolution must be > 0, radii >= 0", who);
    p.lethal_threshold = std::max(std::min(p.lethal_threshold, 100), 0);  // as StaticLayer clamps it
    // Costmap2D::cellDistance
    const double cells = std::max(0.0, std::ceil(p.inflation_radius / p.resolution));
    PP_REQUIRE(h, cells <= MAX_R, PP_ERR_RANGE, "%s: inflation radius of %.0f cells exceeds %d", who, cells, MAX_R);
    R = static_cast<int>(cells);
    PP_CUDA(h, cudaSetDevice(h->device));

    const pp_layer_params &q = h->layer;
    params_changed = !h->layer_valid || q.track_unknown_space != p.track_unknown_space ||
                     q.lethal_threshold != p.lethal_threshold || q.unknown_cost_value != p.unknown_cost_value ||
                     q.trinary_costmap != p.trinary_costmap || q.inflate_unknown != p.inflate_unknown ||
                     q.resolution != p.resolution || q.inscribed_radius != p.inscribed_radius ||
                     q.inflation_radius != p.inflation_radius || q.cost_scaling_factor != p.cost_scaling_factor;
    if (params_changed)
    {
        // cached_costs_[i][j] = computeCost(hypot(i, j)) re-indexed by i*i + j*j
        std::vector<int> tab(static_cast<size_t>(R) * R + 1, -1);
        for (int i = 0; i <= R; ++i)
            for (int j = 0; j <= R; ++j)
            {
                const int d2 = i * i + j * j;
                if (d2 > R * R)
                    continue;
                const int c = compute_cost(std::hypot(static_cast<double>(i), static_cast<double>(j)), p);
                PP_REQUIRE(h, tab[d2] < 0 || tab[d2] == c, PP_ERR_RANGE,
                           "%s: inflation cost is not a function of the squared distance at (%d,%d)", who, i, j);
                tab[d2] = c;
            }
        std::vector<uint8_t> bytes(tab.size());
        for (size_t i = 0; i < tab.size(); ++i)
            bytes[i] = static_cast<uint8_t>(std::max(tab[i], 0));
        PP_CUDA(h, pp_reserve(h->d_infl, h->infl_cap, static_cast<size_t>(MAX_R) * MAX_R + 1));
        PP_CUDA(h, cudaStreamSynchronize(h->stream));  // an earlier launch may still read the table
        PP_CUDA(h, cudaMemcpy(h->d_infl, bytes.data(), bytes.size(), cudaMemcpyHostToDevice));
        h->layer = p;
        h->layer_valid = true;
    }
    return PP_OK;
}

// the handle's costmap becomes a W x H grid that is about to be written in full
pp_status resize_costmap(pp_handle h, uint32_t W, uint32_t H)
{
    if (W != h->W || H != h->H)
        pp_se2_plan_free(h);
    PP_CUDA(h, pp_reserve(h->d_cost, h->cost_cap, static_cast<size_t>(W) * H));
    h->W = W;
    h->H = H;
    h->dirty = false;  // like a full upload: the volume is to be rebuilt
    h->vol_valid = false;
    return PP_OK;
}

void fill_layer_args(IngestArgs &a, const pp_layer_params &p, int R)
{
    a.R = R;
    a.unknown = static_cast<uint8_t>(p.unknown_cost_value);
    a.lethal_threshold = static_cast<uint8_t>(p.lethal_threshold);
    a.track_unknown = p.track_unknown_space, a.trinary = p.trinary_costmap, a.inflate_unknown = p.inflate_unknown;
}
}  // namespace

pp_status pp_costmap_from_occupancy(pp_handle h, const pp_layer_params *params)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_occ, PP_ERR_STATE, "pp_costmap_from_occupancy: no occupancy grid yet");
    pp_layer_params p;
    int R = 0;
    bool params_changed = false;
    pp_status st = prepare_layers(h, params, "pp_costmap_from_occupancy", p, R, params_changed);
    if (st != PP_OK)
        return st;

    const uint32_t W = h->occ_W, H = h->occ_H;
    const bool full = params_changed || h->occ_all || !h->d_cost || h->W != W || h->H != H || h->cost_is_window;
    if (!full && !h->occ_dirty)
        return PP_OK;
    IngestArgs a{};
    if (full)
    {
        st = resize_costmap(h, W, H);
        if (st != PP_OK)
            return st;
        h->cost_is_window = false;
        a.rx0 = 0, a.ry0 = 0, a.rx1 = W, a.ry1 = H;
    }
    else
    {
        a.rx0 = h->occ_x0 > static_cast<uint32_t>(R) ? h->occ_x0 - R : 0;
        a.ry0 = h->occ_y0 > static_cast<uint32_t>(R) ? h->occ_y0 - R : 0;
        a.rx1 = std::min(W, h->occ_x1 + R);
        a.ry1 = std::min(H, h->occ_y1 + R);
        if (!h->dirty)
        {
            h->dirty = true;
            h->dirty_x0 = a.rx0, h->dirty_y0 = a.ry0, h->dirty_x1 = a.rx1, h->dirty_y1 = a.ry1;
        }
        else
        {
            h->dirty_x0 = std::min(h->dirty_x0, a.rx0);
            h->dirty_y0 = std::min(h->dirty_y0, a.ry0);
            h->dirty_x1 = std::max(h->dirty_x1, a.rx1);
            h->dirty_y1 = std::max(h->dirty_y1, a.ry1);
        }
    }
    a.occ = h->d_occ;
    a.cost = h->d_cost;
    a.infl = h->d_infl;
    a.W = a.occ_W = W, a.H = a.occ_H = H;
    a.ox = a.oy = 0;
    a.tile_x0 = a.rx0 / IT_W, a.tile_y0 = a.ry0 / IT_H;
    fill_layer_args(a, p, R);
    const dim3 grid((a.rx1 + IT_W - 1) / IT_W - a.tile_x0, (a.ry1 + IT_H - 1) / IT_H - a.tile_y0);
    ingest_kernel<<<grid, IT_THREADS, 0, h->stream>>>(a);
    PP_CUDA(h, cudaGetLastError());
    h->launches++;
    h->occ_dirty = false;
    h->occ_all = false;
    return PP_OK;
}

// The follower's rolling window (path_follower/config/params.yaml:12-24): see include/pp_b200.h
pp_status pp_costmap_window_from_occupancy(pp_handle h, pp_handle src, int32_t origin_x, int32_t origin_y, uint32_t win_w, uint32_t win_h,
                                           const pp_layer_params *params)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, src && src->d_occ, PP_ERR_STATE, "pp_costmap_window_from_occupancy: the source handle holds no occupancy grid");
    PP_REQUIRE(h, src->device == h->device, PP_ERR_INVALID, "pp_costmap_window_from_occupancy: handles on devices %d and %d", h->device,
               src->device);
    PP_REQUIRE(h, win_w > 0 && win_h > 0 && win_w <= 32768 && win_h <= 32768, PP_ERR_RANGE,
               "pp_costmap_window_from_occupancy: window of %ux%u cells", win_w, win_h);
    PP_REQUIRE(h, origin_x > -(1 << 30) && origin_x < (1 << 30) && origin_y > -(1 << 30) && origin_y < (1 << 30), PP_ERR_RANGE,
               "pp_costmap_window_from_occupancy: origin (%d, %d)", origin_x, origin_y);
    pp_layer_params p;
    int R = 0;
    bool params_changed = false;
    pp_status st = prepare_layers(h, params, "pp_costmap_window_from_occupancy", p, R, params_changed);
    if (st != PP_OK)
        return st;
    if (src != h)
        PP_CUDA(h, cudaStreamSynchronize(src->stream));  // the grid's uploads run on the source handle's stream
    st = resize_costmap(h, win_w, win_h);
    if (st != PP_OK)
        return st;
    h->cost_is_window = true;  // pp_costmap_from_occupancy on this handle starts over
    IngestArgs a{};
    a.occ = src->d_occ;
    a.cost = h->d_cost;
    a.infl = h->d_infl;
    a.W = win_w, a.H = win_h;
    a.occ_W = src->occ_W, a.occ_H = src->occ_H;
    a.ox = origin_x, a.oy = origin_y;
    a.rx0 = 0, a.ry0 = 0, a.rx1 = win_w, a.ry1 = win_h;
    a.tile_x0 = a.tile_y0 = 0;
    fill_layer_args(a, p, R);
    const dim3 grid((win_w + IT_W - 1) / IT_W, (win_h + IT_H - 1) / IT_H);
    ingest_kernel<<<grid, IT_THREADS, 0, h->stream>>>(a);
    PP_CUDA(h, cudaGetLastError());
    h->launches++;
    return PP_OK;
}

// Costmap2D::updateOrigin (the window's own grid starts at 0 and moves in whole cells, truncating toward zero)
// followed by StaticLayer::updateCosts' worldToMap of the window's first cell centre
int32_t pp_window_origin_cell(double robot, double window_size_m, double resolution, double map_origin)
{
    const double wanted = robot - window_size_m / 2.0;
    const int cell_o = static_cast<int>(wanted / resolution);
    const double grid_o = 0.0 + cell_o * resolution;
    const double centre = grid_o + (0 + 0.5) * resolution;
    return static_cast<int32_t>(std::floor((centre - map_origin) / resolution));
}
