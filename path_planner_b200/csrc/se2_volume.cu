// K1: SE2 footprint-collision volume (see include/pp_b200.h).
//
// se2_generic_kernel is the exact table-free path: it evaluates the reference expression
//   x_new = (x + px*cos) - py*sin ,  y_new = (y + px*sin) + py*cos ,  idx = floor(.)
// (collision_checker.cpp:155-158) per footprint point and per tile column / tile row with
// separately rounded IEEE operations (__fmul_rn/__fadd_rn/__fsub_rn: no FMA contraction, like the
// reference build) and cos/sin supplied by the host libm, so indices are bit-identical to the
// reference for every cell, including the cells where float rounding makes the integer offset
// depend on the position. Because x_new depends only on (column, point, bin)
// and y_new only on (row, point, bin), the offsets are tabulated once per tile in shared memory.
// Out-of-map samples are ignored by the reference (:160); here the tile is staged with zeros
// outside the map, which is the identity of max over uint8.
#include "pp_internal.h"
#include "se2_plan.h"

#include <cmath>
#include <algorithm>

namespace
{
constexpr int GT_W = 64;        // tile width  (outputs)
constexpr int GT_H = 16;        // tile height (outputs); the work list of the tiled fast path is in these 64 x 16 tiles
constexpr int GT_THREADS = 256;
constexpr int GT_ROWS_PER_THREAD = GT_H / (GT_THREADS / GT_W);  // 8
constexpr int GT_PCHUNK = 256;  // footprint points tabulated per pass

struct GenericArgs
{
    const uint8_t *cost;
    uint8_t *vol[8];  // destination volumes: own first, then mapped peers (fused exchange)
    int n_vol;
    const float2 *trig;
    const int32_t *fp;
    uint32_t W, H, P;
    int R;
    float ox, oy;
    uint32_t t_begin, t_end, t_per_cta;
    uint32_t x0, y0, x1, y1;
    const uint32_t *worklist;  // optional: pairs (tile_x | tile_y << 16, t)
};

__global__ void __launch_bounds__(GT_THREADS) se2_generic_kernel(GenericArgs a)
{
    extern __shared__ __align__(16) uint8_t smem[];
    const int R = a.R;
    const int pitch = (GT_W + 2 * R + 3) & ~3;
    const int rows = GT_H + 2 * R;
    uint8_t *tile = smem;
    int8_t *xo = reinterpret_cast<int8_t *>(smem + ((pitch * rows + 15) & ~15));
    int8_t *yo = xo + GT_PCHUNK * GT_W;

    uint32_t tile_x, tile_y, t_lo, t_hi;
    if (a.worklist)
    {
        const uint32_t w = a.worklist[2 * blockIdx.x];
        tile_x = w & 0xFFFFu;
        tile_y = w >> 16;
        t_lo = a.worklist[2 * blockIdx.x + 1];
        t_hi = t_lo + 1;
    }
    else
    {
        tile_x = blockIdx.x;
        tile_y = blockIdx.y;
        t_lo = a.t_begin + blockIdx.z * a.t_per_cta;
        t_hi = min(t_lo + a.t_per_cta, a.t_end);
    }
    const int tx0 = static_cast<int>(a.x0 + tile_x * GT_W);
    const int ty0 = static_cast<int>(a.y0 + tile_y * GT_H);
    const int tid = threadIdx.x;

    // stage the tile with its halo; zeros outside the map
    for (int i = tid; i < rows * pitch; i += GT_THREADS)
    {
        const int r = i / pitch, c = i - r * pitch;
        const int gy = ty0 - R + r, gx = tx0 - R + c;
        uint8_t v = 0;
        if (gy >= 0 && gy < static_cast<int>(a.H) && gx >= 0 && gx < static_cast<int>(a.W))
            v = a.cost[static_cast<size_t>(gy) * a.W + gx];
        tile[i] = v;
    }

    const int xl = tid % GT_W;
    const int yg = tid / GT_W;  // 0..3; rows yg, yg+4, ...
    const size_t plane = static_cast<size_t>(a.W) * a.H;

    for (uint32_t t = t_lo; t < t_hi; ++t)
    {
        const float2 cs = a.trig[t];
        unsigned acc[GT_ROWS_PER_THREAD];
#pragma unroll
        for (int k = 0; k < GT_ROWS_PER_THREAD; ++k)
            acc[k] = 0;

        for (uint32_t p0 = 0; p0 < a.P; p0 += GT_PCHUNK)
        {
            const int pc = static_cast<int>(min(static_cast<uint32_t>(GT_PCHUNK), a.P - p0));
            __syncthreads();  // tile staged / previous chunk consumed
            for (int i = tid; i < pc * GT_W; i += GT_THREADS)
            {
                const int p = i / GT_W, l = i - p * GT_W;
                const float px = static_cast<float>(a.fp[2 * (p0 + p)]), py = static_cast<float>(a.fp[2 * (p0 + p) + 1]);
                const int gx = tx0 + l;
                const float x = __fadd_rn(static_cast<float>(gx), a.ox);
                const float xn = __fsub_rn(__fadd_rn(x, __fmul_rn(px, cs.x)), __fmul_rn(py, cs.y));
                int off = __float2int_rd(xn) - gx;
                off = max(-R, min(R, off));  // memory safety only: the host refuses configurations that would need it (pp_se2_prepare_plan)
                xo[i] = static_cast<int8_t>(off);
            }
            for (int i = tid; i < pc * GT_H; i += GT_THREADS)
            {
                const int p = i / GT_H, l = i - p * GT_H;
                const float px = static_cast<float>(a.fp[2 * (p0 + p)]), py = static_cast<float>(a.fp[2 * (p0 + p) + 1]);
                const int gy = ty0 + l;
                const float y = __fadd_rn(static_cast<float>(gy), a.oy);
                const float yn = __fadd_rn(__fadd_rn(y, __fmul_rn(px, cs.y)), __fmul_rn(py, cs.x));
                int off = __float2int_rd(yn) - gy;
                off = max(-R, min(R, off));  // see above
                yo[i] = static_cast<int8_t>(off);
            }
            __syncthreads();
            for (int p = 0; p < pc; ++p)
            {
                const uint8_t *col = tile + (R + xl + xo[p * GT_W + xl]);
#pragma unroll
                for (int k = 0; k < GT_ROWS_PER_THREAD; ++k)
                {
                    const int row = yg + k * (GT_THREADS / GT_W);
                    const int sr = R + row + yo[p * GT_H + row];
                    acc[k] = max(acc[k], static_cast<unsigned>(col[sr * pitch]));
                }
            }
        }
        const int gx = tx0 + xl;
        if (gx < static_cast<int>(a.x1))
        {
#pragma unroll
            for (int k = 0; k < GT_ROWS_PER_THREAD; ++k)
            {
                const int gy = ty0 + yg + k * (GT_THREADS / GT_W);
                if (gy < static_cast<int>(a.y1))
                {
                    const size_t at = plane * t + static_cast<size_t>(gy) * a.W + gx;
                    for (int v = 0; v < a.n_vol; ++v)
                        a.vol[v][at] = static_cast<uint8_t>(acc[k]);
                }
            }
        }
    }
}

size_t generic_smem_bytes(int R)
{
    const int pitch = (GT_W + 2 * R + 3) & ~3;
    const int rows = GT_H + 2 * R;
    return static_cast<size_t>((pitch * rows + 15) & ~15) + static_cast<size_t>(GT_PCHUNK) * (GT_W + GT_H);
}
}  // namespace

// Launch the generic kernel over region [x0,x1) x [y0,y1) for the handle's bins.
static pp_status launch_generic(pp_handle h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, const uint32_t *d_worklist,
                                uint32_t n_work, cudaStream_t stream)
{
    GenericArgs a;
    a.cost = h->d_cost;
    a.n_vol = pp_volume_targets(h, a.vol);
    a.trig = h->d_trig;
    a.fp = h->d_fp;
    a.W = h->W;
    a.H = h->H;
    a.P = static_cast<uint32_t>(h->fp.size() / 2);
    a.R = h->reach + 2;
    a.ox = h->ox;
    a.oy = h->oy;
    a.t_begin = h->t_begin;
    a.t_end = h->t_end;
    a.x0 = x0;
    a.y0 = y0;
    a.x1 = x1;
    a.y1 = y1;
    a.worklist = d_worklist;
    const size_t smem = generic_smem_bytes(a.R);
    PP_REQUIRE(h, smem <= 227 * 1024, PP_ERR_RANGE, "footprint reach %d needs %zu bytes of shared memory", h->reach, smem);
    PP_CUDA(h, cudaFuncSetAttribute(se2_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    dim3 grid;
    if (d_worklist)
    {
        a.t_per_cta = 1;
        grid = dim3(n_work, 1, 1);
    }
    else
    {
        const uint32_t tiles_x = (x1 - x0 + GT_W - 1) / GT_W, tiles_y = (y1 - y0 + GT_H - 1) / GT_H;
        const uint32_t n_t = h->t_end - h->t_begin;
        // enough CTAs to fill 148 SMs several times over, but as many bins per CTA as that allows
        // so that the staged tile is reused
        const uint64_t tiles = static_cast<uint64_t>(tiles_x) * tiles_y;
        uint32_t chunks = static_cast<uint32_t>(std::min<uint64_t>(n_t, (148ull * 16 + tiles - 1) / tiles));
        chunks = std::max(1u, chunks);
        a.t_per_cta = (n_t + chunks - 1) / chunks;
        chunks = (n_t + a.t_per_cta - 1) / a.t_per_cta;
        PP_REQUIRE(h, tiles_y <= 65535 && chunks <= 65535, PP_ERR_RANGE, "grid too large");
        grid = dim3(tiles_x, tiles_y, chunks);
        h->stat_generic_tiles += tiles * n_t;
    }
    se2_generic_kernel<<<grid, GT_THREADS, smem, stream>>>(a);
    h->launches++;
    PP_CUDA(h, cudaGetLastError());
    return PP_OK;
}

pp_status pp_se2_launch(pp_handle h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1)
{
    h->stat_fast_tiles = 0;
    h->stat_generic_tiles = 0;
    h->stat_samples_per_cell = 0.0;
    if (x1 <= x0 || y1 <= y0 || h->bins.empty())
        return PP_OK;
    {
        pp_status s = pp_se2_prepare_plan(h);
        if (s != PP_OK)
            return s;
    }
    if (h->vol_mode == 0)
    {
        pp_status s = pp_se2_launch_fast(h, x0, y0, x1, y1);
        if (s != PP_INTERNAL_DECLINED)
            return s;
    }
    // the generic kernel takes a range of bins: one launch per run of consecutive bins
    size_t i = 0;
    const uint32_t keep_b = h->t_begin, keep_e = h->t_end;
    pp_status s = PP_OK;
    while (i < h->bins.size() && s == PP_OK)
    {
        size_t j = i + 1;
        while (j < h->bins.size() && h->bins[j] == h->bins[j - 1] + 1)
            ++j;
        h->t_begin = h->bins[i];
        h->t_end = h->bins[j - 1] + 1;
        s = launch_generic(h, x0, y0, x1, y1, nullptr, 0, h->stream);
        i = j;
    }
    h->t_begin = keep_b;
    h->t_end = keep_e;
    return s;
}

pp_status pp_se2_generic_worklist(pp_handle h, const uint32_t *d_worklist, uint32_t n_work)
{
    if (n_work == 0)
        return PP_OK;
    h->stat_generic_tiles += n_work;
    // fork: the side stream waits for what is already queued (costmap / work-list uploads, earlier builds)
    PP_CUDA(h, cudaEventRecord(h->ev_fork, h->stream));
    PP_CUDA(h, cudaStreamWaitEvent(h->aux, h->ev_fork, 0));
    return launch_generic(h, 0, 0, h->W, h->H, d_worklist, n_work, h->aux);
}

pp_status pp_se2_join_aux(pp_handle h)
{
    PP_CUDA(h, cudaEventRecord(h->ev_join, h->aux));
    PP_CUDA(h, cudaStreamWaitEvent(h->stream, h->ev_join, 0));
    return PP_OK;
}

extern "C" {

pp_status pp_se2_volume_set_mode(pp_handle h, int mode)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, mode == 0 || mode == 1, PP_ERR_INVALID, "pp_se2_volume_set_mode: mode must be 0 or 1");
    h->vol_mode = mode;
    return PP_OK;
}

pp_status pp_se2_volume_stats(pp_handle h, uint64_t *fast_tiles, uint64_t *generic_tiles, double *samples_per_cell)
{
    if (!h)
        return PP_ERR_INVALID;
    if (fast_tiles)
        *fast_tiles = h->stat_fast_tiles;
    if (generic_tiles)
        *generic_tiles = h->stat_generic_tiles;
    if (samples_per_cell)
        *samples_per_cell = h->stat_samples_per_cell;
    return PP_OK;
}

pp_status pp_se2_volume_build(pp_handle h, uint32_t n_theta, float ox, float oy, float ot, uint32_t t_begin,
                              uint32_t t_end)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, t_begin <= t_end && t_end <= n_theta, PP_ERR_INVALID, "bins [%u,%u) outside [0,%u)", t_begin, t_end, n_theta);
    std::vector<uint32_t> bins(t_end - t_begin);
    for (uint32_t t = t_begin; t < t_end; ++t)
        bins[t - t_begin] = t;
    return pp_se2_volume_build_bins(h, n_theta, ox, oy, ot, bins.data(), static_cast<uint32_t>(bins.size()));
}

pp_status pp_se2_volume_build_bins(pp_handle h, uint32_t n_theta, float ox, float oy, float ot, const uint32_t *bins_in,
                                   uint32_t n_bins)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, bins_in || n_bins == 0, PP_ERR_INVALID, "pp_se2_volume_build_bins: null bin list");
    std::vector<uint32_t> bins(bins_in, bins_in + n_bins);
    std::sort(bins.begin(), bins.end());
    bins.erase(std::unique(bins.begin(), bins.end()), bins.end());
    PP_REQUIRE(h, bins.empty() || bins.back() < n_theta, PP_ERR_INVALID, "pp_se2_volume_build_bins: bin %u outside [0,%u)",
               bins.empty() ? 0u : bins.back(), n_theta);
    const uint32_t t_begin = bins.empty() ? 0u : bins.front(), t_end = bins.empty() ? 0u : bins.back() + 1;
    PP_REQUIRE(h, h->d_cost, PP_ERR_STATE, "pp_se2_volume_build: upload a costmap first");
    PP_REQUIRE(h, !h->fp.empty(), PP_ERR_STATE, "pp_se2_volume_build: set a footprint first");
    PP_REQUIRE(h, n_theta > 0 && n_theta <= 4096, PP_ERR_RANGE, "n_theta %u outside 1..4096", n_theta);
    PP_REQUIRE(h, ox >= -1.0f && ox <= 1.0f && oy >= -1.0f && oy <= 1.0f, PP_ERR_INVALID,
               "sample offsets must lie in [-1,1] (got %g, %g)", ox, oy);
    PP_CUDA(h, cudaSetDevice(h->device));
    const size_t plane = static_cast<size_t>(h->W) * h->H;
    // peers may have this volume mapped (pp_se2_volume_ipc_open): it must not move under them
    PP_REQUIRE(h, h->n_peer == 0 || plane * n_theta <= h->vol_cap, PP_ERR_STATE,
               "pp_se2_volume_build: the volume is mapped by peers and too small for %u bins; reserve the final size before exporting it", n_theta);
    PP_CUDA(h, pp_reserve(h->d_vol, h->vol_cap, plane * n_theta));
    if (n_theta != h->n_theta || ox != h->ox || oy != h->oy || ot != h->ot)
        pp_se2_plan_free(h);
    h->n_theta = n_theta;
    h->ox = ox;
    h->oy = oy;
    h->ot = ot;
    h->t_begin = t_begin;
    h->t_end = t_end;
    h->bins.swap(bins);

    // heading -> (cos, sin) with the host libm, formed exactly like the reference does:
    // theta = motion.theta * bin_size (astar_hybrid.cpp:283), bin_size = 2.0f*(float)M_PI/size_theta
    // (motion_hybrid.cpp:10); float cos/sin as in collision_checker.cpp:155-156.
    // The table only depends on (n_theta, ot): it is recomputed and uploaded when those change, from a
    // vector the handle keeps alive, so a rebuild with unchanged headings costs no copy and no host wait.
    if (h->host_trig.size() != n_theta || h->trig_ot != ot || !h->d_trig)
    {
        // the previous upload (if any) reads the old vector: finish it before the vector is replaced
        PP_CUDA(h, cudaStreamSynchronize(h->stream));
        std::vector<float2> trig(n_theta);
        const float bin = 2.0f * static_cast<float>(M_PI) / static_cast<float>(n_theta);
        for (uint32_t t = 0; t < n_theta; ++t)
        {
            const float tf = static_cast<float>(t) + ot;
            const float theta = tf * bin;
            trig[t].x = cosf(theta);
            trig[t].y = sinf(theta);
        }
        h->host_trig.swap(trig);
        h->trig_ot = ot;
        PP_CUDA(h, pp_reserve(h->d_trig, h->trig_cap, static_cast<size_t>(n_theta)));
        PP_CUDA(h, cudaMemcpyAsync(h->d_trig, h->host_trig.data(), sizeof(float2) * n_theta, cudaMemcpyHostToDevice, h->stream));
    }

    const bool banded = h->band_y1 > h->band_y0;
    pp_status s = pp_se2_launch(h, 0, banded ? std::min(h->band_y0, h->H) : 0, h->W, banded ? std::min(h->band_y1, h->H) : h->H);
    if (s != PP_OK)
        return s;
    h->vol_valid = true;
    h->dirty = false;
    return PP_OK;
}

pp_status pp_se2_volume_set_rows(pp_handle h, uint32_t y_begin, uint32_t y_end)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, y_end == 0 || y_begin < y_end, PP_ERR_INVALID, "pp_se2_volume_set_rows: empty band [%u,%u)", y_begin, y_end);
    h->band_y0 = y_end ? y_begin : 0;
    h->band_y1 = y_end;
    return PP_OK;
}

pp_status pp_se2_volume_refresh(pp_handle h)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->vol_valid, PP_ERR_STATE, "pp_se2_volume_refresh: no volume built for the current costmap/footprint");
    if (!h->dirty)
        return PP_OK;
    PP_CUDA(h, cudaSetDevice(h->device));
    // an output cell reads inputs within reach+1 cells (rotation keeps |p|, floor and the sample
    // offset add at most one more), so those are the outputs a changed input can affect
    const uint32_t grow = static_cast<uint32_t>(h->reach + 2);
    const uint32_t x0 = h->dirty_x0 > grow ? h->dirty_x0 - grow : 0;
    const uint32_t y0 = h->dirty_y0 > grow ? h->dirty_y0 - grow : 0;
    const uint32_t x1 = std::min(h->W, h->dirty_x1 + grow);
    const uint32_t y1 = std::min(h->H, h->dirty_y1 + grow);
    pp_status s = pp_se2_launch(h, x0, y0, x1, y1);
    if (s != PP_OK)
        return s;
    h->dirty = false;
    return PP_OK;
}

pp_status pp_se2_volume_download(pp_handle h, uint32_t t_begin, uint32_t t_end, uint8_t *host_out, int host_is_pinned)
{
    if (!h)
        return PP_ERR_INVALID;
    (void)host_is_pinned;  // cudaMemcpyAsync handles both; pinned memory makes it a true async DMA
    PP_REQUIRE(h, h->d_vol && host_out, PP_ERR_STATE, "pp_se2_volume_download: no volume or null output");
    PP_REQUIRE(h, t_begin <= t_end && t_end <= h->n_theta, PP_ERR_INVALID, "bins [%u,%u) outside [0,%u)", t_begin, t_end, h->n_theta);
    PP_CUDA(h, cudaSetDevice(h->device));
    const size_t plane = static_cast<size_t>(h->W) * h->H;
    PP_CUDA(h, cudaMemcpyAsync(host_out, h->d_vol + plane * t_begin, plane * (t_end - t_begin), cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    return PP_OK;
}

pp_status pp_se2_volume_download_rows(pp_handle h, uint32_t t_begin, uint32_t t_end, uint32_t y_begin, uint32_t y_end,
                                      uint8_t *host_out)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_vol && host_out, PP_ERR_STATE, "pp_se2_volume_download_rows: no volume or null output");
    PP_REQUIRE(h, t_begin <= t_end && t_end <= h->n_theta, PP_ERR_INVALID, "bins [%u,%u) outside [0,%u)", t_begin, t_end, h->n_theta);
    PP_REQUIRE(h, y_begin <= y_end && y_end <= h->H, PP_ERR_INVALID, "rows [%u,%u) outside [0,%u)", y_begin, y_end, h->H);
    if (t_begin == t_end || y_begin == y_end)
        return PP_OK;
    PP_CUDA(h, cudaSetDevice(h->device));
    const size_t plane = static_cast<size_t>(h->W) * h->H, band = static_cast<size_t>(h->W) * (y_end - y_begin);
    // one strided copy: a "row" of it is the band of one bin
    PP_CUDA(h, cudaMemcpy2DAsync(host_out, band, h->d_vol + plane * t_begin + static_cast<size_t>(y_begin) * h->W, plane, band,
                                 t_end - t_begin, cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    return PP_OK;
}

pp_status pp_se2_volume_ptr(pp_handle h, const uint8_t **device_ptr_out, uint64_t *bytes_out)
{
    if (!h || !device_ptr_out)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_vol, PP_ERR_STATE, "pp_se2_volume_ptr: no volume");
    *device_ptr_out = h->d_vol;
    if (bytes_out)
        *bytes_out = static_cast<uint64_t>(h->W) * h->H * h->n_theta;
    return PP_OK;
}

}  // extern "C"
