// Handle lifetime, error reporting, K0 costmap upload (pinned staging, async, dirty region) and
// footprint set-up. See include/pp_b200.h for the contract of each entry point.
#include "pp_internal.h"

#include <cmath>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <algorithm>

pp_status pp_fail(pp_handle h, pp_status code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (h)
        h->err = buf;
    else
        fprintf(stderr, "pp_b200: %s\n", buf);
    return code;
}

pp_status pp_ensure_stage(pp_handle h, size_t bytes)
{
    if (bytes <= h->stage_cap)
        return PP_OK;
    // the staging buffer may still be the source of an in-flight copy
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    if (h->h_stage)
        PP_CUDA(h, cudaFreeHost(h->h_stage));
    h->h_stage = nullptr;
    h->stage_cap = 0;
    PP_CUDA(h, cudaHostAlloc(reinterpret_cast<void **>(&h->h_stage), bytes, cudaHostAllocDefault));
    h->stage_cap = bytes;
    return PP_OK;
}

extern "C" {

const char *pp_version(void) { return "pp_b200 0.1 (sm_100a)"; }

pp_status pp_create(pp_handle *out, int device)
{
    if (!out)
        return PP_ERR_INVALID;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return pp_fail(nullptr, PP_ERR_CUDA, "no CUDA device: %s (this library has no CPU path)",
                       e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= count)
        return pp_fail(nullptr, PP_ERR_INVALID, "device %d out of range (have %d)", device, count);
    pp_ctx *h = new pp_ctx();
    h->device = device;
    if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&h->aux, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreate(&h->ev_start)) != cudaSuccess || (e = cudaEventCreate(&h->ev_stop)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaHostAlloc(reinterpret_cast<void **>(&h->h_counter), 64, cudaHostAllocDefault)) != cudaSuccess)
    {
        pp_fail(nullptr, PP_ERR_CUDA, "pp_create: %s", cudaGetErrorString(e));
        delete h;
        return PP_ERR_CUDA;
    }
    if (cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || h->sm_count < 1)
        h->sm_count = 148;
    *out = h;
    return PP_OK;
}

pp_status pp_destroy(pp_handle h)
{
    if (!h)
        return PP_OK;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    if (h->aux)
        cudaStreamSynchronize(h->aux);
    pp_se2_plan_free(h);
    pp_comm_destroy(h);
    pp_plan_log_free(h);
    if (h->h_cost_mirror)
        cudaFreeHost(h->h_cost_mirror);
    cudaFree(h->d_cost);
    cudaFree(h->d_occ);
    cudaFree(h->d_infl);
    cudaFree(h->d_img);
    cudaFree(h->d_poses);
    cudaFree(h->d_edge_out);
    cudaFree(h->d_flush);
    cudaFree(h->d_fp);
    cudaFree(h->d_vol);
    cudaFree(h->d_trig);
    cudaFree(h->d_field);
    cudaFree(h->d_tile_flags);
    cudaFree(h->d_travel);
    cudaFree(h->d_pts);
    cudaFree(h->d_col_counts);
    cudaFree(h->d_queries);
    cudaFree(h->d_topk_idx);
    cudaFree(h->d_topk_d2);
    cudaFree(h->d_topk_cnt);
    cudaFreeHost(h->h_stage);
    cudaFreeHost(h->h_teb);
    cudaFreeHost(h->h_vol_mirror);
    cudaFreeHost(h->h_field_mirror);
    cudaFreeHost(h->h_counter);
    cudaEventDestroy(h->ev_fork);
    cudaEventDestroy(h->ev_join);
    cudaStreamDestroy(h->aux);
    cudaEventDestroy(h->ev_start);
    cudaEventDestroy(h->ev_stop);
    cudaStreamDestroy(h->stream);
    delete h;
    return PP_OK;
}

const char *pp_last_error(pp_handle h) { return h ? h->err.c_str() : "null handle"; }

pp_status pp_sync(pp_handle h)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_CUDA(h, cudaSetDevice(h->device));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    return PP_OK;
}

pp_status pp_host_alloc(void **out, size_t bytes)
{
    if (!out)
        return PP_ERR_INVALID;
    cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault);
    return e == cudaSuccess ? PP_OK : pp_fail(nullptr, PP_ERR_CUDA, "cudaHostAlloc(%zu): %s", bytes, cudaGetErrorString(e));
}

pp_status pp_host_free(void *p)
{
    cudaError_t e = cudaFreeHost(p);
    return e == cudaSuccess ? PP_OK : PP_ERR_CUDA;
}

pp_status pp_timer_start(pp_handle h)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_CUDA(h, cudaSetDevice(h->device));
    PP_CUDA(h, cudaEventRecord(h->ev_start, h->stream));
    return PP_OK;
}

pp_status pp_timer_stop(pp_handle h, float *elapsed_ms)
{
    if (!h || !elapsed_ms)
        return PP_ERR_INVALID;
    PP_CUDA(h, cudaSetDevice(h->device));
    PP_CUDA(h, cudaEventRecord(h->ev_stop, h->stream));
    PP_CUDA(h, cudaEventSynchronize(h->ev_stop));
    PP_CUDA(h, cudaEventElapsedTime(elapsed_ms, h->ev_start, h->ev_stop));
    return PP_OK;
}

pp_status pp_l2_flush(pp_handle h, size_t bytes)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_CUDA(h, cudaSetDevice(h->device));
    PP_CUDA(h, pp_reserve(h->d_flush, h->flush_cap, bytes));
    PP_CUDA(h, cudaMemsetAsync(h->d_flush, 0x5a, bytes, h->stream));
    return PP_OK;
}

pp_status pp_launch_count(pp_handle h, uint64_t *count_out)
{
    if (!h || !count_out)
        return PP_ERR_INVALID;
    *count_out = h->launches;
    return PP_OK;
}

// ---- K0 ---------------------------------------------------------------------------------------

pp_status pp_costmap_upload(pp_handle h, const uint8_t *host, uint32_t W, uint32_t H, int host_is_pinned)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, host && W > 0 && H > 0, PP_ERR_INVALID, "pp_costmap_upload: null grid or empty size");
    PP_REQUIRE(h, W <= 32768 && H <= 32768, PP_ERR_RANGE, "pp_costmap_upload: %ux%u exceeds 32768x32768", W, H);
    PP_CUDA(h, cudaSetDevice(h->device));
    const size_t bytes = static_cast<size_t>(W) * H;
    if (W != h->W || H != h->H)
    {
        h->vol_valid = false;
        pp_se2_plan_free(h);
    }
    PP_CUDA(h, pp_reserve(h->d_cost, h->cost_cap, bytes));
    h->W = W;
    h->H = H;
    const uint8_t *src = host;
    if (!host_is_pinned)
    {
        pp_status s = pp_ensure_stage(h, bytes);
        if (s != PP_OK)
            return s;
        // the previous async copy out of the staging buffer must be done before it is overwritten
        PP_CUDA(h, cudaStreamSynchronize(h->stream));
        std::memcpy(h->h_stage, host, bytes);
        src = h->h_stage;
    }
    PP_CUDA(h, cudaMemcpyAsync(h->d_cost, src, bytes, cudaMemcpyHostToDevice, h->stream));
    h->dirty = false;  // a full upload invalidates the volume as a whole; a rebuild is expected
    h->vol_valid = false;
    return PP_OK;
}

pp_status pp_costmap_update_rect(pp_handle h, uint32_t x0, uint32_t y0, uint32_t w, uint32_t hgt, const uint8_t *src,
                                 size_t src_pitch)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_cost, PP_ERR_STATE, "pp_costmap_update_rect: no costmap uploaded");
    PP_REQUIRE(h, src && w > 0 && hgt > 0 && src_pitch >= w, PP_ERR_INVALID, "pp_costmap_update_rect: bad source");
    PP_REQUIRE(h, static_cast<uint64_t>(x0) + w <= h->W && static_cast<uint64_t>(y0) + hgt <= h->H, PP_ERR_INVALID,
               "pp_costmap_update_rect: rect (%u,%u)+(%u,%u) outside %ux%u", x0, y0, w, hgt, h->W, h->H);
    PP_CUDA(h, cudaSetDevice(h->device));
    const size_t bytes = static_cast<size_t>(w) * hgt;
    pp_status s = pp_ensure_stage(h, std::max(bytes, static_cast<size_t>(h->W) * h->H));
    if (s != PP_OK)
        return s;
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    for (uint32_t r = 0; r < hgt; ++r)
        std::memcpy(h->h_stage + static_cast<size_t>(r) * w, src + static_cast<size_t>(r) * src_pitch, w);
    PP_CUDA(h, cudaMemcpy2DAsync(h->d_cost + static_cast<size_t>(y0) * h->W + x0, h->W, h->h_stage, w, w, hgt,
                                 cudaMemcpyHostToDevice, h->stream));
    if (!h->dirty)
    {
        h->dirty = true;
        h->dirty_x0 = x0;
        h->dirty_y0 = y0;
        h->dirty_x1 = x0 + w;
        h->dirty_y1 = y0 + hgt;
    }
    else
    {
        h->dirty_x0 = std::min(h->dirty_x0, x0);
        h->dirty_y0 = std::min(h->dirty_y0, y0);
        h->dirty_x1 = std::max(h->dirty_x1, x0 + w);
        h->dirty_y1 = std::max(h->dirty_y1, y0 + hgt);
    }
    return PP_OK;
}

pp_status pp_costmap_download(pp_handle h, uint8_t *host_out)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_cost && host_out, PP_ERR_STATE, "pp_costmap_download: no costmap or null output");
    PP_CUDA(h, cudaSetDevice(h->device));
    PP_CUDA(h, cudaMemcpyAsync(host_out, h->d_cost, static_cast<size_t>(h->W) * h->H, cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    return PP_OK;
}

// ---- footprint ----------------------------------------------------------------------------------

static pp_status commit_footprint(pp_handle h)
{
    const uint32_t P = static_cast<uint32_t>(h->fp.size() / 2);
    PP_REQUIRE(h, P > 0, PP_ERR_INVALID, "footprint has no points");
    PP_REQUIRE(h, P <= 4096, PP_ERR_RANGE, "footprint has %u points; the limit is 4096", P);
    double r2 = 0.0;
    for (uint32_t p = 0; p < P; ++p)
    {
        double x = h->fp[2 * p], y = h->fp[2 * p + 1];
        r2 = std::max(r2, x * x + y * y);
    }
    h->reach = static_cast<int>(std::ceil(std::sqrt(r2)));
    if (const char *bias = getenv("PP_DEBUG_REACH_BIAS"))  // test hook: a deliberately short reach must be refused by the build
        h->reach = std::max(0, h->reach + atoi(bias));
    PP_REQUIRE(h, h->reach <= 120, PP_ERR_RANGE, "footprint reach %d cells exceeds the limit of 120", h->reach);
    PP_CUDA(h, cudaSetDevice(h->device));
    PP_CUDA(h, pp_reserve(h->d_fp, h->fp_cap, static_cast<size_t>(2) * P));
    PP_CUDA(h, cudaMemcpyAsync(h->d_fp, h->fp.data(), sizeof(int32_t) * 2 * P, cudaMemcpyHostToDevice, h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));  // h->fp may be rewritten by the next call
    h->vol_valid = false;
    pp_se2_plan_free(h);
    return PP_OK;
}

pp_status pp_footprint_from_boundary(pp_handle h, double resolution, const float *boundary, int nb, int hole)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, boundary && resolution > 0.0, PP_ERR_INVALID, "pp_footprint_from_boundary: null boundary or resolution <= 0");
    PP_REQUIRE(h, nb == 1 || nb == 2 || nb == 4, PP_ERR_INVALID,
               "pp_footprint_from_boundary: boundary must have 1, 2 or 4 numbers (got %d)", nb);
    std::vector<int32_t> pts;
    auto rim_or_all = [&](int i0, int i1, int j0, int j1) {
        for (int i = i0; i <= i1; ++i)
            for (int j = j0; j <= j1; ++j)
                if (!hole || i == i0 || i == i1 || j == j0 || j == j1)
                {
                    pts.push_back(i);
                    pts.push_back(j);
                }
    };
    if (nb == 1)
    {
        PP_REQUIRE(h, boundary[0] > 0, PP_ERR_INVALID, "circle radius must be > 0");
        const int r = static_cast<int>(std::ceil(boundary[0] / resolution));
        const int inner = std::max(0, r - 1);
        for (int i = -r; i <= r; ++i)
            for (int j = -r; j <= r; ++j)
            {
                const int d = i * i + j * j;
                if (d <= r * r && (!hole || d >= inner * inner))
                {
                    pts.push_back(i);
                    pts.push_back(j);
                }
            }
    }
    else if (nb == 2)
    {
        PP_REQUIRE(h, boundary[0] > 0 && boundary[1] > 0, PP_ERR_INVALID, "rectangle width/height must be > 0");
        const int hw = static_cast<int>(std::ceil(0.5 * boundary[0] / resolution));
        const int hh = static_cast<int>(std::ceil(0.5 * boundary[1] / resolution));
        rim_or_all(-hw, hw, -hh, hh);
    }
    else
    {
        PP_REQUIRE(h, boundary[0] <= 0 && boundary[1] <= 0 && boundary[2] >= 0 && boundary[3] >= 0, PP_ERR_INVALID,
                   "box needs x_min,y_min <= 0 <= x_max,y_max");
        rim_or_all(static_cast<int>(std::floor(boundary[0] / resolution)), static_cast<int>(std::ceil(boundary[2] / resolution)),
                   static_cast<int>(std::floor(boundary[1] / resolution)), static_cast<int>(std::ceil(boundary[3] / resolution)));
    }
    h->fp.swap(pts);
    return commit_footprint(h);
}

pp_status pp_footprint_set(pp_handle h, const int32_t *offsets_xy, uint32_t P)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, offsets_xy && P > 0, PP_ERR_INVALID, "pp_footprint_set: empty footprint");
    PP_REQUIRE(h, P <= 4096, PP_ERR_RANGE, "footprint has %u points; the limit is 4096", P);
    h->fp.assign(offsets_xy, offsets_xy + 2 * static_cast<size_t>(P));
    return commit_footprint(h);
}

pp_status pp_footprint_get(pp_handle h, int32_t *out, uint32_t max_points, uint32_t *P_out)
{
    if (!h || !P_out)
        return PP_ERR_INVALID;
    const uint32_t P = static_cast<uint32_t>(h->fp.size() / 2);
    *P_out = P;
    if (out)
        std::memcpy(out, h->fp.data(), sizeof(int32_t) * 2 * std::min(P, max_points));
    return PP_OK;
}

}  // extern "C"
