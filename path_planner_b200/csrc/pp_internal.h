// Internal state of a pp_handle. Not part of the ABI.
#pragma once
#include "../../include/pp_b200.h"

#include <cuda_runtime.h>
#include <cstdio>
#include <string>
#include <vector>

struct Se2Plan;  // se2_plan.h

struct pp_ctx
{
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t aux = nullptr;  // side stream: the exact work-list kernel runs beside the tiled kernel
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
    std::string err;
    uint64_t launches = 0;  // kernels launched by this library on this handle
    uint8_t *d_flush = nullptr;
    size_t flush_cap = 0;

    // K0 costmap
    uint8_t *d_cost = nullptr;
    size_t cost_cap = 0;
    uint32_t W = 0, H = 0;
    uint8_t *h_stage = nullptr;  // pinned staging
    size_t stage_cap = 0;
    bool dirty = false;
    uint32_t dirty_x0 = 0, dirty_y0 = 0, dirty_x1 = 0, dirty_y1 = 0;  // half-open

    // footprint
    std::vector<int32_t> fp;  // x0,y0,x1,y1,...
    int32_t *d_fp = nullptr;
    size_t fp_cap = 0;
    int reach = 0;  // ceil(max |p|) in cells

    // K1 volume
    uint8_t *d_vol = nullptr;
    size_t vol_cap = 0;
    uint32_t n_theta = 0, t_begin = 0, t_end = 0;  // t_begin/t_end: the hull of `bins`
    std::vector<uint32_t> bins;                    // heading bins of the current build, ascending
    float ox = 0.5f, oy = 0.5f, ot = 0.0f;
    bool vol_valid = false;
    float2 *d_trig = nullptr;  // (cos, sin) per heading bin, from the host libm
    size_t trig_cap = 0;
    std::vector<float2> host_trig;  // what d_trig holds (kept alive: the upload reads it asynchronously)
    float trig_ot = 0.0f;
    int vol_mode = 0;
    uint64_t stat_fast_tiles = 0, stat_generic_tiles = 0;
    double stat_samples_per_cell = 0.0;
    Se2Plan *plan = nullptr;

    // K2 heuristic
    float *d_field = nullptr;
    size_t field_cap = 0;
    uint32_t *d_tile_flags = nullptr;  // two generations of per-tile activity flags + counters
    size_t tile_flags_cap = 0;
    float *d_travel = nullptr;  // 2 x 256 travel-cost table (straight, diagonal)
    uint32_t *h_counter = nullptr;  // pinned
    uint32_t heur_rounds = 0;
    uint64_t heur_tile_visits = 0, heur_sweeps = 0;

    // N1 plan: pinned host mirrors of the tables
    uint8_t *h_vol_mirror = nullptr;
    size_t vol_mirror_cap = 0;
    float *h_field_mirror = nullptr;
    size_t field_mirror_cap = 0;
    uint8_t *h_cost_mirror = nullptr;  // exact analytic expansion: the costmap on the host
    size_t cost_mirror_cap = 0;
    pp_plan_stats_t plan_stats = {};
    bool plan_log_on = false;
    void *plan_log = nullptr;  // std::vector<ppb::Query>, owned by plan_api.cu (pp_plan_log_free)

    // K3 TEB
    double *d_pts = nullptr;
    size_t pts_cap = 0;
    uint32_t n_pts = 0;
    uint32_t *d_col_counts = nullptr;  // exclusive prefix of the points per costmap column (pp_obstacles_extract)
    uint32_t pts_cols = 0;             // > 0: d_pts is column-ordered over this many columns ...
    double pts_col_origin = 0.0, pts_col_res = 0.0;  // ... with centres at origin + (c + 0.5) * res
    size_t col_cap = 0;
    double *d_queries = nullptr;
    size_t q_cap = 0;
    uint32_t *d_topk_idx = nullptr;
    double *d_topk_d2 = nullptr;
    uint32_t *d_topk_cnt = nullptr;
    size_t topk_cap = 0, topk_q_cap = 0;
    uint8_t *h_teb = nullptr;  // pinned landing buffer of the TEB results (every TEB call synchronises before it returns)
    size_t teb_host_cap = 0;

    // N2 ingestion
    int8_t *d_occ = nullptr;
    size_t occ_cap = 0;
    uint32_t occ_W = 0, occ_H = 0;
    bool occ_dirty = false, occ_all = false;  // a rectangle / the whole grid changed since the last costmap
    bool cost_is_window = false;              // the costmap is a rolling window, not the occupancy grid's own
    uint32_t occ_x0 = 0, occ_y0 = 0, occ_x1 = 0, occ_y1 = 0;  // half-open
    uint8_t *d_img = nullptr;  // map image as uploaded
    size_t img_cap = 0;
    uint8_t *d_infl = nullptr;  // inflation cost by squared cell distance, 0..R*R
    size_t infl_cap = 0;
    pp_layer_params layer = {};
    bool layer_valid = false;

    // N3 obstacle edges
    int robot_nv = -1;  // -1 unset, 0 circle, else polygon vertex count
    double robot_verts[2 * PP_ROBOT_MAX_VERTS] = {};
    double robot_inscribed = 0.0, robot_circum = 0.0;
    uint32_t topk_Q = 0, topk_K = 0;  // shape of the association left on the device by pp_radius_topk
    double *d_poses = nullptr;
    size_t poses_cap = 0;
    double *d_edge_out = nullptr;
    size_t edge_cap = 0;

    // multi-GPU
    uint8_t *peer_vol[8] = {};  // peer_vol[r] = rank r's volume mapped here (own volume at [rank]); null if not opened
    int n_peer = 0;             // world size of the mapping
    bool push = false;          // builds store to every mapped volume
    uint32_t band_y0 = 0, band_y1 = 0;  // row band of pp_se2_volume_build (0,0 = whole map)
    void *comm = nullptr;  // ncclComm_t
    // pp_se2_volume_build_sharded: the exchange runs on streams of its own beside the build
    cudaStream_t comm_stream = nullptr, push_stream[8] = {};
    cudaEvent_t ev_comm = nullptr, ev_chunk[16] = {}, ev_push[8] = {};
    int *d_token = nullptr;
    int few_bins_override = -1;  // -1: by the number of bins of a build; 0 / 1: forced (one plan for all chunks of a sharded build)
    int rank = 0, world = 1;
};

pp_status pp_fail(pp_handle h, pp_status code, const char *fmt, ...);

#define PP_CUDA(h, call)                                                                             \
    do                                                                                               \
    {                                                                                                \
        cudaError_t _e = (call);                                                                     \
        if (_e != cudaSuccess)                                                                       \
            return pp_fail((h), PP_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), \
                           __FILE__, __LINE__);                                                      \
    } while (0)

#define PP_REQUIRE(h, cond, code, ...)            \
    do                                            \
    {                                             \
        if (!(cond))                              \
            return pp_fail((h), (code), __VA_ARGS__); \
    } while (0)

// grow-only device buffer helper
template <class T>
inline cudaError_t pp_reserve(T *&ptr, size_t &cap, size_t need)
{
    if (need <= cap && ptr)
        return cudaSuccess;
    if (ptr)
    {
        cudaError_t e = cudaFree(ptr);
        if (e != cudaSuccess)
            return e;
        ptr = nullptr;
        cap = 0;
    }
    cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&ptr), need * sizeof(T));
    if (e == cudaSuccess)
        cap = need;
    return e;
}

// pinned staging buffer of at least `bytes` (pp_api.cu)
pp_status pp_ensure_stage(pp_handle h, size_t bytes);
// implemented per kernel file
pp_status pp_se2_launch(pp_handle h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1);
void pp_se2_plan_free(pp_handle h);
void pp_comm_destroy(pp_handle h);
void pp_plan_log_free(pp_handle h);
// fills up to 8 destination volumes for a build (own first); returns how many
int pp_volume_targets(pp_handle h, uint8_t **out);
