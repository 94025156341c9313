/* pp_b200.h -- C ABI of libpp_b200.so: the costmap-parallel hot path of liangheming/path_planner,
 * built from scratch for NVIDIA B200 (sm_100a).
 *
 * The reference has no plugin / FFI seam for this path: it is reached through three source-level
 * C++ surfaces. Each entry point below names the reference interface it replaces,
 * as file:line relative to the reference repository root.
 *
 * Conventions
 *   - plain C types only; every function returns pp_status (0 = PP_OK) unless stated;
 *   - pp_last_error(h) gives the text of the last failure on that handle;
 *   - the caller owns all host buffers; the library owns all device buffers for the lifetime of
 *     the handle; one handle per (thread, device) -- a handle is not thread-safe;
 *   - all work is issued on the handle's own CUDA stream; calls that return data in host
 *     buffers synchronise that stream before returning, the others are asynchronous;
 *   - there is no CPU fallback: without a CUDA device pp_create fails.
 */
#ifndef PP_B200_H
#define PP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pp_ctx *pp_handle;
typedef int pp_status;

enum
{
    PP_OK = 0,
    PP_ERR_INVALID = 1, /* bad argument */
    PP_ERR_STATE = 2,   /* call order: e.g. volume build before costmap upload */
    PP_ERR_CUDA = 3,    /* a CUDA runtime call failed; text in pp_last_error */
    PP_ERR_RANGE = 4,   /* size beyond a documented limit */
    PP_ERR_COMM = 5     /* NCCL / peer-access failure */
};

/* cost values the path depends on (costmap_2d/cost_values.h, used at collision_checker.cpp:138,163,165) */
#define PP_COST_FREE 0
#define PP_COST_INSCRIBED 253
#define PP_COST_LETHAL 254
#define PP_COST_NO_INFORMATION 255

const char *pp_version(void);
pp_status pp_create(pp_handle *out, int device);
pp_status pp_destroy(pp_handle h);
const char *pp_last_error(pp_handle h);
/* Block until everything issued on the handle's stream has finished. */
pp_status pp_sync(pp_handle h);

/* Page-locked host memory for callers that want the asynchronous copies to be true DMA. */
pp_status pp_host_alloc(void **out, size_t bytes);
pp_status pp_host_free(void *p);

/* Measurement helpers (no counterpart in the reference):
 *   pp_timer_start/stop  CUDA-event stopwatch on the handle's stream (what bench.py times with)
 *   pp_l2_flush          overwrite a scratch buffer of `bytes` (> L2 size) on the handle's stream
 *                        so that the next kernel starts with a cold L2
 *   pp_launch_count      number of this library's kernels launched on the handle so far */
pp_status pp_timer_start(pp_handle h);
pp_status pp_timer_stop(pp_handle h, float *elapsed_ms); /* synchronises */
pp_status pp_l2_flush(pp_handle h, size_t bytes);
pp_status pp_launch_count(pp_handle h, uint64_t *count_out);

/* ---- K0: costmap upload ---------------------------------------------------------------------
 * Replaces the direct reads of costmap_2d::Costmap2D the reference does at
 * collision_checker.cpp:129-138,162,208 and teb_follower.cpp:143-155: the grid
 * (Costmap2D::getCharMap(), row-major my*W+mx) is copied to the device once per update instead.
 * The reference takes no lock while costmap_2d mutates the grid on its own thread; callers
 * should copy under costmap->getMutex().
 *   pp_costmap_upload       whole grid; `host_is_pinned` != 0 skips the staging copy (the copy is
 *                           then asynchronous: keep the buffer unchanged until pp_sync or a call
 *                           that returns data)
 *   pp_costmap_update_rect  rows [y0,y0+h) x cols [x0,x0+w) from `src` (src_pitch bytes per row);
 *                           the rectangle is merged into the handle's dirty region
 *   pp_costmap_download     device grid back to the host (tests)
 */
pp_status pp_costmap_upload(pp_handle h, const uint8_t *host, uint32_t W, uint32_t H, int host_is_pinned);
pp_status pp_costmap_update_rect(pp_handle h, uint32_t x0, uint32_t y0, uint32_t w, uint32_t hgt, const uint8_t *src,
                                 size_t src_pitch);
pp_status pp_costmap_download(pp_handle h, uint8_t *host_out);

/* ---- footprint ------------------------------------------------------------------------------
 * Replaces CollisionChecker::initBoundary (collision_checker.cpp:3-81, dispatch :72-81) and
 * setHole (collision_checker.h:14). `boundary` holds 1 number (circle radius), 2 (width, height)
 * or 4 (x_min, y_min, x_max, y_max) in metres; offsets are integer cells, produced x-outer /
 * y-inner exactly as the reference loops do. Unlike the reference, which appends to its point
 * list on every call, these calls REPLACE the footprint.
 * pp_footprint_set takes an arbitrary offset list (any rasterised polygon), P <= 4096.
 */
pp_status pp_footprint_from_boundary(pp_handle h, double resolution, const float *boundary, int n_boundary, int hole);
pp_status pp_footprint_set(pp_handle h, const int32_t *offsets_xy, uint32_t P);
pp_status pp_footprint_get(pp_handle h, int32_t *offsets_xy_out, uint32_t max_points, uint32_t *P_out);

/* ---- K1: SE2 footprint-collision volume -----------------------------------------------------
 * Replaces per-pose CollisionChecker::checkAndGetFootprintCost (collision_checker.cpp:144-172)
 * as called from AstarHybrid::getNeighbors (astar_hybrid.cpp:283) and isInputValid (:37,:40).
 * V[t][iy][ix] = (uint8) cost that function returns at x = fl(ix+ox), y = fl(iy+oy),
 * theta = fl(fl(t+ot) * bin), bin = 2.0f*(float)M_PI/(float)n_theta (motion_hybrid.cpp:10);
 * the collision flag it returns equals V >= 253. Layout [t][y][x], x fastest, one byte per cell.
 * Bit-exact with the reference's float code path (cosf/sinf from the host libm, no FMA).
 *   pp_se2_volume_build    compute heading bins [t_begin, t_end) of an n_theta-bin volume
 *   pp_se2_volume_refresh  recompute only the dirty region (grown by the footprint reach) for
 *                          the bins of the last build, then clear the dirty region
 *   pp_se2_volume_download copy bins [t_begin, t_end) to `host_out` ((t_end-t_begin)*H*W bytes)
 *   pp_se2_volume_download_rows  the same for map rows [y_begin, y_end) only: host_out receives
 *                          (t_end-t_begin) bands of (y_end-y_begin)*W bytes, one after the other
 *                          (what a row-band-sharded rank hands to its host; see pp_se2_volume_set_rows)
 *   pp_se2_volume_ptr      device address of bin 0 (for zero-copy consumers and NCCL)
 */
pp_status pp_se2_volume_build(pp_handle h, uint32_t n_theta, float ox, float oy, float ot, uint32_t t_begin,
                              uint32_t t_end);
/* The same for an arbitrary set of heading bins (any order, duplicates ignored): what a rank of a
 * heading-sharded job builds when it owns the bins t and t + n_theta/2 together, which a centrally
 * symmetric footprint lets it evaluate once. pp_se2_volume_refresh afterwards refreshes these bins. */
pp_status pp_se2_volume_build_bins(pp_handle h, uint32_t n_theta, float ox, float oy, float ot, const uint32_t *bins,
                                   uint32_t n_bins);
pp_status pp_se2_volume_refresh(pp_handle h);
pp_status pp_se2_volume_download(pp_handle h, uint32_t t_begin, uint32_t t_end, uint8_t *host_out,
                                 int host_is_pinned);
pp_status pp_se2_volume_download_rows(pp_handle h, uint32_t t_begin, uint32_t t_end, uint32_t y_begin, uint32_t y_end,
                                      uint8_t *host_out);
pp_status pp_se2_volume_ptr(pp_handle h, const uint8_t **device_ptr_out, uint64_t *bytes_out);
/* Which kernel variant pp_se2_volume_build uses: 0 = automatic, 1 = force the generic exact
 * gather everywhere (used by tests to cross-check the tiled fast path against it). */
pp_status pp_se2_volume_set_mode(pp_handle h, int mode);
/* Counters of the last build: tiles done by the fast path / by the exact generic path, and the
 * mean number of shared-memory samples per output of the fast path. */
pp_status pp_se2_volume_stats(pp_handle h, uint64_t *fast_tiles, uint64_t *generic_tiles, double *samples_per_cell);

/* ---- K1b: batched footprint cost at arbitrary poses ------------------------------------------
 * Replaces CollisionChecker::checkAndGetFootprintCost (collision_checker.cpp:144-172) and
 * checkFootPrintCollision (:126-142, called per analytic-expansion sample at
 * astar_hybrid.cpp:325) for poses that are not bin-aligned. poses_xyt = n x (x, y, theta), x and y
 * in cell units, theta in radians. cost_out[i] is the float the reference stores (0..255),
 * flag_out[i] its boolean (also the value checkFootPrintCollision returns).
 */
pp_status pp_footprint_cost_batch(pp_handle h, const float *poses_xyt, uint32_t n, float *cost_out,
                                  uint8_t *flag_out);

/* ---- K2: obstacle-aware 2D heuristic field ---------------------------------------------------
 * Replaces AstarHybrid::cacheObstacleHeuristic (astar_hybrid.cpp:97-165) and the lazy resume in
 * getObstacleHeuristic (:184-239): the whole goal-rooted field is built at once, so every later
 * lookup is the table hit at :182-183. F[y][x] = converged value of the reference recurrence
 * (h(goal) = 1e-5f, 8-connected, travel cost of :151, border cells and cost >= 253 never
 * entered); unreachable cells are 0, like the reference's unseen entries.
 * `field_out` may be NULL (keep the field on the device; see pp_heuristic_ptr).
 */
pp_status pp_heuristic_build(pp_handle h, uint32_t goal_x, uint32_t goal_y, float cost_penalty, float *field_out);
pp_status pp_heuristic_ptr(pp_handle h, const float **device_ptr_out);
pp_status pp_heuristic_stats(pp_handle h, uint32_t *rounds, uint64_t *tile_visits);

/* pp_heuristic_build_noalloc -- no counterpart in the reference. The same field as
 * pp_heuristic_build for callers that are not allowed to allocate device memory (hard real-time
 * planner threads, plugins inside a process that has given all device memory to its own pool,
 * processes that forbid allocation after start-up): it needs no handle and calls no cudaMalloc,
 * cudaFree or other allocating routine. The caller passes the costmap and the output field as
 * DEVICE pointers (W*H bytes in, W*H floats out) and the stream to run on (a cudaStream_t, or
 * NULL for the default stream); the kernel's scratch (per-tile keys and work lists, travel-cost table)
 * lives in a fixed ~260 KB array inside the library image, held by every GPU that loads the
 * library until the process exits.
 *   limit    ceil(W/32) * ceil(H/32) <= 16384 tiles (e.g. 4096 x 4096); larger maps are refused
 *            with PP_ERR_RANGE and a message on stderr
 *   device   the CUDA device that is current in the calling thread
 *   set-up   the first call on each device looks up the array's device address (and, like any first
 *            use of this library on a device, makes the CUDA runtime load the library's kernels, which
 *            the driver may serve with allocations of its own): make one warm-up call per device during
 *            start-up, before the real-time phase and before any stream capture on that device. First
 *            calls from several threads at once are safe (they store the same address)
 *   sharing  one array per device, cleared at the start of every call; calls whose GPU work
 *            overlaps race on it, so serialise them. Returns after the field is complete. */
pp_status pp_heuristic_build_noalloc(const uint8_t *d_cost, uint32_t W, uint32_t H, uint32_t goal_x, uint32_t goal_y,
                                     float cost_penalty, float *d_field_out, void *cuda_stream);

/* ---- N1: Hybrid-A* plan over the GPU tables --------------------------------------------------
 * Replaces AstarHybrid::createPath (astar_hybrid.cpp:392-515) as PathPlannerROS::goalSetCallback
 * calls it (path_planner_ros.cpp:136): builds the SE2 volume for the start pose's heading phase and
 * the heuristic field for the goal on the GPU, mirrors both in pinned host memory and runs the
 * reference's search on the host with O(1) table lookups (path_planner_b200/host/hybrid_astar.h).
 * The parameters are SearchInfo's (planner/commons.h:26-48; defaults config/params.yaml:32-45).
 * Poses are (x [m], y [m], yaw [rad]) in the map frame; the path comes back goal -> start, like the
 * reference's before path_planner_ros.cpp:140 reverses it. *found_out is 0 when the search fails
 * (that is a result, not an error). *n_points_out is the length of the whole path; at most max_points
 * poses are written, so a larger value means the output was truncated. Needs an uploaded costmap and a footprint.
 * Footprint checks use the pose's cell centre and nearest heading bin. */
typedef struct pp_plan_params
{
    int32_t motion_type; /* 1 Dubins, 2 Reeds-Shepp */
    uint32_t n_theta;
    float minimum_turning_radius, cost_penalty, change_penalty, non_straight_penalty, reverse_penalty;
    float analytic_expansion_ratio, analytic_expansion_tolerance, tolerance;
    uint32_t max_iterations, max_on_approach_iterations;
    uint32_t flags; /* PP_PLAN_* */
} pp_plan_params;
/* PP_PLAN_EXACT_ANALYTIC: the samples of every analytic expansion (astar_hybrid.cpp:315-329) are decided at
 * their CONTINUOUS poses like the reference's checkFootPrintCollision (collision_checker.cpp:126-142) instead
 * of at the cell centre / nearest bin: the volume filters, flagged samples are re-checked exactly, and an
 * expansion that survives is confirmed by one batched pp_footprint_cost_batch launch over all its samples.
 * PP_PLAN_VALIDATE_PATH: after a successful search every pose of the path is evaluated exactly at its
 * continuous position in one batched launch. If a pose collides there (the primitives of the search are
 * decided by the volume: see DESIGN.md for measured rates), its search node is excluded and the search is
 * repeated, up to three times; pp_plan_stats reports the repeats and what the last path still holds. */
#define PP_PLAN_EXACT_ANALYTIC 1u
#define PP_PLAN_VALIDATE_PATH 2u
pp_status pp_plan(pp_handle h, const pp_plan_params *params, double resolution, double origin_x, double origin_y,
                  const float *start_xyt, const float *goal_xyt, float *path_xyt_out, uint32_t max_points,
                  uint32_t *n_points_out, int32_t *iterations_out, int32_t *found_out, double *tables_ms_out,
                  double *search_ms_out);

/* Counters of the last pp_plan on this handle. */
typedef struct pp_plan_stats_t
{
    uint64_t lookups;                    /* volume lookups of the search */
    uint32_t analytic_attempts;          /* analytic expansions tried */
    uint32_t analytic_confirm_batches;   /* PP_PLAN_EXACT_ANALYTIC: batched exact confirmations launched */
    uint32_t exact_host_checks;          /* ... samples the volume flagged that were re-checked exactly */
    uint32_t false_collisions_overruled; /* ... of those, found free at the continuous pose */
    uint32_t false_free_caught;          /* ... samples the volume called free that the exact batch rejected */
    uint32_t path_poses_checked;         /* PP_PLAN_VALIDATE_PATH: poses of the path evaluated exactly ... */
    uint32_t path_poses_in_collision;    /* ... and how many of them collide at their continuous pose (last search) */
    uint32_t replans;                    /* ... searches repeated with the colliding poses' nodes excluded (at most 3) */
} pp_plan_stats_t;
pp_status pp_plan_stats(pp_handle h, pp_plan_stats_t *out);
/* Quantisation statistics: with the log enabled, pp_plan records every footprint question of the search --
 * the continuous pose (cells, cells, radians) the reference would have evaluated, whether it came from a
 * motion primitive / end pose (kind 0, astar_hybrid.cpp:37,40,283) or an analytic-expansion sample (kind 1,
 * :325), and what the volume answered. Feeding the poses to pp_footprint_cost_batch gives the exact answers. */
pp_status pp_plan_set_query_log(pp_handle h, int enable);
pp_status pp_plan_query_log(pp_handle h, float *poses_xyt_out, uint8_t *kind_out, uint8_t *flag_out, uint8_t *cost_out,
                            uint32_t max_queries, uint32_t *n_out);

/* ---- K3: TEB obstacle association -----------------------------------------------------------
 * pp_obstacles_extract replaces TebFollower::addObstacles (teb_follower.cpp:140-169): all cells
 * with cost >= 253, as world points (cell centres), filtered by the +-1.2*max_follow_range box
 * and the behind-the-robot rule, in the reference's x-outer / y-inner order. The points stay on
 * the device for pp_radius_topk; pts_out (max_points x 2 doubles) may be NULL.
 * pp_radius_topk replaces KDTree2D::buildIndex + radiusSearch + keep-first-K
 * (teb_optimizer.h:29, teb_optimizer.cpp:294-311; nanoflann.hpp:1747-1758): per query all points
 * with d2 < radius_arg (strict; radius_arg is compared against SQUARED distances, exactly as the
 * reference passes it), ascending by d2, first min(K, found) kept, K <= 64. pts_xy == NULL uses
 * the points of the last pp_obstacles_extract. Equal distances are ordered by point index (the
 * reference's order among ties is unspecified).
 */
pp_status pp_obstacles_extract(pp_handle h, double resolution, double origin_x, double origin_y, double robot_x,
                               double robot_y, double robot_theta, double max_follow_range,
                               double behind_ignore_range, double *pts_out, uint32_t max_points, uint32_t *n_out);
pp_status pp_radius_topk(pp_handle h, const double *pts_xy, uint32_t N, const double *queries_xy, uint32_t Q,
                         double radius_arg, uint32_t K, uint32_t *idx_out, double *d2_out, uint32_t *cnt_out);

/* ---- N3: TEB obstacle edges, batched (the step after K3) ---------------------
 * pp_robot_model_set replaces RobotModelHelper::createRobotModel
 * (src/path_follower/src/teb_path_follower/robot_model.cpp:68-95): 1 number = circle radius, 2 =
 * width and height of a rectangle, an even count >= 6 = polygon vertices x0,y0,x1,y1,...; the
 * inscribed / circumscribed radii of PolygonRobotModel::computeRadius (:15-37) are returned.
 * pp_obstacle_edges evaluates, for every (pose q, kept obstacle k) of the association that the last
 * pp_radius_topk left on the device (same Q and K), what g2o evaluates per EdgeObstacle and
 * Levenberg-Marquardt iteration (edges built at teb_optimizer.cpp:283-313): computeError()
 * (g2o_edges/edge_obstacle.hpp:19-26 -> RobotModel::distanceToPoint, robot_model.cpp:3-6,39-66,
 * and two penaltyBelow terms, commons.h:53-62) into error_out[Q][K][2], and the central-difference
 * Jacobian g2o computes for an edge without an analytic one (step 1e-9 per pose dimension through
 * Vertex2E::oplusImpl, commons.cpp:75-80) into jacobian_out[Q][K][2][3] (may be NULL). Slots
 * k >= cnt[q] are zero. Poses are (x, y, theta) doubles. Floating point: agrees with the
 * reference to 1e-12 (errors) and 1e-6 (Jacobians, whose step amplifies last-bit differences of
 * hypot/sincos by 5e8); tests/test_gpu_teb_edges.py states the tolerances. */
#define PP_ROBOT_MAX_VERTS 32
typedef struct pp_obstacle_edge_params
{
    double obstacle_collision_dist;   /* commons.h:219, 0.1 */
    double margin_obstacle_collision; /* commons.h:220, 0.05 */
    double obstacle_inflation_dist;   /* commons.h:221, 0.2 */
    double margin_obstacle_inflation; /* commons.h:222, 0.1 */
} pp_obstacle_edge_params;
pp_status pp_robot_model_set(pp_handle h, const double *numbers, uint32_t n, double *inscribed_out, double *circum_out);
pp_status pp_obstacle_edges(pp_handle h, const double *poses_xyt, uint32_t Q, uint32_t K,
                            const pp_obstacle_edge_params *params, double *error_out, double *jacobian_out);

/* ---- multi-GPU: heading-sharded volume --------------------------------------------------------
 * No counterpart in the reference (single process, single thread). Rank g of `world` builds bins
 * [g*n_theta/world, (g+1)*n_theta/world) and pp_se2_volume_allgather exchanges the slabs in place
 * with ncclAllGather so that every rank holds the whole volume.
 */
#define PP_COMM_ID_BYTES 128
pp_status pp_comm_unique_id(void *id_out);
pp_status pp_comm_init(pp_handle h, const void *id, int rank, int world);
pp_status pp_se2_volume_allgather(pp_handle h);
/* Heading-sharded build with the exchange hidden behind it (no counterpart in the reference). The
 * n_theta/2 pairs of bins (u, u + n_theta/2) are dealt out in contiguous runs, rank g owning pairs
 * [g*U/world, (g+1)*U/world) (single bins when n_theta is odd); a rank's run is cut into n_chunks
 * parts, and while part k+1 is being built on the handle's stream, part k of EVERY rank travels on a
 * second stream: by default as one group of ncclBroadcast calls (one per owner and half of the
 * volume; the parts need not be equal), with PP_SHARD_PUSH through the copy engines into the peers'
 * mapped volumes (pp_se2_volume_ipc_open) followed by a one-word all-reduce. When the call's work
 * is done on the handle's stream every rank holds the whole volume. PP_SHARD_EXCHANGE_ONLY skips
 * the builds (to time the exchange alone on an already built volume). Needs pp_comm_init. */
#define PP_SHARD_PUSH 1u
#define PP_SHARD_EXCHANGE_ONLY 2u
pp_status pp_se2_volume_build_sharded(pp_handle h, uint32_t n_theta, float ox, float oy, float ot, uint32_t n_chunks,
                                      uint32_t flags);

/* Fused build + exchange (no counterpart in the reference). Instead of building a slab and then
 * gathering, the build kernels store every output word into the volume of EVERY rank over
 * NVLink/NVSwitch peer mappings, so the slabs arrive while they are being computed and no separate
 * collective runs. Set-up, once per handle: every rank calls pp_se2_volume_reserve (allocates the
 * full-size volume; later builds of the same size do not move it) and pp_se2_volume_ipc_export,
 * the 64-byte handles are exchanged by any means, and every rank calls pp_se2_volume_ipc_open with
 * all of them (indexed by rank). pp_se2_volume_set_push(h, 1) then makes pp_se2_volume_build /
 * _refresh write to all mapped volumes. A rank's data is complete on every peer once that rank's
 * stream has been synchronised; a barrier across ranks after that makes the whole volume valid
 * everywhere. At most 8 ranks. */
#define PP_IPC_HANDLE_BYTES 64
pp_status pp_se2_volume_reserve(pp_handle h, uint32_t n_theta);
pp_status pp_se2_volume_ipc_export(pp_handle h, void *handle_out);
pp_status pp_se2_volume_ipc_open(pp_handle h, const void *handles, int rank, int world);
pp_status pp_se2_volume_set_push(pp_handle h, int enable);
/* Row-band sharding (no counterpart in the reference): restrict pp_se2_volume_build to the map rows
 * [y_begin, y_end) (all columns, the bins given to the build). Whole 16-row tiles that overlap the
 * band are computed, so neighbouring bands may both write a few identical rows. y_end == 0 restores
 * the whole map. Together with pp_se2_volume_set_push this shards the volume spatially across GPUs:
 * every rank pays the per-tile table build only for its own band, which scales linearly, whereas
 * heading-bin sharding repeats that cost on every rank. */
pp_status pp_se2_volume_set_rows(pp_handle h, uint32_t y_begin, uint32_t y_end);

/* ---- N2: costmap ingestion (the step before the hot path) ----------------
 * Today costmap_2d builds the grid on the CPU and the hot path only receives the result
 * (pp_costmap_upload). These entry points build it on the GPU instead, so that a map change goes
 * image/occupancy -> costmap -> SE2 volume refresh without leaving the device.
 *
 * pp_occupancy_from_image replaces the loop of MapManagerROS::loadMap
 * (src/map_manager/src/map_manager_ros.cpp:93-108): `image` is the W x H 8-bit grayscale raster
 * (row 0 at the top), thresholds are compared with the 0..255 pixel value as the reference does,
 * the grid is stored bottom row first (map_manager_ros.h:15) as nav_msgs/OccupancyGrid data
 * (0..100). pp_occupancy_upload / _update_rect take that data directly (the `/map` topic and its
 * updates, path_planner/config/params.yaml:18); _download returns it.
 *
 * pp_costmap_from_occupancy replaces the two costmap_2d layers the reference configures
 * (params.yaml:15-26; costmap_2d itself is not part of the reference tree): StaticLayer
 * (interpretValue, written over the master grid) and InflationLayer (computeCost over the cell
 * distance to the nearest lethal cell, out to ceil(inflation_radius / resolution) cells, with the
 * NO_INFORMATION rule of its updateCosts). Distances are exact Euclidean nearest-obstacle
 * distances; costmap_2d's queue hands sources over through four-neighbours and can settle on a
 * farther obstacle in rare configurations, where this result is the higher (safer) cost: cell by
 * cell the result is never lower than costmap_2d's, and equal except there (7 of 128 417 cells on the
 * reference's own map; parity with costmap_2d itself is unpinned, it is not in the reference tree) -- see
 * DESIGN.md. The result becomes the handle's costmap exactly as if it had been passed to
 * pp_costmap_upload; after pp_occupancy_update_rect only the changed rectangle grown by the
 * inflation radius is recomputed and recorded as the dirty region for pp_se2_volume_refresh.
 * Limits: inflation radius <= 32 cells. */
typedef struct pp_layer_params
{
    int track_unknown_space;    /* params.yaml:17 true */
    int lethal_threshold;       /* costmap_2d default 100 */
    int unknown_cost_value;     /* costmap_2d default -1 */
    int trinary_costmap;        /* costmap_2d default 1 */
    int inflate_unknown;        /* costmap_2d default 0 */
    double resolution;          /* metres per cell */
    double inscribed_radius;    /* metres; from the costmap footprint (params.yaml:6 gives 0.25) */
    double inflation_radius;    /* params.yaml:22, 0.5 */
    double cost_scaling_factor; /* params.yaml:21, 8.0 */
} pp_layer_params;
pp_status pp_occupancy_from_image(pp_handle h, const uint8_t *image, uint32_t W, uint32_t H, double occupied_thresh,
                                  double free_thresh, int negate);
pp_status pp_occupancy_upload(pp_handle h, const int8_t *occ, uint32_t W, uint32_t H);
pp_status pp_occupancy_update_rect(pp_handle h, uint32_t x0, uint32_t y0, uint32_t w, uint32_t hgt, const int8_t *src,
                                   size_t src_pitch);
pp_status pp_occupancy_download(pp_handle h, int8_t *occ_out);
pp_status pp_costmap_from_occupancy(pp_handle h, const pp_layer_params *params);

/* The follower's costmap is a rolling window (src/path_follower/config/params.yaml:12-24: rolling_window true,
 * width = height = 10 m at 0.05 m, StaticLayer + InflationLayer): costmap_2d moves the window's origin with the
 * robot (Costmap2D::updateOrigin), the static layer copies the part of the map under it into the master grid
 * (cells beyond the map keep the default value, NO_INFORMATION with track_unknown_space) and the inflation layer
 * sees the window's own cells only. pp_costmap_window_from_occupancy does that in one launch: the costmap of `h`
 * becomes the win_w x win_h window whose cell (0, 0) lies over cell (origin_x, origin_y) of the occupancy grid
 * held by `src` (the same handle or another one on the same device; the origin may be negative or beyond the
 * grid). pp_obstacles_extract, pp_radius_topk and the SE2 volume of `h` then work on the window. Same
 * nearest-obstacle inflation as above (parity with costmap_2d itself unpinned). pp_window_origin_cell gives the
 * origin along one axis as costmap_2d computes it from the robot's position, the window's size in metres, the
 * resolution and the map's origin. */
pp_status pp_costmap_window_from_occupancy(pp_handle h, pp_handle src, int32_t origin_x, int32_t origin_y, uint32_t win_w,
                                           uint32_t win_h, const pp_layer_params *params);
int32_t pp_window_origin_cell(double robot, double window_size_m, double resolution, double map_origin);

#ifdef __cplusplus
}
#endif
#endif /* PP_B200_H */
