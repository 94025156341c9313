/* TEST INFRASTRUCTURE ONLY (oracle). CPU restatement of the reference's algorithms for the
 * costmap-parallel hot path. Only tests/, __graft_entry__.smoke() and bench.py's CPU baseline may
 * use it; the product library never links or loads it.
 *
 * Pinning status: the reference has no tests, golden vectors or fixtures, so this
 * restatement is pinned against the reference's OWN sources compiled unmodified into
 * oracle/_ref/libpp_ref.so (oracle/Makefile) -- tests/test_oracle.py compares the two on seeded
 * inputs and against the fixtures in tests/golden/ that gen_golden.py produced from _ref.
 */
#ifndef PP_ORACLE_PORT_H
#define PP_ORACLE_PORT_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

int port_footprint_points(double resolution, const float *boundary, int nb, int hole, int32_t *out_xy, int max_pts);

int port_se2_volume(const uint8_t *cost, uint32_t W, uint32_t H, const int32_t *pts_xy, int P, uint32_t n_theta,
                    float ox, float oy, float ot, uint32_t t0, uint32_t t1, uint32_t y0, uint32_t y1, uint8_t *out,
                    int nthreads, int trig_double);

int port_footprint_cost_batch(const uint8_t *cost, uint32_t W, uint32_t H, const int32_t *pts_xy, int P,
                              const float *poses_xyt, uint32_t n, float *cost_out, uint8_t *flag_out,
                              uint8_t *collision_out, int trig_double);

int port_heuristic(const uint8_t *cost, uint32_t W, uint32_t H, float start_x, float start_y, float goal_x,
                   float goal_y, float cost_penalty, const float *queries_xy, uint32_t n_queries, float *query_out,
                   float *table_out);

int port_heuristic_fixed_point(const uint8_t *cost, uint32_t W, uint32_t H, uint32_t goal_x, uint32_t goal_y,
                               float cost_penalty, float *field_out);

int port_extract_obstacles(const uint8_t *cost, uint32_t W, uint32_t H, double res, double origin_x, double origin_y,
                           double robot_x, double robot_y, double robot_theta, double max_follow_range,
                           double behind_ignore_range, double *pts_out, uint32_t max_pts);

int port_radius_topk(const double *pts_xy, uint32_t N, const double *q_xy, uint32_t Q, double radius_arg, uint32_t K,
                     uint32_t *idx_out, double *d2_out, uint32_t *cnt_out, uint32_t *found_out);

/* ---- ingestion before the hot path (oracle/ingest.c) ------------------ */
typedef struct
{
    int track_unknown_space;    /* params.yaml:17 true */
    int lethal_threshold;       /* costmap_2d default 100 */
    int unknown_cost_value;     /* costmap_2d default -1 */
    int trinary_costmap;        /* costmap_2d default true */
    int inflate_unknown;        /* costmap_2d default false */
    double resolution;          /* metres per cell */
    double inscribed_radius;    /* metres, from the costmap footprint (params.yaml:6 -> 0.25) */
    double inflation_radius;    /* params.yaml:22 0.5 */
    double cost_scaling_factor; /* params.yaml:21 8.0 */
} port_layer_params;

int port_occupancy_from_image(const uint8_t *image, uint32_t W, uint32_t H, double occ_th, double free_th, int negate,
                              int8_t *occ_out);
int port_static_layer(const int8_t *occ, uint32_t W, uint32_t H, const port_layer_params *p, uint8_t *cost_out);
uint32_t port_cell_inflation_radius(const port_layer_params *p);
int port_inflation_cost_table(const port_layer_params *p, uint8_t *table);
int port_inflate_propagate(uint8_t *master, uint32_t W, uint32_t H, const port_layer_params *p);
int port_inflate_nearest(uint8_t *master, uint32_t W, uint32_t H, const port_layer_params *p);

/* ---- obstacle edges after the TEB association (oracle/teb_edges.c) ---- */
int port_robot_model(const double *numbers, int n, double *verts_xy, int max_verts, double *inscribed, double *circum);
int port_obstacle_edges(const double *poses_xyt, uint32_t Q, const double *pts_xy, const uint32_t *idx,
                        const uint32_t *cnt, uint32_t K, const double *robot_numbers, int n_robot, const double *info4,
                        double *err_out, double *jac_out);

#ifdef __cplusplus
}
#endif
#endif
