// K1b: footprint cost at arbitrary continuous poses (see include/pp_b200.h).
// One warp per pose; lanes stride over the footprint points, evaluate the reference expression
// (collision_checker.cpp:155-158) with separately rounded float operations and host-libm cos/sin
// (computed per pose in pp_footprint_cost_batch), gather from the device costmap and combine with
// a warp max (__reduce_max_sync). The collision flag is cost >= 253 (:165), which is also what
// checkFootPrintCollision (:126-142) returns for the same pose.
#include "pp_internal.h"

#include <cmath>
#include <vector>

namespace
{
struct PoseTrig
{
    float x, y, c, s;
};

__global__ void __launch_bounds__(256) footprint_batch_kernel(const uint8_t *__restrict__ cost, uint32_t W, uint32_t H,
                                                              const int32_t *__restrict__ fp, uint32_t P,
                                                              const PoseTrig *__restrict__ poses, uint32_t n,
                                                              float *__restrict__ cost_out, uint8_t *__restrict__ flag_out)
{
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31;
    if (warp >= n)
        return;
    const PoseTrig q = poses[warp];
    unsigned m = 0;
    for (uint32_t p = lane; p < P; p += 32)
    {
        const float px = static_cast<float>(fp[2 * p]), py = static_cast<float>(fp[2 * p + 1]);
        const float xn = __fsub_rn(__fadd_rn(q.x, __fmul_rn(px, q.c)), __fmul_rn(py, q.s));
        const float yn = __fadd_rn(__fadd_rn(q.y, __fmul_rn(px, q.s)), __fmul_rn(py, q.c));
        // The reference converts floor(.) to unsigned; x86-64 gcc does that through a 64-bit
        // signed truncation, so negative values wrap to huge indices and fail "< size". A signed
        // conversion plus a sign test gives the same accept/reject decision.
        const float fx = floorf(xn), fy = floorf(yn);
        if (fx >= 0.0f && fy >= 0.0f && fx < static_cast<float>(W) && fy < static_cast<float>(H))
        {
            const uint32_t xi = static_cast<uint32_t>(fx), yi = static_cast<uint32_t>(fy);
            m = max(m, static_cast<unsigned>(cost[static_cast<size_t>(yi) * W + xi]));
        }
    }
    m = __reduce_max_sync(0xffffffffu, m);
    if (lane == 0)
    {
        cost_out[warp] = static_cast<float>(m);
        flag_out[warp] = m >= PP_COST_INSCRIBED ? 1 : 0;
    }
}
}  // namespace

extern "C" pp_status pp_footprint_cost_batch(pp_handle h, const float *poses_xyt, uint32_t n, float *cost_out,
                                             uint8_t *flag_out)
{
    if (!h)
        return PP_ERR_INVALID;
    PP_REQUIRE(h, h->d_cost, PP_ERR_STATE, "pp_footprint_cost_batch: upload a costmap first");
    PP_REQUIRE(h, !h->fp.empty(), PP_ERR_STATE, "pp_footprint_cost_batch: set a footprint first");
    if (n == 0)
        return PP_OK;
    PP_REQUIRE(h, poses_xyt && cost_out && flag_out, PP_ERR_INVALID, "pp_footprint_cost_batch: null buffer");
    PP_CUDA(h, cudaSetDevice(h->device));
    std::vector<PoseTrig> host(n);
    for (uint32_t i = 0; i < n; ++i)
    {
        const float th = poses_xyt[3 * i + 2];
        host[i].x = poses_xyt[3 * i];
        host[i].y = poses_xyt[3 * i + 1];
        host[i].c = cosf(th);  // host libm, float overloads: what the reference build resolves to
        host[i].s = sinf(th);
    }
    PoseTrig *d_poses = nullptr;
    float *d_cost_out = nullptr;
    uint8_t *d_flag = nullptr;
    PP_CUDA(h, cudaMallocAsync(reinterpret_cast<void **>(&d_poses), sizeof(PoseTrig) * n, h->stream));
    PP_CUDA(h, cudaMallocAsync(reinterpret_cast<void **>(&d_cost_out), sizeof(float) * n, h->stream));
    PP_CUDA(h, cudaMallocAsync(reinterpret_cast<void **>(&d_flag), n, h->stream));
    PP_CUDA(h, cudaMemcpyAsync(d_poses, host.data(), sizeof(PoseTrig) * n, cudaMemcpyHostToDevice, h->stream));
    const uint32_t warps_per_block = 8;
    footprint_batch_kernel<<<(n + warps_per_block - 1) / warps_per_block, warps_per_block * 32, 0, h->stream>>>(
        h->d_cost, h->W, h->H, h->d_fp, static_cast<uint32_t>(h->fp.size() / 2), d_poses, n, d_cost_out, d_flag);
    h->launches++;
    PP_CUDA(h, cudaGetLastError());
    PP_CUDA(h, cudaMemcpyAsync(cost_out, d_cost_out, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaMemcpyAsync(flag_out, d_flag, n, cudaMemcpyDeviceToHost, h->stream));
    PP_CUDA(h, cudaFreeAsync(d_poses, h->stream));
    PP_CUDA(h, cudaFreeAsync(d_cost_out, h->stream));
    PP_CUDA(h, cudaFreeAsync(d_flag, h->stream));
    PP_CUDA(h, cudaStreamSynchronize(h->stream));
    return PP_OK;
}
