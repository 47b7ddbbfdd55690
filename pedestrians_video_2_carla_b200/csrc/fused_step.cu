// p2c_fused_step: ProjectionModule.forward -> Normalizer -> loc_2d / loc_3d / loc_2d_3d, forward AND
// the gradient w.r.t. the pose input, for one batch (modules/flow/pose_lifting.py:121-195,
// modules/flow/base.py:440-469).
//
// Generic-tree path (this file): FK forward kernel -> loss kernel (normalise + masked MSE partials) ->
// finalize (fixed-order double reduction) -> loss-gradient kernel -> FK backward kernel.
// The single-pass kernel for rotation_6d input lives in fused_fast.cu and is dispatched from here.
#include "common.cuh"

extern "C" int p2c_fused_fast_supported(const p2c_fused_step_t* p);
extern "C" long p2c_fused_fast_workspace_floats(void);
extern "C" int p2c_fused_fast(const p2c_fused_step_t* p, void* stream);
// the pose-per-lane tile kernel (fused_tile.cu): preferred whenever it supports the call
extern "C" int p2c_fused_tile_supported(const p2c_fused_step_t* p);
extern "C" int p2c_fused_tile_launches(const p2c_fused_step_t* p);
extern "C" int p2c_fused_tile(const p2c_fused_step_t* p, void* stream);

namespace {

constexpr int WARPS = 4;
constexpr int LOSS_BLOCKS = 592;  // 148 SMs x 4: fixed grid => fixed summation order

struct LossArgs {
    long M;
    int J, hips, neck, normalize, mask, gt_c;
    const float* proj;
    const float* abs_loc;
    const float* gt2;
    const float* gt3;
    float* proj_norm;
    float* partials;  // LOSS_BLOCKS x 4
    const float* losses;
    float w2, w3;
    float* g_proj;
    float* g_abs_loc;
};

struct Norm {
    float sx, sy, dx, dy, scale;
    bool ok;
};

__device__ __forceinline__ Norm norm_of(const float* p, int hips, int neck) {
    Norm n;
    n.sx = __ldg(p + hips * 3);
    n.sy = __ldg(p + hips * 3 + 1);
    n.dx = __ldg(p + neck * 3) - n.sx;
    n.dy = __ldg(p + neck * 3 + 1) - n.sy;
    n.scale = sqrtf(n.dx * n.dx + n.dy * n.dy);
    n.ok = n.scale > 0.f && n.scale <= 3.402823466e38f;
    return n;
}

// normalised prediction of joint j + whether gradient flows to it (see loss_kernels.cu normalize_*)
__device__ __forceinline__ void pred_2d(const float* p, int j, const Norm& n, bool normalize, float& x, float& y,
                                        float& q, bool& live_x, bool& live_y) {
    q = p[j * 3 + 2];
    if (!normalize) {
        x = p[j * 3]; y = p[j * 3 + 1];
        live_x = live_y = true;
        return;
    }
    const float rx = (p[j * 3] - n.sx) / n.scale, ry = (p[j * 3 + 1] - n.sy) / n.scale;
    live_x = fabsf(rx) <= 3.402823466e38f;
    live_y = fabsf(ry) <= 3.402823466e38f;
    x = live_x ? rx : 0.f;
    y = live_y ? ry : 0.f;
    q = nan_to_zero(q);
    if (!(q >= 1e-5f)) { x = 0.f; y = 0.f; live_x = live_y = false; }
}

__device__ __forceinline__ bool kept_2d(const float* g, int j, int hips, int mask) {
    return !mask || j == hips || (g[0] != 0.f && g[1] != 0.f);
}

__global__ void __launch_bounds__(WARPS * 32) loss_fwd_kernel(const LossArgs a) {
    const int lane = threadIdx.x & 31;
    float s2 = 0.f, c2 = 0.f, s3 = 0.f;
    const long wstride = (long)gridDim.x * WARPS;
    for (long m = (long)blockIdx.x * WARPS + (threadIdx.x >> 5); m < a.M; m += wstride) {
        const float* p = a.proj + m * a.J * 3;
        const Norm n = norm_of(p, a.hips, a.neck);
        for (int j = lane; j < a.J; j += 32) {
            float x, y, q;
            bool lx, ly;
            pred_2d(p, j, n, a.normalize, x, y, q, lx, ly);
            if (a.proj_norm) {
                float* o = a.proj_norm + (m * a.J + j) * 3;
                o[0] = x; o[1] = y; o[2] = q;
            }
            const float* g = a.gt2 + (m * a.J + j) * a.gt_c;
            if (kept_2d(g, j, a.hips, a.mask)) {
                const float ex = x - g[0], ey = y - g[1];
                s2 = fmaf(ex, ex, fmaf(ey, ey, s2));
                c2 += 2.f;
            }
            const float* t = a.abs_loc + (m * a.J + j) * 3;
            const float* g3 = a.gt3 + (m * a.J + j) * 3;
            const float e0 = t[0] - g3[0], e1 = t[1] - g3[1], e2 = t[2] - g3[2];
            s3 = fmaf(e0, e0, fmaf(e1, e1, fmaf(e2, e2, s3)));
        }
    }
    __shared__ float sh[WARPS][3];
    s2 = warp_sum(s2); c2 = warp_sum(c2); s3 = warp_sum(s3);
    if (lane == 0) { sh[threadIdx.x >> 5][0] = s2; sh[threadIdx.x >> 5][1] = c2; sh[threadIdx.x >> 5][2] = s3; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float x = 0.f, y = 0.f, z = 0.f;
        for (int w = 0; w < WARPS; ++w) { x += sh[w][0]; y += sh[w][1]; z += sh[w][2]; }
        float* o = a.partials + blockIdx.x * 4;
        o[0] = x; o[1] = y; o[2] = z; o[3] = 0.f;
    }
}

}  // namespace

// shared with fused_carla.cu: partials (nblocks x 4: sum_sq_2d, count_2d, sum_sq_3d, -) -> losses[8]
__global__ void p2c_loss_finalize_kernel(const float* __restrict__ partials, int nblocks, double count3, float w2,
                                         float w3, float* __restrict__ losses, const XchgArgs xchg) {
    double s2 = 0.0, c2 = 0.0, s3 = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 32) {
        s2 += (double)partials[i * 4];
        c2 += (double)partials[i * 4 + 1];
        s3 += (double)partials[i * 4 + 2];
    }
    for (int o = 16; o > 0; o >>= 1) {
        s2 += __shfl_xor_sync(FULL_MASK, s2, o);
        c2 += __shfl_xor_sync(FULL_MASK, c2, o);
        s3 += __shfl_xor_sync(FULL_MASK, s3, o);
    }
    if (threadIdx.x == 0) {
        const float l2 = (float)(s2 / c2), l3 = (float)(s3 / count3);
        losses[0] = l2; losses[1] = l3; losses[2] = w2 * l2 + w3 * l3;
        losses[3] = (float)s2; losses[4] = (float)c2; losses[5] = (float)s3; losses[6] = (float)count3;
        losses[7] = 0.f;
        exchange_publish(xchg, (float)s2, (float)c2, (float)s3, (float)count3);
    }
}

namespace {

// rot_3d on the generic path: {loss, sum_sq, count} of the 9-channel reduction -> losses[8..11], weight into losses[2];
// `scale` receives weight_rot (the upstream factor of the reduction's gradient kernel)
__global__ void rot_merge_kernel(const float* __restrict__ rout, float wr, float* __restrict__ losses,
                                 float* __restrict__ scale) {
    losses[8] = rout[0]; losses[9] = rout[1]; losses[10] = rout[2]; losses[11] = 0.f;
    losses[2] += wr * rout[0];
    *scale = wr;
}

// d(w2*loc_2d + w3*loc_3d) / d(proj, abs_loc)
__global__ void __launch_bounds__(WARPS * 32) loss_bwd_kernel(const LossArgs a) {
    const int lane = threadIdx.x & 31;
    const float k2 = a.w2 * 2.0f / __ldg(a.losses + 4);
    const float k3 = a.w3 * 2.0f / __ldg(a.losses + 6);
    const long wstride = (long)gridDim.x * WARPS;
    for (long m = (long)blockIdx.x * WARPS + (threadIdx.x >> 5); m < a.M; m += wstride) {
        const float* p = a.proj + m * a.J * 3;
        float* gp = a.g_proj + m * a.J * 3;
        const Norm n = norm_of(p, a.hips, a.neck);
        float sgx = 0.f, sgy = 0.f, sgn = 0.f;
        for (int j = lane; j < a.J; j += 32) {
            float x, y, q;
            bool lx, ly;
            pred_2d(p, j, n, a.normalize, x, y, q, lx, ly);
            const float* g = a.gt2 + (m * a.J + j) * a.gt_c;
            float gx = 0.f, gy = 0.f;
            if (kept_2d(g, j, a.hips, a.mask)) {
                gx = lx ? k2 * (x - g[0]) : 0.f;
                gy = ly ? k2 * (y - g[1]) : 0.f;
            }
            if (a.normalize) {
                if (!n.ok) { gx = 0.f; gy = 0.f; }
                sgx += gx; sgy += gy;
                sgn += gx * x + gy * y;
                gx = n.ok ? gx / n.scale : 0.f;
                gy = n.ok ? gy / n.scale : 0.f;
            }
            gp[j * 3] = gx; gp[j * 3 + 1] = gy; gp[j * 3 + 2] = 0.f;
            const float* t = a.abs_loc + (m * a.J + j) * 3;
            const float* g3 = a.gt3 + (m * a.J + j) * 3;
            float* gl = a.g_abs_loc + (m * a.J + j) * 3;
            gl[0] = k3 * (t[0] - g3[0]); gl[1] = k3 * (t[1] - g3[1]); gl[2] = k3 * (t[2] - g3[2]);
        }
        if (a.normalize) {
            sgx = warp_sum(sgx); sgy = warp_sum(sgy); sgn = warp_sum(sgn);
            __syncwarp();
            if (lane == 0 && n.ok) {
                const float gscale = -sgn / n.scale;
                const float ux = n.dx / n.scale, uy = n.dy / n.scale;
                float hx = -sgx / n.scale - gscale * ux, hy = -sgy / n.scale - gscale * uy;
                float kx = gscale * ux, ky = gscale * uy;
                if (a.hips == a.neck) { hx += kx; hy += ky; kx = 0.f; ky = 0.f; }
                gp[a.hips * 3] += hx; gp[a.hips * 3 + 1] += hy;
                gp[a.neck * 3] += kx; gp[a.neck * 3 + 1] += ky;
            }
        }
    }
}

inline long align4(long x) { return (x + 3) & ~3L; }

}  // namespace

extern "C" int64_t p2c_fused_step_workspace_floats(int64_t n_poses, int32_t n_bones, int32_t input_6d) {
    long w = 16 + p2c_fused_fast_workspace_floats();        // zero-invariant header + fast-path partials
    w += LOSS_BLOCKS * 4;                                   // loss partials (generic path)
    w += p2c_mse_workspace_floats(n_poses) + 4;             // rot_3d partials + result (generic path)
    w += align4((long)n_poses * n_bones * 9);               // g_abs_rot (generic path with rot_3d)
    w += 2 * align4((long)n_poses * n_bones * 3);           // g_proj, g_abs_loc (generic path)
    if (input_6d) w += 2 * align4((long)n_poses * n_bones * 9);  // matrices, g_matrices
    return w;
}

extern "C" int p2c_fused_fast_launches(const p2c_fused_step_t* p);

extern "C" int p2c_fused_step_launches(const p2c_fused_step_t* p) {
    if (!p) return P2C_E_NULL;
    if (p->n_clips == 0) return 0;
    if (p2c_fused_tile_supported(p)) return p2c_fused_tile_launches(p);
    if (p2c_fused_fast_supported(p)) return p2c_fused_fast_launches(p);
    // fk_fwd, loss_fwd, finalize, loss_bwd, fk_bwd (+ rot6d fwd/bwd) (+ rot_3d: partial, finalize, merge, gradient)
    return 5 + (p->input_6d ? 2 : 0) + (p->gt_abs_rot ? 4 : 0);
}

extern "C" int p2c_fused_step(const p2c_fused_step_t* p, void* stream) {
    if (!p) return P2C_E_NULL;
    if (p->tree.n_bones < 1 || p->tree.n_bones > P2C_MAX_BONES) return P2C_E_BONES;
    const int B = p->tree.n_bones;
    if (p->n_clips < 0 || p->clip_len < 1 || p->gt_2d_channels < 2) return P2C_E_BADARG;
    if (p->hips < 0 || p->hips >= B || p->neck < 0 || p->neck >= B) return P2C_E_BADARG;
    if (p->n_clips == 0) return 0;
    if (!p->pose_in || !p->ref_loc || !p->gt_2d || !p->gt_3d || !p->g_pose_in || !p->workspace || !p->losses)
        return P2C_E_NULL;
    if (p->exchange && (p->exchange_world < 1 || p->exchange_world > XCHG_MAX_RANKS || p->exchange_rank < 0 ||
                        p->exchange_rank >= p->exchange_world || (reinterpret_cast<uintptr_t>(p->exchange) & 15)))
        return P2C_E_BADARG;
    if (p2c_fused_tile_supported(p)) return p2c_fused_tile(p, stream);
    if (p2c_fused_fast_supported(p)) return p2c_fused_fast(p, stream);

    if (!p->rel_rot && p->mode == P2C_MODE_POSE_CHANGES) return P2C_E_NULL;
    if (!p->abs_rot || !p->abs_loc || !p->proj) return P2C_E_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    const long M = (long)p->n_clips * p->clip_len;
    float* ws = p->workspace + 16 + p2c_fused_fast_workspace_floats();  // the header stays zero (fast path's counters)
    float* partials = ws; ws += LOSS_BLOCKS * 4;
    float* g_proj = ws; ws += align4(M * B * 3);
    float* g_abs_loc = ws; ws += align4(M * B * 3);
    const float* mats = p->pose_in;
    float* g_mats = p->g_pose_in;
    if (p->input_6d) {
        float* m = ws; ws += align4(M * B * 9);
        g_mats = ws; ws += align4(M * B * 9);
        int rc = p2c_rot6d_to_matrix_fwd(p->pose_in, m, M * B, stream);
        if (rc) return rc;
        mats = m;
    }
    p2c_fk_fwd_t f{};
    f.tree = p->tree; f.n_clips = p->n_clips; f.clip_len = p->clip_len; f.mode = p->mode;
    f.pose_in = mats; f.ref_loc = p->ref_loc; f.ref_rot = p->ref_rot; f.ref_index = p->ref_index;
    f.world_loc = p->world_loc; f.world_rot = p->world_rot; f.camera = p->camera;
    f.rel_rot = p->rel_rot; f.abs_rot = p->abs_rot; f.abs_loc = p->abs_loc; f.proj = p->proj;
    int rc = p2c_fk_fwd(&f, stream);
    if (rc) return rc;

    LossArgs a{};
    a.M = M; a.J = B; a.hips = p->hips; a.neck = p->neck; a.normalize = p->normalize;
    a.mask = p->mask_missing_joints; a.gt_c = p->gt_2d_channels;
    a.proj = p->proj; a.abs_loc = p->abs_loc; a.gt2 = p->gt_2d; a.gt3 = p->gt_3d; a.proj_norm = p->proj_norm;
    a.partials = partials; a.losses = p->losses; a.w2 = p->weight_2d; a.w3 = p->weight_3d;
    a.g_proj = g_proj; a.g_abs_loc = g_abs_loc;
    long blocks = (M + WARPS - 1) / WARPS;
    if (blocks > LOSS_BLOCKS) blocks = LOSS_BLOCKS;
    loss_fwd_kernel<<<(int)blocks, WARPS * 32, 0, st>>>(a);
    P2C_LAUNCH_CHECK();
    p2c_loss_finalize_kernel<<<1, 32, 0, st>>>(partials, (int)blocks, (double)M * B * 3, p->weight_2d, p->weight_3d,
                                               p->losses, XchgArgs{p->exchange, p->exchange_rank, p->exchange_world});
    P2C_LAUNCH_CHECK();
    loss_bwd_kernel<<<(int)blocks, WARPS * 32, 0, st>>>(a);
    P2C_LAUNCH_CHECK();

    float* g_abs_rot = nullptr;
    if (p->gt_abs_rot) {
        // rot_3d on the generic path: the 9-channel masked-MSE reduction (no mask) and its gradient as separate kernels,
        // merged into the step's loss block; the gradient enters the FK backward as g_abs_rot
        float* rpart = ws; ws += p2c_mse_workspace_floats(M);
        float* rout = ws; ws += 4;
        g_abs_rot = ws; ws += align4(M * B * 9);
        rc = p2c_masked_mse_fwd(p->abs_rot, p->gt_abs_rot, M, B, 9, 9, 9, -1, rpart, rout, stream);
        if (rc) return rc;
        rot_merge_kernel<<<1, 1, 0, st>>>(rout, p->weight_rot, p->losses, rout + 3);
        P2C_LAUNCH_CHECK();
        rc = p2c_masked_mse_bwd(p->abs_rot, p->gt_abs_rot, rout, rout + 3, g_abs_rot, M, B, 9, 9, 9, -1, stream);
        if (rc) return rc;
    }
    p2c_fk_bwd_t b{};
    b.g_abs_rot = g_abs_rot;
    b.tree = p->tree; b.n_clips = p->n_clips; b.clip_len = p->clip_len; b.mode = p->mode;
    b.pose_in = mats; b.ref_loc = p->ref_loc; b.ref_rot = p->ref_rot; b.ref_index = p->ref_index;
    b.world_loc = p->world_loc; b.world_rot = p->world_rot; b.camera = p->camera;
    b.rel_rot = p->rel_rot; b.abs_rot = p->abs_rot; b.abs_loc = p->abs_loc;
    b.g_abs_loc = g_abs_loc; b.g_proj = g_proj; b.g_pose_in = g_mats;
    rc = p2c_fk_bwd(&b, stream);
    if (rc) return rc;
    if (p->input_6d) {
        rc = p2c_rot6d_to_matrix_bwd(p->pose_in, g_mats, p->g_pose_in, M * B, stream);
        if (rc) return rc;
    }
    return 0;
}
