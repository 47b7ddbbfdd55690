// p2c_fused_tile: the pose-per-lane fused step for rotation_6d input (the pose_lifting training path).
//
// Same contract as fused_fast.cu (rotation_6d_to_matrix -> relative-rotation scan -> FK -> projection -> hips/neck
// normaliser -> loc_2d / loc_3d / loc_2d_3d and d(w2*loc_2d + w3*loc_3d)/d(6-D input) in one kernel), different
// mapping.  fused_fast.cu puts one bone on every lane and one pose on every warp: tree traffic goes over shuffles,
// 26 of 32 lanes work, FK by pointer jumping multiplies three full matrices per bone, subtree sums cost a 6-value
// warp scan: ~1000 warp instructions per pose, issue bound at a third of the HBM roofline.  Here a CTA owns a TILE
// of up to 64 consecutive poses (whole clips), all resident in shared memory, and
//   * every LANE is a POSE, every warp task is one bone (tree phases) or one bone pair (element-wise phases):
//     no shuffles, no redundant products (FK is exactly one 3x3 product per bone), all 32 lanes busy;
//   * the tree is walked level by level, tasks of a level spread over the warps, CTA barrier between levels;
//   * the two frame scans (forward product scan, backward suffix sum) switch to (clip, bone, column) threads that
//     run serially over the frames of their clip.
// HBM traffic is TMA only: cp.async.bulk global->shared for the tile's inputs, shared->global for every output
// region; abs_loc / projection are produced in place over the target records they were compared with, the gradient in
// place over the 6-D records (re-fetched, L2 hot, once the region has served as scratch), so a 32-pose CARLA tile
// is 104 KB and two CTAs share an SM.
//
// Shared-memory regions of a tile (PS = poses rounded up to 32):
//   D   PS*B*6   6-D records (AoS)            -> scratch, SoA rows [0,3B): abs_loc t_b, then subtree sum of g x t
//                                                         SoA rows [3B,6B): g_b = dL/d abs_loc[b], then subtree sum of g
//                                             -> 6-D records again (second TMA fetch) -> gradient records (in place)
//   G3  PS*B*3   3-D target records           -> abs_loc records (in place)
//   G2  PS*B*3   2-D target records (3 ch)    -> projection records (in place)
//   R   PS*B*9   rel_rot records (scan output; Gram-Schmidt output in relative_rot mode)
//   A   PS*B*9   C[t] (Gram-Schmidt output, scan input) -> abs_rot records -> word 0..2 of slot b: y_b, then z_b
// AoS pose records are even word counts (B even), so a lane reads its own record with 64-bit (128-bit for the 6-D
// records) accesses that are bank-conflict free across the 32 poses of a warp (record strides 117 / 39 / 39 vector
// units: odd); SoA scratch rows are conflict free by construction.
//
// The backward is the torque form derived in fused_fast.cu:  m_b = sum_{d in subtree(b)} g_d x (t_d - t_b),
// y_b = A_parent m_b, z[t] = sum_{s >= t} y[s], omega = R[t] z[t], tangent-form Gram-Schmidt backward.
#include <stdlib.h>

#include "common.cuh"

namespace {

constexpr int MAX_LEVELS = 24;
constexpr int COUNT_BLOCKS = 296;
constexpr int MAX_THREADS = 832;

struct TileArgs {
    int N, L, B, P, PS, slices, scan;   // clips, frames per clip, bones, poses per tile, P rounded up to 32, PS/32
    int hips, neck, normalize, mask, nlevels, npairs;
    int tree_warps;   // warps that take part in the level-synchronous tree phases
    int8_t parent[P2C_MAX_BONES];
    int8_t order[P2C_MAX_BONES];          // bones sorted by tree level
    int8_t lvl_start[MAX_LEVELS + 1];     // order[lvl_start[l] .. lvl_start[l+1]) = bones of level l
    int8_t child_start[P2C_MAX_BONES + 1];
    int8_t child_list[P2C_MAX_BONES];
    const float* d6;
    const float* ref_loc;
    const float* ref_rot;
    const int32_t* ref_index;
    const float* world_loc;
    const float* world_rot;
    p2c_camera_t cam;
    const float* gt2;
    const float* gt3;
    float w2, w3;
    float* rel_rot;
    float* abs_rot;
    float* abs_loc;
    float* proj;
    float* proj_norm;
    float* g_d6;
    float* partials;       // gridDim.x x 2
    unsigned int* ticket;  // self-resetting
    int* count2;           // self-resetting, filled by count_kernel when mask is on
    float* losses;
    double count3;
};

struct GS {
    V3 b1, b2, b3;
    float inv1, inv2, s;
};

__device__ __forceinline__ float dot3(const V3& a, const V3& b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ V3 cross3(const V3& a, const V3& b) {
    return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

// pytorch3d rotation_6d_to_matrix; F.normalize(v) = v / max(|v|, 1e-12) written as v * rsqrt(max(|v|^2, 1e-24))
__device__ __forceinline__ GS gram_schmidt(const float* a) {
    GS r;
    const V3 a1{a[0], a[1], a[2]};
    const V3 a2{a[3], a[4], a[5]};
    r.inv1 = rsqrtf(fmaxf(dot3(a1, a1), 1e-24f));
    r.b1 = V3{a1.x * r.inv1, a1.y * r.inv1, a1.z * r.inv1};
    r.s = dot3(r.b1, a2);
    const V3 u{fmaf(-r.s, r.b1.x, a2.x), fmaf(-r.s, r.b1.y, a2.y), fmaf(-r.s, r.b1.z, a2.z)};
    r.inv2 = rsqrtf(fmaxf(dot3(u, u), 1e-24f));
    r.b2 = V3{u.x * r.inv2, u.y * r.inv2, u.z * r.inv2};
    r.b3 = cross3(r.b1, r.b2);
    return r;
}

// Backward of gram_schmidt in tangent form (see fused_fast.cu): only the component of dL/dC along the tangent space
// of SO(3) at C matters, omega = vee(dL/dC . C^T).
__device__ __forceinline__ void gram_schmidt_bwd_tangent(const GS& r, const V3& w, float* out) {
    const float c3 = fmaf(w.x * r.s, r.inv2, w.y) * r.inv1, c2 = -w.z * r.inv1, e3 = -w.x * r.inv2;
    out[0] = fmaf(c3, r.b3.x, c2 * r.b2.x);
    out[1] = fmaf(c3, r.b3.y, c2 * r.b2.y);
    out[2] = fmaf(c3, r.b3.z, c2 * r.b2.z);
    out[3] = e3 * r.b3.x;
    out[4] = e3 * r.b3.y;
    out[5] = e3 * r.b3.z;
}

// 64-bit shared accesses as PTX: the two parity branches of lds9/sts9 read the same nine floats, and with plain
// float2 casts the optimiser merges them into ONE sequence (keeping one branch's alignment assumption), which
// faults on the other parity.  Opaque asm keeps the branches apart.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ float2 lds64(const float* p) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(smem_u32(p)) : "memory");
    return v;
}
__device__ __forceinline__ void sts64(float* p, float x, float y) {
    asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(smem_u32(p)), "f"(x), "f"(y) : "memory");
}
// 9-word bone record b of an AoS pose record whose base is 8-byte aligned: 64-bit accesses around the odd word
__device__ __forceinline__ M3 lds9(const float* pose, int b) {
    const float* p = pose + 9 * b;
    M3 r;
    if (b & 1) {
        r.m[0] = p[0];
        const float2 x0 = lds64(p + 1), x1 = lds64(p + 3), x2 = lds64(p + 5), x3 = lds64(p + 7);
        r.m[1] = x0.x; r.m[2] = x0.y; r.m[3] = x1.x; r.m[4] = x1.y;
        r.m[5] = x2.x; r.m[6] = x2.y; r.m[7] = x3.x; r.m[8] = x3.y;
    } else {
        const float2 x0 = lds64(p), x1 = lds64(p + 2), x2 = lds64(p + 4), x3 = lds64(p + 6);
        r.m[0] = x0.x; r.m[1] = x0.y; r.m[2] = x1.x; r.m[3] = x1.y;
        r.m[4] = x2.x; r.m[5] = x2.y; r.m[6] = x3.x; r.m[7] = x3.y;
        r.m[8] = p[8];
    }
    return r;
}
__device__ __forceinline__ void sts9(float* pose, int b, const M3& a) {
    float* p = pose + 9 * b;
    if (b & 1) {
        p[0] = a.m[0];
        sts64(p + 1, a.m[1], a.m[2]); sts64(p + 3, a.m[3], a.m[4]);
        sts64(p + 5, a.m[5], a.m[6]); sts64(p + 7, a.m[7], a.m[8]);
    } else {
        sts64(p, a.m[0], a.m[1]); sts64(p + 2, a.m[2], a.m[3]);
        sts64(p + 4, a.m[4], a.m[5]); sts64(p + 6, a.m[6], a.m[7]);
        p[8] = a.m[8];
    }
}

// utils/tensors.py:29-40 count of kept (joint, channel) pairs of the 2-D targets (integer => order independent)
__global__ void __launch_bounds__(256) tile_count_kernel(const float* __restrict__ gt2, long poses, int B, int gt_c,
                                                        int hips, int* __restrict__ count) {
    int c = 0;
    const long n = poses * B;
    for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long)gridDim.x * 256) {
        const int j = (int)(i % B);
        const float* g = gt2 + i * gt_c;
        if (j == hips || (__ldg(g) != 0.f && __ldg(g + 1) != 0.f)) c += 2;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(FULL_MASK, c, o);
    __shared__ int sh[8];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < 8; ++w) t += sh[w];
        if (t) atomicAdd(count, t);
    }
}

// ---- TMA / mbarrier primitives (cp.async.bulk, sm_90+)
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// shared-memory plan (floats); every region starts 16-byte aligned (PS is a multiple of 32)
struct TilePlan {
    int d, g3, g2, r, a, nrm, part, row, red, bars, total;
};
__device__ __host__ __forceinline__ TilePlan tile_plan(int PS, int B, int npairs) {
    TilePlan s;
    int o = 0;
    s.d = o; o += PS * B * 6;
    s.g3 = o; o += PS * B * 3;
    s.g2 = o; o += PS * B * 3;
    s.r = o; o += PS * B * 9;
    s.a = o; o += PS * B * 9;
    s.nrm = o; o += 8 * PS;            // sx, sy, dx, dy, inv, ok per pose
    s.part = o; o += npairs * 3 * PS;  // per bone pair: sum gx, sum gy, sum (gx x + gy y) per pose
    s.row = o; o += 2 * PS;            // skeleton-table row per pose, double buffered (next tile's rows prefetched)
    s.red = o; o += 2 * 32;            // loss partials per warp
    s.bars = o; o += 8;                // three 8-byte mbarriers
    s.total = o;
    return s;
}

struct Proj {
    V3 c;       // camera-space point
    float iz;   // 1 / c.z
    V3 p;       // (u, v, 1/z): what P3dPoseProjection.forward returns
};
__device__ __forceinline__ M3 ldg9(const float* g) {
    M3 r;
#pragma unroll
    for (int i = 0; i < 9; ++i) r.m[i] = __ldg(g + i);
    return r;
}
// p3d_pose_projection.py:137-152
__device__ __forceinline__ Proj project(const TileArgs& a, const V3& t, long m) {
    Proj r;
    V3 wp{t.y, -t.x, t.z};
    if (a.world_rot) wp = v3_mul_m(wp, ldg9(a.world_rot + m * 9));
    if (a.world_loc) {
        wp.x += __ldg(a.world_loc + m * 3); wp.y += __ldg(a.world_loc + m * 3 + 1); wp.z += __ldg(a.world_loc + m * 3 + 2);
    }
    const float* cr = a.cam.R;
    r.c.x = fmaf(wp.z, cr[6], fmaf(wp.y, cr[3], wp.x * cr[0])) + a.cam.T[0];
    r.c.y = fmaf(wp.z, cr[7], fmaf(wp.y, cr[4], wp.x * cr[1])) + a.cam.T[1];
    r.c.z = fmaf(wp.z, cr[8], fmaf(wp.y, cr[5], wp.x * cr[2])) + a.cam.T[2];
    r.iz = __fdividef(1.0f, r.c.z);
    r.p = V3{fmaf(-a.cam.f * r.c.x, r.iz, a.cam.px), fmaf(-a.cam.f * r.c.y, r.iz, a.cam.py), r.iz};
    return r;
}
// pixel-space gradient (gx, gy) -> gradient wrt the p3d-space point
__device__ __forceinline__ V3 project_bwd(const TileArgs& a, const Proj& r, float gx, float gy, long m) {
    const float fgx = a.cam.f * gx * r.iz, fgy = a.cam.f * gy * r.iz;
    const V3 gc{-fgx, -fgy, fmaf(fgx, r.c.x, fgy * r.c.y) * r.iz};
    const float* cr = a.cam.R;
    V3 gw{fmaf(gc.z, cr[2], fmaf(gc.y, cr[1], gc.x * cr[0])), fmaf(gc.z, cr[5], fmaf(gc.y, cr[4], gc.x * cr[3])),
          fmaf(gc.z, cr[8], fmaf(gc.y, cr[7], gc.x * cr[6]))};
    if (a.world_rot) gw = v3_mul_mt(gw, ldg9(a.world_rot + m * 9));
    return V3{-gw.y, gw.x, gw.z};
}

struct LossCtx {   // per-lane state of one element-wise task
    float sx, sy, inv, k2, k3;
    bool okn;
    float sgx, sgy, sgn;   // partial sums of this task for the normaliser backward
    float s2, s3;          // loss sums
};

// one bone of one pose: projection, normaliser, loc_2d / loc_3d terms, dL/d abs_loc (without the hips/neck terms of
// the normaliser backward, which the torque pass adds).  Returns the projection record through `pout`.
__device__ __forceinline__ V3 bone_loss(const TileArgs& a, int b, const V3& t, float g2x, float g2y, const V3& g3, long m,
                                        bool valid, LossCtx& cx, V3& pout, V3& nout) {
    const Proj pr = project(a, t, m);
    pout = pr.p;
    float x = pr.p.x, y = pr.p.y;
    bool lx = true, ly = true;
    if (a.normalize) {
        const float rx = (pr.p.x - cx.sx) * cx.inv, ry = (pr.p.y - cx.sy) * cx.inv;
        lx = fabsf(rx) <= 3.402823466e38f;
        ly = fabsf(ry) <= 3.402823466e38f;
        x = lx ? rx : 0.f;
        y = ly ? ry : 0.f;
        const float q = nan_to_zero(pr.p.z);
        if (!(q >= 1e-5f)) { x = 0.f; y = 0.f; lx = ly = false; }
        nout = V3{x, y, q};
    }
    const bool kept = valid && (!a.mask || b == a.hips || (g2x != 0.f && g2y != 0.f));
    float gx = 0.f, gy = 0.f;
    if (kept) {
        const float ex = x - g2x, ey = y - g2y;
        cx.s2 = fmaf(ex, ex, fmaf(ey, ey, cx.s2));
        gx = lx ? cx.k2 * ex : 0.f;
        gy = ly ? cx.k2 * ey : 0.f;
    }
    if (a.normalize) {
        if (!cx.okn) { gx = 0.f; gy = 0.f; }
        cx.sgx += gx;
        cx.sgy += gy;
        cx.sgn = fmaf(gx, x, fmaf(gy, y, cx.sgn));
        const float ie = cx.okn ? cx.inv : 0.f;
        gx *= ie;
        gy *= ie;
    }
    const V3 e{t.x - g3.x, t.y - g3.y, t.z - g3.z};
    if (valid) cx.s3 = fmaf(e.x, e.x, fmaf(e.y, e.y, fmaf(e.z, e.z, cx.s3)));
    const V3 gp = project_bwd(a, pr, gx, gy, m);
    return V3{fmaf(cx.k3, e.x, gp.x), fmaf(cx.k3, e.y, gp.y), fmaf(cx.k3, e.z, gp.z)};
}

__device__ __forceinline__ void named_barrier(int id, int threads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// BT / SL: bones and 32-pose slices per tile at compile time (0 = read from the arguments)
template <int BT, int SL>
__global__ void __launch_bounds__(MAX_THREADS, 1) tile_kernel(const TileArgs a) {
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NT = blockDim.x, W = NT >> 5;
    const int B = BT ? BT : a.B, slices = SL ? SL : a.slices, npairs = B / 2;
    const int B3 = B * 3, B6 = B * 6, B9 = B * 9, PS = 32 * slices, L = a.L;
    const TilePlan sm = tile_plan(PS, B, npairs);
    float* D = smem + sm.d;
    float* G3 = smem + sm.g3;     // G3 and G2 are adjacent: together they hold the re-fetched 6-D records
    float* G2 = smem + sm.g2;
    float* R = smem + sm.r;
    float* A = smem + sm.a;
    float* NRM = smem + sm.nrm;
    float* PART = smem + sm.part;
    int* ROW = reinterpret_cast<int*>(smem + sm.row);   // two buffers of PS
    float* red = smem + sm.red;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + sm.bars);   // [0] 6-D in, [1] targets in, [2] 6-D again
    float* TS = D;               // SoA scratch rows [0, 3B): t_b, later subtree sums of g x t
    float* GV = D + B3 * PS;     // SoA scratch rows [3B, 6B): g_b, later subtree sums of g
    const bool tma = tid == (W - 1) * 32;   // the thread that issues every bulk copy (a warp without tree work)
    const int TW = a.tree_warps;            // warps that walk the tree; the others skip to the next CTA barrier
    const long total = (long)a.N * L;
    const long ntiles = (total + a.P - 1) / a.P;

    auto load_d6 = [&](long tile) {
        const long m0 = tile * a.P;
        const uint32_t bytes = (uint32_t)((int)min((long)a.P, total - m0) * B6) * 4u;
        mbar_expect_tx(bars, bytes);
        bulk_g2s(D, a.d6 + m0 * B6, bytes, bars);
    };
    auto load_targets = [&](long tile) {
        const long m0 = tile * a.P;
        const uint32_t bytes = (uint32_t)((int)min((long)a.P, total - m0) * B3) * 4u;
        mbar_expect_tx(bars + 1, 2u * bytes);
        bulk_g2s(G3, a.gt3 + m0 * B3, bytes, bars + 1);
        bulk_g2s(G2, a.gt2 + m0 * B3, bytes, bars + 1);
    };
    auto fill_rows = [&](long tile, int buf) {   // skeleton-table row of every pose (lanes past the tile repeat its last)
        const long m0 = tile * a.P;
        const int np = (int)min((long)a.P, total - m0);
        for (int i = lane; i < PS; i += 32) {
            const long clip = (m0 + min(i, np - 1)) / L;
            ROW[buf * PS + i] = a.ref_index ? __ldg(a.ref_index + clip) : (int)clip;
        }
    };

    if (tma) {
        mbar_init(bars, 1);
        mbar_init(bars + 1, 1);
        mbar_init(bars + 2, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if ((long)blockIdx.x < ntiles) {
            load_d6(blockIdx.x);
            load_targets(blockIdx.x);
        }
    }
    if (warp == 0 && (long)blockIdx.x < ntiles) fill_rows(blockIdx.x, 0);
    __syncthreads();

    const float k2 = a.w2 * 2.0f / (a.mask ? (float)__ldcg(a.count2) : (float)(total * B * 2));
    const float k3 = a.w3 * 2.0f / (float)a.count3;
    float s2 = 0.f, s3 = 0.f;
    unsigned it = 0;

    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const long m0 = tile * a.P, nxt = tile + gridDim.x;
        const int np = (int)min((long)a.P, total - m0);
        const int* row = ROW + (it & 1) * PS;
        while (!mbar_try_wait(bars, it & 1u)) {}
        __syncthreads();   // #1 (first tile: the row table; later tiles: ordered by the barriers of the previous tile)

        // ---- phase A: 6-D -> C[t] (rotation_6d_to_matrix), two bones per task
        {
            float* dst = a.scan ? A : R;
            const int nt = slices * npairs;
            for (int task = warp; task < nt; task += W) {
                const int slice = task >= npairs ? 1 : 0, q = task - slice * npairs;
                const int col = slice * 32 + lane, pl = min(col, np - 1);
                const float4* dp = reinterpret_cast<const float4*>(D + pl * B6 + q * 12);
                const float4 u0 = dp[0], u1 = dp[1], u2 = dp[2];
                const float d[12] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w, u2.x, u2.y, u2.z, u2.w};
                const GS r0 = gram_schmidt(d), r1 = gram_schmidt(d + 6);
                if (col < np) {
                    float2* o = reinterpret_cast<float2*>(dst + pl * B9 + q * 18);
                    o[0] = make_float2(r0.b1.x, r0.b1.y); o[1] = make_float2(r0.b1.z, r0.b2.x);
                    o[2] = make_float2(r0.b2.y, r0.b2.z); o[3] = make_float2(r0.b3.x, r0.b3.y);
                    o[4] = make_float2(r0.b3.z, r1.b1.x); o[5] = make_float2(r1.b1.y, r1.b1.z);
                    o[6] = make_float2(r1.b2.x, r1.b2.y); o[7] = make_float2(r1.b2.z, r1.b3.x);
                    o[8] = make_float2(r1.b3.y, r1.b3.z);
                }
            }
        }
        if (!a.scan) fence_async_smem();
        __syncthreads();   // #2
        // ---- phase B: R[t] = C[t] . R[t-1] (p3d_pose.py:98-114 looped by layers/projection.py:190-193); one thread
        //      per (clip, bone, column of R), serial over the frames of the clip
        if (a.scan) {
            const int items = (np / L) * B3;
            for (int idx = tid; idx < items; idx += NT) {
                const int c = idx / B3, rem = idx - c * B3, b = rem / 3, j = rem - b * 3;
                const float* rr = a.ref_rot + ((long)row[c * L] * B + b) * 9 + j;
                float p0 = __ldg(rr), p1 = __ldg(rr + 3), p2 = __ldg(rr + 6);
                const float* src = A + (c * L) * B9 + b * 9;
                float* dst = R + (c * L) * B9 + b * 9 + j;
#pragma unroll 4
                for (int f = 0; f < L; ++f) {
                    const float n0 = fmaf(src[2], p2, fmaf(src[1], p1, src[0] * p0));
                    const float n1 = fmaf(src[5], p2, fmaf(src[4], p1, src[3] * p0));
                    const float n2 = fmaf(src[8], p2, fmaf(src[7], p1, src[6] * p0));
                    dst[0] = n0; dst[3] = n1; dst[6] = n2;
                    p0 = n0; p1 = n1; p2 = n2;
                    src += B9;
                    dst += B9;
                }
            }
            fence_async_smem();
            __syncthreads();   // #3
        }
        if (tma && a.rel_rot) {
            bulk_s2g(a.rel_rot + m0 * B9, R, (uint32_t)(np * B9) * 4u);
            bulk_commit();
        }

        // ---- phase C: FK, level by level (p3d_pose.py:116-184): A_b = R_b . A_parent, t_b = loc_b . A_parent + t_parent;
        //      only the first TW warps take part (named barrier 1 between levels)
        if (warp < TW) {
            for (int lv = 0; lv < a.nlevels; ++lv) {
                const int s0 = a.lvl_start[lv], cnt = a.lvl_start[lv + 1] - s0;
                const int nt = slices * cnt;
                for (int task = warp; task < nt; task += TW) {
                    const int slice = task >= cnt ? 1 : 0, b = a.order[s0 + task - slice * cnt], par = a.parent[b];
                    const int col = slice * 32 + lane, pl = min(col, np - 1);
                    const M3 Rb = lds9(R + pl * B9, b);
                    const float* rl = a.ref_loc + ((long)row[col] * B + b) * 3;
                    const V3 loc{__ldg(rl), __ldg(rl + 1), __ldg(rl + 2)};
                    M3 Ab = Rb;
                    V3 tb = loc;
                    if (par >= 0) {
                        const M3 Ap = lds9(A + pl * B9, par);
                        const V3 q = v3_mul_m(loc, Ap);
                        tb = V3{q.x + TS[(par * 3) * PS + col], q.y + TS[(par * 3 + 1) * PS + col],
                                q.z + TS[(par * 3 + 2) * PS + col]};
                        Ab = m3_mul(Rb, Ap);
                    }
                    if (col < np) sts9(A + pl * B9, b, Ab);
                    TS[(b * 3) * PS + col] = tb.x;
                    TS[(b * 3 + 1) * PS + col] = tb.y;
                    TS[(b * 3 + 2) * PS + col] = tb.z;
                }
                named_barrier(1, TW * 32);
            }
            // hips/neck shift and scale of every pose (extractor.py:23-36, hips_neck_extractor.py:6-13)
            if (a.normalize && warp < slices) {
                const int col = warp * 32 + lane;
                const long m = m0 + min(col, np - 1);
                const V3 th{TS[(a.hips * 3) * PS + col], TS[(a.hips * 3 + 1) * PS + col], TS[(a.hips * 3 + 2) * PS + col]};
                const V3 tn{TS[(a.neck * 3) * PS + col], TS[(a.neck * 3 + 1) * PS + col], TS[(a.neck * 3 + 2) * PS + col]};
                const Proj ph = project(a, th, m), pn = project(a, tn, m);
                const float dx = pn.p.x - ph.p.x, dy = pn.p.y - ph.p.y;
                const float scale = sqrtf(fmaf(dx, dx, dy * dy));
                NRM[col] = ph.p.x;
                NRM[PS + col] = ph.p.y;
                NRM[2 * PS + col] = dx;
                NRM[3 * PS + col] = dy;
                NRM[4 * PS + col] = __fdividef(1.0f, scale);
                NRM[5 * PS + col] = (scale > 0.f && scale <= 3.402823466e38f) ? 1.f : 0.f;
            }
            fence_async_smem();
        }
        __syncthreads();   // #4
        if (tma && a.abs_rot) {
            bulk_s2g(a.abs_rot + m0 * B9, A, (uint32_t)(np * B9) * 4u);
            bulk_commit();
        }

        // ---- projection, normaliser, loss terms, dL/d abs_loc; two bones per task; abs_loc and projection records are
        //      written over the target records they were compared with
        while (!mbar_try_wait(bars + 1, it & 1u)) {}
        {
            const int nt = slices * npairs;
            for (int task = warp; task < nt; task += W) {
                const int slice = task >= npairs ? 1 : 0, q = task - slice * npairs;
                const int col = slice * 32 + lane, pl = min(col, np - 1);
                const bool valid = col < np;
                const long m = m0 + pl;
                LossCtx cx;
                cx.k2 = k2; cx.k3 = k3; cx.s2 = s2; cx.s3 = s3;
                cx.sgx = cx.sgy = cx.sgn = 0.f;
                cx.sx = cx.sy = 0.f; cx.inv = 1.f; cx.okn = true;
                if (a.normalize) {
                    cx.sx = NRM[col]; cx.sy = NRM[PS + col]; cx.inv = NRM[4 * PS + col]; cx.okn = NRM[5 * PS + col] != 0.f;
                }
                float2* g3p = reinterpret_cast<float2*>(G3 + pl * B3 + q * 6);
                float2* g2p = reinterpret_cast<float2*>(G2 + pl * B3 + q * 6);
                const float2 h0 = g3p[0], h1 = g3p[1], h2 = g3p[2];
                const float2 e0 = g2p[0], e1 = g2p[1], e2 = g2p[2];
                const int b0 = 2 * q, b1 = 2 * q + 1;
                const V3 t0{TS[(b0 * 3) * PS + col], TS[(b0 * 3 + 1) * PS + col], TS[(b0 * 3 + 2) * PS + col]};
                const V3 t1{TS[(b1 * 3) * PS + col], TS[(b1 * 3 + 1) * PS + col], TS[(b1 * 3 + 2) * PS + col]};
                V3 p0, p1, n0, n1;
                const V3 ga = bone_loss(a, b0, t0, e0.x, e0.y, V3{h0.x, h0.y, h1.x}, m, valid, cx, p0, n0);
                const V3 gb = bone_loss(a, b1, t1, e1.y, e2.x, V3{h1.y, h2.x, h2.y}, m, valid, cx, p1, n1);
                s2 = cx.s2; s3 = cx.s3;
                if (valid) {
                    g3p[0] = make_float2(t0.x, t0.y); g3p[1] = make_float2(t0.z, t1.x); g3p[2] = make_float2(t1.y, t1.z);
                    g2p[0] = make_float2(p0.x, p0.y); g2p[1] = make_float2(p0.z, p1.x); g2p[2] = make_float2(p1.y, p1.z);
                    if (a.proj_norm && a.normalize) {   // off the benchmark path: straight to global
                        float* o = a.proj_norm + (m * B + b0) * 3;
                        o[0] = n0.x; o[1] = n0.y; o[2] = n0.z; o[3] = n1.x; o[4] = n1.y; o[5] = n1.z;
                    }
                }
                GV[(b0 * 3) * PS + col] = ga.x; GV[(b0 * 3 + 1) * PS + col] = ga.y; GV[(b0 * 3 + 2) * PS + col] = ga.z;
                GV[(b1 * 3) * PS + col] = gb.x; GV[(b1 * 3 + 1) * PS + col] = gb.y; GV[(b1 * 3 + 2) * PS + col] = gb.z;
                if (a.normalize) {
                    PART[(q * 3) * PS + col] = cx.sgx;
                    PART[(q * 3 + 1) * PS + col] = cx.sgy;
                    PART[(q * 3 + 2) * PS + col] = cx.sgn;
                }
            }
        }
        fence_async_smem();
        if (tma) bulk_wait_read_all();   // rel_rot / abs_rot stores have read R and A: A is about to take y_b
        __syncthreads();   // #5
        if (tma) {
            if (a.abs_loc) bulk_s2g(a.abs_loc + m0 * B3, G3, (uint32_t)(np * B3) * 4u);
            if (a.proj) bulk_s2g(a.proj + m0 * B3, G2, (uint32_t)(np * B3) * 4u);
            bulk_commit();
            bulk_wait_read_all();
            // the 6-D records come back (L2 hot) over the two target regions; phase E waits for them
            mbar_expect_tx(bars + 2, (uint32_t)(np * B6) * 4u);
            bulk_g2s(G3, a.d6 + m0 * B6, (uint32_t)(np * B6) * 4u, bars + 2);
        }

        // ---- torque backward, leaves first:  S_b = (g_b x t_b, g_b) + sum over children S_c  (in place over t_b, g_b),
        //      m_b = S_b[0:3] - S_b[3:6] x t_b,  y_b = A_parent m_b  -> words 0..2 of slot b of the A records
        if (warp < TW) {
            for (int lv = a.nlevels - 1; lv >= 0; --lv) {
                const int s0 = a.lvl_start[lv], cnt = a.lvl_start[lv + 1] - s0;
                const int nt = slices * cnt;
                for (int task = warp; task < nt; task += TW) {
                    const int slice = task >= cnt ? 1 : 0, b = a.order[s0 + task - slice * cnt], par = a.parent[b];
                    const int col = slice * 32 + lane, pl = min(col, np - 1);
                    const V3 t{TS[(b * 3) * PS + col], TS[(b * 3 + 1) * PS + col], TS[(b * 3 + 2) * PS + col]};
                    V3 g{GV[(b * 3) * PS + col], GV[(b * 3 + 1) * PS + col], GV[(b * 3 + 2) * PS + col]};
                    if (a.normalize && (b == a.hips || b == a.neck)) {
                        // normaliser backward through shift (hips) and scale (|neck - hips|): sums over all bones
                        float sgx = 0.f, sgy = 0.f, sgn = 0.f;
                        for (int q = 0; q < npairs; ++q) {
                            sgx += PART[(q * 3) * PS + col];
                            sgy += PART[(q * 3 + 1) * PS + col];
                            sgn += PART[(q * 3 + 2) * PS + col];
                        }
                        const float inv = NRM[5 * PS + col] != 0.f ? NRM[4 * PS + col] : 0.f;
                        const float gs = -sgn * inv * inv;   // dL/dscale . (1/scale)
                        const float kx = gs * NRM[2 * PS + col], ky = gs * NRM[3 * PS + col];
                        float cx = 0.f, cy = 0.f;
                        if (b == a.neck) { cx += kx; cy += ky; }
                        if (b == a.hips) { cx -= fmaf(sgx, inv, kx); cy -= fmaf(sgy, inv, ky); }
                        const long m = m0 + pl;
                        const V3 gp = project_bwd(a, project(a, t, m), cx, cy, m);
                        g.x += gp.x; g.y += gp.y; g.z += gp.z;
                    }
                    float S[6] = {g.y * t.z - g.z * t.y, g.z * t.x - g.x * t.z, g.x * t.y - g.y * t.x, g.x, g.y, g.z};
                    for (int ci = a.child_start[b]; ci < a.child_start[b + 1]; ++ci) {
                        const int c = a.child_list[ci];
                        S[0] += TS[(c * 3) * PS + col]; S[1] += TS[(c * 3 + 1) * PS + col]; S[2] += TS[(c * 3 + 2) * PS + col];
                        S[3] += GV[(c * 3) * PS + col]; S[4] += GV[(c * 3 + 1) * PS + col]; S[5] += GV[(c * 3 + 2) * PS + col];
                    }
                    if (par >= 0) {   // the root's sums are not read by anyone
                        TS[(b * 3) * PS + col] = S[0]; TS[(b * 3 + 1) * PS + col] = S[1]; TS[(b * 3 + 2) * PS + col] = S[2];
                        GV[(b * 3) * PS + col] = S[3]; GV[(b * 3 + 1) * PS + col] = S[4]; GV[(b * 3 + 2) * PS + col] = S[5];
                    }
                    const V3 mv{S[0] - (S[4] * t.z - S[5] * t.y), S[1] - (S[5] * t.x - S[3] * t.z),
                                S[2] - (S[3] * t.y - S[4] * t.x)};
                    V3 yv = mv;
                    if (par >= 0) yv = v3_mul_mt(mv, lds9(A + pl * B9, par));
                    if (col < np) {
                        float* o = A + pl * B9 + b * 9;
                        o[0] = yv.x; o[1] = yv.y; o[2] = yv.z;
                    }
                }
                if (lv > 0) named_barrier(1, TW * 32);
            }
            fence_async_smem();   // the scratch rows are about to be overwritten by the next tile's 6-D records (TMA)
        }
        __syncthreads();   // #6
        if (nxt < ntiles) {   // the D region and the other row buffer are free: prefetch for the next tile
            if (tma) load_d6(nxt);
            if (warp == W - 1) fill_rows(nxt, (it + 1) & 1);
        }
        // ---- phase D: z[t] = sum_{s >= t} y[s] per clip (the reverse-time scan in tangent form is a suffix sum)
        if (a.scan) {
            const int items = (np / L) * B3;
            for (int idx = tid; idx < items; idx += NT) {
                const int c = idx / B3, rem = idx - c * B3, b = rem / 3, j = rem - b * 3;
                float* y = A + (c * L + L - 1) * B9 + b * 9 + j;
                float acc = 0.f;
#pragma unroll 4
                for (int f = L - 1; f >= 0; --f) {
                    acc += *y;
                    *y = acc;
                    y -= B9;
                }
            }
        }
        while (!mbar_try_wait(bars + 2, it & 1u)) {}
        __syncthreads();   // #7
        // ---- phase E: omega[t] = R[t] z[t]; tangent-form Gram-Schmidt backward -> gradient records over the 6-D records
        {
            const int nt = slices * npairs;
            for (int task = warp; task < nt; task += W) {
                const int slice = task >= npairs ? 1 : 0, q = task - slice * npairs;
                const int col = slice * 32 + lane, pl = min(col, np - 1);
                float4* dp = reinterpret_cast<float4*>(G3 + pl * B6 + q * 12);
                const float4 u0 = dp[0], u1 = dp[1], u2 = dp[2];
                const float d[12] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w, u2.x, u2.y, u2.z, u2.w};
                const float2* rp = reinterpret_cast<const float2*>(R + pl * B9 + q * 18);
                const float2 r0 = rp[0], r1 = rp[1], r2 = rp[2], r3 = rp[3], r4 = rp[4], r5 = rp[5], r6 = rp[6], r7 = rp[7],
                             r8 = rp[8];
                const float* zp = A + pl * B9 + q * 18;
                const float2 z01 = *reinterpret_cast<const float2*>(zp);
                const float z2 = zp[2], z3 = zp[9];
                const float2 z45 = *reinterpret_cast<const float2*>(zp + 10);
                M3 Ra, Rb;
                Ra.m[0] = r0.x; Ra.m[1] = r0.y; Ra.m[2] = r1.x; Ra.m[3] = r1.y; Ra.m[4] = r2.x; Ra.m[5] = r2.y;
                Ra.m[6] = r3.x; Ra.m[7] = r3.y; Ra.m[8] = r4.x;
                Rb.m[0] = r4.y; Rb.m[1] = r5.x; Rb.m[2] = r5.y; Rb.m[3] = r6.x; Rb.m[4] = r6.y; Rb.m[5] = r7.x;
                Rb.m[6] = r7.y; Rb.m[7] = r8.x; Rb.m[8] = r8.y;
                const V3 oa = v3_mul_mt(V3{z01.x, z01.y, z2}, Ra), ob = v3_mul_mt(V3{z3, z45.x, z45.y}, Rb);
                float o[12];
                gram_schmidt_bwd_tangent(gram_schmidt(d), oa, o);
                gram_schmidt_bwd_tangent(gram_schmidt(d + 6), ob, o + 6);
                if (col < np) {
                    dp[0] = make_float4(o[0], o[1], o[2], o[3]);
                    dp[1] = make_float4(o[4], o[5], o[6], o[7]);
                    dp[2] = make_float4(o[8], o[9], o[10], o[11]);
                }
            }
        }
        fence_async_smem();
        __syncthreads();   // #8
        if (tma) {
            bulk_s2g(a.g_d6 + m0 * B6, G3, (uint32_t)(np * B6) * 4u);
            bulk_commit();
            bulk_wait_read_all();   // the gradient has left the target regions: the next tile's targets may land
            if (nxt < ntiles) load_targets(nxt);
        }
    }
    if (tma) bulk_wait_all();

    // ---- loss sums: fixed-order reduction, the last CTA to finish folds the per-CTA partials
    s2 = warp_sum(s2);
    s3 = warp_sum(s3);
    if (lane == 0) { red[warp * 2] = s2; red[warp * 2 + 1] = s3; }
    __syncthreads();
    __shared__ bool last;
    if (tid == 0) {
        float x = 0.f, y = 0.f;
        for (int w = 0; w < W; ++w) { x += red[w * 2]; y += red[w * 2 + 1]; }
        a.partials[blockIdx.x * 2] = x;
        a.partials[blockIdx.x * 2 + 1] = y;
        __threadfence();
        last = (atomicAdd(a.ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (last && warp == 0) {
        __threadfence();
        double d2 = 0.0, d3 = 0.0;
        for (int i = lane; i < (int)gridDim.x; i += 32) {
            d2 += (double)__ldcg(a.partials + i * 2);
            d3 += (double)__ldcg(a.partials + i * 2 + 1);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            d2 += __shfl_xor_sync(FULL_MASK, d2, o);
            d3 += __shfl_xor_sync(FULL_MASK, d3, o);
        }
        if (lane == 0) {
            const double c2 = a.mask ? (double)*a.count2 : (double)total * B * 2;
            const float l2 = (float)(d2 / c2), l3 = (float)(d3 / a.count3);
            a.losses[0] = l2; a.losses[1] = l3; a.losses[2] = a.w2 * l2 + a.w3 * l3;
            a.losses[3] = (float)d2; a.losses[4] = (float)c2; a.losses[5] = (float)d3; a.losses[6] = (float)a.count3;
            a.losses[7] = 0.f;
            *a.ticket = 0u;   // self-resetting workspace header: ready for the next launch
            *a.count2 = 0;
        }
    }
}

int g_sms = 0;
int sm_count() {
    if (!g_sms) {
        int dev = 0, n = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
            g_sms = n;
        else
            g_sms = 148;
    }
    return g_sms;
}

// level order + child lists of the bone tree (parents precede children: pose.py:41-47)
bool tree_tables(const p2c_tree_t& t, TileArgs& a) {
    const int B = t.n_bones;
    if (B < 2 || B > P2C_MAX_BONES || (B & 1)) return false;   // AoS records must be even word counts
    int level[P2C_MAX_BONES], maxl = 0;
    if (t.parent[0] != -1) return false;
    level[0] = 0;
    for (int b = 1; b < B; ++b) {
        const int p = t.parent[b];
        if (p < 0 || p >= b) return false;
        level[b] = level[p] + 1;
        if (level[b] > maxl) maxl = level[b];
    }
    if (maxl + 1 > MAX_LEVELS) return false;
    a.nlevels = maxl + 1;
    int n = 0;
    for (int l = 0; l <= maxl; ++l) {
        a.lvl_start[l] = (int8_t)n;
        for (int b = 0; b < B; ++b)
            if (level[b] == l) a.order[n++] = (int8_t)b;
    }
    a.lvl_start[maxl + 1] = (int8_t)n;
    a.tree_warps = 1;   // widest level (x slices, set by the launcher)
    for (int l = 0; l <= maxl; ++l)
        if (a.lvl_start[l + 1] - a.lvl_start[l] > a.tree_warps) a.tree_warps = a.lvl_start[l + 1] - a.lvl_start[l];
    n = 0;
    for (int b = 0; b < B; ++b) {
        a.parent[b] = (int8_t)t.parent[b];
        a.child_start[b] = (int8_t)n;
        for (int c = b + 1; c < B; ++c)
            if (t.parent[c] == b) a.child_list[n++] = (int8_t)c;
    }
    a.child_start[B] = (int8_t)n;
    a.B = B;
    a.npairs = B / 2;
    return true;
}

struct TileLaunch {
    int ok, P, PS, slices, threads, grid;
    size_t smem;
};

TileLaunch make_launch(const p2c_fused_step_t* p) {
    TileLaunch t{};
    const int B = p->tree.n_bones, L = p->clip_len;
    const long total = (long)p->n_clips * L;
    if (p->mode == P2C_MODE_POSE_CHANGES) {
        if (L > 64) return t;
        t.P = L <= 32 ? (32 / L) * L : L;   // whole clips: as many as fit the 32 lanes, or one clip over two slices
    } else {
        t.P = 32;
    }
    if ((long)t.P > total) t.P = (int)total;
    // every tile moves by TMA: 16-byte aligned records on both sides => even pose counts per tile (B*3 floats per
    // pose record is an even, not a multiple-of-four, count) and 16-byte aligned tensors; anything else takes the
    // bone-per-lane kernels of fused_fast.cu, which also copy cooperatively
    if ((t.P & 1) || (total & 1)) return t;
    const void* ptrs[] = {p->pose_in, p->gt_2d, p->gt_3d, p->g_pose_in, p->rel_rot, p->abs_rot, p->abs_loc, p->proj};
    for (const void* q : ptrs)
        if (reinterpret_cast<uintptr_t>(q) & 15) return t;
    t.PS = (t.P + 31) & ~31;
    t.slices = t.PS / 32;
    t.smem = (size_t)tile_plan(t.PS, B, B / 2).total * sizeof(float);
    const size_t cap = 227 * 1024;
    if (t.smem + 1024 > cap) return t;
    int warps = (B / 2) * t.slices;
    if (warps < 4) warps = 4;
    if (warps > MAX_THREADS / 32) warps = MAX_THREADS / 32;
    t.threads = warps * 32;
    int per_sm = (int)((228 * 1024) / (t.smem + 1024));
    if (per_sm * t.threads > 1664) per_sm = 1664 / t.threads;   // 78 registers per thread (launch bounds 832 x 1)
    if (per_sm > 4) per_sm = 4;   // the workspace holds 2 x 592 loss partials
    if (per_sm < 1) per_sm = 1;
    const long ntiles = (total + t.P - 1) / t.P;
    const long slots = (long)sm_count() * per_sm;
    t.grid = (int)(ntiles < slots ? ntiles : slots);
    t.ok = 1;
    return t;
}

bool env_on(const char* name) {
    const char* v = getenv(name);
    return v && v[0] == '1';
}

}  // namespace

extern "C" int p2c_fused_tile_supported(const p2c_fused_step_t* p) {
    // test hooks: the older kernel variants stay reachable
    if (env_on("P2C_DISABLE_TILE") || env_on("P2C_DISABLE_FAST") || env_on("P2C_FORCE_SLIM") || env_on("P2C_DISABLE_TMA"))
        return 0;
    if (!p->input_6d) return 0;   // arbitrary 3x3 input is not assumed to be a rotation: exact generic path
    if (p->mode != P2C_MODE_POSE_CHANGES && p->mode != P2C_MODE_RELATIVE_ROT) return 0;
    if (p->mode == P2C_MODE_POSE_CHANGES && !p->ref_rot) return 0;
    if (p->gt_2d_channels != 3) return 0;   // the projection records replace the 2-D target records in place
    TileArgs a{};
    if (!tree_tables(p->tree, a)) return 0;
    return make_launch(p).ok;
}

extern "C" int p2c_fused_tile_launches(const p2c_fused_step_t* p) { return p->mask_missing_joints ? 2 : 1; }

extern "C" int p2c_fused_tile(const p2c_fused_step_t* p, void* stream) {
    TileArgs a{};
    if (!tree_tables(p->tree, a)) return P2C_E_UNSUPPORTED;
    const TileLaunch t = make_launch(p);
    if (!t.ok) return P2C_E_UNSUPPORTED;
    const int B = p->tree.n_bones;
    a.N = p->n_clips; a.L = p->clip_len; a.P = t.P; a.PS = t.PS; a.slices = t.slices;
    a.tree_warps *= t.slices;
    if (a.tree_warps > t.threads / 32) a.tree_warps = t.threads / 32;
    a.scan = p->mode == P2C_MODE_POSE_CHANGES;
    a.hips = p->hips; a.neck = p->neck; a.normalize = p->normalize; a.mask = p->mask_missing_joints;
    a.d6 = p->pose_in; a.ref_loc = p->ref_loc; a.ref_rot = p->ref_rot; a.ref_index = p->ref_index;
    a.world_loc = p->world_loc; a.world_rot = p->world_rot; a.cam = p->camera;
    a.gt2 = p->gt_2d; a.gt3 = p->gt_3d; a.w2 = p->weight_2d; a.w3 = p->weight_3d;
    a.rel_rot = p->rel_rot; a.abs_rot = p->abs_rot; a.abs_loc = p->abs_loc; a.proj = p->proj;
    a.proj_norm = p->proj_norm; a.g_d6 = p->g_pose_in;
    a.ticket = reinterpret_cast<unsigned int*>(p->workspace);
    a.count2 = reinterpret_cast<int*>(p->workspace) + 1;
    a.partials = p->workspace + 16;
    a.losses = p->losses;
    const long total = (long)p->n_clips * p->clip_len;
    a.count3 = (double)total * B * 3;
    cudaStream_t st = (cudaStream_t)stream;
    if (a.mask) {
        long blocks = (total * B + 255) / 256;
        if (blocks > COUNT_BLOCKS) blocks = COUNT_BLOCKS;
        tile_count_kernel<<<(int)blocks, 256, 0, st>>>(p->gt_2d, total, B, p->gt_2d_channels, p->hips, a.count2);
        P2C_LAUNCH_CHECK();
    }
    typedef void (*kern_t)(const TileArgs);
    const int ki = B == 26 ? (t.slices == 1 ? 0 : 1) : 2;   // CARLA_SKELETON gets compile-time bone and slice counts
    const kern_t k = ki == 0 ? tile_kernel<26, 1> : ki == 1 ? tile_kernel<26, 2> : tile_kernel<0, 0>;
    static bool configured[3] = {false, false, false};
    if (!configured[ki]) {
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024);
        if (e != cudaSuccess) return (int)e;
        configured[ki] = true;
    }
    k<<<t.grid, t.threads, t.smem, st>>>(a);
    P2C_LAUNCH_CHECK();
    return 0;
}
