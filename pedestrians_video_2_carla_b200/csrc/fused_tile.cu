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
//
// Two launch shapes.  MANY tiles per CTA: 13 task warps (bones in (even, odd) pairs per warp), two CTAs
// per SM, tiles drawn from a grid-wide counter, mask count by a pre-pass kernel, loss partials folded by the last CTA
// to take a ticket.  At most ONE tile per CTA (the 256 x 16 step): 26 task warps (a bone each) + a service warp that
// publishes the CTA's loss partial as self-validating entries and, in CTA 0, folds the grid's; mask count inside the
// kernel; steps chained by programmatic dependent launch with the set-up and L2 prefetches ahead of the dependency
// wait.  Narrow phases (frame scan: 5 warps, tree walk: 5 or 15) are as long as the instruction count of their busiest
// warp scheduler: the shift / scale of the normaliser is formed beside the walk by a spare warp, the loss fold beside
// the backward phases by the service warp.
#include <stdlib.h>

#include "common.cuh"
#include "host_util.h"
#include "tile_common.cuh"

#ifndef P2C_ENTRY_PREFETCH
#define P2C_ENTRY_PREFETCH 1   // 0: experiment, no L2 prefetch of the first tile ahead of the dependency wait
#endif
#ifndef P2C_POLL_FOLD
#define P2C_POLL_FOLD 1   // 0: experiment, the service warp of the LAST CTA folds (ticket scheme)
#endif
#ifndef P2C_EARLY_NORM
#define P2C_EARLY_NORM 1   // 0: experiment, shift / scale after the forward walk
#endif
#ifndef P2C_EARLY_NORM_MANY
#define P2C_EARLY_NORM_MANY 1   // 0: experiment, early shift / scale only in the one-tile kernels
#endif
#ifndef P2C_MERGE_YE
#define P2C_MERGE_YE 1   // 0: experiment, y in a pass of its own in the one-tile kernels too
#endif
#ifndef P2C_SERVICE_WARP
#define P2C_SERVICE_WARP 1   // 0: experiment, the loss fold stays in task warp W - 2
#endif
#ifndef P2C_LONE_TASKS
// one-bone-per-warp passes (the one-tile-per-CTA kernels): 2 = the body compiled once per parity behind a warp-uniform
// branch, 0 = parity looked at inside the accessors (measured: 14.20 -> 14.02 us at 256 x 16)
#define P2C_LONE_TASKS 2
#endif

namespace {

constexpr int MAX_THREADS = 832;                   // worker threads: at most 26 warps with bone / column tasks
constexpr int LAUNCH_THREADS = MAX_THREADS + 32;   // + the service warp of the one-tile-per-CTA CARLA kernels

struct TileArgs {
    int N, L, B, P, PS, slices, scan;   // clips, frames per clip, bones, poses per tile, P rounded up to 32, PS/32
    int hips, neck, normalize, mask;
    int count_inline; // every CTA owns exactly one tile: the mask count happens inside the kernel (no pre-pass)
    int clock_cta;    // debug (P2C_TILE_CTA=k): the CTA that records (default 0)
    int clocks;       // debug (P2C_TILE_CLOCKS=n): CTA 0 records clock64() at every phase boundary of its n-th tile
    int8_t parent[P2C_MAX_BONES];
    const float* d6;
    const float* ref_loc;
    const float* ref_rot;
    const int32_t* ref_index;
    const float* world_loc;
    const float* world_rot;
    p2c_camera_t cam;
    const float* gt2;
    const float* gt3;
    const float* gt_rot;   // targets['absolute_pose_rot'] (N,L,B,3,3) or NULL: rot_3d fused (loss/rot_3d.py:9-37)
    float w2, w3, wr;
    float* rel_rot;
    float* abs_rot;
    float* abs_loc;
    float* proj;
    float* proj_norm;
    float* g_d6;
    float* partials;       // tiles x 4: sum_sq_2d, sum_sq_3d, sum_sq_rot, -
    unsigned int* ticket;  // self-resetting
    int* count2;           // self-resetting, filled by count_kernel when mask is on
    unsigned int* arrive;  // self-resetting: CTAs that have added their part of count2 (count_inline)
    unsigned int* next;    // self-resetting: tiles handed out beyond the first of every CTA (dynamic)
    int warp_fold;         // few tiles: the loss partials are folded by one warp inside the tile loop
    int dynamic;           // more tiles than CTAs: tiles beyond a CTA's first are drawn from `next`
    float* losses;
    double count3;
    XchgArgs xchg;         // data-parallel ranks: where the finishing thread publishes the four loss sums
    // per-launch constants worked out by the launcher: read from the constant bank where they are used instead of
    // being computed at the top of the kernel and held in registers across the tile loop (the kernel runs at its
    // register cap: every long-lived value is a spill somewhere else)
    int total, ntiles, clips_full, whole_clips;
    float k2_all, k3, kr;
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
    r.inv1 = frsqrt(fmaxf(dot3(a1, a1), 1e-24f));
    r.b1 = V3{a1.x * r.inv1, a1.y * r.inv1, a1.z * r.inv1};
    r.s = dot3(r.b1, a2);
    const V3 u{fmaf(-r.s, r.b1.x, a2.x), fmaf(-r.s, r.b1.y, a2.y), fmaf(-r.s, r.b1.z, a2.z)};
    r.inv2 = frsqrt(fmaxf(dot3(u, u), 1e-24f));
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

// utils/tensors.py:29-40 count of kept (joint, channel) pairs of the 2-D targets (integer => order independent).
// The 3-channel records are read as the flat float4 stream they are: four joints per 48-byte group (the tile path
// only takes 16-byte aligned tensors with an even bone and pose count, so the joint count is a multiple of four).
__global__ void __launch_bounds__(256) tile_count_kernel(const float* __restrict__ gt2, unsigned joints, unsigned B,
                                                        unsigned gt_c, unsigned hips, int* __restrict__ count) {
    // programmatic dependent launch: the step kernel may start its prologue, input loads and forward phases now; it
    // executes griddepcontrol.wait (= this grid complete and flushed) before it reads the count
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    int c = 0;
    const float4* q4 = reinterpret_cast<const float4*>(gt2);
    const unsigned groups = joints >> 2;
#pragma unroll 2
    for (unsigned q = blockIdx.x * 256u + threadIdx.x; q < groups; q += gridDim.x * 256u) {
        const float4 v0 = __ldg(q4 + 3 * (size_t)q), v1 = __ldg(q4 + 3 * (size_t)q + 1), v2 = __ldg(q4 + 3 * (size_t)q + 2);
        const unsigned j = (q << 2) % B;   // bone of the group's first joint
        auto wrap = [B](unsigned h) {      // (h < B + 3 and B >= 2: two conditional subtractions reduce it)
            h -= h >= B ? B : 0u;
            h -= h >= B ? B : 0u;
            return h;
        };
        const unsigned h0 = j, h1 = wrap(j + 1), h2 = wrap(j + 2), h3 = wrap(j + 3);
        if (h0 == hips || (v0.x != 0.f && v0.y != 0.f)) c += 2;
        if (h1 == hips || (v0.w != 0.f && v1.x != 0.f)) c += 2;
        if (h2 == hips || (v1.z != 0.f && v1.w != 0.f)) c += 2;
        if (h3 == hips || (v2.y != 0.f && v2.z != 0.f)) c += 2;
    }
    c = __reduce_add_sync(FULL_MASK, c);
    __shared__ int sh[8];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < 8; ++w) t += sh[w];
        if (t) atomicAdd(count, t);
    }
}

// shared-memory plan (floats); every region starts 16-byte aligned (PS is a multiple of 32)
struct TilePlan {
    int d, g3, g2, r, a, nrm, part, row, red, bars, total;
};
__device__ __host__ __forceinline__ TilePlan tile_plan(int PS, int B) {
    TilePlan s;
    int o = 0;
    s.d = o; o += PS * B * 6;
    s.g3 = o; o += PS * B * 3;
    s.g2 = o; o += PS * B * 3;
    s.r = o; o += PS * B * 9;
    s.a = o; o += PS * B * 9;
    s.nrm = o; o += 10 * PS;           // sx, sy, dx, dy, inv, ok, then the three normaliser sums, per pose
    s.part = o; o += B * 3 * PS;       // per bone: gx, gy, gx x + gy y per pose (normaliser backward), then Sx rows
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
template <bool WORLD>
__device__ __forceinline__ Proj project(const TileArgs& a, const V3& t, long m) {
    Proj r;
    V3 wp{t.y, -t.x, t.z};
    if (WORLD && a.world_rot) wp = v3_mul_m(wp, ldg9(a.world_rot + m * 9));
    if (WORLD && a.world_loc) {
        wp.x += __ldg(a.world_loc + m * 3); wp.y += __ldg(a.world_loc + m * 3 + 1); wp.z += __ldg(a.world_loc + m * 3 + 2);
    }
    const float* cr = a.cam.R;
    r.c.x = fmaf(wp.z, cr[6], fmaf(wp.y, cr[3], wp.x * cr[0])) + a.cam.T[0];
    r.c.y = fmaf(wp.z, cr[7], fmaf(wp.y, cr[4], wp.x * cr[1])) + a.cam.T[1];
    r.c.z = fmaf(wp.z, cr[8], fmaf(wp.y, cr[5], wp.x * cr[2])) + a.cam.T[2];
    r.iz = frcp(r.c.z);
    r.p = V3{fmaf(-a.cam.f * r.c.x, r.iz, a.cam.px), fmaf(-a.cam.f * r.c.y, r.iz, a.cam.py), r.iz};
    return r;
}
// pixel-space gradient (gx, gy) -> gradient wrt the p3d-space point
template <bool WORLD>
__device__ __forceinline__ V3 project_bwd(const TileArgs& a, const Proj& r, float gx, float gy, long m) {
    const float fgx = a.cam.f * gx * r.iz, fgy = a.cam.f * gy * r.iz;
    const V3 gc{-fgx, -fgy, fmaf(fgx, r.c.x, fgy * r.c.y) * r.iz};
    const float* cr = a.cam.R;
    V3 gw{fmaf(gc.z, cr[2], fmaf(gc.y, cr[1], gc.x * cr[0])), fmaf(gc.z, cr[5], fmaf(gc.y, cr[4], gc.x * cr[3])),
          fmaf(gc.z, cr[8], fmaf(gc.y, cr[7], gc.x * cr[6]))};
    if (WORLD && a.world_rot) gw = v3_mul_mt(gw, ldg9(a.world_rot + m * 9));
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
template <bool WORLD>
__device__ __forceinline__ V3 bone_loss(const TileArgs& a, int b, const V3& t, float g2x, float g2y, const V3& g3, long m,
                                        bool valid, LossCtx& cx, V3& pout, V3& nout) {
    const Proj pr = project<WORLD>(a, t, m);
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
    const V3 gp = project_bwd<WORLD>(a, pr, gx, gy, m);
    return V3{fmaf(cx.k3, e.x, gp.x), fmaf(cx.k3, e.y, gp.y), fmaf(cx.k3, e.z, gp.z)};
}

__device__ long long g_tile_clocks[64];
__device__ long long g_warp_clocks[32 * 16];   // debug: arrival of every warp at the barriers of the recorded tile
__device__ unsigned long long g_cta_times[2 * 1024];   // debug (P2C_TILE_CLOCKS): globaltimer at entry / exit of every CTA
#ifdef P2C_CLOCKS_BUILD
__device__ __forceinline__ unsigned long long globaltimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#endif
// phase clocks are a debug build (P2C_BUILD_CLOCKS=1 python -c 'import __graft_entry__ as g; g.build()'): the checks
// cost 2 % of the step when compiled in
#ifndef P2C_CLOCKS_BUILD
#define CTA_CLOCK(k)
#define TILE_CLOCK(k)
#define WARP_CLOCK(k)
#define FOLD_CLOCK(k)
#define DEBUG_CLOCK(stmt)
#else
#define FOLD_CLOCK(k) \
    if ((threadIdx.x & 31) == 0) g_tile_clocks[k] = clock64();
#define WARP_CLOCK(k)                                                                                         \
    if (a.clocks && (int)blockIdx.x == a.clock_cta && (int)it == a.clocks - 1 && (threadIdx.x & 31) == 0)      \
        g_warp_clocks[(threadIdx.x >> 5) * 16 + k] = clock64();
#define DEBUG_CLOCK(stmt) stmt
#define CTA_CLOCK(k) \
    if (a.clocks && (int)blockIdx.x == a.clock_cta && threadIdx.x == 0) g_tile_clocks[k] = clock64();
#define TILE_CLOCK(k)                                                              \
    if (a.clocks && (int)blockIdx.x == a.clock_cta && (int)it == a.clocks - 1 && tid == 0) g_tile_clocks[k] = clock64();
#endif

// ---- loss totals (rarely executed, kept out of line: the double-precision divisions are long instruction sequences)
// draw a ticket with RELEASE semantics (this thread's partial store is ordered before it).  __threadfence() would
// also do, but it is a sequentially consistent fence followed by an L1 invalidate (MEMBAR.SC.GPU + CCTL.IVALL), which
// stalls the whole SM's memory pipe for ~2k cycles in the middle of the backward phases
__device__ __forceinline__ unsigned ticket_release(unsigned int* p) {
    unsigned t;
    asm volatile("atom.add.release.gpu.global.u32 %0, [%1], 1;" : "=r"(t) : "l"(p) : "memory");
    return t;
}
struct FoldArgs {   // (by value: a reference to the kernel parameters would force a local copy of all of them)
    float* losses;
    const float* partials;
    unsigned int* header;   // ticket, -, count2, arrive, next
    double count3, all2;
    float w2, w3, wr;
    int mask, rot;
    int count_known;   // >= 0: the masked-joint count is already in hand (the CTA read it when it polled the grid's count)
    unsigned next_seq; // != 0: written to header[5] (the launch sequence of the poll-fold scheme)
    XchgArgs xchg;
};
__device__ __noinline__ void finish_losses(const FoldArgs a, double d2, double d3, double dr) {   // one thread
    const double c2 = !a.mask ? a.all2
                      : a.count_known >= 0 ? (double)a.count_known
                                           : (double)__ldcg(reinterpret_cast<const int*>(a.header) + 2);
    const float l2 = (float)(d2 / c2), l3 = (float)(d3 / a.count3);
    float total = a.w2 * l2 + a.w3 * l3;
    if (a.rot) {   // rot_3d: 9 matrix entries per bone where loc_3d has 3 coordinates
        const double cr = a.count3 * 3.0;
        const float lr = (float)(dr / cr);
        total += a.wr * lr;
        a.losses[8] = lr; a.losses[9] = (float)dr; a.losses[10] = (float)cr; a.losses[11] = 0.f;
    }
    a.losses[0] = l2; a.losses[1] = l3; a.losses[2] = total;
    a.losses[3] = (float)d2; a.losses[4] = (float)c2; a.losses[5] = (float)d3; a.losses[6] = (float)a.count3;
    a.losses[7] = 0.f;
    exchange_stash(a.xchg, (float)d2, (float)c2, (float)d3, (float)a.count3);   // (sent by the next step's kernel)
    a.header[0] = 0u;   // self-resetting workspace header: ready for the next launch
    a.header[2] = 0u;
    a.header[3] = 0u;
    a.header[4] = 0u;
    if (a.next_seq) a.header[5] = a.next_seq;
}
// few tiles (a.warp_fold): the warp that stored this CTA's last partial waits for its ticket and, if it is the grid's
// last, folds right there, behind the backward phases of the other warps
__device__ __noinline__ void warp_fold_if_last(const FoldArgs a, int ntiles, unsigned ticket) {
    const int lane = threadIdx.x & 31;
    if (__shfl_sync(FULL_MASK, ticket, 0) != gridDim.x - 1) return;
    FOLD_CLOCK(51)
    // (acquire side of the ticket chain: the partials are read with ld.cg, past L1 -- no need for the sequentially
    // consistent fence + L1 invalidate of __threadfence())
    asm volatile("fence.acq_rel.gpu;" ::: "memory");
    double d2 = 0.0, d3 = 0.0, dr = 0.0;
    for (int i = lane; i < ntiles; i += 32) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(a.partials) + i);
        d2 += (double)v.x;
        d3 += (double)v.y;
        dr += (double)v.z;
    }
    FOLD_CLOCK(52)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        d2 += __shfl_xor_sync(FULL_MASK, d2, o);
        d3 += __shfl_xor_sync(FULL_MASK, d3, o);
        dr += __shfl_xor_sync(FULL_MASK, dr, o);
    }
    FOLD_CLOCK(53)
    if (lane == 0) finish_losses(a, d2, d3, dr);
    FOLD_CLOCK(54)
}

// BT / SL: bones and 32-pose slices per tile at compile time (0 = read from the arguments); WORLD: a per-frame world
// transform may be given (ZeroTrajectory, the only trajectory model of the reference, gives none); LONE: the launch
// has at most one tile per CTA (only selects which forward walk is compiled in: the straight-line walks are most of
// the code, and a kernel that runs every phase once per launch pays for its size in instruction fetches); ROT: rot_3d is
// fused (targets['absolute_pose_rot'] given) -- its own instances, so the plain step carries neither its code nor its
// registers (measured: +6 % on every size when the term was a run-time branch)
// threads per CTA: one bone per warp and slice when the launch has at most one tile per CTA (latency), else two
__host__ __device__ constexpr int tile_threads(int bones, int slices, bool lone) {
    return ((lone ? bones : bones / 2) * slices < 4 ? 4
            : (lone ? bones : bones / 2) * slices > MAX_THREADS / 32 ? MAX_THREADS / 32
                                                                      : (lone ? bones : bones / 2) * slices) * 32;
}

template <int BT, int SL, bool WORLD, bool LONE, bool ROT>
__global__ void __launch_bounds__(LAUNCH_THREADS, 1) tile_kernel(const TileArgs a) {
    extern __shared__ __align__(16) float smem[];
    // (CARLA instances: the CTA size is part of the instance -- tile_threads() -- so warp counts and loop strides fold)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // SERVICE (one tile per CTA, CARLA): one more warp than there are task warps.  It stores the tile's loss partial,
    // draws the CTA's ticket and, in the grid's last CTA, folds the partials -- a chain of L2 round trips (store
    // acknowledged -> ticket -> partials -> totals stored: ~5 k cycles) that sat in a task warp and made the folding CTA,
    // and with it the step, ~2 us longer than the others.  The service warp takes no part in the barriers after the loss
    // pass (the task warps meet at a named barrier of their own), so the chain runs beside the backward phases.
    constexpr bool SERVICE = BT == 26 && LONE && P2C_SERVICE_WARP;
    const int WT = BT == 26 ? tile_threads(BT, SL, LONE) : (int)blockDim.x;   // threads with bone / column tasks
    const int W = WT >> 5, NT = WT + (SERVICE ? 32 : 0);
    const int svc = SERVICE ? W : W - 2;   // the warp that stores the partial, draws the ticket and folds
    auto sync_tasks = [&]() {              // barrier of the task warps once the service warp has left
        if (SERVICE) named_barrier(7, W * 32);
        else __syncthreads();
    };
    const int B = BT ? BT : a.B, slices = SL ? SL : a.slices;
    const int B3 = B * 3, B6 = B * 6, B9 = B * 9, PS = 32 * slices, L = a.L;
    const TilePlan sm = tile_plan(PS, B);
    CTA_CLOCK(16)
    DEBUG_CLOCK(if (a.clocks && threadIdx.x == 0 && blockIdx.x < 1024) g_cta_times[blockIdx.x * 2] = globaltimer();)
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
    const int total = a.total;              // (the launcher keeps N*L*B*9 below 2^31 elements... offsets are size_t)
    const int ntiles = a.ntiles;

    auto load_d6 = [&](int tile) {
        const int m0 = tile * a.P;
        const uint32_t bytes = (uint32_t)(min(a.P, total - m0) * B6) * 4u;
        mbar_expect_tx(bars, bytes);
        bulk_g2s(D, a.d6 + (size_t)m0 * B6, bytes, bars);
    };
    auto load_targets = [&](int tile) {
        const int m0 = tile * a.P;
        const uint32_t bytes = (uint32_t)(min(a.P, total - m0) * B3) * 4u;
        mbar_expect_tx(bars + 1, 2u * bytes);
        bulk_g2s(G3, a.gt3 + (size_t)m0 * B3, bytes, bars + 1);
        bulk_g2s(G2, a.gt2 + (size_t)m0 * B3, bytes, bars + 1);
    };
    auto fill_rows = [&](int tile, int buf) {   // skeleton-table row of every pose (lanes past the tile repeat its last)
        const int m0 = tile * a.P;
        const int np = min(a.P, total - m0);
        for (int i = lane; i < PS; i += 32) {
            const int clip = (m0 + min(i, np - 1)) / L;
            ROW[buf * PS + i] = a.ref_index ? __ldg(a.ref_index + clip) : clip;
        }
    };

    if (tma) {
        mbar_init(bars, 1);
        mbar_init(bars + 1, 1);
        mbar_init(bars + 2, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // Set-up that touches no global memory comes BEFORE the wait for the previous kernel of the stream: the CTAs of a
    // chained step are resident (and idle in that wait) before the step in front of them has finished.
#if P2C_ENTRY_PREFETCH
    // ... and so do L2 PREFETCHES of the first tile's records: they return no data (L2 is the point of coherence: if
    // the previous kernel still writes these lines, the loads after the wait see its writes), they only start the trip
    // from HBM early, so that the TMA loads issued after the wait find the records in L2.
    if (tma && (int)blockIdx.x < ntiles) {
        const int m0 = blockIdx.x * a.P, np0 = min(a.P, total - m0);
        const uint32_t b6 = (uint32_t)(np0 * B6) * 4u, b3 = (uint32_t)(np0 * B3) * 4u;
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a.d6 + (size_t)m0 * B6), "r"(b6) : "memory");
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a.gt3 + (size_t)m0 * B3), "r"(b3) : "memory");
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a.gt2 + (size_t)m0 * B3), "r"(b3) : "memory");
        if (a.ref_index) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.ref_index + m0 / L) : "memory");   // (row table)
    }
#endif
    // suffix scan over the frames of a clip (phase E): bit k of scan_ok<slice> = "the lane 2^k to the right holds a later
    // frame of MY clip" -- a function of the lane alone, so the rounds of the scan test one bit instead of comparing
    // (kept in shared memory: two more live registers cost more in spills than a read per task)
    __shared__ unsigned scan_ok_s[64];
    if (warp == 0) {
        const int lane_clip = lane / L, lane_frame = lane - lane_clip * L;   // (one slice: whole clips side by side)
        unsigned ok0 = (unsigned)lane_clip << 8, ok1 = 0u;   // (bits 8..: the lane's clip within a one-slice tile)
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const int off = 1 << k;
            if (lane + off < 32) {
                if ((slices == 1 ? lane_frame : lane) + off < L) ok0 |= 1u << k;
                if (32 + lane + off < L) ok1 |= 1u << k;
            }
        }
        scan_ok_s[lane] = ok0;
        scan_ok_s[32 + lane] = ok1;
    }
    __shared__ signed char par_s[P2C_MAX_BONES];      // parents (a shared-memory read per task instead of arithmetic)
    if (tid < B) par_s[tid] = (signed char)(BT == 26 ? carla_parent(tid) : a.parent[tid]);
    __shared__ int cta_count;
    __shared__ int count_s;   // the grid's masked-joint count as this CTA read it (one tile per CTA)
    __shared__ float k2_shared;
    __shared__ int last_s;
    if (tid == 0) { cta_count = 0; last_s = 0; count_s = -1; }
    __syncthreads();   // (mbarriers initialised, tables and counters in place)
    if (a.count_inline || !a.mask) {
        // No count pre-pass in front of this launch (one tile per CTA, or the mask is off): the kernel is launched with
        // programmatic stream serialization, so the previous kernel of the stream may still be running (a one-tile-
        // per-CTA step triggers its dependents at its very top).  Everything in global memory -- the inputs, the
        // self-resetting workspace header -- may come from it: wait until it is complete and flushed.
        asm volatile("griddepcontrol.wait;" ::: "memory");
        CTA_CLOCK(23)
        // (one tile per CTA) let the next kernel's CTAs take the SMs as this grid's CTAs retire
        if (a.count_inline) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    }
    // POLLFOLD (one tile per CTA, service warp): no ticket.  Every CTA publishes its loss partial as self-validating
    // 8-byte (value, launch sequence) pairs; the service warp of CTA 0 polls the entries in tile order, folds them as
    // they turn valid and writes the totals.  The chain behind the grid's last partial is one store and one read
    // (~2 k cycles with the double-precision divisions) instead of store -> acknowledged -> ticket -> fence -> read
    // (7.5 k cycles measured: the folding CTA ended 1.4 us after the others).  A one-way dependency: producers never
    // wait for the folder, so nothing can hang whatever the placement of the CTAs.
    constexpr bool POLLFOLD = SERVICE && P2C_POLL_FOLD;
    unsigned seq_tag = 0u;
    if (POLLFOLD && warp == svc && lane == 0) {
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(seq_tag) : "l"(a.ticket + 5) : "memory");
        seq_tag += 1u;   // (never 0: a zeroed workspace holds no valid entry)
        if (seq_tag == 0u) seq_tag = 1u;
    }
    // data-parallel ranks: the sums the previous step left in this rank's exchange block go out to the peers now, so
    // that their NVLink round trip is over long before this kernel's own tail
    // (by the service warp where there is one: a task warp's next proxy fence would wait for the peer stores' NVLink
    // round trip, and with it CTA 0 at its next barrier -- 0.5 us per step at N > 1)
    if (blockIdx.x == 0 && tid == svc * 32) exchange_send_pending(a.xchg);
    if (tma) {
        if ((int)blockIdx.x < ntiles) {
            load_d6(blockIdx.x);
            load_targets(blockIdx.x);
        }
    }
    // (the first tile's row table: visible to the other warps after barrier #1)
    if (warp == 0 && (int)blockIdx.x < ntiles) fill_rows(blockIdx.x, 0);

    float k2 = a.k2_all;   // (mask on: replaced below once the count is known)
    // ---- loss sums: one partial pair per TILE (fixed order inside the tile), so the totals do not depend on which
    //      CTA ran which tile; the last CTA to take a ticket folds them in tile order at the end of the kernel
    __shared__ int next_s[2];
    DEBUG_CLOCK(long long svc_t0 = 0;)
    unsigned my_ticket = 0xffffffffu;   // (drawn by one lane after this CTA's last partial store, looked at later)
    auto fold_args = [&]() {
        FoldArgs f;
        f.losses = a.losses; f.partials = a.partials; f.header = a.ticket;
        f.count3 = a.count3; f.all2 = (double)total * B * 2; f.w2 = a.w2; f.w3 = a.w3; f.mask = a.mask;
        f.wr = a.wr; f.rot = a.gt_rot != nullptr;
        f.count_known = -1; f.next_seq = 0u;
        f.xchg = a.xchg;
        return f;
    };
    auto finish = [&](double d2, double d3, double dr) { finish_losses(fold_args(), d2, d3, dr); };
    auto fold_if_last = [&]() { warp_fold_if_last(fold_args(), ntiles, my_ticket); };
    auto take_ticket = [&]() { my_ticket = ticket_release(a.ticket); };
    constexpr bool lone_sums = BT == 26 && LONE && SL == 1, lone_sums_early = lone_sums;
    // (even, odd) bone pairs per warp in the element-wise passes where the warps have two or more tasks each: the record
    // accessors then know the parity of their bone at compile time (tile_common.cuh)
    constexpr int PAIRS = BT == 26 ? (LONE ? P2C_LONE_TASKS : 1) : 0;
    const float k3 = a.k3;
    const float kr = a.kr;   // rot_3d: mean over 9 entries per bone
    __shared__ float red_r[32];
    unsigned it = 0;
    CTA_CLOCK(17)
    const int clips_full = a.clips_full;
    const bool whole_clips = a.whole_clips != 0;   // pose_changes mode; relative_rot tiles are 32 free poses

    // tiles beyond the first come from a grid-wide counter (SMs do not run at one speed: up to a quarter between the
    // fastest and the slowest CTA under a static split); the ticket is drawn at the top of a tile and first needed
    // after the forward phases
    int nxt = ntiles;
    for (int tile = blockIdx.x; tile < ntiles; tile = nxt, ++it) {
        const int m0 = tile * a.P;
        unsigned drawn = 0;   // (the next tile of this CTA: drawn after the forward walk, see barrier #4)
        if (ROT && tma) {   // the tile's rot_3d target records on their way into L2 while the forward phases run
            const uint32_t bytes = (uint32_t)(min(a.P, total - m0) * B9) * 4u;
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a.gt_rot + (size_t)m0 * B9), "r"(bytes) : "memory");
        }
        float s2 = 0.f, s3 = 0.f, sr = 0.f;
        const int np = min(a.P, total - m0);
        const int* row = ROW + (it & 1) * PS;
        const int tile_clips = np == a.P ? clips_full : np / L;   // (only the last tile can be short)
        TILE_CLOCK(0)
        while (!mbar_try_wait(bars, it & 1u)) {}
        TILE_CLOCK(1)
        WARP_CLOCK(1)
        __syncthreads();   // #1 (first tile: the row table; later tiles: ordered by the barriers of the previous tile)

        TILE_CLOCK(2)
        // the reference rotations that start the scan of this thread's first (clip, bone, column) item: asked for now,
        // used after phase A (an L2 round trip on the SMs far from the table's slice)
        float q0 = 0.f, q1 = 0.f, q2 = 0.f;
        const bool split_scan = slices == 2 && np > 32;   // a clip over two slices: its halves are scanned side by side
        if (a.scan && tid < tile_clips * B3) {   // (first-half items only when the scan is split: tile_clips == 1)
            const int cl = tid / B3, rem = tid - cl * B3;
            const float* rr = a.ref_rot + (size_t)row[cl * L] * B9 + (rem / 3) * 9 + rem % 3;
            q0 = __ldg(rr); q1 = __ldg(rr + 3); q2 = __ldg(rr + 6);
        }
        // ---- phase A: 6-D -> C[t] (rotation_6d_to_matrix), one bone per task
        {
            float* dst = a.scan ? A : R;
            bone_tasks<PAIRS>(warp, W, slices * B, [&](int task, auto par) {
                constexpr int PAR = decltype(par)::value;
                const int slice = task >= B ? 1 : 0, b = task - slice * B;
                const int col = slice * 32 + lane, pl = min(col, np - 1);
                float d[6];
                lds6<PAR>(D + pl * B6, b, d);
                const GS r = gram_schmidt(d);
                M3 C;
                C.m[0] = r.b1.x; C.m[1] = r.b1.y; C.m[2] = r.b1.z;
                C.m[3] = r.b2.x; C.m[4] = r.b2.y; C.m[5] = r.b2.z;
                C.m[6] = r.b3.x; C.m[7] = r.b3.y; C.m[8] = r.b3.z;
                if (col < np) sts9<PAR>(dst + pl * B9, b, C);
            });
        }
        if (!a.scan) fence_async_smem();
        WARP_CLOCK(2)
        __syncthreads();   // #2
        TILE_CLOCK(3)
        // (Measured alternatives for the one-tile kernels, all slower than the column mapping below -- a warp issues an
        // FFMA every 2 cycles at best, so what counts is the instruction count on the busiest scheduler, not the length
        // of the dependency chain: a thread per ELEMENT (three neighbouring lanes per column, 3 reads + 3 shuffles + 3
        // FMAs + 1 store per frame, 16 warps): 3.1 k cycles against 2.8 k; a LANE PER BONE forming the full 3x3 product
        // (two warps, every word read once): 3.75 k.)
        // ---- phase B: R[t] = C[t] . R[t-1] (p3d_pose.py:98-114 looped by layers/projection.py:190-193); one thread
        //      per (clip, bone, column of R), serial over the frames of the clip
        if (a.scan) {
            // A clip over two slices (33..64 frames; the tile is that ONE clip): the two halves are scanned side by side --
            // frames 0..31 from the reference rotation, frames 32.. from the identity (a column of the prefix product only
            // needs the same column of its predecessor) -- and the second half is multiplied by R[31] afterwards, one
            // bone per warp task: 32 dependent steps instead of 64.  (FOUR chains of 16 frames measured slower, 357.9
            // against 352.5 us at 4096 x 64: ten scan warps and a fix-up pass with two more products per bone.)
            const int items = (split_scan ? 2 : tile_clips) * B3;
#pragma unroll 1
            for (int idx = tid; idx < items; idx += NT) {
                const int cl = idx / B3, rem = idx - cl * B3, b = rem / 3, j = rem - b * 3;
                const bool second = split_scan && cl == 1;
                const int f0 = split_scan ? cl * 32 : cl * L;            // first frame (lane column) of this item
                const int nf = split_scan ? (cl ? np - 32 : 32) : L;     // frames it runs over
                float p0 = q0, p1 = q1, p2 = q2;
                if (second) {
                    p0 = j == 0 ? 1.f : 0.f; p1 = j == 1 ? 1.f : 0.f; p2 = j == 2 ? 1.f : 0.f;
                } else if (idx != tid) {
                    const float* rr = a.ref_rot + ((size_t)row[f0] * B + b) * 9 + j;
                    p0 = __ldg(rr); p1 = __ldg(rr + 3); p2 = __ldg(rr + 6);
                }
                const float* src = A + f0 * B9 + b * 9;
                float* dst = R + f0 * B9 + b * 9 + j;
                // two frames per trip; the records of the next two frames are in flight while these products run
                float c[9], e[9];
#pragma unroll
                for (int i = 0; i < 9; ++i) c[i] = src[i];
                int f = 0;
#pragma unroll 1
                for (; f + 2 <= nf; f += 2) {
#pragma unroll
                    for (int i = 0; i < 9; ++i) e[i] = src[B9 + i];
                    const float n0 = fmaf(c[2], p2, fmaf(c[1], p1, c[0] * p0));
                    const float n1 = fmaf(c[5], p2, fmaf(c[4], p1, c[3] * p0));
                    const float n2 = fmaf(c[8], p2, fmaf(c[7], p1, c[6] * p0));
                    dst[0] = n0; dst[3] = n1; dst[6] = n2;
                    src += 2 * B9;
                    if (f + 2 < nf) {
#pragma unroll
                        for (int i = 0; i < 9; ++i) c[i] = src[i];
                    }
                    p0 = fmaf(e[2], n2, fmaf(e[1], n1, e[0] * n0));
                    p1 = fmaf(e[5], n2, fmaf(e[4], n1, e[3] * n0));
                    p2 = fmaf(e[8], n2, fmaf(e[7], n1, e[6] * n0));
                    dst[B9] = p0; dst[B9 + 3] = p1; dst[B9 + 6] = p2;
                    dst += 2 * B9;
                    DEBUG_CLOCK(if (a.clocks && (int)blockIdx.x == a.clock_cta && (int)it == a.clocks - 1 && tid == 0 && f < 32) g_tile_clocks[40 + (f >> 1)] = clock64();)
                }
                if (f < nf) {   // odd frame count: the last frame
                    dst[0] = fmaf(c[2], p2, fmaf(c[1], p1, c[0] * p0));
                    dst[3] = fmaf(c[5], p2, fmaf(c[4], p1, c[3] * p0));
                    dst[6] = fmaf(c[8], p2, fmaf(c[7], p1, c[6] * p0));
                }
            }
        }
        // bone offsets of the skeleton of every clip the tile touches, one 3B-word record per clip, in the upper half of
        // the scratch region: the forward walk reads them from there (lanes of one clip read one address: a broadcast)
        // instead of going to the table in global memory
        // (tiles of whole clips start at a clip boundary: no division by the run-time clip length per tile)
        int clip0, last_clip;
        if (whole_clips) {
            clip0 = tile * clips_full;
            last_clip = np == a.P ? clips_full - 1 : (np - 1) / L;
        } else {
            clip0 = m0 / L;
            last_clip = (m0 + np - 1) / L - clip0;
        }
        const int nclip = last_clip + 1;
        // clip (counted from the tile's first) of the pose in column col = slot pl (columns past the tile repeat its last)
        auto clip_of = [&](int col, int pl) {
            (void)col;
            if (!whole_clips) return (m0 + pl) / L - clip0;
            return slices == 1 ? min((int)(scan_ok_s[lane] >> 8), last_clip) : 0;
        };
        {
            const int busy = (a.scan && tile_clips * B3 <= 160 && NT >= 256) ? 160 : 0;   // threads that ran the scan
            for (int i = tid - busy; i >= 0 && i < nclip * B3; i += NT - busy) {
                const int cl = i / B3, k = i - cl * B3;
                GV[i] = __ldg(a.ref_loc + (size_t)row[max((clip0 + cl) * L - m0, 0)] * B3 + k);
            }
            if (a.mask && a.count_inline) {
                // utils/tensors.py:29-40 (one tile per CTA): the threads without a scan item count the kept (joint,
                // channel) pairs of this CTA's tile; the grid total goes through one 64-bit atomic per CTA (count in
                // the low word, arrivals in the high word; integer => order independent), sent now and read back
                // during the forward walk
                const int nc = NT - busy, ci = tid - busy;
                if (ci >= 0) {
                    while (!mbar_try_wait(bars + 1, it & 1u)) {}
                    int c = 0;
                    for (int i = ci; i < np * B; i += nc) {
                        const int j = i % B;
                        const float* g = G2 + i * 3;
                        if (j == a.hips || (g[0] != 0.f && g[1] != 0.f)) c += 2;
                    }
                    c = __reduce_add_sync(FULL_MASK, c);
                    if (lane == 0 && c) atomicAdd(&cta_count, c);
                    named_barrier(1, nc);
                    if (ci == 0)
                        atomicAdd(reinterpret_cast<unsigned long long*>(a.count2),
                                  (1ull << 32) | (unsigned long long)(unsigned)cta_count);
                }
            }
        }
        if (a.scan && split_scan) {
            __syncthreads();   // #2b: both halves of the clip are in R
            // R[t] = (C[t] .. C[32]) . R[31] for the second half, one bone per task
            bone_tasks<(PAIRS == 1 && SL != 1) ? P2C_LONE_TASKS : PAIRS>(warp, W, B, [&](int b, auto par) {
                constexpr int PAR = decltype(par)::value;
                const int col = 32 + lane;
                if (col < np) {
                    const M3 Q = lds9<PAR>(R + col * B9, b), P31 = lds9<PAR>(R + 31 * B9, b);
                    sts9<PAR>(R + col * B9, b, m3_mul(Q, P31));
                }
            });
        }
        fence_async_smem();
        WARP_CLOCK(3)
        __syncthreads();   // #3
        TILE_CLOCK(4)
        if (tma && a.rel_rot) {
            bulk_s2g(a.rel_rot + (size_t)m0 * B9, R, (uint32_t)(np * B9) * 4u);
            bulk_commit();
        }

        // ---- phase C: FK (p3d_pose.py:116-184): A_b = R_b . A_parent, t_b = loc_b . A_parent + t_parent.
        //      The phase is bound by shared-memory wavefronts, so every record is read once: with the CARLA tree (known at
        //      compile time: straight-line code, immediate offsets, 64-bit conflict-free record accesses) ONE warp per
        //      slice and limb group forms the full 3x3 products; the trunk-and-left-arm group signals (bar.arrive, it
        //      never waits) when hips and spine01 are done and the groups that hang off them wait on those named
        //      barriers.  Any other tree, and one tile per CTA (latency counts, not wavefronts: three short instruction
        //      streams beat one long one): three warps per slice, one per column of A (a column needs only the same
        //      column of the parent), each walking the whole tree in bone order.
        constexpr bool limbs = BT == 26 && !LONE;
        constexpr bool lone_limbs = BT == 26 && LONE && SL == 1;   // limb groups of three column warps each
        // (one tile per CTA) hips and neck come early in the walk (trunk, first bone of the head group): a warp without
        // a column forms the normaliser's shift and scale from them WHILE the limbs are still walked -- one phase and
        // one barrier (#4b) less on the CTA's critical path
        const bool early_norm = P2C_EARLY_NORM && (lone_limbs || (limbs && P2C_EARLY_NORM_MANY)) && a.normalize &&
                                a.neck == 8 && a.hips <= 3;
        const int fk_warps = limbs ? 5 * slices : lone_limbs ? 15 : 3 * slices;
        if (limbs && warp < fk_warps) {
            const int grp = warp / slices, col = (warp - grp * slices) * 32 + lane, pl = min(col, np - 1);
            const bool valid = col < np;
            const float* rb = R + pl * B9;
            float* ab = A + pl * B9;
            const float* lc = GV + clip_of(col, pl) * B3;   // this pose's skeleton offsets
            float P[9], tp[3];   // A and t of the previous bone of the run
            auto rec_ld = [&](const float* base, int b, float* x) {   // 9-word record b of an even-word pose record
                const float* q = base + 9 * b;
                if (b & 1) {
                    x[0] = q[0];
                    const float2 v0 = lds64(q + 1), v1 = lds64(q + 3), v2 = lds64(q + 5), v3 = lds64(q + 7);
                    x[1] = v0.x; x[2] = v0.y; x[3] = v1.x; x[4] = v1.y; x[5] = v2.x; x[6] = v2.y; x[7] = v3.x; x[8] = v3.y;
                } else {
                    const float2 v0 = lds64(q), v1 = lds64(q + 2), v2 = lds64(q + 4), v3 = lds64(q + 6);
                    x[0] = v0.x; x[1] = v0.y; x[2] = v1.x; x[3] = v1.y; x[4] = v2.x; x[5] = v2.y; x[6] = v3.x; x[7] = v3.y;
                    x[8] = q[8];
                }
            };
            auto rec_st = [&](float* base, int b, const float* x) {
                float* q = base + 9 * b;
                if (b & 1) {
                    q[0] = x[0];
                    sts64(q + 1, x[1], x[2]); sts64(q + 3, x[3], x[4]); sts64(q + 5, x[5], x[6]); sts64(q + 7, x[7], x[8]);
                } else {
                    sts64(q, x[0], x[1]); sts64(q + 2, x[2], x[3]); sts64(q + 4, x[4], x[5]); sts64(q + 6, x[6], x[7]);
                    q[8] = x[8];
                }
            };
            auto bone = [&](int b, int par, bool head) {
                if (head || par != b - 1) {
                    if (par >= 0) {
                        rec_ld(ab, par, P);
                        tp[0] = TS[(par * 3) * PS + col]; tp[1] = TS[(par * 3 + 1) * PS + col];
                        tp[2] = TS[(par * 3 + 2) * PS + col];
                    } else {   // root: the parent transform is the identity
#pragma unroll
                        for (int i = 0; i < 9; ++i) P[i] = (i == 0 || i == 4 || i == 8) ? 1.f : 0.f;
                        tp[0] = tp[1] = tp[2] = 0.f;
                    }
                }
                float x[9], n[9];
                rec_ld(rb, b, x);
                const float l0 = lc[3 * b], l1 = lc[3 * b + 1], l2 = lc[3 * b + 2];
#pragma unroll
                for (int i = 0; i < 3; ++i)
#pragma unroll
                    for (int j = 0; j < 3; ++j)
                        n[3 * i + j] = fmaf(x[3 * i + 2], P[6 + j], fmaf(x[3 * i + 1], P[3 + j], x[3 * i] * P[j]));
#pragma unroll
                for (int j = 0; j < 3; ++j) tp[j] = fmaf(l2, P[6 + j], fmaf(l1, P[3 + j], l0 * P[j])) + tp[j];
#pragma unroll
                for (int i = 0; i < 9; ++i) P[i] = n[i];
                if (valid) rec_st(ab, b, n);
                TS[(3 * b) * PS + col] = tp[0]; TS[(3 * b + 1) * PS + col] = tp[1]; TS[(3 * b + 2) * PS + col] = tp[2];
            };
#define P2C_RUN(B0, B1)                                                                   \
    _Pragma("unroll") for (int b = B0; b < B1; ++b) {                                      \
        bone(b, carla_parent(b), b == B0);                                                 \
        DEBUG_CLOCK(if (B0 < 8 && a.clocks && (int)blockIdx.x == a.clock_cta && (int)it == a.clocks - 1 && tid == 0) g_tile_clocks[32 + b] = clock64();) \
    }
            if (grp == 0) {
                P2C_RUN(0, 2)
                named_arrive(3, 3 * slices * 32);   // hips
                P2C_RUN(2, 4)
                named_arrive(4, 3 * slices * 32);   // spine01
                P2C_RUN(4, 8)               // left arm
            } else if (grp == 1) {
                named_barrier(4, 3 * slices * 32);
                P2C_RUN(8, 9)               // neck
                if (early_norm) named_arrive(8 + (warp - grp * slices), 2 * 32);   // (per slice: this warp + the spare one)
                P2C_RUN(9, 12)              // head, eyes
            } else if (grp == 2) {
                named_barrier(4, 3 * slices * 32);
                P2C_RUN(12, 16)             // right arm
            } else if (grp == 3) {
                named_barrier(3, 3 * slices * 32);
                P2C_RUN(16, 21)             // right leg
            } else {
                named_barrier(3, 3 * slices * 32);
                P2C_RUN(21, 26)             // left leg
            }
#undef P2C_RUN
            fence_async_smem();
            DEBUG_CLOCK(if (a.clocks && (int)blockIdx.x == a.clock_cta && tid == 0) g_tile_clocks[21] = clock64();)
        } else if (!limbs && warp < fk_warps) {
            const int cgrp = lone_limbs ? warp / 3 : 0, cw = lone_limbs ? warp - 3 * cgrp : warp;
            const int jc = cw % 3, col = (cw / 3) * 32 + lane, pl = min(col, np - 1);
            const bool valid = col < np;
            const float* rb = R + pl * B9;
            float* ab = A + pl * B9 + jc;
            const float* lc = GV + clip_of(col, pl) * B3;   // this pose's skeleton offsets
            float* tr = TS + jc * PS + col;
            float c0 = 0.f, c1 = 0.f, c2 = 0.f, ct = 0.f;   // column jc of A_{b-1}, component jc of t_{b-1}
            // two bones per trip (B is even); the inputs of the next bone (its rel_rot record, its offset, its parent)
            // are fetched into the other register set while this one is computed
            float u[12], w[12];
            int pu, pw;
            auto fetch = [&](float* x, int b) {
                if (BT == 26) {   // (b is a compile-time constant in the unrolled walks: no parity branch)
                    const M3 r = lds9(rb, b);
#pragma unroll
                    for (int i = 0; i < 9; ++i) x[i] = r.m[i];
                } else {
#pragma unroll
                    for (int i = 0; i < 9; ++i) x[i] = rb[9 * b + i];
                }
                x[9] = lc[3 * b]; x[10] = lc[3 * b + 1]; x[11] = lc[3 * b + 2];
            };
            auto bone = [&](const float* x, int b, int par) {
                float p0 = c0, p1 = c1, p2 = c2, tp = ct;
                if (par != b - 1 || par < 0) {
                    if (par >= 0) {
                        const float* ap = A + pl * B9 + par * 9 + jc;
                        p0 = ap[0]; p1 = ap[3]; p2 = ap[6];
                        tp = TS[(par * 3 + jc) * PS + col];
                    } else {   // root: the parent transform is the identity
                        p0 = jc == 0 ? 1.f : 0.f; p1 = jc == 1 ? 1.f : 0.f; p2 = jc == 2 ? 1.f : 0.f;
                        tp = 0.f;
                    }
                }
                c0 = fmaf(x[2], p2, fmaf(x[1], p1, x[0] * p0));
                c1 = fmaf(x[5], p2, fmaf(x[4], p1, x[3] * p0));
                c2 = fmaf(x[8], p2, fmaf(x[7], p1, x[6] * p0));
                ct = fmaf(x[11], p2, fmaf(x[10], p1, x[9] * p0)) + tp;
                if (valid) { ab[9 * b] = c0; ab[9 * b + 3] = c1; ab[9 * b + 6] = c2; }
                tr[(3 * b) * PS] = ct;
            };
            if (BT == 26) {   // (one tile per CTA: latency matters, three short instruction streams beat one long one)
#define P2C_RUNC(B0, B1)                                  \
    _Pragma("unroll") for (int b = B0; b < B1; ++b) {     \
        fetch(u, b);                                      \
        bone(u, b, carla_parent(b));                      \
        DEBUG_CLOCK(if (B0 < 8 && a.clocks && (int)blockIdx.x == a.clock_cta && (int)it == a.clocks - 1 && tid == 0) g_tile_clocks[32 + b] = clock64();) \
    }
                if (!lone_limbs) {
                    P2C_RUNC(0, 26)
                } else if (cgrp == 0) {   // the same limb groups as above, three column warps each
                    P2C_RUNC(0, 2)
                    named_arrive(3, 9 * 32);   // hips
                    P2C_RUNC(2, 4)
                    named_arrive(4, 9 * 32);   // spine01
                    P2C_RUNC(4, 8)
                } else if (cgrp == 1) {
                    named_barrier(4, 9 * 32);
                    P2C_RUNC(8, 9)
                    if (early_norm) named_arrive(8, 4 * 32);   // neck (and, through barrier 4, hips) are in place
                    P2C_RUNC(9, 12)
                } else if (cgrp == 2) {
                    named_barrier(4, 9 * 32);
                    P2C_RUNC(12, 16)
                } else if (cgrp == 3) {
                    named_barrier(3, 9 * 32);
                    P2C_RUNC(16, 21)
                } else {
                    named_barrier(3, 9 * 32);
                    P2C_RUNC(21, 26)
                }
#undef P2C_RUNC
            } else {
                fetch(u, 0);
                pu = a.parent[0];
    #pragma unroll 1
                for (int b = 0; b < B; b += 2) {
                    fetch(w, b + 1);
                    pw = a.parent[b + 1];
                    bone(u, b, pu);
                    if (b + 2 < B) {
                        fetch(u, b + 2);
                        pu = a.parent[b + 2];
                    }
                    bone(w, b + 1, pw);
                }
            }
            fence_async_smem();
        } else if (early_norm && (lone_limbs ? warp == fk_warps + 1 : warp < fk_warps + slices)) {
            // (many tiles: warps fk_warps .. fk_warps + slices - 1, one per slice; one tile per CTA: the warp behind the
            // lane that polls the grid's count)
            const int sl = lone_limbs ? 0 : warp - fk_warps;
            if (lone_limbs) named_barrier(8, 4 * 32);
            else named_barrier(8 + sl, 2 * 32);
            const int col = sl * 32 + lane, m = m0 + min(col, np - 1);
            const V3 th{TS[(a.hips * 3) * PS + col], TS[(a.hips * 3 + 1) * PS + col], TS[(a.hips * 3 + 2) * PS + col]};
            const V3 tn{TS[(a.neck * 3) * PS + col], TS[(a.neck * 3 + 1) * PS + col], TS[(a.neck * 3 + 2) * PS + col]};
            const Proj ph = project<WORLD>(a, th, m), pn = project<WORLD>(a, tn, m);
            const float dx = pn.p.x - ph.p.x, dy = pn.p.y - ph.p.y;
            const float scale = fsqrt(fmaf(dx, dx, dy * dy));
            NRM[col] = ph.p.x;
            NRM[PS + col] = ph.p.y;
            NRM[2 * PS + col] = dx;
            NRM[3 * PS + col] = dy;
            NRM[4 * PS + col] = frcp(scale);
            NRM[5 * PS + col] = (scale > 0.f && scale <= 3.402823466e38f) ? 1.f : 0.f;
        } else if (a.mask && a.count_inline && tid == fk_warps * 32) {
            // one lane of a warp without a column polls until every CTA of the grid has sent its count (all CTAs are
            // resident: one tile each, at most one CTA per SM)
            unsigned long long v;
            do {
                v = *reinterpret_cast<volatile unsigned long long*>(a.count2);
            } while ((unsigned)(v >> 32) < gridDim.x);
            k2_shared = a.w2 * 2.0f / (float)(unsigned)(v & 0xffffffffull);
            count_s = (int)(unsigned)(v & 0xffffffffull);
            DEBUG_CLOCK(if (a.clocks && (int)blockIdx.x == a.clock_cta) g_tile_clocks[22] = clock64();)
        }
        WARP_CLOCK(4)
        __syncthreads();   // #4
        // The CTA's next tile comes from the grid-wide counter as LATE as its first use allows (after the loss pass:
        // barrier #5): a tile that has been drawn is out of every other CTA's reach until its owner gets to it, and
        // the slowest owners set the length of the kernel's tail.
        if (tma && a.dynamic) drawn = atomicAdd(a.next, 1u);
        TILE_CLOCK(5)
        if (tma && a.abs_rot) {
            bulk_s2g(a.abs_rot + (size_t)m0 * B9, A, (uint32_t)(np * B9) * 4u);
            bulk_commit();
        }
        // hips/neck shift and scale of every pose (extractor.py:23-36, hips_neck_extractor.py:6-13)
        if (a.normalize && !early_norm && warp < slices) {
            const int col = warp * 32 + lane;
            const int m = m0 + min(col, np - 1);
            const V3 th{TS[(a.hips * 3) * PS + col], TS[(a.hips * 3 + 1) * PS + col], TS[(a.hips * 3 + 2) * PS + col]};
            const V3 tn{TS[(a.neck * 3) * PS + col], TS[(a.neck * 3 + 1) * PS + col], TS[(a.neck * 3 + 2) * PS + col]};
            const Proj ph = project<WORLD>(a, th, m), pn = project<WORLD>(a, tn, m);
            const float dx = pn.p.x - ph.p.x, dy = pn.p.y - ph.p.y;
            const float scale = fsqrt(fmaf(dx, dx, dy * dy));
            NRM[col] = ph.p.x;
            NRM[PS + col] = ph.p.y;
            NRM[2 * PS + col] = dx;
            NRM[3 * PS + col] = dy;
            NRM[4 * PS + col] = frcp(scale);
            NRM[5 * PS + col] = (scale > 0.f && scale <= 3.402823466e38f) ? 1.f : 0.f;
        }
        while (!mbar_try_wait(bars + 1, it & 1u)) {}
        WARP_CLOCK(5)
        if (!early_norm) __syncthreads();   // #4b (shift / scale in place; early_norm: they were before barrier #4)
        if (a.mask && a.count_inline) k2 = k2_shared;
        if (a.mask && !a.count_inline && it == 0) {
            asm volatile("griddepcontrol.wait;" ::: "memory");   // the count pre-pass (launched just before) is complete
            k2 = a.w2 * 2.0f / (float)__ldcg(a.count2);
        }
        TILE_CLOCK(6)

        // ---- projection, normaliser, loss terms, dL/d abs_loc; one bone per task; abs_loc and projection records are
        //      written over the target records they were compared with
        {
            bone_tasks<PAIRS>(warp, W, slices * B, [&](int task, auto par) {
                constexpr int PAR = decltype(par)::value;
                const int slice = task >= B ? 1 : 0, b = task - slice * B;
                const int col = slice * 32 + lane, pl = min(col, np - 1);
                const bool valid = col < np;
                const int m = m0 + pl;
                LossCtx cx;
                cx.k2 = k2; cx.k3 = k3; cx.s2 = s2; cx.s3 = s3;
                cx.sgx = cx.sgy = cx.sgn = 0.f;
                cx.sx = cx.sy = 0.f; cx.inv = 1.f; cx.okn = true;
                if (a.normalize) {
                    cx.sx = NRM[col]; cx.sy = NRM[PS + col]; cx.inv = NRM[4 * PS + col]; cx.okn = NRM[5 * PS + col] != 0.f;
                }
                const V3 g3 = lds3(G3 + pl * B3, b), g2 = lds3(G2 + pl * B3, b);
                const V3 t{TS[(b * 3) * PS + col], TS[(b * 3 + 1) * PS + col], TS[(b * 3 + 2) * PS + col]};
                V3 p, n;
                const V3 g = bone_loss<WORLD>(a, b, t, g2.x, g2.y, g3, m, valid, cx, p, n);
                s2 = cx.s2; s3 = cx.s3;
                if (valid) {
                    sts3<PAR>(G3 + pl * B3, b, t);
                    sts3<PAR>(G2 + pl * B3, b, p);
                    if (a.proj_norm && a.normalize) {   // off the benchmark path: straight to global
                        float* o = a.proj_norm + ((size_t)m * B + b) * 3;
                        o[0] = n.x; o[1] = n.y; o[2] = n.z;
                    }
                }
                GV[(b * 3) * PS + col] = g.x; GV[(b * 3 + 1) * PS + col] = g.y; GV[(b * 3 + 2) * PS + col] = g.z;
                if (a.normalize) {
                    PART[(b * 3) * PS + col] = cx.sgx;
                    PART[(b * 3 + 1) * PS + col] = cx.sgy;
                    PART[(b * 3 + 2) * PS + col] = cx.sgn;
                }
            });
        }
        s2 = warp_sum(s2);
        s3 = warp_sum(s3);
        if (lane == 0) { red[warp * 2] = s2; red[warp * 2 + 1] = s3; }
        fence_async_smem();
        if (tma) {
            bulk_wait_read_all();   // rel_rot / abs_rot stores have read R and A
            next_s[it & 1] = a.dynamic ? (int)min(drawn + gridDim.x, (unsigned)ntiles) : ntiles;
        }
        WARP_CLOCK(6)
        __syncthreads();   // #5
        nxt = next_s[it & 1];
        TILE_CLOCK(7)
        auto store_partial = [&]() {
            // the tile's loss partial (a fixed shuffle tree over the per-warp sums: W <= 32); on this CTA's last tile
            // also its ticket, drawn now and looked at after the next phases (an L2 round trip later)
            float x = lane < W ? red[lane * 2] : 0.f, y = lane < W ? red[lane * 2 + 1] : 0.f;
            float z = (ROT && lane < W) ? red_r[lane] : 0.f;
            x = warp_sum(x);
            y = warp_sum(y);
            if (ROT) z = warp_sum(z);
            if (lane == 0) {
                reinterpret_cast<float4*>(a.partials)[tile] = make_float4(x, y, z, 0.f);
                if (nxt >= ntiles) take_ticket();
            }
        };
        if (SERVICE && warp == svc) {
            // (rot_3d fused: its sum is formed in the next pass -- the task warps arrive at barrier 6 when it is there)
            if (ROT) named_barrier(6, (W + 1) * 32);
            DEBUG_CLOCK(if (lane == 0) svc_t0 = clock64();)
            if (POLLFOLD) {
                float x = lane < W ? red[lane * 2] : 0.f, y = lane < W ? red[lane * 2 + 1] : 0.f;
                float z = (ROT && lane < W) ? red_r[lane] : 0.f;
                x = warp_sum(x);
                y = warp_sum(y);
                if (ROT) z = warp_sum(z);
                const unsigned tag = __shfl_sync(FULL_MASK, seq_tag, 0);
                unsigned long long* ent = reinterpret_cast<unsigned long long*>(a.partials);   // 4 x 8 bytes per tile
                if (lane == 0) {
                    auto put = [&](int k, float v) {
                        const unsigned long long w = ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v);
                        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(ent + 4 * (size_t)tile + k), "l"(w) : "memory");
                    };
                    put(0, x);
                    put(1, y);
                    if (ROT) put(2, z);
                }
                if (blockIdx.x == 0) {
                    // fold in tile order: lane l takes tiles l, l + 32, ... (the order of warp_fold_if_last: same totals)
                    double d2 = 0.0, d3 = 0.0, dr = 0.0;
                    for (int i = lane; i < ntiles; i += 32) {
                        unsigned long long wx, wy, wz = (unsigned long long)tag << 32;
                        do {
                            asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(wx) : "l"(ent + 4 * (size_t)i) : "memory");
                            asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(wy) : "l"(ent + 4 * (size_t)i + 1) : "memory");
                            if (ROT)
                                asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(wz) : "l"(ent + 4 * (size_t)i + 2) : "memory");
                        } while ((unsigned)(wx >> 32) != tag || (unsigned)(wy >> 32) != tag || (unsigned)(wz >> 32) != tag);
                        d2 += (double)__uint_as_float((unsigned)wx);
                        d3 += (double)__uint_as_float((unsigned)wy);
                        dr += (double)__uint_as_float((unsigned)wz);
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        d2 += __shfl_xor_sync(FULL_MASK, d2, o);
                        d3 += __shfl_xor_sync(FULL_MASK, d3, o);
                        dr += __shfl_xor_sync(FULL_MASK, dr, o);
                    }
                    // the entries are spent: no later launch may take them for its own, whatever the caller does to
                    // the header in between (a header zeroed again restarts the sequence)
                    for (int i = lane; i < ntiles; i += 32) {
#pragma unroll
                        for (int k = 0; k < (ROT ? 3 : 2); ++k)
                            asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(ent + 4 * (size_t)i + k), "l"(0ull) : "memory");
                    }
                    if (lane == 0) {
                        FoldArgs f = fold_args();
                        f.count_known = a.mask ? count_s : -1;
                        f.next_seq = tag;
                        finish_losses(f, d2, d3, dr);
                        DEBUG_CLOCK(g_tile_clocks[50] = svc_t0; g_tile_clocks[54] = clock64(); g_tile_clocks[51] = g_tile_clocks[52] = g_tile_clocks[53] = svc_t0;)
                    }
                }
                break;
            }
            store_partial();
            DEBUG_CLOCK(if (lane == 0 && my_ticket == gridDim.x - 1) { g_tile_clocks[50] = svc_t0; g_tile_clocks[55] = clock64(); })
            if (a.warp_fold) fold_if_last();
            break;   // one tile per CTA: nothing else for this warp (the kernel's tail waits for it)
        }
        if (!SERVICE && !ROT && warp == svc) store_partial();   // (rot_3d fused: its sum is formed in the next pass)
        if (tma) {
            if (a.abs_loc) bulk_s2g(a.abs_loc + (size_t)m0 * B3, G3, (uint32_t)(np * B3) * 4u);
            if (a.proj) bulk_s2g(a.proj + (size_t)m0 * B3, G2, (uint32_t)(np * B3) * 4u);
            bulk_commit();   // (the regions are free again once these stores have read them: see after barrier #5b)
        }
        // ---- normaliser backward through shift (hips) and scale (|neck - hips|): it needs the sums over all bones
        //      (three warps per slice, one per sum), and lands on g_hips and g_neck before the subtree sums are formed
        if (a.normalize) {
            if (warp < 3 * slices) {
                const int k = warp % 3, col = (warp / 3) * 32 + lane;
                float s0_ = 0.f, s1_ = 0.f;
                if (BT == 26) {
                    // every read in flight before the first addition (the additions keep their order): three warps
                    // hold up the whole CTA here, and a read-add-read-add chain cost 1.2 - 2.4 k cycles per tile
                    float v[26];
#pragma unroll
                    for (int q = 0; q < 26; ++q) v[q] = PART[(q * 3 + k) * PS + col];
#pragma unroll
                    for (int q = 0; q < 26; q += 2) {
                        s0_ += v[q];
                        s1_ += v[q + 1];
                    }
                } else {
#pragma unroll 1
                    for (int q = 0; q + 1 < B; q += 2) {
                        s0_ += PART[(q * 3 + k) * PS + col];
                        s1_ += PART[(q * 3 + 3 + k) * PS + col];
                    }
                }
                NRM[(6 + k) * PS + col] = s0_ + s1_;   // (B is even)
            }
            WARP_CLOCK(7)
            sync_tasks();   // #5a
            TILE_CLOCK(10)
        }
        // ---- own parts of the subtree sums, one bone per task: g_b x t_b -> rows XR (the former partial-sum rows)
        float* XR = PART;
        {
            bone_tasks<ROT ? PAIRS : 0>(warp, W, slices * B, [&](int task, auto par) {
                constexpr int PAR = decltype(par)::value;
                const int slice = task >= B ? 1 : 0, b = task - slice * B;
                const int col = slice * 32 + lane;
                const V3 t{TS[(b * 3) * PS + col], TS[(b * 3 + 1) * PS + col], TS[(b * 3 + 2) * PS + col]};
                V3 g{GV[(b * 3) * PS + col], GV[(b * 3 + 1) * PS + col], GV[(b * 3 + 2) * PS + col]};
                if (a.normalize && (b == a.hips || b == a.neck)) {
                    const int m = m0 + min(col, np - 1);
                    const float sgx = NRM[6 * PS + col], sgy = NRM[7 * PS + col], sgn = NRM[8 * PS + col];
                    const float inv = NRM[5 * PS + col] != 0.f ? NRM[4 * PS + col] : 0.f;
                    const float gs = -sgn * inv * inv;   // dL/dscale . (1/scale)
                    const float kx = gs * NRM[2 * PS + col], ky = gs * NRM[3 * PS + col];
                    float cx = 0.f, cy = 0.f;
                    if (b == a.neck) { cx += kx; cy += ky; }
                    if (b == a.hips) { cx -= fmaf(sgx, inv, kx); cy -= fmaf(sgy, inv, ky); }
                    const V3 gp = project_bwd<WORLD>(a, project<WORLD>(a, t, m), cx, cy, m);
                    g.x += gp.x; g.y += gp.y; g.z += gp.z;
                    GV[(b * 3) * PS + col] = g.x; GV[(b * 3 + 1) * PS + col] = g.y; GV[(b * 3 + 2) * PS + col] = g.z;
                }
                V3 x{g.y * t.z - g.z * t.y, g.z * t.x - g.x * t.z, g.x * t.y - g.y * t.x};
                if (ROT) {
                    // rot_3d (loss/rot_3d.py:9-37): MSE of the absolute rotation matrices.  Turning the chain at a joint
                    // by w turns every row of A_d in its subtree like a point (row x w), so the loss gradient G_d =
                    // k_r (A_d - Agt_d) acts as the torque  tau_d = sum_rows G_d,i x A_d,i = k_r sum_rows A_d,i x Agt_d,i,
                    // which joins the own part of the subtree sums.  The target record comes straight from global
                    // memory (prefetched into L2 at the top of the tile; 36 bytes per lane, L1 keeps the lines).
                    const int pl = min(col, np - 1);
                    const M3 A_ = lds9<PAR>(A + pl * B9, b);
                    const float* gq = a.gt_rot + ((size_t)(m0 + pl) * B + b) * 9;
                    float q[9];
#pragma unroll
                    for (int i = 0; i < 9; ++i) q[i] = __ldg(gq + i);
                    float e2 = 0.f;
#pragma unroll
                    for (int i = 0; i < 9; ++i) {
                        const float e = A_.m[i] - q[i];
                        e2 = fmaf(e, e, e2);
                    }
                    if (col < np) sr += e2;
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
                        const float ax = A_.m[3 * i], ay = A_.m[3 * i + 1], az = A_.m[3 * i + 2];
                        const float qx = q[3 * i], qy = q[3 * i + 1], qz = q[3 * i + 2];
                        x.x = fmaf(kr, ay * qz - az * qy, x.x);
                        x.y = fmaf(kr, az * qx - ax * qz, x.y);
                        x.z = fmaf(kr, ax * qy - ay * qx, x.z);
                    }
                }
                XR[(b * 3) * PS + col] = x.x;
                XR[(b * 3 + 1) * PS + col] = x.y;
                XR[(b * 3 + 2) * PS + col] = x.z;
            });
        }
        if (ROT) {
            sr = warp_sum(sr);
            if (lane == 0) red_r[warp] = sr;
        }
        WARP_CLOCK(8)
        if (SERVICE && ROT) named_arrive(6, (W + 1) * 32);   // (red_r is complete: the service warp may go on)
        sync_tasks();   // #5b
        if (tma) {
            // the 6-D records come back (L2 hot) over the two target regions, as soon as the stores of abs_loc and
            // projection have read them: asked for HERE, where the issuing thread's warp has no part in the subtree
            // walks (the wait for the stores used to hold the whole CTA at barrier #5a); phase E waits for the records
            bulk_wait_read_all();
            mbar_expect_tx(bars + 2, (uint32_t)(np * B6) * 4u);
            bulk_g2s(G3, a.d6 + (size_t)m0 * B6, (uint32_t)(np * B6) * 4u, bars + 2);
        }
        TILE_CLOCK(11)
        if (!SERVICE && ROT && warp == svc) store_partial();
        // few tiles: the warp that drew this CTA's ticket folds the grid's partials if it was the last -- NOW, when it has
        // no part in the subtree walks below, so that the fold (an L2 round trip, and with data-parallel ranks the peer
        // stores of the exchange, an NVLink round trip) sits behind the rest of the backward pass of this CTA
        const bool fold_early = W - 2 >= (lone_sums_early ? 15 : 6 * slices);
        if (!SERVICE && fold_early && nxt >= ntiles && a.warp_fold && warp == svc) fold_if_last();

        // ---- subtree sums, leaves first (the bone order has parents before children, so a walk in reverse bone order
        //      meets every child before its parent):  Sx_b = sum_{d in subtree(b)} g_d x t_d (rows XR), Sg_b = sum g_d
        //      (in place over g_b).  The six components are independent: one warp each per slice, no barrier inside
        //      the walk, the running value stays in a register along chains.
        if (lone_sums) {
            // (one tile per CTA) the limbs in parallel, three warps each (component k of both sums); the limb groups
            // leave the sum of their head bone in its own row and signal, the trunk group adds those rows when it
            // reaches spine01 (head, right arm) and hips (legs)
            if (warp < 15) {
                const int grp = warp / 3, k = warp - 3 * grp;
                float* rx = XR + k * PS + lane;
                float* rg = GV + k * PS + lane;
                acc_t cx = 0.f, cg = 0.f;
#define P2C_SUMS(HI, LO)                                                                      \
    _Pragma("unroll") for (int c = HI; c >= LO; --c) {                                        \
        acc_t vx = rx[c * 3 * PS], vg = rg[c * 3 * PS];                                       \
        bool changed = false;                                                                 \
        if (c + 1 <= HI && carla_parent(c + 1) == c) { vx += cx; vg += cg; changed = true; }  \
        if (LO == 0 && c == 3) {                                                              \
            named_barrier(5, 15 * 32);                                                        \
            vx += rx[8 * 3 * PS] + rx[12 * 3 * PS];                                           \
            vg += rg[8 * 3 * PS] + rg[12 * 3 * PS];                                           \
        }                                                                                     \
        if (LO == 0 && c == 1) {                                                              \
            vx += rx[16 * 3 * PS] + rx[21 * 3 * PS];                                          \
            vg += rg[16 * 3 * PS] + rg[21 * 3 * PS];                                          \
        }                                                                                     \
        if (changed || (LO == 0 && (c == 3 || c == 1))) { rx[c * 3 * PS] = (float)vx; rg[c * 3 * PS] = (float)vg; } \
        if (c > LO) {                                                                         \
            if (carla_parent(c) == c - 1) { cx = vx; cg = vg; }                               \
            else { rx[carla_parent(c) * 3 * PS] = (float)(rx[carla_parent(c) * 3 * PS] + vx);   \
                   rg[carla_parent(c) * 3 * PS] = (float)(rg[carla_parent(c) * 3 * PS] + vg); } \
        }                                                                                     \
    }
                if (grp == 0) {
                    P2C_SUMS(7, 0)
                } else {
                    if (grp == 1) { P2C_SUMS(11, 8) }
                    else if (grp == 2) { P2C_SUMS(15, 12) }
                    else if (grp == 3) { P2C_SUMS(20, 16) }
                    else { P2C_SUMS(25, 21) }
                    named_arrive(5, 15 * 32);
                }
#undef P2C_SUMS
            }
        } else if (warp < 6 * slices) {
            const int k = warp % 6, col = (warp / 6) * 32 + lane;
            float* rowp = (k < 3 ? XR + k * PS : GV + (k - 3) * PS) + col;
            if (BT == 26) {
                acc_t carry = 0.f;
#pragma unroll
                for (int c = 25; c >= 0; --c) {
                    const bool have = c + 1 < 26 && carla_parent(c + 1) == c;   // the next bone is this one's child
                    acc_t v = rowp[c * 3 * PS];
                    if (have) {
                        v += carry;
                        rowp[c * 3 * PS] = (float)v;
                    }
                    if (c > 0) {
                        if (carla_parent(c) == c - 1) carry = v;
                        else rowp[carla_parent(c) * 3 * PS] = (float)(rowp[carla_parent(c) * 3 * PS] + v);
                    }
                }
            } else {
                float carry = 0.f;
                bool have = false;
                float nv = rowp[(B - 1) * 3 * PS];
                int npar = a.parent[B - 1];
    #pragma unroll 1
                for (int c = B - 1; c >= 1; --c) {
                    float v = nv;
                    const int par = npar;
                    nv = rowp[(c - 1) * 3 * PS];   // (row c-1 is not touched in this trip unless it is the parent; see below)
                    npar = a.parent[c - 1];
                    if (have) {
                        v += carry;
                        rowp[c * 3 * PS] = v;
                    }
                    if (par == c - 1) {
                        carry = v;
                        have = true;
                    } else {
                        rowp[par * 3 * PS] += v;
                        have = false;
                    }
                }
                if (have) rowp[0] = nv + carry;
            }
        }
        // (two warps without a column when W >= 6 slices + 2: both jobs wait on global memory, off the critical path)
        if (nxt < ntiles && warp == W - 1) fill_rows(nxt, (it + 1) & 1);
        if (!SERVICE && !fold_early && nxt >= ntiles && a.warp_fold && warp == svc) fold_if_last();   // (the ticket was drawn after the loss pass)
        WARP_CLOCK(9)
        sync_tasks();   // #6
        TILE_CLOCK(8)
        // ---- per bone, all independent:  m_b = Sx_b - Sg_b x t_b,  y_b = A_parent m_b  (in place over Sx_b).  After this
        //      pass the scratch rows (t, Sg) and the abs_rot records are dead, so the next tile's 6-D records can start
        //      their way in while the gradient pass runs.
        // (one tile per CTA, one slice: bone b's y and its gradient pass run in the SAME warp -- y stays in registers, no
        // pass of its own, no barrier #7: MERGE_YE)
        constexpr bool MERGE_YE = BT == 26 && LONE && SL == 1 && P2C_MERGE_YE;
        auto y_of = [&](int b, int col, int pl) {
            const int par = par_s[b];
            const V3 sx{XR[(b * 3) * PS + col], XR[(b * 3 + 1) * PS + col], XR[(b * 3 + 2) * PS + col]};
            const V3 sg{GV[(b * 3) * PS + col], GV[(b * 3 + 1) * PS + col], GV[(b * 3 + 2) * PS + col]};
            const V3 t{TS[(b * 3) * PS + col], TS[(b * 3 + 1) * PS + col], TS[(b * 3 + 2) * PS + col]};
            const V3 mv{sx.x - (sg.y * t.z - sg.z * t.y), sx.y - (sg.z * t.x - sg.x * t.z),
                        sx.z - (sg.x * t.y - sg.y * t.x)};
            V3 y = mv;
            if (par >= 0) {   // (the parent's parity is a run-time value: two access sequences behind a branch)
                if (par & 1) y = v3_mul_mt(mv, lds9<1>(A + pl * B9, par));
                else y = v3_mul_mt(mv, lds9<0>(A + pl * B9, par));
            }
            if (col >= np) y = V3{0.f, 0.f, 0.f};
            return y;
        };
        if (!MERGE_YE) {
            const int nt = slices * B;
#pragma unroll 1
            for (int task = warp; task < nt; task += W) {
                const int sl = task >= B ? 1 : 0, b = task - sl * B;
                const int par = par_s[b];
                const int col = sl * 32 + lane, pl = min(col, np - 1);
                const V3 sx{XR[(b * 3) * PS + col], XR[(b * 3 + 1) * PS + col], XR[(b * 3 + 2) * PS + col]};
                const V3 sg{GV[(b * 3) * PS + col], GV[(b * 3 + 1) * PS + col], GV[(b * 3 + 2) * PS + col]};
                const V3 t{TS[(b * 3) * PS + col], TS[(b * 3 + 1) * PS + col], TS[(b * 3 + 2) * PS + col]};
                const V3 mv{sx.x - (sg.y * t.z - sg.z * t.y), sx.y - (sg.z * t.x - sg.x * t.z),
                            sx.z - (sg.x * t.y - sg.y * t.x)};
                V3 y = mv;
                if (par >= 0) {
                    if (par & 1) y = v3_mul_mt(mv, lds9<1>(A + pl * B9, par));
                    else y = v3_mul_mt(mv, lds9<0>(A + pl * B9, par));
                }
                if (col >= np) y = V3{0.f, 0.f, 0.f};
                XR[(b * 3) * PS + col] = y.x; XR[(b * 3 + 1) * PS + col] = y.y; XR[(b * 3 + 2) * PS + col] = y.z;
            }
        }
        // (the scratch rows are about to be overwritten by TMA -- when a tile follows; on the CTA's last tile the fence
        // would only make the thread that folded the grid's loss sums wait for its global stores: 3 k cycles)
        if (nxt < ntiles) fence_async_smem();
        WARP_CLOCK(10)
        if (!MERGE_YE) sync_tasks();   // #7
        TILE_CLOCK(9)
        if (tma && nxt < ntiles) load_d6(nxt);
        while (!mbar_try_wait(bars + 2, it & 1u)) {}   // the 6-D records are back
        // ---- z[t] = sum_{s >= t} y[s] over the frames of the clip (lanes are frames: a segmented suffix scan over
        //      shuffles),  omega = R[t] z[t],  tangent-form Gram-Schmidt backward -> gradient records over the 6-D
        //      records (which now live in the two target regions)
        {
            // (a clip over two slices: one bone per warp, both slices in turn -- no pairs, the warps have one task each)
            bone_tasks<(PAIRS == 1 && SL != 1) ? P2C_LONE_TASKS : PAIRS>(warp, W, B, [&](int b, auto par) {
                constexpr int PAR = decltype(par)::value;
                V3 later{0.f, 0.f, 0.f};   // sum over the later slice of the same clip (clips longer than 32 frames)
#pragma unroll 1
                for (int sl = slices - 1; sl >= 0; --sl) {
                    const int col = sl * 32 + lane, pl = min(col, np - 1);
                    const bool valid = col < np;
                    V3 z = MERGE_YE ? y_of(b, col, pl)
                                    : V3{XR[(b * 3) * PS + col], XR[(b * 3 + 1) * PS + col], XR[(b * 3 + 2) * PS + col]};
                    if (a.scan) {
                        const unsigned ok = scan_ok_s[sl * 32 + lane];
#pragma unroll
                        for (int k = 0; k < 5; ++k) {
                            const int off = 1 << k;
                            if (off >= L) break;
                            const float vx = __shfl_down_sync(FULL_MASK, z.x, off), vy = __shfl_down_sync(FULL_MASK, z.y, off),
                                        vz = __shfl_down_sync(FULL_MASK, z.z, off);
                            if (ok & (1u << k)) { z.x += vx; z.y += vy; z.z += vz; }
                        }
                        if (slices > 1) {
                            z.x += later.x; z.y += later.y; z.z += later.z;
                            later = V3{__shfl_sync(FULL_MASK, z.x, 0), __shfl_sync(FULL_MASK, z.y, 0),
                                       __shfl_sync(FULL_MASK, z.z, 0)};
                        }
                    }
                    float d[6], o[6];
                    lds6<PAR>(G3 + pl * B6, b, d);
                    const V3 om = v3_mul_mt(z, lds9<PAR>(R + pl * B9, b));
                    gram_schmidt_bwd_tangent(gram_schmidt(d), om, o);
                    if (valid) sts6<PAR>(G3 + pl * B6, b, o);
                }
            });
        }
        fence_async_smem();
        WARP_CLOCK(11)
        sync_tasks();   // #8
        TILE_CLOCK(12)
        if (tma) {
            bulk_s2g(a.g_d6 + (size_t)m0 * B6, G3, (uint32_t)(np * B6) * 4u);
            bulk_commit();
            bulk_wait_read_all();             // the gradient has left the target regions: the next tile's targets may land
            if (nxt < ntiles) load_targets(nxt);
        }
    }
    CTA_CLOCK(18)
    if (tma) bulk_wait_all();
    CTA_CLOCK(19)

    if (it == 0 && warp == 0) {   // (a CTA without a tile still takes its ticket)
        if (lane == 0) take_ticket();
        if (a.warp_fold) fold_if_last();
    }
    if (!a.warp_fold && my_ticket == gridDim.x - 1) last_s = 1;
    __syncthreads();
    if (last_s) {
        // every CTA has stored its partials and taken its ticket: fold them in tile order, in double, all threads
        __threadfence();
        double d2 = 0.0, d3 = 0.0, dr = 0.0;
        for (int i = tid; i < ntiles; i += NT) {
            const float4 v = __ldcg(reinterpret_cast<const float4*>(a.partials) + i);
            d2 += (double)v.x;
            d3 += (double)v.y;
            dr += (double)v.z;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            d2 += __shfl_xor_sync(FULL_MASK, d2, o);
            d3 += __shfl_xor_sync(FULL_MASK, d3, o);
            dr += __shfl_xor_sync(FULL_MASK, dr, o);
        }
        double* dred = reinterpret_cast<double*>(R);   // (no bulk copy still reads the relative-rotation region)
        if (lane == 0) { dred[warp * 3] = d2; dred[warp * 3 + 1] = d3; dred[warp * 3 + 2] = dr; }
        __syncthreads();
        if (tid == 0) {
            d2 = 0.0; d3 = 0.0; dr = 0.0;
            for (int w = 0; w < W; ++w) { d2 += dred[w * 3]; d3 += dred[w * 3 + 1]; dr += dred[w * 3 + 2]; }
            finish(d2, d3, dr);
        }
    }
    CTA_CLOCK(20)
    DEBUG_CLOCK(if (a.clocks && threadIdx.x == 0 && blockIdx.x < 1024) g_cta_times[blockIdx.x * 2 + 1] = globaltimer();)
}

int sm_count() { return p2c_sm_count(); }   // of the current device (host_util.cu)

// parents of the bone tree (parents precede children: pose.py:41-47)
bool tree_tables(const p2c_tree_t& t, TileArgs& a) {
    const int B = t.n_bones;
    if (B < 2 || B > P2C_MAX_BONES || (B & 1)) return false;   // AoS records must be even word counts
    if (t.parent[0] != -1) return false;
    for (int b = 1; b < B; ++b)
        if (t.parent[b] < 0 || t.parent[b] >= b) return false;
    for (int b = 0; b < B; ++b) a.parent[b] = (int8_t)t.parent[b];
    a.B = B;
    return true;
}

struct TileLaunch {
    int ok, P, PS, slices, threads, grid, count_inline, dynamic;
    size_t smem;
};

TileLaunch make_launch(const p2c_fused_step_t* p, const TileArgs& a) {   // a: tree tables filled in
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
    if (total > (1L << 26)) return t;   // pose and joint indices are 32-bit inside the kernels
    const void* ptrs[] = {p->pose_in, p->gt_2d, p->gt_3d, p->g_pose_in, p->rel_rot, p->abs_rot, p->abs_loc, p->proj,
                          p->gt_abs_rot};
    for (const void* q : ptrs)
        if (reinterpret_cast<uintptr_t>(q) & 15) return t;
    if (reinterpret_cast<uintptr_t>(p->workspace) & 15) return t;   // (64-bit counter word, float4 tile partials)
    t.PS = (t.P + 31) & ~31;
    t.slices = t.PS / 32;
    const long ntiles = (total + t.P - 1) / t.P;
    // many tiles: two bones per warp and two CTAs per SM; at most one tile per SM (the 256 x 16 step): one bone per
    // warp, which halves the latency of the element-wise passes
    const bool lone = ntiles <= sm_count();
    t.threads = tile_threads(B, t.slices, lone);
    const int warps = t.threads / 32;
    if (warps < 6 * t.slices) return t;   // the subtree-sum walks take six warps per slice (tiny trees: other kernels)
    t.smem = (size_t)tile_plan(t.PS, B).total * sizeof(float);
    const size_t cap = 227 * 1024;
    if (t.smem + 1024 > cap) return t;
    int per_sm = (int)((228 * 1024) / (t.smem + 1024));
    if (per_sm * t.threads > 896) per_sm = 896 / t.threads;   // 72 registers per thread (launch bounds 832 x 1)
    if (per_sm > 4) per_sm = 4;
    if (per_sm < 1) per_sm = 1;
    if (p2c_env().tile_one_cta) per_sm = 1;   // debug: no co-resident CTA
    const long slots = (long)sm_count() * per_sm;
    t.grid = (int)(ntiles < slots ? ntiles : slots);
    t.dynamic = ntiles > t.grid ? 1 : 0;
    t.count_inline = lone ? 1 : 0;   // grid == ntiles <= SMs: every CTA is resident, the in-kernel wait cannot hang
    t.ok = 1;
    return t;
}

}  // namespace

extern "C" int p2c_fused_tile_supported(const p2c_fused_step_t* p) {
    // test hooks: the older kernel variants stay reachable
    const P2cEnv& env = p2c_env();
    if (env.disable_tile || env.disable_fast || env.force_slim || env.disable_tma) return 0;
    if (!p->input_6d) return 0;   // arbitrary 3x3 input is not assumed to be a rotation: exact generic path
    if (p->mode != P2C_MODE_POSE_CHANGES && p->mode != P2C_MODE_RELATIVE_ROT) return 0;
    if (p->mode == P2C_MODE_POSE_CHANGES && !p->ref_rot) return 0;
    if (p->gt_2d_channels != 3) return 0;   // the projection records replace the 2-D target records in place
    TileArgs a{};
    if (!tree_tables(p->tree, a)) return 0;
    return make_launch(p, a).ok;
}

// debug: the phase clocks recorded by CTA 0 (P2C_TILE_CLOCKS=1); host buffer of 32 values
extern "C" int p2c_debug_cta_times(unsigned long long* out) {
    return (int)cudaMemcpyFromSymbol(out, g_cta_times, sizeof(unsigned long long) * 2048);
}
extern "C" int p2c_debug_warp_clocks(long long* out) {
    return (int)cudaMemcpyFromSymbol(out, g_warp_clocks, sizeof(long long) * 512);
}
extern "C" int p2c_debug_tile_clocks(long long* out) {
    return (int)cudaMemcpyFromSymbol(out, g_tile_clocks, sizeof(long long) * 64);
}

extern "C" int p2c_fused_tile_launches(const p2c_fused_step_t* p) {
    TileArgs a{};
    if (!tree_tables(p->tree, a)) return 1;
    return (p->mask_missing_joints && !make_launch(p, a).count_inline) ? 2 : 1;
}

extern "C" int p2c_fused_tile(const p2c_fused_step_t* p, void* stream) {
    TileArgs a{};
    if (!tree_tables(p->tree, a)) return P2C_E_UNSUPPORTED;
    const TileLaunch t = make_launch(p, a);
    if (!t.ok) return P2C_E_UNSUPPORTED;
    const int B = p->tree.n_bones;
    a.N = p->n_clips; a.L = p->clip_len; a.P = t.P; a.PS = t.PS; a.slices = t.slices;
    a.clock_cta = p2c_env().tile_cta;
    a.clocks = p2c_env().tile_clocks;   // n: record the n-th tile of CTA 0
    a.scan = p->mode == P2C_MODE_POSE_CHANGES;
    a.hips = p->hips; a.neck = p->neck; a.normalize = p->normalize; a.mask = p->mask_missing_joints;
    a.d6 = p->pose_in; a.ref_loc = p->ref_loc; a.ref_rot = p->ref_rot; a.ref_index = p->ref_index;
    a.world_loc = p->world_loc; a.world_rot = p->world_rot; a.cam = p->camera;
    a.gt2 = p->gt_2d; a.gt3 = p->gt_3d; a.w2 = p->weight_2d; a.w3 = p->weight_3d;
    a.gt_rot = p->gt_abs_rot; a.wr = p->gt_abs_rot ? p->weight_rot : 0.f;
    a.rel_rot = p->rel_rot; a.abs_rot = p->abs_rot; a.abs_loc = p->abs_loc; a.proj = p->proj;
    a.proj_norm = p->proj_norm; a.g_d6 = p->g_pose_in;
    a.ticket = reinterpret_cast<unsigned int*>(p->workspace);
    a.count2 = reinterpret_cast<int*>(p->workspace) + 2;             // (8-byte aligned: count2 and arrive are also
    a.arrive = reinterpret_cast<unsigned int*>(p->workspace) + 3;    //  updated together as one 64-bit word)
    a.next = reinterpret_cast<unsigned int*>(p->workspace) + 4;
    a.count_inline = t.count_inline;
    a.dynamic = t.dynamic;
    a.warp_fold = ((long)p->n_clips * p->clip_len + t.P - 1) / t.P <= 1024;
    // one partial pair per tile: at most n_poses floats, inside the scratch area every caller provides
    // (p2c_fused_step_workspace_floats: 16 header words, then >= 6 * n_bones * n_poses floats of scratch)
    a.partials = p->workspace + 16;
    a.losses = p->losses;
    a.xchg = XchgArgs{p->exchange, p->exchange_rank, p->exchange_world};
    const long total = (long)p->n_clips * p->clip_len;
    a.count3 = (double)total * B * 3;
    a.total = (int)total;
    a.ntiles = (int)((total + t.P - 1) / t.P);
    a.clips_full = t.P / p->clip_len;
    a.whole_clips = a.clips_full * p->clip_len == t.P ? 1 : 0;
    a.k2_all = p->weight_2d * 2.0f / ((float)total * (float)(B * 2));
    a.k3 = p->weight_3d * 2.0f / (float)a.count3;
    a.kr = a.wr * 2.0f / (float)(a.count3 * 3.0);
    cudaStream_t st = (cudaStream_t)stream;
    if (a.mask && !a.count_inline) {
        long blocks = (total * B / 4 + 511) / 512;   // two 48-byte groups per thread and trip
        if (blocks > 148 * 8) blocks = 148 * 8;
        tile_count_kernel<<<(int)blocks, 256, 0, st>>>(p->gt_2d, (unsigned)(total * B), (unsigned)B,
                                                       (unsigned)p->gt_2d_channels, (unsigned)p->hips, a.count2);
        P2C_LAUNCH_CHECK();
    }
    typedef void (*kern_t)(const TileArgs);
    const bool world = p->world_loc || p->world_rot;
    bool carla = B == 26;
    for (int b = 0; b < B && carla; ++b) carla = p->tree.parent[b] == carla_parent(b);
    // CARLA_SKELETON: compile-time tree and slices, one instance per launch shape (limb walk needs 5 warps per slice)
    const bool limb_ok = t.threads / 32 > 5 * t.slices;
    const int ki = world ? 5 : !carla ? 4 : t.count_inline ? (t.slices == 1 ? 2 : 3) : !limb_ok ? 4 : (t.slices == 1 ? 0 : 1);
    const bool rot = p->gt_abs_rot != nullptr;
    const int threads = t.threads + ((P2C_SERVICE_WARP && (ki == 2 || ki == 3)) ? 32 : 0);   // + the service warp
    const kern_t k = rot ? (ki == 0   ? tile_kernel<26, 1, false, false, true>
                            : ki == 1 ? tile_kernel<26, 2, false, false, true>
                            : ki == 2 ? tile_kernel<26, 1, false, true, true>
                            : ki == 3 ? tile_kernel<26, 2, false, true, true>
                            : ki == 4 ? tile_kernel<0, 0, false, false, true>
                                      : tile_kernel<0, 0, true, false, true>)
                         : (ki == 0   ? tile_kernel<26, 1, false, false, false>
                            : ki == 1 ? tile_kernel<26, 2, false, false, false>
                            : ki == 2 ? tile_kernel<26, 1, false, true, false>
                            : ki == 3 ? tile_kernel<26, 2, false, true, false>
                            : ki == 4 ? tile_kernel<0, 0, false, false, false>
                                      : tile_kernel<0, 0, true, false, false>);
    if (const int e = p2c_allow_big_smem(reinterpret_cast<const void*>(k), 227 * 1024 - 1024)) return e;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(t.grid);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = t.smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    attr[1].id = cudaLaunchAttributeCooperative;
    attr[1].val.cooperative = 1;
    cfg.attrs = attr;
    // programmatic dependent launch: with the count pre-pass in front, this kernel overlaps it up to the loss pass; with
    // one tile per CTA (no pre-pass) the CTAs may be placed while the previous kernel of the stream drains, and wait
    // (griddepcontrol.wait, before their first global access) until it is complete and flushed.
    // One tile per CTA with the mask on: the CTAs wait for each other's joint counts inside the kernel, so the launch
    // is COOPERATIVE (the whole grid is resident at once or the launch fails: two such kernels on two streams cannot
    // starve each other).
    cfg.numAttrs = (a.mask && a.count_inline && !p2c_env().tile_no_coop) ? 2 : 1;
    cudaError_t le = cudaLaunchKernelEx(&cfg, k, a);
    if (le != cudaSuccess && cfg.numAttrs == 2) {   // (the two attributes together are refused: cooperative only)
        (void)cudaGetLastError();
        cfg.attrs = attr + 1;
        cfg.numAttrs = 1;
        le = cudaLaunchKernelEx(&cfg, k, a);
    }
    if (le != cudaSuccess) return (int)le;
    P2C_LAUNCH_CHECK();
    return 0;
}
