// fk_tile: the module API's FK + projection, forward AND exact backward, pose per lane.
//
// Behind p2c_fk_fwd / p2c_fk_bwd (include/p2c_b200.h), i.e. behind ops.fk_clip -> ProjectionModule.forward,
// P3dPose.forward / relative_to_absolute, the reference-skeleton tables and the device target generator -- the calls
// the UNMODIFIED reference flows make with 3x3 matrices (modules/movements/movements.py:105-118 ->
// modules/layers/projection.py:73-123,144-195 -> walker_control/p3d_pose.py:98-213, p3d_pose_projection.py:115-152).
// Nothing is assumed about the matrices: the reference's own test feeds rand*0.8 "changes"
// (tests/walker_control/test_p3d_pose.py:93-95), so the forward is plain matrix products and the backward the full
// matrix chain rule (no rotation shortcut, unlike the fused 6-D step of fused_tile.cu).
//
// Mapping (same idea as fused_tile.cu): a CTA owns a tile of up to 32 consecutive poses -- whole clips side by side in
// the lanes (clip_len <= 32), or 32 frames of ONE clip (clip_len > 32: the CTA walks its clip tile by tile and carries
// the scan state across, so any clip length runs) -- every LANE is a POSE, a warp task is a bone (element-wise passes),
// a limb (tree walks) or a (clip, bone, column) item (frame scans).  HBM traffic is TMA only: cp.async.bulk in for the
// tile's records, out for every output region; records stay AoS in shared memory (64-bit conflict-free accesses).
// One tile per CTA, three (forward) / two (backward) CTAs per SM: the hardware CTA scheduler balances the tiles.
#include <initializer_list>

#include "common.cuh"
#include "host_util.h"
#include "tile_common.cuh"

namespace {

constexpr int FT = 256;   // threads per CTA
constexpr int FW = FT / 32;

struct FkTile {
    int N, L, B, scan;
    int P;          // poses per CTA when a CTA owns whole clips (clip_len <= 32) or free poses (relative_rot)
    int nsub;       // tiles per clip when clip_len > 32 in pose_changes mode (the CTA owns ONE clip), else 1
    int loc_rows;   // skeleton-offset rows staged per tile
    int8_t parent[P2C_MAX_BONES];
    const float* pose_in;
    const float* ref_loc;
    const float* ref_rot;
    const int32_t* ref_index;
    const float* world_loc;
    const float* world_rot;
    p2c_camera_t cam;
    float* rel_rot;
    float* abs_rot;
    float* abs_loc;
    float* proj;
    // backward: saved forward tensors, upstream gradients, result
    const float* s_rel_rot;
    const float* s_abs_rot;
    const float* s_abs_loc;
    const float* g_rel_rot;
    const float* g_abs_rot;
    const float* g_abs_loc;
    const float* g_proj;
    float* g_pose_in;
};

struct World {   // per-pose world transform (utils/world.py output), identity when absent
    float r[9], l[3];
};
__device__ __forceinline__ World load_world(const FkTile& a, long m) {
    World w;
#pragma unroll
    for (int i = 0; i < 9; ++i) w.r[i] = (i == 0 || i == 4 || i == 8) ? 1.f : 0.f;
    w.l[0] = w.l[1] = w.l[2] = 0.f;
    if (a.world_rot) {
#pragma unroll
        for (int i = 0; i < 9; ++i) w.r[i] = __ldg(a.world_rot + m * 9 + i);
    }
    if (a.world_loc) {
#pragma unroll
        for (int i = 0; i < 3; ++i) w.l[i] = __ldg(a.world_loc + m * 3 + i);
    }
    return w;
}
// p3d_pose_projection.py:137-152: (x,y,z) -> (y,-x,z) -> . world_rot + world_loc -> camera -> (u, v, 1/z)
__device__ __forceinline__ V3 project_pt(const p2c_camera_t& cam, const World& w, const V3& t, V3* cpt) {
    const float x0 = t.y, x1 = -t.x, x2 = t.z;
    const float wx = fmaf(x2, w.r[6], fmaf(x1, w.r[3], x0 * w.r[0])) + w.l[0];
    const float wy = fmaf(x2, w.r[7], fmaf(x1, w.r[4], x0 * w.r[1])) + w.l[1];
    const float wz = fmaf(x2, w.r[8], fmaf(x1, w.r[5], x0 * w.r[2])) + w.l[2];
    const float* cr = cam.R;
    V3 c;
    c.x = fmaf(wz, cr[6], fmaf(wy, cr[3], wx * cr[0])) + cam.T[0];
    c.y = fmaf(wz, cr[7], fmaf(wy, cr[4], wx * cr[1])) + cam.T[1];
    c.z = fmaf(wz, cr[8], fmaf(wy, cr[5], wx * cr[2])) + cam.T[2];
    if (cpt) *cpt = c;
    const float iz = frcp(c.z);
    return V3{fmaf(-cam.f * c.x, iz, cam.px), fmaf(-cam.f * c.y, iz, cam.py), iz};
}
// upstream (gu, gv, gq) on (u, v, 1/z) -> gradient wrt the p3d-space point
__device__ __forceinline__ V3 project_pt_bwd(const p2c_camera_t& cam, const World& w, const V3& t, const V3& g) {
    V3 c;
    (void)project_pt(cam, w, t, &c);
    const float iz = frcp(c.z);
    const float gcx = -cam.f * g.x * iz, gcy = -cam.f * g.y * iz;
    const float gcz = (cam.f * fmaf(g.x, c.x, g.y * c.y) - g.z) * iz * iz;
    const float* cr = cam.R;
    const float gwx = fmaf(gcz, cr[2], fmaf(gcy, cr[1], gcx * cr[0]));
    const float gwy = fmaf(gcz, cr[5], fmaf(gcy, cr[4], gcx * cr[3]));
    const float gwz = fmaf(gcz, cr[8], fmaf(gcy, cr[7], gcx * cr[6]));
    const float g0 = fmaf(gwz, w.r[2], fmaf(gwy, w.r[1], gwx * w.r[0]));
    const float g1 = fmaf(gwz, w.r[5], fmaf(gwy, w.r[4], gwx * w.r[3]));
    const float g2 = fmaf(gwz, w.r[8], fmaf(gwy, w.r[7], gwx * w.r[6]));
    return V3{-g1, g0, g2};
}

// 9-word record b of an even-word AoS pose record as 64-bit accesses (b may be a run-time value: the parity branch is
// warp-uniform)
__device__ __forceinline__ void rec_ld(const float* base, int b, float* x) {
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
}
__device__ __forceinline__ void rec_st(float* base, int b, const float* x) {
    float* q = base + 9 * b;
    if (b & 1) {
        q[0] = x[0];
        sts64(q + 1, x[1], x[2]); sts64(q + 3, x[3], x[4]); sts64(q + 5, x[5], x[6]); sts64(q + 7, x[7], x[8]);
    } else {
        sts64(q, x[0], x[1]); sts64(q + 2, x[2], x[3]); sts64(q + 4, x[4], x[5]); sts64(q + 6, x[6], x[7]);
        q[8] = x[8];
    }
}

__device__ __forceinline__ void bulk_wait_read_1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }

struct TileSpan {
    long m0;    // first pose of the tile
    int np;     // poses in it
    int lt;     // frames of one clip inside the tile
    int clips;  // whole clips in the tile (scan mode)
};
__device__ __forceinline__ TileSpan tile_span(const FkTile& a, int sub) {
    TileSpan s;
    const long total = (long)a.N * a.L;
    if (a.nsub > 1) {
        s.m0 = (long)blockIdx.x * a.L + sub * 32;
        s.np = min(32, a.L - sub * 32);
        s.lt = s.np;
        s.clips = 1;
    } else {
        s.m0 = (long)blockIdx.x * a.P;
        const long left = total - s.m0;
        s.np = (int)(left < a.P ? left : a.P);
        s.lt = a.L;
        s.clips = a.scan ? s.np / a.L : 0;
    }
    return s;
}

// ------------------------------------------------------------------------------------------------ forward
// shared memory (floats): X 32*B9 (input changes -> abs_rot records) | R 32*B9 (rel_rot records -> abs_loc + projection
// records once the walk has consumed them) | TS 3B*32 (abs_loc, SoA rows) | LOC rows*3B | CARRY B9 | ROW 32 | barrier
template <int BT>
__global__ void __launch_bounds__(FT, 3) fk_tile_fwd_kernel(const FkTile a) {
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int B = BT ? BT : a.B, B3 = B * 3, B9 = B * 9, L = a.L;
    float* X = smem;
    float* R = X + 32 * B9;
    float* TS = R + 32 * B9;
    float* LOC = TS + 32 * B3;
    float* CARRY = LOC + a.loc_rows * B3;
    int* ROW = reinterpret_cast<int*>(CARRY + B9);
    uint64_t* bar = reinterpret_cast<uint64_t*>(ROW + 32);
    const bool tma = tid == FT - 32;   // first lane of the last warp (a warp without a limb of its own)
    if (tma) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    for (int sub = 0; sub < a.nsub; ++sub) {
        const TileSpan sp = tile_span(a, sub);
        const long m0 = sp.m0;
        const int np = sp.np;
        float* IN = a.scan ? X : R;
        if (tma) {
            const uint32_t bytes = (uint32_t)(np * B9) * 4u;
            mbar_expect_tx(bar, bytes);
            bulk_g2s(IN, a.pose_in + (size_t)m0 * B9, bytes, bar);
        }
        const long clip0 = m0 / L;
        const int nclip = (int)((m0 + np - 1) / L - clip0) + 1;
        if (warp == 1) {
            const long clip = (m0 + min(lane, np - 1)) / L;
            ROW[lane] = a.ref_index ? __ldg(a.ref_index + clip) : (int)clip;
        }
        for (int i = tid; i < nclip * B3; i += FT) {   // bone offsets of every skeleton the tile touches
            const int cl = i / B3, k = i - cl * B3;
            const long clip = clip0 + cl;
            const long row = a.ref_index ? (long)__ldg(a.ref_index + clip) : clip;
            LOC[i] = __ldg(a.ref_loc + row * B3 + k);
        }
        while (!mbar_try_wait(bar, sub & 1u)) {}
        __syncthreads();
        // ---- R[t] = C[t] . R[t-1] (p3d_pose.py:98-114 looped by layers/projection.py:190-193): one thread per (clip,
        //      bone, column of R), serial over the frames the tile holds of that clip
        if (a.scan) {
            const int items = sp.clips * B3;
#pragma unroll 1
            for (int idx = tid; idx < items; idx += FT) {
                const int cl = idx / B3, rem = idx - cl * B3, b = rem / 3, j = rem - b * 3;
                float p0, p1, p2;
                if (sub == 0) {
                    const float* rr = a.ref_rot + ((size_t)ROW[cl * sp.lt] * B + b) * 9 + j;
                    p0 = __ldg(rr); p1 = __ldg(rr + 3); p2 = __ldg(rr + 6);
                } else {
                    p0 = CARRY[b * 9 + j]; p1 = CARRY[b * 9 + 3 + j]; p2 = CARRY[b * 9 + 6 + j];
                }
                const float* src = X + (cl * sp.lt) * B9 + b * 9;
                float* dst = R + (cl * sp.lt) * B9 + b * 9 + j;
                float c[9];
#pragma unroll
                for (int i = 0; i < 9; ++i) c[i] = src[i];
#pragma unroll 1
                for (int f = 0; f < sp.lt; ++f) {
                    float e[9];
                    if (f + 1 < sp.lt) {
#pragma unroll
                        for (int i = 0; i < 9; ++i) e[i] = src[B9 + i];
                    }
                    const float n0 = fmaf(c[2], p2, fmaf(c[1], p1, c[0] * p0));
                    const float n1 = fmaf(c[5], p2, fmaf(c[4], p1, c[3] * p0));
                    const float n2 = fmaf(c[8], p2, fmaf(c[7], p1, c[6] * p0));
                    dst[0] = n0; dst[3] = n1; dst[6] = n2;
                    p0 = n0; p1 = n1; p2 = n2;
                    src += B9;
                    dst += B9;
                    if (f + 1 < sp.lt) {
#pragma unroll
                        for (int i = 0; i < 9; ++i) c[i] = e[i];
                    }
                }
                if (a.nsub > 1) { CARRY[b * 9 + j] = p0; CARRY[b * 9 + 3 + j] = p1; CARRY[b * 9 + 6 + j] = p2; }
            }
        }
        fence_async_smem();
        __syncthreads();
        if (tma && a.rel_rot) {
            bulk_s2g(a.rel_rot + (size_t)m0 * B9, R, (uint32_t)(np * B9) * 4u);
            bulk_commit();
        }
        // ---- FK (p3d_pose.py:116-184): A_b = R_b . A_parent, t_b = loc_b . A_parent + t_parent.  CARLA tree: one warp
        //      per limb forming full 3x3 products in straight-line code, the trunk warp signals hips / spine01 on named
        //      barriers; any other tree: one warp per column of A walking the bones in order.
        {
            const int pl = min(lane, np - 1);
            const bool valid = lane < np;
            const float* rb = R + pl * B9;
            const float* lc = LOC + (int)((m0 + pl) / L - clip0) * B3;
            if (BT == 26 && warp < 5) {
                float* ab = X + pl * B9;
                float P[9], tp[3];
                auto bone = [&](int b, int par, bool head) {
                    if (head || par != b - 1) {
                        if (par >= 0) {
                            rec_ld(ab, par, P);
                            tp[0] = TS[(par * 3) * 32 + lane]; tp[1] = TS[(par * 3 + 1) * 32 + lane];
                            tp[2] = TS[(par * 3 + 2) * 32 + lane];
                        } else {
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
                    TS[(3 * b) * 32 + lane] = tp[0]; TS[(3 * b + 1) * 32 + lane] = tp[1]; TS[(3 * b + 2) * 32 + lane] = tp[2];
                };
#define FK_RUN(B0, B1) \
    _Pragma("unroll") for (int b = B0; b < B1; ++b) bone(b, carla_parent(b), b == B0);
                if (warp == 0) {
                    FK_RUN(0, 2)
                    named_arrive(3, 96);   // hips
                    FK_RUN(2, 4)
                    named_arrive(4, 96);   // spine01
                    FK_RUN(4, 8)           // left arm
                } else if (warp == 1) {
                    named_barrier(4, 96);
                    FK_RUN(8, 12)          // neck, head, eyes
                } else if (warp == 2) {
                    named_barrier(4, 96);
                    FK_RUN(12, 16)         // right arm
                } else if (warp == 3) {
                    named_barrier(3, 96);
                    FK_RUN(16, 21)         // right leg
                } else {
                    named_barrier(3, 96);
                    FK_RUN(21, 26)         // left leg
                }
#undef FK_RUN
            } else if (BT != 26 && warp < 3) {
                const int jc = warp;
                float* ab = X + pl * B9 + jc;
                float c0 = 0.f, c1 = 0.f, c2 = 0.f, ct = 0.f;
#pragma unroll 1
                for (int b = 0; b < B; ++b) {
                    const int par = a.parent[b];
                    float p0 = c0, p1 = c1, p2 = c2, tp = ct;
                    if (par != b - 1 || par < 0) {
                        if (par >= 0) {
                            const float* ap = X + pl * B9 + par * 9 + jc;
                            p0 = ap[0]; p1 = ap[3]; p2 = ap[6];
                            tp = TS[(par * 3 + jc) * 32 + lane];
                        } else {
                            p0 = jc == 0 ? 1.f : 0.f; p1 = jc == 1 ? 1.f : 0.f; p2 = jc == 2 ? 1.f : 0.f;
                            tp = 0.f;
                        }
                    }
                    float x[9];
                    rec_ld(rb, b, x);
                    c0 = fmaf(x[2], p2, fmaf(x[1], p1, x[0] * p0));
                    c1 = fmaf(x[5], p2, fmaf(x[4], p1, x[3] * p0));
                    c2 = fmaf(x[8], p2, fmaf(x[7], p1, x[6] * p0));
                    ct = fmaf(lc[3 * b + 2], p2, fmaf(lc[3 * b + 1], p1, lc[3 * b] * p0)) + tp;
                    if (valid) { ab[9 * b] = c0; ab[9 * b + 3] = c1; ab[9 * b + 6] = c2; }
                    TS[(3 * b + jc) * 32 + lane] = ct;
                }
            }
        }
        fence_async_smem();
        __syncthreads();
        if (tma) {
            if (a.abs_rot) {
                bulk_s2g(a.abs_rot + (size_t)m0 * B9, X, (uint32_t)(np * B9) * 4u);
                bulk_commit();
                bulk_wait_read_1();   // the rel_rot store (the older group) has read R: the region is free
            } else {
                bulk_wait_read_all();
            }
        }
        __syncthreads();
        // ---- abs_loc and projection records (AoS) over the dead rel_rot region, one bone per warp task
        float* T3 = R;
        float* PJ = R + 32 * B3;
        if (a.abs_loc || a.proj) {
            const int pl = min(lane, np - 1);
            World w;
            if (a.proj) w = load_world(a, m0 + pl);
#pragma unroll 1
            for (int b = warp; b < B; b += FW) {
                const V3 t{TS[(3 * b) * 32 + lane], TS[(3 * b + 1) * 32 + lane], TS[(3 * b + 2) * 32 + lane]};
                if (lane < np) {
                    sts3(T3 + pl * B3, b, t);
                    if (a.proj) sts3(PJ + pl * B3, b, project_pt(a.cam, w, t, nullptr));
                }
            }
            fence_async_smem();
            __syncthreads();
            if (tma) {
                if (a.abs_loc) bulk_s2g(a.abs_loc + (size_t)m0 * B3, T3, (uint32_t)(np * B3) * 4u);
                if (a.proj) bulk_s2g(a.proj + (size_t)m0 * B3, PJ, (uint32_t)(np * B3) * 4u);
                bulk_commit();
            }
        }
        if (sub + 1 < a.nsub) {   // the next tile of this clip overwrites every region: all stores must have read them
            if (tma) bulk_wait_read_all();
            __syncthreads();
        }
    }
    if (tma) bulk_wait_all();
}

// ------------------------------------------------------------------------------------------------ backward
// Exact chain rule of the forward above (row vectors):  A_b = R_b A_p,  t_b = l_b A_p + t_p,  R[t] = C[t] R[t-1]:
//   G_t[b]  = g_abs_loc[b] + projection backward of g_proj[b]          G_A[b] = g_abs_rot[b]
//   leaves -> root:   G_A[p] += R_b^T G_A[b] + l_b^T (x) G_t[b],   G_t[p] += G_t[b]
//   G_R[b]  = G_A[b] A_p^T + g_rel_rot[b]
//   frames, last -> first:   Gtot[t] = G_R[t] + C[t+1]^T Gtot[t+1],   G_C[t] = Gtot[t] R[t-1]^T   (R[-1] = reference)
// shared memory (floats): RR 32*B9 (saved rel_rot) | AA 32*B9 (upstream g_abs_loc / g_proj + saved abs_loc records ->
// saved abs_rot -> the input changes C) | GA 32*B9 (G_A -> G_R -> Gtot -> G_C, all in place) | GT 3B*32 (G_t, SoA) |
// LOC rows*3B | CARRY B9 | PREV B9 | ROW 32 | barriers
template <int BT>
__global__ void __launch_bounds__(FT, 2) fk_tile_bwd_kernel(const FkTile a) {
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int B = BT ? BT : a.B, B3 = B * 3, B9 = B * 9, L = a.L;
    float* RR = smem;
    float* AA = RR + 32 * B9;
    float* GA = AA + 32 * B9;
    float* GT = GA + 32 * B9;
    float* LOC = GT + 32 * B3;
    float* CARRY = LOC + a.loc_rows * B3;
    float* PREV = CARRY + B9;
    int* ROW = reinterpret_cast<int*>(PREV + B9);
    uint64_t* bars = reinterpret_cast<uint64_t*>(ROW + 32);   // [0] rel_rot + upstream records, [1] abs_rot, [2] changes
    const bool tma = tid == FT - 32;
    if (tma) {
        mbar_init(bars, 1);
        mbar_init(bars + 1, 1);
        mbar_init(bars + 2, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const float* rel_saved = a.scan ? a.s_rel_rot : a.pose_in;   // relative_rot mode: rel_rot IS the input
    for (int it = 0; it < a.nsub; ++it) {
        const int sub = a.nsub - 1 - it;   // the later tiles of a clip first: the frame recursion runs backwards
        const TileSpan sp = tile_span(a, sub);
        const long m0 = sp.m0;
        const int np = sp.np;
        const uint32_t ph = it & 1u;
        float* S3 = AA;   // three 32*B3 record regions: g_abs_loc, g_proj, saved abs_loc
        if (tma) {
            const uint32_t b9 = (uint32_t)(np * B9) * 4u, b3 = (uint32_t)(np * B3) * 4u;
            uint32_t bytes = b9 + (a.g_abs_rot ? b9 : 0u) + (a.g_abs_loc ? b3 : 0u) + (a.g_proj ? 2u * b3 : 0u);
            mbar_expect_tx(bars, bytes);
            bulk_g2s(RR, rel_saved + (size_t)m0 * B9, b9, bars);
            if (a.g_abs_rot) bulk_g2s(GA, a.g_abs_rot + (size_t)m0 * B9, b9, bars);
            if (a.g_abs_loc) bulk_g2s(S3, a.g_abs_loc + (size_t)m0 * B3, b3, bars);
            if (a.g_proj) {
                bulk_g2s(S3 + 32 * B3, a.g_proj + (size_t)m0 * B3, b3, bars);
                bulk_g2s(S3 + 64 * B3, a.s_abs_loc + (size_t)m0 * B3, b3, bars);
            }
        }
        const long clip0 = m0 / L;
        const int nclip = (int)((m0 + np - 1) / L - clip0) + 1;
        if (warp == 1) {
            const long clip = (m0 + min(lane, np - 1)) / L;
            ROW[lane] = a.ref_index ? __ldg(a.ref_index + clip) : (int)clip;
        }
        for (int i = tid; i < nclip * B3; i += FT) {
            const int cl = i / B3, k = i - cl * B3;
            const long clip = clip0 + cl;
            const long row = a.ref_index ? (long)__ldg(a.ref_index + clip) : clip;
            LOC[i] = __ldg(a.ref_loc + row * B3 + k);
        }
        if (a.scan && sub > 0) {   // R[t-1] of the tile's first frame lives in the previous tile of the clip
            for (int i = tid; i < B9; i += FT) PREV[i] = __ldg(rel_saved + (size_t)(m0 - 1) * B9 + i);
        }
        if (!a.g_abs_rot) {
            for (int i = tid; i < 32 * B9; i += FT) GA[i] = 0.f;
        }
        while (!mbar_try_wait(bars, ph)) {}
        __syncthreads();
        // ---- G_t = g_abs_loc + projection backward, SoA rows; one bone per warp task
        {
            const int pl = min(lane, np - 1);
            World w;
            if (a.g_proj) w = load_world(a, m0 + pl);
#pragma unroll 1
            for (int b = warp; b < B; b += FW) {
                V3 g{0.f, 0.f, 0.f};
                if (a.g_abs_loc) g = lds3(S3 + pl * B3, b);
                if (a.g_proj) {
                    const V3 gp = lds3(S3 + 32 * B3 + pl * B3, b), t = lds3(S3 + 64 * B3 + pl * B3, b);
                    const V3 q = project_pt_bwd(a.cam, w, t, gp);
                    g.x += q.x; g.y += q.y; g.z += q.z;
                }
                if (lane >= np) g = V3{0.f, 0.f, 0.f};
                GT[(3 * b) * 32 + lane] = g.x; GT[(3 * b + 1) * 32 + lane] = g.y; GT[(3 * b + 2) * 32 + lane] = g.z;
            }
        }
        fence_async_smem();
        __syncthreads();
        if (tma) {
            const uint32_t b9 = (uint32_t)(np * B9) * 4u;
            mbar_expect_tx(bars + 1, b9);
            bulk_g2s(AA, a.s_abs_rot + (size_t)m0 * B9, b9, bars + 1);
        }
        // ---- leaves -> root, one warp per column j of G_A (a column only needs the same column of the children) and
        //      component j of G_t:  G_A[p][:,j] += R_b^T G_A[b][:,j] + l_b G_t[b][j],  G_t[p][j] += G_t[b][j]
        if (warp < 3 && lane < np) {   // (lanes past the tile own no record)
            const int jc = warp;
            const int pl = lane;
            const float* rb = RR + pl * B9;
            float* ga = GA + pl * B9 + jc;
            const float* lc = LOC + (int)((m0 + pl) / L - clip0) * B3;
#pragma unroll 1
            for (int b = B - 1; b >= 1; --b) {
                const int par = BT == 26 ? carla_parent(b) : a.parent[b];
                float x[9];
                rec_ld(rb, b, x);
                const float g0 = ga[9 * b], g1 = ga[9 * b + 3], g2 = ga[9 * b + 6];
                const float gt = GT[(3 * b + jc) * 32 + lane];
                // (R_b^T g)[i] = sum_k R_b[k][i] g[k]
                const float c0 = fmaf(x[6], g2, fmaf(x[3], g1, x[0] * g0)) + lc[3 * b] * gt;
                const float c1 = fmaf(x[7], g2, fmaf(x[4], g1, x[1] * g0)) + lc[3 * b + 1] * gt;
                const float c2 = fmaf(x[8], g2, fmaf(x[5], g1, x[2] * g0)) + lc[3 * b + 2] * gt;
                ga[9 * par] += c0; ga[9 * par + 3] += c1; ga[9 * par + 6] += c2;
                GT[(3 * par + jc) * 32 + lane] += gt;
            }
        }
        while (!mbar_try_wait(bars + 1, ph)) {}
        __syncthreads();
        // ---- G_R[b] = G_A[b] A_parent^T (+ upstream on rel_rot), in place; one bone per warp task
        {
            const int pl = min(lane, np - 1);
#pragma unroll 1
            for (int b = warp; b < B && lane < np; b += FW) {
                const int par = BT == 26 ? carla_parent(b) : a.parent[b];
                float g[9], r[9];
                rec_ld(GA + pl * B9, b, g);
                if (par >= 0) {
                    float p[9];
                    rec_ld(AA + pl * B9, par, p);
#pragma unroll
                    for (int i = 0; i < 3; ++i)
#pragma unroll
                        for (int k = 0; k < 3; ++k)
                            r[3 * i + k] = fmaf(g[3 * i + 2], p[3 * k + 2], fmaf(g[3 * i + 1], p[3 * k + 1], g[3 * i] * p[3 * k]));
                } else {
#pragma unroll
                    for (int i = 0; i < 9; ++i) r[i] = g[i];
                }
                if (a.g_rel_rot) {   // (rare: a loss on relative_pose_rot itself) straight from global
                    const float* u = a.g_rel_rot + ((size_t)(m0 + pl) * B + b) * 9;
#pragma unroll
                    for (int i = 0; i < 9; ++i) r[i] += __ldg(u + i);
                }
                rec_st(GA + pl * B9, b, r);
            }
        }
        fence_async_smem();
        __syncthreads();
        if (a.scan) {
            if (tma) {
                const uint32_t b9 = (uint32_t)(np * B9) * 4u;
                mbar_expect_tx(bars + 2, b9);
                bulk_g2s(AA, a.pose_in + (size_t)m0 * B9, b9, bars + 2);   // the changes C[t] over the dead abs_rot records
            }
            while (!mbar_try_wait(bars + 2, ph)) {}
            __syncthreads();
            // ---- Gtot[t] = G_R[t] + C[t+1]^T Gtot[t+1], last frame first; one thread per (clip, bone, column)
            {
                const int items = sp.clips * B3;
#pragma unroll 1
                for (int idx = tid; idx < items; idx += FT) {
                    const int cl = idx / B3, rem = idx - cl * B3, b = rem / 3, j = rem - b * 3;
                    float* gq = GA + (cl * sp.lt + sp.lt - 1) * B9 + b * 9 + j;
                    const float* cq = AA + (cl * sp.lt + sp.lt - 1) * B9 + b * 9;
                    float h0 = 0.f, h1 = 0.f, h2 = 0.f;   // (C[t+1]^T Gtot[t+1])[:, j]
                    if (it > 0) { h0 = CARRY[b * 9 + j]; h1 = CARRY[b * 9 + 3 + j]; h2 = CARRY[b * 9 + 6 + j]; }
#pragma unroll 1
                    for (int f = sp.lt - 1; f >= 0; --f) {
                        const float g0 = gq[0] + h0, g1 = gq[3] + h1, g2 = gq[6] + h2;
                        gq[0] = g0; gq[3] = g1; gq[6] = g2;
                        float c[9];
#pragma unroll
                        for (int i = 0; i < 9; ++i) c[i] = cq[i];
                        h0 = fmaf(c[6], g2, fmaf(c[3], g1, c[0] * g0));
                        h1 = fmaf(c[7], g2, fmaf(c[4], g1, c[1] * g0));
                        h2 = fmaf(c[8], g2, fmaf(c[5], g1, c[2] * g0));
                        gq -= B9;
                        cq -= B9;
                    }
                    if (a.nsub > 1) { CARRY[b * 9 + j] = h0; CARRY[b * 9 + 3 + j] = h1; CARRY[b * 9 + 6 + j] = h2; }
                }
            }
            __syncthreads();
            // ---- G_C[t] = Gtot[t] R[t-1]^T in place; one bone per warp task; R[t-1] is the neighbouring lane's record,
            //      the previous tile's last (PREV) or the clip's reference rotation
            {
                const int pl = min(lane, np - 1);
                const int fr = pl % sp.lt;   // frame inside the tile's part of its clip
#pragma unroll 1
                for (int b = warp; b < B && lane < np; b += FW) {
                    float g[9], p[9], r[9];
                    rec_ld(GA + pl * B9, b, g);
                    if (fr > 0) {
                        rec_ld(RR + (pl - 1) * B9, b, p);
                    } else if (sub > 0) {
#pragma unroll
                        for (int i = 0; i < 9; ++i) p[i] = PREV[b * 9 + i];
                    } else {
                        const float* rr = a.ref_rot + ((size_t)ROW[pl] * B + b) * 9;
#pragma unroll
                        for (int i = 0; i < 9; ++i) p[i] = __ldg(rr + i);
                    }
#pragma unroll
                    for (int i = 0; i < 3; ++i)
#pragma unroll
                        for (int k = 0; k < 3; ++k)
                            r[3 * i + k] = fmaf(g[3 * i + 2], p[3 * k + 2], fmaf(g[3 * i + 1], p[3 * k + 1], g[3 * i] * p[3 * k]));
                    rec_st(GA + pl * B9, b, r);
                }
            }
            fence_async_smem();
            __syncthreads();
        }
        if (tma) {
            bulk_s2g(a.g_pose_in + (size_t)m0 * B9, GA, (uint32_t)(np * B9) * 4u);
            bulk_commit();
            if (it + 1 < a.nsub) bulk_wait_read_all();
        }
        if (it + 1 < a.nsub) __syncthreads();
    }
    if (tma) bulk_wait_all();
}

bool carla_tree(const p2c_tree_t& t) {
    if (t.n_bones != 26) return false;
    for (int b = 0; b < 26; ++b)
        if (t.parent[b] != carla_parent(b)) return false;
    return true;
}

struct FkPlan {
    int ok, grid;
    size_t smem;
};

// fills the geometry fields of `a`; ok = 0: the tile kernels decline (odd record sizes, unaligned tensors, odd bone
// count...) and the generic warp-per-clip kernels of fk_kernels.cu run instead
FkPlan fk_plan(FkTile& a, const p2c_tree_t& tree, int n_clips, int clip_len, int mode, bool backward,
               std::initializer_list<const void*> tensors) {
    FkPlan pl{};
    const int B = tree.n_bones;
    if (B < 2 || B > P2C_MAX_BONES || (B & 1)) return pl;   // AoS records are read with 64-bit accesses
    if (tree.parent[0] != -1) return pl;
    for (int b = 1; b < B; ++b)
        if (tree.parent[b] < 0 || tree.parent[b] >= b) return pl;
    for (int b = 0; b < B; ++b) a.parent[b] = (int8_t)tree.parent[b];
    const long total = (long)n_clips * clip_len;
    if (total <= 0 || total > (1L << 26)) return pl;
    a.N = n_clips; a.L = clip_len; a.B = B; a.scan = mode == P2C_MODE_POSE_CHANGES;
    if (a.scan && clip_len > 32) {
        a.P = 0;
        a.nsub = (clip_len + 31) / 32;
        pl.grid = n_clips;
    } else {
        a.P = a.scan ? (32 / clip_len) * clip_len : 32;
        if ((long)a.P > total) a.P = (int)total;
        a.nsub = 1;
        pl.grid = (int)((total + a.P - 1) / a.P);
    }
    // every tile moves by TMA: 16-byte aligned offsets and sizes for the 9- and 3-word bone records
    auto ok16 = [B](long poses) { return (poses * B * 9) % 4 == 0 && (poses * B * 3) % 4 == 0; };
    if (!ok16(total)) return pl;
    if (a.nsub > 1 ? !(ok16(clip_len) && ok16(32)) : !ok16(a.P)) return pl;
    for (const void* q : tensors)
        if (reinterpret_cast<uintptr_t>(q) & 15) return pl;
    a.loc_rows = a.scan ? (clip_len > 32 ? 1 : 32 / clip_len) : (clip_len >= 32 ? 2 : 32 / clip_len + 2);
    if (a.loc_rows > 33) a.loc_rows = 33;
    const size_t floats = (size_t)(backward ? 3 : 2) * 32 * B * 9 + 32 * B * 3 + (size_t)a.loc_rows * B * 3 +
                          (backward ? 2 : 1) * B * 9 + 32 + 8;
    pl.smem = floats * sizeof(float);
    if (pl.smem > 200 * 1024) return pl;
    pl.ok = 1;
    return pl;
}

}  // namespace

// called by p2c_fk_fwd / p2c_fk_bwd (fk_kernels.cu): 1 = launched, 0 = declined, anything else = an error code
extern "C" int p2c_fk_tile_fwd(const p2c_fk_fwd_t* p, void* stream, int* rc) {
    *rc = 0;
    if (p2c_env().disable_fwd_tile) return 0;
    FkTile a{};
    const FkPlan pl = fk_plan(a, p->tree, p->n_clips, p->clip_len, p->mode, false,
                              {p->pose_in, p->rel_rot, p->abs_rot, p->abs_loc, p->proj});
    if (!pl.ok) return 0;
    a.pose_in = p->pose_in; a.ref_loc = p->ref_loc; a.ref_rot = p->ref_rot; a.ref_index = p->ref_index;
    a.world_loc = p->proj ? p->world_loc : nullptr; a.world_rot = p->proj ? p->world_rot : nullptr; a.cam = p->camera;
    a.rel_rot = p->rel_rot; a.abs_rot = p->abs_rot; a.abs_loc = p->abs_loc; a.proj = p->proj;
    typedef void (*kern_t)(const FkTile);
    const kern_t k = carla_tree(p->tree) ? fk_tile_fwd_kernel<26> : fk_tile_fwd_kernel<0>;
    if ((*rc = p2c_allow_big_smem(reinterpret_cast<const void*>(k), 200 * 1024))) return -1;
    k<<<pl.grid, FT, pl.smem, (cudaStream_t)stream>>>(a);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { *rc = (int)e; return -1; }
    return 1;
}

extern "C" int p2c_fk_tile_bwd(const p2c_fk_bwd_t* p, void* stream, int* rc) {
    *rc = 0;
    if (p2c_env().disable_fwd_tile) return 0;
    if (p->g_ref_rot || p->g_ref_loc || p->g_world_loc || p->g_world_rot) return 0;   // (the generic kernel has them)
    FkTile a{};
    const FkPlan pl = fk_plan(a, p->tree, p->n_clips, p->clip_len, p->mode, true,
                              {p->pose_in, p->rel_rot, p->abs_rot, p->abs_loc, p->g_abs_rot, p->g_abs_loc, p->g_proj,
                               p->g_pose_in});
    if (!pl.ok) return 0;
    a.pose_in = p->pose_in; a.ref_loc = p->ref_loc; a.ref_rot = p->ref_rot; a.ref_index = p->ref_index;
    a.world_loc = p->g_proj ? p->world_loc : nullptr; a.world_rot = p->g_proj ? p->world_rot : nullptr; a.cam = p->camera;
    a.s_rel_rot = p->rel_rot; a.s_abs_rot = p->abs_rot; a.s_abs_loc = p->abs_loc;
    a.g_rel_rot = p->g_rel_rot; a.g_abs_rot = p->g_abs_rot; a.g_abs_loc = p->g_abs_loc; a.g_proj = p->g_proj;
    a.g_pose_in = p->g_pose_in;
    typedef void (*kern_t)(const FkTile);
    const kern_t k = carla_tree(p->tree) ? fk_tile_bwd_kernel<26> : fk_tile_bwd_kernel<0>;
    if ((*rc = p2c_allow_big_smem(reinterpret_cast<const void*>(k), 200 * 1024))) return -1;
    k<<<pl.grid, FT, pl.smem, (cudaStream_t)stream>>>(a);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { *rc = (int)e; return -1; }
    return 1;
}
