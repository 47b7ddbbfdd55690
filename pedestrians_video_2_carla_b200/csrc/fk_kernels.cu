// Generic-tree forward kinematics + projection, forward and backward (any tree <= 32 bones).
//
// Mapping: one warp per clip (pose_changes: frames are a sequential 3x3 product scan per bone) or per
// pose (relative_rot: frames independent); lane b = bone b.  The tree is walked level by level, each
// lane pulling its parent's 12 floats (R_abs, t_abs) with warp shuffles; poses move between HBM and the
// lanes as coalesced float2 tiles staged in a per-warp shared-memory buffer.
//
// Replaces: walker_control/p3d_pose.py:98-184 (__move_to_relative, __transform_descendants,
// relative_to_absolute, forward), the frame loops of modules/layers/projection.py:144-195 and
// walker_control/p3d_pose_projection.py:115-152, plus torch autograd's backward through them.
#include "common.cuh"

namespace {

constexpr int WARPS = 4;
constexpr int TILE = P2C_MAX_BONES * 9;  // floats per warp staging tile

struct FkArgs {
    TreeDev tree;
    int N, L;
    const float* pose_in;
    const float* ref_loc;
    const float* ref_rot;
    const int32_t* ref_index;
    const float* world_loc;
    const float* world_rot;
    p2c_camera_t cam;
    // forward outputs / saved tensors for backward
    float* rel_rot;
    float* abs_rot;
    float* abs_loc;
    float* proj;
    // backward only
    const float* g_rel_rot;
    const float* g_abs_rot;
    const float* g_abs_loc;
    const float* g_proj;
    float* g_pose_in;
    float* g_ref_rot;
    float* g_ref_loc;
    float* g_world_loc;
    float* g_world_rot;
};

struct WorldT {
    M3 rot;
    V3 loc;
};

__device__ __forceinline__ WorldT load_world(const FkArgs& a, long m) {
    WorldT w;
    w.rot = m3_identity();
    w.loc = V3{0.f, 0.f, 0.f};
    if (a.world_rot) w.rot = m3_from(a.world_rot + m * 9);
    if (a.world_loc) w.loc = V3{a.world_loc[m * 3], a.world_loc[m * 3 + 1], a.world_loc[m * 3 + 2]};
    return w;
}

// p3d_pose_projection.py:137-152: (x,y,z) -> (y,-x,z) -> . world_rot + world_loc -> camera -> pixels
__device__ __forceinline__ V3 project_point(const V3& t, const WorldT& w, const p2c_camera_t& cam, V3* cam_pt) {
    V3 wx{t.y, -t.x, t.z};
    V3 wp = v3_mul_m(wx, w.rot);
    wp.x += w.loc.x; wp.y += w.loc.y; wp.z += w.loc.z;
    M3 R = m3_from(cam.R);
    V3 c = v3_mul_m(wp, R);
    c.x += cam.T[0]; c.y += cam.T[1]; c.z += cam.T[2];
    if (cam_pt) *cam_pt = c;
    float iz = 1.0f / c.z;
    return V3{cam.px - cam.f * c.x * iz, cam.py - cam.f * c.y * iz, iz};
}

// backward of project_point: upstream (gu, gv, gq) -> d/dt ; optionally d/d world (per point, unreduced)
__device__ __forceinline__ V3 project_point_bwd(const V3& t, const WorldT& w, const p2c_camera_t& cam, const V3& g,
                                                V3* g_wp_out, V3* wx_out) {
    V3 c;
    (void)project_point(t, w, cam, &c);
    float iz = 1.0f / c.z;
    V3 gc;
    gc.x = -cam.f * g.x * iz;
    gc.y = -cam.f * g.y * iz;
    gc.z = (cam.f * (g.x * c.x + g.y * c.y) - g.z) * iz * iz;
    M3 R = m3_from(cam.R);
    V3 gwp = v3_mul_mt(gc, R);
    V3 gwx = v3_mul_mt(gwp, w.rot);
    if (g_wp_out) *g_wp_out = gwp;
    if (wx_out) *wx_out = V3{t.y, -t.x, t.z};
    return V3{-gwx.y, gwx.x, gwx.z};
}

template <bool SCAN>
__global__ void __launch_bounds__(WARPS * 32) fk_fwd_kernel(const FkArgs a) {
    __shared__ __align__(16) float smem[WARPS][TILE];
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    float* s = smem[wib];
    const int B = a.tree.n_bones;
    const bool act = lane < B;
    const int par = act ? a.tree.parent[lane] : -1;
    const int dep = act ? a.tree.depth[lane] : -1;
    const int psrc = par < 0 ? 0 : par;
    const long units = SCAN ? (long)a.N : (long)a.N * a.L;
    const long wstride = (long)gridDim.x * WARPS;
    for (long unit = (long)blockIdx.x * WARPS + wib; unit < units; unit += wstride) {
        const long n = SCAN ? unit : unit / a.L;
        const int t0 = SCAN ? 0 : (int)(unit % a.L);
        const int nt = SCAN ? a.L : 1;
        const long row = a.ref_index ? (long)a.ref_index[n] : n;
        V3 l{0.f, 0.f, 0.f};
        M3 R = m3_identity();
        if (act) {
            const float* rl = a.ref_loc + (row * B + lane) * 3;
            l = V3{__ldg(rl), __ldg(rl + 1), __ldg(rl + 2)};
            if (SCAN) {
                const float* rr = a.ref_rot + (row * B + lane) * 9;
#pragma unroll
                for (int i = 0; i < 9; ++i) R.m[i] = __ldg(rr + i);
            }
        }
        for (int t = t0; t < t0 + nt; ++t) {
            const long m = n * a.L + t;
            warp_tile_load(s, a.pose_in + m * B * 9, B * 9, lane);
            __syncwarp();
            M3 C = m3_identity();
            if (act) C = m3_from(s + lane * 9);
            __syncwarp();
            R = SCAN ? m3_mul(C, R) : C;
            if (a.rel_rot) {
                if (act) m3_to(s + lane * 9, R);
                __syncwarp();
                warp_tile_store(a.rel_rot + m * B * 9, s, B * 9, lane);
                __syncwarp();
            }
            // level-order walk: abs_rot[b] = rel_rot[b] . abs_rot[parent]; abs_loc[b] = l[b] . abs_rot[parent] + abs_loc[parent]
            M3 A = R;
            V3 tl = l;
            for (int d = 1; d <= a.tree.max_depth; ++d) {
                M3 Ap = m3_shfl(A, psrc);
                V3 tp = v3_shfl(tl, psrc);
                if (dep == d) {
                    A = m3_mul(R, Ap);
                    V3 q = v3_mul_m(l, Ap);
                    tl = V3{q.x + tp.x, q.y + tp.y, q.z + tp.z};
                }
            }
            if (a.abs_rot) {
                if (act) m3_to(s + lane * 9, A);
                __syncwarp();
                warp_tile_store(a.abs_rot + m * B * 9, s, B * 9, lane);
                __syncwarp();
            }
            if (a.abs_loc) {
                if (act) { s[lane * 3] = tl.x; s[lane * 3 + 1] = tl.y; s[lane * 3 + 2] = tl.z; }
                __syncwarp();
                warp_tile_store(a.abs_loc + m * B * 3, s, B * 3, lane);
                __syncwarp();
            }
            if (a.proj) {
                WorldT w = load_world(a, m);
                V3 p = project_point(tl, w, a.cam, nullptr);
                if (act) { s[lane * 3] = p.x; s[lane * 3 + 1] = p.y; s[lane * 3 + 2] = p.z; }
                __syncwarp();
                warp_tile_store(a.proj + m * B * 3, s, B * 3, lane);
                __syncwarp();
            }
        }
    }
}

__device__ __forceinline__ M3 tile_m3(float* s, const float* g, int B, int lane, bool act) {
    M3 r = m3_zero();
    if (g) {
        warp_tile_load(s, g, B * 9, lane);
        __syncwarp();
        if (act) r = m3_from(s + lane * 9);
        __syncwarp();
    }
    return r;
}
__device__ __forceinline__ V3 tile_v3(float* s, const float* g, int B, int lane, bool act) {
    V3 r{0.f, 0.f, 0.f};
    if (g) {
        warp_tile_load(s, g, B * 3, lane);
        __syncwarp();
        if (act) r = V3{s[lane * 3], s[lane * 3 + 1], s[lane * 3 + 2]};
        __syncwarp();
    }
    return r;
}

template <bool SCAN>
__global__ void __launch_bounds__(WARPS * 32) fk_bwd_kernel(const FkArgs a) {
    __shared__ __align__(16) float smem[WARPS][TILE];
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    float* s = smem[wib];
    const int B = a.tree.n_bones;
    const bool act = lane < B;
    const int par = act ? a.tree.parent[lane] : -1;
    const int dep = act ? a.tree.depth[lane] : -1;
    const int psrc = par < 0 ? 0 : par;
    int child[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) child[k] = act ? a.tree.child[lane][k] : -1;
    const float* rel_saved = a.rel_rot ? a.rel_rot : a.pose_in;  // mode 1: rel_rot == pose_in
    const long units = SCAN ? (long)a.N : (long)a.N * a.L;
    const long wstride = (long)gridDim.x * WARPS;
    for (long unit = (long)blockIdx.x * WARPS + wib; unit < units; unit += wstride) {
        const long n = SCAN ? unit : unit / a.L;
        const int t0 = SCAN ? 0 : (int)(unit % a.L);
        const int nt = SCAN ? a.L : 1;
        const long row = a.ref_index ? (long)a.ref_index[n] : n;
        V3 l{0.f, 0.f, 0.f};
        if (act) {
            const float* rl = a.ref_loc + (row * B + lane) * 3;
            l = V3{__ldg(rl), __ldg(rl + 1), __ldg(rl + 2)};
        }
        M3 Gcarry = m3_zero();
        V3 Gl{0.f, 0.f, 0.f};
        M3 Rcur = tile_m3(s, rel_saved + (n * a.L + t0 + nt - 1) * B * 9, B, lane, act);
        for (int t = t0 + nt - 1; t >= t0; --t) {
            const long m = n * a.L + t;
            const M3 R = Rcur;
            M3 A = tile_m3(s, a.abs_rot + m * B * 9, B, lane, act);
            M3 Ap = m3_shfl(A, psrc);
            if (par < 0) Ap = m3_identity();
            M3 GA = tile_m3(s, a.g_abs_rot ? a.g_abs_rot + m * B * 9 : nullptr, B, lane, act);
            V3 Gt = tile_v3(s, a.g_abs_loc ? a.g_abs_loc + m * B * 3 : nullptr, B, lane, act);
            if (a.g_proj) {
                V3 gp = tile_v3(s, a.g_proj + m * B * 3, B, lane, act);
                V3 tl = tile_v3(s, a.abs_loc + m * B * 3, B, lane, act);
                WorldT w = load_world(a, m);
                V3 gwp, wx;
                V3 gt = project_point_bwd(tl, w, a.cam, gp, &gwp, &wx);
                if (!act) { gt = V3{0.f, 0.f, 0.f}; gwp = gt; wx = gt; }
                Gt.x += gt.x; Gt.y += gt.y; Gt.z += gt.z;
                if (a.g_world_loc || a.g_world_rot) {  // world_pos = wx . Wrot + Wloc, summed over bones
                    float gl[3] = {warp_sum(gwp.x), warp_sum(gwp.y), warp_sum(gwp.z)};
                    M3 o = m3_zero();
                    m3_acc_outer(o, wx, gwp);
#pragma unroll
                    for (int i = 0; i < 9; ++i) o.m[i] = warp_sum(o.m[i]);
                    if (lane == 0) {
                        if (a.g_world_loc)
                            for (int i = 0; i < 3; ++i) a.g_world_loc[m * 3 + i] = gl[i];
                        if (a.g_world_rot)
                            for (int i = 0; i < 9; ++i) a.g_world_rot[m * 9 + i] = o.m[i];
                    }
                }
            }
            // leaves -> root: G_A[parent] += R^T . G_A + l^T (x) G_t ; G_t[parent] += G_t
            for (int d = a.tree.max_depth; d >= 1; --d) {
                M3 CA = m3_mul_at(R, GA);
                m3_acc_outer(CA, l, Gt);
                for (int k = 0; k < a.tree.max_children; ++k) {
                    const int src = child[k] < 0 ? 0 : child[k];
                    M3 ca = m3_shfl(CA, src);
                    V3 ct = v3_shfl(Gt, src);
                    if (child[k] >= 0 && dep == d - 1) {
                        m3_acc(GA, ca);
                        Gt.x += ct.x; Gt.y += ct.y; Gt.z += ct.z;
                    }
                }
            }
            M3 Gtot = m3_mul_bt(GA, Ap);  // d/d rel_rot[b] = G_A . abs_rot[parent]^T
            if (a.g_ref_loc) {
                V3 q = v3_mul_mt(Gt, Ap);
                Gl.x += q.x; Gl.y += q.y; Gl.z += q.z;
            }
            if (a.g_rel_rot) {
                M3 gr = tile_m3(s, a.g_rel_rot + m * B * 9, B, lane, act);
                m3_acc(Gtot, gr);
            }
            if (SCAN) {
                m3_acc(Gtot, Gcarry);
                M3 prev;
                if (t > 0) {
                    prev = tile_m3(s, rel_saved + (m - 1) * B * 9, B, lane, act);
                } else {
                    prev = m3_zero();
                    if (act) {
                        const float* rr = a.ref_rot + (row * B + lane) * 9;
#pragma unroll
                        for (int i = 0; i < 9; ++i) prev.m[i] = __ldg(rr + i);
                    }
                }
                M3 C = tile_m3(s, a.pose_in + m * B * 9, B, lane, act);
                M3 GC = m3_mul_bt(Gtot, prev);  // rel_rot[t] = C . rel_rot[t-1]
                Gcarry = m3_mul_at(C, Gtot);
                Rcur = prev;
                if (act) m3_to(s + lane * 9, GC);
            } else {
                if (act) m3_to(s + lane * 9, Gtot);
            }
            __syncwarp();
            warp_tile_store(a.g_pose_in + m * B * 9, s, B * 9, lane);
            __syncwarp();
        }
        if (SCAN && a.g_ref_rot && act) {
            float* o = a.g_ref_rot + (n * B + lane) * 9;
#pragma unroll
            for (int i = 0; i < 9; ++i) o[i] = Gcarry.m[i];
        }
        if (a.g_ref_loc && act) {
            float* o = a.g_ref_loc + (n * B + lane) * 3;
            o[0] = Gl.x; o[1] = Gl.y; o[2] = Gl.z;
        }
    }
}

// ---- P3dPoseProjection.forward alone: one thread per point -------------------------------------
__global__ void project_fwd_kernel(const float* __restrict__ x, const float* __restrict__ loc,
                                   const float* __restrict__ rot, const p2c_camera_t cam, float* __restrict__ out,
                                   long m_total, int J) {
    const long total = m_total * J;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long m = i / J;
        WorldT w;
        w.rot = m3_from(rot + m * 9);
        w.loc = V3{loc[m * 3], loc[m * 3 + 1], loc[m * 3 + 2]};
        V3 t{x[i * 3], x[i * 3 + 1], x[i * 3 + 2]};
        V3 p = project_point(t, w, cam, nullptr);
        out[i * 3] = p.x; out[i * 3 + 1] = p.y; out[i * 3 + 2] = p.z;
    }
}

// one warp per pose so the per-pose world gradients reduce with shuffles (deterministic)
__global__ void __launch_bounds__(WARPS * 32) project_bwd_kernel(const float* __restrict__ x,
                                                                const float* __restrict__ loc,
                                                                const float* __restrict__ rot, const p2c_camera_t cam,
                                                                const float* __restrict__ g_out, float* __restrict__ g_x,
                                                                float* __restrict__ g_loc, float* __restrict__ g_rot,
                                                                long m_total, int J) {
    const int lane = threadIdx.x & 31;
    const long wstride = (long)gridDim.x * WARPS;
    for (long m = (long)blockIdx.x * WARPS + (threadIdx.x >> 5); m < m_total; m += wstride) {
        WorldT w;
        w.rot = m3_from(rot + m * 9);
        w.loc = V3{loc[m * 3], loc[m * 3 + 1], loc[m * 3 + 2]};
        V3 sl{0.f, 0.f, 0.f};
        M3 sr = m3_zero();
        for (int j0 = 0; j0 < J; j0 += 32) {
            const int j = j0 + lane;
            if (j < J) {
                const long i = m * J + j;
                V3 t{x[i * 3], x[i * 3 + 1], x[i * 3 + 2]};
                V3 g{g_out[i * 3], g_out[i * 3 + 1], g_out[i * 3 + 2]};
                V3 gwp, wx;
                V3 gt = project_point_bwd(t, w, cam, g, &gwp, &wx);
                if (g_x) { g_x[i * 3] = gt.x; g_x[i * 3 + 1] = gt.y; g_x[i * 3 + 2] = gt.z; }
                sl.x += gwp.x; sl.y += gwp.y; sl.z += gwp.z;
                m3_acc_outer(sr, wx, gwp);
            }
        }
        if (g_loc) {
            float v[3] = {warp_sum(sl.x), warp_sum(sl.y), warp_sum(sl.z)};
            if (lane == 0)
                for (int i = 0; i < 3; ++i) g_loc[m * 3 + i] = v[i];
        }
        if (g_rot) {
#pragma unroll
            for (int i = 0; i < 9; ++i) sr.m[i] = warp_sum(sr.m[i]);
            if (lane == 0)
                for (int i = 0; i < 9; ++i) g_rot[m * 9 + i] = sr.m[i];
        }
    }
}

// ---- utils/world.py:16-63: one thread per clip, sequential over frames --------------------------
__global__ void world_scan_fwd_kernel(const float* __restrict__ dl, const float* __restrict__ dr,
                                      const float* __restrict__ il, const float* __restrict__ ir,
                                      float* __restrict__ wl, float* __restrict__ wr, int N, int L) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    V3 loc{0.f, 0.f, 0.f};
    M3 rot = m3_identity();
    if (il) loc = V3{il[n * 3], il[n * 3 + 1], il[n * 3 + 2]};
    if (ir) rot = m3_from(ir + (long)n * 9);
    for (int t = 0; t < L; ++t) {
        const long m = (long)n * L + t;
        if (dr) rot = m3_mul(rot, m3_from(dr + m * 9));
        if (dl) { loc.x += dl[m * 3]; loc.y += dl[m * 3 + 1]; loc.z += dl[m * 3 + 2]; }
        wl[m * 3] = loc.x; wl[m * 3 + 1] = loc.y; wl[m * 3 + 2] = loc.z;
        m3_to(wr + m * 9, rot);
    }
}

__global__ void world_scan_bwd_kernel(const float* __restrict__ dr, const float* __restrict__ wr,
                                      const float* __restrict__ ir, const float* __restrict__ g_wl,
                                      const float* __restrict__ g_wr, float* __restrict__ g_dl,
                                      float* __restrict__ g_dr, int N, int L) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    V3 cl{0.f, 0.f, 0.f};
    M3 cr = m3_zero();
    for (int t = L - 1; t >= 0; --t) {
        const long m = (long)n * L + t;
        if (g_wl) { cl.x += g_wl[m * 3]; cl.y += g_wl[m * 3 + 1]; cl.z += g_wl[m * 3 + 2]; }
        if (g_dl) { g_dl[m * 3] = cl.x; g_dl[m * 3 + 1] = cl.y; g_dl[m * 3 + 2] = cl.z; }
        if (g_dr) {
            if (g_wr) m3_acc(cr, m3_from(g_wr + m * 9));
            // world_rot[t] = world_rot[t-1] . dr[t]
            M3 prev = (t > 0) ? m3_from(wr + (m - 1) * 9) : (ir ? m3_from(ir + (long)n * 9) : m3_identity());
            M3 g = m3_mul_at(prev, cr);
            m3_to(g_dr + m * 9, g);
            cr = m3_mul_bt(cr, m3_from(dr + m * 9));
        }
    }
}

int grid_for_warps(long units) {
    long blocks = (units + WARPS - 1) / WARPS;
    const long cap = 148L * 16 * 8;  // persistent-ish: enough CTAs to fill 148 SMs many times over
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace

// the pose-per-lane tile kernels (fk_tile.cu): preferred whenever they take the call
extern "C" int p2c_fk_tile_fwd(const p2c_fk_fwd_t* p, void* stream, int* rc);
extern "C" int p2c_fk_tile_bwd(const p2c_fk_bwd_t* p, void* stream, int* rc);

extern "C" int p2c_fk_fwd(const p2c_fk_fwd_t* p, void* stream) {
    if (!p) return P2C_E_NULL;
    FkArgs a{};
    int rc = build_tree(p->tree, a.tree);
    if (rc) return rc;
    if (p->n_clips < 0 || p->clip_len < 1) return P2C_E_BADARG;
    if (p->mode != P2C_MODE_POSE_CHANGES && p->mode != P2C_MODE_RELATIVE_ROT) return P2C_E_BADARG;
    if (p->n_clips == 0) return 0;  // empty batch: nothing to read or write (empty tensors have NULL storage)
    if (!p->pose_in || !p->ref_loc) return P2C_E_NULL;
    if (p->mode == P2C_MODE_POSE_CHANGES && !p->ref_rot) return P2C_E_NULL;
    a.N = p->n_clips; a.L = p->clip_len;
    a.pose_in = p->pose_in; a.ref_loc = p->ref_loc; a.ref_rot = p->ref_rot; a.ref_index = p->ref_index;
    a.world_loc = p->world_loc; a.world_rot = p->world_rot; a.cam = p->camera;
    a.rel_rot = p->rel_rot; a.abs_rot = p->abs_rot; a.abs_loc = p->abs_loc; a.proj = p->proj;
    cudaStream_t st = (cudaStream_t)stream;
    {
        int trc = 0;
        const int took = p2c_fk_tile_fwd(p, stream, &trc);
        if (took < 0) return trc;
        if (took > 0) return 0;
    }
    if (p->mode == P2C_MODE_POSE_CHANGES)
        fk_fwd_kernel<true><<<grid_for_warps(a.N), WARPS * 32, 0, st>>>(a);
    else
        fk_fwd_kernel<false><<<grid_for_warps((long)a.N * a.L), WARPS * 32, 0, st>>>(a);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_fk_bwd(const p2c_fk_bwd_t* p, void* stream) {
    if (!p) return P2C_E_NULL;
    FkArgs a{};
    int rc = build_tree(p->tree, a.tree);
    if (rc) return rc;
    if (p->n_clips < 0 || p->clip_len < 1) return P2C_E_BADARG;
    if (p->mode != P2C_MODE_POSE_CHANGES && p->mode != P2C_MODE_RELATIVE_ROT) return P2C_E_BADARG;
    if (p->n_clips == 0) return 0;
    if (!p->pose_in || !p->ref_loc || !p->abs_rot || !p->g_pose_in) return P2C_E_NULL;
    if (p->mode == P2C_MODE_POSE_CHANGES && (!p->ref_rot || !p->rel_rot)) return P2C_E_NULL;
    if (p->g_proj && !p->abs_loc) return P2C_E_NULL;
    if ((p->g_world_loc || p->g_world_rot) && !p->g_proj) return P2C_E_BADARG;
    if ((p->g_ref_rot || p->g_ref_loc) && p->ref_index) return P2C_E_BADARG;
    if (p->g_ref_rot && p->mode != P2C_MODE_POSE_CHANGES) return P2C_E_BADARG;
    if (p->g_ref_loc && p->mode == P2C_MODE_RELATIVE_ROT && p->clip_len != 1) return P2C_E_UNSUPPORTED;
    if (p->n_clips == 0) return 0;
    a.N = p->n_clips; a.L = p->clip_len;
    a.pose_in = p->pose_in; a.ref_loc = p->ref_loc; a.ref_rot = p->ref_rot; a.ref_index = p->ref_index;
    a.world_loc = p->world_loc; a.world_rot = p->world_rot; a.cam = p->camera;
    a.rel_rot = const_cast<float*>(p->rel_rot); a.abs_rot = const_cast<float*>(p->abs_rot);
    a.abs_loc = const_cast<float*>(p->abs_loc);
    a.g_rel_rot = p->g_rel_rot; a.g_abs_rot = p->g_abs_rot; a.g_abs_loc = p->g_abs_loc; a.g_proj = p->g_proj;
    a.g_pose_in = p->g_pose_in; a.g_ref_rot = p->g_ref_rot; a.g_ref_loc = p->g_ref_loc;
    a.g_world_loc = p->g_world_loc; a.g_world_rot = p->g_world_rot;
    cudaStream_t st = (cudaStream_t)stream;
    {
        int trc = 0;
        const int took = p2c_fk_tile_bwd(p, stream, &trc);
        if (took < 0) return trc;
        if (took > 0) return 0;
    }
    if (p->mode == P2C_MODE_POSE_CHANGES)
        fk_bwd_kernel<true><<<grid_for_warps(a.N), WARPS * 32, 0, st>>>(a);
    else
        fk_bwd_kernel<false><<<grid_for_warps((long)a.N * a.L), WARPS * 32, 0, st>>>(a);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_project_fwd(const float* x, const float* loc, const float* rot, const p2c_camera_t* cam,
                               float* out, int64_t m, int32_t j, void* stream) {
    if (!x || !loc || !rot || !cam || !out) return P2C_E_NULL;
    if (m < 0 || j < 1) return P2C_E_BADARG;
    if (m == 0) return 0;
    long total = (long)m * j;
    long blocks = (total + 255) / 256;
    if (blocks > 148L * 32) blocks = 148L * 32;
    project_fwd_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(x, loc, rot, *cam, out, (long)m, j);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_project_bwd(const float* x, const float* loc, const float* rot, const p2c_camera_t* cam,
                               const float* g_out, float* g_x, float* g_loc, float* g_rot, int64_t m, int32_t j,
                               void* stream) {
    if (!x || !loc || !rot || !cam || !g_out) return P2C_E_NULL;
    if (m < 0 || j < 1) return P2C_E_BADARG;
    if (m == 0) return 0;
    project_bwd_kernel<<<grid_for_warps((long)m), WARPS * 32, 0, (cudaStream_t)stream>>>(x, loc, rot, *cam, g_out, g_x,
                                                                                        g_loc, g_rot, (long)m, j);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_world_scan_fwd(const float* loc_change, const float* rot_change, const float* init_loc,
                                  const float* init_rot, float* world_loc, float* world_rot, int32_t n_clips,
                                  int32_t clip_len, void* stream) {
    if (!world_loc || !world_rot) return P2C_E_NULL;
    if (n_clips < 0 || clip_len < 1) return P2C_E_BADARG;
    if (n_clips == 0) return 0;
    world_scan_fwd_kernel<<<(n_clips + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        loc_change, rot_change, init_loc, init_rot, world_loc, world_rot, n_clips, clip_len);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_world_scan_bwd(const float* rot_change, const float* world_rot, const float* init_rot,
                                  const float* g_world_loc, const float* g_world_rot, float* g_loc_change,
                                  float* g_rot_change, int32_t n_clips, int32_t clip_len, void* stream) {
    if (g_rot_change && (!rot_change || !world_rot)) return P2C_E_NULL;
    if (n_clips < 0 || clip_len < 1) return P2C_E_BADARG;
    if (n_clips == 0) return 0;
    world_scan_bwd_kernel<<<(n_clips + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        rot_change, world_rot, init_rot, g_world_loc, g_world_rot, g_loc_change, g_rot_change, n_clips, clip_len);
    P2C_LAUNCH_CHECK();
    return 0;
}
