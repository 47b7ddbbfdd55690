// p2c_fused_fast: the single-pass fused step for rotation_6d input (the pose_lifting training path):
//   rotation_6d_to_matrix (movements.py:113-116) -> relative-rotation scan over frames (layers/projection.py:170-195,
//   p3d_pose.py:98-114) -> FK over the bone tree (p3d_pose.py:116-184) -> P3dPoseProjection.forward
//   (p3d_pose_projection.py:115-152) -> Normalizer(HipsNeckExtractor) (normalizer.py:20-41) -> loc_2d / loc_3d /
//   loc_2d_3d (loss/loc_2d.py, loc_3d.py, loc_2d_3d.py) AND d(w2*loc_2d + w3*loc_3d)/d(6-D input), in ONE kernel
//   (plus a tiny mask-count pre-pass when mask_missing_joints is on).
//
// Mapping: one CTA per clip (persistent loop over clips), one warp per frame, lane b = bone b.
//   phase A  every warp: 6-D -> C[t] (Gram-Schmidt)                        -> smem Rs[t]
//   phase B  26 threads: R[t] = C[t] . R[t-1] serial over the clip's frames (in place), then rel_rot leaves as one
//            coalesced block copy
//   phase C  every warp, its frames: FK by pointer jumping over warp shuffles (ceil(log2(depth+1)) rounds instead
//            of one round per tree level), projection, normaliser, loss terms, loss gradient, and the tree
//            backward as a prefix sum over the DFS-ordered lanes                -> smem Ys[t] = R[t]^T dL/dR[t]
//   phase D  suffix sum of Ys over the frames (the reverse-time scan reduces to additions, see below)
//   phase E  every warp: dL/dC[t] = R[t] Z[t] R[t-1]^T, Gram-Schmidt backward -> grad wrt 6-D, coalesced store
//
// The backward uses that every matrix on this path is a rotation (C[t] comes out of Gram-Schmidt, the reference
// rotations out of Euler angles, products of rotations are rotations):
//   dL/dA_b   = A_b . M_b,   M_b = sum_{d in subtree(b)} (t_d - t_b)^T (x) g_d        (g_d = dL/d abs_loc[d])
//   Y_b[t]   := R_b[t]^T dL/dR_b[t]|direct = A_p . M_b . A_p^T                         (p = parent of b)
//   dL/dR[t]  (with the contributions of later frames) = R[t] . sum_{s >= t} Y[s]
// so subtree sums are differences of an inclusive prefix scan over lanes (the bone axis is a DFS pre-order, every
// subtree is a contiguous lane range) and the reverse-time product scan is a plain suffix sum.  Arbitrary
// (non-rotation) 3x3 input takes the exact generic path in fk_kernels.cu / fused_step.cu instead.
#include <stdlib.h>

#include "common.cuh"
#include "host_util.h"

namespace {

constexpr int MAX_ROUNDS = 5;
constexpr int IDENT_LANE = 31;  // holds the identity transform for lanes without an ancestor at the jump distance
constexpr int COUNT_BLOCKS = 296;

struct FastArgs {
    int N, L, GL, B, slot;  // clips, frames per clip, frames per CTA group, bones, B*9
    int scan;               // 1: pose_changes, 0: relative_rot
    int hips, neck, normalize, mask, gt_c, nrounds;
    int8_t parent[P2C_MAX_BONES];
    int8_t sub_end[P2C_MAX_BONES];
    int32_t anc_packed[P2C_MAX_BONES];  // 5 bits per round: lane of the 2^k-th ancestor, IDENT_LANE if none
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
    XchgArgs xchg;   // data-parallel ranks: where the finishing thread publishes the four loss sums
};

struct GS {
    V3 b1, b2, b3, a2;
    float inv1, inv2, s;
    bool deg1, deg2;
};

__device__ __forceinline__ float dot3(const V3& a, const V3& b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ V3 cross3(const V3& a, const V3& b) {
    return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

// pytorch3d rotation_6d_to_matrix; F.normalize(v) = v / max(|v|, 1e-12) written as v * rsqrt(max(|v|^2, 1e-24))
__device__ __forceinline__ GS gram_schmidt(const float* a) {
    GS r;
    const V3 a1{a[0], a[1], a[2]};
    r.a2 = V3{a[3], a[4], a[5]};
    const float n1 = dot3(a1, a1);
    r.deg1 = !(n1 > 1e-24f);
    r.inv1 = rsqrtf(fmaxf(n1, 1e-24f));
    r.b1 = V3{a1.x * r.inv1, a1.y * r.inv1, a1.z * r.inv1};
    r.s = dot3(r.b1, r.a2);
    const V3 u{fmaf(-r.s, r.b1.x, r.a2.x), fmaf(-r.s, r.b1.y, r.a2.y), fmaf(-r.s, r.b1.z, r.a2.z)};
    const float n2 = dot3(u, u);
    r.deg2 = !(n2 > 1e-24f);
    r.inv2 = rsqrtf(fmaxf(n2, 1e-24f));
    r.b2 = V3{u.x * r.inv2, u.y * r.inv2, u.z * r.inv2};
    r.b3 = cross3(r.b1, r.b2);
    return r;
}


// Backward of gram_schmidt in tangent form.  C(d6) is a rotation for every d6, so only the component of dL/dC
// along the tangent space {[k]x C} matters: omega = vee(dL/dC . C^T), dL = omega . k, and the Gram-Schmidt
// differential gives  k1 = -(da2.b3 - s da1.b3 / n1) / n2,  k2 = da1.b3 / n1,  k3 = -da1.b2 / n1.
// (Degenerate inputs |a1| or |u| < 1e-12, where F.normalize stops normalising and C is no rotation, get the same
// formula; the exact generic path covers arbitrary matrices.)
__device__ __forceinline__ void gram_schmidt_bwd_tangent(const GS& r, const V3& w, float* out) {
    const float c3 = fmaf(w.x * r.s, r.inv2, w.y) * r.inv1, c2 = -w.z * r.inv1, e3 = -w.x * r.inv2;
    out[0] = fmaf(c3, r.b3.x, c2 * r.b2.x);
    out[1] = fmaf(c3, r.b3.y, c2 * r.b2.y);
    out[2] = fmaf(c3, r.b3.z, c2 * r.b2.z);
    out[3] = e3 * r.b3.x;
    out[4] = e3 * r.b3.y;
    out[5] = e3 * r.b3.z;
}

__device__ __forceinline__ void load6(const float* __restrict__ p, float* a) {
    const float2* p2 = reinterpret_cast<const float2*>(p);  // 24-byte records: always 8-byte aligned
    const float2 x = __ldg(p2), y = __ldg(p2 + 1), z = __ldg(p2 + 2);
    a[0] = x.x; a[1] = x.y; a[2] = y.x; a[3] = y.y; a[4] = z.x; a[5] = z.y;
}

// CTA-wide coalesced smem -> global copy of a flat staging region; the widest vector both sides are aligned for
template <int NT>
__device__ __forceinline__ void block_copy_out(float* __restrict__ g, const float* __restrict__ s, int n, int tid) {
    const uintptr_t both = reinterpret_cast<uintptr_t>(g) | (uintptr_t)__cvta_generic_to_shared(s);
    if (((both & 15) == 0) && ((n & 3) == 0)) {
        float4* g4 = reinterpret_cast<float4*>(g);
        const float4* s4 = reinterpret_cast<const float4*>(s);
        const int n4 = n >> 2;
#pragma unroll 4
        for (int i = tid; i < n4; i += NT) __stcs(g4 + i, s4[i]);
    } else if (((both & 7) == 0) && ((n & 1) == 0)) {
        float2* g2 = reinterpret_cast<float2*>(g);
        const float2* s2 = reinterpret_cast<const float2*>(s);
        const int n2 = n >> 1;
#pragma unroll 4
        for (int i = tid; i < n2; i += NT) __stcs(g2 + i, s2[i]);
    } else {
        for (int i = tid; i < n; i += NT) g[i] = s[i];
    }
}

// utils/tensors.py:29-40 count of kept (joint, channel) pairs of the 2-D targets (integer => order independent)
__global__ void __launch_bounds__(256) count_kernel(const float* __restrict__ gt2, long poses, int B, int gt_c, int hips,
                                                   int* __restrict__ count) {
    int c = 0;
    const int lane = threadIdx.x & 31;
    for (long p = (long)blockIdx.x * 8 + (threadIdx.x >> 5); p < poses; p += (long)gridDim.x * 8) {
        const float* g = gt2 + p * B * gt_c;
        for (int j = lane; j < B; j += 32)
            if (j == hips || (__ldg(g + j * gt_c) != 0.f && __ldg(g + j * gt_c + 1) != 0.f)) c += 2;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(FULL_MASK, c, o);
    __shared__ int sh[8];
    if (lane == 0) sh[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < 8; ++w) t += sh[w];
        if (t) atomicAdd(count, t);
    }
}

__device__ __host__ __forceinline__ int round4(int x) { return (x + 3) & ~3; }

// shared-memory plan of one CTA (floats); every region starts 16-byte aligned
struct Smem {
    int rs, as, ts, ps, ns, ys, ident, red, total;
};
__device__ __host__ __forceinline__ Smem smem_plan(int gl, int B, int warps, int want_norm) {
    Smem s;
    int o = 0;
    s.rs = o; o += round4(gl * B * 9);   // C[t] -> R[t]: the rel_rot records of the group
    s.as = o; o += round4(gl * B * 9);   // abs_rot records
    s.ts = o; o += round4(gl * B * 3);   // abs_loc records   } after their copy-out the two regions together hold
    s.ps = o; o += round4(gl * B * 3);   // projection records } the grad-wrt-6-D records (gl * B * 6)
    s.ns = o; o += want_norm ? round4(gl * B * 3) : 0;
    s.ys = o; o += round4(gl * B * 3);   // y[t] (torque in the parent frame), then its suffix sum over frames
    s.ident = o; o += 12;
    s.red = o; o += round4(warps * 2);
    s.total = o;
    return s;
}

template <int WARPS, int BT>
__global__ void __launch_bounds__(WARPS * 32, (WARPS == 8 ? 2 : 1)) fast_kernel(const FastArgs a) {
    extern __shared__ __align__(16) float smem[];
    constexpr int NT = WARPS * 32;
    const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
    const int B = BT ? BT : a.B;
    const int B3 = B * 3, B6 = B * 6, slot = B * 9;
    const bool act = lane < B;
    const int lb = act ? lane : B - 1;   // clamped bone index: inactive lanes load valid memory and are ignored
    const Smem sm = smem_plan(a.GL, B, WARPS, a.proj_norm != nullptr);
    float* Rs = smem + sm.rs;
    float* As = smem + sm.as;
    float* Ts = smem + sm.ts;
    float* Ps = smem + sm.ps;
    float* Ns = smem + sm.ns;
    float* Ys = smem + sm.ys;
    float* Gd = smem + sm.ts;            // grad wrt 6-D records; reuses Ts+Ps once those have been copied out
    float* ident = smem + sm.ident;
    float* red = smem + sm.red;
    if (tid < 12) ident[tid] = (tid == 0 || tid == 4 || tid == 8) ? 1.f : 0.f;

    const int par = act ? a.parent[lane] : -1;
    const int psrc = par < 0 ? IDENT_LANE : par;
    const int sub_end = act ? a.sub_end[lane] : lane;
    const int ancp = act ? a.anc_packed[lane] : 0x1ffffff;  // all rounds -> IDENT_LANE
    const long total_poses = (long)a.N * a.L;
    const long groups = a.scan ? (long)a.N : (total_poses + a.GL - 1) / a.GL;
    const float k2 = a.w2 * 2.0f / (a.mask ? (float)__ldcg(a.count2) : (float)(total_poses * B * 2));
    const float k3 = a.w3 * 2.0f / (float)a.count3;
    const M3 camR = m3_from(a.cam.R);
    const float camf = a.cam.f;
    float s2 = 0.f, s3 = 0.f;

    for (long grp = blockIdx.x; grp < groups; grp += gridDim.x) {
        const long m0 = grp * a.GL;
        const int nf = (int)min((long)a.GL, total_poses - m0);

        // ---- phase A: 6-D -> C[t] (rotation_6d_to_matrix)
        for (int f = wib; f < nf; f += WARPS) {
            float d[6];
            load6(a.d6 + ((m0 + f) * B + lb) * 6, d);
            const GS r = gram_schmidt(d);
            if (act) {
                float* o = Rs + f * slot + lane * 9;
                o[0] = r.b1.x; o[1] = r.b1.y; o[2] = r.b1.z;
                o[3] = r.b2.x; o[4] = r.b2.y; o[5] = r.b2.z;
                o[6] = r.b3.x; o[7] = r.b3.y; o[8] = r.b3.z;
            }
        }
        __syncthreads();
        // ---- phase B: R[t] = C[t] . R[t-1] per bone (p3d_pose.py:98-114 looped by layers/projection.py:190-193)
        if (a.scan) {
            if (tid < B) {
                const long row = a.ref_index ? (long)__ldg(a.ref_index + grp) : grp;
                M3 prev;
                const float* rr = a.ref_rot + (row * B + tid) * 9;
#pragma unroll
                for (int i = 0; i < 9; ++i) prev.m[i] = __ldg(rr + i);
                float* o = Rs + tid * 9;
                M3 cur = m3_from(o);
                for (int f = 0; f < nf; ++f) {
                    M3 nxt;
                    if (f + 1 < nf) nxt = m3_from(o + slot);  // next C[t] in flight while this product runs
                    prev = m3_mul(cur, prev);
                    m3_to(o, prev);
                    cur = nxt;
                    o += slot;
                }
            }
            __syncthreads();
        }
        if (a.rel_rot) block_copy_out<NT>(a.rel_rot + m0 * slot, Rs, nf * slot, tid);

        // ---- phase C: per frame FK + projection + loss + torque backward
        for (int f = wib; f < nf; f += WARPS) {
            const long m = m0 + f;
            const long clip = a.scan ? grp : m / a.L;
            const long row = a.ref_index ? (long)__ldg(a.ref_index + clip) : clip;
            const float* q3 = a.gt3 + (m * B + lb) * 3;
            const V3 g3{__ldg(q3), __ldg(q3 + 1), __ldg(q3 + 2)};
            const float* q2 = a.gt2 + (m * B + lb) * a.gt_c;
            const float g2x = __ldg(q2), g2y = __ldg(q2 + 1);
            const float* rl = a.ref_loc + (row * B + lb) * 3;
            V3 t{__ldg(rl), __ldg(rl + 1), __ldg(rl + 2)};
            if (!act) t = V3{0.f, 0.f, 0.f};
            M3 A = m3_from(act ? Rs + f * slot + lane * 9 : ident);
            // FK by pointer jumping: (A, t) <- (A, t) o (A, t)[2^k-th ancestor]; abs = rel o parent o ... o root
            for (int k = 0; k < a.nrounds; ++k) {
                const int src = (ancp >> (5 * k)) & 31;
                const M3 Am = m3_shfl(A, src);
                const V3 tm = v3_shfl(t, src);
                const V3 q = v3_mul_m(t, Am);
                t = V3{q.x + tm.x, q.y + tm.y, q.z + tm.z};
                A = m3_mul(A, Am);
            }
            // projection (p3d_pose_projection.py:137-152)
            V3 c;
            {
                V3 wp{t.y, -t.x, t.z};
                if (a.world_rot) wp = v3_mul_m(wp, m3_from(a.world_rot + m * 9));
                if (a.world_loc) { wp.x += __ldg(a.world_loc + m * 3); wp.y += __ldg(a.world_loc + m * 3 + 1); wp.z += __ldg(a.world_loc + m * 3 + 2); }
                c = v3_mul_m(wp, camR);
                c.x += a.cam.T[0]; c.y += a.cam.T[1]; c.z += a.cam.T[2];
            }
            const float iz = (1.0f / c.z);
            const V3 p{fmaf(-camf * c.x, iz, a.cam.px), fmaf(-camf * c.y, iz, a.cam.py), iz};
            if (act) {
                m3_to(As + f * slot + lane * 9, A);
                float* o = Ts + f * B3 + lane * 3;
                o[0] = t.x; o[1] = t.y; o[2] = t.z;
                o = Ps + f * B3 + lane * 3;
                o[0] = p.x; o[1] = p.y; o[2] = p.z;
            }
            // Normalizer(HipsNeckExtractor) + loc_2d + loc_3d and their gradient wrt (proj, abs_loc)
            float x = p.x, y = p.y;
            bool lx = true, ly = true;
            float inv = 1.f, dx = 0.f, dy = 0.f;
            bool okn = true;
            if (a.normalize) {
                const float sx = __shfl_sync(FULL_MASK, p.x, a.hips), sy = __shfl_sync(FULL_MASK, p.y, a.hips);
                dx = __shfl_sync(FULL_MASK, p.x, a.neck) - sx;
                dy = __shfl_sync(FULL_MASK, p.y, a.neck) - sy;
                const float scale = sqrtf(fmaf(dx, dx, dy * dy));
                okn = scale > 0.f && scale <= 3.402823466e38f;
                inv = (1.0f / scale);
                const float rx = (p.x - sx) * inv, ry = (p.y - sy) * inv;
                lx = fabsf(rx) <= 3.402823466e38f;
                ly = fabsf(ry) <= 3.402823466e38f;
                x = lx ? rx : 0.f;
                y = ly ? ry : 0.f;
                const float q = nan_to_zero(p.z);
                if (!(q >= 1e-5f)) { x = 0.f; y = 0.f; lx = ly = false; }
                if (a.proj_norm && act) {
                    float* o = Ns + f * B3 + lane * 3;
                    o[0] = x; o[1] = y; o[2] = q;
                }
            }
            const bool kept = act && (!a.mask || lane == a.hips || (g2x != 0.f && g2y != 0.f));
            float gx = 0.f, gy = 0.f;
            if (kept) {
                const float ex = x - g2x, ey = y - g2y;
                s2 = fmaf(ex, ex, fmaf(ey, ey, s2));
                gx = lx ? k2 * ex : 0.f;
                gy = ly ? k2 * ey : 0.f;
            }
            if (a.normalize) {
                if (!okn) { gx = 0.f; gy = 0.f; inv = 0.f; }
                const float sgx = warp_sum(gx), sgy = warp_sum(gy), sgn = warp_sum(fmaf(gx, x, gy * y));
                gx *= inv;
                gy *= inv;
                // d/d(hips, neck) through shift and scale: scale = |neck - hips|, u = (neck - hips) / scale
                const float gs = -sgn * inv * inv;   // dL/dscale . (1/scale)
                const float kx = gs * dx, ky = gs * dy;
                if (lane == a.neck) { gx += kx; gy += ky; }
                if (lane == a.hips) { gx -= fmaf(sgx, inv, kx); gy -= fmaf(sgy, inv, ky); }
            }
            V3 g;  // dL / d abs_loc[lane]
            {
                const V3 e{t.x - g3.x, t.y - g3.y, t.z - g3.z};
                if (act) s3 = fmaf(e.x, e.x, fmaf(e.y, e.y, fmaf(e.z, e.z, s3)));
                // backward of the projection: (gx, gy, 0) -> camera point -> world point -> p3d point
                const float fgx = camf * gx * iz, fgy = camf * gy * iz;
                const V3 gc{-fgx, -fgy, fmaf(fgx, c.x, fgy * c.y) * iz};
                V3 gw = v3_mul_mt(gc, camR);
                if (a.world_rot) gw = v3_mul_mt(gw, m3_from(a.world_rot + m * 9));
                g = V3{fmaf(k3, e.x, -gw.y), fmaf(k3, e.y, gw.x), fmaf(k3, e.z, gw.z)};
                if (!act) g = V3{0.f, 0.f, 0.f};
            }
            // torque backward: m_b = sum_{d in subtree(b)} g_d x (t_d - t_b); subtree sums of (g x t, g) are
            // differences of an inclusive prefix scan over the DFS-ordered lanes
            float w[6] = {g.y * t.z - g.z * t.y, g.z * t.x - g.x * t.z, g.x * t.y - g.y * t.x, g.x, g.y, g.z};
            float ex_[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) ex_[i] = w[i];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    const float v = __shfl_up_sync(FULL_MASK, w[i], o);
                    if (lane >= o) w[i] += v;
                }
            }
            float s[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) s[i] = __shfl_sync(FULL_MASK, w[i], sub_end) - (w[i] - ex_[i]);
            const V3 mv{s[0] - (s[4] * t.z - s[5] * t.y), s[1] - (s[5] * t.x - s[3] * t.z), s[2] - (s[3] * t.y - s[4] * t.x)};
            const M3 Ap = m3_shfl(A, psrc);       // IDENT_LANE keeps the identity for the root
            const V3 yv = v3_mul_mt(mv, Ap);      // torque rotated into the parent frame
            if (act) {
                float* o = Ys + f * B3 + lane * 3;
                o[0] = yv.x; o[1] = yv.y; o[2] = yv.z;
            }
        }
        __syncthreads();
        if (a.abs_rot) block_copy_out<NT>(a.abs_rot + m0 * slot, As, nf * slot, tid);
        if (a.abs_loc) block_copy_out<NT>(a.abs_loc + m0 * B3, Ts, nf * B3, tid);
        if (a.proj) block_copy_out<NT>(a.proj + m0 * B3, Ps, nf * B3, tid);
        if (a.proj_norm) block_copy_out<NT>(a.proj_norm + m0 * B3, Ns, nf * B3, tid);
        // ---- phase D: z[t] = sum_{s >= t} y[s]  (the reverse-time scan in tangent form is a plain suffix sum)
        if (a.scan) {
            for (int col = tid; col < B3; col += NT) {
                float acc = 0.f;
                for (int f = nf - 1; f >= 0; --f) {
                    acc += Ys[f * B3 + col];
                    Ys[f * B3 + col] = acc;
                }
            }
        }
        __syncthreads();
        // ---- phase E: omega[t] = R[t] z[t]; tangent-form Gram-Schmidt backward -> grad wrt the 6-D input
        for (int f = wib; f < nf; f += WARPS) {
            float d[6], o6[6];
            load6(a.d6 + ((m0 + f) * B + lb) * 6, d);
            const float* zp = Ys + f * B3 + lb * 3;
            const V3 z{zp[0], zp[1], zp[2]};
            const V3 om = v3_mul_mt(z, m3_from(Rs + f * slot + lb * 9));
            gram_schmidt_bwd_tangent(gram_schmidt(d), om, o6);
            if (act) {
                float* o = Gd + f * B6 + lane * 6;
#pragma unroll
                for (int i = 0; i < 6; ++i) o[i] = o6[i];
            }
        }
        __syncthreads();
        block_copy_out<NT>(a.g_d6 + m0 * B6, Gd, nf * B6, tid);
        // (the next group's phase A writes Rs only; Gd/Ts/Ps are next written in phase C, two barriers later)
    }

    // ---- loss sums: fixed-order reduction, the last CTA to finish folds the per-CTA partials
    s2 = warp_sum(s2);
    s3 = warp_sum(s3);
    if (lane == 0) { red[wib * 2] = s2; red[wib * 2 + 1] = s3; }
    __syncthreads();
    __shared__ bool last;
    if (tid == 0) {
        float x = 0.f, y = 0.f;
        for (int w = 0; w < WARPS; ++w) { x += red[w * 2]; y += red[w * 2 + 1]; }
        a.partials[blockIdx.x * 2] = x;
        a.partials[blockIdx.x * 2 + 1] = y;
        __threadfence();
        last = (atomicAdd(a.ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (last && wib == 0) {
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
            const double c2 = a.mask ? (double)*a.count2 : (double)total_poses * B * 2;
            const float l2 = (float)(d2 / c2), l3 = (float)(d3 / a.count3);
            a.losses[0] = l2; a.losses[1] = l3; a.losses[2] = a.w2 * l2 + a.w3 * l3;
            a.losses[3] = (float)d2; a.losses[4] = (float)c2; a.losses[5] = (float)d3; a.losses[6] = (float)a.count3;
            a.losses[7] = 0.f;
            exchange_publish(a.xchg, (float)d2, (float)c2, (float)d3, (float)a.count3);
            *a.ticket = 0u;   // self-resetting workspace header: ready for the next launch
            *a.count2 = 0;
        }
    }
}
// ------------------------------------------------------------------------------------------------------------------
// Staged variant (the default when a clip group fits): the same phases, but every HBM access is a TMA bulk copy.
//   inputs  : cp.async.bulk global -> shared for the NEXT clip group (6-D records, 3-D targets, 2-D targets), double
//             buffered and signalled through an mbarrier, issued by one thread while the CTA works on the current group
//   outputs : cp.async.bulk shared -> global of the flat staging regions (rel_rot, abs_rot, abs_loc, projection,
//             normalised projection, grad wrt 6-D), issued by one thread after the barrier that completes the region
// so no warp ever waits on a global load and the LSU only sees shared-memory traffic.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
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

struct SmemStaged {
    int in_g3, in_g2, stage;  // input stage k starts at k * stage: 6-D records at +0, 3-D targets, 2-D targets
    int rs, as, ts, ps, ns, gd, ys, ident, red, bars, total;
};
// slim = 1: one input stage, abs_loc / projection records written over the target records they were computed from,
// gradient records over the abs_rot records: 55 KB per 16-frame CARLA clip, four CTAs per SM
__device__ __host__ __forceinline__ SmemStaged staged_plan(int gl, int B, int gt_c, int warps, int want_norm,
                                                           int slim = 0) {
    SmemStaged s;
    int o = 0;
    s.in_g3 = round4(gl * B * 6);
    s.in_g2 = s.in_g3 + round4(gl * B * 3);
    s.stage = s.in_g2 + round4(gl * B * (gt_c > 3 ? gt_c : 3));
    o = (slim ? 1 : 2) * s.stage;
    s.rs = o; o += round4(gl * B * 9);
    s.as = o; o += round4(gl * B * 9);
    if (slim) {
        s.ts = s.in_g3;
        s.ps = s.in_g2;
    } else {
        s.ts = o; o += round4(gl * B * 3);
        s.ps = o; o += round4(gl * B * 3);
    }
    s.ns = o; o += want_norm ? round4(gl * B * 3) : 0;
    if (slim) {
        s.gd = s.as;
    } else {
        s.gd = o; o += round4(gl * B * 6);
    }
    s.ys = o; o += round4(gl * B * 3);
    s.ident = o; o += 12;
    s.red = o; o += round4(warps * 2);
    s.bars = o; o += 4;  // two 8-byte mbarriers
    s.total = o;
    return s;
}

// one flat region: TMA when source, destination and size are 16-byte multiples, else a cooperative copy by all threads
__device__ __forceinline__ bool bulk_ok(const void* g, long off_floats, int n_floats) {
    return (((reinterpret_cast<uintptr_t>(g) + (uintptr_t)off_floats * 4) & 15) == 0) && ((n_floats & 3) == 0);
}

template <int WARPS, int BT, bool SLIM>
__global__ void __launch_bounds__(WARPS * 32, (SLIM ? 4 : (WARPS == 8 ? 2 : 1))) staged_kernel(const FastArgs a) {
    extern __shared__ __align__(16) float smem[];
    constexpr int NT = WARPS * 32;
    const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
    const int B = BT ? BT : a.B;
    const int B3 = B * 3, B6 = B * 6, slot = B * 9, Bc = B * a.gt_c;
    const bool act = lane < B;
    const int lb = act ? lane : B - 1;
    const SmemStaged sm = staged_plan(a.GL, B, a.gt_c, WARPS, a.proj_norm != nullptr, SLIM ? 1 : 0);
    float* Rs = smem + sm.rs;
    float* As = smem + sm.as;
    float* Ts = smem + sm.ts;
    float* Ps = smem + sm.ps;
    float* Ns = smem + sm.ns;
    float* Gd = smem + sm.gd;
    float* Ys = smem + sm.ys;
    float* ident = smem + sm.ident;
    float* red = smem + sm.red;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + sm.bars);
    if (tid < 12) ident[tid] = (tid == 0 || tid == 4 || tid == 8) ? 1.f : 0.f;
    if (tid == 0) {
        mbar_init(bars, 1);
        mbar_init(bars + 1, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int par = act ? a.parent[lane] : -1;
    const int psrc = par < 0 ? IDENT_LANE : par;
    const int sub_end = act ? a.sub_end[lane] : lane;
    const int ancp = act ? a.anc_packed[lane] : 0x1ffffff;
    const long total_poses = (long)a.N * a.L;
    const long groups = a.scan ? (long)a.N : (total_poses + a.GL - 1) / a.GL;
    const float k2 = a.w2 * 2.0f / (a.mask ? (float)__ldcg(a.count2) : (float)(total_poses * B * 2));
    const float k3 = a.w3 * 2.0f / (float)a.count3;
    const M3 camR = m3_from(a.cam.R);
    const float camf = a.cam.f;
    float s2 = 0.f, s3 = 0.f;

    auto frames_of = [&](long g) { return (int)min((long)a.GL, total_poses - g * a.GL); };
    auto in_bulk = [&](long g) {
        const long m0 = g * a.GL;
        const int nf = frames_of(g);
        return bulk_ok(a.d6, m0 * B6, nf * B6) && bulk_ok(a.gt3, m0 * B3, nf * B3) && bulk_ok(a.gt2, m0 * Bc, nf * Bc);
    };
    auto issue_load = [&](long g, int st) {  // one thread
        const long m0 = g * a.GL;
        const int nf = frames_of(g);
        mbar_expect_tx(bars + st, (uint32_t)(nf * (B6 + B3 + Bc)) * 4u);
        float* in = smem + st * sm.stage;
        bulk_g2s(in, a.d6 + m0 * B6, (uint32_t)(nf * B6) * 4u, bars + st);
        bulk_g2s(in + sm.in_g3, a.gt3 + m0 * B3, (uint32_t)(nf * B3) * 4u, bars + st);
        bulk_g2s(in + sm.in_g2, a.gt2 + m0 * Bc, (uint32_t)(nf * Bc) * 4u, bars + st);
    };
    auto store_out = [&](float* g, const float* s, long off, int n) {  // all threads call; after a barrier
        if (g == nullptr) return;
        if (bulk_ok(g, off, n)) {
            if (tid == 0) bulk_s2g(g + off, s, (uint32_t)n * 4u);
        } else {
            block_copy_out<NT>(g + off, s, n, tid);
        }
    };

    unsigned uses0 = 0, uses1 = 0;  // completed TMA fills per stage (mbarrier phase parity)
    int it = 0;
    if (!SLIM && (long)blockIdx.x < groups && in_bulk(blockIdx.x) && tid == 0) issue_load(blockIdx.x, 0);

    for (long grp = blockIdx.x; grp < groups; grp += gridDim.x, ++it) {
        const int st = SLIM ? 0 : (it & 1);
        const long m0 = grp * a.GL;
        const int nf = frames_of(grp);
        float* d6s = smem + st * sm.stage;
        float* g3s = d6s + sm.in_g3;
        float* g2s = d6s + sm.in_g2;
        if (SLIM) {  // single stage: its readers passed the closing barrier, the stores out of it were waited for
            if (in_bulk(grp) && tid == 0) issue_load(grp, 0);
        } else {     // prefetch the next group of this CTA into the other stage
            const long nxt = grp + gridDim.x;
            if (nxt < groups && in_bulk(nxt) && tid == 0) issue_load(nxt, st ^ 1);
        }
        if (in_bulk(grp)) {
            const unsigned u = st ? uses1 : uses0;
            while (!mbar_try_wait(bars + st, u & 1u)) {}
            if (st) ++uses1; else ++uses0;
        } else {  // unaligned group (odd clip length / odd pose offset): synchronous cooperative fill
            for (int i = tid; i < nf * B6; i += NT) d6s[i] = __ldg(a.d6 + m0 * B6 + i);
            for (int i = tid; i < nf * B3; i += NT) g3s[i] = __ldg(a.gt3 + m0 * B3 + i);
            for (int i = tid; i < nf * Bc; i += NT) g2s[i] = __ldg(a.gt2 + m0 * Bc + i);
            __syncthreads();
        }
        V3 lref{0.f, 0.f, 0.f};
        if (a.scan) {  // one skeleton per clip: its bone offsets stay in registers for the whole group
            const long row = a.ref_index ? (long)__ldg(a.ref_index + grp) : grp;
            const float* rl = a.ref_loc + (row * B + lb) * 3;
            lref = V3{__ldg(rl), __ldg(rl + 1), __ldg(rl + 2)};
        }

        // ---- phase A: 6-D -> C[t] (rotation_6d_to_matrix)
        for (int f = wib; f < nf; f += WARPS) {
            const float2* dp = reinterpret_cast<const float2*>(d6s + f * B6 + lb * 6);
            const float2 u0 = dp[0], u1 = dp[1], u2 = dp[2];
            const float d[6] = {u0.x, u0.y, u1.x, u1.y, u2.x, u2.y};
            const GS r = gram_schmidt(d);
            if (act) {
                float* o = Rs + f * slot + lane * 9;
                o[0] = r.b1.x; o[1] = r.b1.y; o[2] = r.b1.z;
                o[3] = r.b2.x; o[4] = r.b2.y; o[5] = r.b2.z;
                o[6] = r.b3.x; o[7] = r.b3.y; o[8] = r.b3.z;
            }
        }
        fence_async_smem();   // relative_rot mode: these records are what the TMA store below reads
        __syncthreads();
        // ---- phase B: R[t] = C[t] . R[t-1] per bone
        if (a.scan) {
            if (tid < B) {
                const long row = a.ref_index ? (long)__ldg(a.ref_index + grp) : grp;
                M3 prev;
                const float* rr = a.ref_rot + (row * B + tid) * 9;
#pragma unroll
                for (int i = 0; i < 9; ++i) prev.m[i] = __ldg(rr + i);
                float* o = Rs + tid * 9;
                M3 cur = m3_from(o);
                for (int f = 0; f < nf; ++f) {
                    M3 nxt;
                    if (f + 1 < nf) nxt = m3_from(o + slot);
                    prev = m3_mul(cur, prev);
                    m3_to(o, prev);
                    cur = nxt;
                    o += slot;
                }
                fence_async_smem();
            }
            __syncthreads();
        }
        store_out(a.rel_rot, Rs, m0 * slot, nf * slot);
        if (tid == 0) bulk_commit();

        // ---- phase C: per frame FK + projection + loss + torque backward
        for (int f = wib; f < nf; f += WARPS) {
            const long m = m0 + f;
            const float* q3 = g3s + f * B3 + lb * 3;
            const V3 g3{q3[0], q3[1], q3[2]};
            const float* q2 = g2s + (f * B + lb) * a.gt_c;
            const float g2x = q2[0], g2y = q2[1];
            V3 t = lref;
            if (!a.scan) {
                const long clip = m / a.L;
                const long row = a.ref_index ? (long)__ldg(a.ref_index + clip) : clip;
                const float* rl = a.ref_loc + (row * B + lb) * 3;
                t = V3{__ldg(rl), __ldg(rl + 1), __ldg(rl + 2)};
            }
            if (!act) t = V3{0.f, 0.f, 0.f};
            M3 A = m3_from(act ? Rs + f * slot + lane * 9 : ident);
            for (int k = 0; k < a.nrounds; ++k) {
                const int src = (ancp >> (5 * k)) & 31;
                const M3 Am = m3_shfl(A, src);
                const V3 tm = v3_shfl(t, src);
                const V3 q = v3_mul_m(t, Am);
                t = V3{q.x + tm.x, q.y + tm.y, q.z + tm.z};
                A = m3_mul(A, Am);
            }
            V3 c;
            {
                V3 wp{t.y, -t.x, t.z};
                if (a.world_rot) wp = v3_mul_m(wp, m3_from(a.world_rot + m * 9));
                if (a.world_loc) { wp.x += __ldg(a.world_loc + m * 3); wp.y += __ldg(a.world_loc + m * 3 + 1); wp.z += __ldg(a.world_loc + m * 3 + 2); }
                c = v3_mul_m(wp, camR);
                c.x += a.cam.T[0]; c.y += a.cam.T[1]; c.z += a.cam.T[2];
            }
            const float iz = (1.0f / c.z);
            const V3 p{fmaf(-camf * c.x, iz, a.cam.px), fmaf(-camf * c.y, iz, a.cam.py), iz};
            if (act) {
                m3_to(As + f * slot + lane * 9, A);
                float* o = Ts + f * B3 + lane * 3;
                o[0] = t.x; o[1] = t.y; o[2] = t.z;
                o = Ps + f * B3 + lane * 3;
                o[0] = p.x; o[1] = p.y; o[2] = p.z;
            }
            float x = p.x, y = p.y;
            bool lx = true, ly = true;
            float inv = 1.f, dx = 0.f, dy = 0.f;
            bool okn = true;
            if (a.normalize) {
                const float sx = __shfl_sync(FULL_MASK, p.x, a.hips), sy = __shfl_sync(FULL_MASK, p.y, a.hips);
                dx = __shfl_sync(FULL_MASK, p.x, a.neck) - sx;
                dy = __shfl_sync(FULL_MASK, p.y, a.neck) - sy;
                const float scale = sqrtf(fmaf(dx, dx, dy * dy));
                okn = scale > 0.f && scale <= 3.402823466e38f;
                inv = (1.0f / scale);
                const float rx = (p.x - sx) * inv, ry = (p.y - sy) * inv;
                lx = fabsf(rx) <= 3.402823466e38f;
                ly = fabsf(ry) <= 3.402823466e38f;
                x = lx ? rx : 0.f;
                y = ly ? ry : 0.f;
                const float q = nan_to_zero(p.z);
                if (!(q >= 1e-5f)) { x = 0.f; y = 0.f; lx = ly = false; }
                if (a.proj_norm && act) {
                    float* o = Ns + f * B3 + lane * 3;
                    o[0] = x; o[1] = y; o[2] = q;
                }
            }
            const bool kept = act && (!a.mask || lane == a.hips || (g2x != 0.f && g2y != 0.f));
            float gx = 0.f, gy = 0.f;
            if (kept) {
                const float ex = x - g2x, ey = y - g2y;
                s2 = fmaf(ex, ex, fmaf(ey, ey, s2));
                gx = lx ? k2 * ex : 0.f;
                gy = ly ? k2 * ey : 0.f;
            }
            if (a.normalize) {
                if (!okn) { gx = 0.f; gy = 0.f; inv = 0.f; }
                const float sgx = warp_sum(gx), sgy = warp_sum(gy), sgn = warp_sum(fmaf(gx, x, gy * y));
                gx *= inv;
                gy *= inv;
                const float gs = -sgn * inv * inv;
                const float kx = gs * dx, ky = gs * dy;
                if (lane == a.neck) { gx += kx; gy += ky; }
                if (lane == a.hips) { gx -= fmaf(sgx, inv, kx); gy -= fmaf(sgy, inv, ky); }
            }
            V3 g;
            {
                const V3 e{t.x - g3.x, t.y - g3.y, t.z - g3.z};
                if (act) s3 = fmaf(e.x, e.x, fmaf(e.y, e.y, fmaf(e.z, e.z, s3)));
                const float fgx = camf * gx * iz, fgy = camf * gy * iz;
                const V3 gc{-fgx, -fgy, fmaf(fgx, c.x, fgy * c.y) * iz};
                V3 gw = v3_mul_mt(gc, camR);
                if (a.world_rot) gw = v3_mul_mt(gw, m3_from(a.world_rot + m * 9));
                g = V3{fmaf(k3, e.x, -gw.y), fmaf(k3, e.y, gw.x), fmaf(k3, e.z, gw.z)};
                if (!act) g = V3{0.f, 0.f, 0.f};
            }
            float w[6] = {g.y * t.z - g.z * t.y, g.z * t.x - g.x * t.z, g.x * t.y - g.y * t.x, g.x, g.y, g.z};
            float ex_[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) ex_[i] = w[i];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    const float v = __shfl_up_sync(FULL_MASK, w[i], o);
                    if (lane >= o) w[i] += v;
                }
            }
            float s[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) s[i] = __shfl_sync(FULL_MASK, w[i], sub_end) - (w[i] - ex_[i]);
            const V3 mv{s[0] - (s[4] * t.z - s[5] * t.y), s[1] - (s[5] * t.x - s[3] * t.z), s[2] - (s[3] * t.y - s[4] * t.x)};
            const M3 Ap = m3_shfl(A, psrc);
            const V3 yv = v3_mul_mt(mv, Ap);
            if (act) {
                float* o = Ys + f * B3 + lane * 3;
                o[0] = yv.x; o[1] = yv.y; o[2] = yv.z;
            }
        }
        fence_async_smem();
        if (tid == 0) bulk_wait_read_all();   // the previous group's grad store has finished reading Gd
        __syncthreads();
        store_out(a.abs_rot, As, m0 * slot, nf * slot);
        store_out(a.abs_loc, Ts, m0 * B3, nf * B3);
        store_out(a.proj, Ps, m0 * B3, nf * B3);
        store_out(a.proj_norm, Ns, m0 * B3, nf * B3);
        if (tid == 0) bulk_commit();
        if (SLIM && tid == 0) bulk_wait_read_all();   // abs_rot records are about to be overwritten by the gradient
        // ---- phase D: z[t] = sum_{s >= t} y[s]
        if (a.scan) {
            for (int col = tid; col < B3; col += NT) {
                float acc = 0.f;
                for (int f = nf - 1; f >= 0; --f) {
                    acc += Ys[f * B3 + col];
                    Ys[f * B3 + col] = acc;
                }
            }
            __syncthreads();
        } else if (SLIM) {
            __syncthreads();
        }
        // ---- phase E: omega[t] = R[t] z[t]; tangent-form Gram-Schmidt backward -> grad wrt the 6-D input
        for (int f = wib; f < nf; f += WARPS) {
            const float2* dp = reinterpret_cast<const float2*>(d6s + f * B6 + lb * 6);
            const float2 u0 = dp[0], u1 = dp[1], u2 = dp[2];
            const float d[6] = {u0.x, u0.y, u1.x, u1.y, u2.x, u2.y};
            float o6[6];
            const float* zp = Ys + f * B3 + lb * 3;
            const V3 z{zp[0], zp[1], zp[2]};
            const V3 om = v3_mul_mt(z, m3_from(Rs + f * slot + lb * 9));
            gram_schmidt_bwd_tangent(gram_schmidt(d), om, o6);
            if (act) {
                float2* o = reinterpret_cast<float2*>(Gd + f * B6 + lane * 6);
                o[0] = make_float2(o6[0], o6[1]);
                o[1] = make_float2(o6[2], o6[3]);
                o[2] = make_float2(o6[4], o6[5]);
            }
        }
        fence_async_smem();
        if (tid == 0) bulk_wait_read_all();   // rel_rot / abs_* / projection stores have finished reading their regions
        __syncthreads();
        store_out(a.g_d6, Gd, m0 * B6, nf * B6);
        if (tid == 0) bulk_commit();
    }
    if (tid == 0) bulk_wait_all();

    // ---- loss sums: fixed-order reduction, the last CTA to finish folds the per-CTA partials
    s2 = warp_sum(s2);
    s3 = warp_sum(s3);
    if (lane == 0) { red[wib * 2] = s2; red[wib * 2 + 1] = s3; }
    __syncthreads();
    __shared__ bool last;
    if (tid == 0) {
        float x = 0.f, y = 0.f;
        for (int w = 0; w < WARPS; ++w) { x += red[w * 2]; y += red[w * 2 + 1]; }
        a.partials[blockIdx.x * 2] = x;
        a.partials[blockIdx.x * 2 + 1] = y;
        __threadfence();
        last = (atomicAdd(a.ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (last && wib == 0) {
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
            const double c2 = a.mask ? (double)*a.count2 : (double)total_poses * B * 2;
            const float l2 = (float)(d2 / c2), l3 = (float)(d3 / a.count3);
            a.losses[0] = l2; a.losses[1] = l3; a.losses[2] = a.w2 * l2 + a.w3 * l3;
            a.losses[3] = (float)d2; a.losses[4] = (float)c2; a.losses[5] = (float)d3; a.losses[6] = (float)a.count3;
            a.losses[7] = 0.f;
            exchange_publish(a.xchg, (float)d2, (float)c2, (float)d3, (float)a.count3);
            *a.ticket = 0u;
            *a.count2 = 0;
        }
    }
}

struct Plan {
    int ok, warps, grid, gl, staged, slim;
    size_t smem;
};

int sm_count() { return p2c_sm_count(); }   // of the current device (host_util.cu)

size_t smem_bytes(int gl, int B, int warps, int want_norm) {
    return (size_t)smem_plan(gl, B, warps, want_norm).total * sizeof(float);
}

// the bone axis must be a DFS pre-order (every subtree a contiguous index range): pose.py:41-47 produces one
bool preorder_tables(const p2c_tree_t& t, FastArgs& a) {
    const int B = t.n_bones;
    if (B < 1 || B >= IDENT_LANE + 1) return false;  // lane 31 is reserved for the identity
    int depth[P2C_MAX_BONES], size[P2C_MAX_BONES], stack[P2C_MAX_BONES], top = 0, maxd = 0;
    if (t.parent[0] != -1) return false;
    stack[top++] = 0;
    depth[0] = 0;
    for (int b = 1; b < B; ++b) {
        const int p = t.parent[b];
        if (p < 0 || p >= b) return false;
        while (top > 0 && stack[top - 1] != p) --top;
        if (top == 0) return false;
        stack[top++] = b;
        depth[b] = depth[p] + 1;
        if (depth[b] > maxd) maxd = depth[b];
    }
    for (int b = 0; b < B; ++b) size[b] = 1;
    for (int b = B - 1; b > 0; --b) size[t.parent[b]] += size[b];
    int rounds = 0;
    while ((1 << rounds) < maxd + 1) ++rounds;
    if (rounds > MAX_ROUNDS) return false;
    a.nrounds = rounds;
    for (int b = 0; b < P2C_MAX_BONES; ++b) {
        a.parent[b] = (int8_t)(b < B ? t.parent[b] : -1);
        a.sub_end[b] = (int8_t)(b < B ? b + size[b] - 1 : b);
        int packed = 0;
        for (int k = 0; k < MAX_ROUNDS; ++k) {
            int anc = b < B ? b : -1;
            for (int s = 0; s < (1 << k) && anc >= 0; ++s) anc = t.parent[anc];
            packed |= ((anc >= 0 && b < B) ? anc : IDENT_LANE) << (5 * k);
        }
        a.anc_packed[b] = packed;
    }
    return true;
}

Plan make_plan(const p2c_fused_step_t* p) {
    Plan pl{};
    const int B = p->tree.n_bones;
    const long total = (long)p->n_clips * p->clip_len;
    pl.gl = p->mode == P2C_MODE_POSE_CHANGES ? p->clip_len : (int)(total < 16 ? total : 16);
    const size_t cap = 227 * 1024;
    const int want_norm = p->proj_norm != nullptr;
    const size_t t8 = (size_t)staged_plan(pl.gl, B, p->gt_2d_channels, 8, want_norm).total * sizeof(float);
    const size_t t16 = (size_t)staged_plan(pl.gl, B, p->gt_2d_channels, 16, want_norm).total * sizeof(float);
    const size_t s8 = smem_bytes(pl.gl, B, 8, want_norm), s16 = smem_bytes(pl.gl, B, 16, want_norm);
    const P2cEnv& env = p2c_env();
    const bool allow_staged = !env.disable_tma;  // test hook: the register-staged variant
    const long groups_ = p->mode == P2C_MODE_POSE_CHANGES ? p->n_clips : (total + pl.gl - 1) / pl.gl;
    const size_t tslim = (size_t)staged_plan(pl.gl, B, p->gt_2d_channels, 8, want_norm, 1).total * sizeof(float);
    if (allow_staged && !env.disable_slim && p->gt_2d_channels == 3 && 4 * (tslim + 1024) <= 228 * 1024 &&
        (groups_ >= 3L * sm_count() || env.force_slim)) {
        pl.warps = 8; pl.smem = tslim; pl.staged = 1; pl.slim = 1;   // many clips: 4 CTAs (32 warps) per SM
    } else if (allow_staged && (2 * (t8 + 1024) <= cap || pl.gl <= 8) && t8 + 1024 <= cap) {
        pl.warps = 8; pl.smem = t8; pl.staged = 1;       // TMA-staged, two CTAs per SM
    } else if (allow_staged && t16 + 1024 <= cap) {
        pl.warps = 16; pl.smem = t16; pl.staged = 1;     // TMA-staged, one 16-warp CTA per SM
    } else if (2 * (s8 + 1024) <= cap || pl.gl <= 8) {
        pl.warps = 8; pl.smem = s8;
    } else {
        pl.warps = 16; pl.smem = s16;
    }
    if (pl.smem + 1024 > cap) return pl;
    const long groups = p->mode == P2C_MODE_POSE_CHANGES ? p->n_clips : (total + pl.gl - 1) / pl.gl;
    int per_sm = pl.warps == 8 ? (int)(cap / (pl.smem + 1024)) : 1;
    if (pl.slim) per_sm = 4;      // __launch_bounds__(256, 4): 64 registers per thread
    else if (per_sm > 2) per_sm = 2;   // __launch_bounds__(256, 2): 128 registers per thread
    if (per_sm < 1) per_sm = 1;
    const long slots = (long)sm_count() * per_sm;
    pl.grid = (int)(groups < slots ? groups : slots);
    pl.ok = 1;
    return pl;
}

}  // namespace

extern "C" int p2c_fused_fast_supported(const p2c_fused_step_t* p) {
    if (p2c_env().disable_fast) return 0;  // test hook: force the exact generic composite
    if (p->gt_abs_rot) return 0;           // rot_3d is fused in the tile kernel only; other shapes: generic composite
    if (!p->input_6d) return 0;  // arbitrary 3x3 input is not assumed to be a rotation: exact generic path
    if (p->mode != P2C_MODE_POSE_CHANGES && p->mode != P2C_MODE_RELATIVE_ROT) return 0;
    if (p->mode == P2C_MODE_POSE_CHANGES && !p->ref_rot) return 0;
    FastArgs a{};
    if (!preorder_tables(p->tree, a)) return 0;
    return make_plan(p).ok;
}

extern "C" int p2c_fused_fast_launches(const p2c_fused_step_t* p) { return p->mask_missing_joints ? 2 : 1; }

extern "C" long p2c_fused_fast_workspace_floats(void) { return 16 + 2L * 148 * 4; }

extern "C" int p2c_fused_fast(const p2c_fused_step_t* p, void* stream) {
    FastArgs args{};
    if (!preorder_tables(p->tree, args)) return P2C_E_UNSUPPORTED;
    const Plan pl = make_plan(p);
    if (!pl.ok) return P2C_E_UNSUPPORTED;
    const int B = p->tree.n_bones;
    args.N = p->n_clips; args.L = p->clip_len; args.GL = pl.gl; args.B = B; args.slot = B * 9;
    args.scan = p->mode == P2C_MODE_POSE_CHANGES;
    args.hips = p->hips; args.neck = p->neck; args.normalize = p->normalize; args.mask = p->mask_missing_joints;
    args.gt_c = p->gt_2d_channels;
    args.d6 = p->pose_in; args.ref_loc = p->ref_loc; args.ref_rot = p->ref_rot; args.ref_index = p->ref_index;
    args.world_loc = p->world_loc; args.world_rot = p->world_rot; args.cam = p->camera;
    args.gt2 = p->gt_2d; args.gt3 = p->gt_3d; args.w2 = p->weight_2d; args.w3 = p->weight_3d;
    args.rel_rot = p->rel_rot; args.abs_rot = p->abs_rot; args.abs_loc = p->abs_loc; args.proj = p->proj;
    args.proj_norm = p->proj_norm; args.g_d6 = p->g_pose_in;
    args.ticket = reinterpret_cast<unsigned int*>(p->workspace);
    args.count2 = reinterpret_cast<int*>(p->workspace) + 1;
    args.partials = p->workspace + 16;
    args.losses = p->losses;
    args.xchg = XchgArgs{p->exchange, p->exchange_rank, p->exchange_world};
    const long total = (long)p->n_clips * p->clip_len;
    args.count3 = (double)total * B * 3;
    cudaStream_t st = (cudaStream_t)stream;
    if (args.mask) {
        long blocks = (total + 7) / 8;
        if (blocks > COUNT_BLOCKS) blocks = COUNT_BLOCKS;
        count_kernel<<<(int)blocks, 256, 0, st>>>(p->gt_2d, total, B, p->gt_2d_channels, p->hips, args.count2);
        P2C_LAUNCH_CHECK();
    }
    typedef void (*kern_t)(const FastArgs);
    const bool carla = (B == 26);   // CARLA_SKELETON: bone count known at compile time (unrolled copies, immediates)
    kern_t k;
    if (pl.staged)
        k = pl.slim ? (carla ? staged_kernel<8, 26, true> : staged_kernel<8, 0, true>)
            : pl.warps == 8 ? (carla ? staged_kernel<8, 26, false> : staged_kernel<8, 0, false>)
                            : (carla ? staged_kernel<16, 26, false> : staged_kernel<16, 0, false>);
    else
        k = pl.warps == 8 ? (carla ? fast_kernel<8, 26> : fast_kernel<8, 0>)
                          : (carla ? fast_kernel<16, 26> : fast_kernel<16, 0>);
    if (const int e = p2c_allow_big_smem(reinterpret_cast<const void*>(k), 227 * 1024 - 1024)) return e;
    k<<<pl.grid, pl.warps * 32, pl.smem, st>>>(args);
    P2C_LAUNCH_CHECK();
    return 0;
}
