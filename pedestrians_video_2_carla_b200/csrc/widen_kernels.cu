// the callers either side of the fused step.
//   cum_pose_changes scan : loss/cum_pose_changes.py:9-56 (per-bone cumulative product of the pose changes over frames)
//   Normalizer(dim=2|3) + ReferenceSkeletonsDeNormalizer.from_abs / from_projection:
//                           transforms/pose/normalization/normalizer.py:20-41, extractor.py:23-36,
//                           denormalizer.py:13-37, reference_skeletons_denormalizer.py:41-91
//                           (modules/layers/projection.py:125-142: absolute_loc(_rot) movements output)
// Both are streaming passes: every input float is read once and every output float written once.
#include "common.cuh"

namespace {

constexpr int WARPS = 4;
constexpr float FMAX = 3.402823466e38f;

// ---- cum_pose_changes: P[n,t,b] = P[n,t-1,b] . C[n,t,b], P[n,-1,b] = I --------------------------
// thread = (clip, bone); for a fixed frame the 26 bones of a clip are 936 contiguous bytes, so a warp's loads of one
// frame touch one or two contiguous runs and every sector fetched is fully used.
// Direct form (any alignment): a warp's nine 36-byte-strided loads touch 32 sectors each.
__global__ void __launch_bounds__(128) cum_changes_fwd_kernel(const float* __restrict__ c, float* __restrict__ out,
                                                              long units, int L, int B) {
    const long u = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= units) return;
    const long n = u / B;
    const int b = (int)(u - n * B);
    M3 p = m3_identity();
#pragma unroll 4
    for (int t = 0; t < L; ++t) {
        const long e = ((n * L + t) * B + b) * 9;
        p = m3_mul(p, m3_from(c + e));
        m3_to(out + e, p);
    }
}

// Staged form (B * 9 even, 8-byte aligned tensors): a CTA owns CK clips; per frame their CK rows of B * 9 floats come in
// as coalesced float2 (prefetched into registers one frame ahead), the (clip, bone) threads update their 3x3 in shared
// memory in place (stride 9 words: conflict free), and the rows leave as coalesced float2.  ncu on the direct form:
// 32 sectors per request, stalled on lg_throttle.
constexpr int CK = 8;
constexpr int CT = 256;
constexpr int CPF = 4;   // float2 per thread and frame: CK * 16 * 9 / 2 = 576 float2 <= CT * CPF (B <= 32)

__global__ void __launch_bounds__(CT) cum_changes_fwd_staged_kernel(const float* __restrict__ c,
                                                                   float* __restrict__ out, int N, int L, int B,
                                                                   int ck) {
    extern __shared__ float2 tile2[];                 // ck * row2 float2 (ck <= CK clips per CTA)
    float* tile = reinterpret_cast<float*>(tile2);
    const int row2 = B * 9 / 2;                        // float2 per (clip, frame) row
    const int n0 = blockIdx.x * ck;
    const int nk = min(ck, N - n0);
    const int total2 = nk * row2;
    const int k = threadIdx.x / B, b = threadIdx.x - k * B;
    const bool worker = threadIdx.x < nk * B;
    float2 pre[CPF];
    long base[CPF];   // float offset of this thread's r-th float2 in frame 0 (the frame adds t * B * 9): no per-frame division
#pragma unroll
    for (int r = 0; r < CPF; ++r) {
        const int i = threadIdx.x + r * CT;
        const int kk = i / row2, w = i - kk * row2;
        base[r] = (i < total2) ? (long)(n0 + kk) * L * B * 9 + 2 * w : -1;
    }
    const int frame = B * 9;
    auto fetch = [&](int t) {
#pragma unroll
        for (int r = 0; r < CPF; ++r)
            if (base[r] >= 0) pre[r] = __ldg(reinterpret_cast<const float2*>(c + base[r] + (long)t * frame));
    };
    M3 p = m3_identity();
    fetch(0);
    for (int t = 0; t < L; ++t) {
#pragma unroll
        for (int r = 0; r < CPF; ++r)
            if (base[r] >= 0) tile2[threadIdx.x + r * CT] = pre[r];
        __syncthreads();
        if (t + 1 < L) fetch(t + 1);
        if (worker) {
            float* m = tile + (k * B + b) * 9;
            p = m3_mul(p, m3_from(m));
            m3_to(m, p);
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < CPF; ++r)
            if (base[r] >= 0)
                *reinterpret_cast<float2*>(out + base[r] + (long)t * frame) = tile2[threadIdx.x + r * CT];
        __syncthreads();
    }
}

// G[t] = g_out[t] + G[t+1] . C[t+1]^T ;  g_C[t] = P[t-1]^T . G[t]
__global__ void __launch_bounds__(128) cum_changes_bwd_kernel(const float* __restrict__ c,
                                                              const float* __restrict__ cum,
                                                              const float* __restrict__ g_out,
                                                              float* __restrict__ g_c, long units, int L, int B) {
    const long u = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= units) return;
    const long n = u / B;
    const int b = (int)(u - n * B);
    M3 g = m3_zero();
    for (int t = L - 1; t >= 0; --t) {
        const long e = ((n * L + t) * B + b) * 9;
        m3_acc(g, m3_from(g_out + e));
        const M3 prev = (t > 0) ? m3_from(cum + e - (long)B * 9) : m3_identity();
        m3_to(g_c + e, m3_mul_at(prev, g));
        g = m3_mul_bt(g, m3_from(c + e));
    }
}


// Staged backward: the same CTA / row mapping as the staged forward, three input tiles per frame (C[t], g_out[t],
// P[t-1]), walking the frames backwards; g_C[t] replaces g_out[t] in its tile and leaves as coalesced float2.
__global__ void __launch_bounds__(CT) cum_changes_bwd_staged_kernel(const float* __restrict__ c,
                                                                   const float* __restrict__ cum,
                                                                   const float* __restrict__ g_out,
                                                                   float* __restrict__ g_c, int N, int L, int B,
                                                                   int ck) {
    extern __shared__ float2 tile2[];                 // 3 tiles of ck * row2 float2
    const int row2 = B * 9 / 2;
    const int n0 = blockIdx.x * ck;
    const int nk = min(ck, N - n0);
    const int total2 = nk * row2;
    float2* t_c = tile2;
    float2* t_g = tile2 + ck * row2;
    float2* t_p = tile2 + 2 * ck * row2;
    const int k = threadIdx.x / B, b = threadIdx.x - k * B;
    const bool worker = threadIdx.x < nk * B;
    float2 pc[CPF], pg[CPF], pp[CPF];
    long base[CPF];
#pragma unroll
    for (int r = 0; r < CPF; ++r) {
        const int i = threadIdx.x + r * CT;
        const int kk = i / row2, w = i - kk * row2;
        base[r] = (i < total2) ? (long)(n0 + kk) * L * B * 9 + 2 * w : -1;
    }
    const int frame = B * 9;
    auto fetch = [&](int t) {
#pragma unroll
        for (int r = 0; r < CPF; ++r)
            if (base[r] >= 0) {
                const long e = base[r] + (long)t * frame;
                pc[r] = __ldg(reinterpret_cast<const float2*>(c + e));
                pg[r] = __ldg(reinterpret_cast<const float2*>(g_out + e));
                if (t > 0) pp[r] = __ldg(reinterpret_cast<const float2*>(cum + e - frame));
            }
    };
    M3 g = m3_zero();
    fetch(L - 1);
    for (int t = L - 1; t >= 0; --t) {
#pragma unroll
        for (int r = 0; r < CPF; ++r)
            if (base[r] >= 0) {
                const int i = threadIdx.x + r * CT;
                t_c[i] = pc[r]; t_g[i] = pg[r]; t_p[i] = pp[r];
            }
        __syncthreads();
        if (t > 0) fetch(t - 1);
        if (worker) {
            const int o = (k * B + b) * 9;
            float* mg = reinterpret_cast<float*>(t_g) + o;
            m3_acc(g, m3_from(mg));
            const M3 prev = (t > 0) ? m3_from(reinterpret_cast<const float*>(t_p) + o) : m3_identity();
            m3_to(mg, m3_mul_at(prev, g));
            g = m3_mul_bt(g, m3_from(reinterpret_cast<const float*>(t_c) + o));
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < CPF; ++r)
            if (base[r] >= 0)
                *reinterpret_cast<float2*>(g_c + base[r] + (long)t * frame) = t_g[threadIdx.x + r * CT];
        __syncthreads();
    }
}

// ---- Normalizer (optional) followed by DeNormalizer.from_reference (optional) ------------------
// one warp per pose, lanes stride over joints.  DIM = channels that are shifted and scaled (2: projections, 3:
// absolute locations); C = channels per joint (DIM, or 3 with DIM 2: channel 2 is carried through).
struct Renorm {
    long M;
    int J, C, L, hips, neck, autonormalize;
    float near_zero;
    const float* ref_shift;   // (rows, DIM) or NULL
    const float* ref_scale;   // (rows) or NULL
    const int* ref_index;     // (clips) or NULL: row = clip
};

template <int DIM>
__global__ void __launch_bounds__(WARPS * 32) renorm_fwd_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                               Renorm a) {
    const int lane = threadIdx.x & 31;
    const int C = a.C;
    const long wstride = (long)gridDim.x * WARPS;
    for (long m = (long)blockIdx.x * WARPS + (threadIdx.x >> 5); m < a.M; m += wstride) {
        const float* p = in + m * a.J * C;
        float* o = out + m * a.J * C;
        float s[DIM], scale = 1.f, rs = 1.f, rh[DIM];
#pragma unroll
        for (int k = 0; k < DIM; ++k) { s[k] = 0.f; rh[k] = 0.f; }
        if (a.autonormalize) {
            float q = 0.f;
#pragma unroll
            for (int k = 0; k < DIM; ++k) {
                s[k] = __ldg(p + a.hips * C + k);
                const float d = __ldg(p + a.neck * C + k) - s[k];
                q = fmaf(d, d, q);
            }
            scale = sqrtf(q);
        }
        if (a.ref_scale) {
            const long clip = m / a.L;
            const long row = a.ref_index ? (long)__ldg(a.ref_index + clip) : clip;
            rs = __ldg(a.ref_scale + row);
#pragma unroll
            for (int k = 0; k < DIM; ++k) rh[k] = __ldg(a.ref_shift + row * DIM + k);
        }
        for (int j = lane; j < a.J; j += 32) {
            float y[DIM];
#pragma unroll
            for (int k = 0; k < DIM; ++k)
                y[k] = a.autonormalize ? nan_to_zero((p[j * C + k] - s[k]) / scale) : p[j * C + k];
            if (DIM == 2 && C > 2) {
                float q = p[j * C + 2];
                if (a.autonormalize) {
                    q = nan_to_zero(q);
                    if (!(q >= a.near_zero)) {
#pragma unroll
                        for (int k = 0; k < DIM; ++k) y[k] = 0.f;
                    }
                }
                o[j * C + 2] = q;
            }
#pragma unroll
            for (int k = 0; k < DIM; ++k) o[j * C + k] = a.ref_scale ? fmaf(y[k], rs, rh[k]) : y[k];
        }
    }
}

template <int DIM>
__global__ void __launch_bounds__(WARPS * 32) renorm_bwd_kernel(const float* __restrict__ in,
                                                               const float* __restrict__ g_out,
                                                               float* __restrict__ g_in, Renorm a) {
    const int lane = threadIdx.x & 31;
    const int C = a.C;
    const long wstride = (long)gridDim.x * WARPS;
    for (long m = (long)blockIdx.x * WARPS + (threadIdx.x >> 5); m < a.M; m += wstride) {
        const float* p = in + m * a.J * C;
        const float* g = g_out + m * a.J * C;
        float* o = g_in + m * a.J * C;
        float rs = 1.f;
        if (a.ref_scale) {
            const long clip = m / a.L;
            rs = __ldg(a.ref_scale + (a.ref_index ? (long)__ldg(a.ref_index + clip) : clip));
        }
        if (!a.autonormalize) {
            for (int j = lane; j < a.J; j += 32) {
#pragma unroll
                for (int k = 0; k < DIM; ++k) o[j * C + k] = g[j * C + k] * rs;
                if (DIM == 2 && C > 2) o[j * C + 2] = g[j * C + 2];
            }
            continue;
        }
        float s[DIM], d[DIM], q2 = 0.f;
#pragma unroll
        for (int k = 0; k < DIM; ++k) {
            s[k] = __ldg(p + a.hips * C + k);
            d[k] = __ldg(p + a.neck * C + k) - s[k];
            q2 = fmaf(d[k], d[k], q2);
        }
        const float scale = sqrtf(q2);
        const bool ok = scale > 0.f && scale <= FMAX;
        float sg[DIM], sgn = 0.f;
#pragma unroll
        for (int k = 0; k < DIM; ++k) sg[k] = 0.f;
        for (int j = lane; j < a.J; j += 32) {
            float gy[DIM], r[DIM];
            bool zeroed = false;
            if (DIM == 2 && C > 2) {
                const float qraw = p[j * C + 2];
                const bool fin = fabsf(qraw) <= FMAX;
                zeroed = !((fin ? qraw : 0.f) >= a.near_zero);
                o[j * C + 2] = fin ? g[j * C + 2] : 0.f;
            }
#pragma unroll
            for (int k = 0; k < DIM; ++k) {
                r[k] = (p[j * C + k] - s[k]) / scale;
                gy[k] = g[j * C + k] * rs;
                // nan_to_num passes no gradient to non-finite entries
                if (!(fabsf(r[k]) <= FMAX) || zeroed || !ok) gy[k] = 0.f;
                o[j * C + k] = ok ? gy[k] / scale : 0.f;
                sg[k] += gy[k];
                sgn = fmaf(gy[k], nan_to_zero(r[k]), sgn);
            }
        }
#pragma unroll
        for (int k = 0; k < DIM; ++k) sg[k] = warp_sum(sg[k]);
        sgn = warp_sum(sgn);
        __syncwarp();
        if (lane == 0 && ok) {
            const float gscale = -sgn / scale;  // d / d scale
#pragma unroll
            for (int k = 0; k < DIM; ++k) {
                const float un = gscale * d[k] / scale;  // through scale = |neck - hips|
                if (a.hips == a.neck) {
                    o[a.hips * C + k] += -sg[k] / scale;
                } else {
                    o[a.hips * C + k] += -sg[k] / scale - un;
                    o[a.neck * C + k] += un;
                }
            }
        }
    }
}


// ---- quaternion_to_matrix (pytorch3d 0.6.0 rotation_conversions.quaternion_to_matrix, real part first) -------------
// The north star's quaternion entry next to the Euler / 6-D ones.  M = I + (2 / |q|^2) A(q), A quadratic in q; not
// normalised first (pytorch3d does not either), so the gradient has the -(2/|q|^2)^2 q <G, A> term of the scale.
// 16 B in / 36 B out per record: quaternions are one float4 per thread, matrices go through shared memory (stride 9
// is conflict free) and leave as coalesced float4.
constexpr int QT = 256;

__device__ __forceinline__ void quat_A(float r, float i, float j, float k, float* A) {
    A[0] = -(j * j + k * k); A[1] = i * j - k * r;    A[2] = i * k + j * r;
    A[3] = i * j + k * r;    A[4] = -(i * i + k * k); A[5] = j * k - i * r;
    A[6] = i * k - j * r;    A[7] = j * k + i * r;    A[8] = -(i * i + j * j);
}

__global__ void __launch_bounds__(QT) quat_fwd_kernel(const float* __restrict__ q, float* __restrict__ out, long n) {
    __shared__ __align__(16) float s[QT * 9];
    const long first = (long)blockIdx.x * QT;
    const int cnt = (int)min((long)QT, n - first);
    if (threadIdx.x < cnt) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(q) + first + threadIdx.x);
        float A[9];
        quat_A(v.x, v.y, v.z, v.w, A);
        const float two_s = 2.0f / (v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w);
#pragma unroll
        for (int e = 0; e < 9; ++e) s[threadIdx.x * 9 + e] = ((e & 3) == 0 ? 1.f : 0.f) + two_s * A[e];
    }
    __syncthreads();
    float* g = out + first * 9;  // QT * 9 floats per CTA: 16-byte aligned whenever `out` is
    const int count = cnt * 9;
    const int n4 = ((reinterpret_cast<uintptr_t>(g) & 15) == 0) ? (count >> 2) : 0;
    for (int t = threadIdx.x; t < n4; t += QT) reinterpret_cast<float4*>(g)[t] = reinterpret_cast<const float4*>(s)[t];
    for (int t = (n4 << 2) + threadIdx.x; t < count; t += QT) g[t] = s[t];
}

__global__ void __launch_bounds__(QT) quat_bwd_kernel(const float* __restrict__ q, const float* __restrict__ g_out,
                                                      float* __restrict__ g_q, long n) {
    __shared__ __align__(16) float s[QT * 9];
    const long first = (long)blockIdx.x * QT;
    const int cnt = (int)min((long)QT, n - first);
    const float* g = g_out + first * 9;
    const int count = cnt * 9;
    const int n4 = ((reinterpret_cast<uintptr_t>(g) & 15) == 0) ? (count >> 2) : 0;
    for (int t = threadIdx.x; t < n4; t += QT) reinterpret_cast<float4*>(s)[t] = __ldg(reinterpret_cast<const float4*>(g) + t);
    for (int t = (n4 << 2) + threadIdx.x; t < count; t += QT) s[t] = __ldg(g + t);
    __syncthreads();
    if (threadIdx.x >= cnt) return;
    const float4 v = __ldg(reinterpret_cast<const float4*>(q) + first + threadIdx.x);
    const float r = v.x, i = v.y, j = v.z, k = v.w;
    float G[9], A[9];
#pragma unroll
    for (int e = 0; e < 9; ++e) G[e] = s[threadIdx.x * 9 + e];
    quat_A(r, i, j, k, A);
    const float two_s = 2.0f / (r * r + i * i + j * j + k * k);
    float ga = 0.f;
#pragma unroll
    for (int e = 0; e < 9; ++e) ga = fmaf(G[e], A[e], ga);
    const float dr = -k * G[1] + j * G[2] + k * G[3] - i * G[5] - j * G[6] + i * G[7];
    const float di = j * (G[1] + G[3]) + k * (G[2] + G[6]) - 2.f * i * (G[4] + G[8]) + r * (G[7] - G[5]);
    const float dj = -2.f * j * (G[0] + G[8]) + i * (G[1] + G[3]) + r * (G[2] - G[6]) + k * (G[5] + G[7]);
    const float dk = -2.f * k * (G[0] + G[4]) + r * (G[3] - G[1]) + i * (G[2] + G[6]) + j * (G[5] + G[7]);
    const float c = two_s * two_s * ga;  // d(2/|q|^2)/dq = -(2/|q|^2)^2 q
    float4 o;
    o.x = two_s * dr - c * r; o.y = two_s * di - c * i; o.z = two_s * dj - c * j; o.w = two_s * dk - c * k;
    reinterpret_cast<float4*>(g_q)[first + threadIdx.x] = o;
}

// ---- C == DIM (absolute locations, 2-channel projections): chunked flat variant --------------------------------
// A CTA takes 32 consecutive poses (32 * J * DIM floats: a 16-byte multiple for even J * DIM), one thread per pose
// derives the pose's constants into shared memory, then the whole chunk streams through as float4 with the pose and
// channel of every float recovered from its flat index.  The per-element work is one FMA: with A = rs / scale and
// B_k = rh_k - s_k * A, (x - s_k) / scale * rs + rh_k = x * A + B_k (the per-element IEEE division of the first
// version kept the kernel instruction bound at 35 % of the HBM rate: ncu, issue slots 72 % busy).  nan_to_zero keeps
// its meaning: a degenerate scale (0, inf, NaN) or a non-finite input gives y = 0, i.e. the bare reference shift.
constexpr int RP = 32;
constexpr int RTH = 128;

template <int DIM>
struct PoseConst {
    float s[DIM], rh[DIM], d[DIM], hx[DIM], kx[DIM], bk[DIM];
    float scale, rs, inv, A;
    int ok;
};

template <int DIM>
__device__ __forceinline__ void pose_consts(PoseConst<DIM>& c, const float* p, long m, const Renorm& a) {
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < DIM; ++k) {
        c.s[k] = 0.f; c.rh[k] = 0.f; c.d[k] = 0.f; c.hx[k] = 0.f; c.kx[k] = 0.f;
        if (a.autonormalize) {
            c.s[k] = p[a.hips * DIM + k];      // plain loads: p may point into shared memory
            c.d[k] = p[a.neck * DIM + k] - c.s[k];
            q = fmaf(c.d[k], c.d[k], q);
        }
    }
    c.scale = a.autonormalize ? sqrtf(q) : 1.f;
    c.ok = c.scale > 0.f && c.scale <= FMAX;
    c.inv = c.ok ? 1.0f / c.scale : 0.f;
    c.rs = 1.f;
    if (a.ref_scale) {
        const long clip = m / a.L;
        const long row = a.ref_index ? (long)__ldg(a.ref_index + clip) : clip;
        c.rs = __ldg(a.ref_scale + row);
        if (a.ref_shift) {
#pragma unroll
            for (int k = 0; k < DIM; ++k) c.rh[k] = __ldg(a.ref_shift + row * DIM + k);
        }
    }
    c.A = c.ok ? c.rs / c.scale : 0.f;
#pragma unroll
    for (int k = 0; k < DIM; ++k) c.bk[k] = c.ok ? fmaf(-c.s[k], c.A, c.rh[k]) : c.rh[k];
    // (a non-finite hips or neck coordinate makes the scale non-finite: ok = 0, A = 0, y = 0 for the whole pose, which
    //  is what nan_to_zero leaves of (x - s) / scale in the reference)
}

template <int DIM>
__device__ __forceinline__ float pick(const float* v, int k) {
    return k == 0 ? v[0] : (k == 1 ? v[1] : v[DIM - 1]);
}

template <int DIM>
__global__ void __launch_bounds__(RTH) renorm_flat_fwd_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                             Renorm a) {
    __shared__ PoseConst<DIM> cs[RP];
    const int JC = a.J * DIM;
    const float inv_jc = 1.0f / (float)JC;
    for (long chunk = blockIdx.x; chunk * RP < a.M; chunk += gridDim.x) {
        const long m0 = chunk * RP;
        const int np = (int)min((long)RP, a.M - m0);
        if (threadIdx.x < np) pose_consts<DIM>(cs[threadIdx.x], in + (m0 + threadIdx.x) * JC, m0 + threadIdx.x, a);
        __syncthreads();
        const float* src = in + m0 * JC;
        float* dst = out + m0 * JC;
        const int nf = np * JC;
        const bool vec = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
        const int n4 = vec ? (nf >> 2) : 0;
        auto value = [&](float x, float A, float B, float shift) {
            // y = 0 (-> the bare shift) for what nan_to_zero cuts: a non-finite x
            return (a.autonormalize && !(fabsf(x) <= FMAX)) ? shift : fmaf(x, A, B);
        };
        for (int i = threadIdx.x; i < n4; i += RTH) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(src) + i);
            float r[4] = {v.x, v.y, v.z, v.w};
            // floor(e / JC) through the reciprocal: exact, the quotient is >= 0.5 / JC away from an integer boundary
            int pose = (int)(((float)(4 * i) + 0.5f) * inv_jc);
            int rem = 4 * i - pose * JC;
            int k = rem % DIM;
            // the pose's constants in registers, reloaded only where the float4 crosses into the next pose
            float A = cs[pose].A, bk[DIM], rh[DIM];
#pragma unroll
            for (int d = 0; d < DIM; ++d) { bk[d] = cs[pose].bk[d]; rh[d] = cs[pose].rh[d]; }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                r[q] = value(r[q], A, pick<DIM>(bk, k), pick<DIM>(rh, k));
                k = (k + 1 == DIM) ? 0 : k + 1;
                if (++rem == JC) {
                    rem = 0;
                    pose = min(pose + 1, np - 1);
                    A = cs[pose].A;
#pragma unroll
                    for (int d = 0; d < DIM; ++d) { bk[d] = cs[pose].bk[d]; rh[d] = cs[pose].rh[d]; }
                }
            }
            reinterpret_cast<float4*>(dst)[i] = make_float4(r[0], r[1], r[2], r[3]);
        }
        for (int e = (n4 << 2) + threadIdx.x; e < nf; e += RTH) {
            const int pose = e / JC, k = (e - pose * JC) % DIM;
            const PoseConst<DIM>& c = cs[pose];
            dst[e] = value(__ldg(src + e), c.A, pick<DIM>(c.bk, k), pick<DIM>(c.rh, k));
        }
        __syncthreads();
    }
}

// Backward: the chunk's x and g rows are staged in shared memory by ONE coalesced float4 pass (all loads of a thread
// in flight together); the per-pose sums over joints and the output pass then read shared memory.  (The first version
// re-read global memory in a warp-per-pose reduction: eight dependent load / shuffle rounds per warp and chunk.)
template <int DIM>
__global__ void __launch_bounds__(RTH) renorm_flat_bwd_kernel(const float* __restrict__ in,
                                                             const float* __restrict__ g_out,
                                                             float* __restrict__ g_in, Renorm a) {
    __shared__ PoseConst<DIM> cs[RP];
    extern __shared__ __align__(16) float stage[];    // x rows, then g rows: 2 * RP * JC floats
    const int JC = a.J * DIM;
    float* sx = stage;
    float* sg_ = stage + RP * JC;
    const float inv_jc = 1.0f / (float)JC;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long chunk = blockIdx.x; chunk * RP < a.M; chunk += gridDim.x) {
        const long m0 = chunk * RP;
        const int np = (int)min((long)RP, a.M - m0);
        const float* src = in + m0 * JC;
        const float* gsrc = g_out + m0 * JC;
        float* dst = g_in + m0 * JC;
        const int nf = np * JC;
        const bool vec = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(gsrc) |
                           reinterpret_cast<uintptr_t>(dst)) & 15) == 0 && ((RP * JC) & 3) == 0;
        const int n4 = vec ? (nf >> 2) : 0;
        for (int i = threadIdx.x; i < n4; i += RTH) {
            reinterpret_cast<float4*>(sx)[i] = __ldg(reinterpret_cast<const float4*>(src) + i);
            reinterpret_cast<float4*>(sg_)[i] = __ldg(reinterpret_cast<const float4*>(gsrc) + i);
        }
        for (int e = (n4 << 2) + threadIdx.x; e < nf; e += RTH) { sx[e] = __ldg(src + e); sg_[e] = __ldg(gsrc + e); }
        __syncthreads();
        if (threadIdx.x < np) pose_consts<DIM>(cs[threadIdx.x], sx + threadIdx.x * JC, m0 + threadIdx.x, a);
        __syncthreads();
        // the upstream gradient of one normalised element (zero where nan_to_zero or a degenerate scale cut it);
        // r = (x - s_k) / scale through the pose's reciprocal
        auto upstream = [&](float x, float g, float sk, float inv, float rs, float& r) {
            r = (x - sk) * inv;
            return (fabsf(r) <= FMAX) ? g * rs : 0.f;
        };
        if (a.autonormalize) {
            for (int pose = warp; pose < np; pose += RTH / 32) {   // sums over the joints of a pose: one warp each
                PoseConst<DIM>& c = cs[pose];
                const float inv = c.inv, rs = c.ok ? c.rs : 0.f;
                float sg[DIM], sgn = 0.f;
#pragma unroll
                for (int k = 0; k < DIM; ++k) sg[k] = 0.f;
                for (int e = lane; e < JC; e += 32) {
                    const int k = e % DIM;
                    float r;
                    const float gy = upstream(sx[pose * JC + e], sg_[pose * JC + e], pick<DIM>(c.s, k), inv, rs, r);
#pragma unroll
                    for (int kk = 0; kk < DIM; ++kk) sg[kk] += (kk == k) ? gy : 0.f;
                    sgn = fmaf(gy, nan_to_zero(r), sgn);
                }
#pragma unroll
                for (int k = 0; k < DIM; ++k) sg[k] = warp_sum(sg[k]);
                sgn = warp_sum(sgn);
                if (lane == 0 && c.ok) {
                    const float gscale = -sgn * inv;
#pragma unroll
                    for (int k = 0; k < DIM; ++k) {
                        const float un = gscale * c.d[k] * inv;
                        c.hx[k] = (a.hips == a.neck) ? -sg[k] * inv : -sg[k] * inv - un;
                        c.kx[k] = (a.hips == a.neck) ? 0.f : un;
                    }
                }
            }
            __syncthreads();
        }
        // output pass: the pose's constants live in registers and are reloaded only where a float4 crosses into the
        // next pose; channel and joint of an element advance incrementally (no division per element)
        struct PoseReg { float s[DIM], hx[DIM], kx[DIM], inv, rs; };
        auto load_reg = [&](PoseReg& q, int pose) {
            const PoseConst<DIM>& c = cs[pose];
#pragma unroll
            for (int k = 0; k < DIM; ++k) { q.s[k] = c.s[k]; q.hx[k] = c.hx[k]; q.kx[k] = c.kx[k]; }
            q.inv = c.inv;
            q.rs = (a.autonormalize && !c.ok) ? 0.f : c.rs;
        };
        auto value = [&](float x, float g, const PoseReg& q, int joint, int k) {
            if (!a.autonormalize) return g * q.rs;
            float r;
            const float gy = upstream(x, g, pick<DIM>(q.s, k), q.inv, q.rs, r);
            float o = gy * q.inv;
            if (joint == a.hips) o += pick<DIM>(q.hx, k);
            if (joint == a.neck) o += pick<DIM>(q.kx, k);
            return o;
        };
        for (int i = threadIdx.x; i < n4; i += RTH) {
            const float4 v = reinterpret_cast<const float4*>(sx)[i];
            const float4 w = reinterpret_cast<const float4*>(sg_)[i];
            const float x[4] = {v.x, v.y, v.z, v.w}, g[4] = {w.x, w.y, w.z, w.w};
            float r[4];
            int pose = (int)(((float)(4 * i) + 0.5f) * inv_jc);
            const int rem = 4 * i - pose * JC;
            int joint = rem / DIM, k = rem - joint * DIM;
            PoseReg q;
            load_reg(q, pose);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                r[e] = value(x[e], g[e], q, joint, k);
                if (++k == DIM) {
                    k = 0;
                    if (++joint == a.J) { joint = 0; load_reg(q, min(++pose, np - 1)); }
                }
            }
            reinterpret_cast<float4*>(dst)[i] = make_float4(r[0], r[1], r[2], r[3]);
        }
        for (int e = (n4 << 2) + threadIdx.x; e < nf; e += RTH) {
            const int pose = e / JC, rem = e - pose * JC;
            PoseReg q;
            load_reg(q, pose);
            dst[e] = value(sx[e], sg_[e], q, rem / DIM, rem % DIM);
        }
        __syncthreads();
    }
}

int chunk_grid(long m) {
    long blocks = (m + RP - 1) / RP;
    if (blocks > 148L * 32) blocks = 148L * 32;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

// ---- Normalizer with the BBoxExtractor / HipsNeckBBoxFallbackExtractor (bbox_extractor.py:6-18,
//      hips_neck_bbox_fallback_extractor.py:10-46; utils/tensors.py:12-26) -- forward only: these normalise dataset
//      INPUTS (`hips_neck_bbox` is the reference's default CLI transform, base_datamodule.py:61,276).
//      extractor 1: shift = centre of the box around the detected points, scale = |y_min - y_centre|;
//      extractor 2: hips / neck as usual, but scale = 0.5748 * the box scale when hips or neck are undetected
//      (the reference's fallback SHIFT is written into a temporary and has no effect, :27-31 -- reproduced).
__global__ void __launch_bounds__(WARPS * 32) normalize_extractor_kernel(const float* __restrict__ in,
                                                                        float* __restrict__ out, long M, int J, int C,
                                                                        int hips, int neck, int extractor,
                                                                        float near_zero) {
    const int lane = threadIdx.x & 31;
    const float inf = __builtin_huge_valf();
    const long wstride = (long)gridDim.x * WARPS;
    for (long m = (long)blockIdx.x * WARPS + (threadIdx.x >> 5); m < M; m += wstride) {
        const float* p = in + m * J * C;
        float* o = out + m * J * C;
        float lox = inf, loy = inf, hix = -inf, hiy = -inf;
        for (int j = lane; j < J; j += 32) {
            const float x = p[j * C], y = p[j * C + 1];
            if (!(x < near_zero && y < near_zero)) {   // detected
                lox = fminf(lox, x); hix = fmaxf(hix, x);
                loy = fminf(loy, y); hiy = fmaxf(hiy, y);
            }
        }
        for (int s = 16; s > 0; s >>= 1) {
            lox = fminf(lox, __shfl_xor_sync(FULL_MASK, lox, s)); hix = fmaxf(hix, __shfl_xor_sync(FULL_MASK, hix, s));
            loy = fminf(loy, __shfl_xor_sync(FULL_MASK, loy, s)); hiy = fmaxf(hiy, __shfl_xor_sync(FULL_MASK, hiy, s));
        }
        const float cx = (lox + hix) / 2.f, cy = (loy + hiy) / 2.f;
        // |(cx, y_min) - (cx, cy)|_2 as torch.linalg.norm forms it
        const float ddx = cx - cx, ddy = loy - cy;
        const float bb_scale = sqrtf(ddx * ddx + ddy * ddy);
        float sx, sy, scale;
        if (extractor == 1) {
            sx = cx; sy = cy; scale = bb_scale;
        } else {
            const float hx = __ldg(p + hips * C), hy = __ldg(p + hips * C + 1);
            const float nx = __ldg(p + neck * C), ny = __ldg(p + neck * C + 1);
            sx = hx; sy = hy;
            scale = sqrtf((nx - hx) * (nx - hx) + (ny - hy) * (ny - hy));
            const bool missing_hips = hx < near_zero && hy < near_zero;
            const bool missing_neck = nx < near_zero && ny < near_zero;
            if (missing_hips || missing_neck) scale = bb_scale * 0.5748f;
        }
        for (int j = lane; j < J; j += 32) {
            float nx = nan_to_zero((p[j * C] - sx) / scale);
            float ny = nan_to_zero((p[j * C + 1] - sy) / scale);
            if (C > 2) {
                const float q = nan_to_zero(p[j * C + 2]);
                if (!(q >= near_zero)) { nx = 0.f; ny = 0.f; }
                o[j * C + 2] = q;
            }
            o[j * C] = nx;
            o[j * C + 1] = ny;
        }
    }
}

// ---- Extractor.get_shift_scale alone (extractor.py:23-36 and the three extractors): what Normalizer.shift / .scale
//      report after a call (normalizer.py:43-49, read by data/base/mixins/dataset/projection_2d_mixin.py:229-230).
//      extractor 0: hips / neck over `dim` channels; 1 / 2: the bounding-box extractors above (dim 2).
__global__ void __launch_bounds__(WARPS * 32) shift_scale_kernel(const float* __restrict__ in, float* __restrict__ shift,
                                                                float* __restrict__ scale, long M, int J, int C, int dim,
                                                                int hips, int neck, int extractor, float near_zero) {
    const int lane = threadIdx.x & 31;
    const float inf = __builtin_huge_valf();
    const long wstride = (long)gridDim.x * WARPS;
    for (long m = (long)blockIdx.x * WARPS + (threadIdx.x >> 5); m < M; m += wstride) {
        const float* p = in + m * J * C;
        float s[3] = {0.f, 0.f, 0.f}, sc;
        float d2 = 0.f;
        for (int k = 0; k < dim; ++k) {
            s[k] = __ldg(p + hips * C + k);
            const float d = __ldg(p + neck * C + k) - s[k];
            d2 = fmaf(d, d, d2);
        }
        sc = sqrtf(d2);
        if (extractor != 0) {
            float lox = inf, loy = inf, hix = -inf, hiy = -inf;
            for (int j = lane; j < J; j += 32) {
                const float x = p[j * C], y = p[j * C + 1];
                if (!(x < near_zero && y < near_zero)) {
                    lox = fminf(lox, x); hix = fmaxf(hix, x);
                    loy = fminf(loy, y); hiy = fmaxf(hiy, y);
                }
            }
            for (int o = 16; o > 0; o >>= 1) {
                lox = fminf(lox, __shfl_xor_sync(FULL_MASK, lox, o)); hix = fmaxf(hix, __shfl_xor_sync(FULL_MASK, hix, o));
                loy = fminf(loy, __shfl_xor_sync(FULL_MASK, loy, o)); hiy = fmaxf(hiy, __shfl_xor_sync(FULL_MASK, hiy, o));
            }
            const float cx = (lox + hix) / 2.f, cy = (loy + hiy) / 2.f;
            const float ddx = cx - cx, ddy = loy - cy;   // (as torch.linalg.norm forms it: NaN boxes stay NaN)
            const float bb_scale = sqrtf(ddx * ddx + ddy * ddy);
            if (extractor == 1) {
                s[0] = cx; s[1] = cy; sc = bb_scale;
            } else {
                const bool missing_hips = s[0] < near_zero && s[1] < near_zero;
                const float nx = __ldg(p + neck * C), ny = __ldg(p + neck * C + 1);
                const bool missing_neck = nx < near_zero && ny < near_zero;
                if (missing_hips || missing_neck) sc = bb_scale * 0.5748f;
            }
        }
        if (lane < dim) shift[m * dim + lane] = s[lane];
        if (lane == 0) scale[m] = sc;
    }
}

int warp_grid(long units) {
    long blocks = (units + WARPS - 1) / WARPS;
    if (blocks > 148L * 64) blocks = 148L * 64;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

int renorm_args(Renorm& a, int64_t m, int32_t j, int32_t c, int32_t dim, int32_t hips, int32_t neck,
                int32_t autonormalize, float near_zero, const float* ref_shift, const float* ref_scale,
                const int32_t* ref_index, int32_t clip_len) {
    if (m < 0 || j < 1 || (dim != 2 && dim != 3) || !(c == dim || (dim == 2 && c == 3))) return P2C_E_BADARG;
    if (hips < 0 || hips >= j || neck < 0 || neck >= j) return P2C_E_BADARG;
    if ((ref_shift == nullptr) != (ref_scale == nullptr)) return P2C_E_NULL;
    if (ref_scale && clip_len < 1) return P2C_E_BADARG;
    a.M = (long)m; a.J = j; a.C = c; a.L = clip_len > 0 ? clip_len : 1; a.hips = hips; a.neck = neck;
    a.autonormalize = autonormalize ? 1 : 0; a.near_zero = near_zero;
    a.ref_shift = ref_shift; a.ref_scale = ref_scale; a.ref_index = ref_index;
    return 0;
}

}  // namespace

// pose-per-lane variants (lane_kernels.cu): 1 = launched, 0 = declined, -1 = launch error
extern "C" int p2c_renorm_lane(const float* in, const float* g_out, float* out, int64_t m, int32_t j, int32_t dim,
                               int32_t hips, int32_t neck, int32_t autonormalize, const float* ref_shift,
                               const float* ref_scale, const int32_t* ref_index, int32_t clip_len, int backward,
                               void* stream);

extern "C" int p2c_cum_changes_fwd(const float* changes, float* cumulative, int32_t n_clips, int32_t clip_len,
                                   int32_t n_bones, void* stream) {
    if (n_clips < 0 || clip_len < 1 || n_bones < 1) return P2C_E_BADARG;
    const long units = (long)n_clips * n_bones;
    if (units == 0) return 0;  // empty batch: nothing to read or write (pointers of empty tensors are NULL)
    if (!changes || !cumulative) return P2C_E_NULL;
    const bool aligned = ((reinterpret_cast<uintptr_t>(changes) | reinterpret_cast<uintptr_t>(cumulative)) & 7) == 0;
    if (aligned && (n_bones * 9) % 2 == 0 && CK * n_bones <= CT && CK * n_bones * 9 / 2 <= CT * CPF) {
        int ck = CK;   // fewer clips per CTA while the grid would leave SMs idle
        while (ck > 1 && (n_clips + ck - 1) / ck < 2 * 148) ck >>= 1;
        const size_t smem = (size_t)ck * n_bones * 9 * sizeof(float);
        cum_changes_fwd_staged_kernel<<<(unsigned)((n_clips + ck - 1) / ck), CT, smem, (cudaStream_t)stream>>>(
            changes, cumulative, n_clips, clip_len, n_bones, ck);
        P2C_LAUNCH_CHECK();
        return 0;
    }
    cum_changes_fwd_kernel<<<(unsigned)((units + 127) / 128), 128, 0, (cudaStream_t)stream>>>(changes, cumulative,
                                                                                             units, clip_len, n_bones);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_cum_changes_bwd(const float* changes, const float* cumulative, const float* g_cumulative,
                                   float* g_changes, int32_t n_clips, int32_t clip_len, int32_t n_bones,
                                   void* stream) {
    if (n_clips < 0 || clip_len < 1 || n_bones < 1) return P2C_E_BADARG;
    const long units = (long)n_clips * n_bones;
    if (units == 0) return 0;
    if (!changes || !cumulative || !g_cumulative || !g_changes) return P2C_E_NULL;
    const bool aligned = ((reinterpret_cast<uintptr_t>(changes) | reinterpret_cast<uintptr_t>(cumulative) |
                           reinterpret_cast<uintptr_t>(g_cumulative) | reinterpret_cast<uintptr_t>(g_changes)) & 7) == 0;
    if (aligned && (n_bones * 9) % 2 == 0 && CK * n_bones <= CT && CK * n_bones * 9 / 2 <= CT * CPF) {
        int ck = CK;
        while (ck > 1 && (n_clips + ck - 1) / ck < 2 * 148) ck >>= 1;
        const size_t smem = (size_t)3 * ck * n_bones * 9 * sizeof(float);
        cum_changes_bwd_staged_kernel<<<(unsigned)((n_clips + ck - 1) / ck), CT, smem, (cudaStream_t)stream>>>(
            changes, cumulative, g_cumulative, g_changes, n_clips, clip_len, n_bones, ck);
        P2C_LAUNCH_CHECK();
        return 0;
    }
    cum_changes_bwd_kernel<<<(unsigned)((units + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        changes, cumulative, g_cumulative, g_changes, units, clip_len, n_bones);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_renormalize_fwd(const float* sample, float* out, int64_t m, int32_t j, int32_t c, int32_t dim,
                                   int32_t hips, int32_t neck, int32_t autonormalize, float near_zero,
                                   const float* ref_shift, const float* ref_scale, const int32_t* ref_index,
                                   int32_t clip_len, void* stream) {
    Renorm a;
    const int rc = renorm_args(a, m, j, c, dim, hips, neck, autonormalize, near_zero, ref_shift, ref_scale, ref_index,
                               clip_len);
    if (rc) return rc;
    if (m == 0) return 0;
    if (!sample || !out) return P2C_E_NULL;
    if (c == dim) {   // no carried channel: the chunked float4 variant (the pose-per-lane forward measured the same)
        if (dim == 2)
            renorm_flat_fwd_kernel<2><<<chunk_grid(m), RTH, 0, (cudaStream_t)stream>>>(sample, out, a);
        else
            renorm_flat_fwd_kernel<3><<<chunk_grid(m), RTH, 0, (cudaStream_t)stream>>>(sample, out, a);
        P2C_LAUNCH_CHECK();
        return 0;
    }
    if (dim == 2)
        renorm_fwd_kernel<2><<<warp_grid(m), WARPS * 32, 0, (cudaStream_t)stream>>>(sample, out, a);
    else
        renorm_fwd_kernel<3><<<warp_grid(m), WARPS * 32, 0, (cudaStream_t)stream>>>(sample, out, a);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_renormalize_bwd(const float* sample, const float* g_out, float* g_sample, int64_t m, int32_t j,
                                   int32_t c, int32_t dim, int32_t hips, int32_t neck, int32_t autonormalize,
                                   float near_zero, const float* ref_scale, const int32_t* ref_index,
                                   int32_t clip_len, void* stream) {
    Renorm a;
    // the reference shift does not enter the gradient; pass the scale pointer for both so the pair check holds
    const int rc = renorm_args(a, m, j, c, dim, hips, neck, autonormalize, near_zero, ref_scale, ref_scale, ref_index,
                               clip_len);
    if (rc) return rc;
    if (m == 0) return 0;
    if (!sample || !g_out || !g_sample) return P2C_E_NULL;
    if (c == dim) {
        const int took = p2c_renorm_lane(sample, g_out, g_sample, m, j, dim, hips, neck, autonormalize, nullptr, ref_scale,
                                         ref_index, clip_len, 1, stream);
        if (took > 0) return 0;
        if (took < 0) return (int)cudaErrorLaunchFailure;
    }
    if (c == dim && (size_t)2 * RP * j * dim * sizeof(float) <= 40 * 1024) {   // else: the warp-per-pose kernels
        a.ref_shift = nullptr;   // the shift does not enter the gradient
        if (dim == 2)
            renorm_flat_bwd_kernel<2><<<chunk_grid(m), RTH, (size_t)2 * RP * j * 2 * sizeof(float), (cudaStream_t)stream>>>(
                sample, g_out, g_sample, a);
        else
            renorm_flat_bwd_kernel<3><<<chunk_grid(m), RTH, (size_t)2 * RP * j * 3 * sizeof(float), (cudaStream_t)stream>>>(
                sample, g_out, g_sample, a);
        P2C_LAUNCH_CHECK();
        return 0;
    }
    if (dim == 2)
        renorm_bwd_kernel<2><<<warp_grid(m), WARPS * 32, 0, (cudaStream_t)stream>>>(sample, g_out, g_sample, a);
    else
        renorm_bwd_kernel<3><<<warp_grid(m), WARPS * 32, 0, (cudaStream_t)stream>>>(sample, g_out, g_sample, a);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_quaternion_to_matrix_fwd(const float* quat, float* out, int64_t n, void* stream) {
    if (n < 0) return P2C_E_BADARG;
    if (n == 0) return 0;
    if (!quat || !out) return P2C_E_NULL;
    if (reinterpret_cast<uintptr_t>(quat) & 15) return P2C_E_UNSUPPORTED;  // quaternions are read as float4
    quat_fwd_kernel<<<(unsigned)((n + QT - 1) / QT), QT, 0, (cudaStream_t)stream>>>(quat, out, (long)n);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_quaternion_to_matrix_bwd(const float* quat, const float* g_out, float* g_quat, int64_t n,
                                            void* stream) {
    if (n < 0) return P2C_E_BADARG;
    if (n == 0) return 0;
    if (!quat || !g_out || !g_quat) return P2C_E_NULL;
    if ((reinterpret_cast<uintptr_t>(quat) | reinterpret_cast<uintptr_t>(g_quat)) & 15) return P2C_E_UNSUPPORTED;
    quat_bwd_kernel<<<(unsigned)((n + QT - 1) / QT), QT, 0, (cudaStream_t)stream>>>(quat, g_out, g_quat, (long)n);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_normalize_extractor_fwd(const float* sample, float* out, int64_t m, int32_t j, int32_t c,
                                           int32_t hips, int32_t neck, int32_t extractor, float near_zero,
                                           void* stream) {
    if (m < 0 || j < 1 || (c != 2 && c != 3) || (extractor != 1 && extractor != 2)) return P2C_E_BADARG;
    if (hips < 0 || hips >= j || neck < 0 || neck >= j) return P2C_E_BADARG;
    if (m == 0) return 0;
    if (!sample || !out) return P2C_E_NULL;
    normalize_extractor_kernel<<<warp_grid(m), WARPS * 32, 0, (cudaStream_t)stream>>>(sample, out, (long)m, j, c, hips,
                                                                                   neck, extractor, near_zero);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_extractor_shift_scale(const float* sample, float* shift, float* scale, int64_t m, int32_t j, int32_t c,
                                         int32_t dim, int32_t hips, int32_t neck, int32_t extractor, float near_zero,
                                         void* stream) {
    if (m < 0 || j < 1 || (dim != 2 && dim != 3) || c < dim || extractor < 0 || extractor > 2) return P2C_E_BADARG;
    if (extractor != 0 && dim != 2) return P2C_E_BADARG;
    if (hips < 0 || hips >= j || neck < 0 || neck >= j) return P2C_E_BADARG;
    if (m == 0) return 0;
    if (!sample || !shift || !scale) return P2C_E_NULL;
    shift_scale_kernel<<<warp_grid(m), WARPS * 32, 0, (cudaStream_t)stream>>>(sample, shift, scale, (long)m, j, c, dim, hips,
                                                                           neck, extractor, near_zero);
    P2C_LAUNCH_CHECK();
    return 0;
}
