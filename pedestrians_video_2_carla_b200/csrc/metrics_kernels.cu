// Evaluation metrics on the buffers the fused step leaves in HBM:
//   MPJPE  metrics/mpjpe.py:29-45   sum over (pose, joint) of |pred - gt|_2 of absolute_pose_loc
//   MRPE   metrics/mrpe.py:38-76    sum over poses of |(world_loc + hips)_pred - (world_loc + hips)_gt|_2
//   PCK    metrics/pck.py:54-97     joints whose normalised distance is below a threshold; normalisation by the
//                                   target's hips-neck distance ('hn') or bounding-box diagonal ('bbox',
//                                   utils/tensors.py:12-26); missing-joints mask utils/tensors.py:29-40
// One warp per pose, lanes stride over joints; per-CTA partials in a caller-owned workspace, folded in fixed order in
// double by a one-warp kernel => run-to-run bitwise deterministic.  Streaming: each operand is read once.
#include "common.cuh"

namespace {

constexpr int WARPS = 4;
constexpr int MAX_BLOCKS = 148 * 16;
constexpr float FINF = __builtin_huge_valf();

struct MetricArgs {
    const float *pred, *gt, *pred_world, *gt_world, *mask_src;
    long M;
    int J, C, Cm, hips, neck, mask_hips, kind, norm;
    float threshold, near_zero;
};

__device__ __forceinline__ float warp_min(float v) {
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(FULL_MASK, v, o));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(FULL_MASK, v, o));
    return v;
}

__global__ void __launch_bounds__(WARPS * 32) metric_partial_kernel(MetricArgs a, float* __restrict__ partials) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int C = a.C;
    float s0 = 0.f, s1 = 0.f;  // (sum, count) or (correct, total)
    if (a.kind == 0) {  // MPJPE has no per-pose structure: a flat grid-stride loop over joints
        const long total = a.M * a.J;
        const long tid = (long)blockIdx.x * (WARPS * 32) + threadIdx.x, nthreads = (long)gridDim.x * (WARPS * 32);
        long done = 0;
        if (C == 3 && ((reinterpret_cast<uintptr_t>(a.pred) | reinterpret_cast<uintptr_t>(a.gt)) & 15) == 0) {
            // four joints = 48 bytes = three float4 per operand and thread: six 16-byte loads in flight per thread
            const long groups = total >> 2;
            const float4* p4 = reinterpret_cast<const float4*>(a.pred);
            const float4* g4 = reinterpret_cast<const float4*>(a.gt);
            for (long i = tid; i < groups; i += nthreads) {
                const float4 p0 = __ldg(p4 + 3 * i), p1 = __ldg(p4 + 3 * i + 1), p2 = __ldg(p4 + 3 * i + 2);
                const float4 g0 = __ldg(g4 + 3 * i), g1 = __ldg(g4 + 3 * i + 1), g2 = __ldg(g4 + 3 * i + 2);
                const float d[12] = {p0.x - g0.x, p0.y - g0.y, p0.z - g0.z, p0.w - g0.w, p1.x - g1.x, p1.y - g1.y,
                                     p1.z - g1.z, p1.w - g1.w, p2.x - g2.x, p2.y - g2.y, p2.z - g2.z, p2.w - g2.w};
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    s0 += sqrtf(fmaf(d[3 * j + 2], d[3 * j + 2], fmaf(d[3 * j + 1], d[3 * j + 1], d[3 * j] * d[3 * j])));
                s1 += 4.f;
            }
            done = groups << 2;
        }
        for (long i = done + tid; i < total; i += nthreads) {
            float q = 0.f;
            for (int c = 0; c < C; ++c) { const float d = __ldg(a.pred + i * C + c) - __ldg(a.gt + i * C + c); q = fmaf(d, d, q); }
            s0 += sqrtf(q);
            s1 += 1.f;
        }
    }
    for (long m = (long)blockIdx.x * WARPS + warp; a.kind != 0 && m < a.M; m += (long)gridDim.x * WARPS) {
        const float* p = a.pred + m * a.J * C;
        const float* g = a.gt + m * a.J * C;
        if (a.kind == 1) {  // MRPE
            if (lane == 0) {
                float q = 0.f;
                for (int c = 0; c < C; ++c) {
                    const float pw = a.pred_world ? a.pred_world[m * C + c] : 0.f;
                    const float gw = a.gt_world ? a.gt_world[m * C + c] : 0.f;
                    const float d = (pw + p[a.hips * C + c]) - (gw + g[a.hips * C + c]);
                    q = fmaf(d, d, q);
                }
                s0 += sqrtf(q);
                s1 += 1.f;
            }
        } else if (a.J <= 32 && C <= 3) {  // PCK, lane = joint: every operand is read once into registers
            const bool act = lane < a.J;
            float gv[3] = {0.f, 0.f, 0.f}, pv[3] = {0.f, 0.f, 0.f};
#pragma unroll
            for (int c = 0; c < 3; ++c)
                if (c < C && act) { gv[c] = __ldg(g + lane * C + c); pv[c] = __ldg(p + lane * C + c); }
            bool kept = act;
            if (a.mask_src && act && lane != a.mask_hips) {
                const float* ms = a.mask_src + (m * a.J + lane) * a.Cm;
                for (int c = 0; c < a.Cm; ++c) kept = kept && (__ldg(ms + c) != 0.f);
            }
            float q = 0.f;
            if (a.norm == 0) {  // 'hn'
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float d = __shfl_sync(FULL_MASK, gv[c], a.neck) - __shfl_sync(FULL_MASK, gv[c], a.hips);
                    if (c < C) q = fmaf(d, d, q);
                }
            } else {  // 'bbox' around the detected points of the target (all of x, y < near_zero = missing)
                const bool detected = act && !(gv[0] < a.near_zero && gv[1] < a.near_zero);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float d = warp_max(detected ? gv[c] : -FINF) - warp_min(detected ? gv[c] : FINF);
                    if (c < C) q = fmaf(d, d, q);
                }
            }
            float norm = sqrtf(q);
            if (norm < a.near_zero) { kept = false; norm = 1.f; }   // mask[normalize < near_zero] = 0
            if (kept) {
                float e2 = 0.f;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float d = (pv[c] - gv[c]) / norm;
                    if (c < C) e2 = fmaf(d, d, e2);
                }
                if (sqrtf(e2) < a.threshold) s0 += 1.f;
                s1 += 1.f;
            }
        } else {  // PCK, any joint count
            float norm;
            if (a.norm == 0) {  // 'hn': |neck - hips| of the target
                float q = 0.f;
                for (int c = 0; c < C; ++c) { const float d = g[a.neck * C + c] - g[a.hips * C + c]; q = fmaf(d, d, q); }
                norm = sqrtf(q);
            } else {  // 'bbox': diagonal of the box around the target's detected points (all of x, y < near_zero = missing)
                float q = 0.f;
                for (int c = 0; c < C; ++c) {
                    float lo = FINF, hi = -FINF;
                    for (int j = lane; j < a.J; j += 32) {
                        const bool missing = g[j * C] < a.near_zero && g[j * C + 1] < a.near_zero;
                        if (!missing) { lo = fminf(lo, g[j * C + c]); hi = fmaxf(hi, g[j * C + c]); }
                    }
                    const float d = warp_max(hi) - warp_min(lo);
                    q = fmaf(d, d, q);
                }
                norm = sqrtf(q);
            }
            const bool pose_ok = !(norm < a.near_zero);   // mask[normalize < near_zero] = 0 (NaN norms stay in)
            if (norm < a.near_zero) norm = 1.f;
            const float* ms = a.mask_src ? a.mask_src + m * a.J * a.Cm : nullptr;
            for (int j = lane; j < a.J; j += 32) {
                bool kept = pose_ok;
                if (ms && j != a.mask_hips) {
                    for (int c = 0; c < a.Cm; ++c) kept = kept && (ms[j * a.Cm + c] != 0.f);
                }
                if (kept) {
                    float q = 0.f;
                    for (int c = 0; c < C; ++c) { const float d = (p[j * C + c] - g[j * C + c]) / norm; q = fmaf(d, d, q); }
                    if (sqrtf(q) < a.threshold) s0 += 1.f;
                    s1 += 1.f;
                }
            }
        }
    }
    __shared__ float sa[WARPS], sb[WARPS];
    s0 = warp_sum(s0);
    s1 = warp_sum(s1);
    if (lane == 0) { sa[warp] = s0; sb[warp] = s1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float x = 0.f, y = 0.f;
        for (int w = 0; w < WARPS; ++w) { x += sa[w]; y += sb[w]; }
        partials[blockIdx.x * 2] = x;
        partials[blockIdx.x * 2 + 1] = y;
    }
}

__global__ void metric_finalize_kernel(const float* __restrict__ partials, int nblocks, double* __restrict__ out,
                                       int accumulate) {
    double s = 0.0, c = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 32) { s += (double)partials[i * 2]; c += (double)partials[i * 2 + 1]; }
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(FULL_MASK, s, o);
        c += __shfl_xor_sync(FULL_MASK, c, o);
    }
    if (threadIdx.x == 0) {
        out[0] = (accumulate ? out[0] : 0.0) + s;
        out[1] = (accumulate ? out[1] : 0.0) + c;
    }
}

int run(const MetricArgs& a, float* partials, double* out, int accumulate, void* stream) {
    if (!partials || !out) return P2C_E_NULL;
    long blocks = (a.M + WARPS - 1) / WARPS;
    if (blocks > MAX_BLOCKS) blocks = MAX_BLOCKS;
    if (blocks < 1) blocks = 1;
    metric_partial_kernel<<<(int)blocks, WARPS * 32, 0, (cudaStream_t)stream>>>(a, partials);
    P2C_LAUNCH_CHECK();
    metric_finalize_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(partials, (int)blocks, out, accumulate);
    P2C_LAUNCH_CHECK();
    return 0;
}

}  // namespace

extern "C" int64_t p2c_metric_workspace_floats(void) { return 2 * MAX_BLOCKS; }

extern "C" int p2c_mpjpe_sum(const float* pred, const float* gt, int64_t m, int32_t j, int32_t c, float* partials,
                             double* out, int32_t accumulate, void* stream) {
    if (m < 0 || j < 1 || c < 1) return P2C_E_BADARG;
    if (m > 0 && (!pred || !gt)) return P2C_E_NULL;
    MetricArgs a{};
    a.pred = pred; a.gt = gt; a.M = (long)m; a.J = j; a.C = c; a.kind = 0;
    return run(a, partials, out, accumulate, stream);
}

extern "C" int p2c_mrpe_sum(const float* pred, const float* gt, const float* pred_world, const float* gt_world,
                            int64_t m, int32_t j, int32_t c, int32_t hips, float* partials, double* out,
                            int32_t accumulate, void* stream) {
    if (m < 0 || j < 1 || c < 1 || hips < 0 || hips >= j) return P2C_E_BADARG;
    if (m > 0 && (!pred || !gt)) return P2C_E_NULL;
    MetricArgs a{};
    a.pred = pred; a.gt = gt; a.pred_world = pred_world; a.gt_world = gt_world;
    a.M = (long)m; a.J = j; a.C = c; a.hips = hips; a.kind = 1;
    return run(a, partials, out, accumulate, stream);
}

extern "C" int p2c_pck_count(const float* pred, const float* gt, const float* mask_src, int64_t m, int32_t j,
                             int32_t c, int32_t c_mask, int32_t mask_hips, int32_t hips, int32_t neck,
                             int32_t norm_kind, float threshold, float near_zero, float* partials, double* out,
                             int32_t accumulate, void* stream) {
    if (m < 0 || j < 1 || c < 2 || (norm_kind != 0 && norm_kind != 1)) return P2C_E_BADARG;
    if (hips < 0 || hips >= j || neck < 0 || neck >= j || mask_hips >= j) return P2C_E_BADARG;
    if (mask_src && c_mask < 1) return P2C_E_BADARG;
    if (m > 0 && (!pred || !gt)) return P2C_E_NULL;
    MetricArgs a{};
    a.pred = pred; a.gt = gt; a.mask_src = mask_src; a.M = (long)m; a.J = j; a.C = c; a.Cm = c_mask;
    a.hips = hips; a.neck = neck; a.mask_hips = mask_hips; a.kind = 2; a.norm = norm_kind;
    a.threshold = threshold; a.near_zero = near_zero;
    return run(a, partials, out, accumulate, stream);
}
