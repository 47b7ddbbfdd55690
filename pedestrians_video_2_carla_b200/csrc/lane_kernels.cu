// Pose-per-lane streaming kernel for the per-pose passes either side of the step: Normalizer / DeNormalizer /
// ReferenceSkeletonsDeNormalizer (transforms/pose/normalization/normalizer.py:20-41, denormalizer.py:13-37,
// reference_skeletons_denormalizer.py:41-91); used for the BACKWARD (the forward measured equal to the chunked float4
// kernel of widen_kernels.cu; a pose-per-lane PCK measured slower than the joint-per-lane one -- three staged tiles per
// warp leave too few warps per SM -- and was dropped).
//
// The warp-per-pose versions (widen_kernels.cu, metrics_kernels.cu) spend ~80 set-up instructions per pose (hips / neck
// reads, the scale, a warp reduction per sum) on 26 of 32 lanes for 78 floats of payload: instruction bound at 24-29 % of
// the HBM roofline.  Here a warp owns a TILE of 32 consecutive poses: the tile's records are staged into shared memory
// by coalesced 128-bit copies (every load of the tile in flight together), then every LANE works on ITS OWN pose --
// constants, sums over joints and the output pass are lane-local loops: no shuffles, no per-pose set-up shared by a
// warp, all 32 lanes busy.  Records are read as 128-bit (2-channel) / 64-bit (3-channel) accesses whose lane stride
// (13 / 39 vector units for 26 joints) is bank-conflict free; results are written in place and leave by the same
// coalesced copies.  Taken when sample and result have the same channel count and the joint count is even; anything
// else stays on the warp-per-pose kernels.
#include "common.cuh"
#include "host_util.h"
#include "tile_common.cuh"

namespace {

constexpr float LFMAX = 3.402823466e38f;

// tile <-> global, one warp: flat coalesced copies, widest vector the pointers allow
__device__ __forceinline__ void warp_copy_in(float* s, const float* g, int n, int lane) {
    if (((reinterpret_cast<uintptr_t>(g) & 15) == 0) && ((n & 3) == 0)) {
        const float4* g4 = reinterpret_cast<const float4*>(g);
        float4* s4 = reinterpret_cast<float4*>(s);
        for (int i = lane; i < (n >> 2); i += 32) s4[i] = __ldcs(g4 + i);
    } else if (((reinterpret_cast<uintptr_t>(g) & 7) == 0) && ((n & 1) == 0)) {
        const float2* g2 = reinterpret_cast<const float2*>(g);
        float2* s2 = reinterpret_cast<float2*>(s);
        for (int i = lane; i < (n >> 1); i += 32) s2[i] = __ldcs(g2 + i);
    } else {
        for (int i = lane; i < n; i += 32) s[i] = __ldcs(g + i);
    }
}
__device__ __forceinline__ void warp_copy_out(float* g, const float* s, int n, int lane) {
    if (((reinterpret_cast<uintptr_t>(g) & 15) == 0) && ((n & 3) == 0)) {
        float4* g4 = reinterpret_cast<float4*>(g);
        const float4* s4 = reinterpret_cast<const float4*>(s);
        for (int i = lane; i < (n >> 2); i += 32) __stcs(g4 + i, s4[i]);
    } else if (((reinterpret_cast<uintptr_t>(g) & 7) == 0) && ((n & 1) == 0)) {
        float2* g2 = reinterpret_cast<float2*>(g);
        const float2* s2 = reinterpret_cast<const float2*>(s);
        for (int i = lane; i < (n >> 1); i += 32) __stcs(g2 + i, s2[i]);
    } else {
        for (int i = lane; i < n; i += 32) __stcs(g + i, s[i]);
    }
}

struct LaneRenorm {
    long M;
    int J, L, hips, neck, autonormalize;
    const float* in;
    const float* g_out;   // backward only
    float* out;           // forward: result; backward: g_in
    const float* ref_shift;
    const float* ref_scale;
    const int* ref_index;
};

// visits the record of one pose as (element index, value pair / quad) groups with compile-time channel patterns
template <int DIM, typename F>
__device__ __forceinline__ void for_record(float* rec, int J, F f) {
    if (DIM == 2) {
        for (int u = 0; u < (J >> 1); ++u) {   // two joints per 128-bit access
            float4 v = lds128(rec + 4 * u);
            float e[4] = {v.x, v.y, v.z, v.w};
            f(e[0], 2 * u, 0); f(e[1], 2 * u, 1); f(e[2], 2 * u + 1, 0); f(e[3], 2 * u + 1, 1);
            sts128(rec + 4 * u, e[0], e[1], e[2], e[3]);
        }
    } else {
        for (int u = 0; u < (J >> 1); ++u) {   // two joints = three 64-bit accesses
            float2 a = lds64(rec + 6 * u), b = lds64(rec + 6 * u + 2), c = lds64(rec + 6 * u + 4);
            f(a.x, 2 * u, 0); f(a.y, 2 * u, 1); f(b.x, 2 * u, 2); f(b.y, 2 * u + 1, 0); f(c.x, 2 * u + 1, 1); f(c.y, 2 * u + 1, 2);
            sts64(rec + 6 * u, a.x, a.y); sts64(rec + 6 * u + 2, b.x, b.y); sts64(rec + 6 * u + 4, c.x, c.y);
        }
    }
}

template <int DIM, bool BWD, int WPB>
__global__ void __launch_bounds__(WPB * 32) renorm_lane_kernel(const LaneRenorm a) {
    extern __shared__ __align__(16) float smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int JC = a.J * DIM;
    float* X = smem + (size_t)warp * (BWD ? 2 : 1) * 32 * JC;   // this warp's tile of samples
    float* G = X + 32 * JC;                                    // (backward) upstream gradients -> result, in place
    const long tiles = (a.M + 31) >> 5;
    for (long tile = (long)blockIdx.x * WPB + warp; tile < tiles; tile += (long)gridDim.x * WPB) {
        const long m0 = tile << 5;
        const int np = (int)min(32L, a.M - m0);
        const int nf = np * JC;
        warp_copy_in(X, a.in + m0 * JC, nf, lane);
        if (BWD) warp_copy_in(G, a.g_out + m0 * JC, nf, lane);
        __syncwarp();
        if (lane < np) {
            float* x = X + lane * JC;
            float rs = 1.f, rh[DIM];
#pragma unroll
            for (int k = 0; k < DIM; ++k) rh[k] = 0.f;
            if (a.ref_scale) {
                const long clip = (m0 + lane) / a.L;
                const long row = a.ref_index ? (long)__ldg(a.ref_index + clip) : clip;
                rs = __ldg(a.ref_scale + row);
                if (!BWD) {
#pragma unroll
                    for (int k = 0; k < DIM; ++k) rh[k] = __ldg(a.ref_shift + row * DIM + k);
                }
            }
            float s[DIM], d[DIM], q2 = 0.f;
#pragma unroll
            for (int k = 0; k < DIM; ++k) {
                s[k] = x[a.hips * DIM + k];
                d[k] = x[a.neck * DIM + k] - s[k];
                q2 = fmaf(d[k], d[k], q2);
            }
            const float scale = sqrtf(q2);
            const float inv = 1.0f / scale;   // (scale 0 / NaN: inf / NaN, cut by nan_to_zero like the reference's division)
            if (!BWD) {
                if (a.autonormalize) {
                    for_record<DIM>(x, a.J, [&](float& v, int, int k) {
                        const float y = nan_to_zero((v - s[k]) * inv);
                        v = a.ref_scale ? fmaf(y, rs, rh[k]) : y;
                    });
                } else if (a.ref_scale) {
                    for_record<DIM>(x, a.J, [&](float& v, int, int k) { v = fmaf(v, rs, rh[k]); });
                }
            } else {
                float* g = G + lane * JC;
                if (!a.autonormalize) {
                    for_record<DIM>(g, a.J, [&](float& v, int, int) { v *= rs; });
                } else {
                    const bool ok = scale > 0.f && scale <= LFMAX;
                    float sg[DIM], sgn = 0.f;
#pragma unroll
                    for (int k = 0; k < DIM; ++k) sg[k] = 0.f;
                    // one pass: the element's own term in place, the three sums on the way (x is read through the
                    // element index: the two records share the layout)
                    for_record<DIM>(g, a.J, [&](float& v, int j, int k) {
                        const float r = (x[j * DIM + k] - s[k]) * inv;
                        float gy = v * rs;
                        if (!(fabsf(r) <= LFMAX) || !ok) gy = 0.f;   // nan_to_num passes no gradient to non-finite entries
                        v = ok ? gy * inv : 0.f;
                        sg[k] += gy;
                        sgn = fmaf(gy, nan_to_zero(r), sgn);
                    });
                    if (ok) {
                        const float gscale = -sgn * inv;   // d / d scale
#pragma unroll
                        for (int k = 0; k < DIM; ++k) {
                            const float un = gscale * d[k] * inv;   // through scale = |neck - hips|
                            if (a.hips == a.neck) {
                                g[a.hips * DIM + k] += -sg[k] * inv;
                            } else {
                                g[a.hips * DIM + k] += -sg[k] * inv - un;
                                g[a.neck * DIM + k] += un;
                            }
                        }
                    }
                }
            }
        }
        __syncwarp();
        warp_copy_out(a.out + m0 * JC, BWD ? G : X, nf, lane);
        __syncwarp();
    }
}

template <int DIM, bool BWD>
int launch_renorm(const LaneRenorm& a, void* stream) {
    constexpr int WPB = BWD ? 2 : 4;
    const size_t smem = (size_t)WPB * (BWD ? 2 : 1) * 32 * a.J * DIM * sizeof(float);
    if (smem > 96 * 1024) return 0;
    auto k = renorm_lane_kernel<DIM, BWD, WPB>;
    if (p2c_allow_big_smem(reinterpret_cast<const void*>(k), 96 * 1024)) return 0;
    const long tiles = (a.M + 31) >> 5;
    long blocks = (tiles + WPB - 1) / WPB;
    const long cap = (long)p2c_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    k<<<(int)blocks, WPB * 32, smem, (cudaStream_t)stream>>>(a);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace

// 1 = launched, 0 = declined (the caller falls back to its warp-per-pose kernel), -1 = launch error
extern "C" int p2c_renorm_lane(const float* in, const float* g_out, float* out, int64_t m, int32_t j, int32_t dim,
                               int32_t hips, int32_t neck, int32_t autonormalize, const float* ref_shift,
                               const float* ref_scale, const int32_t* ref_index, int32_t clip_len, int backward,
                               void* stream) {
    if ((j & 1) || m < 1) return 0;
    if ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(g_out)) & 15) return 0;
    LaneRenorm a{};
    a.M = (long)m; a.J = j; a.L = clip_len > 0 ? clip_len : 1; a.hips = hips; a.neck = neck;
    a.autonormalize = autonormalize ? 1 : 0;
    a.in = in; a.g_out = g_out; a.out = out; a.ref_shift = ref_shift; a.ref_scale = ref_scale; a.ref_index = ref_index;
    if (dim == 2) return backward ? launch_renorm<2, true>(a, stream) : launch_renorm<2, false>(a, stream);
    if (dim == 3) return backward ? launch_renorm<3, true>(a, stream) : launch_renorm<3, false>(a, stream);
    return 0;
}

