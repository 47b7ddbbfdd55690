// Hips/neck normaliser and (masked) MSE, forward + backward.
//   Normalizer(HipsNeckExtractor): transforms/pose/normalization/normalizer.py:20-41, extractor.py:23-36,
//                                  hips_neck_extractor.py:6-13; nan_to_zero utils/tensors.py:43-53
//   Loc2DPoseLoss / loc_3d with nn.MSELoss('mean'): loss/base_pose_loss.py:36-66, loss/loc_2d.py:69-89,
//                                  loss/loc_3d.py:12-40; mask utils/tensors.py:29-40
// Reductions are two-stage with a fixed grid and fixed summation order => bitwise run-to-run deterministic.
#include "common.cuh"

namespace {

constexpr int WARPS = 4;
constexpr int MSE_BLOCKS = 592;   // 148 SMs x 4
constexpr int MSE_THREADS = 256;

// one warp per pose; lanes stride over joints
__global__ void __launch_bounds__(WARPS * 32) normalize_fwd_kernel(const float* __restrict__ in,
                                                                  float* __restrict__ out, long M, int J, int C,
                                                                  int hips, int neck, float near_zero) {
    const int lane = threadIdx.x & 31;
    const long wstride = (long)gridDim.x * WARPS;
    for (long m = (long)blockIdx.x * WARPS + (threadIdx.x >> 5); m < M; m += wstride) {
        const float* p = in + m * J * C;
        float* o = out + m * J * C;
        const float sx = __ldg(p + hips * C), sy = __ldg(p + hips * C + 1);
        const float dx = __ldg(p + neck * C) - sx, dy = __ldg(p + neck * C + 1) - sy;
        const float scale = sqrtf(dx * dx + dy * dy);
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

__global__ void __launch_bounds__(WARPS * 32) normalize_bwd_kernel(const float* __restrict__ in,
                                                                  const float* __restrict__ g_out,
                                                                  float* __restrict__ g_in, long M, int J, int C,
                                                                  int hips, int neck, float near_zero) {
    const int lane = threadIdx.x & 31;
    const long wstride = (long)gridDim.x * WARPS;
    for (long m = (long)blockIdx.x * WARPS + (threadIdx.x >> 5); m < M; m += wstride) {
        const float* p = in + m * J * C;
        const float* g = g_out + m * J * C;
        float* o = g_in + m * J * C;
        const float sx = __ldg(p + hips * C), sy = __ldg(p + hips * C + 1);
        const float dx = __ldg(p + neck * C) - sx, dy = __ldg(p + neck * C + 1) - sy;
        const float scale = sqrtf(dx * dx + dy * dy);
        const bool ok = scale > 0.f && fabsf(scale) <= 3.402823466e38f;
        float sgx = 0.f, sgy = 0.f, sgn = 0.f;
        for (int j = lane; j < J; j += 32) {
            const float rx = (p[j * C] - sx) / scale, ry = (p[j * C + 1] - sy) / scale;
            float gx = g[j * C], gy = g[j * C + 1];
            if (!(fabsf(rx) <= 3.402823466e38f)) gx = 0.f;  // nan_to_num passes no gradient to non-finite entries
            if (!(fabsf(ry) <= 3.402823466e38f)) gy = 0.f;
            if (C > 2) {
                const float qraw = p[j * C + 2];
                const bool fin = fabsf(qraw) <= 3.402823466e38f;
                const float q = fin ? qraw : 0.f;
                if (!(q >= near_zero)) { gx = 0.f; gy = 0.f; }
                o[j * C + 2] = fin ? g[j * C + 2] : 0.f;
            }
            if (!ok) { gx = 0.f; gy = 0.f; }
            o[j * C] = ok ? gx / scale : 0.f;
            o[j * C + 1] = ok ? gy / scale : 0.f;
            sgx += gx; sgy += gy;
            sgn += gx * nan_to_zero(rx) + gy * nan_to_zero(ry);
        }
        sgx = warp_sum(sgx); sgy = warp_sum(sgy); sgn = warp_sum(sgn);
        __syncwarp();
        if (lane == 0 && ok) {
            const float gscale = -sgn / scale;            // d/d scale
            const float ux = dx / scale, uy = dy / scale;  // d scale / d (neck - hips)
            float hx = -sgx / scale - gscale * ux, hy = -sgy / scale - gscale * uy;
            float kx = gscale * ux, ky = gscale * uy;
            if (hips == neck) { hx += kx; hy += ky; kx = 0.f; ky = 0.f; }
            o[hips * C] += hx; o[hips * C + 1] += hy;
            o[neck * C] += kx; o[neck * C + 1] += ky;
        }
    }
}

// utils/tensors.py:29-40: joint kept iff every used gt channel != 0, hips always kept
__device__ __forceinline__ bool joint_kept(const float* gt, int c_used, int j, int mask_hips) {
    if (mask_hips < 0 || j == mask_hips) return true;
    bool k = true;
    for (int c = 0; c < c_used; ++c) k = k && (gt[c] != 0.f);
    return k;
}

__global__ void __launch_bounds__(MSE_THREADS) mse_partial_kernel(const float* __restrict__ pred,
                                                                 const float* __restrict__ gt, long total, int J,
                                                                 int c_pred, int c_gt, int c_used, int mask_hips,
                                                                 float* __restrict__ partials) {
    float s = 0.f, cnt = 0.f;
    for (long i = (long)blockIdx.x * MSE_THREADS + threadIdx.x; i < total; i += (long)gridDim.x * MSE_THREADS) {
        const int j = (int)(i % J);
        const float* g = gt + i * c_gt;
        if (joint_kept(g, c_used, j, mask_hips)) {
            const float* p = pred + i * c_pred;
            for (int c = 0; c < c_used; ++c) {
                const float d = p[c] - g[c];
                s = fmaf(d, d, s);
            }
            cnt += (float)c_used;
        }
    }
    __shared__ float ss[MSE_THREADS / 32], sc[MSE_THREADS / 32];
    s = warp_sum(s);
    cnt = warp_sum(cnt);
    if ((threadIdx.x & 31) == 0) { ss[threadIdx.x >> 5] = s; sc[threadIdx.x >> 5] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f, b = 0.f;
        for (int w = 0; w < MSE_THREADS / 32; ++w) { a += ss[w]; b += sc[w]; }
        partials[blockIdx.x * 2] = a;
        partials[blockIdx.x * 2 + 1] = b;
    }
}

__global__ void mse_finalize_kernel(const float* __restrict__ partials, int nblocks, float* __restrict__ out) {
    // single warp, fixed order, double accumulation
    double s = 0.0, c = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 32) { s += (double)partials[i * 2]; c += (double)partials[i * 2 + 1]; }
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(FULL_MASK, s, o);
        c += __shfl_xor_sync(FULL_MASK, c, o);
    }
    if (threadIdx.x == 0) {
        out[0] = (float)(s / c);  // 0/0 -> NaN like nn.MSELoss on an empty selection
        out[1] = (float)s;
        out[2] = (float)c;
    }
}

__global__ void __launch_bounds__(MSE_THREADS) mse_bwd_kernel(const float* __restrict__ pred,
                                                             const float* __restrict__ gt,
                                                             const float* __restrict__ out,
                                                             const float* __restrict__ g_loss,
                                                             float* __restrict__ g_pred, long total, int J, int c_pred,
                                                             int c_gt, int c_used, int mask_hips) {
    const float k = (g_loss ? __ldg(g_loss) : 1.0f) * 2.0f / __ldg(out + 2);
    for (long i = (long)blockIdx.x * MSE_THREADS + threadIdx.x; i < total; i += (long)gridDim.x * MSE_THREADS) {
        const int j = (int)(i % J);
        const float* g = gt + i * c_gt;
        const bool kept = joint_kept(g, c_used, j, mask_hips);
        const float* p = pred + i * c_pred;
        float* o = g_pred + i * c_pred;
        for (int c = 0; c < c_pred; ++c) o[c] = (kept && c < c_used) ? k * (p[c] - g[c]) : 0.f;
    }
}

int mse_blocks(long total) {
    long b = (total + MSE_THREADS - 1) / MSE_THREADS;
    if (b > MSE_BLOCKS) b = MSE_BLOCKS;
    if (b < 1) b = 1;
    return (int)b;
}
int warp_grid(long units) {
    long blocks = (units + WARPS - 1) / WARPS;
    if (blocks > 148L * 64) blocks = 148L * 64;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace

extern "C" int p2c_hips_neck_normalize_fwd(const float* sample, float* out, int64_t m, int32_t j, int32_t c,
                                           int32_t hips, int32_t neck, float near_zero, void* stream) {
    if (!sample || !out) return P2C_E_NULL;
    if (m < 0 || j < 1 || (c != 2 && c != 3) || hips < 0 || hips >= j || neck < 0 || neck >= j) return P2C_E_BADARG;
    if (m == 0) return 0;
    normalize_fwd_kernel<<<warp_grid(m), WARPS * 32, 0, (cudaStream_t)stream>>>(sample, out, (long)m, j, c, hips, neck,
                                                                             near_zero);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_hips_neck_normalize_bwd(const float* sample, const float* g_out, float* g_sample, int64_t m,
                                           int32_t j, int32_t c, int32_t hips, int32_t neck, float near_zero,
                                           void* stream) {
    if (!sample || !g_out || !g_sample) return P2C_E_NULL;
    if (m < 0 || j < 1 || (c != 2 && c != 3) || hips < 0 || hips >= j || neck < 0 || neck >= j) return P2C_E_BADARG;
    if (m == 0) return 0;
    normalize_bwd_kernel<<<warp_grid(m), WARPS * 32, 0, (cudaStream_t)stream>>>(sample, g_out, g_sample, (long)m, j, c,
                                                                             hips, neck, near_zero);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int64_t p2c_mse_workspace_floats(int64_t m) {
    (void)m;
    return 2 * MSE_BLOCKS;
}

extern "C" int p2c_masked_mse_fwd(const float* pred, const float* gt, int64_t m, int32_t j, int32_t c_pred,
                                  int32_t c_gt, int32_t c_used, int32_t mask_hips, float* partials, float* out,
                                  void* stream) {
    if (!pred || !gt || !partials || !out) return P2C_E_NULL;
    if (m < 0 || j < 1 || c_used < 1 || c_used > c_pred || c_used > c_gt || mask_hips >= j) return P2C_E_BADARG;
    const long total = (long)m * j;
    const int nb = mse_blocks(total);
    mse_partial_kernel<<<nb, MSE_THREADS, 0, (cudaStream_t)stream>>>(pred, gt, total, j, c_pred, c_gt, c_used,
                                                                     mask_hips, partials);
    P2C_LAUNCH_CHECK();
    mse_finalize_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(partials, nb, out);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_masked_mse_bwd(const float* pred, const float* gt, const float* out, const float* g_loss,
                                  float* g_pred, int64_t m, int32_t j, int32_t c_pred, int32_t c_gt, int32_t c_used,
                                  int32_t mask_hips, void* stream) {
    if (!pred || !gt || !out || !g_pred) return P2C_E_NULL;
    if (m < 0 || j < 1 || c_used < 1 || c_used > c_pred || c_used > c_gt || mask_hips >= j) return P2C_E_BADARG;
    if (m == 0) return 0;
    const long total = (long)m * j;
    mse_bwd_kernel<<<mse_blocks(total), MSE_THREADS, 0, (cudaStream_t)stream>>>(pred, gt, out, g_loss, g_pred, total, j,
                                                                                c_pred, c_gt, c_used, mask_hips);
    P2C_LAUNCH_CHECK();
    return 0;
}
