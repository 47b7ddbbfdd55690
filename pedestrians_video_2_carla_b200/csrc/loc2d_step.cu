// p2c_loc2d_step: the autoencoder flow's step on this path (BASELINE configs[2]): the model emits 2-D poses directly
// (modules/flow/autoencoder.py:102-112), the loss is Loc2DPoseLoss (loss/loc_2d.py:69-89 over loss/base_pose_loss.py:36-66
// with the missing-joints mask of utils/tensors.py:29-40), and loss.backward() needs d loss / d prediction.
//
// One COOPERATIVE launch does both: every CTA keeps the masked errors of its slice of the joints in shared memory while
// the grid agrees on the sums (per-CTA partials in a fixed order, folded in double by every CTA the same way), then
// writes the gradient 2 (pred - gt) / count from there.  HBM traffic = the algorithmic minimum: predictions and targets
// read once, gradient written once (624 B per 26-joint pose).  Run-to-run bitwise deterministic (fixed grid, fixed order).
// Batches too large for the errors to stay on chip (> ~145 k poses of 26 joints) take the two-pass kernels of
// loss_kernels.cu (the wrapper in ops.py picks).
#include <mutex>

#include "common.cuh"
#include "host_util.h"

namespace {

#ifndef P2C_LOC2D_THREADS
#define P2C_LOC2D_THREADS 1024
#endif
constexpr int LT = P2C_LOC2D_THREADS;
constexpr int MAX_CTAS = 1024;

struct Loc2dArgs {
    const float* pred;   // (joints, 2)
    const float* gt;     // (joints, c_gt)
    float* g_pred;       // (joints, 2)
    long joints;
    int J, c_gt, mask_hips, per_cta;
    float upstream;
    float* partials;     // 2 x grid
    unsigned* sync;      // [0] arrivals, [1] departures: self-resetting
    float* out;          // loss, sum_sq, count, 0
    XchgArgs xchg;
};

__device__ __forceinline__ bool kept2(float gx, float gy, int j, int hips) {
    return hips < 0 || j == hips || (gx != 0.f && gy != 0.f);
}

__global__ void __launch_bounds__(LT, 1) loc2d_step_kernel(const Loc2dArgs a) {
    extern __shared__ __align__(16) float esm[];   // masked errors of this CTA's joints, (x, y) pairs
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // Launched with programmatic stream serialization (p2c_loc2d_step): the previous kernel of the stream may still be
    // running.  Everything in global memory may come from it (predictions, the self-resetting barrier words): wait until
    // it is complete and flushed, then let the NEXT kernel's launch overlap this one.  (A no-op in a plain launch.)
    // (ahead of the wait: L2 prefetches of this CTA's slice -- they return no data, L2 is the point of coherence, they
    // only start the trip from HBM while the previous kernel drains)
    if (tid == 0) {
        const long pj0 = (long)blockIdx.x * a.per_cta, pleft = a.joints - pj0;
        const long pn = pleft <= 0 ? 0 : (pleft < a.per_cta ? pleft : a.per_cta);
        const unsigned bytes = (unsigned)((pn * 2 * sizeof(float)) & ~15ul);
        if (bytes) {
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a.pred + pj0 * 2), "r"(bytes) : "memory");
            if (a.c_gt == 2)
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a.gt + pj0 * 2), "r"(bytes) : "memory");
        }
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (blockIdx.x == 0 && tid == LT - 32) exchange_send_pending(a.xchg);   // the previous step's sums, off the tail
    const long j0 = (long)blockIdx.x * a.per_cta;   // (per_cta is even: pairs never straddle CTAs)
    long left = a.joints - j0;
    const int nj = left <= 0 ? 0 : (left < a.per_cta ? (int)left : a.per_cta);
    float s = 0.f, c = 0.f;
    const int pairs = nj >> 1;
    const float4* p4 = reinterpret_cast<const float4*>(a.pred + j0 * 2);
    float4* e4 = reinterpret_cast<float4*>(esm);
#pragma unroll 4   // (the loads of four trips in flight together: a CTA per SM needs ~64 KB outstanding to stream)
    for (int q = tid; q < pairs; q += LT) {
        const float4 p = __ldcs(p4 + q);   // streaming: read once
        float g0x, g0y, g1x, g1y;
        if (a.c_gt == 2) {
            const float4 g = __ldcs(reinterpret_cast<const float4*>(a.gt + j0 * 2) + q);
            g0x = g.x; g0y = g.y; g1x = g.z; g1y = g.w;
        } else {   // 3-channel target records: six words per pair, 8-byte aligned (even joint index)
            const float2* g2 = reinterpret_cast<const float2*>(a.gt + (j0 + 2 * (long)q) * 3);
            const float2 u = __ldcs(g2), v = __ldcs(g2 + 1), w = __ldcs(g2 + 2);
            g0x = u.x; g0y = u.y; g1x = v.y; g1y = w.x;
        }
        const int b0 = (int)((j0 + 2 * (long)q) % a.J), b1 = b0 + 1 == a.J ? 0 : b0 + 1;
        float4 e = make_float4(0.f, 0.f, 0.f, 0.f);
        if (kept2(g0x, g0y, b0, a.mask_hips)) {
            e.x = p.x - g0x; e.y = p.y - g0y;
            s = fmaf(e.x, e.x, fmaf(e.y, e.y, s));
            c += 2.f;
        }
        if (kept2(g1x, g1y, b1, a.mask_hips)) {
            e.z = p.z - g1x; e.w = p.w - g1y;
            s = fmaf(e.z, e.z, fmaf(e.w, e.w, s));
            c += 2.f;
        }
        e4[q] = e;
    }
    if ((nj & 1) && tid == 0) {   // odd joint count: the very last joint
        const long i = j0 + nj - 1;
        const float px = a.pred[i * 2], py = a.pred[i * 2 + 1], gx = a.gt[i * a.c_gt], gy = a.gt[i * a.c_gt + 1];
        float ex = 0.f, ey = 0.f;
        if (kept2(gx, gy, (int)(i % a.J), a.mask_hips)) {
            ex = px - gx; ey = py - gy;
            s = fmaf(ex, ex, fmaf(ey, ey, s));
            c += 2.f;
        }
        esm[(nj - 1) * 2] = ex; esm[(nj - 1) * 2 + 1] = ey;
    }
    __shared__ float rs[LT / 32], rc[LT / 32];
    __shared__ float k_shared;
    s = warp_sum(s);
    c = warp_sum(c);
    if (lane == 0) { rs[warp] = s; rc[warp] = c; }
    __syncthreads();
    if (tid == 0) {
        float x = 0.f, y = 0.f;
        for (int w = 0; w < LT / 32; ++w) { x += rs[w]; y += rc[w]; }
        a.partials[blockIdx.x * 2] = x;
        a.partials[blockIdx.x * 2 + 1] = y;
        // grid barrier (all CTAs are resident: cooperative launch): arrive with release, wait with acquire
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(a.sync) : "memory");
        unsigned seen;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(a.sync) : "memory");
        } while (seen < gridDim.x);
    }
    __syncthreads();
    if (warp == 0) {   // every CTA folds all partials the same way: identical k everywhere
        double ds = 0.0, dc = 0.0;
        for (int i = lane; i < (int)gridDim.x; i += 32) {
            const float2 v = __ldcg(reinterpret_cast<const float2*>(a.partials) + i);
            ds += (double)v.x;
            dc += (double)v.y;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            ds += __shfl_xor_sync(FULL_MASK, ds, o);
            dc += __shfl_xor_sync(FULL_MASK, dc, o);
        }
        if (lane == 0) {
            k_shared = (float)((double)a.upstream * 2.0 / dc);   // count 0: inf / NaN like nn.MSELoss on an empty selection
            if (blockIdx.x == 0) {
                a.out[0] = (float)(ds / dc); a.out[1] = (float)ds; a.out[2] = (float)dc; a.out[3] = 0.f;
                exchange_stash(a.xchg, (float)ds, (float)dc, 0.f, 0.f);   // (sent by the next step's kernel)
            }
            // the last CTA to leave resets the barrier words for the next launch
            unsigned gone;
            asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(gone) : "l"(a.sync + 1) : "memory");
            if (gone == gridDim.x - 1) { a.sync[0] = 0u; a.sync[1] = 0u; }
        }
    }
    __syncthreads();
    const float k = k_shared;
    float4* o4 = reinterpret_cast<float4*>(a.g_pred + j0 * 2);
#pragma unroll 4
    for (int q = tid; q < pairs; q += LT) {
        const float4 e = e4[q];
        __stcs(o4 + q, make_float4(k * e.x, k * e.y, k * e.z, k * e.w));
    }
    if ((nj & 1) && tid == 0) {
        const long i = j0 + nj - 1;
        a.g_pred[i * 2] = k * esm[(nj - 1) * 2];
        a.g_pred[i * 2 + 1] = k * esm[(nj - 1) * 2 + 1];
    }
}

struct Loc2dPlan {
    int ok, grid, per_cta;
    size_t smem;
};

Loc2dPlan loc2d_plan(int64_t m, int32_t j) {
    Loc2dPlan pl{};
    const long joints = (long)m * j;
    if (joints <= 0 || joints >= (1L << 31)) return pl;
    int grid = p2c_sm_count();
    if (grid > MAX_CTAS) grid = MAX_CTAS;
    const long min_per = 2L * LT;   // at least one pair per thread before more CTAs are worth their barrier
    if ((joints + min_per - 1) / min_per < grid) grid = (int)((joints + min_per - 1) / min_per);
    long per = (joints + grid - 1) / grid;
    per = (per + 1) & ~1L;
    pl.smem = (size_t)per * 2 * sizeof(float);
    if (pl.smem > 200 * 1024) return pl;
    pl.grid = (int)((joints + per - 1) / per);
    pl.per_cta = (int)per;
    pl.ok = 1;
    return pl;
}

}  // namespace

extern "C" int64_t p2c_loc2d_step_workspace_floats(void) { return 16 + 2 * MAX_CTAS; }

extern "C" int p2c_loc2d_step_supported(int64_t m, int32_t j, int32_t c_gt) {
    if (m < 1 || j < 1 || (c_gt != 2 && c_gt != 3)) return 0;
    return loc2d_plan(m, j).ok;
}

extern "C" int p2c_loc2d_step(const float* pred, const float* gt, int64_t m, int32_t j, int32_t c_gt, int32_t mask_hips,
                              float upstream, float* g_pred, float* workspace, float* out, float* exchange,
                              int32_t exchange_rank, int32_t exchange_world, void* stream) {
    if (m < 0 || j < 1 || (c_gt != 2 && c_gt != 3) || mask_hips >= j) return P2C_E_BADARG;
    if (m == 0) return 0;
    if (!pred || !gt || !g_pred || !workspace || !out) return P2C_E_NULL;
    if ((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(g_pred)) & 15) return P2C_E_UNSUPPORTED;
    if (reinterpret_cast<uintptr_t>(gt) & (c_gt == 2 ? 15 : 7)) return P2C_E_UNSUPPORTED;
    if (exchange && (exchange_world < 1 || exchange_world > XCHG_MAX_RANKS || exchange_rank < 0 ||
                     exchange_rank >= exchange_world))
        return P2C_E_BADARG;
    const Loc2dPlan pl = loc2d_plan(m, j);
    if (!pl.ok) return P2C_E_UNSUPPORTED;
    Loc2dArgs a{};
    a.pred = pred; a.gt = gt; a.g_pred = g_pred; a.joints = (long)m * j; a.J = j; a.c_gt = c_gt; a.mask_hips = mask_hips;
    a.per_cta = pl.per_cta; a.upstream = upstream;
    a.sync = reinterpret_cast<unsigned*>(workspace); a.partials = workspace + 16; a.out = out;
    a.xchg = XchgArgs{exchange, exchange_rank, exchange_world};
    if (const int e = p2c_allow_big_smem(reinterpret_cast<const void*>(loc2d_step_kernel), 200 * 1024)) return e;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(pl.grid);
    cfg.blockDim = dim3(LT);
    cfg.dynamicSmemBytes = pl.smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeCooperative;   // the CTAs wait for each other: all resident at once or no launch
    attr[0].val.cooperative = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // chained steps: see the top of the kernel
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    cudaError_t le = cudaLaunchKernelEx(&cfg, loc2d_step_kernel, a);
    if (le != cudaSuccess) {   // (the two attributes together are refused: cooperative only)
        (void)cudaGetLastError();
        cfg.numAttrs = 1;
        le = cudaLaunchKernelEx(&cfg, loc2d_step_kernel, a);
    }
    if (le != cudaSuccess) return (int)le;
    P2C_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------------ no-allocation entry
// p2c_loc2d_step_static: the same kernel with its scratch (barrier words + per-CTA partial sums) in a file-scope device
// array instead of a caller-allocated workspace, for callers that cannot allocate (CUDA graph capture, serving loops).
// Every GPU that loads the library holds the array until exit; every call on a device shares it.
namespace {

constexpr int STATIC_SCRATCH_FLOATS = 2176;   // 16 barrier words + 2 x MAX_CTAS partials = 2064, padded to 8704 bytes
static_assert(STATIC_SCRATCH_FLOATS >= 16 + 2 * MAX_CTAS, "scratch too small for the largest grid");
constexpr long STATIC_MAX_JOINTS = 148L * 25600L;   // 3 788 800: what one B200 keeps on chip (148 CTAs x 200 KB of errors)

static __device__ __align__(256) float g_loc2d_scratch[STATIC_SCRATCH_FLOATS];

constexpr int STATIC_MAX_DEVICES = 64;
float* g_scratch_ptr[STATIC_MAX_DEVICES];   // device address of g_loc2d_scratch, per device (each device has its own copy)
std::mutex g_scratch_mutex;

// In host code g_loc2d_scratch is only a handle: the device address comes from cudaGetSymbolAddress, asked once per
// device with that device current.
int static_scratch(float** out) {
    const int dev = p2c_current_device();
    if (dev >= STATIC_MAX_DEVICES) return P2C_E_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_ptr[dev]) {
        float* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_loc2d_scratch);
        if (e != cudaSuccess) return (int)e;
        if (!ptr) return P2C_E_NULL;
        g_scratch_ptr[dev] = ptr;
    }
    *out = g_scratch_ptr[dev];
    return 0;
}

}  // namespace

extern "C" int64_t p2c_loc2d_step_static_max_joints(void) { return STATIC_MAX_JOINTS; }

extern "C" int p2c_loc2d_step_static(const float* pred, const float* gt, int64_t m, int32_t j, int32_t c_gt,
                                     int32_t mask_hips, float upstream, float* g_pred, float* out, void* stream) {
    if (m < 0 || j < 1 || (c_gt != 2 && c_gt != 3) || mask_hips >= j) return P2C_E_BADARG;
    if (m == 0) return 0;
    if (!pred || !gt || !g_pred || !out) return P2C_E_NULL;
    if (m > STATIC_MAX_JOINTS || (long)m * j > STATIC_MAX_JOINTS) return P2C_E_TOO_LARGE;
    if ((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(g_pred)) & 15) return P2C_E_UNSUPPORTED;
    if (reinterpret_cast<uintptr_t>(gt) & (c_gt == 2 ? 15 : 7)) return P2C_E_UNSUPPORTED;
    const Loc2dPlan pl = loc2d_plan(m, j);
    if (!pl.ok) return P2C_E_TOO_LARGE;   // (a GPU with fewer SMs than a B200 keeps fewer errors on chip)
    if (16 + 2 * pl.grid > STATIC_SCRATCH_FLOATS) return P2C_E_TOO_LARGE;
    float* scratch = nullptr;
    if (const int e = static_scratch(&scratch)) return e;
    // the array keeps the last call's contents: the barrier words must start at zero (a launch that failed half way
    // would leave them set); the partials are overwritten before they are read
    const cudaError_t me = cudaMemsetAsync(scratch, 0, 16 * sizeof(float), (cudaStream_t)stream);
    if (me != cudaSuccess) return (int)me;
    Loc2dArgs a{};
    a.pred = pred; a.gt = gt; a.g_pred = g_pred; a.joints = (long)m * j; a.J = j; a.c_gt = c_gt; a.mask_hips = mask_hips;
    a.per_cta = pl.per_cta; a.upstream = upstream;
    a.sync = reinterpret_cast<unsigned*>(scratch); a.partials = scratch + 16; a.out = out;
    a.xchg = XchgArgs{nullptr, 0, 1};
    if (const int e = p2c_allow_big_smem(reinterpret_cast<const void*>(loc2d_step_kernel), 200 * 1024)) return e;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(pl.grid);
    cfg.blockDim = dim3(LT);
    cfg.dynamicSmemBytes = pl.smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t le = cudaLaunchKernelEx(&cfg, loc2d_step_kernel, a);
    if (le != cudaSuccess) return (int)le;
    P2C_LAUNCH_CHECK();
    return 0;
}
