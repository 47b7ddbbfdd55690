// Device helpers shared by the pose-per-lane tile kernels (fused_tile.cu: the fused training step; fk_tile.cu: the
// module API's forward / backward): MUFU math, 64/128-bit shared-memory record accesses, TMA bulk copies + mbarriers,
// named barriers, the compile-time CARLA tree.
#pragma once
#include "common.cuh"

namespace {

// Reciprocal, reciprocal square root and square root: the MUFU approximation (rel. error ~2^-22) refined by ONE Newton
// step (2-3 FMAs), i.e. within ~1 ulp of the correctly rounded result.  The raw approximations cost the gradient
// parity (per-bone error against the float64 oracle 1.1e-4 -> 3.8e-5 on the trunk bones at 256x16, 7.7e-5 -> 1e-5 on
// the legs; profiles/r02_grad_error_table.txt): the 2-D error terms (x - target) are small differences of O(1)
// normalised coordinates, so a 2e-7 relative error of 1/z or 1/scale is a 1e-5 relative error of the term.
// Experiment builds (tools/variants.py): P2C_APPROX_MATH = raw MUFU, P2C_IEEE_MATH = div.rn / sqrt.rn.
__device__ __forceinline__ float mufu_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float mufu_rsqrt(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
#if defined(P2C_IEEE_MATH)
__device__ __forceinline__ float frcp(float x) { return __fdiv_rn(1.f, x); }
__device__ __forceinline__ float frsqrt(float x) { return __fdiv_rn(1.f, __fsqrt_rn(x)); }
__device__ __forceinline__ float fsqrt(float x) { return __fsqrt_rn(x); }
#elif defined(P2C_APPROX_MATH)
__device__ __forceinline__ float frcp(float x) { return mufu_rcp(x); }
__device__ __forceinline__ float frsqrt(float x) { return mufu_rsqrt(x); }
__device__ __forceinline__ float fsqrt(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
#elif defined(P2C_MIXED_MATH)   // experiment: which of the three matters (P2C_MIXED_MATH = bit mask rcp 1, rsqrt 2, sqrt 4)
__device__ __forceinline__ float frcp(float x) { return (P2C_MIXED_MATH & 1) ? __fdiv_rn(1.f, x) : mufu_rcp(x); }
__device__ __forceinline__ float frsqrt(float x) { return (P2C_MIXED_MATH & 2) ? __fdiv_rn(1.f, __fsqrt_rn(x)) : mufu_rsqrt(x); }
__device__ __forceinline__ float fsqrt(float x) { return (P2C_MIXED_MATH & 4) ? __fsqrt_rn(x) : x * mufu_rsqrt(fmaxf(x, 1e-37f)); }
#else
__device__ __forceinline__ float frcp(float x) {   // (x = 0 or inf: the seed is already inf / 0 and is returned)
    const float r = mufu_rcp(x);
    const float e = fmaf(-x, r, 1.f);
    return fabsf(e) <= 1.f ? fmaf(r, e, r) : r;
}
__device__ __forceinline__ float frsqrt(float x) {   // callers clamp x to >= 1e-24
    const float y = mufu_rsqrt(x);
    const float e = fmaf(-x * y, y, 1.f);
    return fmaf(0.5f * y, e, y);
}
__device__ __forceinline__ float fsqrt(float x) {    // x >= 0; 0 -> 0, inf / NaN propagate
    const float y = mufu_rsqrt(fmaxf(x, 1e-37f));
    const float s = x * y;
    const float e = fmaf(-s, s, x);
    return fabsf(x) <= 3.0e38f ? fmaf(0.5f * y, e, s) : x;
}
#endif
#ifdef P2C_SUMS_DOUBLE   // experiment build: the running values of the subtree-sum walks in double
typedef double acc_t;
#else
typedef float acc_t;
#endif

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
__device__ __forceinline__ float lds32(const float* p) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(smem_u32(p)) : "memory");
    return v;
}
__device__ __forceinline__ void sts32(float* p, float x) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(smem_u32(p)), "f"(x) : "memory");
}
// 9-word bone record b of an AoS pose record (b is warp-uniform at every call site: the parity branch does not diverge)
// (64-bit accesses: half the wavefronts of scalar ones; the pose stride, 117 x 8 bytes, is conflict free)
// PAR: the parity of b when the caller knows it at compile time (0 / 1), -1 = look at b.  A run-time parity is compiled
// as BOTH access sequences under predicates plus the moves that merge their registers (17 issue slots for a 5-access
// record load): the task loops of the kernels therefore hand out bones in (even, odd) pairs where they can.
template <int PAR>
__device__ __forceinline__ bool odd_record(int b) { return PAR < 0 ? (b & 1) != 0 : PAR != 0; }
template <int PAR = -1>
__device__ __forceinline__ M3 lds9(const float* pose, int b) {
    const float* p = pose + 9 * b;
    M3 r;
    if (odd_record<PAR>(b)) {
        r.m[0] = p[0];
        const float2 v0 = lds64(p + 1), v1 = lds64(p + 3), v2 = lds64(p + 5), v3 = lds64(p + 7);
        r.m[1] = v0.x; r.m[2] = v0.y; r.m[3] = v1.x; r.m[4] = v1.y; r.m[5] = v2.x; r.m[6] = v2.y; r.m[7] = v3.x; r.m[8] = v3.y;
    } else {
        const float2 v0 = lds64(p), v1 = lds64(p + 2), v2 = lds64(p + 4), v3 = lds64(p + 6);
        r.m[0] = v0.x; r.m[1] = v0.y; r.m[2] = v1.x; r.m[3] = v1.y; r.m[4] = v2.x; r.m[5] = v2.y; r.m[6] = v3.x; r.m[7] = v3.y;
        r.m[8] = p[8];
    }
    return r;
}
template <int PAR = -1>
__device__ __forceinline__ void sts9(float* pose, int b, const M3& a) {
    float* p = pose + 9 * b;
    if (odd_record<PAR>(b)) {
        p[0] = a.m[0];
        sts64(p + 1, a.m[1], a.m[2]); sts64(p + 3, a.m[3], a.m[4]);
        sts64(p + 5, a.m[5], a.m[6]); sts64(p + 7, a.m[7], a.m[8]);
    } else {
        sts64(p, a.m[0], a.m[1]); sts64(p + 2, a.m[2], a.m[3]);
        sts64(p + 4, a.m[4], a.m[5]); sts64(p + 6, a.m[6], a.m[7]);
        p[8] = a.m[8];
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

// CARLA_SKELETON parents (the reference's skeleton table, utils/unreal.py: crl_root, hips, spine, spine01, left arm,
// neck/head/eyes, right arm, right leg, left leg); folds to a constant in the fully unrolled walks of the BT == 26
// kernels, which the launcher only picks when the call's tree equals this table
__host__ __device__ __forceinline__ constexpr int carla_parent(int b) {
    return b == 0 ? -1 : b == 8 || b == 12 ? 3 : b == 11 ? 9 : b == 16 || b == 21 ? 1 : b - 1;
}

__device__ __forceinline__ void named_arrive(int id, int threads) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void named_barrier(int id, int threads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ float4 lds128(const float* p) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "r"(smem_u32(p))
                 : "memory");
    return v;
}
__device__ __forceinline__ void sts128(float* p, float x, float y, float z, float w) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(p)), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
// 6-word bone record b of a 16-byte aligned AoS pose record (6-D / gradient records): 128 + 64 bits, order by parity
template <int PAR = -1>
__device__ __forceinline__ void lds6(const float* pose, int b, float* d) {
    const float* p = pose + 6 * b;
    if (odd_record<PAR>(b)) {
        const float2 u = lds64(p);
        const float4 v = lds128(p + 2);
        d[0] = u.x; d[1] = u.y; d[2] = v.x; d[3] = v.y; d[4] = v.z; d[5] = v.w;
    } else {
        const float4 v = lds128(p);
        const float2 u = lds64(p + 4);
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w; d[4] = u.x; d[5] = u.y;
    }
}
template <int PAR = -1>
__device__ __forceinline__ void sts6(float* pose, int b, const float* d) {
    float* p = pose + 6 * b;
    if (odd_record<PAR>(b)) {
        sts64(p, d[0], d[1]);
        sts128(p + 2, d[2], d[3], d[4], d[5]);
    } else {
        sts128(p, d[0], d[1], d[2], d[3]);
        sts64(p + 4, d[4], d[5]);
    }
}
// 3-word bone record b of an 8-byte aligned AoS pose record (locations, projections, 3-channel targets)
__device__ __forceinline__ V3 lds3(const float* pose, int b) {
    const float* p = pose + 3 * b;
    return V3{p[0], p[1], p[2]};
}
template <int PAR = -1>
__device__ __forceinline__ void sts3(float* pose, int b, const V3& v) {
    float* p = pose + 3 * b;
    if (odd_record<PAR>(b)) {
        p[0] = v.x;
        sts64(p + 1, v.y, v.z);
    } else {
        sts64(p, v.x, v.y);
        p[2] = v.z;
    }
}

}  // namespace

// Bone tasks of the element-wise passes: task t of nt goes to warp t mod W (one bone per task), or -- PAIRS -- tasks
// (2u, 2u + 1) go to warp u mod W, so that every call of the body knows the parity of its bone at compile time (nt and
// the bone count are even: the parity of a task is the parity of its bone).  The body is f(task, Parity<..>{}).
template <int PAR>
struct Parity {
    static constexpr int value = PAR;
};
// MODE 0: one task per trip, parity looked at inside the accessors; 1: pairs; 2: one task per trip, the body compiled
// twice and picked by a (warp-uniform) branch on the task's parity
template <int MODE, typename F>
__device__ __forceinline__ void bone_tasks(int warp, int W, int nt, F f) {
    if (warp >= W) return;   // (a warp beyond the W task warps -- the service warp of fused_tile.cu -- has no tasks)
    if (MODE == 1) {
#pragma unroll 1
        for (int t = 2 * warp; t < nt; t += 2 * W) {
            f(t, Parity<0>{});
            f(t + 1, Parity<1>{});
        }
    } else if (MODE == 2) {
#pragma unroll 1
        for (int t = warp; t < nt; t += W) {
            if (t & 1) f(t, Parity<1>{});
            else f(t, Parity<0>{});
        }
    } else {
#pragma unroll 1
        for (int t = warp; t < nt; t += W) f(t, Parity<-1>{});
    }
}
