// Shared device helpers for the p2c_b200 kernels (sm_100a, CUDA cores only: 3x3 / 1x3.3x3 work).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "p2c_b200.h"

#define P2C_LAUNCH_CHECK()                        \
    do {                                          \
        cudaError_t e__ = cudaGetLastError();     \
        if (e__ != cudaSuccess) return (int)e__;  \
    } while (0)

#define FULL_MASK 0xffffffffu

// 3x3 row-major matrix in registers.
struct M3 {
    float m[9];
};
struct V3 {
    float x, y, z;
};

__device__ __forceinline__ M3 m3_identity() {
    M3 r;
    r.m[0] = 1.f; r.m[1] = 0.f; r.m[2] = 0.f;
    r.m[3] = 0.f; r.m[4] = 1.f; r.m[5] = 0.f;
    r.m[6] = 0.f; r.m[7] = 0.f; r.m[8] = 1.f;
    return r;
}
__device__ __forceinline__ M3 m3_zero() {
    M3 r;
#pragma unroll
    for (int i = 0; i < 9; ++i) r.m[i] = 0.f;
    return r;
}
// C = A . B
__device__ __forceinline__ M3 m3_mul(const M3& a, const M3& b) {
    M3 c;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            c.m[i * 3 + j] = fmaf(a.m[i * 3 + 2], b.m[6 + j], fmaf(a.m[i * 3 + 1], b.m[3 + j], a.m[i * 3] * b.m[j]));
    return c;
}
// C = A . B^T
__device__ __forceinline__ M3 m3_mul_bt(const M3& a, const M3& b) {
    M3 c;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            c.m[i * 3 + j] =
                fmaf(a.m[i * 3 + 2], b.m[j * 3 + 2], fmaf(a.m[i * 3 + 1], b.m[j * 3 + 1], a.m[i * 3] * b.m[j * 3]));
    return c;
}
// C = A^T . B
__device__ __forceinline__ M3 m3_mul_at(const M3& a, const M3& b) {
    M3 c;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            c.m[i * 3 + j] = fmaf(a.m[6 + i], b.m[6 + j], fmaf(a.m[3 + i], b.m[3 + j], a.m[i] * b.m[j]));
    return c;
}
__device__ __forceinline__ void m3_acc(M3& a, const M3& b) {
#pragma unroll
    for (int i = 0; i < 9; ++i) a.m[i] += b.m[i];
}
// v . A  (row vector times matrix)
__device__ __forceinline__ V3 v3_mul_m(const V3& v, const M3& a) {
    V3 r;
    r.x = fmaf(v.z, a.m[6], fmaf(v.y, a.m[3], v.x * a.m[0]));
    r.y = fmaf(v.z, a.m[7], fmaf(v.y, a.m[4], v.x * a.m[1]));
    r.z = fmaf(v.z, a.m[8], fmaf(v.y, a.m[5], v.x * a.m[2]));
    return r;
}
// v . A^T
__device__ __forceinline__ V3 v3_mul_mt(const V3& v, const M3& a) {
    V3 r;
    r.x = fmaf(v.z, a.m[2], fmaf(v.y, a.m[1], v.x * a.m[0]));
    r.y = fmaf(v.z, a.m[5], fmaf(v.y, a.m[4], v.x * a.m[3]));
    r.z = fmaf(v.z, a.m[8], fmaf(v.y, a.m[7], v.x * a.m[6]));
    return r;
}
// a += u^T (x) v   (outer product, 3x1 . 1x3)
__device__ __forceinline__ void m3_acc_outer(M3& a, const V3& u, const V3& v) {
    a.m[0] = fmaf(u.x, v.x, a.m[0]); a.m[1] = fmaf(u.x, v.y, a.m[1]); a.m[2] = fmaf(u.x, v.z, a.m[2]);
    a.m[3] = fmaf(u.y, v.x, a.m[3]); a.m[4] = fmaf(u.y, v.y, a.m[4]); a.m[5] = fmaf(u.y, v.z, a.m[5]);
    a.m[6] = fmaf(u.z, v.x, a.m[6]); a.m[7] = fmaf(u.z, v.y, a.m[7]); a.m[8] = fmaf(u.z, v.z, a.m[8]);
}
__device__ __forceinline__ M3 m3_shfl(const M3& a, int src) {
    M3 r;
#pragma unroll
    for (int i = 0; i < 9; ++i) r.m[i] = __shfl_sync(FULL_MASK, a.m[i], src);
    return r;
}
__device__ __forceinline__ V3 v3_shfl(const V3& a, int src) {
    V3 r;
    r.x = __shfl_sync(FULL_MASK, a.x, src);
    r.y = __shfl_sync(FULL_MASK, a.y, src);
    r.z = __shfl_sync(FULL_MASK, a.z, src);
    return r;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
    return v;
}

// torch.nan_to_num(x, nan=0, posinf=0, neginf=0)  (utils/tensors.py:43-53)
__device__ __forceinline__ float nan_to_zero(float x) { return (fabsf(x) <= 3.402823466e38f) ? x : 0.f; }

// ---- warp-cooperative tile movement -------------------------------------------------------------
// Per-pose records are 8-byte but not 16-byte multiples (B*9*4 = 936, B*3*4 = 312 for B = 26), so the
// widest vector that is aligned for every pose is float2 when the record has an even float count.
// Lanes move consecutive words => full-sector coalescing; the lane<->bone redistribution goes through
// a per-warp shared-memory tile (stride 9 or 3 words per lane: odd => bank-conflict free).
__device__ __forceinline__ void warp_tile_load(float* __restrict__ s, const float* __restrict__ g, int nfloats,
                                               int lane) {
    if (((nfloats & 1) == 0) && ((reinterpret_cast<uintptr_t>(g) & 7) == 0)) {
        const float2* g2 = reinterpret_cast<const float2*>(g);
        float2* s2 = reinterpret_cast<float2*>(s);
        for (int i = lane; i < (nfloats >> 1); i += 32) s2[i] = __ldg(g2 + i);
    } else {
        for (int i = lane; i < nfloats; i += 32) s[i] = __ldg(g + i);
    }
}
__device__ __forceinline__ void warp_tile_store(float* __restrict__ g, const float* __restrict__ s, int nfloats,
                                                int lane) {
    if (((nfloats & 1) == 0) && ((reinterpret_cast<uintptr_t>(g) & 7) == 0)) {
        float2* g2 = reinterpret_cast<float2*>(g);
        const float2* s2 = reinterpret_cast<const float2*>(s);
        for (int i = lane; i < (nfloats >> 1); i += 32) g2[i] = s2[i];
    } else {
        for (int i = lane; i < nfloats; i += 32) g[i] = s[i];
    }
}
__device__ __forceinline__ M3 m3_from(const float* s) {
    M3 r;
#pragma unroll
    for (int i = 0; i < 9; ++i) r.m[i] = s[i];
    return r;
}
__device__ __forceinline__ void m3_to(float* s, const M3& a) {
#pragma unroll
    for (int i = 0; i < 9; ++i) s[i] = a.m[i];
}

// ---- loss-sum exchange block (include/p2c_b200.h, p2c_exchange_*): the path's only cross-GPU step, done by the step
//      kernels themselves.  Layout of a rank's block (32-bit words): [0,32) 16 peer-mapped base pointers, one per rank
//      (slot r = rank r's block as THIS rank addresses it); [32] number of steps this rank has finished; [33] number of
//      steps it has sent; [36,40) / [40,44) the sums of the last two finished steps (slot = step & 1), waiting to be
//      sent; [64, ...) DEPTH x 16 entries of 8 words {sum_sq_2d, seq, count_2d, seq, sum_sq_3d, seq, count_3d, seq},
//      entry [seq % DEPTH][r] = rank r's sums of its step number `seq`.  Every value travels with the sequence number
//      in ONE aligned 8-byte store (single-copy atomic): an entry whose four sequence words equal `seq` is complete,
//      whatever order the stores land in -- no release fence, the sending thread never waits for an acknowledgement.
constexpr int XCHG_DEPTH = 16;
constexpr int XCHG_MAX_RANKS = 16;
constexpr int XCHG_HEADER_WORDS = 64;
constexpr int XCHG_ENTRY_WORDS = 8;
constexpr size_t XCHG_BLOCK_BYTES = (XCHG_HEADER_WORDS + XCHG_DEPTH * XCHG_MAX_RANKS * XCHG_ENTRY_WORDS) * 4;

struct XchgArgs {   // by value in the kernel arguments; block == nullptr: no exchange
    float* block;
    int rank, world;
};

// The thread that finishes a step's loss sums leaves them in the rank's OWN block (two plain local stores): nothing
// crosses NVLink on the step's tail.
__device__ __forceinline__ void exchange_stash(const XchgArgs x, float s0, float s1, float s2, float s3) {
    if (!x.block) return;
    unsigned* hdr = reinterpret_cast<unsigned*>(x.block);
    const unsigned seq = __ldcg(hdr + 32) + 1u;   // (only the finishing thread of this stream's steps writes it)
    *reinterpret_cast<float4*>(x.block + 36 + 4 * (seq & 1u)) = make_float4(s0, s1, s2, s3);
    hdr[32] = seq;
}
// Sends what is waiting (the sums of the step before) into entry [seq][rank] of EVERY rank's block: four relaxed 8-byte
// (value, seq) stores per rank, over NVLink for the peers.  Called by one thread at the START of the next step's kernel
// -- the stores then have the whole step to land (issued on the tail they cost 1.2 us per 14 us step: a kernel is
// not complete before its remote writes are acknowledged) -- and by p2c_exchange_flush behind the last step.
__device__ __forceinline__ void exchange_send_pending(const XchgArgs x) {
    if (!x.block) return;
    unsigned* hdr = reinterpret_cast<unsigned*>(x.block);
    const unsigned last = __ldcg(hdr + 32), sent = __ldcg(hdr + 33);
    if (last == sent) return;
    const unsigned long long* table = reinterpret_cast<const unsigned long long*>(x.block);
    unsigned long long peer[XCHG_MAX_RANKS];   // every peer pointer requested before the first store goes out
#pragma unroll
    for (int r = 0; r < XCHG_MAX_RANKS; ++r) peer[r] = r < x.world ? __ldcg(table + r) : 0ull;
    // (the block keeps the last TWO finished steps: a stashing kernel followed by one that publishes at once loses nothing)
    for (unsigned seq = (last - sent > 1u) ? last - 1u : last;; ++seq) {
        const float4 v4 = __ldcg(reinterpret_cast<const float4*>(x.block + 36 + 4 * (seq & 1u)));
        const size_t off = XCHG_HEADER_WORDS + ((size_t)(seq % XCHG_DEPTH) * XCHG_MAX_RANKS + x.rank) * XCHG_ENTRY_WORDS;
        const float v[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
        for (int r = 0; r < XCHG_MAX_RANKS; ++r) {
            if (r < x.world) {
                float* slot = reinterpret_cast<float*>(peer[r]) + off;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    asm volatile("st.relaxed.sys.global.v2.b32 [%0], {%1, %2};" ::"l"(slot + 2 * k),
                                 "r"(__float_as_uint(v[k])), "r"(seq)
                                 : "memory");
            }
        }
        if (seq == last) break;
    }
    hdr[33] = last;
}
// (kernels that are not on the latency-critical path publish at once)
__device__ __forceinline__ void exchange_publish(const XchgArgs x, float s0, float s1, float s2, float s3) {
    exchange_stash(x, s0, s1, s2, s3);
    exchange_send_pending(x);
}

// Host-side tree preprocessing shared by the launchers.
struct TreeDev {
    int n_bones;
    int max_depth;
    int max_children;
    int8_t parent[P2C_MAX_BONES];
    int8_t depth[P2C_MAX_BONES];
    int8_t child[P2C_MAX_BONES][4];  // up to 4 children per bone are handled by the shuffle gather
};

static inline int build_tree(const p2c_tree_t& t, TreeDev& d) {
    if (t.n_bones < 1 || t.n_bones > P2C_MAX_BONES) return P2C_E_BONES;
    d.n_bones = t.n_bones;
    d.max_depth = 0;
    d.max_children = 0;
    int nchild[P2C_MAX_BONES];
    for (int b = 0; b < P2C_MAX_BONES; ++b) {
        d.parent[b] = -1;
        d.depth[b] = 0;
        nchild[b] = 0;
        for (int k = 0; k < 4; ++k) d.child[b][k] = -1;
    }
    for (int b = 0; b < t.n_bones; ++b) {
        int p = t.parent[b];
        if (b == 0) {
            if (p != -1) return P2C_E_BONES;
            continue;
        }
        if (p < 0 || p >= b) return P2C_E_BONES;  // DFS order: a parent precedes its children
        d.parent[b] = (int8_t)p;
        d.depth[b] = (int8_t)(d.depth[p] + 1);
        if (d.depth[b] > d.max_depth) d.max_depth = d.depth[b];
        if (nchild[p] >= 4) return P2C_E_UNSUPPORTED;
        d.child[p][nchild[p]++] = (int8_t)b;
        if (nchild[p] > d.max_children) d.max_children = nchild[p];
    }
    return 0;
}
