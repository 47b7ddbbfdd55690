// p2c_draw_sparse_changes: the random draws of the synthetic-target generator on the device
// (Carla2D3DIterableDataset.generate_batch, data/carla/datasets/carla_2d3d_dataset.py:145-183): per frame
// `changes_per_frame` DISTINCT bones get Euler changes uniform in +-max_change (the reference: np.random.choice without
// replacement + torch.rand, :157-162), the world yaw gets an initial and a per-frame change (:166-173), every clip an
// (age, gender) skeleton (:181-182).  The reference draws on the host in a Python loop over batch x clip; here one warp
// per pose draws from a counter-based generator (Philox4x32-10, keyed by the seed, counter = (pose, draw, stream,
// offset)): no state, no host round trip, any launch shape gives the same numbers, and tests/ restates it bit for bit.
// The random STREAM is not the reference's (numpy's Mersenne twister has no device form); the distribution is.
#include "common.cuh"

namespace {

struct U4 {
    unsigned x, y, z, w;
};

__device__ __forceinline__ U4 philox4x32_10(unsigned c0, unsigned c1, unsigned c2, unsigned c3, unsigned k0, unsigned k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const unsigned hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const unsigned n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return U4{c0, c1, c2, c3};
}
// 24 random bits -> [0, 1) exactly representable; (2u - 1) * amp without fused multiply-add (bit-reproducible on the host)
__device__ __forceinline__ float unit(unsigned r) { return (float)(r >> 8) * (1.0f / 16777216.0f); }
__device__ __forceinline__ float symmetric(unsigned r, float amp) {
    return __fmul_rn(__fsub_rn(__fmul_rn(unit(r), 2.0f), 1.0f), amp);
}

struct DrawArgs {
    unsigned long long seed, offset;
    int N, L, B, k, n_types;
    float max_change, max_world, max_world_initial;
    float* euler;        // (N, L, B, 3)
    float* world_euler;  // (N, L, 3)
    int32_t* skeleton;   // (N)
};

constexpr int DW = 8;   // warps per CTA

__global__ void __launch_bounds__(DW * 32) draw_kernel(const DrawArgs a) {
    const int lane = threadIdx.x & 31;
    const long poses = (long)a.N * a.L;
    const unsigned k0 = (unsigned)a.seed, k1 = (unsigned)(a.seed >> 32);
    const unsigned o0 = (unsigned)a.offset, o1 = (unsigned)(a.offset >> 32);
    for (long m = (long)blockIdx.x * DW + (threadIdx.x >> 5); m < poses; m += (long)gridDim.x * DW) {
        // every lane runs the same draws (same counters): the chosen set is warp-uniform without any exchange
        unsigned chosen = 0u;
        float ax = 0.f, ay = 0.f, az = 0.f;
        for (int i = 0; i < a.k; ++i) {
            const U4 r = philox4x32_10((unsigned)m, (unsigned)i, o0, o1, k0, k1);   // stream 0 (o-words carry the offset)
            // the (r.x mod (B - i))-th bone not chosen yet
            int pick = (int)(r.x % (unsigned)(a.B - i));
            int bone = 0;
            for (int b = 0; b < a.B; ++b) {
                if (chosen & (1u << b)) continue;
                if (pick == 0) { bone = b; break; }
                --pick;
            }
            chosen |= 1u << bone;
            if (lane == bone) {
                ax = symmetric(r.y, a.max_change); ay = symmetric(r.z, a.max_change); az = symmetric(r.w, a.max_change);
            }
        }
        if (lane < a.B) {
            float* o = a.euler + (m * a.B + lane) * 3;
            o[0] = ax; o[1] = ay; o[2] = az;
        }
        if (lane == 0) {
            const long clip = m / a.L;
            const int frame = (int)(m - clip * a.L);
            const U4 w = philox4x32_10((unsigned)m, 0xFFFFFFFFu, o0 ^ 0x5bd1e995u, o1, k0, k1);   // stream 1: world yaw
            float yaw = 0.f;
            if (frame == 0) {
                if (a.max_world_initial > 0.f) yaw = symmetric(w.x, a.max_world_initial);
            } else if (a.max_world != 0.f) {
                yaw = symmetric(w.y, a.max_world);
            }
            float* wo = a.world_euler + m * 3;
            wo[0] = 0.f; wo[1] = 0.f; wo[2] = yaw;
            if (frame == 0 && a.skeleton) a.skeleton[clip] = (int)(w.z % (unsigned)a.n_types);
        }
    }
}

}  // namespace

extern "C" int p2c_draw_sparse_changes(uint64_t seed, uint64_t offset, int32_t n_clips, int32_t clip_len, int32_t n_bones,
                                       int32_t changes_per_frame, float max_change, float max_world_change,
                                       float max_initial_world_change, int32_t n_types, float* euler,
                                       float* world_euler, int32_t* skeleton_index, void* stream) {
    if (n_clips < 0 || clip_len < 1 || n_bones < 1 || n_bones > P2C_MAX_BONES) return P2C_E_BADARG;
    if (changes_per_frame < 0 || changes_per_frame > n_bones || n_types < 1) return P2C_E_BADARG;
    if (n_clips == 0) return 0;
    if (!euler || !world_euler) return P2C_E_NULL;
    if ((long)n_clips * clip_len >= (1L << 32)) return P2C_E_UNSUPPORTED;   // the pose index is one counter word
    DrawArgs a{};
    a.seed = seed; a.offset = offset; a.N = n_clips; a.L = clip_len; a.B = n_bones; a.k = changes_per_frame;
    a.n_types = n_types; a.max_change = max_change; a.max_world = max_world_change;
    a.max_world_initial = max_initial_world_change; a.euler = euler; a.world_euler = world_euler; a.skeleton = skeleton_index;
    long blocks = ((long)n_clips * clip_len + DW - 1) / DW;
    if (blocks > 148L * 16) blocks = 148L * 16;
    draw_kernel<<<(int)blocks, DW * 32, 0, (cudaStream_t)stream>>>(a);
    P2C_LAUNCH_CHECK();
    return 0;
}
