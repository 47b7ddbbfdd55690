// Rotation conversions at the reference's call sites (pytorch3d 0.6.0 semantics):
//   rotation_6d_to_matrix  (modules/movements/movements.py:113-116)        fwd + hand-derived bwd
//   euler_angles_to_matrix(.,"XYZ") / matrix_to_euler_angles(.,"XYZ")      (walker_control/p3d_pose.py:54,78,259)
// Elementwise, HBM-bound: records are 24 B in / 36 B out, so each CTA stages its 256 records through
// shared memory and moves them as coalesced float4.
#include "common.cuh"

namespace {

constexpr int RT = 256;  // records per CTA

__device__ __forceinline__ void block_copy_in(float* s, const float* g, long nfloats_total, long base, int count) {
    // base is a multiple of 4 floats for every CTA (RT*6, RT*9 are multiples of 4) => float4 aligned
    const float4* g4 = reinterpret_cast<const float4*>(g + base);
    float4* s4 = reinterpret_cast<float4*>(s);
    const int n4 = ((reinterpret_cast<uintptr_t>(g + base) & 15) == 0) ? (count >> 2) : 0;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) s4[i] = __ldg(g4 + i);
    for (int i = (n4 << 2) + threadIdx.x; i < count; i += blockDim.x) s[i] = __ldg(g + base + i);
}
__device__ __forceinline__ void block_copy_out(float* g, const float* s, long base, int count) {
    float4* g4 = reinterpret_cast<float4*>(g + base);
    const float4* s4 = reinterpret_cast<const float4*>(s);
    const int n4 = ((reinterpret_cast<uintptr_t>(g + base) & 15) == 0) ? (count >> 2) : 0;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) g4[i] = s4[i];
    for (int i = (n4 << 2) + threadIdx.x; i < count; i += blockDim.x) g[base + i] = s[i];
}

struct Gs {  // Gram-Schmidt intermediates of rotation_6d_to_matrix
    V3 b1, b2, b3, u;
    float n1, n2, s;
};

__device__ __forceinline__ Gs rot6d(const float* a) {
    Gs r;
    const float eps = 1e-12f;  // F.normalize default
    r.n1 = fmaxf(sqrtf(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]), eps);
    r.b1 = V3{a[0] / r.n1, a[1] / r.n1, a[2] / r.n1};
    r.s = r.b1.x * a[3] + r.b1.y * a[4] + r.b1.z * a[5];
    r.u = V3{a[3] - r.s * r.b1.x, a[4] - r.s * r.b1.y, a[5] - r.s * r.b1.z};
    r.n2 = fmaxf(sqrtf(r.u.x * r.u.x + r.u.y * r.u.y + r.u.z * r.u.z), eps);
    r.b2 = V3{r.u.x / r.n2, r.u.y / r.n2, r.u.z / r.n2};
    r.b3 = V3{r.b1.y * r.b2.z - r.b1.z * r.b2.y, r.b1.z * r.b2.x - r.b1.x * r.b2.z,
              r.b1.x * r.b2.y - r.b1.y * r.b2.x};
    return r;
}

__device__ __forceinline__ V3 cross(const V3& a, const V3& b) {
    return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ float dot(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

__global__ void __launch_bounds__(RT) rot6d_fwd_kernel(const float* __restrict__ d6, float* __restrict__ out, long n) {
    __shared__ __align__(16) float s[RT * 9];
    const long first = (long)blockIdx.x * RT;
    const int cnt = (int)min((long)RT, n - first);
    block_copy_in(s, d6, n * 6, first * 6, cnt * 6);
    __syncthreads();
    float a[6];
    const bool act = threadIdx.x < cnt;
    if (act)
#pragma unroll
        for (int i = 0; i < 6; ++i) a[i] = s[threadIdx.x * 6 + i];
    __syncthreads();
    if (act) {
        Gs r = rot6d(a);
        float* o = s + threadIdx.x * 9;
        o[0] = r.b1.x; o[1] = r.b1.y; o[2] = r.b1.z;
        o[3] = r.b2.x; o[4] = r.b2.y; o[5] = r.b2.z;
        o[6] = r.b3.x; o[7] = r.b3.y; o[8] = r.b3.z;
    }
    __syncthreads();
    block_copy_out(out, s, first * 9, cnt * 9);
}

__global__ void __launch_bounds__(RT) rot6d_bwd_kernel(const float* __restrict__ d6, const float* __restrict__ g_out,
                                                       float* __restrict__ g_d6, long n) {
    __shared__ __align__(16) float s[RT * 9];
    __shared__ __align__(16) float sa[RT * 6];
    const long first = (long)blockIdx.x * RT;
    const int cnt = (int)min((long)RT, n - first);
    block_copy_in(s, g_out, n * 9, first * 9, cnt * 9);
    block_copy_in(sa, d6, n * 6, first * 6, cnt * 6);
    __syncthreads();
    const bool act = threadIdx.x < cnt;
    if (act) {
        float a[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) a[i] = sa[threadIdx.x * 6 + i];
        const float* g = s + threadIdx.x * 9;
        V3 g1{g[0], g[1], g[2]}, g2{g[3], g[4], g[5]}, g3{g[6], g[7], g[8]};
        Gs r = rot6d(a);
        const float eps = 1e-12f;
        V3 a2{a[3], a[4], a[5]};
        // b3 = b1 x b2
        V3 t = cross(r.b2, g3);
        V3 gb1{g1.x + t.x, g1.y + t.y, g1.z + t.z};
        t = cross(g3, r.b1);
        V3 gb2{g2.x + t.x, g2.y + t.y, g2.z + t.z};
        // b2 = u / max(|u|, eps)
        V3 gu;
        if (r.n2 > eps) {
            float d = dot(r.b2, gb2);
            gu = V3{(gb2.x - r.b2.x * d) / r.n2, (gb2.y - r.b2.y * d) / r.n2, (gb2.z - r.b2.z * d) / r.n2};
        } else {
            gu = V3{gb2.x / eps, gb2.y / eps, gb2.z / eps};
        }
        // u = a2 - (b1.a2) b1
        float du = dot(r.b1, gu);
        V3 ga2{gu.x - du * r.b1.x, gu.y - du * r.b1.y, gu.z - du * r.b1.z};
        gb1.x -= r.s * gu.x + du * a2.x;
        gb1.y -= r.s * gu.y + du * a2.y;
        gb1.z -= r.s * gu.z + du * a2.z;
        // b1 = a1 / max(|a1|, eps)
        V3 ga1;
        if (r.n1 > eps) {
            float d = dot(r.b1, gb1);
            ga1 = V3{(gb1.x - r.b1.x * d) / r.n1, (gb1.y - r.b1.y * d) / r.n1, (gb1.z - r.b1.z * d) / r.n1};
        } else {
            ga1 = V3{gb1.x / eps, gb1.y / eps, gb1.z / eps};
        }
        float* o = sa + threadIdx.x * 6;
        o[0] = ga1.x; o[1] = ga1.y; o[2] = ga1.z; o[3] = ga2.x; o[4] = ga2.y; o[5] = ga2.z;
    }
    __syncthreads();
    block_copy_out(g_d6, sa, first * 6, cnt * 6);
}

// R = Rx(e0) . Ry(e1) . Rz(e2)
__global__ void euler_xyz_to_matrix_kernel(const float* __restrict__ e, float* __restrict__ out, long n) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        float sx, cx, sy, cy, sz, cz;
        sincosf(e[i * 3], &sx, &cx);
        sincosf(e[i * 3 + 1], &sy, &cy);
        sincosf(e[i * 3 + 2], &sz, &cz);
        float* o = out + i * 9;
        o[0] = cy * cz;                 o[1] = -cy * sz;                o[2] = sy;
        o[3] = sx * sy * cz + cx * sz;  o[4] = -sx * sy * sz + cx * cz; o[5] = -sx * cy;
        o[6] = -cx * sy * cz + sx * sz; o[7] = cx * sy * sz + sx * cz;  o[8] = cx * cy;
    }
}

// (atan2(-M12, M22), asin(M02), atan2(-M01, M00))
__global__ void matrix_to_euler_xyz_kernel(const float* __restrict__ m, float* __restrict__ e, long n) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const float* a = m + i * 9;
        e[i * 3] = atan2f(-a[5], a[8]);
        e[i * 3 + 1] = asinf(a[2]);
        e[i * 3 + 2] = atan2f(-a[1], a[0]);
    }
}

}  // namespace

extern "C" int p2c_rot6d_to_matrix_fwd(const float* d6, float* out, int64_t n, void* stream) {
    if (!d6 || !out) return P2C_E_NULL;
    if (n < 0) return P2C_E_BADARG;
    if (n == 0) return 0;
    rot6d_fwd_kernel<<<(unsigned)((n + RT - 1) / RT), RT, 0, (cudaStream_t)stream>>>(d6, out, (long)n);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_rot6d_to_matrix_bwd(const float* d6, const float* g_out, float* g_d6, int64_t n, void* stream) {
    if (!d6 || !g_out || !g_d6) return P2C_E_NULL;
    if (n < 0) return P2C_E_BADARG;
    if (n == 0) return 0;
    rot6d_bwd_kernel<<<(unsigned)((n + RT - 1) / RT), RT, 0, (cudaStream_t)stream>>>(d6, g_out, g_d6, (long)n);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_euler_xyz_to_matrix(const float* euler, float* out, int64_t n, void* stream) {
    if (!euler || !out) return P2C_E_NULL;
    if (n < 0) return P2C_E_BADARG;
    if (n == 0) return 0;
    long blocks = (n + 255) / 256;
    if (blocks > 148L * 16) blocks = 148L * 16;
    euler_xyz_to_matrix_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(euler, out, (long)n);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int p2c_matrix_to_euler_xyz(const float* mat, float* euler, int64_t n, void* stream) {
    if (!mat || !euler) return P2C_E_NULL;
    if (n < 0) return P2C_E_BADARG;
    if (n == 0) return 0;
    long blocks = (n + 255) / 256;
    if (blocks > 148L * 16) blocks = 148L * 16;
    matrix_to_euler_xyz_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(mat, euler, (long)n);
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" const char* p2c_version(void) { return "p2c_b200 0.1 (sm_100a)"; }
