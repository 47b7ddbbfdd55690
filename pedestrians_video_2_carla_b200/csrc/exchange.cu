// p2c_exchange_*: the block through which the ranks of a data-parallel job see each other's loss sums (clips
// shard by batch, the only cross-GPU data are the four sums behind loc_2d / loc_3d).  The step kernels publish into
// the block themselves (common.cuh exchange_publish: peer stores over NVLink, no collective kernel, nothing waits);
// this file holds the host side -- allocation of an IPC-shareable block, mapping the peers' blocks, the peer table --
// and the reader.
#include <string.h>

#include "common.cuh"

namespace {

// lane r reads rank r's entry of step `seq`; the totals are formed in rank order by one lane (deterministic)
__global__ void exchange_read_kernel(const float* __restrict__ block, int world, unsigned seq, float* __restrict__ out) {
    const int lane = threadIdx.x;
    if (seq == 0u) seq = __ldcg(reinterpret_cast<const unsigned*>(block) + 32);   // the last step this rank published
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    int ok = 0;
    if (lane < world && seq != 0u) {
        const float* slot = block + XCHG_HEADER_WORDS + ((size_t)(seq % XCHG_DEPTH) * XCHG_MAX_RANKS + lane) * XCHG_ENTRY_WORDS;
        unsigned val[4], tag[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)   // every (value, seq) pair is one 8-byte load of one 8-byte store: never torn
            asm volatile("ld.relaxed.sys.global.v2.b32 {%0, %1}, [%2];" : "=r"(val[k]), "=r"(tag[k]) : "l"(slot + 2 * k) : "memory");
        if (tag[0] == seq && tag[1] == seq && tag[2] == seq && tag[3] == seq) {
            v = make_float4(__uint_as_float(val[0]), __uint_as_float(val[1]), __uint_as_float(val[2]), __uint_as_float(val[3]));
            ok = 1;
        }
    }
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int arrived = 0;
    for (int r = 0; r < world; ++r) {   // (every lane runs the same loop: shuffles need the whole warp)
        s0 += (double)__shfl_sync(FULL_MASK, v.x, r);
        s1 += (double)__shfl_sync(FULL_MASK, v.y, r);
        s2 += (double)__shfl_sync(FULL_MASK, v.z, r);
        s3 += (double)__shfl_sync(FULL_MASK, v.w, r);
        arrived += __shfl_sync(FULL_MASK, ok, r);
    }
    if (lane == 0) {
        out[0] = (float)s0; out[1] = (float)s1; out[2] = (float)s2; out[3] = (float)s3;
        out[4] = (float)arrived;
        out[5] = s1 > 0.0 ? (float)(s0 / s1) : 0.f;   // global loc_2d (sum over ranks / count over ranks)
        out[6] = s3 > 0.0 ? (float)(s2 / s3) : 0.f;   // global loc_3d
        out[7] = (float)seq;
    }
}

__global__ void exchange_flush_kernel(const XchgArgs x) { exchange_send_pending(x); }

}  // namespace

extern "C" int p2c_exchange_flush(void* block, int32_t rank, int32_t world, void* stream) {
    if (!block) return P2C_E_NULL;
    if (world < 1 || world > XCHG_MAX_RANKS || rank < 0 || rank >= world) return P2C_E_BADARG;
    exchange_flush_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(XchgArgs{reinterpret_cast<float*>(block), rank, world});
    P2C_LAUNCH_CHECK();
    return 0;
}

extern "C" int64_t p2c_exchange_block_bytes(void) { return (int64_t)XCHG_BLOCK_BYTES; }
extern "C" int32_t p2c_exchange_max_ranks(void) { return XCHG_MAX_RANKS; }

extern "C" int p2c_exchange_create(void** block, unsigned char* ipc_handle) {
    if (!block) return P2C_E_NULL;
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, XCHG_BLOCK_BYTES);
    if (e != cudaSuccess) return (int)e;
    e = cudaMemset(p, 0, XCHG_BLOCK_BYTES);
    if (e == cudaSuccess && ipc_handle) {
        cudaIpcMemHandle_t h;
        e = cudaIpcGetMemHandle(&h, p);
        if (e == cudaSuccess) memcpy(ipc_handle, &h, sizeof(h));
    }
    if (e != cudaSuccess) {
        cudaFree(p);
        return (int)e;
    }
    *block = p;
    return 0;
}

extern "C" int p2c_exchange_open(const unsigned char* ipc_handle, void** peer_block) {
    if (!ipc_handle || !peer_block) return P2C_E_NULL;
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle, sizeof(h));
    const cudaError_t e = cudaIpcOpenMemHandle(peer_block, h, cudaIpcMemLazyEnablePeerAccess);
    return e == cudaSuccess ? 0 : (int)e;
}

extern "C" int p2c_exchange_close(void* peer_block) {
    if (!peer_block) return 0;
    const cudaError_t e = cudaIpcCloseMemHandle(peer_block);
    return e == cudaSuccess ? 0 : (int)e;
}

extern "C" int p2c_exchange_destroy(void* block) {
    if (!block) return 0;
    const cudaError_t e = cudaFree(block);
    return e == cudaSuccess ? 0 : (int)e;
}

extern "C" int p2c_exchange_set_peers(void* block, void* const* peer_blocks, int32_t world) {
    if (!block || !peer_blocks) return P2C_E_NULL;
    if (world < 1 || world > XCHG_MAX_RANKS) return P2C_E_BADARG;
    unsigned long long table[XCHG_MAX_RANKS];
    for (int r = 0; r < XCHG_MAX_RANKS; ++r) table[r] = r < world ? (unsigned long long)(uintptr_t)peer_blocks[r] : 0ull;
    for (int r = 0; r < world; ++r)
        if (!table[r]) return P2C_E_NULL;
    const cudaError_t e = cudaMemcpy(block, table, sizeof(table), cudaMemcpyHostToDevice);
    return e == cudaSuccess ? 0 : (int)e;
}

extern "C" int p2c_exchange_read(const void* block, int32_t world, uint32_t seq, float* out, void* stream) {
    if (!block || !out) return P2C_E_NULL;
    if (world < 1 || world > XCHG_MAX_RANKS) return P2C_E_BADARG;
    exchange_read_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float*>(block), world, seq, out);
    P2C_LAUNCH_CHECK();
    return 0;
}
