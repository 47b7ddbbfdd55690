// Host-side state shared by the launchers: per-DEVICE caches (SM count, kernels whose dynamic shared-memory cap has been
// raised) and the debug / test environment switches, read once per process instead of on every launch.
// A process may drive several GPUs (one context per device): nothing here is keyed by process alone.
#include <stdlib.h>

#include <mutex>
#include <utility>
#include <vector>

#include "host_util.h"

namespace {

constexpr int MAX_DEVICES = 64;
int g_sm_count[MAX_DEVICES];   // 0 = not asked yet
std::mutex g_mutex;
std::vector<std::pair<const void*, int>> g_configured;   // (kernel, device) pairs already configured
P2cEnv g_env;
bool g_env_read = false;

bool flag(const char* name) {
    const char* v = getenv(name);
    return v && v[0] == '1';
}
int number(const char* name) {
    const char* v = getenv(name);
    return v ? atoi(v) : 0;
}
void read_env() {
    g_env.disable_tile = flag("P2C_DISABLE_TILE");
    g_env.disable_fast = flag("P2C_DISABLE_FAST");
    g_env.force_slim = flag("P2C_FORCE_SLIM");
    g_env.disable_slim = flag("P2C_DISABLE_SLIM");
    g_env.disable_tma = flag("P2C_DISABLE_TMA");
    g_env.tile_one_cta = flag("P2C_TILE_ONE_CTA");
    g_env.tile_no_coop = flag("P2C_TILE_NO_COOP");
    g_env.disable_fwd_tile = flag("P2C_DISABLE_FWD_TILE");
    g_env.tile_cta = number("P2C_TILE_CTA");
    g_env.tile_clocks = number("P2C_TILE_CLOCKS");
    g_env_read = true;
}

}  // namespace

int p2c_current_device() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) dev = 0;
    return dev;
}

int p2c_sm_count() {
    const int dev = p2c_current_device();
    if (dev >= MAX_DEVICES) return 148;
    int n = g_sm_count[dev];
    if (!n) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        g_sm_count[dev] = n;   // (idempotent: a race writes the same value)
    }
    return n;
}

int p2c_allow_big_smem(const void* kernel, size_t bytes) {
    const int dev = p2c_current_device();
    std::lock_guard<std::mutex> lock(g_mutex);
    for (const auto& kd : g_configured)
        if (kd.first == kernel && kd.second == dev) return 0;
    const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) return (int)e;
    g_configured.emplace_back(kernel, dev);
    return 0;
}

const P2cEnv& p2c_env() {
    if (!g_env_read) {
        std::lock_guard<std::mutex> lock(g_mutex);
        if (!g_env_read) read_env();
    }
    return g_env;
}

// test hook: the parity tests switch kernel variants inside one process (monkeypatch.setenv) and call this afterwards
extern "C" void p2c_debug_reload_env(void) {
    std::lock_guard<std::mutex> lock(g_mutex);
    read_env();
}
