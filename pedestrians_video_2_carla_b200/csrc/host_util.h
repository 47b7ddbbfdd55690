// Host-side helpers shared by the launchers (host_util.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

// debug / test switches (environment, read once per process; p2c_debug_reload_env() re-reads them)
struct P2cEnv {
    bool disable_tile, disable_fast, force_slim, disable_slim, disable_tma, tile_one_cta, tile_no_coop, disable_fwd_tile;
    int tile_cta, tile_clocks;
};

int p2c_current_device();
int p2c_sm_count();                                          // SMs of the CURRENT device
int p2c_allow_big_smem(const void* kernel, size_t bytes);    // once per (kernel, device); 0 or a cudaError_t
const P2cEnv& p2c_env();
extern "C" void p2c_debug_reload_env(void);
