// CARLA_SKELETON-specialised single-pass fused step (placeholder until the specialised kernel lands:
// reports "not supported" so p2c_fused_step takes the generic-tree path).
#include "common.cuh"

extern "C" int p2c_fused_carla_supported(const p2c_fused_step_t* p) {
    (void)p;
    return 0;
}
extern "C" int p2c_fused_carla(const p2c_fused_step_t* p, void* stream) {
    (void)p; (void)stream;
    return P2C_E_UNSUPPORTED;
}
extern "C" int p2c_fused_carla_launches(const p2c_fused_step_t* p) {
    (void)p;
    return 0;
}
