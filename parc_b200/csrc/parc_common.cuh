// Host-side helpers shared by the C-ABI translation units: error text, launch checks, grid sizing.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/parc_b200.h"

namespace parc {

void set_error(const char* fmt, ...);
int check_launch(const char* what);

int launch_env_step(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_terrain_t* terrain,
                    const parc_env_config_t* cfg, const parc_env_buffers_t* env, uint32_t parts, const int64_t* env_ids,
                    long long max_rows, const int* env_ids32, const int* n_rows_dev, void* stream, long long first_env = 0);

int validate_env_step(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_terrain_t* terrain,
                      const parc_env_config_t* cfg, const parc_env_buffers_t* env, uint32_t parts);

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// B200 has 148 SMs; grids are sized in whole waves of the SM count where a kernel is grid-strided.
constexpr int kNumSMs = 148;

inline bool model_ok(const parc_char_model_t* m, const char* who) {
    if (m == nullptr) { set_error("%s: model is NULL", who); return false; }
    if (m->num_bodies < 1 || m->num_bodies > PARC_MAX_BODIES) {
        set_error("%s: num_bodies %d outside [1, %d]", who, m->num_bodies, PARC_MAX_BODIES);
        return false;
    }
    if (m->dof_size < 0 || m->dof_size > PARC_MAX_DOFS) {
        set_error("%s: dof_size %d outside [0, %d]", who, m->dof_size, PARC_MAX_DOFS);
        return false;
    }
    for (int j = 1; j < m->num_bodies; ++j) {
        if (m->parent[j] < 0 || m->parent[j] >= j) {
            set_error("%s: parent[%d] = %d must precede its child", who, j, m->parent[j]);
            return false;
        }
    }
    return true;
}

}  // namespace parc

#define PARC_REQUIRE(cond, ...)                              \
    do {                                                     \
        if (!(cond)) {                                       \
            parc::set_error(__VA_ARGS__);                    \
            return PARC_ERR_INVALID_ARG;                     \
        }                                                    \
    } while (0)
