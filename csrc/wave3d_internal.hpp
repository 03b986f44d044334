// wave3d_internal.hpp — declarations shared by the CUDA translation units.
#pragma once
#include "wave3d.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <string>

namespace w3d {

void set_last_error(const std::string& s);
bool check(cudaError_t e, const char* what);  // false + last_error on failure

#define W3D_TRY(call)                                  \
    do {                                               \
        if (!::w3d::check((call), #call)) return -1;   \
    } while (0)

// Division of the stencil sums by the constant h^2.
//  0: h^2 is a power of two -> exact multiply by 1/h^2
//  1: three-op Markstein (faithful, not proven correctly rounded)
//  2: five-op Markstein with correctly rounded 1/h^2 -> equals IEEE division
//  3: IEEE division
int choose_div_mode(const wave3d_problem* p);

// Pre-encoded arguments of one fused-step launch (maps are encoded once per field
// by the runtime and reused for every step).
struct StepLaunch {
    CUtensorMap map_halo[2];  // per field: halo box (tz+4, ty+2, 1), used when the field is the stencil source
    CUtensorMap map_core[2];  // per field: core box (tz, ty, 1), used when the field is updated in place
    const wave3d_problem* prob;
    const wave3d_box* box;
    const wave3d_layout* lay;
    const wave3d_plan* plan;
    const double* tabs[6];
    double* field[2];   // layer n lives in field n & 1
    int n;              // (first) layer computed by the launch
    int nsteps;         // 1: one layer per launch; > 1: layers n .. n+nsteps-1 in one persistent cooperative launch
    int first;
    wave3d_peers peers[2];  // per written field; zero-initialised when absent
    uint64_t* err;       // table base (2 words per layer)
    uint32_t* scratch;
    uint32_t* epoch;     // multi-layer launches: one word per (tile, chunk) item, zeroed before the solve
    uint32_t* stepdone;  // multi-layer launches: 3*(timesteps+2) words (items per layer, then x-face items
                         // per layer and side), zeroed before the solve
    uint32_t* chunkdone; // multi-layer launches: 8 * 4 * WAVE3D_MAX_CHUNKS words (y/z face strips done per x-chunk)
    int yz_chunk_flags;  // y/z neighbours use the same x-cuts: hand-shake per (face, x-chunk)
    int noerr;           // timing aid: skip the error evaluation
    const double* desc;  // device pointer to {arg_t, coef} or nullptr
    // streamed multi-layer launch: descriptor table in device memory (NaN until uploaded), mapped host
    // memory for the layer maxima and their ready flags
    const double* desc_tab;
    double* host_res;
    uint32_t* host_flag;
    double spin_timeout_s;
    int div_mode;        // -1: choose
    cudaStream_t stream;
};

int encode_field_map(CUtensorMap* map, const double* base, const wave3d_layout* lay, int box_w, int box_h);
int launch_step(const StepLaunch& a);
int prepare_step_kernels(const wave3d_plan* plan, int div_mode);
// resident CTAs per SM for a plan candidate (queries the driver; <=0 if unknown)
int step_occupancy(int rj, int warps, int smem_bytes);
int plan_smem_bytes(int ty, int tz, int s1, int s0);

}  // namespace w3d
