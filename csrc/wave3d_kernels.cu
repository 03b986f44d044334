// wave3d_kernels.cu — sm_100a device code of the wave3d solver.
//
// Hot kernel: k_step<RJ,FIRST,DIV> — one launch per time layer that fuses what the
// reference spreads over K2/K3 (update), K4 (boundary planes), K5/K6 + 12 memcpys +
// MPI p2p (halo exchange), K7 + host loop (error max) and two memsets
// (ref: cuda_sol.cpp:384-421, cuda_sol_kernels.cu:24-268).
//
// Design (B200-first, not a translation):
//   * The stencil source layer u^{n-1} and the in-place layer u^{n-2} stream through
//     shared memory as y-z tiles, one x plane at a time, moved by TMA bulk tensor
//     loads (cp.async.bulk.tensor.3d -> UTMALDG) into mbarrier-guarded rings; threads
//     never issue a global load on the hot path.
//   * Each thread owns 2 z-cells x RJ rows, reads its 7-point neighbourhood from the
//     three resident planes and writes u^n with one 16-byte store over u^{n-2}.
//   * The analytic solution is separable: f = ((c*sx[i])*sy[j])*sz[k] with 1-D sine
//     tables -> bit-identical to the reference's per-cell trig at 3 multiplies.
//     max|u-f| is an in-register max; max|u-f|/|f| uses a monotone filter so the
//     exact IEEE division runs only for candidates of the maximum.
//   * Face cells are stored a second time, straight into the neighbour GPU's ghost
//     cells (peer-mapped memory over NVLink; own memory for the periodic self-wrap).
//     The last CTA of the grid publishes an epoch flag per neighbour with
//     st.release.sys; the next step's CTAs ld.acquire.sys-spin on their inbound flags.
//     No pack/unpack kernels, no host staging, no MPI/NCCL on this path.
//   * Per-layer maxima land in err[2n..2n+1] via atomicMax on the u64 bit pattern
//     (values are >= 0; NaN never wins, like fmax) and are forwarded to rank 0's
//     table with a system-scope atomic (the reference's deferred MPI_Reduce).
#include "wave3d_internal.hpp"

#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>

namespace w3d {

static thread_local std::string g_last_error;
void set_last_error(const std::string& s) { g_last_error = s; }
bool check(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    g_last_error = std::string(what) + ": " + cudaGetErrorString(e);
    return false;
}

// ----------------------------------------------------------------- PTX helpers

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    return done != 0;
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
        : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ------------------------------------------------------------- step parameters

struct alignas(64) StepParams {
    CUtensorMap map_halo[2];  // per field: halo box (tz+4, ty+2, 1) — used when the field is the stencil source
    CUtensorMap map_core[2];  // per field: core box (tz, ty, 1)    — used when the field is updated in place
    double* field[2];         // layer n lives in field n & 1
    const double* sx;
    const double* sy;
    const double* sz;
    const double* rsx;
    const double* rsy;
    const double* rsz;
    double* dst[2][6];        // per written field: neighbour ghost cell that receives my face cell (0,0)
    int64_t sa[6];
    int64_t sb[6];
    uint32_t* flag_out[6];
    const uint32_t* flag_in[6];
    uint64_t* err;       // table base: layer n -> err[2n], err[2n+1]
    uint64_t* root_err;  // rank 0's table base or nullptr
    uint32_t* scratch;
    uint32_t* epoch;     // multi-step mode: last completed layer per (tile, chunk) item
    uint32_t* stepdone;  // multi-step mode: completed items per layer
    uint32_t* facedone;  // multi-step mode: completed x-face items per layer (2 words per layer)
    const double* desc;  // optional uploaded step descriptor {arg_t, coef} (single-step mode)
    const double* desc_tab;   // multi-step streamed mode: descriptor table [layer][2] in device memory, NaN until uploaded
    double* host_res;         // multi-step streamed mode: mapped host memory, layer maxima [layer][2]
    uint32_t* host_flag;      // multi-step streamed mode: mapped host memory, 1 when host_res[layer] is valid
    int64_t ps, pitch;
    double coef, arg_t;
    double a_t, tau, twopi;  // multi-step mode: cos(fma(a_t*n, tau, twopi)) per layer
    double hx2, hy2, hz2, rx, ry, rz;
    unsigned long long timeout_ns;
    int n0, nsteps;          // layers n0 .. n0+nsteps-1 are computed by this launch
    int oz, X, Y, Z;
    int lo1, hi1, lo2, hi2;
    int elo0, ehi0;
    int nty, ntz, nchunks, ty, tz;
    int xcut[18];        // x-chunk boundaries by POSITION along x
    signed char pos_of[16];  // processing index (queue order within a layer) -> position
    signed char pi_of[16];   // position -> processing index
    int self_wrap;       // the x neighbours are this rank itself (periodic wrap inside one box)
    int s1, s0;
    int send_x0, send_x1;
    uint32_t push_mask;  // bit f set: dst[.][f] != nullptr
    int dbg;             // write phase timestamps into g_dbg_ts
    int fence_mode;      // 0: no peers, gpu-scope fences only; 1: gpu per CTA + one sys fence per layer; 2: sys in every CTA
};

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Division of a stencil sum by the constant d = h^2 with r = rn(1/d).
template <int DIV>
__device__ __forceinline__ double div_const(double x, double d, double r) {
    if (DIV == 0) {
        return __dmul_rn(x, r);  // d is a power of two: exact
    } else if (DIV == 1) {
        double q = __dmul_rn(x, r);
        double e = __fma_rn(-d, q, x);
        return __fma_rn(e, r, q);
    } else if (DIV == 2) {
        // Markstein: q1 is faithful after one correction; with r = RN(1/d) a second
        // exact-remainder correction of a faithful quotient is correctly rounded.
        double q = __dmul_rn(x, r);
        double e = __fma_rn(-d, q, x);
        q = __fma_rn(e, r, q);
        e = __fma_rn(-d, q, x);
        return __fma_rn(e, r, q);
    } else {
        return __ddiv_rn(x, d);
    }
}

// ---- shared-memory access by 32-bit shared address (keeps LDS, never generic LD)
__device__ __forceinline__ double2 lds2(uint32_t a) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ double lds1(uint32_t a) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

struct ErrAcc {
    double mabs, mrel, best, thr;
};

// max that ignores a NaN candidate (the reference's fmax / host '>' semantics) in 3 instructions
__device__ __forceinline__ double max_keep(double m, double v) { return v > m ? v : m; }

// Error bookkeeping of one row (1 or 2 cells). cand approximates a/|f| to ~1e-15 relative via
// reciprocal tables; the exact IEEE division runs only when a cell can still be the maximum
// (band 2^-40 >> table error), so the result equals max over all cells of the reference's
// `abs_error / fabs(f)` (ref: cuda_sol_kernels.cu:42-45).
template <bool TWO>
__device__ __forceinline__ void err_row(ErrAcc& e, double ua, double ub, double fxy, double rxy, double sz0, double sz1,
                                        double rz0, double rz1) {
    const double fa = __dmul_rn(fxy, sz0);
    const double aa = fabs(__dsub_rn(ua, fa));
    e.mabs = max_keep(e.mabs, aa);
    const double ca = __dmul_rn(__dmul_rn(aa, rxy), rz0);
    double fb = 0.0, ab = 0.0, cb = 0.0;
    if (TWO) {
        fb = __dmul_rn(fxy, sz1);
        ab = fabs(__dsub_rn(ub, fb));
        e.mabs = max_keep(e.mabs, ab);
        cb = __dmul_rn(__dmul_rn(ab, rxy), rz1);
    }
    if (ca >= e.thr || (TWO && cb >= e.thr)) {
        e.mrel = fmax(e.mrel, __ddiv_rn(aa, fabs(fa)));
        e.best = fmax(e.best, ca);
        if (TWO) {
            e.mrel = fmax(e.mrel, __ddiv_rn(ab, fabs(fb)));
            e.best = fmax(e.best, cb);
        }
        e.thr = e.best * (1.0 - 0x1p-40);
    }
}

__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ void store_pair(double* p, double a, double b, bool two) {
    if (two && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
        *reinterpret_cast<double2*>(p) = make_double2(a, b);
    } else {
        p[0] = a;
        if (two) p[1] = b;
    }
}

// Face cells are stored a second time, straight into the neighbour's ghost cells (peer memory).
// x faces: whole planes, hit on at most two planes per rank -> out of line; y and z faces are
// stored inline in the march through per-thread pointers.
__device__ __noinline__ void push_row_x(const StepParams& p, int w, int i, int j, int kk, double ua, double ub,
                                        bool two) {
    if ((p.push_mask & 1u) && i == p.send_x0) store_pair(p.dst[w][0] + j * p.sa[0] + kk, ua, ub, two);
    if ((p.push_mask & 2u) && i == p.send_x1) store_pair(p.dst[w][1] + j * p.sa[1] + kk, ua, ub, two);
}

// Optional per-CTA phase timestamps of the last k_step launch (WAVE3D_DBG_TS=1; scripts/cta_timeline.py)
__device__ unsigned long long g_dbg_ts[2048 * 8];
// Per-CTA phase accumulators (SM clock cycles) of the last launch, same switch: producer lane
// [0] item fetch (queue atomic), [1] dependency poll, [2] acquire fences, [3] waits for a free ring slot,
// [4] publishing finished items, [5] items, [6] items that had to wait, [7] poll cycles of rank-face items,
// [8] rank-face items, [9] kernel cycles; compute warp 0: [10] waits for the next item id, [11] waits for
// the first two planes of an item, [12] waits for planes inside the march, [13] item set-up, [14] march.
__device__ unsigned long long g_dbg_acc[2048 * 16];

constexpr int kMaxWarps = 7;    // compute warps per CTA (one more warp is the TMA producer): 256 threads, 2 CTAs per SM
constexpr int kTZ = 64;         // tile width in z: one warp row = 32 lanes x 2 cells
constexpr int kW1 = kTZ + 4;    // halo slot row length in doubles (2 halo columns each side keep 16-byte alignment)

// MULTI = false: one layer per launch (layer p.n0); the last CTA publishes the layer.
// MULTI = true : layers n0 .. n0+nsteps-1 in ONE persistent launch. Items of consecutive layers form one
//   queue; an item of layer n starts as soon as its own and its six face-adjacent items of layer n-1 are
//   complete (per-item epoch words, release/acquire at gpu scope) and — only if it touches a rank face —
//   the neighbour rank has published layer n-1. No grid-wide barrier, no launch gap, no kernel tail
//   between layers: SMs that finish a layer early run ahead into the next one, interior items overlap
//   the neighbours' flag latency. Needs all CTAs co-resident (cooperative launch).
template <int RJ, bool FIRST, int DIV, bool MULTI>
__global__ void __launch_bounds__((kMaxWarps + 1) * 32, 2) k_step(const __grid_constant__ StepParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double s_bcast[3];
    __shared__ double s_red[kMaxWarps][2];
    __shared__ int s_item[8];                // item ids handed from the producer lane to the compute warps
    __shared__ __align__(8) uint64_t s_pbar[8];  // multi-step mode: compute warps -> producer lane, item finished
    __shared__ __align__(8) uint64_t s_ibar[8];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ncw = (int)(blockDim.x >> 5) - 1;  // compute warps
    const uint32_t slot1_bytes = (uint32_t)(((p.ty + 2) * kW1 * 8 + 127) & ~127);
    const uint32_t slot0_bytes = (uint32_t)((p.ty * kTZ * 8 + 127) & ~127);
    const uint32_t ring1 = (smem_u32(smem_raw) + 127u) & ~127u;
    const uint32_t ring0 = ring1 + (uint32_t)p.s1 * slot1_bytes;
    const uint32_t bars = ring0 + (uint32_t)(FIRST ? 0 : p.s0) * slot0_bytes;
    const uint32_t full1 = bars, empty1 = full1 + 8u * (uint32_t)p.s1;
    const uint32_t full0 = empty1 + 8u * (uint32_t)p.s1, empty0 = full0 + 8u * (uint32_t)p.s0;
    const uint32_t ibar = smem_u32(s_ibar);
    const uint32_t pbar = smem_u32(s_pbar);

    // ---- persistent CTAs pulling work from a device-side queue. Every tile column is cut into
    // p.nchunks x-chunks whose lengths shrink towards the end of the column (guided self-scheduling);
    // items are ordered chunk-major, so CTAs that run at the same time march through the same x range
    // of adjacent tiles (halos shared in L2, DRAM pages streamed in full) while SMs that get more
    // memory bandwidth simply take more of the small late items (measured spread between SMs: 20 %).
    const int ntiles = p.nty * p.ntz;
    const int nitems = ntiles * p.nchunks;
    const int ntotal = nitems * p.nsteps;
    const bool dbg = p.dbg && blockIdx.x < 2048;
    if (dbg && tid == 0) g_dbg_ts[blockIdx.x * 8 + 0] = globaltimer_ns();

    if (tid == 0) {
        for (int s = 0; s < 8; s++) {
            mbar_init(ibar + 8u * s, 1);
            mbar_init(pbar + 8u * s, (uint32_t)ncw);
        }
        for (int s = 0; s < p.s1; s++) {
            mbar_init(full1 + 8u * s, 1);
            mbar_init(empty1 + 8u * s, (uint32_t)ncw);
        }
        if (!FIRST)
            for (int s = 0; s < p.s0; s++) {
                mbar_init(full0 + 8u * s, 1);
                mbar_init(empty0 + 8u * s, (uint32_t)ncw);
            }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        // the layer's time argument and coefficient come from the launch parameters or, in per-step
        // (end-to-end) mode, from the descriptor the host uploaded for this step
        const double arg_t = p.desc ? p.desc[0] : p.arg_t;
        const double c = cos(arg_t);
        s_bcast[0] = c;
        s_bcast[1] = __ddiv_rn(1.0, fabs(c));
        s_bcast[2] = p.desc ? p.desc[1] : p.coef;
    }
    __syncthreads();
    if (dbg && tid == 0) g_dbg_ts[blockIdx.x * 8 + 1] = globaltimer_ns();

    if (warp == ncw) {
        // ================= producer warp: one lane feeds both rings with TMA =================
        if (lane == 0) {
            long long a_q = 0, a_poll = 0, a_acq = 0, a_slot = 0, a_pub = 0, a_pollf = 0;
            unsigned a_items = 0, a_waited = 0, a_fitems = 0;
            const long long a_t0 = dbg ? clock64() : 0;
            // Neighbours must have finished the previous layer: their ghost stores are complete (RAW)
            // and they no longer read the ghosts this layer overwrites (WAR). Nothing of this CTA
            // touches global memory before its first plane has arrived, i.e. after this wait.
            // every wait is bounded on its own (a dead peer must not hang the GPU, a long solve must not trip it)
            auto wait_flag = [&](const uint32_t* fl, uint32_t need, bool sys) {
                const unsigned long long t0 = globaltimer_ns();
                while ((sys ? ld_acquire_sys(fl) : ld_acquire_gpu(fl)) < need) {
                    __nanosleep(32);
                    if (globaltimer_ns() - t0 > p.timeout_ns) {
                        atomicExch(&p.scratch[1], 1u);
                        break;
                    }
                }
            };
            if (!MULTI) {
                for (int f = 0; f < 6; f++)
                    if (p.flag_in[f]) wait_flag(p.flag_in[f], (uint32_t)p.n0, true);
                asm volatile("fence.proxy.async;" ::: "memory");
            }
            const uint32_t box1_bytes = (uint32_t)((p.ty + 2) * kW1 * 8);
            const uint32_t box0_bytes = (uint32_t)(p.ty * kTZ * 8);
            int s1i = 0, s0i = 0;
            uint32_t ph1 = 1, ph0 = 1;  // parity of the "previous" phase: passes immediately on first use
            // multi-layer mode: this lane also publishes finished items. It is idle most of the time, so
            // the gpu-scope fence (which has to wait for the SM's outstanding stores) never sits on a
            // compute warp's critical path; it is cumulative over the warps' block-scope arrivals.
            int pub_kq = 0, issued_kq = 0;
            int pub_gid[8];
            auto publish = [&](bool block) {
                const long long c0 = dbg ? clock64() : 0;
                while (MULTI && pub_kq < issued_kq) {
                    const uint32_t bar = pbar + 8u * (pub_kq & 7), par = (uint32_t)((pub_kq >> 3) & 1);
                    if (block) mbar_wait(bar, par);
                    else if (!mbar_test(bar, par)) break;
                    const int gid = pub_gid[pub_kq & 7];
                    const int n = p.n0 + gid / nitems, item = gid % nitems;
                    __threadfence();
                    st_release_gpu(p.epoch + item, (uint32_t)n);
                    // The x faces are complete as soon as the two end chunks of every tile column are: they
                    // come first in the queue, so the x neighbours get their flag almost a whole layer
                    // before they need it (y/z faces span all chunks and are published with the layer).
                    const int cpos = p.pos_of[item / ntiles];
                    if (cpos == 0 && p.flag_out[0] && atomicAdd(p.facedone + 2 * n, 1u) == (uint32_t)ntiles - 1u) {
                        __threadfence_system();
                        st_release_sys(p.flag_out[0], (uint32_t)n + 1u);
                    }
                    if (cpos == p.nchunks - 1 && p.flag_out[1] &&
                        atomicAdd(p.facedone + 2 * n + 1, 1u) == (uint32_t)ntiles - 1u) {
                        __threadfence_system();
                        st_release_sys(p.flag_out[1], (uint32_t)n + 1u);
                    }
                    if (atomicAdd(p.stepdone + n, 1u) == (uint32_t)nitems - 1u) {
                        // layer n is complete on this rank
                        unsigned long long* e = reinterpret_cast<unsigned long long*>(p.err + 2 * (size_t)n);
                        if (p.fence_mode != 0) __threadfence_system();
                        if (p.root_err) {
                            const unsigned long long va = atomicMax(e, 0ull), vr = atomicMax(e + 1, 0ull);
                            unsigned long long* re = reinterpret_cast<unsigned long long*>(p.root_err + 2 * (size_t)n);
                            atomicMax_system(re, va);
                            atomicMax_system(re + 1, vr);
                        }
                        for (int f = 2; f < 6; f++)
                            if (p.flag_out[f]) st_release_sys(p.flag_out[f], (uint32_t)n + 1u);
                        if (p.host_res) {
                            // streamed mode: the layer's result goes straight to the host (mapped pinned memory)
                            const unsigned long long va = atomicMax(e, 0ull), vr = atomicMax(e + 1, 0ull);
                            reinterpret_cast<volatile unsigned long long*>(p.host_res)[2 * (size_t)n] = va;
                            reinterpret_cast<volatile unsigned long long*>(p.host_res)[2 * (size_t)n + 1] = vr;
                            __threadfence_system();
                            st_release_sys(p.host_flag + n, 1u);
                        }
                    }
                    pub_kq++;
                }
                if (dbg) a_pub += clock64() - c0;
            };
            for (int kq = 0;; kq++) {
                // the first item is the CTA's own index (no atomic on the critical path), later ones
                // come from the queue, which starts at gridDim.x
                const long long cq0 = dbg ? clock64() : 0;
                int gid = kq == 0 ? (int)blockIdx.x : (int)(gridDim.x + atomicAdd(&p.scratch[2], 1u));
                if (gid >= ntotal) gid = -1;
                if (dbg) a_q += clock64() - cq0;
                s_item[kq & 7] = gid;
                mbar_arrive(ibar + 8u * (kq & 7));  // release: the compute warps read s_item after this phase
                if (gid < 0) {
                    publish(true);
                    break;
                }
                pub_gid[kq & 7] = gid;
                issued_kq = kq + 1;
                const int n = p.n0 + (MULTI ? gid / nitems : 0);
                const int item = MULTI ? gid % nitems : gid;
                const int w = n & 1;  // layer n is written into field w, the stencil source is field w^1
                const int pidx = item / ntiles, tile = item - pidx * ntiles;
                const int chunk = p.pos_of[pidx];  // position of the chunk along x
                const int xa = p.xcut[chunk];
                const int nplanes = p.xcut[chunk + 1] - xa;
                const int tyi = tile / p.ntz, tzi = tile - tyi * p.ntz;
                const int j0 = p.lo1 + tyi * p.ty;
                const int k0 = p.lo2 + tzi * kTZ;
                if (MULTI) {
                    // RAW on the source layer and WAR on the in-place layer are both covered by "this item
                    // and its face neighbours finished layer n-1"; items on a rank face also need the
                    // neighbour rank's layer n-1 (published as epoch n). All words are polled together
                    // (independent loads, one L2 round trip per poll), then one fence makes it an acquire.
                    const uint32_t* wp[13];
                    uint32_t wn[13];
                    int nw = 0;
                    if (n > p.n0) {
                        const uint32_t need = (uint32_t)(n - 1);
                        wp[nw] = p.epoch + item; wn[nw++] = need;
                        // x neighbours by position; with the periodic wrap inside this box the two end
                        // chunks of a tile column are neighbours too (ghost planes written by each other)
                        int cm = chunk - 1, cp = chunk + 1;
                        if (p.self_wrap) {
                            if (cm < 0) cm = p.nchunks - 1;
                            if (cp == p.nchunks) cp = 0;
                        }
                        if (cm >= 0 && cm != chunk) { wp[nw] = p.epoch + p.pi_of[cm] * ntiles + tile; wn[nw++] = need; }
                        if (cp < p.nchunks && cp != chunk && cp != cm) { wp[nw] = p.epoch + p.pi_of[cp] * ntiles + tile; wn[nw++] = need; }
                        if (tyi > 0) { wp[nw] = p.epoch + item - p.ntz; wn[nw++] = need; }
                        if (tyi + 1 < p.nty) { wp[nw] = p.epoch + item + p.ntz; wn[nw++] = need; }
                        if (tzi > 0) { wp[nw] = p.epoch + item - 1; wn[nw++] = need; }
                        if (tzi + 1 < p.ntz) { wp[nw] = p.epoch + item + 1; wn[nw++] = need; }
                    }
                    const uint32_t needf = (uint32_t)n;
                    const int nepoch = nw;
                    if (chunk == 0 && p.flag_in[0]) { wp[nw] = p.flag_in[0]; wn[nw++] = needf; }
                    if (chunk == p.nchunks - 1 && p.flag_in[1]) { wp[nw] = p.flag_in[1]; wn[nw++] = needf; }
                    if (tyi == 0 && p.flag_in[2]) { wp[nw] = p.flag_in[2]; wn[nw++] = needf; }
                    if (tyi == p.nty - 1 && p.flag_in[3]) { wp[nw] = p.flag_in[3]; wn[nw++] = needf; }
                    if (tzi == 0 && p.flag_in[4]) { wp[nw] = p.flag_in[4]; wn[nw++] = needf; }
                    if (tzi == p.ntz - 1 && p.flag_in[5]) { wp[nw] = p.flag_in[5]; wn[nw++] = needf; }
                    if (dbg) {
                        a_items++;
                        if (nw > nepoch) a_fitems++;
                    }
                    if (nw) {
                        const unsigned long long t0 = globaltimer_ns();
                        const long long cp0 = dbg ? clock64() : 0;
                        bool waited = false;
                        for (;;) {
                            bool ok = true;
                            for (int q = 0; q < nw; q++)
                                ok = ok && (*reinterpret_cast<const volatile uint32_t*>(wp[q]) >= wn[q]);
                            if (ok) break;
                            waited = true;
                            publish(false);  // never sit on finished items while waiting for somebody else's
                            __nanosleep(32);
                            if (globaltimer_ns() - t0 > p.timeout_ns) {
                                atomicExch(&p.scratch[1], 1u);
                                break;
                            }
                        }
                        // acquire side: one gpu-scope fence for all epoch words (seven dependent acquire
                        // loads cost ~5 us per item); the few rank-face items re-read their flag words with
                        // a system-scope acquire load
                        const long long cp1 = dbg ? clock64() : 0;
                        if (nepoch) __threadfence();
                        for (int q = nepoch; q < nw; q++) (void)ld_acquire_sys(wp[q]);
                        if (dbg) {
                            a_poll += cp1 - cp0;
                            if (nw > nepoch) a_pollf += cp1 - cp0;
                            a_acq += clock64() - cp1;
                            a_waited += waited ? 1u : 0u;
                        }
                    }
                    asm volatile("fence.proxy.async;" ::: "memory");
                }
                const CUtensorMap* msrc = &p.map_halo[w ^ 1];
                const CUtensorMap* mio = &p.map_core[w];
                const int c0_src = p.oz + k0 - 2, c1_src = j0;  // storage row(j0) - 1 == j0
                const int c0_io = p.oz + k0, c1_io = j0 + 1;
                for (int q = 0; q < nplanes + 2; q++) {
                    // source plane seq q = owned plane xa-1+q = storage plane xa+q
                    publish(false);
                    const long long cs0 = dbg ? clock64() : 0;
                    mbar_wait(empty1 + 8u * s1i, ph1);
                    if (dbg) a_slot += clock64() - cs0;
                    mbar_expect_tx(full1 + 8u * s1i, box1_bytes);
                    tma_load_3d(ring1 + (uint32_t)s1i * slot1_bytes, msrc, c0_src, c1_src, xa + q, full1 + 8u * s1i);
                    if (++s1i == p.s1) { s1i = 0; ph1 ^= 1; }
                    if (!FIRST && q >= 2) {
                        // in-place plane of iteration q-2 = owned plane xa+q-2 = storage plane xa+q-1
                        const long long cs1 = dbg ? clock64() : 0;
                        mbar_wait(empty0 + 8u * s0i, ph0);
                        if (dbg) a_slot += clock64() - cs1;
                        mbar_expect_tx(full0 + 8u * s0i, box0_bytes);
                        tma_load_3d(ring0 + (uint32_t)s0i * slot0_bytes, mio, c0_io, c1_io, xa + q - 1, full0 + 8u * s0i);
                        if (++s0i == p.s0) { s0i = 0; ph0 ^= 1; }
                    }
                }
            }
            if (dbg) {
                unsigned long long* acc = g_dbg_acc + (size_t)blockIdx.x * 16;
                acc[0] = (unsigned long long)a_q; acc[1] = (unsigned long long)a_poll; acc[2] = (unsigned long long)a_acq;
                acc[3] = (unsigned long long)a_slot; acc[4] = (unsigned long long)a_pub; acc[5] = a_items;
                acc[6] = a_waited; acc[7] = (unsigned long long)a_pollf; acc[8] = a_fitems;
                acc[9] = (unsigned long long)(clock64() - a_t0);
            }
        }
    } else {
        // ================================ compute warps ================================
        long long b_item = 0, b_first = 0, b_full = 0, b_setup = 0, b_march = 0;
        const bool dbgw = dbg && tid == 0;
        double cosn = s_bcast[0], rcos = s_bcast[1];
        double coef = s_bcast[2];
        const double hx2 = p.hx2, hy2 = p.hy2, hz2 = p.hz2, rx = p.rx, ry = p.ry, rz = p.rz;
        const int64_t ps = p.ps;
        const bool xpush = (p.push_mask & 3u) != 0;
        ErrAcc ea = {0.0, 0.0, 0.0, 0.0};
        // byte offset of (my first row's upper halo row, my first cell) inside a halo slot / core slot
        const uint32_t toff1 = (uint32_t)(((warp * RJ) * kW1 + 2 + 2 * lane) * 8);
        const uint32_t toff0 = (uint32_t)(((warp * RJ) * kTZ + 2 * lane) * 8);
        // ring cursors: F = next source plane to arrive, B = oldest source plane still held, F0 = in-place ring
        int fi = 0, bi = 0, f0i = 0;
        uint32_t fph = 0, f0ph = 0;
        int cur_n = -1;

        for (int kq = 0;; kq++) {
            const long long ci0 = dbgw ? clock64() : 0;
            mbar_wait(ibar + 8u * (kq & 7), (uint32_t)((kq >> 3) & 1));
            const long long ci1 = dbgw ? clock64() : 0;
            if (dbgw) b_item += ci1 - ci0;
            const int gid = s_item[kq & 7];
            if (gid < 0) break;
            const int n = p.n0 + (MULTI ? gid / nitems : 0);
            const int item = MULTI ? gid % nitems : gid;
            const int w = n & 1;
            const int pidx = item / ntiles, tile = item - pidx * ntiles;
            const int chunk = p.pos_of[pidx];
            const int xa = p.xcut[chunk];
            const int nplanes = p.xcut[chunk + 1] - xa;
            const int j0 = p.lo1 + (tile / p.ntz) * p.ty;
            const int k0 = p.lo2 + (tile % p.ntz) * kTZ;
            double* const* dstw = p.dst[w];
            if (MULTI && n != cur_n) {
                // a new layer: its cosine, and fresh maxima (within a layer the running maxima and the
                // rel-error filter threshold carry over from item to item; re-publishing a maximum is harmless)
                cur_n = n;
                double arg_t = __fma_rn(__dmul_rn(p.a_t, (double)n), p.tau, p.twopi);
                if (p.desc_tab) {
                    // streamed (end-to-end) mode: this layer's inputs are being copied host->device while
                    // the kernel runs; the table holds NaN until the copy of layer n has landed
                    const volatile double* d = p.desc_tab + 2 * (size_t)n;
                    double a0 = d[0], a1 = d[1];
                    if (lane == 0) {
                        const unsigned long long t0 = globaltimer_ns();
                        while ((a0 != a0 || a1 != a1) && globaltimer_ns() - t0 < p.timeout_ns) {
                            __nanosleep(64);
                            a0 = d[0];
                            a1 = d[1];
                        }
                        if (a0 != a0 || a1 != a1) atomicExch(&p.scratch[1], 1u);
                    }
                    arg_t = __shfl_sync(0xffffffffu, a0, 0);
                    coef = __shfl_sync(0xffffffffu, a1, 0);
                }
                cosn = cos(arg_t);
                rcos = __ddiv_rn(1.0, fabs(cosn));
                ea.mabs = ea.mrel = ea.best = ea.thr = 0.0;
            }

            const int kk = k0 + 2 * lane;  // owned k of my first cell
            const bool vk0 = kk <= p.hi2, vk1 = kk + 1 <= p.hi2;
            const int jr = j0 + warp * RJ;  // owned j of my first row
            bool vj[RJ];
            double syq[RJ], rsyq[RJ];
            double* orow[RJ];
            // y faces: at most one of my rows is the rank's first / last owned row
            int yq = -1;
            double* yptr = nullptr;
            int64_t ystep = 0;
            // z faces: my first cell is the z- or z+ face column, or my second cell is the z+ one
            double* zpa = nullptr;  // receives ua (row q at zpa + q * zsb)
            double* zpb = nullptr;  // receives ub
            int64_t zsb = 0, zstep = 0;
            if ((p.push_mask & 16u) && kk == 0) {
                zpa = dstw[4] + xa * p.sa[4] + jr * p.sb[4];
                zsb = p.sb[4];
                zstep = p.sa[4];
            }
            if (p.push_mask & 32u) {
                if (kk == p.Z - 1) {
                    zpa = dstw[5] + xa * p.sa[5] + jr * p.sb[5];
                    zsb = p.sb[5];
                    zstep = p.sa[5];
                } else if (kk + 1 == p.Z - 1) {
                    zpb = dstw[5] + xa * p.sa[5] + jr * p.sb[5];
                    zsb = p.sb[5];
                    zstep = p.sa[5];
                }
            }
#pragma unroll
            for (int q = 0; q < RJ; q++) {
                const int j = jr + q;
                vj[q] = j <= p.hi1 && vk0;
                syq[q] = vj[q] ? p.sy[j] : 0.0;
                rsyq[q] = vj[q] ? p.rsy[j] : 0.0;
                orow[q] = p.field[w] + (int64_t)(xa + 1) * p.ps + (int64_t)(j + 1) * p.pitch + (p.oz + kk);
                if (vj[q] && (p.push_mask & 4u) && j == 0) {
                    yq = q;
                    yptr = dstw[2] + xa * p.sa[2] + kk;
                    ystep = p.sa[2];
                }
                if (vj[q] && (p.push_mask & 8u) && j == p.Y - 1) {
                    yq = q;
                    yptr = dstw[3] + xa * p.sa[3] + kk;
                    ystep = p.sa[3];
                }
            }
            const bool zany = zpa != nullptr || zpb != nullptr;
            const double szk0 = vk0 ? p.sz[kk] : 0.0, szk1 = vk1 ? p.sz[kk + 1] : 0.0;
            const double rsz0 = vk0 ? p.rsz[kk] : 0.0, rsz1 = vk1 ? p.rsz[kk + 1] : 0.0;

            // first two source planes of the segment
            const long long cf0 = dbgw ? clock64() : 0;
            if (dbgw) b_setup += cf0 - ci1;
            mbar_wait(full1 + 8u * fi, fph);
            uint32_t am = ring1 + (uint32_t)fi * slot1_bytes;
            if (++fi == p.s1) { fi = 0; fph ^= 1; }
            mbar_wait(full1 + 8u * fi, fph);
            uint32_t ac = ring1 + (uint32_t)fi * slot1_bytes;
            if (++fi == p.s1) { fi = 0; fph ^= 1; }

            if (dbg && tid == 0 && kq == 0) g_dbg_ts[blockIdx.x * 8 + 2] = globaltimer_ns();  // first data is here
            const long long cm0 = dbgw ? clock64() : 0;
            if (dbgw) b_first += cm0 - cf0;
            for (int t = 0; t < nplanes; t++) {
                const int i = xa + t;  // owned x index of the centre plane
                const long long cw0 = dbgw ? clock64() : 0;
                mbar_wait(full1 + 8u * fi, fph);
                const uint32_t ap = ring1 + (uint32_t)fi * slot1_bytes;
                if (++fi == p.s1) { fi = 0; fph ^= 1; }
                uint32_t a0 = 0;
                if (!FIRST) {
                    mbar_wait(full0 + 8u * f0i, f0ph);
                    a0 = ring0 + (uint32_t)f0i * slot0_bytes;
                }
                if (dbgw) b_full += clock64() - cw0;

                const bool ei = i >= p.elo0 && i <= p.ehi0;
                const double cs = __dmul_rn(cosn, __ldg(p.sx + i));
                const double rcs = __dmul_rn(rcos, __ldg(p.rsx + i));
                const bool pslow = xpush && (i == p.send_x0 || i == p.send_x1);

                double2 rowc[RJ + 2];
#pragma unroll
                for (int q = 0; q < RJ + 2; q++) rowc[q] = lds2(ac + toff1 + (uint32_t)(q * kW1 * 8));

#pragma unroll
                for (int q = 0; q < RJ; q++) {
                    const uint32_t rc = toff1 + (uint32_t)((q + 1) * kW1 * 8);
                    const double2 c = rowc[q + 1], ym = rowc[q], yp = rowc[q + 2];
                    const double zl = lds1(ac + rc - 8u);
                    const double zr = lds1(ac + rc + 16u);
                    const double2 xm = lds2(am + rc);
                    const double2 xp = lds2(ap + rc);
                    double2 g0 = c;
                    if (!FIRST) g0 = lds2(a0 + toff0 + (uint32_t)(q * kTZ * 8));

                    // exact expression order of the reference build
                    const double c2a = __dadd_rn(c.x, c.x), c2b = __dadd_rn(c.y, c.y);
                    const double sa_ = FIRST ? c.x : __dsub_rn(c2a, g0.x);
                    const double sb_ = FIRST ? c.y : __dsub_rn(c2b, g0.y);
                    const double lxa = div_const<DIV>(__dsub_rn(__dadd_rn(xp.x, xm.x), c2a), hx2, rx);
                    const double lxb = div_const<DIV>(__dsub_rn(__dadd_rn(xp.y, xm.y), c2b), hx2, rx);
                    const double lya = div_const<DIV>(__dsub_rn(__dadd_rn(yp.x, ym.x), c2a), hy2, ry);
                    const double lyb = div_const<DIV>(__dsub_rn(__dadd_rn(yp.y, ym.y), c2b), hy2, ry);
                    const double lza = div_const<DIV>(__dsub_rn(__dadd_rn(c.y, zl), c2a), hz2, rz);
                    const double lzb = div_const<DIV>(__dsub_rn(__dadd_rn(zr, c.x), c2b), hz2, rz);
                    const double ua = __fma_rn(coef, __dadd_rn(__dadd_rn(lxa, lya), lza), sa_);
                    const double ub = __fma_rn(coef, __dadd_rn(__dadd_rn(lxb, lyb), lzb), sb_);

                    if (vj[q]) {
                        if (vk1) *reinterpret_cast<double2*>(orow[q]) = make_double2(ua, ub);
                        else *orow[q] = ua;
                        if (pslow) push_row_x(p, w, i, jr + q, kk, ua, ub, vk1);
                        if (q == yq) store_pair(yptr, ua, ub, vk1);
                        if (zany) {
                            if (zpa) zpa[q * zsb] = ua;
                            if (zpb && vk1) zpb[q * zsb] = ub;
                        }
                        if (ei) {
                            const double fxy = __dmul_rn(cs, syq[q]);
                            const double rxy = __dmul_rn(rcs, rsyq[q]);
                            if (vk1) err_row<true>(ea, ua, ub, fxy, rxy, szk0, szk1, rsz0, rsz1);
                            else err_row<false>(ea, ua, ub, fxy, rxy, szk0, szk1, rsz0, rsz1);
                        }
                    }
                    orow[q] += ps;
                }
                if (yq >= 0) yptr += ystep;
                if (zany) {
                    if (zpa) zpa += zstep;
                    if (zpb) zpb += zstep;
                }

                // this warp is done with source plane i-1 and in-place plane i
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(empty1 + 8u * bi);
                    if (!FIRST) mbar_arrive(empty0 + 8u * f0i);
                }
                if (++bi == p.s1) bi = 0;
                if (!FIRST) {
                    if (++f0i == p.s0) { f0i = 0; f0ph ^= 1; }
                }
                am = ac;
                ac = ap;
            }
            if (dbgw) b_march += clock64() - cm0;
            // the last two source planes of the segment are no longer needed either
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(empty1 + 8u * bi);
                mbar_arrive(empty1 + 8u * (bi + 1 == p.s1 ? 0 : bi + 1));
            }
            bi += 2;
            if (bi >= p.s1) bi -= p.s1;

            if (MULTI) {
                // ---- item complete: this layer's maxima, then hand the item to the producer lane
                const double wa = warp_max(ea.mabs), wr = warp_max(ea.mrel);
                if (lane == 0) {
                    unsigned long long* e = reinterpret_cast<unsigned long long*>(p.err + 2 * (size_t)n);
                    // thousands of items per layer share these two words: look before the atomic (same-address
                    // atomics serialise in one L2 slice; after the first items hardly any maximum still grows)
                    const unsigned long long ba = (unsigned long long)__double_as_longlong(wa);
                    const unsigned long long br = (unsigned long long)__double_as_longlong(wr);
                    if (wa > 0.0 && ba > *reinterpret_cast<volatile unsigned long long*>(e)) atomicMax(e, ba);
                    if (wr > 0.0 && br > *reinterpret_cast<volatile unsigned long long*>(e + 1)) atomicMax(e + 1, br);
                    mbar_arrive(pbar + 8u * (kq & 7));  // release at CTA scope; the producer lane publishes
                }
                __syncwarp();
            }
        }

        if (dbg && tid == 0) g_dbg_ts[blockIdx.x * 8 + 3] = globaltimer_ns();
        if (dbgw) {
            unsigned long long* acc = g_dbg_acc + (size_t)blockIdx.x * 16;
            acc[10] = (unsigned long long)b_item; acc[11] = (unsigned long long)b_first; acc[12] = (unsigned long long)b_full;
            acc[13] = (unsigned long long)b_setup; acc[14] = (unsigned long long)b_march;
        }
        // ---- error maxima: warp -> CTA
        const double wa = warp_max(ea.mabs), wr = warp_max(ea.mrel);
        if (lane == 0) {
            s_red[warp][0] = wa;
            s_red[warp][1] = wr;
        }
    }
    __syncthreads();  // also orders every thread's global / peer stores before thread 0's fence
    if (MULTI) {
        // the queue counter restarts when the last CTA leaves (everything else was published per item)
        if (tid == 0) {
            __threadfence();
            if (atomicAdd(&p.scratch[0], 1u) == gridDim.x - 1) {
                p.scratch[0] = 0;
                p.scratch[2] = 0;
            }
        }
        return;
    }
    if (tid == 0) {
        if (dbg) g_dbg_ts[blockIdx.x * 8 + 4] = globaltimer_ns();
        double a = 0.0, r = 0.0;
        for (int w = 0; w < ncw; w++) {
            a = fmax(a, s_red[w][0]);
            r = fmax(r, s_red[w][1]);
        }
        unsigned long long* e = reinterpret_cast<unsigned long long*>(p.err + 2 * (size_t)p.n0);
        if (a > 0.0) atomicMax(e, (unsigned long long)__double_as_longlong(a));
        if (r > 0.0) atomicMax(e + 1, (unsigned long long)__double_as_longlong(r));
        // Order this CTA's stores (ghost cells in peer memory included: a store is performed at the
        // home L2 of its address) before the ticket. The gpu-scope fence is enough here: the last CTA
        // observes every ticket and its system-scope fence + release store make, by cumulativity, all
        // of them visible to the neighbour that acquires the flag. A system-scope fence in every CTA
        // costs ~30 us per CTA on B200 (measured) and is kept only as a debugging mode.
        if (p.fence_mode == 2) __threadfence_system();
        else __threadfence();
        if (dbg) g_dbg_ts[blockIdx.x * 8 + 5] = globaltimer_ns();
        const uint32_t ticket = atomicAdd(&p.scratch[0], 1u);
        if (ticket == gridDim.x - 1) {
            // last CTA of the layer: every face store of this GPU is ordered before this point
            if (p.fence_mode == 0) __threadfence();
            else __threadfence_system();
            p.scratch[0] = 0;
            p.scratch[2] = 0;  // every CTA has drawn its last (empty) item: the queue can restart
            if (p.root_err) {
                const unsigned long long va = atomicMax(e, 0ull), vr = atomicMax(e + 1, 0ull);
                unsigned long long* re = reinterpret_cast<unsigned long long*>(p.root_err + 2 * (size_t)p.n0);
                atomicMax_system(re, va);
                atomicMax_system(re + 1, vr);
            }
            for (int f = 0; f < 6; f++)
                if (p.flag_out[f]) st_release_sys(p.flag_out[f], (uint32_t)p.n0 + 1u);
        }
    }
}

// ------------------------------------------------------------- small kernels

struct TabParams {
    double pi, hx, hy, hz, Lx, Ly, Lz;
    int n[3], o[3];
};

// Sine tables with the reference's argument expressions and libdevice sin
// (ref: cuda_sol_kernels.cu:41,184) plus reciprocals for the rel-error filter.
__global__ void k_tables(TabParams tp, double* sx, double* sy, double* sz, double* rsx, double* rsy, double* rsz) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int axis = blockIdx.y;
    if (t >= tp.n[axis]) return;
    const double g = (double)(t + tp.o[axis]);
    double arg;
    if (axis == 0) arg = __ddiv_rn(__dmul_rn(__dmul_rn(__dadd_rn(tp.pi, tp.pi), tp.hx), g), tp.Lx);
    else if (axis == 1) arg = __ddiv_rn(__dmul_rn(__dmul_rn(tp.pi, tp.hy), g), tp.Ly);
    else arg = __ddiv_rn(__dmul_rn(__dmul_rn(tp.pi, tp.hz), g), tp.Lz);
    const double s = sin(arg);
    const double r = __ddiv_rn(1.0, fabs(s));
    if (axis == 0) { sx[t] = s; rsx[t] = r; }
    else if (axis == 1) { sy[t] = s; rsy[t] = r; }
    else { sz[t] = s; rsz[t] = r; }
}

__global__ void k_layer0(double* u, const double* sx, const double* sy, const double* sz, double arg_t, int X, int Y,
                         int Z, int64_t ps, int64_t pitch, int oz) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    if (k >= Z || j >= Y) return;
    const double c = cos(arg_t);
    const double syj = sy[j], szk = sz[k];
    for (int i = blockIdx.z; i < X; i += gridDim.z) {
        const double f = __dmul_rn(__dmul_rn(__dmul_rn(c, sx[i]), syj), szk);
        u[(int64_t)(i + 1) * ps + (int64_t)(j + 1) * pitch + (oz + k)] = f;
    }
}

__global__ void k_zero_walls(double* u, int X, int Y, int Z, int64_t ps, int64_t pitch, int oz, int mask) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    double* pl = u + (int64_t)(i + 1) * ps;
    if (t < Z) {
        if (mask & 1) pl[(int64_t)1 * pitch + oz + t] = 0.0;
        if (mask & 2) pl[(int64_t)Y * pitch + oz + t] = 0.0;
    }
    if (t < Y) {
        if (mask & 4) pl[(int64_t)(t + 1) * pitch + oz] = 0.0;
        if (mask & 8) pl[(int64_t)(t + 1) * pitch + oz + Z - 1] = 0.0;
    }
}

// face <-> packed buffer / peer ghost copy. mode 0: pack, 1: unpack, 2: push to peer.
struct FaceParams {
    int X, Y, Z, oz;
    int64_t ps, pitch;
    int send_x0, send_x1;
};
__device__ __forceinline__ int64_t sidx(const FaceParams& g, int i, int j, int k) {
    return (int64_t)(i + 1) * g.ps + (int64_t)(j + 1) * g.pitch + (g.oz + k);
}
__global__ void k_face(FaceParams g, double* u, int face, int mode, double* buf, int64_t sa, int64_t sb) {
    const int axis = face >> 1, side = face & 1;
    const int na = axis == 0 ? g.Y : g.X;
    const int nb = axis == 2 ? g.Y : g.Z;
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    const int a = blockIdx.y * blockDim.y + threadIdx.y;
    if (a >= na || b >= nb) return;
    int i, j, k;
    if (axis == 0) { j = a; k = b; }
    else if (axis == 1) { i = a; k = b; }
    else { i = a; j = b; }
    if (mode == 1) {  // unpack into ghost layer
        if (axis == 0) i = side ? g.X : -1;
        else if (axis == 1) j = side ? g.Y : -1;
        else k = side ? g.Z : -1;
        u[sidx(g, i, j, k)] = buf[(int64_t)a * nb + b];
    } else {
        if (axis == 0) i = side ? g.send_x1 : g.send_x0;
        else if (axis == 1) j = side ? g.Y - 1 : 0;
        else k = side ? g.Z - 1 : 0;
        const double v = u[sidx(g, i, j, k)];
        if (mode == 0) buf[(int64_t)a * nb + b] = v;
        else buf[a * sa + b * sb] = v;
    }
}

__global__ void k_publish(wave3d_peers peers, uint32_t epoch) {
    if (threadIdx.x == 0) {
        __threadfence_system();
        for (int f = 0; f < 6; f++)
            if (peers.flag_out[f]) st_release_sys(peers.flag_out[f], epoch);
    }
}

// Stand-alone analytic error of a stored layer (two-stage max; ref K7).
struct FErrParams {
    int lo[3], hi[3];
    int oz;
    int64_t ps, pitch;
    double arg_t;
};
__global__ void __launch_bounds__(256) k_field_error(FErrParams g, const double* u, const double* sx, const double* sy,
                                                      const double* sz, unsigned long long* scratch, double* out) {
    __shared__ double s_a[8], s_r[8];
    __shared__ bool s_last;
    const double c = cos(g.arg_t);
    const int nk = g.hi[2] - g.lo[2] + 1, nj = g.hi[1] - g.lo[1] + 1, ni = g.hi[0] - g.lo[0] + 1;
    const int64_t rows = (int64_t)ni * nj;
    double ma = 0.0, mr = 0.0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // one warp per row, lanes stride along z
    for (int64_t row = (int64_t)blockIdx.x * 8 + warp; row < rows; row += (int64_t)gridDim.x * 8) {
        const int i = g.lo[0] + (int)(row / nj), j = g.lo[1] + (int)(row % nj);
        const double csy = __dmul_rn(__dmul_rn(c, sx[i]), sy[j]);
        const double* r = u + (int64_t)(i + 1) * g.ps + (int64_t)(j + 1) * g.pitch + g.oz;
        for (int k = g.lo[2] + lane; k < g.lo[2] + nk; k += 32) {
            const double f = __dmul_rn(csy, sz[k]);
            const double a = fabs(__dsub_rn(r[k], f));
            ma = fmax(ma, a);
            mr = fmax(mr, __ddiv_rn(a, fabs(f)));
        }
    }
    ma = warp_max(ma);
    mr = warp_max(mr);
    if (lane == 0) { s_a[warp] = ma; s_r[warp] = mr; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; w++) { ma = fmax(ma, s_a[w]); mr = fmax(mr, s_r[w]); }
        scratch[2 + 2 * blockIdx.x] = (unsigned long long)__double_as_longlong(ma);
        scratch[3 + 2 * blockIdx.x] = (unsigned long long)__double_as_longlong(mr);
        __threadfence();
        const unsigned long long ticket = atomicAdd(&scratch[0], 1ull);
        s_last = ticket == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
        __threadfence();
        double a = 0.0, r = 0.0;
        for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) {
            a = fmax(a, __longlong_as_double((long long)((volatile unsigned long long*)scratch)[2 + 2 * b]));
            r = fmax(r, __longlong_as_double((long long)((volatile unsigned long long*)scratch)[3 + 2 * b]));
        }
        a = warp_max(a);
        r = warp_max(r);
        __syncthreads();
        if (lane == 0) { s_a[warp] = a; s_r[warp] = r; }
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int w = 1; w < 8; w++) { a = fmax(a, s_a[w]); r = fmax(r, s_r[w]); }
            out[0] = a;
            out[1] = r;
            scratch[0] = 0;  // ready for the next call
        }
    }
}

// ------------------------------------------------------------------ host side

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(f);
    });
    return fn;
}

int encode_field_map(CUtensorMap* map, const double* base, const wave3d_layout* lay, int box_w, int box_h) {
    EncodeTiledFn enc = get_encode();
    if (!enc) {
        set_last_error("cuTensorMapEncodeTiled entry point not available");
        return -1;
    }
    cuuint64_t dims[3] = {(cuuint64_t)lay->pitch, (cuuint64_t)lay->rows, (cuuint64_t)lay->planes};
    cuuint64_t strides[2] = {(cuuint64_t)lay->pitch * 8, (cuuint64_t)lay->plane_stride * 8};
    cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_last_error("cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
        return -1;
    }
    return 0;
}

static bool is_pow2(double d) {
    int e;
    return std::frexp(d, &e) == 0.5;
}
static bool mant_all_ones(double d) {
    uint64_t b;
    memcpy(&b, &d, 8);
    return (b & 0xFFFFFFFFFFFFFull) == 0xFFFFFFFFFFFFFull;
}

int choose_div_mode(const wave3d_problem* p) {
    if (const char* e = getenv("WAVE3D_DIV")) {
        int m = atoi(e);
        if (m >= 0 && m <= 3) return m;
    }
    const double d[3] = {p->hx * p->hx, p->hy * p->hy, p->hz * p->hz};
    if (is_pow2(d[0]) && is_pow2(d[1]) && is_pow2(d[2])) return 0;
    for (double v : d)
        if (mant_all_ones(v) || !(v > 1e-280 && v < 1e280)) return 3;
    return 2;
}

int plan_smem_bytes(int ty, int tz, int s1, int s0) {
    const int slot1 = ((ty + 2) * (tz + 4) * 8 + 127) & ~127;
    const int slot0 = (ty * tz * 8 + 127) & ~127;
    return 128 + s1 * slot1 + s0 * slot0 + 16 * (s1 + s0) + 64;
}

typedef void (*StepKernel)(const StepParams);

// mode: 0 = one layer (n >= 2), 1 = first layer (n == 1), 2 = multi-layer persistent launch
template <int RJ>
static StepKernel pick_rj(int mode, int div) {
    switch (mode * 4 + div) {
        case 0: return k_step<RJ, false, 0, false>;
        case 1: return k_step<RJ, false, 1, false>;
        case 2: return k_step<RJ, false, 2, false>;
        case 3: return k_step<RJ, false, 3, false>;
        case 4: return k_step<RJ, true, 0, false>;
        case 5: return k_step<RJ, true, 1, false>;
        case 6: return k_step<RJ, true, 2, false>;
        case 7: return k_step<RJ, true, 3, false>;
        case 8: return k_step<RJ, false, 0, true>;
        case 9: return k_step<RJ, false, 1, true>;
        case 10: return k_step<RJ, false, 2, true>;
        default: return k_step<RJ, false, 3, true>;
    }
}
static StepKernel pick_kernel(int rj, int mode, int div) {
    if (rj == 1) return pick_rj<1>(mode, div);
    if (rj == 2) return pick_rj<2>(mode, div);
    return pick_rj<4>(mode, div);
}

// cudaFuncSetAttribute is per device; remember what was granted so that launches
// inside a stream capture never have to call it.
static int ensure_smem_attr(const void* k, int smem) {
    static std::mutex mu;
    static std::map<std::pair<int, const void*>, int> granted;
    int dev = 0;
    W3D_TRY(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(mu);
    int& g = granted[{dev, k}];
    if (smem > g) {
        W3D_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        g = smem;
    }
    return 0;
}

int prepare_step_kernels(const wave3d_plan* pl, int div) {
    const int smem = plan_smem_bytes(pl->ty, pl->tz, pl->s1, pl->s0);
    if (ensure_smem_attr((const void*)pick_kernel(pl->rj, 0, div), smem)) return -1;
    if (ensure_smem_attr((const void*)pick_kernel(pl->rj, 2, div), smem)) return -1;
    return ensure_smem_attr((const void*)pick_kernel(pl->rj, 1, div), plan_smem_bytes(pl->ty, pl->tz, pl->s1, 0));
}

int step_occupancy(int rj, int warps, int smem_bytes) {
    StepKernel k = pick_kernel(rj, 2, 0);  // the multi-layer variant needs the most registers
    if (cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, (const void*)k, (warps + 1) * 32, smem_bytes) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return nb;
}

int launch_step(const StepLaunch& a) {
    const wave3d_problem* pr = a.prob;
    const wave3d_box* b = a.box;
    const wave3d_layout* l = a.lay;
    const wave3d_plan* pl = a.plan;
    if (pl->rj != 1 && pl->rj != 2 && pl->rj != 4) {
        set_last_error("plan.rj must be 1, 2 or 4");
        return -1;
    }
    if (pl->warps < 1 || pl->warps > kMaxWarps || pl->ty != pl->rj * pl->warps || pl->tz != 64 || pl->s1 < 3 ||
        pl->s0 < 1 || pl->ctas < 1 || pl->nchunks < 1 || pl->nchunks > 16 || pl->xcut[0] != 0 || pl->xcut[pl->nchunks] != b->n[0]) {
        set_last_error("invalid tile plan");
        return -1;
    }
    {
        int seen = 0;
        for (int c = 0; c < pl->nchunks; c++)
            if (pl->order[c] >= 0 && pl->order[c] < pl->nchunks) seen |= 1 << pl->order[c];
        if (seen != (1 << pl->nchunks) - 1) {
            set_last_error("tile plan: order[] is not a permutation of the chunk positions");
            return -1;
        }
    }
    const bool multi = a.nsteps > 1;
    if (a.nsteps == 1 && a.desc_tab) {
        set_last_error("the streamed descriptor table needs a multi-layer launch");
        return -1;
    }
    if (multi && (a.first || a.n < 2 || !a.epoch || !a.stepdone || a.desc)) {
        set_last_error("multi-layer launch needs n >= 2, epoch/stepdone arrays and no step descriptor");
        return -1;
    }
    StepParams p;
    memset(&p, 0, sizeof p);
    for (int w = 0; w < 2; w++) {
        p.map_halo[w] = a.map_halo[w];
        p.map_core[w] = a.map_core[w];
        p.field[w] = a.field[w];
        for (int f = 0; f < 6; f++) p.dst[w][f] = a.peers[w].dst[f];
    }
    p.sx = a.tabs[0]; p.sy = a.tabs[1]; p.sz = a.tabs[2];
    p.rsx = a.tabs[3]; p.rsy = a.tabs[4]; p.rsz = a.tabs[5];
    const wave3d_peers& pw = a.peers[a.n & 1];
    for (int f = 0; f < 6; f++) {
        p.sa[f] = pw.sa[f];
        p.sb[f] = pw.sb[f];
        p.flag_out[f] = pw.flag_out[f];
        p.flag_in[f] = pw.flag_in[f];
        if (pw.dst[f]) p.push_mask |= 1u << f;
    }
    p.err = a.err;
    p.root_err = pw.root_err;
    p.epoch = a.epoch;
    p.stepdone = a.stepdone;
    p.facedone = a.stepdone ? a.stepdone + (pr->timesteps + 2) : nullptr;
    p.desc_tab = multi ? a.desc_tab : nullptr;
    p.host_res = multi ? a.host_res : nullptr;
    p.host_flag = multi ? a.host_flag : nullptr;
    p.n0 = a.n;
    p.nsteps = multi ? a.nsteps : 1;
    p.a_t = pr->a_t;
    p.tau = pr->tau;
    p.twopi = pr->pi + pr->pi;
    p.scratch = a.scratch;
    p.desc = a.desc;
    p.ps = l->plane_stride;
    p.pitch = l->pitch;
    p.coef = (pr->a2 * pr->tau) * pr->tau;  // ((a2*tau)*tau), *0.5 for the first step
    if (a.first) p.coef = p.coef * 0.5;
    p.arg_t = std::fma(pr->a_t * (double)a.n, pr->tau, pr->pi + pr->pi);
    p.hx2 = pr->hx * pr->hx; p.hy2 = pr->hy * pr->hy; p.hz2 = pr->hz * pr->hz;
    p.rx = 1.0 / p.hx2; p.ry = 1.0 / p.hy2; p.rz = 1.0 / p.hz2;
    const double to = a.spin_timeout_s > 0 ? a.spin_timeout_s : 10.0;
    p.timeout_ns = (unsigned long long)(to * 1e9);
    p.oz = l->oz; p.X = b->n[0]; p.Y = b->n[1]; p.Z = b->n[2];
    p.lo1 = b->lo[1]; p.hi1 = b->hi[1]; p.lo2 = b->lo[2]; p.hi2 = b->hi[2];
    p.elo0 = b->elo[0]; p.ehi0 = b->ehi[0];
    p.nty = pl->nty; p.ntz = pl->ntz; p.nchunks = pl->nchunks; p.ty = pl->ty; p.tz = pl->tz;
    for (int c = 0; c <= pl->nchunks; c++) p.xcut[c] = pl->xcut[c];
    for (int c = 0; c < pl->nchunks; c++) {
        p.pos_of[c] = (signed char)pl->order[c];
        p.pi_of[pl->order[c]] = (signed char)c;
    }
    p.self_wrap = b->nbr[0] == b->rank;
    if (multi && p.self_wrap) {
        // the wrap-around ghost planes are guarded by the per-item epochs of the end chunks, not by a
        // whole-layer flag (which would turn every layer boundary back into a barrier)
        p.flag_in[0] = p.flag_in[1] = nullptr;
    }
    p.s1 = pl->s1; p.s0 = pl->s0;
    p.send_x0 = b->send_x[0]; p.send_x1 = b->send_x[1];
    {
        bool peers = p.push_mask != 0 || p.root_err != nullptr;
        for (int f = 0; f < 6; f++) peers = peers || p.flag_out[f] != nullptr;
        p.fence_mode = peers ? 1 : 0;
        if (const char* e = getenv("WAVE3D_FENCE")) p.fence_mode = atoi(e);
        if (const char* e = getenv("WAVE3D_DBG_TS")) p.dbg = atoi(e);
    }
    const int div = a.div_mode >= 0 ? a.div_mode : choose_div_mode(pr);
    StepKernel k = pick_kernel(pl->rj, multi ? 2 : (a.first ? 1 : 0), div);
    const int smem = plan_smem_bytes(pl->ty, pl->tz, pl->s1, a.first ? 0 : pl->s0);
    if (ensure_smem_attr((const void*)k, smem)) return -1;
    const int grid = pl->ctas;  // persistent CTAs pulling (tile, x-chunk) items from the device-side queue
    const int block = (pl->warps + 1) * 32;  // + the TMA producer warp
    if (multi) {
        // items wait on each other inside the kernel: every CTA must be resident
        void* args[1] = {&p};
        W3D_TRY(cudaLaunchCooperativeKernel((const void*)k, dim3(grid), dim3(block), args, (size_t)smem, a.stream));
    } else {
        k<<<grid, block, smem, a.stream>>>(p);
    }
    W3D_TRY(cudaGetLastError());
    return 0;
}

}  // namespace w3d

// ----------------------------------------------------------------- C entry points
using namespace w3d;

extern "C" {

const char* wave3d_cuda_last_error(void) { return g_last_error.c_str(); }

int wave3d_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}
int wave3d_cuda_available(void) { return wave3d_cuda_device_count() > 0; }

int wave3d_cuda_tables(const wave3d_problem* p, const wave3d_box* b, double* sx, double* sy, double* sz, double* rsx,
                       double* rsy, double* rsz, void* stream) {
    TabParams tp;
    tp.pi = p->pi; tp.hx = p->hx; tp.hy = p->hy; tp.hz = p->hz; tp.Lx = p->Lx; tp.Ly = p->Ly; tp.Lz = p->Lz;
    int nmax = 0;
    for (int a = 0; a < 3; a++) {
        tp.n[a] = b->n[a];
        tp.o[a] = b->o[a];
        nmax = nmax > b->n[a] ? nmax : b->n[a];
    }
    dim3 grid((nmax + 127) / 128, 3);
    k_tables<<<grid, 128, 0, (cudaStream_t)stream>>>(tp, sx, sy, sz, rsx, rsy, rsz);
    W3D_TRY(cudaGetLastError());
    return 0;
}

int wave3d_cuda_layer0(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l, const double* sx,
                       const double* sy, const double* sz, double* u0, void* stream) {
    dim3 block(64, 4);
    dim3 grid((b->n[2] + 63) / 64, (b->n[1] + 3) / 4, b->n[0] < 64 ? b->n[0] : 64);
    const double arg_t = std::fma(p->a_t * 0.0, p->tau, p->pi + p->pi);
    k_layer0<<<grid, block, 0, (cudaStream_t)stream>>>(u0, sx, sy, sz, arg_t, b->n[0], b->n[1], b->n[2],
                                                        l->plane_stride, l->pitch, l->oz);
    W3D_TRY(cudaGetLastError());
    return 0;
}

int wave3d_cuda_zero_walls(const wave3d_box* b, const wave3d_layout* l, double* u, void* stream) {
    int mask = 0;
    if (b->coords[1] == 0) mask |= 1;
    if (b->coords[1] == b->dims[1] - 1) mask |= 2;
    if (b->coords[2] == 0) mask |= 4;
    if (b->coords[2] == b->dims[2] - 1) mask |= 8;
    if (!mask) return 0;
    const int m = b->n[1] > b->n[2] ? b->n[1] : b->n[2];
    dim3 grid((m + 127) / 128, b->n[0]);
    k_zero_walls<<<grid, 128, 0, (cudaStream_t)stream>>>(u, b->n[0], b->n[1], b->n[2], l->plane_stride, l->pitch, l->oz,
                                                          mask);
    W3D_TRY(cudaGetLastError());
    return 0;
}

static int face_launch(const wave3d_box* b, const wave3d_layout* l, double* u, int face, int mode, double* buf,
                       int64_t sa, int64_t sb, cudaStream_t stream) {
    FaceParams g;
    g.X = b->n[0]; g.Y = b->n[1]; g.Z = b->n[2]; g.oz = l->oz;
    g.ps = l->plane_stride; g.pitch = l->pitch;
    g.send_x0 = b->send_x[0]; g.send_x1 = b->send_x[1];
    const int axis = face >> 1;
    const int na = axis == 0 ? g.Y : g.X;
    const int nb = axis == 2 ? g.Y : g.Z;
    dim3 block(64, 4);
    dim3 grid((nb + 63) / 64, (na + 3) / 4);
    k_face<<<grid, block, 0, stream>>>(g, u, face, mode, buf, sa, sb);
    W3D_TRY(cudaGetLastError());
    return 0;
}

int wave3d_cuda_pack_face(const wave3d_box* b, const wave3d_layout* l, const double* u, int face, double* buf,
                          void* stream) {
    return face_launch(b, l, const_cast<double*>(u), face, 0, buf, 0, 0, (cudaStream_t)stream);
}
int wave3d_cuda_unpack_face(const wave3d_box* b, const wave3d_layout* l, double* u, int face, const double* buf,
                            void* stream) {
    return face_launch(b, l, u, face, 1, const_cast<double*>(buf), 0, 0, (cudaStream_t)stream);
}
int wave3d_cuda_push_faces(const wave3d_box* b, const wave3d_layout* l, const double* u, const wave3d_peers* peers,
                           uint32_t epoch, void* stream) {
    for (int f = 0; f < 6; f++) {
        if (!peers->dst[f]) continue;
        if (face_launch(b, l, const_cast<double*>(u), f, 2, peers->dst[f], peers->sa[f], peers->sb[f],
                        (cudaStream_t)stream))
            return -1;
    }
    bool any = false;
    for (int f = 0; f < 6; f++) any = any || peers->flag_out[f];
    if (any) {
        k_publish<<<1, 32, 0, (cudaStream_t)stream>>>(*peers, epoch);
        W3D_TRY(cudaGetLastError());
    }
    return 0;
}

int64_t wave3d_cuda_field_error_scratch_words(const wave3d_box* b) {
    (void)b;
    return 2 + 2 * 1184;  // ticket + pad, then (abs, rel) per block; grid is capped at 8 CTAs x 148 SMs
}

int wave3d_cuda_field_error(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l,
                            const double* const tabs[6], const double* u, int n, uint64_t* scratch, double* out,
                            void* stream) {
    FErrParams g;
    for (int a = 0; a < 3; a++) {
        g.lo[a] = b->elo[a];
        g.hi[a] = b->ehi[a];
    }
    g.oz = l->oz; g.ps = l->plane_stride; g.pitch = l->pitch;
    g.arg_t = std::fma(p->a_t * (double)n, p->tau, p->pi + p->pi);
    const int64_t rows = (int64_t)(g.hi[0] - g.lo[0] + 1) * (g.hi[1] - g.lo[1] + 1);
    int64_t blocks = (rows + 7) / 8;
    if (blocks > 1184) blocks = 1184;
    if (blocks < 1) blocks = 1;
    k_field_error<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(g, u, tabs[0], tabs[1], tabs[2],
                                                                 reinterpret_cast<unsigned long long*>(scratch), out);
    W3D_TRY(cudaGetLastError());
    return 0;
}

int wave3d_cuda_step(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l, const wave3d_plan* plan,
                     const double* const tabs[6], const double* u1, double* u_io, int n, int first,
                     const wave3d_peers* peers, uint64_t* err, uint32_t* scratch, double spin_timeout_s,
                     void* stream) {
    StepLaunch a;
    memset(&a, 0, sizeof a);
    const int w = n & 1;  // the layer being written plays the role of field n & 1
    if (encode_field_map(&a.map_halo[w ^ 1], u1, l, plan->tz + 4, plan->ty + 2)) return -1;
    if (encode_field_map(&a.map_core[w], u_io, l, plan->tz, plan->ty)) return -1;
    a.field[w] = u_io;
    a.field[w ^ 1] = const_cast<double*>(u1);
    a.prob = p; a.box = b; a.lay = l; a.plan = plan;
    for (int i = 0; i < 6; i++) a.tabs[i] = tabs[i];
    a.n = n; a.nsteps = 1; a.first = first;
    if (peers) a.peers[w] = *peers;
    a.err = err; a.scratch = scratch; a.spin_timeout_s = spin_timeout_s;
    a.div_mode = -1;
    a.stream = (cudaStream_t)stream;
    return launch_step(a);
}

}  // extern "C"

namespace w3d {

// ---------------------------------------------------------------------------------------------
// Allocation-free flavour of the stand-alone error evaluation.
//
// wave3d_cuda_field_error() needs a few KB of scratch that its caller has to cudaMalloc (or get
// from torch). Some users of this library are not in a position to allocate device memory at the
// moment they want an error figure: an in-situ monitor that is attached to a production run whose
// job script has already handed every free byte of HBM to the solver or to a pooling allocator
// (a cudaMalloc there can simply fail); a real-time / latency-critical loop that must not take
// the device-wide synchronisation cudaMalloc and cudaFree imply; a plug-in that runs inside a
// host callback or a foreign framework's "no allocation" region. For them the scratch lives in
// the library image itself: one file-scope __device__ array per loaded module, i.e. every GPU
// that loads libwave3d keeps these ~19 KB until the process exits, whether or not this entry
// point is ever used.
//
// The array is shared by all calls on a device and still holds whatever the previous call left
// there, so every call first clears the one word the kernel needs at zero (the CTA ticket) in
// stream order. Two calls whose kernels overlap on the same GPU (different streams, or threads
// racing) would share ticket and partial maxima and corrupt each other — callers serialise them.
//
// A host function cannot use the array's name as a pointer: on the host that name is only a
// handle for the runtime. The device address is looked up with cudaGetSymbolAddress, once per
// device, by the first call made while that device is current; that first call may load the
// module and should therefore be made before any stream capture is started on the device.
constexpr int kStaticErrMaxBlocks = 1184;                      // 8 resident CTAs x 148 SMs
constexpr int kStaticErrWords = 2 + 2 * kStaticErrMaxBlocks;   // ticket + pad, (abs, rel) per CTA
static __device__ __align__(256) unsigned long long g_static_err_scratch[kStaticErrWords];

static int static_err_scratch_ptr(unsigned long long** out) {
    static std::mutex mu;
    static unsigned long long* per_device[64] = {nullptr};
    int dev = -1;
    W3D_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        set_last_error("wave3d_cuda_field_error_static: device index " + std::to_string(dev) + " not supported (max 63)");
        return -1;
    }
    std::lock_guard<std::mutex> lk(mu);
    if (!per_device[dev]) {
        unsigned long long* ptr = nullptr;
        W3D_TRY(cudaGetSymbolAddress((void**)&ptr, g_static_err_scratch));
        if (!ptr) {
            set_last_error("wave3d_cuda_field_error_static: cudaGetSymbolAddress returned a null address");
            return -1;
        }
        per_device[dev] = ptr;
    }
    *out = per_device[dev];
    return 0;
}

}  // namespace w3d

extern "C" int wave3d_debug_copy_timestamps(unsigned long long* host, int nctas) {
    using namespace w3d;
    if (nctas > 2048) nctas = 2048;
    W3D_TRY(cudaDeviceSynchronize());
    W3D_TRY(cudaMemcpyFromSymbol(host, g_dbg_ts, sizeof(unsigned long long) * 8 * (size_t)nctas));
    return 0;
}

extern "C" int wave3d_debug_copy_phases(unsigned long long* host, int nctas, int device) {
    using namespace w3d;
    if (nctas > 2048) nctas = 2048;
    if (device >= 0) W3D_TRY(cudaSetDevice(device));
    W3D_TRY(cudaDeviceSynchronize());
    W3D_TRY(cudaMemcpyFromSymbol(host, g_dbg_acc, sizeof(unsigned long long) * 16 * (size_t)nctas));
    return 0;
}

extern "C" int wave3d_cuda_field_error_static(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l,
                                              const double* const tabs[6], const double* u, int n, int blocks,
                                              double* out, void* stream) {
    using namespace w3d;
    FErrParams g;
    for (int a = 0; a < 3; a++) {
        g.lo[a] = b->elo[a];
        g.hi[a] = b->ehi[a];
    }
    g.oz = l->oz; g.ps = l->plane_stride; g.pitch = l->pitch;
    g.arg_t = std::fma(p->a_t * (double)n, p->tau, p->pi + p->pi);
    const int64_t rows = (int64_t)(g.hi[0] - g.lo[0] + 1) * (g.hi[1] - g.lo[1] + 1);
    int64_t nb = blocks > 0 ? blocks : (rows + 7) / 8;   // same rule as wave3d_cuda_field_error ...
    if (blocks <= 0 && nb > kStaticErrMaxBlocks) nb = kStaticErrMaxBlocks;  // ... including its cap
    if (nb < 1) nb = 1;
    if (nb > kStaticErrMaxBlocks) {
        set_last_error("wave3d_cuda_field_error_static: " + std::to_string((long long)nb) +
                       " CTAs requested but the built-in scratch holds partial maxima for at most " +
                       std::to_string(kStaticErrMaxBlocks) + " (" + std::to_string(kStaticErrWords * 8) +
                       " bytes); use fewer blocks or wave3d_cuda_field_error with caller-allocated scratch");
        return -1;
    }
    unsigned long long* scratch = nullptr;
    if (static_err_scratch_ptr(&scratch)) return -1;
    // the previous call's partial maxima may stay (every CTA overwrites its own slot before the
    // last CTA reads them); the ticket must start at zero
    W3D_TRY(cudaMemsetAsync(scratch, 0, 16, (cudaStream_t)stream));
    k_field_error<<<(int)nb, 256, 0, (cudaStream_t)stream>>>(g, u, tabs[0], tabs[1], tabs[2], scratch, out);
    W3D_TRY(cudaGetLastError());
    return 0;
}
