// wave3d_kernels.cu — sm_100a device code of the wave3d solver.
//
// Hot kernel: k_step<RJ,FIRST,DIV,MULTI> (wave3d_step_kernel.inc) — one launch per time layer, or one
// persistent launch for all layers of a small box, that fuses what the reference spreads over K2/K3
// (update), K4 (boundary planes), K5/K6 + 12 memcpys + MPI p2p (halo exchange), K7 + host loop (error max)
// and two memsets (ref: cuda_sol.cpp:384-421, cuda_sol_kernels.cu:24-268).
//
// Design (B200-first, not a translation):
//   * The stencil source layer u^{n-1} and the in-place layer u^{n-2} stream through shared memory as
//     y-z tiles, one x plane at a time, moved by TMA bulk tensor loads (cp.async.bulk.tensor.3d ->
//     UTMALDG) into mbarrier-guarded rings by a scheduler warp that works one item ahead; the compute
//     warps never issue a global or constant-bank load on the hot path (tables, layer constants and
//     tile geometry reach them through shared memory).
//   * Each thread owns 2 z-cells x RJ rows, reads its 7-point neighbourhood from the three resident
//     planes and writes u^n with one 16-byte store over u^{n-2}.
//   * The analytic solution is separable: f = ((c*sx[i])*sy[j])*sz[k] with 1-D sine tables ->
//     bit-identical to the reference's per-cell trig at 3 multiplies. max|u-f| is an in-register max;
//     max|u-f|/|f| uses a monotone filter with a CTA-shared threshold so the exact IEEE division runs
//     only for candidates of the maximum.
//   * Face cells are stored a second time, straight into the neighbour GPU's ghost cells (peer-mapped
//     memory over NVLink; own memory for the periodic self-wrap). Epoch flags in the receiver's memory
//     (per face, and per x-chunk for y/z faces in the multi-layer launch) are the handshake: published
//     with one system-scope fence by whoever completes a layer / strip, polled by the scheduler warps.
//     No pack/unpack kernels, no host staging, no MPI/NCCL on this path.
//   * Per-layer maxima land in err[2n..2n+1] via atomicMax on the u64 bit pattern (values are >= 0; NaN
//     never wins, like fmax) and are forwarded to rank 0's table with a system-scope atomic (the
//     reference's deferred MPI_Reduce).
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

#include "wave3d_step_kernel.inc"

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

// Layer 0 = the analytic solution on every owned cell. Write-only and bandwidth-bound: each thread owns a
// 16-byte aligned pair of storage columns (one STG.128 per plane where both cells are owned) and walks a strided
// set of planes with the y/z table values in registers.
__global__ void __launch_bounds__(256) k_layer0(double* u, const double* sx, const double* sy, const double* sz,
                                                double arg_t, int X, int Y, int Z, int64_t ps, int64_t pitch, int oz) {
    const int c0 = 2 * (blockIdx.x * blockDim.x + threadIdx.x);  // storage column of my first cell (even)
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k0 = c0 - oz, k1 = k0 + 1;                          // owned z indices of the pair
    const bool v0 = k0 >= 0 && k0 < Z, v1 = k1 >= 0 && k1 < Z;
    if (j >= Y || !(v0 || v1)) return;
    const double c = cos(arg_t);
    const double syj = sy[j];
    const double s0 = v0 ? sz[k0] : 0.0, s1 = v1 ? sz[k1] : 0.0;
    double* row = u + (int64_t)(j + 1) * pitch + c0;
    for (int i = blockIdx.z; i < X; i += gridDim.z) {
        const double cxy = __dmul_rn(__dmul_rn(c, sx[i]), syj);
        double* q = row + (int64_t)(i + 1) * ps;
        if (v0 && v1) *reinterpret_cast<double2*>(q) = make_double2(__dmul_rn(cxy, s0), __dmul_rn(cxy, s1));
        else if (v0) q[0] = __dmul_rn(cxy, s0);
        else q[1] = __dmul_rn(cxy, s1);
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
    // every word of the face's flag block: the per-chunk words and the whole-layer word
    if (threadIdx.x == 0) __threadfence_system();
    __syncwarp();
    for (int f = 0; f < 6; f++)
        if (peers.flag_out[f])
            for (int c = threadIdx.x; c <= kFlagLayer; c += 32) st_relaxed_sys(peers.flag_out[f] + c, epoch);
}

// Stand-alone analytic error of a stored layer (two-stage max; ref K7).
struct FErrParams {
    int lo[3], hi[3];
    int oz;
    int64_t ps, pitch;
    double arg_t;
};
__global__ void __launch_bounds__(256) k_field_error(FErrParams g, const double* u, const double* sx, const double* sy,
                                                      const double* sz, const double* rsx, const double* rsy,
                                                      const double* rsz, unsigned long long* scratch, double* out) {
    __shared__ double s_a[8], s_r[8];
    __shared__ bool s_last;
    const double c = cos(g.arg_t);
    const double rc = __ddiv_rn(1.0, fabs(c));
    const int nk = g.hi[2] - g.lo[2] + 1, nj = g.hi[1] - g.lo[1] + 1, ni = g.hi[0] - g.lo[0] + 1;
    const int64_t rows = (int64_t)ni * nj;
    double ma = 0.0, mr = 0.0, thr = 0.0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // One warp per row, lanes stride along z, four cells per lane in flight. max|u-f|/|f| uses the same monotone
    // filter as the fused kernel: a candidate from the reciprocal tables (~1e-15 relative) and the IEEE division
    // only for cells within 2^-40 of the thread's best — exactly the maximum of the reference's per-cell division.
    for (int64_t row = (int64_t)blockIdx.x * 8 + warp; row < rows; row += (int64_t)gridDim.x * 8) {
        const int i = g.lo[0] + (int)(row / nj), j = g.lo[1] + (int)(row % nj);
        const double csy = __dmul_rn(__dmul_rn(c, sx[i]), sy[j]);
        const double rcsy = __dmul_rn(__dmul_rn(rc, rsx[i]), rsy[j]);
        const double* r = u + (int64_t)(i + 1) * g.ps + (int64_t)(j + 1) * g.pitch + g.oz;
        for (int k = g.lo[2] + lane; k < g.lo[2] + nk; k += 128) {
            double v[4], s[4], q[4];
#pragma unroll
            for (int t = 0; t < 4; t++) {
                const int kk = k + 32 * t;
                const bool in = kk < g.lo[2] + nk;
                v[t] = in ? r[kk] : 0.0;
                s[t] = in ? sz[kk] : 0.0;
                q[t] = in ? rsz[kk] : 0.0;
            }
#pragma unroll
            for (int t = 0; t < 4; t++) {
                if (k + 32 * t >= g.lo[2] + nk) break;
                const double f = __dmul_rn(csy, s[t]);
                const double a = fabs(__dsub_rn(v[t], f));
                ma = fmax(ma, a);
                const double cand = __dmul_rn(__dmul_rn(a, rcsy), q[t]);
                if (cand >= thr) {
                    mr = fmax(mr, __ddiv_rn(a, fabs(f)));
                    thr = fmax(thr, __dmul_rn(cand, 1.0 - 0x1p-40));
                }
            }
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
    return pick_rj<2>(mode, div);
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
    if (pl->rj != 1 && pl->rj != 2) {
        set_last_error("plan.rj must be 1 or 2");
        return -1;
    }
    if (pl->warps < 1 || pl->warps > kMaxWarps || pl->ty != pl->rj * pl->warps || pl->tz != 64 || pl->s1 < 4 || pl->s1 > 16 ||
        pl->s0 < 1 || pl->ctas < 1 || pl->nchunks < 1 || pl->nchunks > kMaxChunks || pl->xcut[0] != 0 ||
        pl->xcut[pl->nchunks] != b->n[0]) {
        set_last_error("invalid tile plan (4 <= s1 <= 16, s0 >= 1, nchunks <= WAVE3D_MAX_CHUNKS)");
        return -1;
    }
    {
        uint64_t seen = 0;
        for (int c = 0; c < pl->nchunks; c++)
            if (pl->order[c] >= 0 && pl->order[c] < pl->nchunks) seen |= 1ull << pl->order[c];
        if (seen != (1ull << pl->nchunks) - 1ull) {
            set_last_error("tile plan: order[] is not a permutation of the chunk positions");
            return -1;
        }
        // The x-face flags and the periodic wrap inside one box are released once the chunks at the two ends
        // are done, so the seam planes have to lie in those chunks; cuts must strictly increase.
        for (int c = 0; c < pl->nchunks; c++)
            if (pl->xcut[c + 1] <= pl->xcut[c]) {
                set_last_error("tile plan: xcut[] must be strictly increasing");
                return -1;
            }
        if (pl->nchunks > 1 && (pl->xcut[1] <= b->send_x[0] || pl->xcut[pl->nchunks - 1] > b->send_x[1])) {
            set_last_error("tile plan: the end chunks must contain the seam planes send_x[0] / send_x[1]");
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
    p.chunkdone = a.chunkdone;
    p.noerr = a.noerr;
    p.yz_chunk_flags = multi && a.chunkdone && a.yz_chunk_flags ? 1 : 0;
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
        // multi-layer launch: the next item's dependencies are checked (and fenced) while the current one is
        // still loading; single-layer launch: only the queue atomic and the table slices have to be hidden,
        // and drawing late keeps the guided hand-out balanced at the end of the launch
        p.ahead = multi ? 6 : 3;
        if (const char* e = getenv("WAVE3D_AHEAD")) p.ahead = atoi(e);
        if (const char* e = getenv("WAVE3D_SCHED")) p.sched_flags = atoi(e);
    }
    const int div = a.div_mode >= 0 ? a.div_mode : choose_div_mode(pr);
    StepKernel k = pick_kernel(pl->rj, multi ? 2 : (a.first ? 1 : 0), div);
    const int smem = plan_smem_bytes(pl->ty, pl->tz, pl->s1, a.first ? 0 : pl->s0);
    if (ensure_smem_attr((const void*)k, smem)) return -1;
    const int grid = pl->ctas;  // persistent CTAs pulling (tile, x-chunk) items from the device-side queue
    const int block = (pl->warps + 1) * 32;  // + the scheduler warp (TMA loads, queue, dependencies, flags)
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
    // pairs of storage columns from column (oz & ~1) on; the pitch is a multiple of 16 doubles, so pairs are 16-byte aligned
    dim3 block(64, 4);
    const int pairs = (l->oz + b->n[2] + 1) / 2;
    dim3 grid((pairs + 63) / 64, (b->n[1] + 3) / 4, b->n[0] < 32 ? b->n[0] : 32);
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
    k_field_error<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(g, u, tabs[0], tabs[1], tabs[2], tabs[3], tabs[4], tabs[5],
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
// handle for the runtime. The device address is looked up with cudaGetSymbolAddress while the
// caller's device is current; the first call on a device may load the module and should therefore
// be made before any stream capture is started on that device.
constexpr int kStaticErrMaxBlocks = 1184;                      // 8 resident CTAs x 148 SMs
constexpr int kStaticErrWords = 2 + 2 * kStaticErrMaxBlocks;   // ticket + pad, (abs, rel) per CTA
static __device__ __align__(256) unsigned long long g_static_err_scratch[kStaticErrWords];

static int static_err_scratch_ptr(unsigned long long** out) {
    // Looked up on every call, with the caller's device current: after the first call on a device this is a
    // table look-up inside the runtime, and the answer stays right when the device was reset (or the module
    // reloaded) in between — a pointer remembered from before a cudaDeviceReset() would dangle.
    unsigned long long* ptr = nullptr;
    W3D_TRY(cudaGetSymbolAddress((void**)&ptr, g_static_err_scratch));
    if (!ptr) {
        set_last_error("wave3d_cuda_field_error_static: cudaGetSymbolAddress returned a null address");
        return -1;
    }
    *out = ptr;
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
    k_field_error<<<(int)nb, 256, 0, (cudaStream_t)stream>>>(g, u, tabs[0], tabs[1], tabs[2], tabs[3], tabs[4], tabs[5],
                                                             scratch, out);
    W3D_TRY(cudaGetLastError());
    return 0;
}
