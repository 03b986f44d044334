/* wave3d.h — C ABI of libwave3d (B200-native 3-D wave-equation solver).
 *
 * One shared library holds everything: problem/CLI spec, block decomposition,
 * report writer, an OpenMP CPU backend (oracle + CPU baseline) and the sm_100a
 * CUDA backend (TMA-fed fused leapfrog + analytic-error + halo-push kernel).
 * The Python package binds this header with ctypes; the `wave3d` CLI links it.
 *
 * Citations `ref:` point into the reference tree (/root/reference) for parity
 * checks; no reference code is reused.
 *
 * Index conventions
 *   owned index   i in [0,X), j in [0,Y), k in [0,Z)  (global = owned + origin)
 *   storage index plane = i + 1, row = j + 1, col = k + oz   (z is contiguous)
 *   ghost layers  plane 0 / X+1, row 0 / Y+1, col oz-1 / oz+Z
 *   faces         0:x-  1:x+  2:y-  3:y+  4:z-  5:z+
 */
#ifndef WAVE3D_H
#define WAVE3D_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WAVE3D_PI_FULL 3.14159265358979323846 /* ref: cuda_sol.cpp:53, cuda_sol_kernels.cu:3 */
#define WAVE3D_PI_TRUNC 3.1415926535          /* ref: mpi_sol.cpp:22, mpi_new.cpp:23, openmp_sol.cpp:20 */

enum { WAVE3D_FACE_XLO = 0, WAVE3D_FACE_XHI, WAVE3D_FACE_YLO, WAVE3D_FACE_YHI, WAVE3D_FACE_ZLO, WAVE3D_FACE_ZHI };

/* ------------------------------------------------------------------ spec -- */

/* Problem definition = the reference CLI `N Np Lx Ly Lz [T] [timesteps]`
 * (ref: cuda_sol.cpp:453-475) plus the derived constants. */
typedef struct wave3d_problem {
    int32_t N;         /* intervals per axis -> (N+1)^3 points */
    int32_t Np;        /* thread / GPU count argument */
    int32_t timesteps; /* default 20 */
    int32_t reserved;
    double Lx, Ly, Lz; /* box lengths; the literal "pi" is accepted */
    double T;          /* default 1 */
    double pi;         /* PI flavour used for every derived constant */
    double a2;         /* 1/(4 pi^2) */
    double a_t;        /* 0.5*sqrt(4/Lx^2 + 1/Ly^2 + 1/Lz^2) */
    double tau;        /* T/timesteps */
    double hx, hy, hz; /* L/N */
    double courant;    /* sqrt(a2)*tau/min(h) -> stdout "C = ..." */
} wave3d_problem;

/* Fill `p` from explicit values and derive the constants. Returns 0, or -1 on
 * invalid input (message in err). */
int wave3d_problem_init(wave3d_problem* p, int N, int Np, double Lx, double Ly, double Lz, double T, int timesteps,
                        double pi, char* err, size_t errlen);

/* Parse the reference's positional arguments (argv[0] is the program name).
 * Unlike the reference (which crashes in std::stoi) a short argv is reported. */
int wave3d_parse_args(int argc, const char* const* argv, double pi, wave3d_problem* p, char* err, size_t errlen);

/* "C = <courant>" formatted like `std::cout << double` (6 significant digits). */
int wave3d_format_courant(const wave3d_problem* p, char* buf, size_t buflen);

/* Format a double like default iostream output (== printf("%g")). */
int wave3d_format_double(double v, char* buf, size_t buflen);

/* ----------------------------------------------------------- decomposition -- */

/* MPI_Dims_create(nranks, 3, {0,0,0})-compatible balanced factorisation,
 * non-increasing (ref: cuda_sol.cpp:478): 8->(2,2,2), 4->(2,2,1), 6->(3,2,1). */
void wave3d_dims_create(int nranks, int dims[3]);

/* Parse "8x1x1" / "2,2,2" into dims; returns 0 on success. */
int wave3d_parse_dims(const char* s, int dims[3]);

/* row-major rank <-> coords, last dim fastest (ref: MPI_Cart_coords, cuda_sol.cpp:483) */
int wave3d_rank_of(const int dims[3], const int coords[3]);
void wave3d_coords_of(const int dims[3], int rank, int coords[3]);

typedef struct wave3d_box {
    int32_t rank, nranks;
    int32_t dims[3];
    int32_t coords[3];
    int32_t n[3];      /* owned points X,Y,Z; remainder goes to the last rank (ref: cuda_sol.cpp:485-489) */
    int32_t o[3];      /* global origin x_0,y_0,z_0 */
    int32_t nbr[6];    /* neighbour rank per face, -1 if none; x is periodic (ref: cuda_sol.cpp:479-496) */
    int32_t lo[3];     /* first/last owned index the stencil updates: all owned x planes (both seam */
    int32_t hi[3];     /*   planes are advanced, ref: mpi_new.cpp:173-174), global 1..N-1 in y and z */
    int32_t elo[3];    /* first/last owned index inside the error set = global interior 1..N-1 */
    int32_t ehi[3];    /*   of every axis (ref: cuda_sol.cpp:388-393) */
    int32_t send_x[2]; /* owned plane sent to x- / x+ neighbour; seam rule: global 1 / N-1 (ref: cuda_sol.cpp:232) */
    int32_t N;
    int32_t reserved;
} wave3d_box;

int wave3d_box_init(int N, const int dims[3], int rank, wave3d_box* box, char* err, size_t errlen);

/* Storage layout of one time layer (padded, z contiguous, rows 128-byte aligned,
 * first updated z cell on a 128-byte boundary so vector and TMA accesses align). */
typedef struct wave3d_layout {
    int64_t pitch;        /* doubles per row (multiple of 16) */
    int64_t rows;         /* Y + 2 */
    int64_t planes;       /* X + 2 */
    int64_t plane_stride; /* rows * pitch */
    int64_t elems;        /* planes * plane_stride */
    int32_t oz;           /* storage column of owned k = 0 */
    int32_t reserved;
} wave3d_layout;

void wave3d_layout_init(const wave3d_box* box, wave3d_layout* lay);

/* ------------------------------------------------------------------ report -- */

typedef struct wave3d_timings {
    double init_ms;     /* allocation + tables (ref: "initialization done in") */
    double solve_ms;    /* layer 0 .. layer timesteps incl. errors + halos (ref: cuda_sol.cpp:382,422-425) */
    double loop_ms;     /* update work: solve - exchange - error */
    double exchange_ms; /* staged transport: time of the exchanges; fused peer path: time CTAs waited for a
                           neighbour rank's flag (mean over the CTAs, max over the ranks) */
    double hostdev_ms;  /* host<->device staging: always 0, nothing is staged through the host */
    double error_ms;    /* cost of the fused error evaluation, measured by wave3d_solver_measure_phases
                           (0 after a plain run) */
} wave3d_timings;

/* Render the rank-0 report exactly in the reference's line format
 * (ref: cuda_sol.cpp:427,435,438-441,572). precision<=0 -> default 6 digits;
 * 17 gives the full-precision side dump. Returns bytes needed. */
size_t wave3d_format_report(const wave3d_problem* p, const wave3d_timings* t, const double* max_abs,
                            const double* max_rel, int precision, int cuda_style, char* buf, size_t buflen);

/* output_N{N}_Np{Np}.txt (ref: README.txt:11, openmp_sol.cpp:229) or, with ng>0,
 * output_N{N}_Np{np}_Ng{ng}_cuda.txt (ref: cuda_sol.cpp:535). */
int wave3d_output_name(int N, int np, int ng, char* buf, size_t buflen);

/* -------------------------------------------------------------- CPU backend -- */
/* Restatement of the reference's `mpi_new` semantics (3 rotating layers become 2
 * in-place layers, error fused in the update) with the CUDA build's exact fp64
 * expression order. Serves as oracle, CPU baseline and as the
 * host twin of every device op. All functions take one rank's box + layout. */

/* sin tables: sx[i]=sin(((2pi*hx)*(i+ox))/Lx), sy[j]=sin(((pi*hy)*(j+oy))/Ly), sz likewise */
void wave3d_cpu_tables(const wave3d_problem* p, const wave3d_box* b, double* sx, double* sy, double* sz);
double wave3d_cpu_coslayer(const wave3d_problem* p, int n); /* cos(fma(a_t*n, tau, pi+pi)) */

void wave3d_cpu_layer0(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l, double* u0, int threads);

/* One update. first!=0: u_out = u1 + 0.5*a2*tau^2*lap(u1)   (layer 1; u_out is a fresh buffer)
 *             first==0: u_io  = 2*u1 - u_io + a2*tau^2*lap(u1) (in place over layer n-2)
 * Also returns max |u-f| and max |u-f|/|f| over this rank's part of the error set. */
void wave3d_cpu_step(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l, const double* u1,
                     double* u_io, int n, int first, double* max_abs, double* max_rel, int threads);

void wave3d_cpu_zero_walls(const wave3d_box* b, const wave3d_layout* l, double* u);

/* Pack the six send faces of `u` into contiguous buffers / unpack received faces
 * into ghost layers. face_elems gives the packed size per face. */
void wave3d_face_elems(const wave3d_box* b, int64_t elems[6]);
void wave3d_cpu_pack_face(const wave3d_box* b, const wave3d_layout* l, const double* u, int face, double* buf);
void wave3d_cpu_unpack_face(const wave3d_box* b, const wave3d_layout* l, double* u, int face, const double* buf);

/* Max abs / rel error of a stored layer against the analytic solution at layer n. */
void wave3d_cpu_field_error(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l, const double* u,
                            int n, double* max_abs, double* max_rel, int threads);

/* Whole solve on the CPU with `nranks` partitions of host memory ("fake peers"):
 * exercises decomposition + halo exchange without MPI or a GPU.
 * max_abs/max_rel: timesteps+1 entries. */
int wave3d_cpu_solve(const wave3d_problem* p, const int dims[3], int threads, double* max_abs, double* max_rel,
                     wave3d_timings* t, char* err, size_t errlen);

/* Independent, deliberately naive solver (all layers stored, 4-D arrays, error
 * post-pass — the shape of ref openmp_sol.cpp). Used to cross-check the fast paths. */
int wave3d_cpu_solve_naive(const wave3d_problem* p, double* max_abs, double* max_rel);

/* ------------------------------------------------------------- CUDA backend -- */

/* 1 if a CUDA device is usable in this process. */
int wave3d_cuda_available(void);
int wave3d_cuda_device_count(void);
const char* wave3d_cuda_last_error(void);

/* Tile plan of the fused kernel for one rank's box. */
#define WAVE3D_MAX_CHUNKS 32 /* x-chunks per tile column */
#define WAVE3D_FLAG_WORDS 40 /* epoch words per inbound face: [c] strip of x-chunk c (y/z faces), [WAVE3D_MAX_CHUNKS] whole layer */
typedef struct wave3d_plan {
    int32_t tz;      /* tile width in z (cells) */
    int32_t ty;      /* tile height in y = rj * warps */
    int32_t rj;      /* rows per warp */
    int32_t warps;   /* warps per CTA */
    int32_t nty, ntz; /* tiles in y and z */
    int32_t nchunks;  /* x-chunks per tile column (<= WAVE3D_MAX_CHUNKS); items = nty*ntz*nchunks, handed out dynamically */
    int32_t s1, s0;  /* ring depth: stencil-source planes (>= 4) / in-place planes */
    int32_t smem_bytes;
    int32_t ctas;    /* persistent CTAs pulling (tile, x-chunk) items from a device-side queue, chunk-major */
    int32_t slots;   /* resident CTAs on the device */
    double efficiency; /* modelled useful fraction */
    int32_t xcut[WAVE3D_MAX_CHUNKS + 2];  /* the chunk at position c along x covers owned planes [xcut[c], xcut[c+1]) */
    int32_t order[WAVE3D_MAX_CHUNKS]; /* queue order within a layer: order[i] = position of the i-th chunk. Zig-zag from both
                          ends (0, m-1, 1, m-2, ...) keeps x-adjacent chunks — the periodic wrap included —
                          close in the queue; lengths shrink along this order */
} wave3d_plan;

/* Choose the tile plan. sms<=0 -> 148. Env WAVE3D_TILE="ty,rj,nchunks[,s1,s0]" overrides (nchunks 0 = auto). */
void wave3d_plan_make(const wave3d_box* b, int sms, wave3d_plan* plan);
/* style 0: guided chunk lengths (one launch per layer); 1: equal chunks of ~32 planes (multi-layer launch;
 * env WAVE3D_CHUNK changes the target length). wave3d_plan_make picks the style by box size. */
void wave3d_plan_make_style(const wave3d_box* b, int sms, int style, wave3d_plan* plan);

/* --- kernel-level entry points (device pointers; stream = cudaStream_t or 0) --- */

/* K1-equivalent tables on the device: sx,sy,sz (libdevice sin, reference argument
 * expressions) and reciprocals rsx,rsy,rsz = 1/|s| used by the rel-error filter. */
int wave3d_cuda_tables(const wave3d_problem* p, const wave3d_box* b, double* sx, double* sy, double* sz, double* rsx,
                       double* rsy, double* rsz, void* stream);

/* Layer 0 on all owned points (ref K1 layer0, cuda_sol_kernels.cu:179-190). */
int wave3d_cuda_layer0(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l, const double* sx,
                       const double* sy, const double* sz, double* u0, void* stream);

/* Peer wiring of one rank for the fused halo push. All pointers are device
 * addresses valid on the launching GPU (own memory, peer-enabled or IPC-mapped).
 * dst[f] = address of the neighbour's ghost cell that receives my owned cell
 * with the two in-face indices 0; sa/sb = neighbour strides (doubles) along the
 * two in-face axes (x faces: j,k; y faces: i,k; z faces: i,j). */
typedef struct wave3d_peers {
    double* dst[6];
    int64_t sa[6];
    int64_t sb[6];
    uint32_t* flag_out[6]; /* neighbour's inbound flag block (WAVE3D_FLAG_WORDS epoch words) for this face pair (NULL: no flag) */
    uint32_t* flag_in[6];  /* my inbound flag blocks (NULL: do not wait) */
    uint64_t* root_err;    /* rank 0's error table (2*(timesteps+1) u64) or NULL */
} wave3d_peers;

/* Scratch words of one rank: [0] CTA ticket, [1] status (0 ok, 1 spin timeout), [2] work queue, rest reserved. */
#define WAVE3D_SCRATCH_WORDS 16

/* The fused step (ref K2/K3 + K4 + K5 + K6 + K7 + memsets + MPI p2p):
 * TMA-fed leapfrog update of layer n from u1 (layer n-1) in place over u_io
 * (layer n-2; for first!=0 u_io is the fresh layer-1 buffer and is only written),
 * analytic error max into err[2n],err[2n+1] (u64 bit patterns, atomicMax), face
 * push into the neighbours' ghost cells and epoch flags. `tabs` = the six tables
 * from wave3d_cuda_tables in order sx,sy,sz,rsx,rsy,rsz. */
int wave3d_cuda_step(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l, const wave3d_plan* plan,
                     const double* const tabs[6], const double* u1, double* u_io, int n, int first,
                     const wave3d_peers* peers, uint64_t* err, uint32_t* scratch, double spin_timeout_s,
                     void* stream);

/* Zero the Dirichlet walls (owned planes global y,z in {0,N}) (ref K4 bits 1-8). */
int wave3d_cuda_zero_walls(const wave3d_box* b, const wave3d_layout* l, double* u, void* stream);

/* Face pack / unpack for the staged transport (ref K5/K6); `push` copies my send
 * faces straight into the neighbours' ghosts (used for layer 0 and as non-fused path). */
int wave3d_cuda_pack_face(const wave3d_box* b, const wave3d_layout* l, const double* u, int face, double* buf,
                          void* stream);
int wave3d_cuda_unpack_face(const wave3d_box* b, const wave3d_layout* l, double* u, int face, const double* buf,
                            void* stream);
int wave3d_cuda_push_faces(const wave3d_box* b, const wave3d_layout* l, const double* u, const wave3d_peers* peers,
                           uint32_t epoch, void* stream);

/* Stand-alone error evaluation of a stored layer (ref K7 calculate_max_errors +
 * host loop, cuda_sol_kernels.cu:50-89) as a two-stage reduction. `scratch` holds
 * 2*nblocks+2 u64 words (wave3d_cuda_field_error_scratch_words) and must be zeroed
 * once before the first use; the kernel leaves it ready for the next call.
 * out[0]=max abs, out[1]=max rel (device doubles). */
int64_t wave3d_cuda_field_error_scratch_words(const wave3d_box* b);
int wave3d_cuda_field_error(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l,
                            const double* const tabs[6], const double* u, int n, uint64_t* scratch, double* out,
                            void* stream);

/* Same evaluation without any allocation, for callers that cannot obtain device memory when they
 * need the figure (a monitor attached to a run that has reserved all HBM, a latency-critical loop
 * that must avoid the device-wide synchronisation of cudaMalloc/cudaFree, code inside a foreign
 * framework's no-allocation region). This entry point has no counterpart in the reference.
 * The scratch is a ~19 KB __device__ array inside the library, present on every GPU that loads it
 * until the process exits. `blocks` = number of CTAs (<= 0: the rule of wave3d_cuda_field_error);
 * more than WAVE3D_FIELD_ERROR_STATIC_MAX_BLOCKS is refused with an error. All calls on a device
 * share that array: calls whose kernels may overlap on one GPU must be serialised by the caller.
 * Every call looks up the array's device address for the current device (so it stays valid across
 * cudaDeviceReset); the first call on a device may load the module — make it before starting a
 * stream capture on that device. out[0]=max abs, out[1]=max rel. */
#define WAVE3D_FIELD_ERROR_STATIC_MAX_BLOCKS 1184
int wave3d_cuda_field_error_static(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l,
                                   const double* const tabs[6], const double* u, int n, int blocks, double* out,
                                   void* stream);

/* ------------------------------------------------------------------ solver -- */
/* A solver owns the ranks that live in this process: all of them (native CLI,
 * single process driving 1..8 GPUs, or several ranks time-sharing one GPU for
 * tests) or exactly one (one process per GPU under torch.distributed). */

enum { WAVE3D_BACKEND_CPU = 0, WAVE3D_BACKEND_CUDA = 1 };
enum { WAVE3D_TRANSPORT_PEER = 0,   /* fused in-kernel stores + flags over peer memory */
       WAVE3D_TRANSPORT_STAGED = 1  /* pack -> caller moves buffers (NCCL / gloo / memcpy) -> unpack */ };

typedef struct wave3d_solver_opts {
    int32_t backend;
    int32_t transport;
    int32_t dims[3];       /* process grid; {0,0,0} -> wave3d_dims_create(nranks) */
    int32_t nranks;        /* world size */
    int32_t first_rank;    /* first rank owned by this process */
    int32_t nlocal;        /* ranks owned by this process */
    int32_t devices[8];    /* CUDA device per local rank (index by local rank % 8); -1 -> local rank % ndev */
    int32_t use_graph;     /* capture the whole solve in one CUDA graph */
    int32_t threads;       /* CPU backend threads */
    int32_t per_step_sync; /* read the step's error maxima back after every step (e2e / progress mode) */
    int32_t verbose;
    int32_t multistep;     /* layers 2..timesteps in ONE persistent dataflow launch per GPU (default 1; needs one rank per GPU) */
    int32_t reserved;
    double spin_timeout_s; /* bound on in-kernel flag waits; <=0 -> 10 s */
} wave3d_solver_opts;

typedef struct wave3d_solver wave3d_solver;

void wave3d_solver_opts_default(wave3d_solver_opts* o);
wave3d_solver* wave3d_solver_create(const wave3d_problem* p, const wave3d_solver_opts* o, char* err, size_t errlen);
void wave3d_solver_destroy(wave3d_solver* s);

/* Multi-process wiring (peer transport): export my rank's memory, import the
 * neighbours'. The blob is plain bytes, safe to all-gather. */
typedef struct wave3d_export {
    unsigned char ipc_handle[64]; /* cudaIpcMemHandle_t of the rank slab */
    int64_t off_field[2];         /* byte offsets inside the slab */
    int64_t off_flags;
    int64_t off_err;
    int64_t slab_bytes;
    wave3d_layout layout;
    wave3d_box box;
    int32_t device;
    int32_t pid;
    int32_t plan_nchunks;                    /* x-cuts of this rank's tile plan: y/z neighbours with equal cuts */
    int32_t plan_xcut[WAVE3D_MAX_CHUNKS + 2]; /*   hand-shake per (face, x-chunk) instead of per layer */
    int32_t reserved;
} wave3d_export;

int wave3d_solver_export(wave3d_solver* s, int local, wave3d_export* out);
/* `all` = exports of every rank in rank order (nranks entries). */
int wave3d_solver_connect(wave3d_solver* s, const wave3d_export* all, char* err, size_t errlen);

/* Geometry queries */
int wave3d_solver_box(const wave3d_solver* s, int local, wave3d_box* box, wave3d_layout* lay, wave3d_plan* plan);
/* Device (or host) pointer of field 0/1 of a local rank; layer n lives in field (n & 1). */
double* wave3d_solver_field(wave3d_solver* s, int local, int which);
void* wave3d_solver_stream(wave3d_solver* s, int local);

/* Phases. reset: zero the error table, the flag words and the work-queue / epoch counters (call before
 * every solve, then barrier across processes). The fields are NOT cleared: layer 0 rewrites every owned
 * cell and every ghost cell that is read is rewritten by a neighbour (or the wrap) before its first use.
 * enqueue_*: asynchronous on the rank streams. */
int wave3d_solver_reset(wave3d_solver* s);
int wave3d_solver_enqueue_layer0(wave3d_solver* s);
int wave3d_solver_enqueue_step(wave3d_solver* s, int n);
/* End-to-end flavour: the step's inputs {time argument of the analytic solution,
 * leapfrog coefficient} (wave3d_step_descriptor) are copied host->device from the
 * caller's pinned buffer on the rank stream and the kernel reads them from there. */
void wave3d_step_descriptor(const wave3d_problem* p, int n, double* desc2);
int wave3d_solver_enqueue_step_e2e(wave3d_solver* s, int n, const double* pinned_desc2);
/* Staged transport hooks: faces of layer n of local rank `local` to/from packed buffers
 * (device pointers for CUDA, host pointers for CPU). */
int wave3d_solver_pack(wave3d_solver* s, int local, int n, int face, double* buf);
int wave3d_solver_unpack(wave3d_solver* s, int local, int n, int face, const double* buf);
/* Intra-process exchange for the staged transport when all ranks are local. */
int wave3d_solver_exchange_local(wave3d_solver* s, int n);
int wave3d_solver_sync(wave3d_solver* s);

/* Whole solve (layer 0, layer 1, ..., timesteps) with device-side timing.
 * With nlocal == nranks this is the complete program; with one rank per
 * process call it on every rank between two barriers. */
int wave3d_solver_run(wave3d_solver* s, wave3d_timings* t);

/* Whole solve plus the reference's four phase totals (ref: cuda_sol.cpp:438-441), attributed by measurement:
 * one extra solve with the error evaluation switched off gives error_ms (the difference), exchange_ms is the
 * time CTAs waited for a neighbour rank's flag (peer transport; mean over CTAs, max over ranks) or the staged
 * exchanges, hostdev_ms is 0 (nothing is staged through the host), loop_ms is the rest. Results as run(). */
int wave3d_solver_measure_phases(wave3d_solver* s, wave3d_timings* t);

/* Streamed solve = the end-to-end flavour of a whole solve through the per-step API: for every layer the
 * step descriptor (wave3d_step_descriptor) is copied host->device from `pinned_desc` (2 doubles per layer,
 * pinned host memory), the fused kernel runs, and the layer's two maxima are copied device->host. The host
 * reads layer n's maxima before it enqueues layer n+1+lag (lag 0: fully synchronous, like the reference's
 * per-layer error read-back; lag 1 hides the host round trip behind the next layer). abs_rel receives
 * 2*(timesteps+1) doubles (max over the local ranks). solve_ms is host wall-clock time. */
int wave3d_solver_run_streamed(wave3d_solver* s, const double* pinned_desc, int lag, double* abs_rel, wave3d_timings* t);

/* Per-layer maxima over the local ranks (timesteps+1 entries each). */
int wave3d_solver_errors(wave3d_solver* s, double* max_abs, double* max_rel);
/* Error maxima of one layer, read back synchronously (per-step mode). */
int wave3d_solver_layer_error(wave3d_solver* s, int n, double* abs_rel2);
/* Post-pass error of the stored layer n (n = timesteps or timesteps-1 after a solve), evaluated by the
 * stand-alone reduction — the shape of the first-generation reference programs' calculate_error(). */
int wave3d_solver_field_error(wave3d_solver* s, int n, double* abs_rel2);
int64_t wave3d_solver_kernel_launches(const wave3d_solver* s);
/* 1 if run()/run_streamed() use the persistent multi-layer launch for layers 2..T (small boxes, one rank per GPU). */
int wave3d_solver_multistep_active(const wave3d_solver* s);
int wave3d_solver_status(wave3d_solver* s); /* 0 ok; 1 in-kernel spin timeout on some rank */

/* Debug aids (WAVE3D_DBG_TS=1): per-CTA phase timestamps (8 words per CTA, globaltimer ns) and per-CTA phase
 * accumulators (16 words per CTA, SM cycles; layout in csrc/wave3d_step_kernel.inc) of the last fused-step
 * launch on `device` (< 0: the current device). The two arrays (384 KB) stay resident on every GPU that loads
 * the library. scripts/cta_timeline.py and scripts/cta_phases.py print them. */
int wave3d_debug_copy_timestamps(unsigned long long* host, int nctas);
int wave3d_debug_copy_phases(unsigned long long* host, int nctas, int device);

/* The reference program end to end: parse, print "C = ...", solve, write the report. */
int wave3d_main(int argc, const char* const* argv);

#ifdef __cplusplus
}
#endif
#endif /* WAVE3D_H */
