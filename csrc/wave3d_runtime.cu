// wave3d_runtime.cu — solver object: rank slabs, tile planner, peer wiring (same
// device, peer-enabled devices of one process, or cudaIpc-mapped memory of other
// processes), stream/graph scheduling, staged-transport hooks, CPU backend glue and
// the reference-compatible program entry `wave3d_main`.
//
// Reference roles replaced here (behaviour only): main() set-up and allocation
// (ref: cuda_sol.cpp:477-575), calculate_start/calculate_num_sol time driver
// (:314-443), exchange() (:230-312), MPI_Reduce of the maxima (:431-432).
#include "wave3d_internal.hpp"

#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <vector>

using namespace w3d;

namespace {

void set_err(char* err, size_t errlen, const std::string& s) {
    if (err && errlen) snprintf(err, errlen, "%s", s.c_str());
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

constexpr size_t kFlagBytes = 1024;  // 6 inbound flag blocks of WAVE3D_FLAG_WORDS epoch words
static_assert(6 * WAVE3D_FLAG_WORDS * 4 <= kFlagBytes, "flag blocks");
constexpr size_t kChunkdoneBytes = 8 * 4 * WAVE3D_MAX_CHUNKS * 4;

struct Rank {
    int rank = 0, local = 0, device = -1;
    wave3d_box box;
    wave3d_layout lay;
    wave3d_plan plan;
    // ---- CUDA
    unsigned char* slab = nullptr;
    size_t slab_bytes = 0;
    size_t off_field[2] = {0, 0}, off_tabs[6] = {0}, off_err = 0, off_flags = 0, off_scratch = 0, off_desc = 0,
           off_ferr = 0, off_epoch = 0, off_stepdone = 0;
    size_t off_chunkdone = 0;
    size_t epoch_bytes = 0, stepdone_bytes = 0;
    uint32_t* chunkdone = nullptr; // multi-layer launch: y/z face strips done per (layer & 7, face, x-chunk)
    bool yz_chunk_flags = false;   // every y/z neighbour cuts x like this rank: per-chunk hand-shake is possible
    uint32_t* epoch = nullptr;     // multi-layer launch: last completed layer per (tile, chunk) item
    uint32_t* stepdone = nullptr;  // multi-layer launch: completed items per layer
    uint64_t* ferr_scratch = nullptr;  // scratch of the stand-alone error evaluation (zeroed with the slab)
    unsigned char* hmap = nullptr; // streamed multi-layer launch: mapped pinned memory (layer maxima + flags)
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copy = nullptr;
    double* field[2] = {nullptr, nullptr};
    double* tabs[6] = {nullptr};
    uint64_t* err = nullptr;
    uint32_t* flags = nullptr;
    uint32_t* scratch = nullptr;
    double* desc = nullptr;  // per-step descriptors (arg_t, coef) uploaded in per-step mode
    CUtensorMap map_halo[2], map_core[2];
    wave3d_peers peers[2];  // indexed by the field being written
    double* sendbuf[6] = {nullptr};
    double* recvbuf[6] = {nullptr};
    cudaEvent_t ev_packed = nullptr, ev_unpacked = nullptr;
    int group = 0;
    // ---- CPU
    std::vector<double> hfield[2];
    // ---- results
    std::vector<double> habs, hrel;
};

struct Group {  // local ranks that share one CUDA device share one stream
    int device = -1;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaGraphExec_t exec = nullptr;
    int64_t graph_launches = 0;  // kernels replayed by one launch of `exec`
    std::vector<int> members;
};

}  // namespace

struct wave3d_solver {
    wave3d_problem prob;
    wave3d_solver_opts opts;
    int dims[3];
    int div_mode = 0;
    bool connected = false;
    bool all_local = false;
    std::vector<Rank> ranks;
    std::vector<Group> groups;
    std::vector<void*> ipc_opened;
    std::vector<std::pair<int, int>> peer_enabled;
    int64_t launches = 0;
    double* pinned = nullptr;  // 2 doubles per layer per local rank (per-step readback) + descriptors
    double exchange_ms = 0.0;
    std::vector<cudaEvent_t> xev;  // event pairs around staged exchanges
    std::vector<cudaEvent_t> sev;  // streamed solve: per layer, per device group
    int xev_used = 0;
    bool fresh = false;  // reset() was called and nothing ran since
    bool noerr = false;  // timing aid of wave3d_solver_measure_phases: kernels skip the error evaluation
    bool ran_once = false;  // a solve has completed on this solver (kernels loaded, caches warm)
    bool eager = false;  // ... and launches are not replayed from the CUDA graph (same launch mode for the A/B pair)
};

namespace {

int use_device(int dev) { return check(cudaSetDevice(dev), "cudaSetDevice") ? 0 : -1; }

cudaStream_t stream_of(wave3d_solver* s, const Rank& r) { return s->groups[r.group].stream; }

int alloc_rank_cuda(wave3d_solver* s, Rank& r) {
    if (use_device(r.device)) return -1;
    const size_t fbytes = align_up((size_t)r.lay.elems * 8, 1024);
    size_t off = 0;
    for (int w = 0; w < 2; w++) {
        r.off_field[w] = off;
        off += fbytes;
    }
    for (int t = 0; t < 6; t++) {
        r.off_tabs[t] = off;
        off += align_up((size_t)r.box.n[t % 3] * 8, 256);
    }
    r.off_err = off;
    off += align_up((size_t)(s->prob.timesteps + 1) * 16, 256);
    r.off_flags = off;
    off += kFlagBytes;
    r.off_scratch = off;
    off += 256;
    r.off_desc = off;
    off += align_up((size_t)(s->prob.timesteps + 1) * 16, 256);
    r.off_ferr = off;
    off += align_up((size_t)wave3d_cuda_field_error_scratch_words(&r.box) * 8 + 16, 256);
    r.off_epoch = off;
    r.epoch_bytes = align_up((size_t)r.plan.nty * r.plan.ntz * r.plan.nchunks * 4, 256);
    off += r.epoch_bytes;
    r.off_stepdone = off;
    r.stepdone_bytes = align_up((size_t)(s->prob.timesteps + 2) * 4 * 3, 256);
    off += r.stepdone_bytes;
    r.off_chunkdone = off;  // directly behind stepdone: one memset clears epoch + stepdone + chunkdone
    off += kChunkdoneBytes;
    r.slab_bytes = off;
    W3D_TRY(cudaMalloc(&r.slab, r.slab_bytes));
    W3D_TRY(cudaMemsetAsync(r.slab, 0, r.slab_bytes, stream_of(s, r)));
    if (prepare_step_kernels(&r.plan, s->div_mode)) return -1;
    for (int w = 0; w < 2; w++) r.field[w] = reinterpret_cast<double*>(r.slab + r.off_field[w]);
    for (int t = 0; t < 6; t++) r.tabs[t] = reinterpret_cast<double*>(r.slab + r.off_tabs[t]);
    r.err = reinterpret_cast<uint64_t*>(r.slab + r.off_err);
    r.flags = reinterpret_cast<uint32_t*>(r.slab + r.off_flags);
    r.scratch = reinterpret_cast<uint32_t*>(r.slab + r.off_scratch);
    r.desc = reinterpret_cast<double*>(r.slab + r.off_desc);
    r.epoch = reinterpret_cast<uint32_t*>(r.slab + r.off_epoch);
    r.ferr_scratch = reinterpret_cast<uint64_t*>(r.slab + r.off_ferr);
    r.stepdone = reinterpret_cast<uint32_t*>(r.slab + r.off_stepdone);
    r.chunkdone = reinterpret_cast<uint32_t*>(r.slab + r.off_chunkdone);
    for (int w = 0; w < 2; w++) {
        if (encode_field_map(&r.map_halo[w], r.field[w], &r.lay, r.plan.tz + 4, r.plan.ty + 2)) return -1;
        if (encode_field_map(&r.map_core[w], r.field[w], &r.lay, r.plan.tz, r.plan.ty)) return -1;
    }
    memset(r.peers, 0, sizeof r.peers);
    if (s->opts.transport == WAVE3D_TRANSPORT_STAGED) {
        int64_t fe[6];
        wave3d_face_elems(&r.box, fe);
        for (int f = 0; f < 6; f++) {
            if (r.box.nbr[f] < 0) continue;
            W3D_TRY(cudaMalloc(&r.sendbuf[f], (size_t)fe[f] * 8));
            W3D_TRY(cudaMalloc(&r.recvbuf[f], (size_t)fe[f] * 8));
        }
        W3D_TRY(cudaEventCreateWithFlags(&r.ev_packed, cudaEventDisableTiming));
        W3D_TRY(cudaEventCreateWithFlags(&r.ev_unpacked, cudaEventDisableTiming));
    }
    return 0;
}

// Ghost-cell address on the neighbour that receives my face f, plus its in-face strides.
void ghost_target(const wave3d_box& qb, const wave3d_layout& ql, double* qfield, int f, double** dst, int64_t* sa,
                  int64_t* sb) {
    const int64_t ps = ql.plane_stride, pz = ql.pitch;
    const int64_t owned0 = ps + pz + ql.oz;  // storage offset of owned (0,0,0)
    switch (f) {
        case 0: *dst = qfield + owned0 + (int64_t)qb.n[0] * ps; *sa = pz; *sb = 1; break;  // lands in q's x+ ghost
        case 1: *dst = qfield + owned0 - ps;                    *sa = pz; *sb = 1; break;  // q's x- ghost
        case 2: *dst = qfield + owned0 + (int64_t)qb.n[1] * pz; *sa = ps; *sb = 1; break;
        case 3: *dst = qfield + owned0 - pz;                    *sa = ps; *sb = 1; break;
        case 4: *dst = qfield + owned0 + qb.n[2];               *sa = ps; *sb = pz; break;
        default: *dst = qfield + owned0 - 1;                    *sa = ps; *sb = pz; break;
    }
}

int enable_peer(wave3d_solver* s, int from, int to) {
    if (from == to) return 0;
    for (auto& pr : s->peer_enabled)
        if (pr.first == from && pr.second == to) return 0;
    int can = 0;
    W3D_TRY(cudaDeviceCanAccessPeer(&can, from, to));
    if (!can) {
        set_last_error("device " + std::to_string(from) + " cannot access peer " + std::to_string(to));
        return -1;
    }
    if (use_device(from)) return -1;
    cudaError_t e = cudaDeviceEnablePeerAccess(to, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
    else if (!check(e, "cudaDeviceEnablePeerAccess")) return -1;
    s->peer_enabled.push_back({from, to});
    return 0;
}

// Wire a rank to neighbour slabs given as base pointers valid on r.device.
void wire_rank(wave3d_solver* s, Rank& r, const std::vector<unsigned char*>& base,
               const std::vector<wave3d_export>& ex) {
    for (int w = 0; w < 2; w++) {
        wave3d_peers& P = r.peers[w];
        memset(&P, 0, sizeof P);
        for (int f = 0; f < 6; f++) {
            const int q = r.box.nbr[f];
            if (q < 0) continue;
            double* qfield = reinterpret_cast<double*>(base[q] + ex[q].off_field[w]);
            ghost_target(ex[q].box, ex[q].layout, qfield, f, &P.dst[f], &P.sa[f], &P.sb[f]);
            P.flag_out[f] = reinterpret_cast<uint32_t*>(base[q] + ex[q].off_flags) + (f ^ 1) * WAVE3D_FLAG_WORDS;
            P.flag_in[f] = r.flags + f * WAVE3D_FLAG_WORDS;
        }
        if (r.rank != 0 && base[0]) P.root_err = reinterpret_cast<uint64_t*>(base[0] + ex[0].off_err);
    }
    // y/z neighbours that cut x exactly like this rank can hand-shake per (face, x-chunk)
    r.yz_chunk_flags = true;
    for (int f = 2; f < 6; f++) {
        const int q = r.box.nbr[f];
        if (q < 0) continue;
        bool same = ex[q].plan_nchunks == r.plan.nchunks;
        for (int c = 0; same && c <= r.plan.nchunks; c++) same = ex[q].plan_xcut[c] == r.plan.xcut[c];
        if (!same) r.yz_chunk_flags = false;
    }
    if (const char* e = getenv("WAVE3D_YZ_CHUNK_FLAGS")) r.yz_chunk_flags = r.yz_chunk_flags && atoi(e) != 0;
    (void)s;
}

void fill_export(const wave3d_solver* s, const Rank& r, wave3d_export* out) {
    memset(out, 0, sizeof *out);
    out->off_field[0] = (int64_t)r.off_field[0];
    out->off_field[1] = (int64_t)r.off_field[1];
    out->off_flags = (int64_t)r.off_flags;
    out->off_err = (int64_t)r.off_err;
    out->slab_bytes = (int64_t)r.slab_bytes;
    out->layout = r.lay;
    out->box = r.box;
    out->device = r.device;
    out->pid = (int)getpid();
    out->plan_nchunks = r.plan.nchunks;
    for (int c = 0; c <= r.plan.nchunks && c < WAVE3D_MAX_CHUNKS + 2; c++) out->plan_xcut[c] = r.plan.xcut[c];
    (void)s;
}

int connect_local(wave3d_solver* s) {
    const int P = (int)s->ranks.size();
    std::vector<unsigned char*> base(P);
    std::vector<wave3d_export> ex(P);
    for (int r = 0; r < P; r++) {
        base[r] = s->ranks[r].slab;
        fill_export(s, s->ranks[r], &ex[r]);
    }
    for (auto& r : s->ranks) {
        for (int f = 0; f < 6; f++)
            if (r.box.nbr[f] >= 0 && enable_peer(s, r.device, s->ranks[r.box.nbr[f]].device)) return -1;
        if (enable_peer(s, r.device, s->ranks[0].device)) return -1;
        wire_rank(s, r, base, ex);
    }
    s->connected = true;
    return 0;
}

int enqueue_layer0_rank(wave3d_solver* s, Rank& r) {
    cudaStream_t st = stream_of(s, r);
    if (use_device(r.device)) return -1;
    if (wave3d_cuda_layer0(&s->prob, &r.box, &r.lay, r.tabs[0], r.tabs[1], r.tabs[2], r.field[0], st)) return -1;
    s->launches++;
    if (s->opts.transport == WAVE3D_TRANSPORT_PEER && s->connected) {
        if (wave3d_cuda_push_faces(&r.box, &r.lay, r.field[0], &r.peers[0], 1u, st)) return -1;
        for (int f = 0; f < 6; f++) s->launches += r.peers[0].dst[f] ? 1 : 0;
        s->launches++;
    }
    return 0;
}

void fill_launch(wave3d_solver* s, Rank& r, StepLaunch& a) {
    memset(&a, 0, sizeof a);
    for (int w = 0; w < 2; w++) {
        a.map_halo[w] = r.map_halo[w];
        a.map_core[w] = r.map_core[w];
        a.field[w] = r.field[w];
        if (s->opts.transport == WAVE3D_TRANSPORT_PEER && s->connected) a.peers[w] = r.peers[w];
    }
    a.prob = &s->prob;
    a.box = &r.box;
    a.lay = &r.lay;
    a.plan = &r.plan;
    for (int t = 0; t < 6; t++) a.tabs[t] = r.tabs[t];
    a.err = r.err;
    a.scratch = r.scratch;
    a.epoch = r.epoch;
    a.stepdone = r.stepdone;
    a.chunkdone = r.chunkdone;
    a.yz_chunk_flags = r.yz_chunk_flags ? 1 : 0;
    a.noerr = s->noerr ? 1 : 0;
    a.spin_timeout_s = s->opts.spin_timeout_s;
    a.div_mode = s->div_mode;
    a.stream = stream_of(s, r);
}

int enqueue_step_rank(wave3d_solver* s, Rank& r, int n, const double* pinned_desc) {
    if (use_device(r.device)) return -1;
    StepLaunch a;
    fill_launch(s, r, a);  // layer n is written into field n&1, the source is the other field
    a.n = n;
    a.nsteps = 1;
    a.first = n == 1;
    if (pinned_desc) {  // end-to-end mode: this step's inputs arrive from pinned host memory
        W3D_TRY(cudaMemcpyAsync(r.desc + 2 * n, pinned_desc, 16, cudaMemcpyHostToDevice, a.stream));
        a.desc = r.desc + 2 * n;
    }
    if (launch_step(a)) return -1;
    s->launches++;
    if (n == 1) {  // layer 0 carried analytic wall values; layer 2 is written in place over it
        if (wave3d_cuda_zero_walls(&r.box, &r.lay, r.field[0], a.stream)) return -1;
        s->launches++;
    }
    return 0;
}

// Layers n0 .. timesteps of one rank in a single persistent cooperative launch (k_step<..., MULTI>).
int enqueue_multi_rank(wave3d_solver* s, Rank& r, int n0, int nsteps, bool streamed = false) {
    if (use_device(r.device)) return -1;
    StepLaunch a;
    fill_launch(s, r, a);
    a.n = n0;
    a.nsteps = nsteps;
    a.first = 0;
    if (streamed) {
        a.desc_tab = r.desc;
        a.host_res = reinterpret_cast<double*>(r.hmap);
        a.host_flag = reinterpret_cast<uint32_t*>(r.hmap + (size_t)(s->prob.timesteps + 2) * 16);
    }
    if (launch_step(a)) return -1;
    s->launches++;
    return 0;
}

// Layers per multi-layer launch: at most 2048 (keeps the item counter far from 2^31 on very long runs) and
// never a left-over single layer (which would need the per-layer kernel).
int multi_chunk(int T) {
    int size = 2048;
    while (T - 1 > size && (T - 1) % size == 1) size--;
    return size;
}

// The multi-layer launch keeps a rank's kernel resident across layers while it waits for its
// neighbours' flags, so it needs the GPU for itself: one local rank per device.
bool multistep_ok(const wave3d_solver* s) {
    if (!s->opts.multistep || s->opts.backend != WAVE3D_BACKEND_CUDA) return false;
    if (s->opts.transport != WAVE3D_TRANSPORT_PEER || s->opts.per_step_sync) return false;
    if (s->prob.timesteps < 3) return false;
    for (auto& g : s->groups)
        if (g.members.size() != 1) return false;
    if (const char* e = getenv("WAVE3D_MULTISTEP")) {
        if (!strcmp(e, "0")) return false;
        if (!strcmp(e, "1")) return true;
    }
    // Measured on B200 (round 2 kernel, bench.py, Gcell/s, multi-layer launch vs one launch per layer):
    //   17 M cells per rank (N=256, 1 GPU): 204 vs 184;  34 M (N=512 on 4 GPUs, 257x257x513): 897 vs 867;
    //   68 M (N=512 on 2 GPUs): 435 vs 474;  135 M (N=512, 1 GPU): 237 vs 244.
    // Below ~50 M cells per rank the launch gap, pipeline fill and kernel tail of one launch per layer
    // dominate; above it the lock-step start of every layer (all CTAs in the same x range: halos hit L2,
    // DRAM pages are streamed in full) is worth more.
    for (auto& r : s->ranks)
        if ((double)r.box.n[0] * r.box.n[1] * r.box.n[2] > 48e6) return false;
    return true;
}

}  // namespace

// ------------------------------------------------------------------- planner

// Boxes below this size run layers 2..T in the persistent multi-layer launch (measured on B200: below
// ~50 M cells per rank the launch gap, pipeline fill and kernel tail of one launch per layer dominate).
static bool small_box(const wave3d_box* b) { return (double)b->n[0] * b->n[1] * b->n[2] <= 48e6; }

// style 0: guided chunk lengths (one launch per layer: long chunks first, short ones last, so that the
//          end of every launch is balanced); style 1: equal chunks of ~32 planes (multi-layer launch: no
//          per-layer tail to balance, fine items keep the dataflow between layers loose, and every rank
//          with the same x extent cuts x the same way, which the per-chunk y/z hand-shake needs).
extern "C" void wave3d_plan_make_style(const wave3d_box* b, int sms, int style, wave3d_plan* plan) {
    memset(plan, 0, sizeof *plan);
    if (sms <= 0) sms = 148;
    const int ny = b->hi[1] - b->lo[1] + 1, nz = b->hi[2] - b->lo[2] + 1, X = b->n[0];
    const int tz = 64;
    const int ntz = (nz + tz - 1) / tz;
    const bool have_gpu = wave3d_cuda_available();
    double best = -1.0;
    int ov[5] = {0, 0, 0, 0, 0};
    int nov = 0;
    if (const char* e = getenv("WAVE3D_TILE")) nov = sscanf(e, "%d,%d,%d,%d,%d", &ov[0], &ov[1], &ov[2], &ov[3], &ov[4]);
    // Chunk length of the multi-layer style: long chunks amortise the per-item fences and the two halo planes,
    // but a CTA should still get 2-3 items per layer or the dataflow between layers stiffens. Measured: 76 tiles
    // (256^3 box): 16 planes 181, 24: ~195, 32: 202, 43: ~198, 64: 175 Gcell/s; 152 tiles (257x257x513 box on 4
    // GPUs): 32: 838, 43: 878, 64: 871 Gcell/s. Rule: ~2.5 items per CTA and layer, never shorter than 32 planes.
    int chunk_env = 0;
    if (const char* e = getenv("WAVE3D_CHUNK")) chunk_env = std::max(2, atoi(e));
    for (int rj = 1; rj <= 2; rj *= 2) {
        for (int warps = 2; warps <= 7; warps++) {
            const int ty = rj * warps;
            if (nov >= 2 && (ty != ov[0] || rj != ov[1])) continue;
            // 6 source planes = 3 planes of look-ahead for the compute warps. Deeper rings put more rows of more
            // planes in flight at the DRAM controllers and cost bandwidth (measured: 8/4 is 6 % slower than 6/4
            // at N=512), shallower ones starve the compute warps.
            int s1 = nov >= 4 ? ov[3] : 6, s0 = nov >= 5 ? ov[4] : 4;
            int smem = plan_smem_bytes(ty, tz, s1, s0);
            while (smem > 106 * 1024 && s1 > 4) {  // keep two CTAs per SM (6 KB of static shared memory on top)
                s1--;
                if (s0 > 2) s0--;
                smem = plan_smem_bytes(ty, tz, s1, s0);
            }
            if (smem > 226 * 1024) continue;
            int occ = 0;
            if (have_gpu) occ = step_occupancy(rj, warps, smem);
            if (occ <= 0) {
                occ = std::min(232448 / (smem + 1024), std::min(64 / (warps + 1), 16 / (warps + 1)));
                if (occ < 1) occ = 1;
            }
            const int nty = (ny + ty - 1) / ty;
            const int ntiles = nty * ntz;
            const int slots = sms * occ;
            // Model (calibrated on B200): a CTA spends (planes + r0) plane-times per item, r0 = extra
            // x-halo planes + item set-up. Losses: masked cells of ragged tiles, halo ring of the source
            // boxes (L2 hits), RJ=1 spends more shared-memory traffic and instructions per cell, few warps
            // per SM cannot cover the LDS/FP64 latencies, unfilled last round.
            const double waste = (double)nty * ty * ntz * tz / ((double)ny * nz);
            const double halo = (double)(ty + 2) * (tz + 4) / ((double)ty * tz) - 1.0;
            const double rjf = rj == 1 ? 1.12 : 1.0;
            const double occf = occ >= 2 ? 1.0 : 1.25;
            const int tlp = occ * warps;
            const double tlpf = tlp >= 14 ? 1.0 : 1.0 + 0.04 * (14 - tlp);
            const double r0 = style ? 0.6 : 1.5;  // plane-times lost per item (look-ahead hides most of it)
            const int ctas = (int)std::min<int64_t>(slots, (int64_t)ntiles * std::max(1, X / 6));
            const double share = (double)ntiles * X / ctas;
            int minlen = std::max(4, std::min(X, std::max(8, X / 32)));
            if (const char* e = getenv("WAVE3D_MINLEN")) minlen = std::max(2, std::min(X, atoi(e)));
            int cuts[WAVE3D_MAX_CHUNKS + 2], ord[WAVE3D_MAX_CHUNKS];
            int m = 0;
            cuts[0] = 0;
            // the two end chunks must hold the seam planes (owned planes 1 and X-2): at least 2 planes each
            const int mmax = std::max(1, std::min(WAVE3D_MAX_CHUNKS, X / 2));
            if (nov >= 3 && ov[2] > 0) {
                m = std::min(mmax, ov[2]);
                for (int c = 1; c <= m; c++) cuts[c] = (int)((int64_t)X * c / m);
                for (int i = 0; i < m; i++) ord[i] = (i % 2 == 0) ? i / 2 : m - 1 - i / 2;
            } else if (style) {
                int chunk_len = chunk_env;
                if (!chunk_len) {
                    const int want = std::max(1, (int)std::lround(2.5 * slots / ntiles));
                    chunk_len = std::max(32, X / want);
                }
                m = std::max(1, std::min(mmax, (X + chunk_len / 2) / chunk_len));
                // tiny boxes: rather short chunks than idle CTAs
                while (!chunk_env && ntiles * m < slots && m < mmax && X / (m + 1) >= 8) m++;
                for (int c = 1; c <= m; c++) cuts[c] = (int)((int64_t)X * c / m);
                for (int i = 0; i < m; i++) ord[i] = (i % 2 == 0) ? i / 2 : m - 1 - i / 2;
            } else {
                // built from the end of the column: two floor-sized chunks, then doubling up to the cap;
                // what is left over goes to the front, where imbalance can still be repaired
                const int cap = std::max(minlen, (int)std::min(X / 2.0, share * 0.6));
                int len[16];
                int left = X, l = minlen, k = 0;
                while (left > 0 && k < 16) {
                    int take = std::min(l, left);
                    if (k == 15 || left - take < minlen) take = left;
                    len[k++] = take;
                    left -= take;
                    if (k >= 2) l = std::min(cap, l * 2);
                }
                m = k;  // len[] is in back-to-front order
                std::sort(len, len + m);
                std::reverse(len, len + m);  // len[i] = length of the i-th chunk in queue order (descending)
                // zig-zag placement along x: queue order visits positions 0, m-1, 1, m-2, ...
                int plen[16];
                for (int i = 0; i < m; i++) {
                    ord[i] = (i % 2 == 0) ? i / 2 : m - 1 - i / 2;
                    plen[ord[i]] = len[i];
                }
                for (int c = 0; c < m; c++) cuts[c + 1] = cuts[c] + plen[c];
            }
            const double tail = style ? 0.0 : 0.5 * (cuts[ord[m - 1] + 1] - cuts[ord[m - 1]]) / std::max(1.0, share);
            const double fill = std::max(1.0, (double)slots / ((double)ntiles * m));
            const double eff = (double)X / (X + m * r0) / (1.0 + tail) / fill /
                               (waste * (1.0 + 0.3 * halo) * rjf * occf * tlpf);
            if (eff > best + 1e-9) {
                best = eff;
                plan->tz = tz; plan->ty = ty; plan->rj = rj; plan->warps = warps;
                plan->nty = nty; plan->ntz = ntz; plan->nchunks = m;
                plan->s1 = s1; plan->s0 = s0; plan->smem_bytes = smem;
                plan->ctas = (int)std::min<int64_t>(ctas, (int64_t)ntiles * m);
                plan->slots = slots; plan->efficiency = eff;
                for (int c = 0; c <= m; c++) plan->xcut[c] = cuts[c];
                for (int c = 0; c < m; c++) plan->order[c] = ord[c];
            }
        }
    }
}

extern "C" void wave3d_plan_make(const wave3d_box* b, int sms, wave3d_plan* plan) {
    wave3d_plan_make_style(b, sms, small_box(b) ? 1 : 0, plan);
}

// -------------------------------------------------------------------- solver

extern "C" {

void wave3d_solver_opts_default(wave3d_solver_opts* o) {
    memset(o, 0, sizeof *o);
    o->backend = WAVE3D_BACKEND_CUDA;
    o->transport = WAVE3D_TRANSPORT_PEER;
    o->nranks = 1;
    o->first_rank = 0;
    o->nlocal = 1;
    for (int i = 0; i < 8; i++) o->devices[i] = -1;
    o->use_graph = 1;
    o->multistep = 1;
    o->spin_timeout_s = 10.0;
}

void wave3d_solver_destroy(wave3d_solver* s) {
    if (!s) return;
    if (s->opts.backend == WAVE3D_BACKEND_CUDA) {
        for (auto& g : s->groups) {
            cudaSetDevice(g.device);
            cudaStreamSynchronize(g.stream);
            if (g.exec) cudaGraphExecDestroy(g.exec);
            if (g.ev0) cudaEventDestroy(g.ev0);
            if (g.ev1) cudaEventDestroy(g.ev1);
            if (g.stream) cudaStreamDestroy(g.stream);
        }
        for (void* p : s->ipc_opened) cudaIpcCloseMemHandle(p);
        for (auto& r : s->ranks) {
            cudaSetDevice(r.device);
            for (int f = 0; f < 6; f++) {
                if (r.sendbuf[f]) cudaFree(r.sendbuf[f]);
                if (r.recvbuf[f]) cudaFree(r.recvbuf[f]);
            }
            if (r.ev_packed) cudaEventDestroy(r.ev_packed);
            if (r.ev_unpacked) cudaEventDestroy(r.ev_unpacked);
            if (r.hmap) cudaFreeHost(r.hmap);
            if (r.copy_stream) cudaStreamDestroy(r.copy_stream);
            if (r.ev_copy) cudaEventDestroy(r.ev_copy);
            if (r.slab) cudaFree(r.slab);
        }
        for (cudaEvent_t e : s->xev) cudaEventDestroy(e);
        for (cudaEvent_t e : s->sev) cudaEventDestroy(e);
        if (s->pinned) cudaFreeHost(s->pinned);
        cudaGetLastError();
    }
    delete s;
}

wave3d_solver* wave3d_solver_create(const wave3d_problem* p, const wave3d_solver_opts* o, char* err, size_t errlen) {
    wave3d_solver* s = new wave3d_solver();
    s->prob = *p;
    s->opts = *o;
    if (s->opts.nranks < 1) s->opts.nranks = 1;
    if (s->opts.nlocal < 1) s->opts.nlocal = 1;
    if (o->dims[0] > 0 && o->dims[1] > 0 && o->dims[2] > 0) {
        memcpy(s->dims, o->dims, sizeof s->dims);
        if (s->dims[0] * s->dims[1] * s->dims[2] != s->opts.nranks) {
            set_err(err, errlen, "process grid does not match the number of ranks");
            delete s;
            return nullptr;
        }
    } else {
        wave3d_dims_create(s->opts.nranks, s->dims);
    }
    if (s->opts.first_rank < 0 || s->opts.first_rank + s->opts.nlocal > s->opts.nranks) {
        set_err(err, errlen, "local rank range outside the world");
        delete s;
        return nullptr;
    }
    s->all_local = s->opts.nlocal == s->opts.nranks;
    s->div_mode = choose_div_mode(p);
    const bool cuda = s->opts.backend == WAVE3D_BACKEND_CUDA;
    int ndev = 0, sms = 148;
    if (cuda) {
        ndev = wave3d_cuda_device_count();
        if (ndev < 1) {
            set_err(err, errlen, "no CUDA device available (backend=cuda)");
            delete s;
            return nullptr;
        }
    }
    s->ranks.resize(s->opts.nlocal);
    for (int l = 0; l < s->opts.nlocal; l++) {
        Rank& r = s->ranks[l];
        r.local = l;
        r.rank = s->opts.first_rank + l;
        if (wave3d_box_init(p->N, s->dims, r.rank, &r.box, err, errlen)) {
            delete s;
            return nullptr;
        }
        wave3d_layout_init(&r.box, &r.lay);
        r.habs.assign(p->timesteps + 1, 0.0);
        r.hrel.assign(p->timesteps + 1, 0.0);
        if (cuda) {
            int d = s->opts.devices[l % 8];
            r.device = d >= 0 ? d : l % ndev;
            if (r.device >= ndev) {
                set_err(err, errlen, "device index out of range");
                delete s;
                return nullptr;
            }
            // one stream per device, shared by the ranks that time-share it
            int g = -1;
            for (size_t i = 0; i < s->groups.size(); i++)
                if (s->groups[i].device == r.device) g = (int)i;
            if (g < 0) {
                Group grp;
                grp.device = r.device;
                s->groups.push_back(grp);
                g = (int)s->groups.size() - 1;
            }
            r.group = g;
            s->groups[g].members.push_back(l);
        } else {
            r.hfield[0].assign(r.lay.elems, 0.0);
            r.hfield[1].assign(r.lay.elems, 0.0);
        }
    }
    if (cuda) {
        auto fail = [&](const char* what) -> wave3d_solver* {
            set_err(err, errlen, std::string(what) + ": " + wave3d_cuda_last_error());
            wave3d_solver_destroy(s);
            return nullptr;
        };
        for (auto& g : s->groups) {
            if (use_device(g.device)) return fail("cudaSetDevice");
            if (!check(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking), "cudaStreamCreate"))
                return fail("stream");
            if (!check(cudaEventCreate(&g.ev0), "cudaEventCreate") || !check(cudaEventCreate(&g.ev1), "cudaEventCreate"))
                return fail("event");
        }
        for (auto& r : s->ranks) {
            if (use_device(r.device)) return fail("cudaSetDevice");
            cudaDeviceProp prop;
            if (check(cudaGetDeviceProperties(&prop, r.device), "props")) sms = prop.multiProcessorCount;
            wave3d_plan_make_style(&r.box, sms, multistep_ok(s) ? 1 : 0, &r.plan);
            if (r.plan.ty == 0) return fail("no feasible tile plan");
            if (alloc_rank_cuda(s, r)) return fail("allocation");
            if (wave3d_cuda_tables(&s->prob, &r.box, r.tabs[0], r.tabs[1], r.tabs[2], r.tabs[3], r.tabs[4], r.tabs[5],
                                   stream_of(s, r)))
                return fail("tables");
            s->launches++;
        }
        if (!check(cudaMallocHost(&s->pinned, sizeof(double) * 4 * (size_t)(p->timesteps + 1) * s->ranks.size()),
                   "cudaMallocHost"))
            return fail("pinned");
        if (s->all_local && s->opts.transport == WAVE3D_TRANSPORT_PEER)
            if (connect_local(s)) return fail("peer wiring");
        for (auto& g : s->groups) {
            cudaSetDevice(g.device);
            if (!check(cudaStreamSynchronize(g.stream), "sync")) return fail("init sync");
        }
    }
    return s;
}

int wave3d_solver_export(wave3d_solver* s, int local, wave3d_export* out) {
    if (!s || local < 0 || local >= (int)s->ranks.size()) return -1;
    Rank& r = s->ranks[local];
    fill_export(s, r, out);
    if (s->opts.backend == WAVE3D_BACKEND_CUDA) {
        if (use_device(r.device)) return -1;
        cudaIpcMemHandle_t h;
        W3D_TRY(cudaIpcGetMemHandle(&h, r.slab));
        static_assert(sizeof(h) <= sizeof(out->ipc_handle), "ipc handle size");
        memcpy(out->ipc_handle, &h, sizeof h);
    }
    return 0;
}

int wave3d_solver_connect(wave3d_solver* s, const wave3d_export* all, char* err, size_t errlen) {
    if (!s || s->opts.backend != WAVE3D_BACKEND_CUDA) return -1;
    const int P = s->opts.nranks;
    std::vector<wave3d_export> ex(all, all + P);
    const int mypid = (int)getpid();
    for (auto& r : s->ranks) {
        if (use_device(r.device)) return -1;
        std::vector<unsigned char*> base(P, nullptr);
        std::vector<int> need;
        for (int f = 0; f < 6; f++)
            if (r.box.nbr[f] >= 0) need.push_back(r.box.nbr[f]);
        need.push_back(0);
        for (int q : need) {
            if (base[q]) continue;
            // local rank of this very process?
            Rank* lr = nullptr;
            for (auto& o : s->ranks)
                if (o.rank == q && ex[q].pid == mypid) lr = &o;
            if (lr) {
                if (enable_peer(s, r.device, lr->device)) {
                    set_err(err, errlen, wave3d_cuda_last_error());
                    return -1;
                }
                base[q] = lr->slab;
                continue;
            }
            cudaIpcMemHandle_t h;
            memcpy(&h, ex[q].ipc_handle, sizeof h);
            void* ptr = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                set_err(err, errlen, std::string("cudaIpcOpenMemHandle(rank ") + std::to_string(q) +
                                         "): " + cudaGetErrorString(e));
                cudaGetLastError();
                return -1;
            }
            s->ipc_opened.push_back(ptr);
            base[q] = static_cast<unsigned char*>(ptr);
        }
        wire_rank(s, r, base, ex);
    }
    s->connected = true;
    return 0;
}

int wave3d_solver_box(const wave3d_solver* s, int local, wave3d_box* box, wave3d_layout* lay, wave3d_plan* plan) {
    if (!s || local < 0 || local >= (int)s->ranks.size()) return -1;
    if (box) *box = s->ranks[local].box;
    if (lay) *lay = s->ranks[local].lay;
    if (plan) *plan = s->ranks[local].plan;
    return 0;
}

double* wave3d_solver_field(wave3d_solver* s, int local, int which) {
    if (!s || local < 0 || local >= (int)s->ranks.size()) return nullptr;
    Rank& r = s->ranks[local];
    return s->opts.backend == WAVE3D_BACKEND_CUDA ? r.field[which & 1] : r.hfield[which & 1].data();
}

void* wave3d_solver_stream(wave3d_solver* s, int local) {
    if (!s || s->opts.backend != WAVE3D_BACKEND_CUDA) return nullptr;
    return stream_of(s, s->ranks[local]);
}

int wave3d_solver_sync(wave3d_solver* s) {
    if (s->opts.backend != WAVE3D_BACKEND_CUDA) return 0;
    for (auto& g : s->groups) {
        if (use_device(g.device)) return -1;
        W3D_TRY(cudaStreamSynchronize(g.stream));
    }
    return 0;
}

int wave3d_solver_reset(wave3d_solver* s) {
    for (auto& r : s->ranks) {
        std::fill(r.habs.begin(), r.habs.end(), 0.0);
        std::fill(r.hrel.begin(), r.hrel.end(), 0.0);
    }
    s->fresh = true;
    if (s->opts.backend != WAVE3D_BACKEND_CUDA) return 0;
    for (auto& r : s->ranks) {
        if (use_device(r.device)) return -1;
        cudaStream_t st = stream_of(s, r);
        W3D_TRY(cudaMemsetAsync(r.err, 0, (size_t)(s->prob.timesteps + 1) * 16, st));
        W3D_TRY(cudaMemsetAsync(r.flags, 0, kFlagBytes, st));
        W3D_TRY(cudaMemsetAsync(r.scratch, 0, 256, st));
        W3D_TRY(cudaMemsetAsync(r.epoch, 0, r.epoch_bytes + r.stepdone_bytes + kChunkdoneBytes, st));
    }
    s->fresh = true;
    return wave3d_solver_sync(s);
}

int wave3d_solver_enqueue_layer0(wave3d_solver* s) {
    s->fresh = false;
    if (s->opts.backend == WAVE3D_BACKEND_CUDA) {
        for (auto& r : s->ranks)
            if (enqueue_layer0_rank(s, r)) return -1;
        return 0;
    }
    for (auto& r : s->ranks) wave3d_cpu_layer0(&s->prob, &r.box, &r.lay, r.hfield[0].data(), s->opts.threads);
    return 0;
}

int wave3d_solver_enqueue_step(wave3d_solver* s, int n) {
    if (n < 1 || n > s->prob.timesteps) return -1;
    if (s->opts.backend == WAVE3D_BACKEND_CUDA) {
        for (auto& r : s->ranks)
            if (enqueue_step_rank(s, r, n, nullptr)) return -1;
        return 0;
    }
    // Ranks are independent within a layer. With one thread per rank (the pure "MPI" flavour of the
    // reference) the local ranks run side by side; with several threads per rank (OpenMP / hybrid
    // flavour) each rank's loop nest is threaded instead.
    const int nl = (int)s->ranks.size();
    const bool side_by_side = nl > 1 && s->opts.threads <= 1;
#pragma omp parallel for schedule(static) if (side_by_side)
    for (int l = 0; l < nl; l++) {
        Rank& r = s->ranks[l];
        wave3d_cpu_step(&s->prob, &r.box, &r.lay, r.hfield[(n & 1) ^ 1].data(), r.hfield[n & 1].data(), n, n == 1,
                        &r.habs[n], &r.hrel[n], side_by_side ? 1 : s->opts.threads);
        if (n == 1) wave3d_cpu_zero_walls(&r.box, &r.lay, r.hfield[0].data());
    }
    return 0;
}

int wave3d_solver_enqueue_step_e2e(wave3d_solver* s, int n, const double* pinned_desc) {
    if (n < 1 || n > s->prob.timesteps || s->opts.backend != WAVE3D_BACKEND_CUDA) return -1;
    for (auto& r : s->ranks)
        if (enqueue_step_rank(s, r, n, pinned_desc)) return -1;
    return 0;
}

void wave3d_step_descriptor(const wave3d_problem* p, int n, double* desc2) {
    desc2[0] = std::fma(p->a_t * (double)n, p->tau, p->pi + p->pi);
    desc2[1] = (p->a2 * p->tau) * p->tau;
    if (n == 1) desc2[1] = desc2[1] * 0.5;
}

int wave3d_solver_pack(wave3d_solver* s, int local, int n, int face, double* buf) {
    Rank& r = s->ranks[local];
    if (s->opts.backend == WAVE3D_BACKEND_CUDA) {
        if (use_device(r.device)) return -1;
        s->launches++;
        return wave3d_cuda_pack_face(&r.box, &r.lay, r.field[n & 1], face, buf, stream_of(s, r));
    }
    wave3d_cpu_pack_face(&r.box, &r.lay, r.hfield[n & 1].data(), face, buf);
    return 0;
}

int wave3d_solver_unpack(wave3d_solver* s, int local, int n, int face, const double* buf) {
    Rank& r = s->ranks[local];
    if (s->opts.backend == WAVE3D_BACKEND_CUDA) {
        if (use_device(r.device)) return -1;
        s->launches++;
        return wave3d_cuda_unpack_face(&r.box, &r.lay, r.field[n & 1], face, buf, stream_of(s, r));
    }
    wave3d_cpu_unpack_face(&r.box, &r.lay, r.hfield[n & 1].data(), face, buf);
    return 0;
}

// Staged exchange between ranks of one process: pack -> device copy -> unpack, ordered
// with events (the non-fused A/B path and the fallback when peer stores are unavailable).
int wave3d_solver_exchange_local(wave3d_solver* s, int n) {
    if (!s->all_local) return -1;
    if (s->opts.backend != WAVE3D_BACKEND_CUDA) {
        std::vector<double> buf;
        for (auto& r : s->ranks) {
            int64_t fe[6];
            wave3d_face_elems(&r.box, fe);
            for (int f = 0; f < 6; f++) {
                const int q = r.box.nbr[f];
                if (q < 0) continue;
                buf.resize(fe[f]);
                wave3d_cpu_pack_face(&r.box, &r.lay, r.hfield[n & 1].data(), f, buf.data());
                Rank& t = s->ranks[q];
                wave3d_cpu_unpack_face(&t.box, &t.lay, t.hfield[n & 1].data(), f ^ 1, buf.data());
            }
        }
        return 0;
    }
    if (s->opts.transport != WAVE3D_TRANSPORT_STAGED) return -1;
    for (auto& r : s->ranks) {
        if (use_device(r.device)) return -1;
        cudaStream_t st = stream_of(s, r);
        // the receivers of the previous exchange must be done reading my send buffers
        for (int f = 0; f < 6; f++)
            if (r.box.nbr[f] >= 0) W3D_TRY(cudaStreamWaitEvent(st, s->ranks[r.box.nbr[f]].ev_unpacked, 0));
        for (int f = 0; f < 6; f++)
            if (r.box.nbr[f] >= 0 && wave3d_solver_pack(s, r.local, n, f, r.sendbuf[f])) return -1;
        W3D_TRY(cudaEventRecord(r.ev_packed, st));
    }
    for (auto& r : s->ranks) {
        if (use_device(r.device)) return -1;
        cudaStream_t st = stream_of(s, r);
        int64_t fe[6];
        wave3d_face_elems(&r.box, fe);
        for (int f = 0; f < 6; f++) {
            const int q = r.box.nbr[f];
            if (q < 0) continue;
            Rank& src = s->ranks[q];
            W3D_TRY(cudaStreamWaitEvent(st, src.ev_packed, 0));
            // my ghost face f receives the neighbour's send face f^1
            W3D_TRY(cudaMemcpyPeerAsync(r.recvbuf[f], r.device, src.sendbuf[f ^ 1], src.device, (size_t)fe[f] * 8, st));
            if (wave3d_solver_unpack(s, r.local, n, f, r.recvbuf[f])) return -1;
        }
        W3D_TRY(cudaEventRecord(r.ev_unpacked, st));
    }
    return 0;
}

static int fetch_errors(wave3d_solver* s) {
    if (s->opts.backend != WAVE3D_BACKEND_CUDA) return 0;
    if (wave3d_solver_sync(s)) return -1;
    const int L = s->prob.timesteps + 1;
    std::vector<uint64_t> tmp(2 * (size_t)L);
    for (auto& r : s->ranks) {
        if (use_device(r.device)) return -1;
        W3D_TRY(cudaMemcpy(tmp.data(), r.err, tmp.size() * 8, cudaMemcpyDeviceToHost));
        for (int n = 0; n < L; n++) {
            memcpy(&r.habs[n], &tmp[2 * n], 8);
            memcpy(&r.hrel[n], &tmp[2 * n + 1], 8);
        }
    }
    return 0;
}

int wave3d_solver_layer_error(wave3d_solver* s, int n, double* out2) {
    out2[0] = out2[1] = 0.0;
    if (s->opts.backend == WAVE3D_BACKEND_CUDA) {
        // pinned readback of the two maxima of layer n from every local rank, then one sync per device
        for (auto& r : s->ranks) {
            if (use_device(r.device)) return -1;
            W3D_TRY(cudaMemcpyAsync(s->pinned + 2 * r.local, r.err + 2 * n, 16, cudaMemcpyDeviceToHost, stream_of(s, r)));
        }
        if (wave3d_solver_sync(s)) return -1;
        for (auto& r : s->ranks) {
            r.habs[n] = s->pinned[2 * r.local];
            r.hrel[n] = s->pinned[2 * r.local + 1];
        }
    }
    for (auto& r : s->ranks) {
        if (r.habs[n] > out2[0]) out2[0] = r.habs[n];
        if (r.hrel[n] > out2[1]) out2[1] = r.hrel[n];
    }
    return 0;
}

int wave3d_solver_errors(wave3d_solver* s, double* max_abs, double* max_rel) {
    if (fetch_errors(s)) return -1;
    for (int n = 0; n <= s->prob.timesteps; n++) {
        double a = 0.0, r_ = 0.0;
        for (auto& r : s->ranks) {
            if (r.habs[n] > a) a = r.habs[n];
            if (r.hrel[n] > r_) r_ = r.hrel[n];
        }
        max_abs[n] = a;
        max_rel[n] = r_;
    }
    return 0;
}

int wave3d_solver_run_streamed(wave3d_solver* s, const double* pinned_desc, int lag, double* abs_rel,
                               wave3d_timings* t) {
    if (!s || s->opts.backend != WAVE3D_BACKEND_CUDA || !pinned_desc || !abs_rel) return -1;
    const int T = s->prob.timesteps, L = (int)s->ranks.size();
    if (lag < 0) lag = 0;
    const bool staged = s->opts.transport == WAVE3D_TRANSPORT_STAGED;
    if (staged && !s->all_local) {
        set_last_error("streamed solve with the staged transport across processes is driven by the caller");
        return -1;
    }
    if (!s->fresh && wave3d_solver_reset(s)) return -1;
    s->fresh = false;
    // one event per layer per device group marks "this layer's maxima are in pinned memory"
    while (s->sev.size() < (size_t)(T + 1) * s->groups.size()) {
        cudaEvent_t e;
        if (use_device(s->groups[s->sev.size() % s->groups.size()].device)) return -1;
        W3D_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        s->sev.push_back(e);
    }
    for (int i = 0; i < 2 * (T + 1); i++) abs_rel[i] = 0.0;
    if (!staged && multistep_ok(s)) {
        // ---- small boxes: layers 2..T run in the persistent multi-layer launch while the host streams the
        // per-layer inputs in (one pinned->device copy per layer on a copy stream; the kernel waits for
        // layer n's descriptor to land) and the kernel streams the per-layer results out (maxima + ready
        // flag written to mapped pinned memory when a layer completes; the host polls the flag).
        const size_t res_bytes = (size_t)(T + 2) * 16, map_bytes = res_bytes + (size_t)(T + 2) * 4;
        for (auto& r : s->ranks) {
            if (use_device(r.device)) return -1;
            if (!r.hmap) {
                W3D_TRY(cudaHostAlloc(&r.hmap, map_bytes, cudaHostAllocMapped | cudaHostAllocPortable));
                W3D_TRY(cudaStreamCreateWithFlags(&r.copy_stream, cudaStreamNonBlocking));
                W3D_TRY(cudaEventCreateWithFlags(&r.ev_copy, cudaEventDisableTiming));
            }
            memset(r.hmap, 0, map_bytes);
        }
        auto t0 = std::chrono::steady_clock::now();
        for (auto& r : s->ranks) {
            if (use_device(r.device)) return -1;
            cudaStream_t st = stream_of(s, r);
            W3D_TRY(cudaMemsetAsync(r.desc, 0xFF, (size_t)(T + 1) * 16, st));  // all-ones = NaN: "not uploaded yet"
            W3D_TRY(cudaEventRecord(r.ev_copy, st));
            W3D_TRY(cudaStreamWaitEvent(r.copy_stream, r.ev_copy, 0));
            if (enqueue_layer0_rank(s, r)) return -1;
            if (enqueue_step_rank(s, r, 1, pinned_desc + 2)) return -1;
            W3D_TRY(cudaMemcpyAsync(s->pinned + 2 * ((size_t)1 * L + r.local), r.err + 2, 16, cudaMemcpyDeviceToHost, st));
        }
        for (size_t g = 0; g < s->groups.size(); g++) {
            if (use_device(s->groups[g].device)) return -1;
            W3D_TRY(cudaEventRecord(s->sev[s->groups.size() + g], s->groups[g].stream));
        }
        const int chunk = multi_chunk(T);
        for (int n0 = 2; n0 <= T; n0 += chunk) {
            const int ns = std::min(chunk, T - n0 + 1);
            for (auto& r : s->ranks)
                if (enqueue_multi_rank(s, r, n0, ns, true)) return -1;
        }
        for (int n = 2; n <= T; n++)
            for (auto& r : s->ranks) {
                if (use_device(r.device)) return -1;
                W3D_TRY(cudaMemcpyAsync(r.desc + 2 * n, pinned_desc + 2 * n, 16, cudaMemcpyHostToDevice, r.copy_stream));
            }
        // layer 1 came through the per-layer path
        for (size_t g = 0; g < s->groups.size(); g++) {
            if (use_device(s->groups[g].device)) return -1;
            W3D_TRY(cudaEventSynchronize(s->sev[s->groups.size() + g]));
        }
        for (auto& r : s->ranks) {
            const double* src = s->pinned + 2 * ((size_t)1 * L + r.local);
            r.habs[1] = src[0];
            r.hrel[1] = src[1];
            abs_rel[2] = std::max(abs_rel[2], src[0]);
            abs_rel[3] = std::max(abs_rel[3], src[1]);
        }
        const auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(120);
        for (int n = 2; n <= T; n++) {
            for (auto& r : s->ranks) {
                volatile uint32_t* flag = reinterpret_cast<volatile uint32_t*>(r.hmap + res_bytes) + n;
                while (*flag == 0) {
                    if (std::chrono::steady_clock::now() > deadline) {
                        set_last_error("streamed solve: layer result did not arrive");
                        return -2;
                    }
                }
                __sync_synchronize();
                const double* src = reinterpret_cast<const double*>(r.hmap) + 2 * (size_t)n;
                r.habs[n] = src[0];
                r.hrel[n] = src[1];
                abs_rel[2 * n] = std::max(abs_rel[2 * n], src[0]);
                abs_rel[2 * n + 1] = std::max(abs_rel[2 * n + 1], src[1]);
            }
        }
        if (wave3d_solver_sync(s)) return -1;
        for (auto& r : s->ranks) {
            if (use_device(r.device)) return -1;
            W3D_TRY(cudaStreamSynchronize(r.copy_stream));
        }
        auto t1 = std::chrono::steady_clock::now();
        if (t) {
            memset(t, 0, sizeof *t);
            t->solve_ms = t->loop_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
        }
        return wave3d_solver_status(s) ? -2 : 0;
    }
    auto collect = [&](int n) -> int {
        for (size_t g = 0; g < s->groups.size(); g++) {
            if (use_device(s->groups[g].device)) return -1;
            W3D_TRY(cudaEventSynchronize(s->sev[(size_t)n * s->groups.size() + g]));
        }
        for (auto& r : s->ranks) {
            const double* src = s->pinned + 2 * ((size_t)n * L + r.local);
            r.habs[n] = src[0];
            r.hrel[n] = src[1];
            if (src[0] > abs_rel[2 * n]) abs_rel[2 * n] = src[0];
            if (src[1] > abs_rel[2 * n + 1]) abs_rel[2 * n + 1] = src[1];
        }
        return 0;
    };
    auto t0 = std::chrono::steady_clock::now();
    for (auto& r : s->ranks)
        if (enqueue_layer0_rank(s, r)) return -1;
    if (staged && wave3d_solver_exchange_local(s, 0)) return -1;
    for (int n = 1; n <= T; n++) {
        for (auto& r : s->ranks)
            if (enqueue_step_rank(s, r, n, pinned_desc + 2 * n)) return -1;
        if (staged && n < T && wave3d_solver_exchange_local(s, n)) return -1;
        for (auto& r : s->ranks) {
            if (use_device(r.device)) return -1;
            W3D_TRY(cudaMemcpyAsync(s->pinned + 2 * ((size_t)n * L + r.local), r.err + 2 * n, 16, cudaMemcpyDeviceToHost,
                                    stream_of(s, r)));
        }
        for (size_t g = 0; g < s->groups.size(); g++) {
            if (use_device(s->groups[g].device)) return -1;
            W3D_TRY(cudaEventRecord(s->sev[(size_t)n * s->groups.size() + g], s->groups[g].stream));
        }
        if (n - lag >= 1 && collect(n - lag)) return -1;
    }
    for (int n = std::max(1, T - lag + 1); n <= T; n++)
        if (collect(n)) return -1;
    if (wave3d_solver_sync(s)) return -1;
    auto t1 = std::chrono::steady_clock::now();
    if (t) {
        memset(t, 0, sizeof *t);
        t->solve_ms = t->loop_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    }
    return wave3d_solver_status(s) ? -2 : 0;
}

// Post-pass error of the stored layer n (it must still be resident: n or n-1 of the last computed layer),
// the shape of the reference's first-generation calculate_error() (ref: mpi_sol.cpp:340-372): evaluated
// by the stand-alone two-stage reduction, maximum over the local ranks.
int wave3d_solver_field_error(wave3d_solver* s, int n, double* abs_rel2) {
    if (!s || n < 0 || n > s->prob.timesteps || !abs_rel2) return -1;
    abs_rel2[0] = abs_rel2[1] = 0.0;
    if (s->opts.backend != WAVE3D_BACKEND_CUDA) {
        for (auto& r : s->ranks) {
            double a = 0.0, b = 0.0;
            wave3d_cpu_field_error(&s->prob, &r.box, &r.lay, r.hfield[n & 1].data(), n, &a, &b, s->opts.threads);
            abs_rel2[0] = std::max(abs_rel2[0], a);
            abs_rel2[1] = std::max(abs_rel2[1], b);
        }
        return 0;
    }
    if (wave3d_solver_sync(s)) return -1;
    for (auto& r : s->ranks) {
        if (use_device(r.device)) return -1;
        cudaStream_t st = stream_of(s, r);
        double* out = r.desc;  // two doubles of device scratch; the descriptor table is idle between solves
        const double* tabs[6] = {r.tabs[0], r.tabs[1], r.tabs[2], r.tabs[3], r.tabs[4], r.tabs[5]};
        if (wave3d_cuda_field_error(&s->prob, &r.box, &r.lay, tabs, r.field[n & 1], n, r.ferr_scratch, out, st)) return -1;
        s->launches++;
        W3D_TRY(cudaMemcpyAsync(s->pinned + 2 * r.local, out, 16, cudaMemcpyDeviceToHost, st));
    }
    if (wave3d_solver_sync(s)) return -1;
    for (auto& r : s->ranks) {
        abs_rel2[0] = std::max(abs_rel2[0], s->pinned[2 * r.local]);
        abs_rel2[1] = std::max(abs_rel2[1], s->pinned[2 * r.local + 1]);
    }
    return 0;
}

int64_t wave3d_solver_kernel_launches(const wave3d_solver* s) { return s ? s->launches : 0; }

int wave3d_solver_multistep_active(const wave3d_solver* s) { return s && multistep_ok(s) ? 1 : 0; }

int wave3d_solver_status(wave3d_solver* s) {
    if (!s || s->opts.backend != WAVE3D_BACKEND_CUDA) return 0;
    int st = 0;
    if (wave3d_solver_sync(s)) return -1;
    for (auto& r : s->ranks) {
        uint32_t v[2] = {0, 0};
        if (use_device(r.device)) return -1;
        if (cudaMemcpy(v, r.scratch, 8, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
        if (v[1]) st = 1;
    }
    return st;
}

// Enqueue the whole solve for one device group, layer by layer, ranks interleaved so
// that ranks sharing a device never wait on a kernel queued behind them.
static int timed_exchange(wave3d_solver* s, int n) {
    // staged transport: bracket every exchange with an event pair on the first device's stream; the
    // pairs are read after the solve, so nothing here blocks the host (ref: total_exchange_time /
    // total_mpi_exchange_time, cuda_sol.cpp:235-243,291-304)
    Group& g = s->groups[0];
    if (use_device(g.device)) return -1;
    while (s->xev.size() < 2 * (size_t)(s->xev_used + 1)) {
        cudaEvent_t e;
        W3D_TRY(cudaEventCreate(&e));
        s->xev.push_back(e);
    }
    W3D_TRY(cudaEventRecord(s->xev[2 * s->xev_used], g.stream));
    if (wave3d_solver_exchange_local(s, n)) return -1;
    if (use_device(g.device)) return -1;
    W3D_TRY(cudaEventRecord(s->xev[2 * s->xev_used + 1], g.stream));
    s->xev_used++;
    return 0;
}

static int enqueue_all(wave3d_solver* s, Group* only) {
    const bool staged = s->opts.transport == WAVE3D_TRANSPORT_STAGED;
    auto members = [&](auto fn) -> int {
        for (auto& r : s->ranks) {
            if (only && &s->groups[r.group] != only) continue;
            if (fn(r)) return -1;
        }
        return 0;
    };
    if (members([&](Rank& r) { return enqueue_layer0_rank(s, r); })) return -1;
    if (staged && timed_exchange(s, 0)) return -1;
    const bool multi = multistep_ok(s);
    for (int n = 1; n <= s->prob.timesteps; n++) {
        if (multi && n == 2) {
            // one launch per 2048 layers keeps the item counter far from 2^31 on very long runs
            const int chunk = multi_chunk(s->prob.timesteps);
            for (int n0 = 2; n0 <= s->prob.timesteps; n0 += chunk) {
                const int ns = std::min(chunk, s->prob.timesteps - n0 + 1);
                if (members([&](Rank& r) { return enqueue_multi_rank(s, r, n0, ns); })) return -1;
            }
            break;
        }
        if (members([&](Rank& r) { return enqueue_step_rank(s, r, n, nullptr); })) return -1;
        if (staged && n < s->prob.timesteps && timed_exchange(s, n)) return -1;
    }
    return 0;
}

int wave3d_solver_run(wave3d_solver* s, wave3d_timings* t) {
    if (t) memset(t, 0, sizeof *t);
    // Across processes the caller resets every rank and then barriers (a neighbour's
    // late flag must not be wiped, an early one must not be lost); otherwise reset here.
    if (!s->fresh && wave3d_solver_reset(s)) return -1;
    s->fresh = false;
    s->exchange_ms = 0.0;
    s->xev_used = 0;
    if (s->opts.backend != WAVE3D_BACKEND_CUDA) {
        if (!s->all_local) return -1;
        auto t0 = std::chrono::steady_clock::now();
        wave3d_solver_enqueue_layer0(s);
        wave3d_solver_exchange_local(s, 0);
        for (int n = 1; n <= s->prob.timesteps; n++) {
            wave3d_solver_enqueue_step(s, n);
            if (n < s->prob.timesteps) wave3d_solver_exchange_local(s, n);
        }
        auto t1 = std::chrono::steady_clock::now();
        if (t) t->solve_ms = t->loop_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
        return 0;
    }
    if (s->opts.transport == WAVE3D_TRANSPORT_PEER && !s->connected && s->opts.nranks > 1) {
        set_last_error("peer transport: ranks are not connected (call wave3d_solver_connect)");
        return -1;
    }
    const bool staged = s->opts.transport == WAVE3D_TRANSPORT_STAGED;
    if (staged && !s->all_local) {
        set_last_error("staged transport across processes is driven by the caller (pack/unpack hooks)");
        return -1;
    }
    const bool per_step = s->opts.per_step_sync != 0;
    // the multi-layer path is a handful of launches: nothing for a graph to save
    const bool graph = s->opts.use_graph && !per_step && !staged && !multistep_ok(s) && !s->noerr && !s->eager;
    if (graph) {
        for (auto& g : s->groups) {
            if (g.exec) continue;
            if (use_device(g.device)) return -1;
            cudaGraph_t gr = nullptr;
            W3D_TRY(cudaStreamBeginCapture(g.stream, cudaStreamCaptureModeThreadLocal));
            const int64_t before = s->launches;
            int rc = enqueue_all(s, &g);
            g.graph_launches = s->launches - before;
            s->launches = before;  // capture is not execution
            cudaError_t e = cudaStreamEndCapture(g.stream, &gr);
            if (rc || !check(e, "cudaStreamEndCapture")) return -1;
            W3D_TRY(cudaGraphInstantiate(&g.exec, gr, 0));
            cudaGraphDestroy(gr);
        }
    }
    for (auto& g : s->groups) {
        if (use_device(g.device)) return -1;
        W3D_TRY(cudaEventRecord(g.ev0, g.stream));
    }
    if (graph) {
        for (auto& g : s->groups) {
            if (use_device(g.device)) return -1;
            W3D_TRY(cudaGraphLaunch(g.exec, g.stream));
        }
        for (auto& g : s->groups) s->launches += g.graph_launches;
    } else if (per_step) {
        double e2[2];
        if (enqueue_all(s, nullptr) == 0) {
            // per-step mode is normally driven by the caller; here: run, then read every layer back
            for (int n = 1; n <= s->prob.timesteps; n++)
                if (wave3d_solver_layer_error(s, n, e2)) return -1;
        } else {
            return -1;
        }
    } else {
        if (enqueue_all(s, nullptr)) return -1;
    }
    for (auto& g : s->groups) {
        if (use_device(g.device)) return -1;
        W3D_TRY(cudaEventRecord(g.ev1, g.stream));
    }
    double solve = 0.0;
    for (auto& g : s->groups) {
        if (use_device(g.device)) return -1;
        W3D_TRY(cudaEventSynchronize(g.ev1));
        float ms = 0.f;
        W3D_TRY(cudaEventElapsedTime(&ms, g.ev0, g.ev1));
        if (ms > solve) solve = ms;  // max over the devices driven by this process
    }
    if (s->xev_used) use_device(s->groups[0].device);
    for (int i = 0; i < s->xev_used; i++) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, s->xev[2 * i], s->xev[2 * i + 1]) == cudaSuccess) s->exchange_ms += ms;
    }
    if (!staged && s->connected) {
        // Fused peer path: the faces travel inside the update kernel; what remains of an "exchange" is the
        // time CTAs had nothing to load because a neighbour rank's flag had not arrived. Every CTA adds its
        // share (ns) to scratch words 4-5; reported: mean over the CTAs, max over the local ranks.
        for (auto& r : s->ranks) {
            unsigned long long ns = 0;
            if (use_device(r.device)) return -1;
            W3D_TRY(cudaMemcpy(&ns, r.scratch + 4, 8, cudaMemcpyDeviceToHost));
            s->exchange_ms = std::max(s->exchange_ms, (double)ns * 1e-6 / std::max(1, r.plan.ctas));
        }
    }
    if (t) {
        t->solve_ms = solve;
        t->exchange_ms = s->exchange_ms;
        t->loop_ms = solve - s->exchange_ms;
    }
    if (wave3d_solver_status(s) != 0) {
        set_last_error("in-kernel halo wait timed out (a neighbour rank did not publish its layer)");
        return -2;
    }
    s->ran_once = true;
    return 0;
}

// The reference reports four phase totals (ref: cuda_sol.cpp:438-441). Here the phases are fused into one
// kernel, so they are attributed by measurement: the solve runs once with the error evaluation switched off
// (same launches, same traffic) and once normally; the difference is the cost of the fused error evaluation,
// the flag waits of the normal run are the exchange time, nothing is staged through the host.
int wave3d_solver_measure_phases(wave3d_solver* s, wave3d_timings* t) {
    if (!s || !t) return -1;
    wave3d_timings t0, t1;
    if (s->opts.backend == WAVE3D_BACKEND_CUDA) {
        // the A/B pair uses the same launch mode (eager), the solve that is reported the normal one
        s->eager = true;
        // a solver that has not run yet pays for module loading and cold caches in its first solve: keep that
        // out of the A/B pair
        int rc = s->ran_once ? 0 : wave3d_solver_run(s, &t0);
        if (rc) {
            s->eager = false;
            return rc;
        }
        s->noerr = true;
        rc = wave3d_solver_run(s, &t0);
        s->noerr = false;
        if (!rc) rc = wave3d_solver_run(s, &t1);
        s->eager = false;
        if (rc) return rc;
    }
    int rc = wave3d_solver_run(s, t);
    if (rc) return rc;
    if (s->opts.backend == WAVE3D_BACKEND_CUDA) {
        t->error_ms = std::max(0.0, t1.solve_ms - t0.solve_ms);
        t->loop_ms = std::max(0.0, t->solve_ms - t->exchange_ms - t->error_ms);
        t->hostdev_ms = 0.0;  // nothing is staged through the host on either CUDA transport
    }
    return 0;
}

// -------------------------------------------------------------- program entry

int wave3d_main(int argc, const char* const* argv) {
    char err[512] = {0};
    wave3d_problem p;
    double pi = WAVE3D_PI_FULL;
    if (const char* e = getenv("WAVE3D_PI"))
        if (!strcmp(e, "trunc")) pi = WAVE3D_PI_TRUNC;
    if (wave3d_parse_args(argc, argv, pi, &p, err, sizeof err)) {
        fprintf(stderr, "%s\n", err);
        return 2;
    }
    char line[128];
    wave3d_format_courant(&p, line, sizeof line);
    printf("%s\n", line);  // ref: cuda_sol.cpp:474
    fflush(stdout);

    wave3d_solver_opts o;
    wave3d_solver_opts_default(&o);
    const char* be = getenv("WAVE3D_BACKEND");
    const int ndev = wave3d_cuda_device_count();
    if (be ? !strcmp(be, "cpu") : ndev == 0) o.backend = WAVE3D_BACKEND_CPU;
    int nranks = 1;
    if (o.backend == WAVE3D_BACKEND_CUDA) nranks = p.Np > 0 ? p.Np : 1;  // Np = number of GPUs / ranks
    else o.threads = p.Np > 0 ? p.Np : 1;                               // Np = threads (ref: openmp_sol.cpp)
    if (const char* e = getenv("WAVE3D_RANKS")) nranks = atoi(e) > 0 ? atoi(e) : nranks;
    if (const char* e = getenv("WAVE3D_DIMS")) {
        if (wave3d_parse_dims(e, o.dims)) {
            fprintf(stderr, "cannot parse WAVE3D_DIMS='%s' (expected e.g. 8x1x1)\n", e);
            return 2;
        }
        nranks = o.dims[0] * o.dims[1] * o.dims[2];
    }
    if (const char* e = getenv("WAVE3D_TRANSPORT"))
        if (!strcmp(e, "staged")) o.transport = WAVE3D_TRANSPORT_STAGED;
    if (const char* e = getenv("WAVE3D_GRAPH")) o.use_graph = atoi(e);
    o.nranks = o.nlocal = nranks;
    o.first_rank = 0;
    if (o.backend == WAVE3D_BACKEND_CUDA && ndev < 1) {
        // WAVE3D_BACKEND=cuda forced on a box without a usable GPU
        fprintf(stderr, "wave3d: no CUDA device available (backend=cuda); set WAVE3D_BACKEND=cpu for the CPU backend\n");
        return 1;
    }
    if (o.backend == WAVE3D_BACKEND_CUDA) {
        char host[256] = "localhost";
        gethostname(host, sizeof host);
        // ref: cuda_sol.cpp:513 prints this line from every MPI rank; here all ranks live in this process
        for (int r = 0; r < nranks; r++)
            printf("Process %d local rank = %d local size = %d hostname = %s (GPU %d)\n", r, r, nranks, host, r % ndev);
        printf("number of CUDA devices on node %s: %d\n", host, ndev);  // ref: cuda_sol.cpp:531
    }
    auto t0 = std::chrono::steady_clock::now();
    wave3d_solver* s = wave3d_solver_create(&p, &o, err, sizeof err);
    if (!s) {
        fprintf(stderr, "wave3d: %s\n", err);
        return 1;
    }
    auto t1 = std::chrono::steady_clock::now();
    wave3d_timings tm;
    int rc = 0;
    const char* prog = getenv("WAVE3D_PROGRESS");
    if (prog && atoi(prog)) {
        // reference-style driver: one line per layer on stdout and the layer's maxima read back
        // before the next layer starts (ref: cuda_sol.cpp:338,385 "calculating layer n")
        memset(&tm, 0, sizeof tm);
        const bool staged = o.backend == WAVE3D_BACKEND_CPU || o.transport == WAVE3D_TRANSPORT_STAGED;
        auto w0 = std::chrono::steady_clock::now();
        rc = wave3d_solver_reset(s) || wave3d_solver_enqueue_layer0(s);
        if (!rc && staged) rc = wave3d_solver_exchange_local(s, 0);
        double e2[2];
        for (int n = 1; !rc && n <= p.timesteps; n++) {
            printf("calculating layer %d\n", n);
            fflush(stdout);
            rc = wave3d_solver_enqueue_step(s, n);
            if (!rc && staged && n < p.timesteps) rc = wave3d_solver_exchange_local(s, n);
            if (!rc) rc = wave3d_solver_layer_error(s, n, e2);
        }
        if (!rc) rc = wave3d_solver_sync(s);
        tm.solve_ms = tm.loop_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w0).count();
    } else {
        // ref: cuda_sol.cpp:338,385 prints "calculating layer n" before every layer. Here the host submits
        // the whole solve at once (one CUDA graph or one persistent launch per GPU) and the device works
        // through the layers on its own, so the lines mark submission, not completion; WAVE3D_PROGRESS=1
        // gives the reference's lock-step behaviour, WAVE3D_QUIET=1 drops the lines.
        const char* quiet = getenv("WAVE3D_QUIET");
        if (!(quiet && atoi(quiet))) {
            for (int n = 1; n <= p.timesteps; n++) printf("calculating layer %d\n", n);
            fflush(stdout);
        }
        // the four phase totals of the reference's report need a second, error-free solve (WAVE3D_PHASES=0
        // skips it: "total error calculation time" then reads 0)
        const char* ph = getenv("WAVE3D_PHASES");
        const bool phases = o.backend == WAVE3D_BACKEND_CUDA && !(ph && !atoi(ph)) && !getenv("WAVE3D_REPEAT");
        rc = phases ? wave3d_solver_measure_phases(s, &tm) : wave3d_solver_run(s, &tm);
        // WAVE3D_REPEAT=k: solve k times and report the last one (warm kernels, loaded modules) — for
        // A/B timing from the command line; the result is the same every time
        if (const char* e = getenv("WAVE3D_REPEAT"))
            for (int i = 1; !rc && i < atoi(e); i++) rc = wave3d_solver_run(s, &tm);
    }
    if (rc) {
        fprintf(stderr, "wave3d: solve failed: %s\n", wave3d_cuda_last_error());
        wave3d_solver_destroy(s);
        return 1;
    }
    tm.init_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    std::vector<double> ea(p.timesteps + 1), er(p.timesteps + 1);
    wave3d_solver_errors(s, ea.data(), er.data());
    const int cuda_style = o.backend == WAVE3D_BACKEND_CUDA;
    char name[256];
    wave3d_output_name(p.N, p.Np, 0, name, sizeof name);
    if (const char* e = getenv("WAVE3D_NAME_STYLE"))  // "cuda": the MPI+CUDA program's own name (ref: cuda_sol.cpp:535)
        if (!strcmp(e, "cuda")) wave3d_output_name(p.N, nranks, ndev > 0 ? ndev : 1, name, sizeof name);
    if (const char* e = getenv("WAVE3D_OUTPUT")) snprintf(name, sizeof name, "%s", e);
    std::vector<char> buf(wave3d_format_report(&p, &tm, ea.data(), er.data(), 0, cuda_style, nullptr, 0));
    wave3d_format_report(&p, &tm, ea.data(), er.data(), 0, cuda_style, buf.data(), buf.size());
    FILE* f = fopen(name, "w");
    if (!f) {
        fprintf(stderr, "wave3d: cannot write %s\n", name);
        wave3d_solver_destroy(s);
        return 1;
    }
    fputs(buf.data(), f);
    fclose(f);
    if (const char* e = getenv("WAVE3D_FULL_PRECISION")) {
        if (atoi(e)) {  // opt-in %.17g side file; the default file keeps the reference format
            std::string n2 = std::string(name) + ".full";
            std::vector<char> b2(wave3d_format_report(&p, &tm, ea.data(), er.data(), 17, cuda_style, nullptr, 0));
            wave3d_format_report(&p, &tm, ea.data(), er.data(), 17, cuda_style, b2.data(), b2.size());
            if (FILE* g = fopen(n2.c_str(), "w")) {
                fputs(b2.data(), g);
                fclose(g);
            }
        }
    }
    const double cells = (double)(p.N + 1) * (p.N + 1) * (p.N + 1) * p.timesteps;
    printf("solve: %.3f ms  (%.2f Gcell-updates/s, %d rank(s), grid %dx%dx%d)  -> %s\n", tm.solve_ms,
           cells / (tm.solve_ms * 1e-3) / 1e9, nranks, s->dims[0], s->dims[1], s->dims[2], name);
    wave3d_solver_destroy(s);
    return 0;
}

}  // extern "C"
