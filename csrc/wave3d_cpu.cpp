// wave3d_cpu.cpp — OpenMP CPU backend: oracle, CPU baseline and host twin of every
// device op. It restates the *semantics* of the reference's second-generation CPU
// solver (ref: mpi_new.cpp:157-179 boundary work, :271-372 time loop; hybrid_new.cpp
// for the OpenMP flavour) with the CUDA build's exact fp64 expression order
//  so that its maxima can be compared digit for digit with the GPU path.
// Differences on purpose: two in-place layers instead of three rotating ones, error
// maxima reduced per thread (no race), 64-bit indexing (Q7), the seam
// planes use the real leapfrog formula (Q1), walls are masked correctly (Q2).
//
// Built with -ffp-contract=off: the only fused operations are the explicit fma()s.
#include "wave3d.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

inline int64_t idx(const wave3d_layout* l, int i, int j, int k) {
    return (int64_t)(i + 1) * l->plane_stride + (int64_t)(j + 1) * l->pitch + (k + l->oz);
}

inline int nthreads(int threads) {
#ifdef _OPENMP
    return threads > 0 ? threads : omp_get_max_threads();
#else
    (void)threads;
    return 1;
#endif
}

// fmax semantics of the reference reduction: NaN never wins (ref: cuda_sol_kernels.cu:60-61,85-86)
inline void upd(double& m, double v) {
    if (v > m) m = v;
}

}  // namespace

extern "C" {

void wave3d_cpu_tables(const wave3d_problem* p, const wave3d_box* b, double* sx, double* sy, double* sz) {
    const double PI = p->pi;
    // arg_x = (((PI+PI)*hx)*g)/Lx, arg_y = ((PI*hy)*g)/Ly  — the parse tree of
    // `2*PI*hx*(i-1+x_0)/Lx` (ref: cuda_sol_kernels.cu:41)
    for (int i = 0; i < b->n[0]; i++) sx[i] = std::sin((((PI + PI) * p->hx) * (double)(i + b->o[0])) / p->Lx);
    for (int j = 0; j < b->n[1]; j++) sy[j] = std::sin(((PI * p->hy) * (double)(j + b->o[1])) / p->Ly);
    for (int k = 0; k < b->n[2]; k++) sz[k] = std::sin(((PI * p->hz) * (double)(k + b->o[2])) / p->Lz);
}

double wave3d_cpu_coslayer(const wave3d_problem* p, int n) {
    return std::cos(std::fma(p->a_t * (double)n, p->tau, p->pi + p->pi));
}

void wave3d_cpu_layer0(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l, double* u0, int threads) {
    std::vector<double> sx(b->n[0]), sy(b->n[1]), sz(b->n[2]);
    wave3d_cpu_tables(p, b, sx.data(), sy.data(), sz.data());
    const double c = wave3d_cpu_coslayer(p, 0);
    (void)threads;
#pragma omp parallel for num_threads(nthreads(threads)) schedule(static)
    for (int i = 0; i < b->n[0]; i++) {
        const double cs = c * sx[i];
        for (int j = 0; j < b->n[1]; j++) {
            const double csy = cs * sy[j];
            double* row = u0 + idx(l, i, j, 0);
            for (int k = 0; k < b->n[2]; k++) row[k] = csy * sz[k];
        }
    }
}

void wave3d_cpu_step(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l, const double* u1,
                     double* u_io, int n, int first, double* max_abs, double* max_rel, int threads) {
    std::vector<double> sx(b->n[0]), sy(b->n[1]), sz(b->n[2]);
    wave3d_cpu_tables(p, b, sx.data(), sy.data(), sz.data());
    const double c = wave3d_cpu_coslayer(p, n);
    const double hx2 = p->hx * p->hx, hy2 = p->hy * p->hy, hz2 = p->hz * p->hz;
    double coef = (p->a2 * p->tau) * p->tau;
    if (first) coef = coef * 0.5;
    const int64_t ps = l->plane_stride, pz = l->pitch;
    const int nt = nthreads(threads);
    std::vector<double> tabs(nt, 0.0), trel(nt, 0.0);
#pragma omp parallel num_threads(nt)
    {
        double mabs = 0.0, mrel = 0.0;
#pragma omp for schedule(static)
        for (int i = b->lo[0]; i <= b->hi[0]; i++) {
            const bool ei = i >= b->elo[0] && i <= b->ehi[0];
            const double cs = c * sx[i];
            for (int j = b->lo[1]; j <= b->hi[1]; j++) {
                const double csy = cs * sy[j];
                const double* g1 = u1 + idx(l, i, j, 0);
                double* g0 = u_io + idx(l, i, j, 0);
                for (int k = b->lo[2]; k <= b->hi[2]; k++) {
                    const double uc = g1[k];
                    const double c2 = uc + uc;
                    const double s = first ? uc : c2 - g0[k];
                    const double lx = ((g1[k + ps] + g1[k - ps]) - c2) / hx2;
                    const double ly = ((g1[k + pz] + g1[k - pz]) - c2) / hy2;
                    const double lz = ((g1[k + 1] + g1[k - 1]) - c2) / hz2;
                    const double u = std::fma(coef, (lx + ly) + lz, s);
                    g0[k] = u;
                    if (ei) {  // y/z update range == y/z error range
                        const double f = csy * sz[k];
                        const double a = std::fabs(u - f);
                        upd(mabs, a);
                        upd(mrel, a / std::fabs(f));
                    }
                }
            }
        }
#ifdef _OPENMP
        const int t = omp_get_thread_num();
#else
        const int t = 0;
#endif
        tabs[t] = mabs;
        trel[t] = mrel;
    }
    double a = 0.0, r = 0.0;
    for (int t = 0; t < nt; t++) {
        upd(a, tabs[t]);
        upd(r, trel[t]);
    }
    if (max_abs) *max_abs = a;
    if (max_rel) *max_rel = r;
}

void wave3d_cpu_zero_walls(const wave3d_box* b, const wave3d_layout* l, double* u) {
    const int X = b->n[0], Y = b->n[1], Z = b->n[2];
    for (int i = 0; i < X; i++) {
        if (b->coords[1] == 0)
            for (int k = 0; k < Z; k++) u[idx(l, i, 0, k)] = 0.0;
        if (b->coords[1] == b->dims[1] - 1)
            for (int k = 0; k < Z; k++) u[idx(l, i, Y - 1, k)] = 0.0;
        if (b->coords[2] == 0)
            for (int j = 0; j < Y; j++) u[idx(l, i, j, 0)] = 0.0;
        if (b->coords[2] == b->dims[2] - 1)
            for (int j = 0; j < Y; j++) u[idx(l, i, j, Z - 1)] = 0.0;
    }
}

void wave3d_cpu_pack_face(const wave3d_box* b, const wave3d_layout* l, const double* u, int face, double* buf) {
    const int X = b->n[0], Y = b->n[1], Z = b->n[2];
    const int axis = face / 2, side = face % 2;
    if (axis == 0) {
        const int i = b->send_x[side];
        for (int j = 0; j < Y; j++) memcpy(buf + (int64_t)j * Z, u + idx(l, i, j, 0), sizeof(double) * Z);
    } else if (axis == 1) {
        const int j = side ? Y - 1 : 0;
        for (int i = 0; i < X; i++) memcpy(buf + (int64_t)i * Z, u + idx(l, i, j, 0), sizeof(double) * Z);
    } else {
        const int k = side ? Z - 1 : 0;
        for (int i = 0; i < X; i++)
            for (int j = 0; j < Y; j++) buf[(int64_t)i * Y + j] = u[idx(l, i, j, k)];
    }
}

void wave3d_cpu_unpack_face(const wave3d_box* b, const wave3d_layout* l, double* u, int face, const double* buf) {
    const int X = b->n[0], Y = b->n[1], Z = b->n[2];
    const int axis = face / 2, side = face % 2;
    if (axis == 0) {
        const int i = side ? X : -1;
        for (int j = 0; j < Y; j++) memcpy(u + idx(l, i, j, 0), buf + (int64_t)j * Z, sizeof(double) * Z);
    } else if (axis == 1) {
        const int j = side ? Y : -1;
        for (int i = 0; i < X; i++) memcpy(u + idx(l, i, j, 0), buf + (int64_t)i * Z, sizeof(double) * Z);
    } else {
        const int k = side ? Z : -1;
        for (int i = 0; i < X; i++)
            for (int j = 0; j < Y; j++) u[idx(l, i, j, k)] = buf[(int64_t)i * Y + j];
    }
}

void wave3d_cpu_field_error(const wave3d_problem* p, const wave3d_box* b, const wave3d_layout* l, const double* u,
                            int n, double* max_abs, double* max_rel, int threads) {
    std::vector<double> sx(b->n[0]), sy(b->n[1]), sz(b->n[2]);
    wave3d_cpu_tables(p, b, sx.data(), sy.data(), sz.data());
    const double c = wave3d_cpu_coslayer(p, n);
    double mabs = 0.0, mrel = 0.0;
    (void)threads;
    for (int i = b->elo[0]; i <= b->ehi[0]; i++) {
        const double cs = c * sx[i];
        for (int j = b->elo[1]; j <= b->ehi[1]; j++) {
            const double csy = cs * sy[j];
            const double* row = u + idx(l, i, j, 0);
            for (int k = b->elo[2]; k <= b->ehi[2]; k++) {
                const double f = csy * sz[k];
                const double a = std::fabs(row[k] - f);
                upd(mabs, a);
                upd(mrel, a / std::fabs(f));
            }
        }
    }
    *max_abs = mabs;
    *max_rel = mrel;
}

int wave3d_cpu_solve(const wave3d_problem* p, const int dims_in[3], int threads, double* max_abs, double* max_rel,
                     wave3d_timings* t, char* err, size_t errlen) {
    int dims[3] = {dims_in ? dims_in[0] : 1, dims_in ? dims_in[1] : 1, dims_in ? dims_in[2] : 1};
    const int P = dims[0] * dims[1] * dims[2];
    std::vector<wave3d_box> box(P);
    std::vector<wave3d_layout> lay(P);
    std::vector<std::vector<double>> f0(P), f1(P);
    auto t_init0 = std::chrono::steady_clock::now();
    for (int r = 0; r < P; r++) {
        if (wave3d_box_init(p->N, dims, r, &box[r], err, errlen)) return -1;
        wave3d_layout_init(&box[r], &lay[r]);
        f0[r].assign(lay[r].elems, 0.0);
        f1[r].assign(lay[r].elems, 0.0);
    }
    auto field = [&](int r, int n) -> double* { return (n & 1) ? f1[r].data() : f0[r].data(); };
    std::vector<double> buf;
    auto exchange = [&](int n) {
        for (int r = 0; r < P; r++) {
            int64_t fe[6];
            wave3d_face_elems(&box[r], fe);
            for (int f = 0; f < 6; f++) {
                const int q = box[r].nbr[f];
                if (q < 0) continue;
                buf.resize(fe[f]);
                wave3d_cpu_pack_face(&box[r], &lay[r], field(r, n), f, buf.data());
                wave3d_cpu_unpack_face(&box[q], &lay[q], field(q, n), f ^ 1, buf.data());
            }
        }
    };
    for (int n = 0; n <= p->timesteps; n++) max_abs[n] = max_rel[n] = 0.0;
    auto t0 = std::chrono::steady_clock::now();
    double loop_ms = 0.0, exch_ms = 0.0;
    auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double, std::milli>(b - a).count();
    };
    {
        auto a = std::chrono::steady_clock::now();
        for (int r = 0; r < P; r++) wave3d_cpu_layer0(p, &box[r], &lay[r], field(r, 0), threads);
        auto c = std::chrono::steady_clock::now();
        exchange(0);
        auto d = std::chrono::steady_clock::now();
        loop_ms += ms(a, c);
        exch_ms += ms(c, d);
    }
    for (int n = 1; n <= p->timesteps; n++) {
        auto a = std::chrono::steady_clock::now();
        for (int r = 0; r < P; r++) {
            double ea, er;
            wave3d_cpu_step(p, &box[r], &lay[r], field(r, n - 1), field(r, n), n, n == 1, &ea, &er, threads);
            upd(max_abs[n], ea);
            upd(max_rel[n], er);
        }
        if (n == 1)  // layer 0 carried analytic (non-zero) wall values; layer 2 overwrites it in place
            for (int r = 0; r < P; r++) wave3d_cpu_zero_walls(&box[r], &lay[r], field(r, 0));
        auto c = std::chrono::steady_clock::now();
        if (n < p->timesteps) exchange(n);
        auto d = std::chrono::steady_clock::now();
        loop_ms += ms(a, c);
        exch_ms += ms(c, d);
    }
    auto t1 = std::chrono::steady_clock::now();
    if (t) {
        memset(t, 0, sizeof(*t));
        t->init_ms = ms(t_init0, t0);
        t->solve_ms = ms(t0, t1);
        t->loop_ms = loop_ms;
        t->exchange_ms = exch_ms;
    }
    return 0;
}

// Independent cross-check in the shape of ref openmp_sol.cpp: every layer stored,
// no ghosts (explicit periodic index), per-cell trig, error post-pass. O(timesteps*N^3)
// memory — small N only.
int wave3d_cpu_solve_naive(const wave3d_problem* p, double* max_abs, double* max_rel) {
    const int N = p->N, M = N + 1, S = p->timesteps;
    const double PI = p->pi;
    const int64_t vol = (int64_t)M * M * M;
    if ((double)vol * (S + 1) * 8 > 6e9) return -1;
    std::vector<double> u((size_t)vol * (S + 1), 0.0);
    auto at = [&](int n, int i, int j, int k) -> double& { return u[(size_t)n * vol + ((int64_t)i * M + j) * M + k]; };
    auto an = [&](int n, int i, int j, int k) {
        const double ct = std::cos(std::fma(p->a_t * (double)n, p->tau, PI + PI));
        const double sxv = std::sin((((PI + PI) * p->hx) * (double)i) / p->Lx);
        const double syv = std::sin(((PI * p->hy) * (double)j) / p->Ly);
        const double szv = std::sin(((PI * p->hz) * (double)k) / p->Lz);
        return ((ct * sxv) * syv) * szv;
    };
    for (int i = 0; i < M; i++)
        for (int j = 0; j < M; j++)
            for (int k = 0; k < M; k++) at(0, i, j, k) = an(0, i, j, k);
    const double hx2 = p->hx * p->hx, hy2 = p->hy * p->hy, hz2 = p->hz * p->hz;
    for (int n = 1; n <= S; n++) {
        double coef = (p->a2 * p->tau) * p->tau;
        if (n == 1) coef = coef * 0.5;
        for (int i = 0; i < M; i++) {
            // periodic in x with the duplicated seam point: neighbours of plane 0 are N-1 and 1,
            // of plane N are N-1 and 1 (ref: openmp_sol.cpp:117-118)
            const int im = i == 0 ? N - 1 : i - 1;
            const int ip = i == N ? 1 : i + 1;
            for (int j = 1; j < N; j++)
                for (int k = 1; k < N; k++) {
                    const double uc = at(n - 1, i, j, k);
                    const double c2 = uc + uc;
                    const double s = n == 1 ? uc : c2 - at(n - 2, i, j, k);
                    const double lx = ((at(n - 1, ip, j, k) + at(n - 1, im, j, k)) - c2) / hx2;
                    const double ly = ((at(n - 1, i, j + 1, k) + at(n - 1, i, j - 1, k)) - c2) / hy2;
                    const double lz = ((at(n - 1, i, j, k + 1) + at(n - 1, i, j, k - 1)) - c2) / hz2;
                    at(n, i, j, k) = std::fma(coef, (lx + ly) + lz, s);
                }
        }
    }
    for (int n = 0; n <= S; n++) {
        double a = 0.0, r = 0.0;
        if (n > 0)
            for (int i = 1; i < N; i++)
                for (int j = 1; j < N; j++)
                    for (int k = 1; k < N; k++) {
                        const double f = an(n, i, j, k);
                        const double e = std::fabs(at(n, i, j, k) - f);
                        upd(a, e);
                        upd(r, e / std::fabs(f));
                    }
        max_abs[n] = a;
        max_rel[n] = r;
    }
    return 0;
}

}  // extern "C"
