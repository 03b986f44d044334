// wave3d_spec.cpp — problem definition, CLI contract, block decomposition, storage
// layout and the rank-0 report. Pure host C++, no CUDA.
//
// Behavioural reference (not code): cuda_sol.cpp:447-519 (CLI, constants, topology,
// extents), cuda_sol.cpp:426-442 and 533-536 (report + file name), SURVEY §2.6.
#include "wave3d.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace {

void set_err(char* err, size_t errlen, const char* fmt, ...) {
    if (!err || errlen == 0) return;
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err, errlen, fmt, ap);
    va_end(ap);
}

bool parse_int(const char* s, int* out) {
    char* end = nullptr;
    long v = strtol(s, &end, 10);
    if (end == s) return false;
    // std::stoi (ref: cuda_sol.cpp:453) accepts trailing junk; so do we.
    *out = (int)v;
    return true;
}

bool parse_len(const char* s, double pi, double* out) {
    if (strcmp(s, "pi") == 0) {  // ref: cuda_sol.cpp:455-460
        *out = pi;
        return true;
    }
    char* end = nullptr;
    double v = strtod(s, &end);
    if (end == s) return false;
    *out = v;
    return true;
}

}  // namespace

extern "C" {

int wave3d_problem_init(wave3d_problem* p, int N, int Np, double Lx, double Ly, double Lz, double T, int timesteps,
                        double pi, char* err, size_t errlen) {
    if (!p) return -1;
    memset(p, 0, sizeof(*p));
    if (N < 3) {
        set_err(err, errlen, "N must be >= 3 (got %d)", N);
        return -1;
    }
    if (timesteps < 1) {
        set_err(err, errlen, "timesteps must be >= 1 (got %d)", timesteps);
        return -1;
    }
    if (!(Lx > 0) || !(Ly > 0) || !(Lz > 0)) {
        set_err(err, errlen, "box lengths must be positive");
        return -1;
    }
    if (!(pi > 3.0 && pi < 3.3)) pi = WAVE3D_PI_FULL;
    p->N = N;
    p->Np = Np;
    p->timesteps = timesteps;
    p->Lx = Lx;
    p->Ly = Ly;
    p->Lz = Lz;
    p->T = T;
    p->pi = pi;
    // Same expression shapes as the reference host code (cuda_sol.cpp:467-471).
    p->a2 = 1 / (4 * pi * pi);
    p->a_t = 0.5 * std::sqrt(4 / (Lx * Lx) + 1 / (Ly * Ly) + 1 / (Lz * Lz));
    p->tau = T / timesteps;
    p->hx = Lx / N;
    p->hy = Ly / N;
    p->hz = Lz / N;
    p->courant = std::sqrt(p->a2) * p->tau / std::min(p->hx, std::min(p->hy, p->hz));
    return 0;
}

int wave3d_parse_args(int argc, const char* const* argv, double pi, wave3d_problem* p, char* err, size_t errlen) {
    if (argc < 6) {
        set_err(err, errlen, "usage: %s N Np Lx Ly Lz [T] [timesteps]   (lengths accept the literal 'pi')",
                argc > 0 ? argv[0] : "wave3d");
        return -1;
    }
    if (!(pi > 3.0 && pi < 3.3)) pi = WAVE3D_PI_FULL;
    int N = 0, Np = 0, timesteps = 20;
    double L[3], T = 1.0;
    if (!parse_int(argv[1], &N) || !parse_int(argv[2], &Np)) {
        set_err(err, errlen, "N and Np must be integers");
        return -1;
    }
    for (int a = 0; a < 3; a++) {
        if (!parse_len(argv[3 + a], pi, &L[a])) {
            set_err(err, errlen, "cannot parse length '%s'", argv[3 + a]);
            return -1;
        }
    }
    if (argc >= 7) {
        char* end = nullptr;
        T = strtod(argv[6], &end);
        if (end == argv[6]) {
            set_err(err, errlen, "cannot parse T '%s'", argv[6]);
            return -1;
        }
    }
    if (argc >= 8 && !parse_int(argv[7], &timesteps)) {
        set_err(err, errlen, "cannot parse timesteps '%s'", argv[7]);
        return -1;
    }
    return wave3d_problem_init(p, N, Np, L[0], L[1], L[2], T, timesteps, pi, err, errlen);
}

int wave3d_format_double(double v, char* buf, size_t buflen) { return snprintf(buf, buflen, "%g", v); }

int wave3d_format_courant(const wave3d_problem* p, char* buf, size_t buflen) {
    return snprintf(buf, buflen, "C = %g", p->courant);
}

// ---------------------------------------------------------------- decomposition

void wave3d_dims_create(int nranks, int dims[3]) {
    dims[0] = dims[1] = dims[2] = 1;
    if (nranks < 1) return;
    std::vector<int> primes;
    int m = nranks;
    for (int f = 2; (long)f * f <= m; f++)
        while (m % f == 0) {
            primes.push_back(f);
            m /= f;
        }
    if (m > 1) primes.push_back(m);
    // largest factor first, always onto the currently smallest dimension
    std::sort(primes.begin(), primes.end(), [](int a, int b) { return a > b; });
    for (int f : primes) {
        int s = 0;
        for (int a = 1; a < 3; a++)
            if (dims[a] < dims[s]) s = a;
        dims[s] *= f;
    }
    std::sort(dims, dims + 3, [](int a, int b) { return a > b; });
}

int wave3d_parse_dims(const char* s, int dims[3]) {
    if (!s) return -1;
    int v[3] = {0, 0, 0};
    int n = 0;
    const char* c = s;
    while (*c && n < 3) {
        char* end = nullptr;
        long x = strtol(c, &end, 10);
        if (end == c || x < 1) return -1;
        v[n++] = (int)x;
        c = end;
        if (*c == 'x' || *c == 'X' || *c == ',' || *c == '*') c++;
        else if (*c) return -1;
    }
    if (n != 3 || *c) return -1;
    dims[0] = v[0];
    dims[1] = v[1];
    dims[2] = v[2];
    return 0;
}

int wave3d_rank_of(const int dims[3], const int coords[3]) {
    return (coords[0] * dims[1] + coords[1]) * dims[2] + coords[2];
}

void wave3d_coords_of(const int dims[3], int rank, int coords[3]) {
    coords[2] = rank % dims[2];
    coords[1] = (rank / dims[2]) % dims[1];
    coords[0] = rank / (dims[2] * dims[1]);
}

int wave3d_box_init(int N, const int dims[3], int rank, wave3d_box* b, char* err, size_t errlen) {
    if (!b) return -1;
    memset(b, 0, sizeof(*b));
    const int nranks = dims[0] * dims[1] * dims[2];
    if (dims[0] < 1 || dims[1] < 1 || dims[2] < 1 || rank < 0 || rank >= nranks) {
        set_err(err, errlen, "bad process grid %dx%dx%d / rank %d", dims[0], dims[1], dims[2], rank);
        return -1;
    }
    b->rank = rank;
    b->nranks = nranks;
    b->N = N;
    wave3d_coords_of(dims, rank, b->coords);
    for (int a = 0; a < 3; a++) {
        b->dims[a] = dims[a];
        const int base = (N + 1) / dims[a];
        b->n[a] = base;
        b->o[a] = b->coords[a] * base;  // origin uses the un-enlarged extent (ref: cuda_sol.cpp:486)
        if (b->coords[a] == dims[a] - 1) b->n[a] += (N + 1) % dims[a];
        // Every rank must own at least two points per axis so that a face and the
        // seam plane next to it are distinct (the reference silently misbehaves).
        if (base < 2) {
            set_err(err, errlen, "axis %d: %d points over %d ranks leaves fewer than 2 per rank", a, N + 1, dims[a]);
            return -1;
        }
    }
    // neighbours: x periodic, y/z open (ref: periods={1,0,0}, cuda_sol.cpp:479)
    for (int a = 0; a < 3; a++) {
        for (int s = 0; s < 2; s++) {
            int c[3] = {b->coords[0], b->coords[1], b->coords[2]};
            c[a] += s ? 1 : -1;
            int r = -1;
            if (a == 0) {
                c[0] = (c[0] + dims[0]) % dims[0];
                r = wave3d_rank_of(dims, c);
            } else if (c[a] >= 0 && c[a] < dims[a]) {
                r = wave3d_rank_of(dims, c);
            }
            b->nbr[2 * a + s] = r;
        }
    }
    for (int a = 0; a < 3; a++) {
        const bool first = b->coords[a] == 0, last = b->coords[a] == dims[a] - 1;
        b->elo[a] = first ? 1 : 0;
        b->ehi[a] = last ? b->n[a] - 2 : b->n[a] - 1;
        if (a == 0) {  // both periodic seam planes (global 0 and N) are advanced
            b->lo[a] = 0;
            b->hi[a] = b->n[a] - 1;
        } else {
            b->lo[a] = b->elo[a];
            b->hi[a] = b->ehi[a];
        }
    }
    // Seam rule: global planes 0 and N are the same physical point, so the wrap-around
    // sends global 1 downward and global N-1 upward (ref: cuda_sol.cpp:232).
    b->send_x[0] = b->coords[0] == 0 ? 1 : 0;
    b->send_x[1] = b->coords[0] == dims[0] - 1 ? b->n[0] - 2 : b->n[0] - 1;
    return 0;
}

void wave3d_layout_init(const wave3d_box* b, wave3d_layout* l) {
    memset(l, 0, sizeof(*l));
    // first updated z cell (owned index lo[2]) sits on storage column 16 => 128-byte aligned
    l->oz = 16 - b->lo[2];
    int64_t need = (int64_t)l->oz + b->n[2] + 1;  // up to and including the z+ ghost column
    l->pitch = (need + 15) / 16 * 16;
    if (const char* e = getenv("WAVE3D_PITCH_PAD")) l->pitch += 16 * (int64_t)atoi(e);  // tuning: extra 128-byte lines per row
    l->rows = (int64_t)b->n[1] + 2;
    l->planes = (int64_t)b->n[0] + 2;
    l->plane_stride = l->rows * l->pitch;
    l->elems = l->planes * l->plane_stride;
}

void wave3d_face_elems(const wave3d_box* b, int64_t elems[6]) {
    elems[0] = elems[1] = (int64_t)b->n[1] * b->n[2];
    elems[2] = elems[3] = (int64_t)b->n[0] * b->n[2];
    elems[4] = elems[5] = (int64_t)b->n[0] * b->n[1];
}

// ---------------------------------------------------------------------- report

int wave3d_output_name(int N, int np, int ng, char* buf, size_t buflen) {
    if (ng > 0) return snprintf(buf, buflen, "output_N%d_Np%d_Ng%d_cuda.txt", N, np, ng);
    return snprintf(buf, buflen, "output_N%d_Np%d.txt", N, np);
}

size_t wave3d_format_report(const wave3d_problem* p, const wave3d_timings* t, const double* max_abs,
                            const double* max_rel, int precision, int cuda_style, char* buf, size_t buflen) {
    std::string out;
    char line[256];
    const int prec = precision > 0 ? precision : 6;
    if (cuda_style) {
        snprintf(line, sizeof line, "initialization done in %ums\n", (unsigned)(t ? t->init_ms : 0.0));
        out += line;
        // the reference prints a float here; float -> 6 significant digits
        snprintf(line, sizeof line, "numerical solution calculated in %gms\n", (double)(float)(t ? t->solve_ms : 0.0));
        out += line;
    } else {
        snprintf(line, sizeof line, "numerical solution calculated in %ums\n", (unsigned)(t ? t->solve_ms : 0.0));
        out += line;
    }
    for (int n = 0; n <= p->timesteps; n++) {
        snprintf(line, sizeof line, "max abs and rel errors on layer %d: %.*g %.*g\n", n, prec, max_abs[n], prec,
                 max_rel[n]);
        out += line;
    }
    if (cuda_style) {
        snprintf(line, sizeof line, "total host-device exchange time: %g ms\n", (double)(float)(t ? t->hostdev_ms : 0.0));
        out += line;
        snprintf(line, sizeof line, "total loop time: %g ms\n", (double)(float)(t ? t->loop_ms : 0.0));
        out += line;
        snprintf(line, sizeof line, "total MPI exchange time: %g ms\n", (double)(float)(t ? t->exchange_ms : 0.0));
        out += line;
        snprintf(line, sizeof line, "total error calculation time: %g ms\n", (double)(float)(t ? t->error_ms : 0.0));
        out += line;
    }
    if (buf && buflen) {
        size_t n = std::min(buflen - 1, out.size());
        memcpy(buf, out.data(), n);
        buf[n] = 0;
    }
    return out.size() + 1;
}

}  // extern "C"
