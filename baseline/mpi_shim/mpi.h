/* mpi.h — single-node MPI stand-in used ONLY to build the reference programs on an image that
 * ships no MPI. Header-only: ranks are OS processes, messages travel through
 * single-slot mailboxes in one shared-memory file. It implements exactly the entry points the
 * reference uses; nothing in the wave3d product links against it.
 *
 * Environment: SHIM_RANK / SHIM_SIZE (fall back to RANK / WORLD_SIZE, then 0 / 1), SHIM_JOB (unique
 * name of the shared file; required when SHIM_SIZE > 1), SHIM_CAP_BYTES (mailbox payload capacity,
 * default 40 MiB; the file is sparse, only used mailboxes are ever touched), SHIM_TIMING_DIR
 * (optional: every rank writes the wall time between its first MPI_Barrier and its first MPI_Reduce,
 * i.e. the reference's timed region, to <dir>/rank_<r>.txt — an outside measurement that leaves the
 * reference sources untouched).
 */
#ifndef W3D_MPI_SHIM_H
#define W3D_MPI_SHIM_H

#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Info;
struct MPI_Status {
    int MPI_SOURCE, MPI_TAG, MPI_ERROR;
};

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_DOUBLE 8
#define MPI_INT 4
#define MPI_MAX 1
#define MPI_PROC_NULL (-2)
#define MPI_INFO_NULL 0
#define MPI_COMM_TYPE_SHARED 1
#define MPI_STATUS_IGNORE ((MPI_Status*)0)

namespace mpishim {

struct Mailbox {
    std::atomic<int> full;
    int tag;
    long long bytes;
    char pad[48];
    /* payload follows */
};

struct Header {
    std::atomic<int> attached;
    std::atomic<int> bar_count;
    std::atomic<int> bar_sense;
    char pad[116];
};

struct State {
    int rank = 0, size = 1;
    size_t cap = 0, stride = 0, total = 0;
    char* base = nullptr;
    char name[256] = {0};
    int dims[3] = {1, 1, 1}, periods[3] = {0, 0, 0};
    int local_sense = 0;
    double t_barrier = -1.0, t_reduce = -1.0;
};

inline State& st() {
    static State s;
    return s;
}

inline double now() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + ts.tv_nsec * 1e-9;
}

inline void die(const char* msg) {
    fprintf(stderr, "[mpi shim rank %d] %s\n", st().rank, msg);
    _exit(97);
}

inline Mailbox* box(int src, int dst) {
    State& s = st();
    return reinterpret_cast<Mailbox*>(s.base + sizeof(Header) + ((size_t)src * s.size + dst) * s.stride);
}

inline void relax(unsigned& spins) {
    if (++spins > 2000) sched_yield();
}

inline void send(const void* buf, long long bytes, int dst, int tag) {
    State& s = st();
    if ((size_t)bytes > s.cap) die("message larger than SHIM_CAP_BYTES");
    Mailbox* m = box(s.rank, dst);
    unsigned spins = 0;
    while (m->full.load(std::memory_order_acquire)) relax(spins);
    m->tag = tag;
    m->bytes = bytes;
    memcpy(reinterpret_cast<char*>(m) + sizeof(Mailbox), buf, (size_t)bytes);
    m->full.store(1, std::memory_order_release);
}

inline void recv(void* buf, long long bytes, int src, int tag) {
    State& s = st();
    Mailbox* m = box(src, s.rank);
    unsigned spins = 0;
    while (!m->full.load(std::memory_order_acquire)) relax(spins);
    if (m->tag != tag) die("tag mismatch (messages between a pair must be consumed in order)");
    if (m->bytes > bytes) die("receive buffer too small");
    memcpy(buf, reinterpret_cast<char*>(m) + sizeof(Mailbox), (size_t)m->bytes);
    m->full.store(0, std::memory_order_release);
}

inline int env_int(const char* a, const char* b, int dflt) {
    const char* v = getenv(a);
    if (!v && b) v = getenv(b);
    return v ? atoi(v) : dflt;
}

}  // namespace mpishim

inline int MPI_Init(int*, char***) {
    using namespace mpishim;
    State& s = st();
    s.rank = env_int("SHIM_RANK", "RANK", 0);
    s.size = env_int("SHIM_SIZE", "WORLD_SIZE", 1);
    const char* cap = getenv("SHIM_CAP_BYTES");
    s.cap = cap ? (size_t)atoll(cap) : (size_t)40 << 20;
    s.stride = (sizeof(Mailbox) + s.cap + 4095) / 4096 * 4096;
    s.total = sizeof(Header) + (size_t)s.size * s.size * s.stride;
    const char* job = getenv("SHIM_JOB");
    if (!job && s.size > 1) die("SHIM_JOB must be set when SHIM_SIZE > 1");
    char jobbuf[64];
    if (!job) {
        snprintf(jobbuf, sizeof jobbuf, "solo%d", (int)getpid());
        job = jobbuf;
    }
    const char* dir = access("/dev/shm", W_OK) == 0 ? "/dev/shm" : "/tmp";
    snprintf(s.name, sizeof s.name, "%s/w3d_mpishim_%s", dir, job);
    int fd = open(s.name, O_RDWR | O_CREAT, 0600);
    if (fd < 0) die("cannot open the shared-memory file");
    if (ftruncate(fd, (off_t)s.total) != 0) die("cannot size the shared-memory file");
    void* p = mmap(nullptr, s.total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) die("mmap failed");
    s.base = static_cast<char*>(p);
    reinterpret_cast<Header*>(s.base)->attached.fetch_add(1);
    return MPI_SUCCESS;
}

inline int MPI_Comm_size(MPI_Comm, int* n) {
    *n = mpishim::st().size;
    return MPI_SUCCESS;
}
inline int MPI_Comm_rank(MPI_Comm, int* r) {
    *r = mpishim::st().rank;
    return MPI_SUCCESS;
}
inline double MPI_Wtime() { return mpishim::now(); }

inline int MPI_Barrier(MPI_Comm) {
    using namespace mpishim;
    State& s = st();
    Header* h = reinterpret_cast<Header*>(s.base);
    s.local_sense ^= 1;
    if (h->bar_count.fetch_add(1) == s.size - 1) {
        h->bar_count.store(0);
        h->bar_sense.store(s.local_sense, std::memory_order_release);
    } else {
        unsigned spins = 0;
        while (h->bar_sense.load(std::memory_order_acquire) != s.local_sense) relax(spins);
    }
    if (s.t_barrier < 0) s.t_barrier = now();
    return MPI_SUCCESS;
}

inline int MPI_Dims_create(int nnodes, int ndims, int* dims) {
    /* balanced factorisation, non-increasing (all entries are 0 = free in the reference) */
    std::vector<int> primes;
    int m = nnodes;
    for (int f = 2; (long)f * f <= m; f++)
        while (m % f == 0) {
            primes.push_back(f);
            m /= f;
        }
    if (m > 1) primes.push_back(m);
    std::sort(primes.rbegin(), primes.rend());
    for (int i = 0; i < ndims; i++) dims[i] = 1;
    for (int f : primes) {
        int sidx = 0;
        for (int i = 1; i < ndims; i++)
            if (dims[i] < dims[sidx]) sidx = i;
        dims[sidx] *= f;
    }
    std::sort(dims, dims + ndims, [](int a, int b) { return a > b; });
    return MPI_SUCCESS;
}

inline int MPI_Cart_create(MPI_Comm, int ndims, const int* dims, const int* periods, int, MPI_Comm* out) {
    mpishim::State& s = mpishim::st();
    for (int i = 0; i < 3; i++) {
        s.dims[i] = i < ndims ? dims[i] : 1;
        s.periods[i] = i < ndims ? periods[i] : 0;
    }
    *out = 1;
    return MPI_SUCCESS;
}

inline int MPI_Cart_coords(MPI_Comm, int rank, int, int* c) {
    mpishim::State& s = mpishim::st();
    c[2] = rank % s.dims[2];
    c[1] = (rank / s.dims[2]) % s.dims[1];
    c[0] = rank / (s.dims[2] * s.dims[1]);
    return MPI_SUCCESS;
}

inline int MPI_Cart_shift(MPI_Comm comm, int dir, int disp, int* src, int* dst) {
    mpishim::State& s = mpishim::st();
    int c[3];
    MPI_Cart_coords(comm, s.rank, 3, c);
    auto at = [&](int v) -> int {
        int cc[3] = {c[0], c[1], c[2]};
        cc[dir] = v;
        if (cc[dir] < 0 || cc[dir] >= s.dims[dir]) {
            if (!s.periods[dir]) return MPI_PROC_NULL;
            cc[dir] = (cc[dir] % s.dims[dir] + s.dims[dir]) % s.dims[dir];
        }
        return (cc[0] * s.dims[1] + cc[1]) * s.dims[2] + cc[2];
    };
    *src = at(c[dir] - disp);
    *dst = at(c[dir] + disp);
    return MPI_SUCCESS;
}

inline int MPI_Comm_split_type(MPI_Comm, int, int, MPI_Info, MPI_Comm* out) {
    *out = 2; /* one node: the shared-memory communicator is the world */
    return MPI_SUCCESS;
}
inline int MPI_Comm_free(MPI_Comm*) { return MPI_SUCCESS; }

inline int MPI_Send(const void* buf, int count, MPI_Datatype dt, int dst, int tag, MPI_Comm) {
    if (dst < 0) return MPI_SUCCESS;
    mpishim::send(buf, (long long)count * dt, dst, tag);
    return MPI_SUCCESS;
}
inline int MPI_Recv(void* buf, int count, MPI_Datatype dt, int src, int tag, MPI_Comm, MPI_Status*) {
    if (src < 0) return MPI_SUCCESS;
    mpishim::recv(buf, (long long)count * dt, src, tag);
    return MPI_SUCCESS;
}
inline int MPI_Sendrecv(const void* sbuf, int scount, MPI_Datatype sdt, int dst, int stag, void* rbuf, int rcount,
                        MPI_Datatype rdt, int src, int rtag, MPI_Comm, MPI_Status*) {
    /* every rank deposits first, then collects: deadlock-free for rings, pairs and self-exchange */
    if (dst >= 0) mpishim::send(sbuf, (long long)scount * sdt, dst, stag);
    if (src >= 0) mpishim::recv(rbuf, (long long)rcount * rdt, src, rtag);
    return MPI_SUCCESS;
}

inline int MPI_Reduce(const void* sbuf, void* rbuf, int count, MPI_Datatype dt, MPI_Op, int root, MPI_Comm) {
    using namespace mpishim;
    State& s = st();
    if (s.t_reduce < 0) s.t_reduce = now();
    if (dt != MPI_DOUBLE) die("MPI_Reduce: only MPI_DOUBLE/MPI_MAX is implemented");
    const int tag = 1 << 20;
    if (s.rank != root) {
        mpishim::send(sbuf, (long long)count * dt, root, tag);
        return MPI_SUCCESS;
    }
    double* out = static_cast<double*>(rbuf);
    memcpy(out, sbuf, (size_t)count * dt);
    std::vector<double> tmp(count);
    for (int r = 0; r < s.size; r++) {
        if (r == root) continue;
        mpishim::recv(tmp.data(), (long long)count * dt, r, tag);
        for (int i = 0; i < count; i++)
            if (tmp[i] > out[i]) out[i] = tmp[i];
    }
    return MPI_SUCCESS;
}

inline int MPI_Finalize() {
    using namespace mpishim;
    State& s = st();
    if (const char* d = getenv("SHIM_TIMING_DIR")) {
        char path[512];
        snprintf(path, sizeof path, "%s/rank_%d.txt", d, s.rank);
        if (FILE* f = fopen(path, "w")) {
            fprintf(f, "%.6f\n", (s.t_barrier >= 0 && s.t_reduce >= 0) ? (s.t_reduce - s.t_barrier) * 1e3 : -1.0);
            fclose(f);
        }
    }
    MPI_Barrier(0);
    if (s.rank == 0) unlink(s.name);
    munmap(s.base, s.total);
    return MPI_SUCCESS;
}

#endif /* W3D_MPI_SHIM_H */
