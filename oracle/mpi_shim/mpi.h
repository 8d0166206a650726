/*
 * oracle/mpi_shim/mpi.h -- the eight MPI calls the reference's Cascade programs use, over fork() and
 * pipes, so that the UNMODIFIED code/mpi_svm_main2.cpp and code/mpi_svm_main3.cpp compile and run on a
 * machine without MPI (this image has no mpi.h / mpirun).  TEST INFRASTRUCTURE ONLY (see smo_oracle.h):
 * it exists to run the reference as the Cascade oracle and as the timed CPU baseline.
 *
 * Model: MPI_Init forks world-1 children (world = $MPI_SHIM_NP, default 2); every ordered pair of
 * ranks has its own pipe, messages carry (tag, bytes) headers and are received in FIFO order per
 * pair, which is all the reference needs (blocking MPI_Send / MPI_Recv in matching program order,
 * MPI_Bcast from rank 0).  A blocking write is a legal MPI_Send (synchronous-mode semantics).
 */
#ifndef MPI_SHIM_H
#define MPI_SHIM_H

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef struct { int unused; } MPI_Status;
#define MPI_COMM_WORLD 0
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_INT 4
#define MPI_DOUBLE 8
#define MPI_UNSIGNED_LONG_LONG 9   /* 8 bytes */
#define MPI_SUCCESS 0

namespace mpi_shim {
inline int &rank() { static int r = 0; return r; }
inline int &world() { static int w = 1; return w; }
inline int (*&fds())[2] { static int (*p)[2] = nullptr; return p; }   /* fds()[src * world + dst] = {read end, write end} */
inline size_t size_of(MPI_Datatype t) { return t == MPI_INT ? 4 : 8; }
inline void xwrite(int fd, const void *buf, size_t n)
{
    const char *p = (const char *)buf;
    while (n) { ssize_t k = write(fd, p, n); if (k <= 0) { perror("mpi_shim write"); _exit(3); } p += k; n -= (size_t)k; }
}
inline void xread(int fd, void *buf, size_t n)
{
    char *p = (char *)buf;
    while (n) { ssize_t k = read(fd, p, n); if (k <= 0) { fprintf(stderr, "mpi_shim: rank %d: peer closed\n", rank()); _exit(3); } p += k; n -= (size_t)k; }
}
} // namespace mpi_shim

inline int MPI_Init(int *, char ***)
{
    using namespace mpi_shim;
    const char *e = getenv("MPI_SHIM_NP");
    world() = e ? atoi(e) : 2;
    if (world() < 1) world() = 1;
    const int w = world();
    fds() = (int (*)[2])malloc(sizeof(int[2]) * w * w);
    for (int i = 0; i < w * w; ++i)
        if (pipe(fds()[i]) != 0) { perror("pipe"); exit(3); }
    fflush(stdout); fflush(stderr);
    for (int r = 1; r < w; ++r) {
        pid_t pid = fork();
        if (pid < 0) { perror("fork"); exit(3); }
        if (pid == 0) { rank() = r; break; }
    }
    return MPI_SUCCESS;
}
inline int MPI_Comm_rank(MPI_Comm, int *r) { *r = mpi_shim::rank(); return MPI_SUCCESS; }
inline int MPI_Comm_size(MPI_Comm, int *s) { *s = mpi_shim::world(); return MPI_SUCCESS; }

inline int MPI_Send(const void *buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm)
{
    using namespace mpi_shim;
    long long hdr[2] = {tag, (long long)count * (long long)size_of(t)};
    const int fd = fds()[rank() * world() + dest][1];
    xwrite(fd, hdr, sizeof hdr);
    xwrite(fd, buf, (size_t)hdr[1]);
    return MPI_SUCCESS;
}
inline int MPI_Recv(void *buf, int count, MPI_Datatype t, int source, int tag, MPI_Comm, MPI_Status *)
{
    using namespace mpi_shim;
    long long hdr[2];
    const int fd = fds()[source * world() + rank()][0];
    xread(fd, hdr, sizeof hdr);
    if (hdr[0] != tag || hdr[1] > (long long)count * (long long)size_of(t)) {
        fprintf(stderr, "mpi_shim: rank %d expected tag %d (<= %lld bytes) from %d, got tag %lld (%lld bytes)\n", rank(), tag,
                (long long)count * (long long)size_of(t), source, hdr[0], hdr[1]);
        _exit(3);
    }
    xread(fd, buf, (size_t)hdr[1]);
    return MPI_SUCCESS;
}
inline int MPI_Bcast(void *buf, int count, MPI_Datatype t, int root, MPI_Comm c)
{
    using namespace mpi_shim;
    if (rank() == root) {
        for (int r = 0; r < world(); ++r)
            if (r != root) MPI_Send(buf, count, t, r, -77, c);
    } else {
        MPI_Recv(buf, count, t, root, -77, c, MPI_STATUS_IGNORE);
    }
    return MPI_SUCCESS;
}
inline int MPI_Abort(MPI_Comm, int code) { fflush(stdout); fflush(stderr); _exit(code ? code : 1); return 0; }
inline int MPI_Finalize()
{
    using namespace mpi_shim;
    fflush(stdout); fflush(stderr);
    if (rank() != 0) _exit(0);
    int st;
    while (wait(&st) > 0) { }
    return MPI_SUCCESS;
}
#endif
