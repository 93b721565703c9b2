/*
 * mpi.h stand-in: the seven MPI calls the reference uses (Init, Comm_rank, Comm_size,
 * Send, Recv, Barrier, Finalize), implemented with fork() and pipes so that
 * "mpirun -np P" becomes the environment variable BSP_REF_RANKS=P (default 1).
 * TEST INFRASTRUCTURE ONLY.  Only the traffic pattern of the reference is supported:
 * ranks 1..P-1 send to rank 0 in order, one barrier after the exchange.
 */
#ifndef BSP_SHIM_MPI_H
#define BSP_SHIM_MPI_H
#include <cstdio>
#include <cstdlib>
#include <unistd.h>
#include <sys/wait.h>

typedef int MPI_Comm;
typedef int MPI_Datatype;
struct MPI_Status { int source, tag; };
#define MPI_COMM_WORLD 0
#define MPI_INT 4
#define MPI_UNSIGNED_LONG_LONG 8
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_SUCCESS 0

namespace bsp_shim {
static int rank = 0, size = 1;
static int up[64][2];   /* child r -> rank 0 */
static int down[64][2]; /* rank 0 -> child r */
static inline void put(int fd, const void *p, size_t n)
{
    const char *c = (const char *)p;
    while (n) { ssize_t w = write(fd, c, n); if (w <= 0) { perror("shim write"); _exit(3); } c += w; n -= (size_t)w; }
}
static inline void get(int fd, void *p, size_t n)
{
    char *c = (char *)p;
    while (n) { ssize_t r = read(fd, c, n); if (r <= 0) { perror("shim read"); _exit(3); } c += r; n -= (size_t)r; }
}
}

static inline int MPI_Init(int *, char ***)
{
    using namespace bsp_shim;
    const char *e = getenv("BSP_REF_RANKS");
    size = e ? atoi(e) : 1;
    if (size < 1 || size > 64) size = 1;
    fflush(stdout);
    for (int r = 1; r < size; r++) { if (pipe(up[r]) || pipe(down[r])) { perror("pipe"); exit(3); } }
    for (int r = 1; r < size; r++) {
        pid_t p = fork();
        if (p < 0) { perror("fork"); exit(3); }
        if (p == 0) { rank = r; break; }
    }
    return MPI_SUCCESS;
}
static inline int MPI_Comm_rank(MPI_Comm, int *r) { *r = bsp_shim::rank; return MPI_SUCCESS; }
static inline int MPI_Comm_size(MPI_Comm, int *s) { *s = bsp_shim::size; return MPI_SUCCESS; }
static inline int MPI_Send(const void *buf, int count, MPI_Datatype t, int, int, MPI_Comm)
{
    bsp_shim::put(bsp_shim::up[bsp_shim::rank][1], buf, (size_t)count * (size_t)t);
    return MPI_SUCCESS;
}
static inline int MPI_Recv(void *buf, int count, MPI_Datatype t, int src, int, MPI_Comm, MPI_Status *)
{
    bsp_shim::get(bsp_shim::up[src][0], buf, (size_t)count * (size_t)t);
    return MPI_SUCCESS;
}
static inline int MPI_Barrier(MPI_Comm)
{
    using namespace bsp_shim;
    char tok = 1;
    if (size == 1) return MPI_SUCCESS;
    if (rank) { put(up[rank][1], &tok, 1); get(down[rank][0], &tok, 1); }
    else {
        for (int r = 1; r < size; r++) get(up[r][0], &tok, 1);
        for (int r = 1; r < size; r++) put(down[r][1], &tok, 1);
    }
    return MPI_SUCCESS;
}
static inline int MPI_Finalize()
{
    using namespace bsp_shim;
    fflush(stdout);
    if (rank == 0) for (int r = 1; r < size; r++) { int st; wait(&st); }
    return MPI_SUCCESS;
}
#endif
