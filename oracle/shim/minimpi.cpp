// Fork-based mini-MPI for the reference oracle build (test infrastructure).
//
// One rank unless LBM_MINIMPI_RANKS=P (P>1): MPI_Init then forks P-1 children
// and every ordered pair (src,dst) gets one single-slot mailbox in anonymous
// shared memory.  A send waits for the slot to be empty, copies, marks it full;
// a receive waits for full, copies out (the sender's byte count wins, because
// FullChannel/solver.hpp:847 posts an over-sized receive), marks it empty.
// Messages between a pair are consumed in program order, which is all the
// reference relies on.  MPI_Isend completes eagerly: both exchange idioms of
// the reference (Isend/Recv/Wait ring and Sendrecv_replace chain) find the slot
// empty when they send, so nothing deadlocks.
#include "mpi.h"
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sched.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>
#include <vector>

namespace {
struct Box {
    std::atomic<int> full;
    int bytes;
    char pad[56];
};
int g_rank = 0, g_size = 1;
size_t g_slot_bytes = 0;
char *g_shared = nullptr;          // [size*size] x (Box + slot)
std::atomic<int> *g_barrier = nullptr;
std::vector<pid_t> g_children;

size_t type_bytes(MPI_Datatype t) { return t == MPI_DOUBLE ? 8 : 4; }
Box *box(int src, int dst) {
    return reinterpret_cast<Box *>(g_shared + (size_t)(src * g_size + dst) * (sizeof(Box) + g_slot_bytes));
}
char *slot(int src, int dst) { return reinterpret_cast<char *>(box(src, dst)) + sizeof(Box); }
void put(int dst, const void *buf, size_t n) {
    if (n > g_slot_bytes) { fprintf(stderr, "minimpi: message of %zu bytes exceeds slot (%zu); raise LBM_MINIMPI_SLOT_MB\n", n, g_slot_bytes); _exit(3); }
    Box *b = box(g_rank, dst);
    while (b->full.load(std::memory_order_acquire)) sched_yield();
    memcpy(slot(g_rank, dst), buf, n);
    b->bytes = (int)n;
    b->full.store(1, std::memory_order_release);
}
void get(int src, void *buf, size_t cap) {
    Box *b = box(src, g_rank);
    while (!b->full.load(std::memory_order_acquire)) sched_yield();
    size_t n = (size_t)b->bytes;
    if (n > cap) n = cap;
    memcpy(buf, slot(src, g_rank), n);
    b->full.store(0, std::memory_order_release);
}
}  // namespace

extern "C" {
int minimpi_rank(void) { return g_rank; }
int minimpi_size(void) { return g_size; }

int MPI_Init(int *, char ***) {
    const char *e = getenv("LBM_MINIMPI_RANKS");
    int want = e ? atoi(e) : 1;
    if (want <= 1) return MPI_SUCCESS;
    const char *mb = getenv("LBM_MINIMPI_SLOT_MB");
    g_slot_bytes = (size_t)(mb ? atoi(mb) : 64) << 20;
    g_size = want;
    size_t total = (size_t)g_size * g_size * (sizeof(Box) + g_slot_bytes) + 4096;
    void *p = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (p == MAP_FAILED) { perror("minimpi mmap"); exit(3); }
    g_barrier = reinterpret_cast<std::atomic<int> *>(p);
    g_barrier[0] = 0;  // arrivals
    g_barrier[1] = 0;  // generation
    g_shared = static_cast<char *>(p) + 4096;
    for (int s = 0; s < g_size; ++s)
        for (int d = 0; d < g_size; ++d) box(s, d)->full = 0;
    fflush(nullptr);
    for (int r = 1; r < g_size; ++r) {
        pid_t pid = fork();
        if (pid < 0) { perror("minimpi fork"); exit(3); }
        if (pid == 0) { g_rank = r; g_children.clear(); return MPI_SUCCESS; }
        g_children.push_back(pid);
    }
    return MPI_SUCCESS;
}
int MPI_Finalize(void) {
    if (g_size > 1) {
        fflush(nullptr);
        if (g_rank != 0) _exit(0);
        for (pid_t c : g_children) { int st; waitpid(c, &st, 0); }
    }
    return MPI_SUCCESS;
}
int MPI_Comm_rank(MPI_Comm, int *rank) { *rank = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm, int *size) { *size = g_size; return MPI_SUCCESS; }
int MPI_Get_processor_name(char *name, int *len) { strcpy(name, "minimpi"); *len = 7; return MPI_SUCCESS; }
int MPI_Abort(MPI_Comm, int code) { fflush(nullptr); _exit(code ? code : 1); }
int MPI_Barrier(MPI_Comm) {
    if (g_size == 1) return MPI_SUCCESS;
    int gen = g_barrier[1].load(std::memory_order_acquire);
    if (g_barrier[0].fetch_add(1, std::memory_order_acq_rel) == g_size - 1) {
        g_barrier[0].store(0, std::memory_order_relaxed);
        g_barrier[1].store(gen + 1, std::memory_order_release);
    } else {
        while (g_barrier[1].load(std::memory_order_acquire) == gen) sched_yield();
    }
    return MPI_SUCCESS;
}
int MPI_Send(const void *buf, int count, MPI_Datatype type, int dest, int, MPI_Comm) {
    put(dest, buf, (size_t)count * type_bytes(type));
    return MPI_SUCCESS;
}
int MPI_Recv(void *buf, int count, MPI_Datatype type, int src, int, MPI_Comm, MPI_Status *) {
    get(src, buf, (size_t)count * type_bytes(type));
    return MPI_SUCCESS;
}
int MPI_Isend(const void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm, MPI_Request *req) {
    if (req) *req = 0;
    return MPI_Send(buf, count, type, dest, tag, comm);
}
int MPI_Wait(MPI_Request *, MPI_Status *) { return MPI_SUCCESS; }
int MPI_Sendrecv_replace(void *buf, int count, MPI_Datatype type, int dest, int sendtag, int src, int recvtag,
                         MPI_Comm comm, MPI_Status *st) {
    MPI_Send(buf, count, type, dest, sendtag, comm);
    return MPI_Recv(buf, count, type, src, recvtag, comm, st);
}
}
