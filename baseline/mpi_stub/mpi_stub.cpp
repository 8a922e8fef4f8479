// Single-node stand-in for an MPI library (see mpi.h in this directory).  Ranks are processes on one machine.
// World size 1 needs nothing; for N > 1 every rank maps the same file (TCLB_MPI_DIR, default /tmp) holding
//   - a sense-reversing barrier and N collective slots (each rank writes its contribution, everybody reads),
//   - one byte ring per ordered pair of ranks for point-to-point messages (header {tag, bytes} + payload, FIFO per pair).
// Isend/Irecv are matched by (source, tag); a progress loop (run inside every Wait*/Test/collective) moves bytes
// between user buffers and the rings in pieces, so a message may be larger than a ring and both sides overlap their
// copies - the same copy-in/copy-out a shared-memory MPI transport does for large messages.
#include "mpi.h"

#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <unistd.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <list>
#include <string>
#include <vector>

namespace {

constexpr size_t COLL_SLOT = 4u << 20;        // bytes one rank can contribute to one collective step
constexpr size_t RING_BYTES = 8u << 20;       // data bytes of one pair ring
constexpr int MAX_RANKS = 64;

struct Ring {
    std::atomic<uint64_t> head;   // bytes written (sender)
    char pad0[56];
    std::atomic<uint64_t> tail;   // bytes read (receiver)
    char pad1[56];
};
struct Shared {
    std::atomic<int> attached;
    std::atomic<int> bar_count;
    std::atomic<int> bar_sense;
    std::atomic<int> aborted;
    char pad[48];
};

int g_rank = 0, g_size = 1;
bool g_init = false;
Shared *g_sh = nullptr;
char *g_coll = nullptr;            // [size][COLL_SLOT]
Ring *g_rings = nullptr;           // [size*size] control blocks, index src*size+dst
char *g_ring_data = nullptr;       // [size*size][RING_BYTES]
int g_local_sense = 0;
std::string g_path;

struct MsgHeader { int tag; int pad; uint64_t bytes; };

struct Req {
    bool active = false, done = false, is_recv = false;
    int peer = -1, tag = 0;
    char *buf = nullptr;
    uint64_t bytes = 0, moved = 0;
    bool header_done = false;
};
std::vector<Req> g_reqs;
std::vector<std::deque<int>> g_sendq;          // per destination: request ids in posting order
std::list<int> g_posted;                        // posted receives not yet matched, in posting order
struct Incoming {
    bool active = false;
    MsgHeader h{};
    uint64_t got = 0;
    int req = -1;                                // receive request being filled, or -1 while unexpected
    std::vector<char> *stash = nullptr;
};
std::vector<Incoming> g_in;                      // per source: message currently streaming out of the ring
struct Unexpected { int src, tag; std::vector<char> *data; bool complete; int claimed_by; };
std::list<Unexpected> g_unexp;

struct DType { size_t size, extent; int base; };
std::vector<DType> g_types;

size_t type_size(MPI_Datatype t) {
    if (t >= 0 && t < (int)g_types.size()) return g_types[t].size;
    return 1;
}
size_t type_extent(MPI_Datatype t) {
    if (t >= 0 && t < (int)g_types.size()) return g_types[t].extent;
    return 1;
}
void init_types() {
    if (!g_types.empty()) return;
    const size_t s[] = {0, 1, 1, 1, 2, 2, 4, 4, sizeof(long), sizeof(unsigned long), 8, 8, 4, 8};
    for (int i = 0; i < 14; i++) g_types.push_back(DType{s[i], s[i], i});
}

int new_req() {
    for (size_t i = 0; i < g_reqs.size(); i++)
        if (!g_reqs[i].active) { g_reqs[i] = Req(); g_reqs[i].active = true; return (int)i; }
    g_reqs.push_back(Req());
    g_reqs.back().active = true;
    return (int)g_reqs.size() - 1;
}

Ring &ring(int src, int dst) { return g_rings[src * g_size + dst]; }
char *ring_data(int src, int dst) { return g_ring_data + (size_t)(src * g_size + dst) * RING_BYTES; }

void ring_write(int dst, const char *p, uint64_t n, uint64_t head) {
    char *d = ring_data(g_rank, dst);
    uint64_t o = head % RING_BYTES, first = n < RING_BYTES - o ? n : RING_BYTES - o;
    memcpy(d + o, p, first);
    if (n > first) memcpy(d, p + first, n - first);
}
void ring_read(int src, char *p, uint64_t n, uint64_t tail) {
    const char *d = ring_data(src, g_rank);
    uint64_t o = tail % RING_BYTES, first = n < RING_BYTES - o ? n : RING_BYTES - o;
    memcpy(p, d + o, first);
    if (n > first) memcpy(p + first, d, n - first);
}

void check_abort() {
    if (g_sh && g_sh->aborted.load(std::memory_order_relaxed)) { fprintf(stderr, "[mpi_stub %d] another rank aborted\n", g_rank); _exit(1); }
}

// one pass of the progress engine; returns true if any byte moved
bool progress() {
    if (g_size == 1) return false;
    bool moved = false;
    // ---- sends ----
    for (int dst = 0; dst < g_size; dst++) {
        std::deque<int> &q = g_sendq[dst];
        while (!q.empty()) {
            Req &r = g_reqs[q.front()];
            Ring &R = ring(g_rank, dst);
            uint64_t head = R.head.load(std::memory_order_relaxed), tail = R.tail.load(std::memory_order_acquire);
            uint64_t space = RING_BYTES - (head - tail);
            if (!r.header_done) {
                if (space < sizeof(MsgHeader)) break;
                MsgHeader h{r.tag, 0, r.bytes};
                ring_write(dst, (const char *)&h, sizeof(h), head);
                head += sizeof(h); space -= sizeof(h);
                R.head.store(head, std::memory_order_release);
                r.header_done = true; moved = true;
            }
            uint64_t n = r.bytes - r.moved; if (n > space) n = space;
            if (n > 0) {
                ring_write(dst, r.buf + r.moved, n, head);
                R.head.store(head + n, std::memory_order_release);
                r.moved += n; moved = true;
            }
            if (r.moved == r.bytes) { r.done = true; q.pop_front(); } else break;
        }
    }
    // ---- receives ----
    for (int src = 0; src < g_size; src++) {
        if (src == g_rank) continue;
        Ring &R = ring(src, g_rank);
        for (;;) {
            uint64_t tail = R.tail.load(std::memory_order_relaxed), head = R.head.load(std::memory_order_acquire);
            uint64_t avail = head - tail;
            Incoming &in = g_in[src];
            if (!in.active) {
                if (avail < sizeof(MsgHeader)) break;
                ring_read(src, (char *)&in.h, sizeof(MsgHeader), tail);
                tail += sizeof(MsgHeader); avail -= sizeof(MsgHeader);
                R.tail.store(tail, std::memory_order_release);
                in.active = true; in.got = 0; in.req = -1; in.stash = nullptr; moved = true;
                for (auto it = g_posted.begin(); it != g_posted.end(); ++it) {
                    Req &r = g_reqs[*it];
                    if (r.peer == src && r.tag == in.h.tag) { in.req = *it; g_posted.erase(it); break; }
                }
                if (in.req < 0) {
                    in.stash = new std::vector<char>(in.h.bytes);
                    g_unexp.push_back(Unexpected{src, in.h.tag, in.stash, false, -1});
                } else if (g_reqs[in.req].bytes < in.h.bytes) {
                    fprintf(stderr, "[mpi_stub %d] message from %d tag %d truncated (%llu > %llu)\n", g_rank, src, in.h.tag,
                            (unsigned long long)in.h.bytes, (unsigned long long)g_reqs[in.req].bytes);
                    MPI_Abort(MPI_COMM_WORLD, 3);
                }
            }
            uint64_t n = in.h.bytes - in.got; if (n > avail) n = avail;
            if (n > 0) {
                char *dstp = in.req >= 0 ? g_reqs[in.req].buf + in.got : in.stash->data() + in.got;
                ring_read(src, dstp, n, tail);
                R.tail.store(tail + n, std::memory_order_release);
                in.got += n; moved = true;
            }
            if (in.got == in.h.bytes) {
                if (in.req >= 0) { g_reqs[in.req].moved = in.h.bytes; g_reqs[in.req].done = true; }
                else {
                    for (auto &u : g_unexp) if (u.data == in.stash) {
                        u.complete = true;
                        if (u.claimed_by >= 0) {
                            Req &r = g_reqs[u.claimed_by];
                            memcpy(r.buf, u.data->data(), u.data->size());
                            r.moved = u.data->size(); r.done = true;
                        }
                    }
                    g_unexp.remove_if([](const Unexpected &u) { if (u.complete && u.claimed_by >= 0) { delete u.data; return true; } return false; });
                }
                in.active = false;
            } else break;
        }
    }
    return moved;
}

void spin(int &idle) {
    if (++idle > 200) { sched_yield(); if ((idle & 1023) == 0) check_abort(); }
}

void barrier() {
    if (g_size == 1) return;
    g_local_sense ^= 1;
    if (g_sh->bar_count.fetch_add(1, std::memory_order_acq_rel) == g_size - 1) {
        g_sh->bar_count.store(0, std::memory_order_relaxed);
        g_sh->bar_sense.store(g_local_sense, std::memory_order_release);
    } else {
        int idle = 0;
        while (g_sh->bar_sense.load(std::memory_order_acquire) != g_local_sense) { if (!progress()) spin(idle); else idle = 0; }
    }
}

template <class T> void reduce_op(T *acc, const T *v, int n, MPI_Op op) {
    for (int i = 0; i < n; i++) {
        switch (op) {
        case MPI_SUM: acc[i] = acc[i] + v[i]; break;
        case MPI_PROD: acc[i] = acc[i] * v[i]; break;
        case MPI_MAX: acc[i] = acc[i] > v[i] ? acc[i] : v[i]; break;
        case MPI_MIN: acc[i] = acc[i] < v[i] ? acc[i] : v[i]; break;
        case MPI_LOR: acc[i] = (T)((acc[i] != 0) || (v[i] != 0)); break;
        case MPI_LAND: acc[i] = (T)((acc[i] != 0) && (v[i] != 0)); break;
        default: break;
        }
    }
}
void reduce_any(void *acc, const void *v, int n, MPI_Datatype t, MPI_Op op) {
    switch (t) {
    case MPI_CHAR: reduce_op((signed char *)acc, (const signed char *)v, n, op); break;
    case MPI_BYTE: case MPI_UNSIGNED_CHAR: reduce_op((unsigned char *)acc, (const unsigned char *)v, n, op); break;
    case MPI_SHORT: reduce_op((short *)acc, (const short *)v, n, op); break;
    case MPI_UNSIGNED_SHORT: reduce_op((unsigned short *)acc, (const unsigned short *)v, n, op); break;
    case MPI_INT: reduce_op((int *)acc, (const int *)v, n, op); break;
    case MPI_UNSIGNED: reduce_op((unsigned *)acc, (const unsigned *)v, n, op); break;
    case MPI_LONG: reduce_op((long *)acc, (const long *)v, n, op); break;
    case MPI_UNSIGNED_LONG: reduce_op((unsigned long *)acc, (const unsigned long *)v, n, op); break;
    case MPI_LONG_LONG: reduce_op((long long *)acc, (const long long *)v, n, op); break;
    case MPI_UNSIGNED_LONG_LONG: reduce_op((unsigned long long *)acc, (const unsigned long long *)v, n, op); break;
    case MPI_FLOAT: reduce_op((float *)acc, (const float *)v, n, op); break;
    case MPI_DOUBLE: reduce_op((double *)acc, (const double *)v, n, op); break;
    default: fprintf(stderr, "[mpi_stub] reduction on datatype %d is not supported\n", t); MPI_Abort(MPI_COMM_WORLD, 4);
    }
}

// every rank publishes `bytes` (<= COLL_SLOT) and may read every slot between the two barriers
void publish(const void *p, size_t bytes) {
    if (bytes > COLL_SLOT) { fprintf(stderr, "[mpi_stub] collective piece of %zu bytes exceeds the slot\n", bytes); MPI_Abort(MPI_COMM_WORLD, 5); }
    if (bytes) memcpy(g_coll + (size_t)g_rank * COLL_SLOT, p, bytes);
    barrier();
}
const char *slot(int r) { return g_coll + (size_t)r * COLL_SLOT; }

}  // namespace

extern "C" {

int MPI_Init(int *, char ***) {
    if (g_init) return MPI_SUCCESS;
    init_types();
    const char *r = getenv("TCLB_MPI_RANK"), *s = getenv("TCLB_MPI_SIZE");
    if (!r) r = getenv("RANK");
    if (!s) s = getenv("WORLD_SIZE");
    g_rank = r ? atoi(r) : 0;
    g_size = s ? atoi(s) : 1;
    if (g_size < 1 || g_size > MAX_RANKS || g_rank < 0 || g_rank >= g_size) { fprintf(stderr, "[mpi_stub] bad rank/size %d/%d\n", g_rank, g_size); exit(2); }
    g_init = true;
    if (g_size == 1) return MPI_SUCCESS;
    const char *dir = getenv("TCLB_MPI_DIR");
    const char *id = getenv("TCLB_MPI_ID");
    if (!id) id = getenv("MASTER_PORT");
    g_path = std::string(dir ? dir : "/tmp") + "/tclb_mpi_stub_" + (id ? id : "0") + "_" + std::to_string(g_size);
    const size_t ctl = 4096, rings_ctl = sizeof(Ring) * g_size * g_size;
    const size_t total = ctl + (size_t)g_size * COLL_SLOT + ((rings_ctl + 4095) / 4096) * 4096 + (size_t)g_size * g_size * RING_BYTES;
    int fd = open(g_path.c_str(), O_RDWR | O_CREAT, 0600);
    if (fd < 0) { perror("[mpi_stub] open"); exit(2); }
    if (ftruncate(fd, (off_t)total) != 0) { perror("[mpi_stub] ftruncate"); exit(2); }
    char *base = (char *)mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (base == MAP_FAILED) { perror("[mpi_stub] mmap"); exit(2); }
    close(fd);
    g_sh = (Shared *)base;
    g_coll = base + ctl;
    g_rings = (Ring *)(g_coll + (size_t)g_size * COLL_SLOT);
    g_ring_data = (char *)g_rings + ((rings_ctl + 4095) / 4096) * 4096;
    g_sendq.assign(g_size, std::deque<int>());
    g_in.assign(g_size, Incoming());
    // the file is fresh (zero-filled) for this job id: wait until every rank has attached, then rank 0 unlinks it
    g_sh->attached.fetch_add(1);
    int idle = 0;
    double t0 = MPI_Wtime();
    while (g_sh->attached.load() < g_size) {
        spin(idle);
        if (MPI_Wtime() - t0 > 120) { fprintf(stderr, "[mpi_stub %d] only %d of %d ranks attached to %s\n", g_rank, g_sh->attached.load(), g_size, g_path.c_str()); exit(2); }
    }
    barrier();
    if (g_rank == 0) unlink(g_path.c_str());
    return MPI_SUCCESS;
}
int MPI_Init_thread(int *a, char ***b, int, int *provided) { if (provided) *provided = MPI_THREAD_SINGLE; return MPI_Init(a, b); }
int MPI_Initialized(int *flag) { *flag = g_init ? 1 : 0; return MPI_SUCCESS; }
int MPI_Finalize(void) { if (g_init && g_size > 1) barrier(); return MPI_SUCCESS; }
int MPI_Abort(MPI_Comm, int code) {
    if (g_sh) g_sh->aborted.store(1);
    fprintf(stderr, "[mpi_stub %d] MPI_Abort(%d)\n", g_rank, code);
    _exit(code ? code : 1);
}
double MPI_Wtime(void) { struct timeval tv; gettimeofday(&tv, nullptr); return tv.tv_sec + 1e-6 * tv.tv_usec; }
int MPI_Get_processor_name(char *name, int *len) {
    if (gethostname(name, MPI_MAX_PROCESSOR_NAME - 1) != 0) strcpy(name, "localhost");
    name[MPI_MAX_PROCESSOR_NAME - 1] = 0;
    *len = (int)strlen(name);
    return MPI_SUCCESS;
}
int MPI_Comm_rank(MPI_Comm c, int *rank) { *rank = c == MPI_COMM_SELF ? 0 : g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm c, int *size) { *size = c == MPI_COMM_SELF ? 1 : g_size; return MPI_SUCCESS; }
int MPI_Comm_remote_size(MPI_Comm, int *size) { *size = 0; return MPI_SUCCESS; }
int MPI_Comm_split(MPI_Comm c, int color, int, MPI_Comm *newcomm) {
    // the reference splits the world by program name (MPMD) and by host name (GPU selection): on one machine with one
    // program every rank has the same colour.  Anything else is outside this stand-in.
    if (c != MPI_COMM_SELF && g_size > 1) {
        int lo = color, hi = color;
        MPI_Allreduce(&color, &lo, 1, MPI_INT, MPI_MIN, c);
        MPI_Allreduce(&color, &hi, 1, MPI_INT, MPI_MAX, c);
        if (lo != hi) { fprintf(stderr, "[mpi_stub] MPI_Comm_split into different colours is not supported\n"); MPI_Abort(c, 6); }
    }
    *newcomm = c;
    return MPI_SUCCESS;
}
int MPI_Comm_dup(MPI_Comm c, MPI_Comm *n) { *n = c; return MPI_SUCCESS; }
int MPI_Comm_free(MPI_Comm *c) { *c = MPI_COMM_NULL; return MPI_SUCCESS; }
int MPI_Comm_group(MPI_Comm, MPI_Group *g) { *g = 1; return MPI_SUCCESS; }
int MPI_Comm_create(MPI_Comm c, MPI_Group, MPI_Comm *n) { *n = c; return MPI_SUCCESS; }
int MPI_Comm_get_attr(MPI_Comm, int keyval, void *attr, int *flag) {
    static int universe = 0, tag_ub = 1 << 30;
    if (keyval == MPI_UNIVERSE_SIZE) { universe = g_size; *(int **)attr = &universe; *flag = 1; }
    else if (keyval == MPI_TAG_UB) { *(int **)attr = &tag_ub; *flag = 1; }
    else *flag = 0;
    return MPI_SUCCESS;
}
int MPI_Comm_get_parent(MPI_Comm *p) { *p = MPI_COMM_NULL; return MPI_SUCCESS; }
int MPI_Comm_spawn(const char *, char *[], int, MPI_Info, int, MPI_Comm, MPI_Comm *, int[]) {
    fprintf(stderr, "[mpi_stub] MPI_Comm_spawn is not available in this stand-in\n");
    return MPI_Abort(MPI_COMM_WORLD, 7);
}
int MPI_Comm_set_errhandler(MPI_Comm, MPI_Errhandler) { return MPI_SUCCESS; }
int MPI_Intercomm_create(MPI_Comm, int, MPI_Comm, int, int, MPI_Comm *inter) { *inter = MPI_COMM_NULL; return MPI_SUCCESS; }
int MPI_Group_excl(MPI_Group g, int n, const int[], MPI_Group *ng) {
    if (n > 0) { fprintf(stderr, "[mpi_stub] MPI_Group_excl with exclusions is not supported\n"); return MPI_Abort(MPI_COMM_WORLD, 8); }
    *ng = g; return MPI_SUCCESS;
}
int MPI_Group_rank(MPI_Group, int *r) { *r = g_rank; return MPI_SUCCESS; }
int MPI_Group_size(MPI_Group, int *s) { *s = g_size; return MPI_SUCCESS; }
int MPI_Group_free(MPI_Group *g) { *g = MPI_GROUP_NULL; return MPI_SUCCESS; }

int MPI_Barrier(MPI_Comm c) { if (c != MPI_COMM_SELF) barrier(); return MPI_SUCCESS; }

int MPI_Bcast(void *buf, int count, MPI_Datatype t, int root, MPI_Comm c) {
    if (g_size == 1 || c == MPI_COMM_SELF) return MPI_SUCCESS;
    const size_t total = (size_t)count * type_size(t);
    for (size_t off = 0; off < total || (total == 0 && off == 0); off += COLL_SLOT) {
        const size_t n = total - off < COLL_SLOT ? total - off : COLL_SLOT;
        publish(g_rank == root ? (const char *)buf + off : nullptr, g_rank == root ? n : 0);
        if (g_rank != root && n) memcpy((char *)buf + off, slot(root), n);
        barrier();
        if (total == 0) break;
    }
    return MPI_SUCCESS;
}
int MPI_Allreduce(const void *send, void *recv, int count, MPI_Datatype t, MPI_Op op, MPI_Comm c) {
    const size_t es = type_size(t), total = (size_t)count * es;
    if (send == MPI_IN_PLACE) send = recv;
    if (g_size == 1 || c == MPI_COMM_SELF) { if (send != recv) memcpy(recv, send, total); return MPI_SUCCESS; }
    const size_t per = (COLL_SLOT / es) * es;
    for (size_t off = 0; off < total; off += per) {
        const size_t n = total - off < per ? total - off : per;
        publish((const char *)send + off, n);
        std::vector<char> acc(slot(0), slot(0) + n);
        for (int r = 1; r < g_size; r++) reduce_any(acc.data(), slot(r), (int)(n / es), t, op);
        barrier();
        memcpy((char *)recv + off, acc.data(), n);
    }
    return MPI_SUCCESS;
}
int MPI_Reduce(const void *send, void *recv, int count, MPI_Datatype t, MPI_Op op, int root, MPI_Comm c) {
    if (g_size == 1 || c == MPI_COMM_SELF) { if (send != MPI_IN_PLACE && send != recv) memcpy(recv, send, (size_t)count * type_size(t)); return MPI_SUCCESS; }
    std::vector<char> tmp((size_t)count * type_size(t));
    if (send == MPI_IN_PLACE) send = recv;
    MPI_Allreduce(send, tmp.data(), count, t, op, c);
    if (g_rank == root) memcpy(recv, tmp.data(), tmp.size());
    return MPI_SUCCESS;
}
int MPI_Gatherv(const void *send, int scount, MPI_Datatype st, void *recv, const int *rcounts, const int *displs, MPI_Datatype rt, int root, MPI_Comm c) {
    const size_t sb = (size_t)scount * type_size(st);
    if (g_size == 1 || c == MPI_COMM_SELF) { memcpy((char *)recv + (size_t)displs[0] * type_size(rt), send, sb); return MPI_SUCCESS; }
    // ranks take turns through slot 0 so that a piece of any size can be streamed
    for (int r = 0; r < g_size; r++) {
        size_t total = sb;
        if (g_rank == r) memcpy(g_coll, &total, sizeof(total));
        barrier();
        memcpy(&total, g_coll, sizeof(total));
        barrier();
        for (size_t off = 0; off < total; off += COLL_SLOT) {
            const size_t n = total - off < COLL_SLOT ? total - off : COLL_SLOT;
            if (g_rank == r) memcpy(g_coll, (const char *)send + off, n);
            barrier();
            if (g_rank == root) memcpy((char *)recv + (size_t)displs[r] * type_size(rt) + off, g_coll, n);
            barrier();
        }
    }
    (void)rcounts;
    return MPI_SUCCESS;
}
int MPI_Gather(const void *send, int scount, MPI_Datatype st, void *recv, int rcount, MPI_Datatype rt, int root, MPI_Comm c) {
    std::vector<int> counts(g_size, rcount), displs(g_size);
    for (int r = 0; r < g_size; r++) displs[r] = r * rcount;
    return MPI_Gatherv(send, scount, st, recv, counts.data(), displs.data(), rt, root, c);
}
int MPI_Allgather(const void *send, int scount, MPI_Datatype st, void *recv, int rcount, MPI_Datatype rt, MPI_Comm c) {
    for (int root = 0; root < (c == MPI_COMM_SELF ? 1 : g_size); root++) MPI_Gather(send, scount, st, recv, rcount, rt, root, c);
    return MPI_SUCCESS;
}
int MPI_Scatterv(const void *send, const int *scounts, const int *displs, MPI_Datatype st, void *recv, int rcount, MPI_Datatype rt, int root, MPI_Comm c) {
    const size_t rb = (size_t)rcount * type_size(rt);
    if (g_size == 1 || c == MPI_COMM_SELF) { memcpy(recv, (const char *)send + (size_t)displs[0] * type_size(st), rb); return MPI_SUCCESS; }
    for (int r = 0; r < g_size; r++) {
        size_t total = rb;
        if (g_rank == r) memcpy(g_coll, &total, sizeof(total));
        barrier();
        memcpy(&total, g_coll, sizeof(total));
        barrier();
        for (size_t off = 0; off < total; off += COLL_SLOT) {
            const size_t n = total - off < COLL_SLOT ? total - off : COLL_SLOT;
            if (g_rank == root) memcpy(g_coll, (const char *)send + (size_t)displs[r] * type_size(st) + off, n);
            barrier();
            if (g_rank == r) memcpy((char *)recv + off, g_coll, n);
            barrier();
        }
    }
    (void)scounts;
    return MPI_SUCCESS;
}

int MPI_Isend(const void *buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm, MPI_Request *req) {
    const int id = new_req();
    Req &r = g_reqs[id];
    r.is_recv = false; r.peer = dest; r.tag = tag; r.buf = (char *)buf; r.bytes = (uint64_t)count * type_size(t);
    *req = id;
    if (dest == MPI_PROC_NULL) { r.done = true; return MPI_SUCCESS; }
    if (dest == g_rank || g_size == 1) {
        // message to myself: hand it to a posted receive or keep a copy
        for (auto it = g_posted.begin(); it != g_posted.end(); ++it) {
            Req &p = g_reqs[*it];
            if (p.peer == g_rank && p.tag == tag) { memcpy(p.buf, buf, r.bytes); p.moved = r.bytes; p.done = true; g_posted.erase(it); r.done = true; return MPI_SUCCESS; }
        }
        auto *v = new std::vector<char>((const char *)buf, (const char *)buf + r.bytes);
        g_unexp.push_back(Unexpected{g_rank, tag, v, true, -1});
        r.done = true;
        return MPI_SUCCESS;
    }
    g_sendq[dest].push_back(id);
    progress();
    return MPI_SUCCESS;
}
int MPI_Irecv(void *buf, int count, MPI_Datatype t, int source, int tag, MPI_Comm, MPI_Request *req) {
    const int id = new_req();
    Req &r = g_reqs[id];
    r.is_recv = true; r.peer = source; r.tag = tag; r.buf = (char *)buf; r.bytes = (uint64_t)count * type_size(t);
    *req = id;
    if (source == MPI_PROC_NULL) { r.done = true; return MPI_SUCCESS; }
    for (auto it = g_unexp.begin(); it != g_unexp.end(); ++it) {
        if (it->src == source && it->tag == tag && it->claimed_by < 0) {
            if (it->complete) {
                memcpy(buf, it->data->data(), it->data->size() < r.bytes ? it->data->size() : r.bytes);
                r.moved = it->data->size(); r.done = true;
                delete it->data;
                g_unexp.erase(it);
            } else it->claimed_by = id;
            return MPI_SUCCESS;
        }
    }
    g_posted.push_back(id);
    return MPI_SUCCESS;
}
static void finish(MPI_Request *req, MPI_Status *st) {
    Req &r = g_reqs[*req];
    if (st) { st->MPI_SOURCE = r.peer; st->MPI_TAG = r.tag; st->MPI_ERROR = MPI_SUCCESS; st->count_bytes = (int)r.moved; }
    r.active = false;
    *req = MPI_REQUEST_NULL;
}
int MPI_Wait(MPI_Request *req, MPI_Status *st) {
    if (*req == MPI_REQUEST_NULL) return MPI_SUCCESS;
    int idle = 0;
    while (!g_reqs[*req].done) { if (!progress()) spin(idle); else idle = 0; }
    finish(req, st);
    return MPI_SUCCESS;
}
int MPI_Test(MPI_Request *req, int *flag, MPI_Status *st) {
    if (*req == MPI_REQUEST_NULL) { *flag = 1; return MPI_SUCCESS; }
    progress();
    *flag = g_reqs[*req].done ? 1 : 0;
    if (*flag) finish(req, st);
    return MPI_SUCCESS;
}
int MPI_Waitall(int n, MPI_Request reqs[], MPI_Status sts[]) {
    for (int i = 0; i < n; i++) MPI_Wait(&reqs[i], sts ? &sts[i] : nullptr);
    return MPI_SUCCESS;
}
int MPI_Waitany(int n, MPI_Request reqs[], int *index, MPI_Status *st) {
    int idle = 0;
    for (;;) {
        bool any = false;
        for (int i = 0; i < n; i++) {
            if (reqs[i] == MPI_REQUEST_NULL) continue;
            any = true;
            if (g_reqs[reqs[i]].done) { *index = i; finish(&reqs[i], st); return MPI_SUCCESS; }
        }
        if (!any) { *index = MPI_UNDEFINED; return MPI_SUCCESS; }
        if (!progress()) spin(idle); else idle = 0;
    }
}
int MPI_Send(const void *buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm c) {
    MPI_Request r; MPI_Isend(buf, count, t, dest, tag, c, &r); return MPI_Wait(&r, nullptr);
}
int MPI_Recv(void *buf, int count, MPI_Datatype t, int source, int tag, MPI_Comm c, MPI_Status *st) {
    MPI_Request r; MPI_Irecv(buf, count, t, source, tag, c, &r); return MPI_Wait(&r, st);
}
int MPI_Get_count(const MPI_Status *st, MPI_Datatype t, int *count) { *count = (int)(st->count_bytes / (type_size(t) ? type_size(t) : 1)); return MPI_SUCCESS; }

int MPI_Type_size(MPI_Datatype t, int *size) { init_types(); *size = (int)type_size(t); return MPI_SUCCESS; }
int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype *nt) {
    init_types();
    g_types.push_back(DType{type_size(old) * count, type_extent(old) * count, old});
    *nt = (int)g_types.size() - 1; return MPI_SUCCESS;
}
int MPI_Type_indexed(int count, const int *bl, const int *displs, MPI_Datatype old, MPI_Datatype *nt) {
    // only gap-free layouts (the reference's particle structs) are representable here
    init_types();
    size_t n = 0; long hi = 0;
    for (int i = 0; i < count; i++) { n += bl[i]; if (displs[i] + bl[i] > hi) hi = displs[i] + bl[i]; }
    g_types.push_back(DType{n * type_size(old), (size_t)hi * type_extent(old), old});
    *nt = (int)g_types.size() - 1; return MPI_SUCCESS;
}
int MPI_Type_create_resized(MPI_Datatype old, MPI_Aint, MPI_Aint extent, MPI_Datatype *nt) {
    init_types();
    g_types.push_back(DType{type_size(old), (size_t)extent, old});
    *nt = (int)g_types.size() - 1; return MPI_SUCCESS;
}
int MPI_Type_get_extent(MPI_Datatype t, MPI_Aint *lb, MPI_Aint *extent) { init_types(); *lb = 0; *extent = (MPI_Aint)type_extent(t); return MPI_SUCCESS; }
int MPI_Type_commit(MPI_Datatype *) { return MPI_SUCCESS; }
int MPI_Type_free(MPI_Datatype *t) { *t = MPI_DATATYPE_NULL; return MPI_SUCCESS; }
int MPI_Error_string(int code, char *str, int *len) { *len = snprintf(str, MPI_MAX_ERROR_STRING, "mpi_stub error %d", code); return MPI_SUCCESS; }

}  // extern "C"
