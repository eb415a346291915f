/*
 * mpi.h -- ranks-as-threads stand-in for the handful of MPI calls the reference uses.
 *
 * TEST INFRASTRUCTURE ONLY (lives under oracle/). It exists so that the reference's
 * MPI sources (stage2/3/4) compile UNMODIFIED in an image that ships no MPI:
 *
 *     g++ -O3 -fopenmp -Ioracle/shim  /root/reference/stage3-openmp+mpi/hybrid.cpp
 *
 * Every "rank" is a std::thread of one process (see shim_mpi_launch below); with no
 * launcher the program is a plain 1-rank job. Only what the reference calls is here:
 *   Init/Finalize/Comm_rank/Comm_size/Barrier/Wtime            (hybrid.cpp:474-500)
 *   Allreduce/Reduce on MPI_DOUBLE with MPI_SUM / MPI_MAX      (hybrid.cpp:404-448,
 *                                                               poisson_mpi_cuda_f.cu:843-967)
 *   Isend/Irecv/Waitall                                        (hybrid.cpp:291-336)
 *   Sendrecv, MPI_PROC_NULL                                    (poisson_mpi_cuda_f.cu:408-425)
 * Reductions combine contributions in rank order 0..P-1, so results are deterministic.
 */
#ifndef ORACLE_SHIM_MPI_H
#define ORACLE_SHIM_MPI_H

#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
struct MPI_Status { int MPI_SOURCE, MPI_TAG, MPI_ERROR; };

#define MPI_COMM_WORLD 0
#define MPI_DOUBLE 8 /* value doubles as the element size */
#define MPI_INT 4
#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_PROC_NULL (-2)
#define MPI_SUCCESS 0
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_STATUSES_IGNORE ((MPI_Status *)0)

namespace shim_mpi {

struct Message {
    int src, dst, tag;
    std::vector<char> bytes;
};

struct World {
    int size = 1;
    std::mutex mu;
    std::condition_variable cv;
    /* barrier */
    int arrived = 0;
    long generation = 0;
    /* reduction scratch: one slot per rank */
    std::vector<std::vector<double>> slots;
    /* mailbox */
    std::deque<Message> mail;
};

inline World &world() {
    static World w;
    return w;
}
inline int &my_rank() {
    static thread_local int r = 0;
    return r;
}

inline void barrier() {
    World &w = world();
    if (w.size == 1) return;
    std::unique_lock<std::mutex> lk(w.mu);
    long gen = w.generation;
    if (++w.arrived == w.size) {
        w.arrived = 0;
        ++w.generation;
        w.cv.notify_all();
    } else {
        w.cv.wait(lk, [&] { return w.generation != gen; });
    }
}

inline void combine(const void *sendbuf, void *recvbuf, int count, MPI_Op op, int root /* -1 = all */) {
    World &w = world();
    const double *in = static_cast<const double *>(sendbuf);
    double *out = static_cast<double *>(recvbuf);
    if (w.size == 1) {
        if (out != in) std::memcpy(out, in, sizeof(double) * count);
        return;
    }
    {
        std::lock_guard<std::mutex> lk(w.mu);
        if ((int)w.slots.size() != w.size) w.slots.assign(w.size, std::vector<double>());
        w.slots[my_rank()].assign(in, in + count);
    }
    barrier();
    if (root < 0 || root == my_rank()) {
        for (int c = 0; c < count; ++c) {
            double acc = w.slots[0][c];
            for (int r = 1; r < w.size; ++r) {
                double v = w.slots[r][c];
                acc = (op == MPI_MAX) ? (v > acc ? v : acc) : acc + v;
            }
            out[c] = acc;
        }
    }
    barrier();
}

inline void post(const void *buf, int bytes, int dst, int tag) {
    World &w = world();
    Message m;
    m.src = my_rank();
    m.dst = dst;
    m.tag = tag;
    m.bytes.assign(static_cast<const char *>(buf), static_cast<const char *>(buf) + bytes);
    {
        std::lock_guard<std::mutex> lk(w.mu);
        w.mail.push_back(std::move(m));
    }
    w.cv.notify_all();
}

inline void take(void *buf, int bytes, int src, int tag) {
    World &w = world();
    std::unique_lock<std::mutex> lk(w.mu);
    for (;;) {
        for (auto it = w.mail.begin(); it != w.mail.end(); ++it) {
            if (it->dst == my_rank() && it->src == src && it->tag == tag) {
                std::memcpy(buf, it->bytes.data(), (size_t)bytes < it->bytes.size() ? (size_t)bytes : it->bytes.size());
                w.mail.erase(it);
                return;
            }
        }
        w.cv.wait(lk);
    }
}

struct Request {
    bool is_recv = false;
    void *buf = nullptr;
    int bytes = 0, peer = 0, tag = 0;
};

} // namespace shim_mpi

typedef shim_mpi::Request MPI_Request;

/* Run fn(rank) on nranks threads that see each other as MPI ranks. */
inline void shim_mpi_launch(int nranks, const std::function<void(int)> &fn) {
    shim_mpi::World &w = shim_mpi::world();
    w.size = nranks;
    w.arrived = 0;
    w.mail.clear();
    w.slots.clear();
    std::vector<std::thread> threads;
    for (int r = 0; r < nranks; ++r)
        threads.emplace_back([r, &fn] {
            shim_mpi::my_rank() = r;
            fn(r);
        });
    for (auto &t : threads) t.join();
    w.size = 1;
}

inline int MPI_Init(int *, char ***) { return MPI_SUCCESS; }
inline int MPI_Finalize() { return MPI_SUCCESS; }
inline int MPI_Comm_rank(MPI_Comm, int *rank) { *rank = shim_mpi::my_rank(); return MPI_SUCCESS; }
inline int MPI_Comm_size(MPI_Comm, int *size) { *size = shim_mpi::world().size; return MPI_SUCCESS; }
inline int MPI_Barrier(MPI_Comm) { shim_mpi::barrier(); return MPI_SUCCESS; }
inline double MPI_Wtime() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
inline int MPI_Allreduce(const void *s, void *r, int count, MPI_Datatype, MPI_Op op, MPI_Comm) {
    shim_mpi::combine(s, r, count, op, -1);
    return MPI_SUCCESS;
}
inline int MPI_Reduce(const void *s, void *r, int count, MPI_Datatype, MPI_Op op, int root, MPI_Comm) {
    shim_mpi::combine(s, r, count, op, root);
    return MPI_SUCCESS;
}
inline int MPI_Isend(const void *buf, int count, MPI_Datatype dt, int dst, int tag, MPI_Comm, MPI_Request *req) {
    *req = MPI_Request();
    if (dst != MPI_PROC_NULL) shim_mpi::post(buf, count * dt, dst, tag); /* eager: never blocks */
    return MPI_SUCCESS;
}
inline int MPI_Irecv(void *buf, int count, MPI_Datatype dt, int src, int tag, MPI_Comm, MPI_Request *req) {
    *req = MPI_Request();
    req->is_recv = (src != MPI_PROC_NULL);
    req->buf = buf;
    req->bytes = count * dt;
    req->peer = src;
    req->tag = tag;
    return MPI_SUCCESS;
}
inline int MPI_Waitall(int n, MPI_Request *reqs, MPI_Status *) {
    for (int i = 0; i < n; ++i)
        if (reqs[i].is_recv) {
            shim_mpi::take(reqs[i].buf, reqs[i].bytes, reqs[i].peer, reqs[i].tag);
            reqs[i].is_recv = false;
        }
    return MPI_SUCCESS;
}
inline int MPI_Sendrecv(const void *sbuf, int scount, MPI_Datatype sdt, int dst, int stag, void *rbuf, int rcount,
                        MPI_Datatype rdt, int src, int rtag, MPI_Comm, MPI_Status *) {
    if (dst != MPI_PROC_NULL) shim_mpi::post(sbuf, scount * sdt, dst, stag);
    if (src != MPI_PROC_NULL) shim_mpi::take(rbuf, rcount * rdt, src, rtag);
    return MPI_SUCCESS;
}

#endif /* ORACLE_SHIM_MPI_H */
