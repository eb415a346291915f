// pb200_comm.cu -- multi-GPU plumbing: one rank per GPU, NCCL over NVLink/NVSwitch.
//
// Replaces the reference's host-staged MPI traffic (stage4-mpi+cuda/poisson_mpi_cuda_f.cu, "S4"):
//   exchange_halos_2d_gpu (S4:331-500: 4 D2H copies, 4 MPI_Sendrecv, 4 H2D copies / memsets, all
//   blocking)            -> one grouped ncclSend/ncclRecv between device buffers, on the solver stream;
//   3 x MPI_Allreduce of one double per iteration (S4:872, 893, 926)
//                        -> 2 x ncclAllReduce on the device-resident scalars (1 and 2 doubles).
// What travels is the boundary of r (after K2), not of p: K1 recomputes p on the ring from r, the
// diagonal class and its own previous ring values, so one exchange per iteration suffices.
//
// NCCL is loaded with dlopen at first use: single-GPU users need no NCCL at all, and inside a
// Python process that already imported torch the same libnccl.so.2 is shared.
#include "pb200_solver.h"

#include <cstring>
#include <dlfcn.h>
#include <unistd.h>
#include <mutex>
#include <nccl.h>

namespace pb200 {

namespace {

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

NcclApi &nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {
            api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) return;
        auto sym = [&](const char *s) { return dlsym(api.handle, s); };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
        api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
        api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
        api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
        api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce && api.Send && api.Recv &&
                 api.GroupStart && api.GroupEnd && api.GetErrorString;
    });
    return api;
}

#define NK(call)                                                                                      \
    do {                                                                                              \
        ncclResult_t r__ = (call);                                                                    \
        if (r__ != ncclSuccess) {                                                                     \
            ::pb200::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, nccl().GetErrorString(r__)); \
            return PB200_ERR_COMM;                                                                    \
        }                                                                                             \
    } while (0)

// column lj of r (rows 1..nx) <-> contiguous buffer
__global__ void pack_cols_kernel(const double *__restrict__ r, int nx, int pitch, int ny, double *__restrict__ lo,
                                 double *__restrict__ hi) {
    const int li = 1 + blockIdx.x * blockDim.x + threadIdx.x;
    if (li > nx) return;
    if (lo) lo[li - 1] = r[(size_t)li * pitch + 1];
    if (hi) hi[li - 1] = r[(size_t)li * pitch + ny];
}
__global__ void unpack_cols_kernel(double *__restrict__ r, int nx, int pitch, int ny, const double *__restrict__ lo,
                                   const double *__restrict__ hi) {
    const int li = 1 + blockIdx.x * blockDim.x + threadIdx.x;
    if (li > nx) return;
    if (lo) r[(size_t)li * pitch + 0] = lo[li - 1];
    if (hi) r[(size_t)li * pitch + ny + 1] = hi[li - 1];
}

} // namespace

struct Comm {
    ncclComm_t nc = nullptr;
    int left = -1, right = -1, down = -1, up = -1; // S4:341-344
    double *send_lo = nullptr, *send_hi = nullptr, *recv_lo = nullptr, *recv_hi = nullptr; // packed columns
    ncclResult_t async_err = ncclSuccess;
};

pb200_status comm_unique_id(uint8_t id_out[128]) {
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    if (!nccl().ok) { set_error("NCCL not available: dlopen(libnccl.so.2) failed"); return PB200_ERR_COMM; }
    ncclUniqueId id;
    NK(nccl().GetUniqueId(&id));
    std::memcpy(id_out, &id, 128);
    return PB200_OK;
}

pb200_status Solver::attach_comm(const uint8_t id_bytes[128]) {
    if (prm.nranks <= 1) return PB200_OK;
    if (!nccl().ok) { set_error("NCCL not available: dlopen(libnccl.so.2) failed"); return PB200_ERR_COMM; }
    CK(cudaSetDevice(dev));
    destroy_comm();
    comm = new Comm();
    ncclUniqueId id;
    std::memcpy(&id, id_bytes, 128);
    NK(nccl().CommInitRank(&comm->nc, prm.nranks, id, prm.rank));
    const int px = prm.rank % Px, py = prm.rank / Px; // S4:338-344
    comm->left = px > 0 ? prm.rank - 1 : -1;
    comm->right = px < Px - 1 ? prm.rank + 1 : -1;
    comm->down = py > 0 ? prm.rank - Px : -1;
    comm->up = py < Py - 1 ? prm.rank + Px : -1;
    const size_t nb = static_cast<size_t>(f.g.nx) * sizeof(double);
    if (comm->down >= 0) { CK(cudaMalloc(&comm->send_lo, nb)); CK(cudaMalloc(&comm->recv_lo, nb)); }
    if (comm->up >= 0) { CK(cudaMalloc(&comm->send_hi, nb)); CK(cudaMalloc(&comm->recv_hi, nb)); }
    if (graph_exec) { cudaGraphExecDestroy(graph_exec); graph_exec = nullptr; }
    return PB200_OK;
}

bool Solver::comm_ready() const { return comm && comm->nc; }
bool Solver::comm_failed() const { return comm && comm->async_err != ncclSuccess; }

void Solver::destroy_comm() {
    if (!comm) return;
    if (comm->nc && nccl().ok) nccl().CommDestroy(comm->nc);
    for (double *q : {comm->send_lo, comm->send_hi, comm->recv_lo, comm->recv_hi})
        if (q) cudaFree(q);
    delete comm;
    comm = nullptr;
}

void Solver::allreduce(int first, int count) {
    ncclResult_t r = nccl().AllReduce(ctl->red + first, ctl->red + first, count, ncclDouble, ncclSum, comm->nc, stream);
    if (r != ncclSuccess && comm->async_err == ncclSuccess) comm->async_err = r;
}

int Solver::halo_kernels_per_iteration() const {
    return (comm && (comm->down >= 0 || comm->up >= 0)) ? 2 : 0;
}

void Solver::exchange_r_halo() {
    const Geom &g = f.g;
    const bool cols = comm->down >= 0 || comm->up >= 0;
    const int tb = 256, nbk = (g.nx + tb - 1) / tb;
    if (cols) pack_cols_kernel<<<nbk, tb, 0, stream>>>(f.r, g.nx, g.pitch, g.ny, comm->send_lo, comm->send_hi);
    NcclApi &n = nccl();
    ncclResult_t r = n.GroupStart();
    auto acc = [&](ncclResult_t x) { if (x != ncclSuccess && r == ncclSuccess) r = x; };
    const size_t P = g.pitch;
    // i-direction neighbours share whole rows: contiguous, sent in place (interior cells lj = 1..ny)
    if (comm->left >= 0) {
        acc(n.Send(f.r + 1 * P + 1, g.ny, ncclDouble, comm->left, comm->nc, stream));
        acc(n.Recv(f.r + 0 * P + 1, g.ny, ncclDouble, comm->left, comm->nc, stream));
    }
    if (comm->right >= 0) {
        acc(n.Send(f.r + (size_t)g.nx * P + 1, g.ny, ncclDouble, comm->right, comm->nc, stream));
        acc(n.Recv(f.r + (size_t)(g.nx + 1) * P + 1, g.ny, ncclDouble, comm->right, comm->nc, stream));
    }
    // j-direction neighbours share columns: packed
    if (comm->down >= 0) {
        acc(n.Send(comm->send_lo, g.nx, ncclDouble, comm->down, comm->nc, stream));
        acc(n.Recv(comm->recv_lo, g.nx, ncclDouble, comm->down, comm->nc, stream));
    }
    if (comm->up >= 0) {
        acc(n.Send(comm->send_hi, g.nx, ncclDouble, comm->up, comm->nc, stream));
        acc(n.Recv(comm->recv_hi, g.nx, ncclDouble, comm->up, comm->nc, stream));
    }
    acc(n.GroupEnd());
    if (r != ncclSuccess && comm->async_err == ncclSuccess) comm->async_err = r;
    if (cols) unpack_cols_kernel<<<nbk, tb, 0, stream>>>(f.r, g.nx, g.pitch, g.ny, comm->recv_lo, comm->recv_hi);
}

// ---------------------------------------------------------------------------------------------
// Peer-memory mode. Every rank exports where its mailbox and its r array live; after the host has handed every rank
// all exports (MPI_Allgather / torch all_gather / shared memory between threads), peer_attach maps them: CUDA IPC
// for ranks in other processes, plain pointers plus peer access for ranks that are threads of this process.
namespace {
struct PeerBlob {
    int magic, pid, dev, rank;
    int nx, ny, pitch, pad;
    void *r_ptr, *mail_ptr;
    cudaIpcMemHandle_t h_r, h_mail;
    void *t_ptr;            // tagged copy of r (resident loop kernel), null when the blocks do not fit that kernel
    cudaIpcMemHandle_t h_t;
};
static_assert(sizeof(PeerBlob) <= 256, "blob fits the 256 bytes of the C ABI");
} // namespace

pb200_status Solver::peer_export(uint8_t blob[256]) {
    if (prm.nranks <= 1) { set_error("peer mode needs nranks > 1"); return PB200_ERR_INVALID; }
    if (prm.nranks > MAX_RANKS) { set_error("peer mode supports up to %d ranks", MAX_RANKS); return PB200_ERR_TOO_LARGE; }
    CK(cudaSetDevice(dev));
    if (!mail) {
        CK(cudaMalloc(&mail, MAIL_DOUBLES * sizeof(double)));
        CK(cudaMemset(mail, 0, MAIL_DOUBLES * sizeof(double)));
    }
    PeerBlob b{};
    b.magic = 0x50423230;
    b.pid = static_cast<int>(getpid());
    b.dev = dev;
    b.rank = prm.rank;
    b.nx = f.g.nx;
    b.ny = f.g.ny;
    b.pitch = f.g.pitch;
    b.r_ptr = f.r;
    b.mail_ptr = mail;
    CK(cudaIpcGetMemHandle(&b.h_r, f.r));
    CK(cudaIpcGetMemHandle(&b.h_mail, mail));
    b.t_ptr = tile_r2;
    if (tile_r2) CK(cudaIpcGetMemHandle(&b.h_t, tile_r2));
    std::memset(blob, 0, 256);
    std::memcpy(blob, &b, sizeof(b));
    return PB200_OK;
}

pb200_status Solver::peer_attach(const uint8_t *blobs, int nranks) {
    if (nranks != prm.nranks || nranks > MAX_RANKS || !mail) { set_error("peer_attach: export first, pass nranks blobs"); return PB200_ERR_INVALID; }
    CK(cudaSetDevice(dev));
    peer_release();
    const int me = static_cast<int>(getpid());
    PeerBlob all[MAX_RANKS];
    for (int q = 0; q < nranks; ++q) {
        std::memcpy(&all[q], blobs + 256 * q, sizeof(PeerBlob));
        if (all[q].magic != 0x50423230 || all[q].rank != q) { set_error("peer_attach: blob %d is not rank %d's export", q, q); return PB200_ERR_INVALID; }
    }
    auto map = [&](const PeerBlob &b, bool want_r, double **mail_out, double **r_out, unsigned long long **t_out) -> pb200_status {
        *r_out = nullptr;
        *t_out = nullptr;
        if (b.rank == prm.rank) {
            *mail_out = mail;
            *r_out = f.r;
            *t_out = tile_r2;
        } else if (b.pid == me) { // a thread of this process: same address space
            if (b.dev != dev) {
                cudaError_t e = cudaDeviceEnablePeerAccess(b.dev, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CK(e);
                cudaGetLastError();
            }
            *mail_out = static_cast<double *>(b.mail_ptr);
            *r_out = static_cast<double *>(b.r_ptr);
            *t_out = static_cast<unsigned long long *>(b.t_ptr);
        } else {
            void *pm = nullptr, *pr = nullptr;
            CK(cudaIpcOpenMemHandle(&pm, b.h_mail, cudaIpcMemLazyEnablePeerAccess));
            ipc_opened.push_back(pm);
            *mail_out = static_cast<double *>(pm);
            if (want_r) {
                CK(cudaIpcOpenMemHandle(&pr, b.h_r, cudaIpcMemLazyEnablePeerAccess));
                ipc_opened.push_back(pr);
                *r_out = static_cast<double *>(pr);
                if (b.t_ptr) {
                    void *pt = nullptr;
                    CK(cudaIpcOpenMemHandle(&pt, b.h_t, cudaIpcMemLazyEnablePeerAccess));
                    ipc_opened.push_back(pt);
                    *t_out = static_cast<unsigned long long *>(pt);
                }
            }
        }
        return PB200_OK;
    };
    const int px = prm.rank % Px, py = prm.rank / Px; // S4:338-344
    const int left = px > 0 ? prm.rank - 1 : -1, right = px < Px - 1 ? prm.rank + 1 : -1;
    const int down = py > 0 ? prm.rank - Px : -1, up = py < Py - 1 ? prm.rank + Px : -1;
    pv = PeerView{};
    pv.nranks = nranks;
    pv.rank = prm.rank;
    for (int q = 0; q < nranks; ++q) {
        const bool nb = (q == left || q == right || q == down || q == up);
        double *rq = nullptr;
        unsigned long long *tq = nullptr;
        pb200_status st = map(all[q], nb, &pv.mail[q], &rq, &tq);
        if (st != PB200_OK) return st;
        if (q == left) { pv.r_left = rq; pv.t_left = tq; pv.nx_left = all[q].nx; }
        if (q == right) { pv.r_right = rq; pv.t_right = tq; }
        if (q == down) { pv.r_down = rq; pv.t_down = tq; pv.ny_down = all[q].ny; pv.pitch_down = all[q].pitch; }
        if (q == up) { pv.r_up = rq; pv.t_up = tq; pv.pitch_up = all[q].pitch; }
    }
    peer_ready = true;
    c.multi = 2;
    plan_launch(); // blocks that fit in shared memory now run the resident loop kernel across the GPUs
    if (graph_exec) { cudaGraphExecDestroy(graph_exec); graph_exec = nullptr; }
    return PB200_OK;
}

void Solver::peer_release() {
    for (void *p : ipc_opened) cudaIpcCloseMemHandle(p);
    ipc_opened.clear();
    peer_ready = false;
    if (c.multi == 2) {
        c.multi = 1;
        use_tile = false;
    }
}

pb200_status Solver::reduce_norms_host(double *) {
    if (comm->async_err != ncclSuccess) { // the loop stopped launching when it saw this
        set_error("NCCL error during the loop: %s", nccl().GetErrorString(comm->async_err));
        comm->async_err = ncclSuccess;
        return PB200_ERR_COMM;
    }
    // in-stream: sums of w^2, r^2, (w-u)^2 and the maxima of |w|, |w-u| (bit pattern of a non-negative double orders like uint64)
    NK(nccl().AllReduce(norms, norms, 3, ncclDouble, ncclSum, comm->nc, stream));
    NK(nccl().AllReduce(norms + 3, norms + 3, 2, ncclUint64, ncclMax, comm->nc, stream));
    if (comm->async_err != ncclSuccess) {
        set_error("NCCL error during the loop: %s", nccl().GetErrorString(comm->async_err));
        return PB200_ERR_COMM;
    }
    return PB200_OK;
}

} // namespace pb200
