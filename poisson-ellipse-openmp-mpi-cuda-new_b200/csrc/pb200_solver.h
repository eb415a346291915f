// pb200_solver.h -- internal C++ interface between the solver, the communicator and the C ABI.
#pragma once
#include "../../include/poisson_b200.h"
#include "pb200_kernels.cuh"
#include "pb200_tile.cuh"

#include <vector>

namespace pb200 {

extern thread_local char g_err[512];
void set_error(const char *fmt, ...);

#define CK(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e__ = (call);                                                                      \
        if (e__ != cudaSuccess) {                                                                      \
            ::pb200::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            return PB200_ERR_CUDA;                                                                     \
        }                                                                                              \
    } while (0)

void choose_process_grid(int size, int &Px, int &Py);
void decompose_2d(int M, int N, int Px, int Py, int rank, int &i0, int &i1, int &j0, int &j1);
pb200_status validate(const pb200_params &p);
Consts make_consts(const pb200_params &p);
bool coef_in_div_range(double v); // zero, or 2^-50 <= |v| <= 2^300
int pitch_for(int ny);
int cpitch_for(int ny);
pb200_status enable_kernel_smem();
struct MarchPlan {
    int ctas;    // grid size: (bands.nbA + bands.nbB) x CTA-windows
    Bands bands; // how the rows are cut
};
MarchPlan march_plan(int resident_blocks, int wn, int nx);
// tile_loop_kernel's cut of an nx x ny block for `sms` CTAs with `smem_max` bytes of dynamic shared memory each:
// false if the block does not fit. R x C = tile size, tr x tc = tiles.
bool tile_plan(int nx, int ny, int sms, size_t smem_max, int &R, int &C, int &tr, int &tc, bool lean = false);
// single-GPU blocks of at most this many cells that do not fit cell_loop_kernel run cell_loop_stream_kernel (measured
// cross-over with the marching kernels)
constexpr double LOOP_STREAM_MAX_POINTS = 1.3e6;

// The device arrays of one block. Hot (touched every iteration): w, r, p[0], p[1] (ping-pong) and one class
// byte per cell. Cold (written once by setup, read only on cells cut by the ellipse): a, b, D.
// Of the reference's d_a..d_Ap (S4:738-749), d_B is folded into the initial r, d_z and d_Ap are never stored.
struct Fields {
    Geom g{};
    size_t elems = 0;
    double *a = nullptr, *b = nullptr, *D = nullptr, *w = nullptr, *r = nullptr, *p[2] = {nullptr, nullptr};
    unsigned char *cls = nullptr; // class bytes, one 128-byte record per warp window and row (pb200_kernels.cuh)
    size_t cls_bytes = 0;
    ~Fields();
    pb200_status allocate(const Geom &geom);
    void release();
};

struct Comm; // pb200_comm.cu

struct Solver {
    pb200_params prm{};
    Consts c{};
    Fields f;
    int dev = -1;
    int Px = 1, Py = 1, i0 = 1, i1 = 1, j0 = 1, j1 = 1;
    cudaStream_t stream = nullptr;
    Ctl *ctl = nullptr;
    Ctl *h_ctl = nullptr; // pinned, 2 slots
    double *part1 = nullptr, *part2 = nullptr, *trace = nullptr, *norms = nullptr;
    cudaEvent_t ev_start = nullptr, ev_stop = nullptr, ev_call0 = nullptr, ev_call1 = nullptr, ev_poll[2] = {nullptr, nullptr};
    cudaGraphExec_t graph_exec = nullptr;
    dim3 k1_grid, k2_grid, norms_grid, init_grid;
    bool safe_div = false; // coefficients outside div_const's verified range: IEEE divides everywhere
    int wn = 1; // warp windows per row
    MarchPlan k1_plan{}, k2_plan{};
    bool use_loop = false; // small block: the whole loop is one cooperative launch (cell_loop_kernel)
    dim3 loop_grid;
    int loop_cells = 1; // cells per thread of cell_loop_kernel; 0: cell_loop_stream_kernel
    // blocks that fit in shared memory: the whole loop in one cooperative launch of tile_loop_kernel, one or several GPUs
    bool use_tile = false;
    int tile_R = 0, tile_C = 0, tile_tr = 0, tile_tc = 0;
    size_t tile_smem = 0;
    unsigned long long *tile_slots = nullptr; // the CTAs' partial sums as tagged words
    unsigned long long *tile_r2 = nullptr;    // tagged copy of r (2 words per cell): the halo of the resident loop kernel
    bool tile_feasible(int sms, size_t smem_max, int &R, int &C, bool lean) const; // on the largest block: the same answer on every rank
    bool tile_lean = false; // r and A p' / z in global memory (blocks of 1.1-2.9 M cells)
    long long launches = 0;
    long long last_iter = 0;
    double t_setup = 0.0;
    bool solved = false;
    bool coefficients_ready = false;
    Comm *comm = nullptr;
    // peer-memory mode (pb200_comm.cu): mailbox of this rank, views of the peers
    double *mail = nullptr;
    PeerView pv{};
    bool peer_ready = false;
    std::vector<void *> ipc_opened;

    ~Solver();
    pb200_status create(const pb200_params &params);
    pb200_status run_setup();
    void plan_launch();
    void plan_tile(int sms);
    void launch_iteration(long long k, bool first);
    void launch_k1(const K1Args &a, bool first);
    void launch_k2(const K2Args &a);
    K1Args k1_args(long long k) const;
    K2Args k2_args(long long k) const;
    void launch_init();
    int kernels_per_iteration() const;
    pb200_status solve(pb200_result *res);
    pb200_status profile_iteration(int iters, double ms_out[2]);
    pb200_status get_solution(double *host, int64_t ld);
    pb200_status get_trace(double *host, int n, int *rows);

    // multi-GPU (pb200_comm.cu)
    pb200_status attach_comm(const uint8_t id[128]);
    bool comm_ready() const;
    bool comm_failed() const; // an NCCL call of the loop returned an error
    void destroy_comm();
    void allreduce(int first, int count); // in place on ctl->red[first .. first+count)
    void exchange_r_halo();               // boundary rows/columns of r -> neighbours' ring
    int halo_kernels_per_iteration() const;
    pb200_status reduce_norms_host(double h_norms[4]);
    pb200_status peer_export(uint8_t blob[256]);
    pb200_status peer_attach(const uint8_t *blobs, int nranks);
    void peer_release();
};

pb200_status comm_unique_id(uint8_t id_out[128]);

} // namespace pb200
