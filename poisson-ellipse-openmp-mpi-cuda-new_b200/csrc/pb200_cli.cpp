// pb200_cli.cpp -- `poisson_b200 [M N] [flags]`: drop-in for the reference binary
// `mpirun -np P ./poisson_mpi_cuda_f [M N]` (stage4-mpi+cuda/poisson_mpi_cuda_f.cu "S4", main at 986-1039).
//
// Same positional arguments (S4:996-1000: used only when both are given, parsed with atoi, default 40 40),
// same stdout lines in the same order and formatting (S4:1002-1003, 901-903, 969-980, 1027-1035), so a
// parser of the reference's output keeps working. One process drives P GPUs (--gpus P, one host thread
// per GPU = one "process" of the banner). Everything the reference hard-codes is a flag; the solution
// dump the reference never had is --dump.
#include "../../include/poisson_b200.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <string>
#include <atomic>
#include <thread>
#include <vector>

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static void usage() {
    std::cerr <<
        "usage: poisson_b200 [M N] [--delta D] [--max-iter K] [--eps E] [--ellipse-k K] [--ellipse-ax A] [--rect A1 B1 A2 B2] [--f F]\n"
        "                    [--rhs indicator|area] [--report-error]\n"
        "                    [--norm weighted|stage0] [--gpus P] [--grid PXxPY] [--device D] [--fixed-iters K]\n"
        "                    [--trace K] [--check-every K] [--no-graph] [--comm peer|nccl] [--dump FILE(.npy|.txt|raw)] [--json FILE|-]\n"
        "Solves -div(k grad u) = f on the rectangle with the ellipse (x/ax)^2+(k y)^2<1 as fictitious domain (PCG).\n"
        "--report-error adds a line with max|w-u| and the L2 error against the analytic solution inside the ellipse.\n"
        "With no flags it behaves as the reference: `poisson_b200 M N`.\n";
}

static bool write_dump(const std::string &path, const std::vector<double> &w, int M, int N, const pb200_params &p) {
    const size_t rows = M + 1, cols = N + 1;
    auto ends = [&](const char *s) { return path.size() >= strlen(s) && path.compare(path.size() - strlen(s), strlen(s), s) == 0; };
    if (ends(".txt")) {
        std::ofstream f(path);
        if (!f) return false;
        f << "# i j x y w\n" << std::setprecision(17);
        const double h1 = (p.B1 - p.A1) / M, h2 = (p.B2 - p.A2) / N;
        for (size_t i = 0; i < rows; ++i)
            for (size_t j = 0; j < cols; ++j)
                f << i << ' ' << j << ' ' << p.A1 + i * h1 << ' ' << p.A2 + j * h2 << ' ' << w[i * cols + j] << '\n';
        return bool(f);
    }
    std::ofstream f(path, std::ios::binary);
    if (!f) return false;
    if (ends(".npy")) { // NPY v1.0 header, C order float64
        std::string dict = "{'descr': '<f8', 'fortran_order': False, 'shape': (" + std::to_string(rows) + ", " + std::to_string(cols) + "), }";
        size_t total = 10 + dict.size() + 1;
        dict.append((64 - total % 64) % 64, ' ');
        dict.push_back('\n');
        const unsigned short hl = static_cast<unsigned short>(dict.size());
        f.write("\x93NUMPY\x01\x00", 8);
        f.write(reinterpret_cast<const char *>(&hl), 2);
        f.write(dict.data(), dict.size());
    }
    f.write(reinterpret_cast<const char *>(w.data()), w.size() * sizeof(double));
    return bool(f);
}

int main(int argc, char **argv) {
    const double t_program_start = now_s(); // S4:994
    pb200_params prm;
    pb200_default_params(&prm);
    int gpus = 1, first_device = 0;
    bool peer_mode = true; // multi-GPU: halos and sums through peer memory inside the kernels (else NCCL calls per iteration)
    bool report_error = false;
    std::string dump, json;
    std::vector<const char *> pos;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto need = [&](int n) {
            if (i + n >= argc) { usage(); std::exit(2); }
        };
        if (a == "--delta") { need(1); prm.delta = atof(argv[++i]); }
        else if (a == "--max-iter") { need(1); prm.max_iter = atoll(argv[++i]); }
        else if (a == "--eps") { need(1); prm.eps = atof(argv[++i]); }
        else if (a == "--ellipse-k") { need(1); prm.ell_k = atof(argv[++i]); }
        else if (a == "--f") { need(1); prm.f_val = atof(argv[++i]); }
        else if (a == "--ellipse-ax") { need(1); prm.ell_ax = atof(argv[++i]); }
        else if (a == "--rhs") { need(1); prm.rhs_mode = std::string(argv[++i]) == "area" ? PB200_RHS_AREA : PB200_RHS_INDICATOR; }
        else if (a == "--report-error") report_error = true;
        else if (a == "--rect") { need(4); prm.A1 = atof(argv[++i]); prm.B1 = atof(argv[++i]); prm.A2 = atof(argv[++i]); prm.B2 = atof(argv[++i]); }
        else if (a == "--norm") { need(1); std::string v = argv[++i]; prm.norm_mode = (v == "stage0" || v == "unweighted") ? PB200_NORM_STAGE0 : PB200_NORM_WEIGHTED; }
        else if (a == "--gpus") { need(1); gpus = atoi(argv[++i]); }
        else if (a == "--grid") { need(1); if (sscanf(argv[++i], "%dx%d", &prm.Px, &prm.Py) != 2) { usage(); return 2; } }
        else if (a == "--device") { need(1); first_device = atoi(argv[++i]); }
        else if (a == "--fixed-iters") { need(1); prm.fixed_iters = atoi(argv[++i]); }
        else if (a == "--trace") { need(1); prm.trace_len = atoi(argv[++i]); }
        else if (a == "--check-every") { need(1); prm.check_every = atoi(argv[++i]); }
        else if (a == "--no-graph") prm.use_graph = 0;
        else if (a == "--comm") { need(1); peer_mode = std::string(argv[++i]) != "nccl"; }
        else if (a == "--dump") { need(1); dump = argv[++i]; }
        else if (a == "--json") { need(1); json = argv[++i]; }
        else if (a == "-h" || a == "--help") { usage(); return 0; }
        else if (a.rfind("--", 0) == 0) { std::cerr << "unknown flag " << a << "\n"; usage(); return 2; }
        else pos.push_back(argv[i]);
    }
    if (pos.size() >= 2) { // S4:997-1000
        prm.M = std::atoi(pos[0]);
        prm.N = std::atoi(pos[1]);
    }
    if (gpus < 1) gpus = 1;
    if (prm.Px > 0 && prm.Py > 0) gpus = prm.Px * prm.Py;
    const int M = prm.M, N = prm.N;
    std::cout << "MPI + CUDA 2D run with " << gpus << " processes; " << "M=" << M << ", N=" << N << std::endl; // S4:1002-1003

    const int ndev = pb200_device_count();
    if (ndev < 1) {
        std::cerr << "poisson_b200: no CUDA device (this solver has no CPU path)\n";
        return 1;
    }
    if (first_device + gpus > ndev) {
        std::cerr << "poisson_b200: " << gpus << " GPUs requested from device " << first_device << ", " << ndev << " present\n";
        return 1;
    }

    const double t_before_solver = now_s(); // S4:1011
    std::vector<pb200_solver *> solvers(gpus, nullptr);
    std::vector<pb200_result> results(gpus);
    std::vector<int> status(gpus, 0);
    std::vector<std::string> errors(gpus);
    std::vector<double> w;
    if (!dump.empty()) w.assign(static_cast<size_t>(M + 1) * (N + 1), 0.0);
    uint8_t uid[128] = {0};
    if (gpus > 1 && pb200_comm_unique_id(uid) != PB200_OK) {
        std::cerr << "poisson_b200: " << pb200_last_error() << "\n";
        return 1;
    }
    std::vector<uint8_t> blobs(static_cast<size_t>(gpus) * 256);
    std::atomic<int> created{0}, exported{0}, failed{0};
    auto rank_main = [&](int rank) {
        pb200_params p = prm;
        p.nranks = gpus;
        p.rank = rank;
        p.device = first_device + rank;
        auto fail = [&](pb200_status st) {
            failed = 1;
            status[rank] = st;
            errors[rank] = pb200_last_error();
        };
        // Every rank reaches each meeting point whether its own step worked or not, and all of them take the same decision
        // there: a rank that failed never leaves the others waiting inside a collective.
        auto meet = [&](std::atomic<int> &counter) {
            ++counter;
            while (counter.load() < gpus) std::this_thread::yield();
            return failed.load() == 0;
        };
        pb200_status st = pb200_create(&p, &solvers[rank]);
        if (st != PB200_OK) fail(st);
        if (gpus > 1 && !meet(created)) {
            if (status[rank] == PB200_OK) { status[rank] = PB200_ERR_STATE; errors[rank] = "another rank failed during set-up"; }
            return;
        }
        if (st != PB200_OK) return;
        if (gpus > 1 && (st = pb200_attach_comm(solvers[rank], uid)) != PB200_OK) return fail(st);
        if (gpus > 1 && peer_mode) { // every rank publishes where its mailbox / residual live, then maps the others'
            st = pb200_peer_export(solvers[rank], blobs.data() + 256 * rank);
            if (st != PB200_OK) fail(st);
            if (!meet(exported)) {
                if (status[rank] == PB200_OK) { status[rank] = PB200_ERR_COMM; errors[rank] = "another rank failed during set-up"; }
                return;
            }
            if ((st = pb200_peer_attach(solvers[rank], blobs.data(), gpus)) != PB200_OK) return fail(st);
        }
        if ((st = pb200_solve(solvers[rank], &results[rank])) != PB200_OK) return fail(st);
        if (!w.empty() && (st = pb200_get_solution(solvers[rank], w.data(), N + 1)) != PB200_OK) return fail(st);
    };
    if (gpus == 1) rank_main(0);
    else {
        std::vector<std::thread> th;
        for (int r = 0; r < gpus; ++r) th.emplace_back(rank_main, r);
        for (auto &t : th) t.join();
    }
    for (int r = 0; r < gpus; ++r)
        if (status[r] != PB200_OK) {
            std::cerr << "poisson_b200: rank " << r << ": " << pb200_status_string(static_cast<pb200_status>(status[r])) << ": " << errors[r] << "\n";
            return 1;
        }
    const pb200_result &res = results[0];
    double t_loop = 0.0, t_setup = 0.0;
    for (const auto &r : results) { t_loop = std::max(t_loop, r.t_solve_s); t_setup = std::max(t_setup, r.t_setup_s); }

    if (res.stop_reason == PB200_STOP_CONVERGED) // S4:900-904
        std::cout << "Converged after " << res.iterations << " iterations (||w(k+1)-w(k)|| < " << prm.delta << ").\n";
    // S4:969-980. The reference's buckets are wall-clock sums around blocking calls; here the loop never
    // blocks, so the whole device loop is reported as GPU compute and the host-staged buckets are 0.
    std::cout << "   GPU compute time (Ap + D^{-1}r, max over ranks) ~ " << t_loop << " s\n";
    std::cout << "   Host<->Device copy time (max over ranks)        ~ " << 0.0 << " s\n";
    std::cout << "   MPI halo exchange time (max over ranks)         ~ " << 0.0 << " s\n";
    std::cout << "   Preconditioner CPU part time (max over ranks)   ~ " << 0.0 << " s\n";
    std::cout << "   Dot products time (max over ranks)              ~ " << 0.0 << " s\n";
    for (auto *s : solvers) pb200_destroy(s);
    const double t_after_solver = now_s(); // S4:1017

    int rc = 0;
    if (!dump.empty() && !write_dump(dump, w, M, N, prm)) {
        std::cerr << "poisson_b200: cannot write " << dump << "\n";
        rc = 1;
    }
    const double t_before_finalize = now_s();
    const double time_init = t_before_solver - t_program_start;      // S4:1022
    const double time_solver = t_after_solver - t_before_solver;     // S4:1023
    const double time_finalize = t_before_finalize - t_after_solver; // S4:1024
    const double elapsed = t_before_finalize - t_program_start;      // S4:1025
    std::cout << "M=" << M << ", N=" << N << " | Iter=" << res.iterations << " | Total Time=" << std::fixed
              << std::setprecision(6) << elapsed << " s\n"; // S4:1028-1031
    std::cout << "   Init time (program)      ~ " << time_init << " s\n";     // S4:1032
    std::cout << "   Solver time (MPI+CUDA)   ~ " << time_solver << " s\n";   // S4:1033
    std::cout << "   Finalization time        ~ " << time_finalize << " s\n"; // S4:1034
    if (report_error) // not a line of the reference (it never compares with the analytic solution it states, R4 p.1-2): opt-in
        std::cout << "   Error vs analytic solution in D: max = " << std::scientific << std::setprecision(6) << res.err_max_in_D
                  << ", L2 = " << res.err_l2_in_D << "\n";

    if (!json.empty()) {
        char buf[1536];
        snprintf(buf, sizeof(buf),
                 "{\"M\": %d, \"N\": %d, \"gpus\": %d, \"iterations\": %lld, \"stop_reason\": %d, \"first_converged_iter\": %d, "
                 "\"diff_norm\": %.17g, \"zr\": %.17g, \"w_norm_E\": %.17g, \"r_norm_E\": %.17g, \"w_max\": %.17g, "
                 "\"err_max_in_D\": %.17g, \"err_l2_in_D\": %.17g, "
                 "\"t_setup_s\": %.6g, \"t_loop_s\": %.6g, \"t_solver_s\": %.6g, \"pt_it_per_s\": %.6g, \"algo_gbps\": %.6g, "
                 "\"kernel_launches\": %lld}\n",
                 M, N, gpus, (long long)res.iterations, res.stop_reason, res.first_converged_iter, res.diff_norm, res.zr,
                 res.w_norm_E, res.r_norm_E, res.w_max, res.err_max_in_D, res.err_l2_in_D, t_setup, t_loop, time_solver,
                 t_loop > 0 ? double(M - 1) * (N - 1) * res.iterations / t_loop : 0.0,
                 t_loop > 0 ? double(M - 1) * (N - 1) * res.iterations / t_loop * 104e-9 : 0.0, (long long)res.kernel_launches);
        if (json == "-") std::cerr << buf;
        else {
            std::ofstream f(json);
            f << buf;
        }
    }
    return rc;
}
