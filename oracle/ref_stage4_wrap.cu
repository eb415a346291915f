/*
 * ref_stage4_wrap.cu -- the UNMODIFIED reference stage4 MPI+CUDA solver
 * (/root/reference/stage4-mpi+cuda/poisson_mpi_cuda_f.cu) built for the B200 box as the
 * "beat THAT kernel" GPU baseline.
 *
 * TEST / MEASUREMENT INFRASTRUCTURE ONLY. The source is #include'd from where it lies
 * (Makefile passes -I/root/reference/stage4-mpi+cuda), `main` renamed, compiled against
 * oracle/shim/mpi.h with nvcc's defaults (so FMA contraction is ON, as in the reference
 * build: Makefile:12-13 there passes no -fmad flag). Ranks are threads sharing GPU 0, as
 * the reference never calls cudaSetDevice.  Usage: ref_stage4 [M N]; $SHIM_MPI_RANKS=P.
 */
#include <mpi.h>
#define main reference_stage4_main
#include "poisson_mpi_cuda_f.cu"
#undef main

int main(int argc, char **argv) {
    const char *e = std::getenv("SHIM_MPI_RANKS");
    int ranks = e ? std::atoi(e) : 1;
    if (ranks <= 1) return reference_stage4_main(argc, argv);
    int rc = 0;
    shim_mpi_launch(ranks, [&](int) { rc |= reference_stage4_main(argc, argv); });
    return rc;
}
