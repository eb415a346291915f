// pb200_lib.cu -- the single translation unit of libpoisson_b200.so. The kernels live in a header
// (pb200_kernels.cuh) and are instantiated exactly once, here; the three parts below share them.
#include "pb200_solver.cu"
#include "pb200_comm.cu"
#include "pb200_capi.cu"
#include "pb200_noalloc.cu"
