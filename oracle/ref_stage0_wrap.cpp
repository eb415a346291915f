/*
 * ref_stage0_wrap.cpp -- the UNMODIFIED reference stage0 serial solver
 * (/root/reference/stage0/Withoutopenmp1.cpp) as a callable library.
 *
 * TEST INFRASTRUCTURE ONLY. The source is #include'd from where it lies (Makefile passes
 * -I/root/reference/stage0) with `main` renamed; nothing is copied. Built WITHOUT -fopenmp,
 * as the stage is meant to be (its pragmas are then ignored). Output goes to oracle/_ref/.
 *
 *   ref0_solve -> solve (Withoutopenmp1.cpp:106-172): unweighted stopping norm
 *                 (Withoutopenmp1.cpp:149-154) and the signed `denom < 1e-15` guard (:128).
 *   ref0_main  -> the reference main (Withoutopenmp1.cpp:176-197): grids 10^2, 20^2, 40^2.
 */
#define main reference_stage0_main
#include "Withoutopenmp1.cpp"
#undef main

extern "C" {
int ref0_solve(int M, int N, double delta, int max_iter) { return solve(M, N, delta, max_iter); }
int ref0_main(void) { return reference_stage0_main(); }
}

#ifdef REF_STAGE0_MAIN
int main() { return reference_stage0_main(); }
#endif
