/*
 * oracle/ctc_oracle.c -- TEST INFRASTRUCTURE ONLY.  See ctc_oracle_impl.h for the contract, the reference
 * call sites it restates (/root/reference/train.py:12,352,457,600,904) and the "parity unpinned" statement.
 *
 * Exports (C ABI, loaded with ctypes by oracle/ctc_oracle.py):
 *   ctc_oracle_compute_f64  -- fp64 log-space: the golden oracle for parity tests
 *   ctc_oracle_compute_f32  -- fp32 log-space: the timed CPU baseline ("port" of the reference-class CPU path)
 *   ctc_oracle_max_threads  -- OpenMP thread count available
 *
 * Build:  make -C oracle        (gcc -O3 -fopenmp -shared -fPIC)
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define REAL double
#define SUFFIX f64
#define R_EXP exp
#define R_LOG log
#define R_LOG1P log1p
#include "ctc_oracle_impl.h"
#undef REAL
#undef SUFFIX
#undef R_EXP
#undef R_LOG
#undef R_LOG1P

#define REAL float
#define SUFFIX f32
#define R_EXP expf
#define R_LOG logf
#define R_LOG1P log1pf
#include "ctc_oracle_impl.h"
#undef REAL
#undef SUFFIX
#undef R_EXP
#undef R_LOG
#undef R_LOG1P

int ctc_oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
