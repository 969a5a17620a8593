/*
 * wkv6_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see wkv6_oracle_impl.h).
 * Builds libwkv6_oracle.so exporting *_f32 (reference operation order, fmaf
 * where nvcc contracts) and *_f64 ("truth") variants of every function.
 * Build: make -C oracle   (gcc -O3 -march=native -fopenmp, no -ffast-math).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stddef.h>

#define REAL float
#define SUF f32
#define EXP expf
#define FMA fmaf
#include "wkv6_oracle_impl.h"
#undef REAL
#undef SUF
#undef EXP
#undef FMA

#define REAL double
#define SUF f64
#define EXP exp
#define FMA fma
#include "wkv6_oracle_impl.h"
#undef REAL
#undef SUF
#undef EXP
#undef FMA

#ifdef _OPENMP
#include <omp.h>
int wkv6_oracle_num_threads(void) { return omp_get_max_threads(); }
void wkv6_oracle_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
#else
int wkv6_oracle_num_threads(void) { return 1; }
void wkv6_oracle_set_threads(int n) { (void)n; }
#endif
