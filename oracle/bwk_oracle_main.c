/*
 * bwk_oracle_main.c -- TEST INFRASTRUCTURE ONLY.
 * Restatement of `program biharmonic_wk_scalar_kernel` (F90:545-580) on top of the C oracle,
 * with the GPU1 block (F90:562-569, commented out upstream) enabled so that both published
 * known-answer norms are printed.   usage: bwk_oracle_main [nelemd nlev qsize]
 */
#include "bwk_oracle.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

static double now(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

int main(int argc, char **argv)
{
    int nelemd = 16, nlev = 72, qsize = 40; /* F90:11-17 */
    if (argc == 4) { nelemd = atoi(argv[1]); nlev = atoi(argv[2]); qsize = atoi(argv[3]); }
    size_t n = (size_t)16 * nlev * qsize * nelemd;
    double dvv[16];
    double *elem = malloc(sizeof(double) * BWKO_ELEM_DOUBLES * nelemd);
    double *qtens = malloc(sizeof(double) * n), *save = malloc(sizeof(double) * n), *grads = malloc(sizeof(double) * 2 * n);

    bwko_initialize_data(nelemd, nlev, qsize, dvv, elem, qtens);
    double t1 = now();
    bwko_biharmonic_wk_scalar_cpu(elem, qtens, dvv, nlev, qsize, 1, nelemd);
    double t2 = now();
    memcpy(save, qtens, sizeof(double) * n);
    printf(" CPU  time: %24.16E\n", t2 - t1);

    bwko_initialize_data(nelemd, nlev, qsize, dvv, elem, qtens);
    t1 = now();
    bwko_biharmonic_wk_scalar_gpu1(elem, qtens, dvv, nlev, qsize, 1, nelemd);
    t2 = now();
    printf(" GPU1 time: %24.16E\n", t2 - t1);
    printf(" GPU1 L2 norm: %24.16E\n", bwko_compute_l2norm(qtens, save, (int64_t)n));

    bwko_initialize_data(nelemd, nlev, qsize, dvv, elem, qtens);
    t1 = now();
    bwko_biharmonic_wk_scalar_gpu2(elem, qtens, grads, dvv, nlev, qsize, nelemd, 1, nelemd);
    t2 = now();
    printf(" GPU2 time: %24.16E\n", t2 - t1);
    printf(" GPU2 L2 norm: %24.16E\n", bwko_compute_l2norm(qtens, save, (int64_t)n));
    free(elem); free(qtens); free(save); free(grads);
    return 0;
}
