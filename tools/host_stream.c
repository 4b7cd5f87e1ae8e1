/* Host memory ceiling for the end-to-end scaling question: P threads each stream-store W MB
 * (the access pattern of the row assembly: _mm_stream_ps into a large array) and, second pass, also read R MB.
 * usage: host_stream threads [MB per thread = 102] [reps = 20]     gcc -O2 -pthread -msse2 */
#include <emmintrin.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <time.h>
typedef struct { float* dst; const float* src; size_t n, nr; int reps; float sink; } job;
static pthread_barrier_t bar;
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }
static void* work(void* p) {
    job* j = (job*)p;
    pthread_barrier_wait(&bar);
    for (int r = 0; r < j->reps; ++r) {
        __m128 acc = _mm_setzero_ps();
        for (size_t i = 0; i < j->nr; i += 4) acc = _mm_add_ps(acc, _mm_load_ps(j->src + i));
        const __m128 z = _mm_setzero_ps();
        for (size_t i = 0; i < j->n; i += 4) _mm_stream_ps(j->dst + i, z);
        _mm_sfence();
        float t[4]; _mm_storeu_ps(t, acc); j->sink += t[0];
    }
    pthread_barrier_wait(&bar);
    return 0;
}
int main(int argc, char** argv) {
    int P = argc > 1 ? atoi(argv[1]) : 32, mb = argc > 2 ? atoi(argv[2]) : 102, reps = argc > 3 ? atoi(argv[3]) : 20;
    for (int pass = 0; pass < 2; ++pass) {
        pthread_t th[256]; job jb[256];
        pthread_barrier_init(&bar, 0, P + 1);
        for (int i = 0; i < P; ++i) {
            jb[i].n = (size_t)mb * 250000; jb[i].nr = pass ? jb[i].n / 4 : 0; jb[i].reps = reps; jb[i].sink = 0;
            jb[i].dst = (float*)aligned_alloc(4096, jb[i].n * 4); jb[i].src = (float*)aligned_alloc(4096, jb[i].n);
            for (size_t k = 0; k < jb[i].n; k += 1024) jb[i].dst[k] = 1.f;
            for (size_t k = 0; k < jb[i].n / 4; k += 1024) ((float*)jb[i].src)[k] = 1.f;
            pthread_create(&th[i], 0, work, &jb[i]);
        }
        pthread_barrier_wait(&bar); double t0 = now(); pthread_barrier_wait(&bar); double dt = now() - t0;
        for (int i = 0; i < P; ++i) pthread_join(th[i], 0);
        double wb = (double)P * mb * 1e6 * reps, rb = pass ? wb / 4 : 0;
        printf("%d threads, %s: %.1f GB/s written%s (%.2f ms per %d x %d MB)\n", P, pass ? "stores + reads 1:4" : "streaming stores only",
               wb / dt / 1e9, pass ? "" : "", dt / reps * 1e3, P, mb);
        if (pass) printf("   + %.1f GB/s read\n", rb / dt / 1e9);
        for (int i = 0; i < P; ++i) { free(jb[i].dst); free((void*)jb[i].src); }
    }
    return 0;
}
