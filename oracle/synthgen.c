/* synthgen.c — CPU generator of the SIFT-shaped synthetic corpus, bit-identical to
 * vectorsearch-rbac_b200/synth.py:sift_like (same counter-based splitmix64 hash, integer arithmetic only).
 *
 * TEST / BENCH INFRASTRUCTURE ONLY (lives under oracle/): `bench.py --impl reference` fills its 10 M x 128 host
 * matrix with it so that the reference arm touches neither the GPU nor librbacknn.so; tests/test_synth.py checks it
 * against the torch generator bit for bit.  The reference itself seeds nothing and reads SIFT from disk
 * (services/read_dataset_function.py:338-339 assigns block / document ids); the values here only have the SIFT
 * shape (integers 0..255 as fp32, skewed, clustered). */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>

static inline uint64_t mix64(uint64_t x)
{
    x *= 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

typedef struct {
    const int64_t *rows;
    size_t a, b;
    int dim;
    int64_t seed;
    int64_t n_clusters;
    float *out;
} gen_job;

static void *gen_thread(void *arg)
{
    gen_job *j = (gen_job *) arg;
    const uint64_t dim = (uint64_t) j->dim;
    const uint64_t s_cl = (uint64_t) (j->seed * 7919 + 1), s_c = (uint64_t) (j->seed * 31 + 2), s_n = (uint64_t) (j->seed * 131 + 3);
    for (size_t i = j->a; i < j->b; i++) {
        uint64_t r = (uint64_t) j->rows[i];
        /* torch: (mix >> 11) as a non-negative int64, then % n_clusters */
        uint64_t cl = (mix64(r ^ s_cl) >> 11) % (uint64_t) j->n_clusters;
        float *o = j->out + i * dim;
        for (uint64_t c = 0; c < dim; c++) {
            uint64_t cb = mix64((cl * dim + c) ^ s_c);
            int64_t base = (int64_t) (((cb & 255) * ((cb >> 8) & 255)) >> 8);
            uint64_t nb = mix64((r * dim + c) ^ s_n);
            int64_t noise = (int64_t) (nb & 31) + (int64_t) ((nb >> 5) & 31) - 31;
            int64_t v = base + noise;
            v = v < 0 ? 0 : (v > 255 ? 255 : v);
            o[c] = (float) v;
        }
    }
    return NULL;
}

/* rows[n] (corpus row numbers) -> out[n][dim] */
int orc_gen_sift(const int64_t *rows, size_t n, int dim, int64_t seed, int64_t n_clusters, float *out, int nthreads)
{
    if (nthreads < 1) nthreads = 1;
    if ((size_t) nthreads > n) nthreads = n ? (int) n : 1;
    pthread_t *th = (pthread_t *) malloc(sizeof(pthread_t) * (size_t) nthreads);
    gen_job *jobs = (gen_job *) malloc(sizeof(gen_job) * (size_t) nthreads);
    if (!th || !jobs) { free(th); free(jobs); return -1; }
    for (int t = 0; t < nthreads; t++) {
        jobs[t] = (gen_job){rows, n * (size_t) t / (size_t) nthreads, n * (size_t) (t + 1) / (size_t) nthreads, dim, seed, n_clusters, out};
        pthread_create(&th[t], NULL, gen_thread, &jobs[t]);
    }
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    free(th);
    free(jobs);
    return 0;
}
