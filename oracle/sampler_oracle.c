/* CPU restatement, in plain C, of DistributedSampler's per-rank index shard.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): tests/, smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load this; the product
 * library never does.
 *
 * Follows the sampler the reference builds at codes/task2/model.py:124, whose
 * arithmetic is third-party torch: torch/utils/data/distributed.py:93-141
 * (sizes, pad, stride) over ATen randperm_cpu + at::mt19937 (init_genrand,
 * 32-bit tempered draws, forward Fisher-Yates with `draw % (n - i)`).
 * Pinned against torch in tests/test_oracle_golden.py.
 */
#include <stdint.h>
#include <stdlib.h>

#define MT_N 624
#define MT_M 397

typedef struct { uint32_t s[MT_N]; int pos; } mt_t;

static void mt_seed(mt_t *g, uint64_t seed) {
    g->s[0] = (uint32_t)(seed & 0xffffffffu);   /* low 32 bits only */
    for (int i = 1; i < MT_N; i++)
        g->s[i] = 1812433253u * (g->s[i - 1] ^ (g->s[i - 1] >> 30)) + (uint32_t)i;
    g->pos = MT_N;
}

static void mt_refill(mt_t *g) {
    uint32_t *s = g->s;
    for (int k = 0; k < MT_N; k++) {
        uint32_t y = (s[k] & 0x80000000u) | (s[(k + 1) % MT_N] & 0x7fffffffu);
        s[k] = s[(k + MT_M) % MT_N] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    g->pos = 0;
}

static uint32_t mt_next(mt_t *g) {
    if (g->pos >= MT_N) mt_refill(g);
    uint32_t y = g->s[g->pos++];
    y ^= y >> 11;
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= y >> 18;
    return y;
}

/* out[0..n) = torch.randperm(n, generator=manual_seed(seed)); n < 2^32/20 */
int oracle_randperm(int64_t n, uint64_t seed, int64_t *out) {
    if (n < 0 || n >= (int64_t)(0xffffffffu / 20)) return -1;
    mt_t g;
    mt_seed(&g, seed);
    for (int64_t i = 0; i < n; i++) out[i] = i;
    for (int64_t i = 0; i + 1 < n; i++) {
        int64_t z = (int64_t)(mt_next(&g) % (uint64_t)(n - i));
        int64_t t = out[i]; out[i] = out[i + z]; out[i + z] = t;
    }
    return 0;
}

/* Writes this rank's indices, returns how many (num_samples) or -1. */
int64_t oracle_sampler_shard(int64_t n, int world, int rank, uint64_t seed, uint64_t epoch,
                             int shuffle, int drop_last, int64_t *out) {
    if (world <= 0 || rank < 0 || rank >= world || n < 0) return -1;
    int64_t num;
    if (drop_last && n % world != 0) {
        int64_t a = n - world;                      /* ceil((n - world) / world) */
        num = a >= 0 ? (a + world - 1) / world : -((-a) / world);
    } else {
        num = (n + world - 1) / world;
    }
    if (num <= 0 || n == 0) return 0;
    int64_t *perm = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
    if (!perm) return -1;
    if (shuffle) {
        if (oracle_randperm(n, seed + epoch, perm) != 0) { free(perm); return -1; }
    } else {
        for (int64_t i = 0; i < n; i++) perm[i] = i;
    }
    for (int64_t k = 0; k < num; k++) {
        int64_t t = rank + (int64_t)world * k;      /* position in the padded list */
        out[k] = perm[t < n ? t : (t - n) % n];
    }
    free(perm);
    return num;
}
