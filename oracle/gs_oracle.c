/*
 * gs_oracle.c -- CPU restatement of GrayScott.jl's per-step hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under grayscott.jl_b200/ may link, load or call
 * this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, and only as the checker / the timed CPU baseline.
 *
 * PARITY UNPINNED BY THE REFERENCE: the reference (Julia) cannot run in this image
 * (no julia, no MPI) and its own tests hold no golden vector for calculate!/exchange!
 * (test/unit/simulation/unit-Simulation.jl:19-32 runs one step and asserts nothing).
 * The oracle is pinned instead by (i) the hand-derived known answers of SURVEY.md
 * Appendix C for exactly that unit-test case, (ii) an independent numpy-scalar
 * restatement in tests/, (iii) init/decomposition invariants.  See DESIGN.md.
 *
 * What is restated (all citations relative to /root/reference):
 *   laplacian            src/simulation/Common.jl:13-18
 *   is_inside            src/simulation/Common.jl:34-47
 *   init_fields (CPU)    src/simulation/Simulation_CPU.jl:14-65
 *   calculate!  (CPU)    src/simulation/Simulation_CPU.jl:77-113
 *   get_fields  (CPU)    src/simulation/Simulation_CPU.jl:125-133
 *   iterate!             src/simulation/public.jl:54-71
 *   init_domain          src/simulation/communication.jl:59-96
 *   get_MPI_faces        src/simulation/communication.jl:109-126
 *   exchange!            src/simulation/communication.jl:138-199
 *
 * Arithmetic rules honoured: Julia evaluates n-ary + as a left
 * fold, promotes Float32 (op) Float64 to Float64, never contracts to FMA, keeps
 * denormals.  Build with -O2 -ffp-contract=off and WITHOUT -ffast-math.
 *
 * Multi-rank runs are done in-process: N subdomains in one address space, faces
 * copied with the index ranges of the three exchange_*! helpers (there is no MPI
 * here).  Boundaries are NOT periodic (MPI.Cart_create default), so ghost cells on
 * the global boundary are never written: u's ghost alternates 1,0,1,0,... with the
 * buffer swap because u = ones(...) and u_temp = zeros(...) include the ghosts.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define GSO_F32 0
#define GSO_F64 1

/* noise modes */
#define GSO_NOISE_SKIP 0    /* no draw at all (only legal when noise == 0)                      */
#define GSO_NOISE_XOSHIRO 1 /* one xoshiro256++ draw per cell, per-thread streams, even at     */
                            /* noise == 0: mirrors Simulation_CPU.jl:103 (task-local RNG)       */
#define GSO_NOISE_PHILOX 2  /* counter-based, keyed on (seed, step, GLOBAL cell): the           */
                            /* generator the B200 backend uses; decomposition independent      */

typedef struct {
    int64_t size[3];   /* proc_sizes   communication.jl:73 */
    int64_t off[3];    /* proc_offsets communication.jl:83-87 (0-based global) */
    int coords[3];
    int nbr[6];        /* west,east,down,up,south,north; -1 = MPI.PROC_NULL  communication.jl:90-92 */
    void *u, *v, *u_temp, *v_temp; /* (sx+2)(sy+2)(sz+2), x fastest   Simulation_CPU.jl:23-28 */
} gso_rank;

typedef struct {
    int dtype;
    int64_t L;
    int dims[3];
    int nranks;
    double Du, Dv, F, k, dt, noise; /* Settings, Structs.jl:4-13 */
    int noise_mode;
    uint64_t seed;
    int64_t step; /* number of completed steps */
    gso_rank *ranks;
} gso_t;

/* ------------------------------------------------------------------ RNGs */

static inline uint64_t rotl64(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }

static inline uint64_t splitmix64(uint64_t *s) {
    uint64_t z = (*s += 0x9e3779b97f4a7c15ULL);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}

/* xoshiro256++ (Blackman & Vigna), the generator behind Julia >= 1.7's task-local
 * default RNG; rand(Float64) keeps the top 53 bits. */
typedef struct { uint64_t s[4]; } xoshiro_t;

static inline uint64_t xoshiro_next(xoshiro_t *g) {
    uint64_t *s = g->s;
    const uint64_t result = rotl64(s[0] + s[3], 23) + s[0];
    const uint64_t t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3];
    s[2] ^= t; s[3] = rotl64(s[3], 45);
    return result;
}

static inline double xoshiro_unit(xoshiro_t *g) {
    return (double)(xoshiro_next(g) >> 11) * 0x1.0p-53;
}

/* Philox4x32-10 (Salmon et al., SC'11; Random123).  Published constants. */
void gso_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* The noise word of GLOBAL cell (gx,gy,gz) at step `step`:
 *   Philox4x32-10(counter = (gx>>2, gy, gz, step), key = seed)[gx & 3]
 * and r = -1 + 2 * word * 2^-32  in [-1, 1)  (exact in Float64).
 * Same definition as grayscott.jl_b200/csrc/gs_philox.cuh. */
double gso_noise_r(uint64_t seed, uint32_t step, uint32_t gx, uint32_t gy, uint32_t gz) {
    uint32_t ctr[4] = { gx >> 2, gy, gz, step };
    uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
    uint32_t out[4];
    gso_philox4x32_10(ctr, key, out);
    return -1.0 + 2.0 * ((double)out[gx & 3] * 0x1.0p-32);
}

/* ------------------------------------------------------------------ domain */

/* MPI.Dims_create(n, [0,0,0]) (communication.jl:64): balanced factorisation,
 * non-increasing.  1->(1,1,1) 2->(2,1,1) 4->(2,2,1) 8->(2,2,2). */
void gso_dims_create(int n, int dims[3]) {
    dims[0] = dims[1] = dims[2] = 1;
    int f[64], nf = 0, m = n;
    for (int p = 2; p * p <= m; ++p) while (m % p == 0) { f[nf++] = p; m /= p; }
    if (m > 1) f[nf++] = m;
    for (int i = nf - 1; i >= 0; --i) { /* largest prime first onto the smallest dim */
        int s = 0;
        for (int d = 1; d < 3; ++d) if (dims[d] < dims[s]) s = d;
        dims[s] *= f[i];
    }
    for (int a = 0; a < 3; ++a) for (int b = a + 1; b < 3; ++b)
        if (dims[b] > dims[a]) { int t = dims[a]; dims[a] = dims[b]; dims[b] = t; }
}

static int rank_of(const int dims[3], int cx, int cy, int cz) {
    if (cx < 0 || cy < 0 || cz < 0 || cx >= dims[0] || cy >= dims[1] || cz >= dims[2]) return -1;
    return (cx * dims[1] + cy) * dims[2] + cz; /* row-major, MPI Cart default (no reorder) */
}

static size_t esize(int dtype) { return dtype == GSO_F32 ? 4 : 8; }

static size_t rank_elems(const gso_rank *r) {
    return (size_t)(r->size[0] + 2) * (size_t)(r->size[1] + 2) * (size_t)(r->size[2] + 2);
}

gso_t *gso_create(int dtype, int64_t L, const int dims[3], double Du, double Dv, double F,
                  double k, double dt, double noise, int noise_mode, uint64_t seed) {
    if (dtype != GSO_F32 && dtype != GSO_F64) return NULL;
    for (int d = 0; d < 3; ++d) if (dims[d] < 1 || L % dims[d] != 0) return NULL; /* K7: InexactError */
    gso_t *h = (gso_t *)calloc(1, sizeof(gso_t));
    h->dtype = dtype; h->L = L;
    memcpy(h->dims, dims, sizeof(int) * 3);
    h->nranks = dims[0] * dims[1] * dims[2];
    h->Du = Du; h->Dv = Dv; h->F = F; h->k = k; h->dt = dt; h->noise = noise;
    h->noise_mode = noise_mode; h->seed = seed; h->step = 0;
    h->ranks = (gso_rank *)calloc((size_t)h->nranks, sizeof(gso_rank));
    for (int cx = 0; cx < dims[0]; ++cx) for (int cy = 0; cy < dims[1]; ++cy) for (int cz = 0; cz < dims[2]; ++cz) {
        gso_rank *r = &h->ranks[rank_of(dims, cx, cy, cz)];
        r->coords[0] = cx; r->coords[1] = cy; r->coords[2] = cz;
        for (int d = 0; d < 3; ++d) {
            r->size[d] = L / dims[d];
            r->off[d] = (L / dims[d]) * r->coords[d];
        }
        r->nbr[0] = rank_of(dims, cx - 1, cy, cz); r->nbr[1] = rank_of(dims, cx + 1, cy, cz); /* west,east  = dim 0 */
        r->nbr[2] = rank_of(dims, cx, cy - 1, cz); r->nbr[3] = rank_of(dims, cx, cy + 1, cz); /* down,up    = dim 1 */
        r->nbr[4] = rank_of(dims, cx, cy, cz - 1); r->nbr[5] = rank_of(dims, cx, cy, cz + 1); /* south,north= dim 2 */
        size_t n = rank_elems(r) * esize(dtype);
        r->u = malloc(n); r->v = malloc(n); r->u_temp = malloc(n); r->v_temp = malloc(n);
    }
    return h;
}

void gso_destroy(gso_t *h) {
    if (!h) return;
    for (int i = 0; i < h->nranks; ++i) {
        free(h->ranks[i].u); free(h->ranks[i].v); free(h->ranks[i].u_temp); free(h->ranks[i].v_temp);
    }
    free(h->ranks); free(h);
}

int gso_nranks(const gso_t *h) { return h->nranks; }
int64_t gso_step_count(const gso_t *h) { return h->step; }
void gso_set_step_count(gso_t *h, int64_t s) { h->step = s; }

void gso_rank_info(const gso_t *h, int rank, int64_t size[3], int64_t off[3], int coords[3], int nbr[6]) {
    const gso_rank *r = &h->ranks[rank];
    memcpy(size, r->size, sizeof(int64_t) * 3); memcpy(off, r->off, sizeof(int64_t) * 3);
    memcpy(coords, r->coords, sizeof(int) * 3); memcpy(nbr, r->nbr, sizeof(int) * 6);
}

/* ------------------------------------------------------------------ init_fields */

#define IDX(i, j, k, nx, ny) ((size_t)(i) + (size_t)(nx) * ((size_t)(j) + (size_t)(ny) * (size_t)(k)))

#define DEFINE_INIT(NAME, T)                                                                   \
    static void NAME(gso_t *h, gso_rank *r) {                                                  \
        const int64_t nx = r->size[0] + 2, ny = r->size[1] + 2;                                \
        const size_t n = rank_elems(r);                                                        \
        T *u = (T *)r->u, *v = (T *)r->v, *ut = (T *)r->u_temp, *vt = (T *)r->v_temp;          \
        for (size_t i = 0; i < n; ++i) { u[i] = (T)1; v[i] = (T)0; ut[i] = (T)0; vt[i] = (T)0; } \
        const int64_t d = 6, minL = h->L / 2 - d, maxL = h->L / 2 + d;                         \
        for (int64_t z = minL; z <= maxL; ++z) for (int64_t y = minL; y <= maxL; ++y)          \
            for (int64_t x = minL; x <= maxL; ++x) {                                           \
                if (x < r->off[0] || x >= r->off[0] + r->size[0]) continue; /* is_inside */    \
                if (y < r->off[1] || y >= r->off[1] + r->size[1]) continue;                    \
                if (z < r->off[2] || z >= r->off[2] + r->size[2]) continue;                    \
                size_t p = IDX(x - r->off[0] + 1, y - r->off[1] + 1, z - r->off[2] + 1, nx, ny); \
                u[p] = (T)0.25; v[p] = (T)0.33;                                                \
            }                                                                                  \
    }
DEFINE_INIT(init_rank_f32, float)
DEFINE_INIT(init_rank_f64, double)

void gso_init_fields(gso_t *h) {
    for (int i = 0; i < h->nranks; ++i) {
        if (h->dtype == GSO_F32) init_rank_f32(h, &h->ranks[i]); else init_rank_f64(h, &h->ranks[i]);
    }
    h->step = 0;
}

/* ------------------------------------------------------------------ calculate! */

/* Float32: Simulation_CPU.jl:82-109 with Julia's promotion. */
static void calc_f32(int64_t sx, int64_t sy, int64_t sz, const int64_t off[3],
                     const float *restrict u, const float *restrict v,
                     float *restrict ut, float *restrict vt,
                     double sDu, double sDv, double sF, double sK, double sdt, double snoise,
                     int noise_mode, uint64_t seed, uint32_t step) {
    const float Du = (float)sDu, Dv = (float)sDv, F = (float)sF, K = (float)sK;
    const float noise = (float)snoise, dt = (float)sdt;           /* convert(T, settings.x) :82-87 */
    const int64_t nx = sx + 2, ny = sy + 2;
    const size_t sj = (size_t)nx, sk = (size_t)nx * (size_t)ny;
#pragma omp parallel
    {
        xoshiro_t g;
        {
            int tid = 0;
#ifdef _OPENMP
            tid = omp_get_thread_num();
#endif
            uint64_t s = seed ^ (0x9E3779B97F4A7C15ULL * (uint64_t)(tid + 1)) ^ ((uint64_t)step << 20);
            for (int q = 0; q < 4; ++q) g.s[q] = splitmix64(&s);
        }
#pragma omp for schedule(static)
        for (int64_t k = 1; k <= sz; ++k)                          /* Threads.@threads for k :92 */
            for (int64_t j = 1; j <= sy; ++j)
                for (int64_t i = 1; i <= sx; ++i) {
                    const size_t p = IDX(i, j, k, nx, ny);
                    const float uc = u[p], vc = v[p];
                    /* laplacian: left fold in Float32, then Float64  Common.jl:14-17 */
                    float su = u[p - 1] + u[p + 1]; su = su + u[p - sj]; su = su + u[p + sj];
                    su = su + u[p - sk]; su = su + u[p + sk];
                    float sv = v[p - 1] + v[p + 1]; sv = sv + v[p - sj]; sv = sv + v[p + sj];
                    sv = sv + v[p - sk]; sv = sv + v[p + sk];
                    const double lap_u = ((double)su - 6.0 * (double)uc) / 6.0;
                    const double lap_v = ((double)sv - 6.0 * (double)vc) / 6.0;
                    const float vv = vc * vc;
                    const float uvv = uc * vv;                         /* u * v^2 in Float32 */
                    double r = 0.0;
                    if (noise_mode == GSO_NOISE_XOSHIRO) r = -1.0 + 2.0 * xoshiro_unit(&g);
                    else if (noise_mode == GSO_NOISE_PHILOX)
                        r = gso_noise_r(seed, step, (uint32_t)(off[0] + i - 1), (uint32_t)(off[1] + j - 1),
                                        (uint32_t)(off[2] + k - 1));
                    double du = (double)Du * lap_u - (double)uvv;
                    du = du + (double)F * (1.0 - (double)uc);
                    if (noise_mode != GSO_NOISE_SKIP) du = du + (double)noise * r;
                    const float fk = F + K;
                    const float fkv = fk * vc;                         /* (F + K) * v in Float32 */
                    double dv = (double)Dv * lap_v + (double)uvv;
                    dv = dv - (double)fkv;
                    ut[p] = (float)((double)uc + du * (double)dt);     /* :108 */
                    vt[p] = (float)((double)vc + dv * (double)dt);     /* :109 */
                }
    }
}

/* Float64: same expression tree, all Float64. */
static void calc_f64(int64_t sx, int64_t sy, int64_t sz, const int64_t off[3],
                     const double *restrict u, const double *restrict v,
                     double *restrict ut, double *restrict vt,
                     double Du, double Dv, double F, double K, double dt, double noise,
                     int noise_mode, uint64_t seed, uint32_t step) {
    const int64_t nx = sx + 2, ny = sy + 2;
    const size_t sj = (size_t)nx, sk = (size_t)nx * (size_t)ny;
#pragma omp parallel
    {
        xoshiro_t g;
        {
            int tid = 0;
#ifdef _OPENMP
            tid = omp_get_thread_num();
#endif
            uint64_t s = seed ^ (0x9E3779B97F4A7C15ULL * (uint64_t)(tid + 1)) ^ ((uint64_t)step << 20);
            for (int q = 0; q < 4; ++q) g.s[q] = splitmix64(&s);
        }
#pragma omp for schedule(static)
        for (int64_t k = 1; k <= sz; ++k)
            for (int64_t j = 1; j <= sy; ++j)
                for (int64_t i = 1; i <= sx; ++i) {
                    const size_t p = IDX(i, j, k, nx, ny);
                    const double uc = u[p], vc = v[p];
                    double su = u[p - 1] + u[p + 1]; su = su + u[p - sj]; su = su + u[p + sj];
                    su = su + u[p - sk]; su = su + u[p + sk];
                    double sv = v[p - 1] + v[p + 1]; sv = sv + v[p - sj]; sv = sv + v[p + sj];
                    sv = sv + v[p - sk]; sv = sv + v[p + sk];
                    const double lap_u = (su - 6.0 * uc) / 6.0;
                    const double lap_v = (sv - 6.0 * vc) / 6.0;
                    const double vv = vc * vc;
                    const double uvv = uc * vv;
                    double r = 0.0;
                    if (noise_mode == GSO_NOISE_XOSHIRO) r = -1.0 + 2.0 * xoshiro_unit(&g);
                    else if (noise_mode == GSO_NOISE_PHILOX)
                        r = gso_noise_r(seed, step, (uint32_t)(off[0] + i - 1), (uint32_t)(off[1] + j - 1),
                                        (uint32_t)(off[2] + k - 1));
                    double du = Du * lap_u - uvv;
                    du = du + F * (1.0 - uc);
                    if (noise_mode != GSO_NOISE_SKIP) du = du + noise * r;
                    const double fk = F + K;
                    const double fkv = fk * vc;
                    double dv = Dv * lap_v + uvv;
                    dv = dv - fkv;
                    ut[p] = uc + du * dt;
                    vt[p] = vc + dv * dt;
                }
    }
}

/* Stateless one-step update on caller arrays in the reference (s+2)^3 layout. */
int gso_calculate_arrays(int dtype, const int64_t size[3], const int64_t off[3], const void *u,
                         const void *v, void *ut, void *vt, double Du, double Dv, double F, double k,
                         double dt, double noise, int noise_mode, uint64_t seed, uint32_t step) {
    if (noise_mode == GSO_NOISE_SKIP && noise != 0.0) return 1;
    if (dtype == GSO_F32)
        calc_f32(size[0], size[1], size[2], off, (const float *)u, (const float *)v, (float *)ut, (float *)vt,
                 Du, Dv, F, k, dt, noise, noise_mode, seed, step);
    else
        calc_f64(size[0], size[1], size[2], off, (const double *)u, (const double *)v, (double *)ut,
                 (double *)vt, Du, Dv, F, k, dt, noise, noise_mode, seed, step);
    return 0;
}

int gso_calculate(gso_t *h) {
    for (int i = 0; i < h->nranks; ++i) {
        gso_rank *r = &h->ranks[i];
        int rc = gso_calculate_arrays(h->dtype, r->size, r->off, r->u, r->v, r->u_temp, r->v_temp, h->Du, h->Dv,
                                      h->F, h->k, h->dt, h->noise, h->noise_mode, h->seed, (uint32_t)h->step);
        if (rc) return rc;
    }
    return 0;
}

/* ------------------------------------------------------------------ exchange! */

/* copy `count` blocks of `blocklen` elements, block stride `stride` elements:
 * the MPI vector datatypes of get_MPI_faces (communication.jl:112-120). */
static void copy_vector(char *dst, const char *src, size_t count, size_t blocklen, size_t stride, size_t es) {
    for (size_t b = 0; b < count; ++b)
        memcpy(dst + b * stride * es, src + b * stride * es, blocklen * es);
}

static void exchange_var(gso_t *h, int which) {
    const size_t es = esize(h->dtype);
    /* phases in the reference order: xy faces (z dir), xz faces (y dir), yz faces (x dir);
     * within a phase: "to north/up/east" then "to south/down/west"   communication.jl:186-198.
     * All ranks sizes are equal (K7), so sender and receiver geometry coincide. */
    for (int phase = 0; phase < 3; ++phase) for (int dirn = 0; dirn < 2; ++dirn)
        for (int rk = 0; rk < h->nranks; ++rk) {
            gso_rank *s = &h->ranks[rk];
            const int64_t sx = s->size[0], sy = s->size[1], sz = s->size[2];
            const size_t nx = (size_t)sx + 2, ny = (size_t)sy + 2;
            int dest; size_t src_idx, dst_idx, count, blocklen, stride;
            if (phase == 0) {          /* exchange_xy! :142-152 ; anchor var[2,1,k] */
                dest = dirn == 0 ? s->nbr[5] : s->nbr[4];
                src_idx = IDX(1, 0, dirn == 0 ? sz : 1, nx, ny);
                dst_idx = IDX(1, 0, dirn == 0 ? 0 : sz + 1, nx, ny);
                count = (size_t)sy + 2; blocklen = (size_t)sx; stride = nx;
            } else if (phase == 1) {   /* exchange_xz! :155-165 ; anchor var[2,j,2] */
                dest = dirn == 0 ? s->nbr[3] : s->nbr[2];
                src_idx = IDX(1, dirn == 0 ? sy : 1, 1, nx, ny);
                dst_idx = IDX(1, dirn == 0 ? 0 : sy + 1, 1, nx, ny);
                count = (size_t)sz; blocklen = (size_t)sx; stride = nx * ny;
            } else {                   /* exchange_yz! :168-178 ; anchor var[i,1,1] */
                dest = dirn == 0 ? s->nbr[1] : s->nbr[0];
                src_idx = IDX(dirn == 0 ? sx : 1, 0, 0, nx, ny);
                dst_idx = IDX(dirn == 0 ? 0 : sx + 1, 0, 0, nx, ny);
                count = ((size_t)sy + 2) * ((size_t)sz + 2); blocklen = 1; stride = nx;
            }
            if (dest < 0) continue;    /* MPI.PROC_NULL: no-op (non-periodic) */
            gso_rank *d = &h->ranks[dest];
            const char *sp = (const char *)(which == 0 ? s->u : s->v);
            char *dp = (char *)(which == 0 ? d->u : d->v);
            copy_vector(dp + dst_idx * es, sp + src_idx * es, count, blocklen, stride, es);
        }
}

void gso_exchange(gso_t *h) { exchange_var(h, 0); exchange_var(h, 1); } /* for var in [u, v] :186 */

void gso_swap(gso_t *h) {  /* public.jl:67-68 */
    for (int i = 0; i < h->nranks; ++i) {
        gso_rank *r = &h->ranks[i];
        void *t = r->u; r->u = r->u_temp; r->u_temp = t;
        t = r->v; r->v = r->v_temp; r->v_temp = t;
    }
    h->step += 1;
}

int gso_step(gso_t *h, int nsteps) { /* iterate!  public.jl:54-71 */
    for (int s = 0; s < nsteps; ++s) {
        gso_exchange(h);
        int rc = gso_calculate(h);
        if (rc) return rc;
        gso_swap(h);
    }
    return 0;
}

/* ------------------------------------------------------------------ access */

static void *field_ptr(gso_rank *r, int which) {
    switch (which) { case 0: return r->u; case 1: return r->v; case 2: return r->u_temp; default: return r->v_temp; }
}

/* which: 0=u 1=v 2=u_temp 3=v_temp; buffers are the reference's ghosted (s+2)^3 layout */
void gso_get_ghosted(gso_t *h, int rank, int which, void *out) {
    gso_rank *r = &h->ranks[rank];
    memcpy(out, field_ptr(r, which), rank_elems(r) * esize(h->dtype));
}
void gso_set_ghosted(gso_t *h, int rank, int which, const void *in) {
    gso_rank *r = &h->ranks[rank];
    memcpy(field_ptr(r, which), in, rank_elems(r) * esize(h->dtype));
}

/* get_fields: interior copies, x fastest, sx*sy*sz   Simulation_CPU.jl:125-133 */
void gso_get_fields(gso_t *h, int rank, void *u_out, void *v_out) {
    gso_rank *r = &h->ranks[rank];
    const size_t es = esize(h->dtype);
    const int64_t sx = r->size[0], sy = r->size[1], sz = r->size[2], nx = sx + 2, ny = sy + 2;
    for (int64_t k = 0; k < sz; ++k) for (int64_t j = 0; j < sy; ++j) {
        size_t src = IDX(1, j + 1, k + 1, nx, ny) * es, dst = ((size_t)j + (size_t)sy * (size_t)k) * (size_t)sx * es;
        memcpy((char *)u_out + dst, (const char *)r->u + src, (size_t)sx * es);
        memcpy((char *)v_out + dst, (const char *)r->v + src, (size_t)sx * es);
    }
}

/* global L^3 arrays (x fastest) assembled from all ranks' interiors: what ADIOS2 would
 * see from IO.write_step! with start = offsets, count = sizes (IO.jl:55-66). */
void gso_gather(gso_t *h, void *u_out, void *v_out) {
    const size_t es = esize(h->dtype);
    const size_t L = (size_t)h->L;
    for (int rk = 0; rk < h->nranks; ++rk) {
        gso_rank *r = &h->ranks[rk];
        const int64_t sx = r->size[0], sy = r->size[1], sz = r->size[2], nx = sx + 2, ny = sy + 2;
        for (int64_t k = 0; k < sz; ++k) for (int64_t j = 0; j < sy; ++j) {
            size_t src = IDX(1, j + 1, k + 1, nx, ny) * es;
            size_t dst = ((size_t)r->off[0] + L * ((size_t)(r->off[1] + j) + L * (size_t)(r->off[2] + k))) * es;
            memcpy((char *)u_out + dst, (const char *)r->u + src, (size_t)sx * es);
            memcpy((char *)v_out + dst, (const char *)r->v + src, (size_t)sx * es);
        }
    }
}

int gso_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void gso_set_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}
