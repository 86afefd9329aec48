// cholmod_b200_shim.c -- the cholmod_common / cholmod_factor level glue of the CHOLMOD/B200 module.
//
// This is the thin C file a CHOLMOD maintainer adds next to CHOLMOD/GPU.  It is
// compiled against the reference's own cholmod.h and exports, with CHOLMOD's naming conventions
// (SURVEY.md §8b):
//
//   cholmod_l_b200_probe            <->  cholmod_l_gpu_probe            (CHOLMOD/GPU/cholmod_gpu.c:148)
//   cholmod_l_b200_allocate         <->  cholmod_l_gpu_allocate         (cholmod_gpu.c:426): Common->maxGpuMemBytes /
//                                        maxGpuMemFraction, CHOLMOD_GPU_MEM_BYTES / CHOLMOD_GPU_MEM_FRACTION become
//                                        the cap on what the module may hold on the device
//   cholmod_l_b200_symbolic         <->  gpu_init (t_cholmod_gpu.c:84): build + upload the plan of a symbolic factor
//   cholmod_l_b200_deallocate       <->  cholmod_l_gpu_deallocate       (cholmod_gpu.c:187)
//   rd_/rs_/cd_/cs_/zd_/zs_cholmod_l_b200_super_numeric <-> the useGPU branch of the six instances of
//                                        TEMPLATE (cholmod_super_numeric_worker)
//                                        (t_cholmod_super_numeric_worker.c:133-1279; the reference's GPU
//                                        branch exists for rd_ and cd_ only)
//   cholmod_l_super_numeric         the public entry (cholmod_super_numeric.c:100) re-implemented so that,
//                                   when this library precedes libcholmod in symbol resolution (link
//                                   order or LD_PRELOAD), an UNMODIFIED cholmod_l_factorize lands in the
//                                   B200 numeric phase.  Inputs outside the module's scope (int32
//                                   indices, malformed arguments) are handed to the next definition of
//                                   cholmod_l_super_numeric (the reference's), exactly as the reference's
//                                   GPU branch only ever covered real/complex double int64.
//   cholmod_l_super_lsolve / _ltsolve  the two supernodal sweeps cholmod_l_solve2 calls
//                                   (cholmod_solve.c:696-708, cholmod_super_solve.c:40, :140): run on the device
//                                   with the factor still resident in HBM when that is enabled (below).
//
// Environment (beside the reference's CHOLMOD_USE_GPU, CHOLMOD_GPU_MEM_BYTES, CHOLMOD_GPU_MEM_FRACTION):
//   CHOLMOD_B200_DEVICE=<d>       device to use (default 0)
//   CHOLMOD_B200_NGPUS=<N>        one factorization over devices d .. d+N-1 of this process (default 1; real double
//                                 symmetric matrices)
//   CHOLMOD_B200_LAZY_LX=1        the numeric phase leaves L->x on the device; host L->x is filled only when host
//                                 code is about to read it (cholmod_l_change_factor, cholmod_l_copy_factor,
//                                 cholmod_l_rcond, a host-side solve).  cholmod_l_solve then runs on the device
//                                 and the 8-59 GB copy of SURVEY.md 8f never happens.  Off by default: a program
//                                 that reads L->x directly, without going through CHOLMOD, would see stale memory.
//   CHOLMOD_B200_TRACE=1          one line on stderr per numeric factorization that ran on the device
//   CHOLMOD_B200_DEVICE_SOLVE=1   device sweeps for cholmod_l_solve without the lazy copy (host L->x is then assumed
//                                 unmodified since the factorization)
//
// A failure of the device path is reported as CHOLMOD_GPU_PROBLEM; it is NOT retried on the CPU.
//
// Threads: as in CHOLMOD, one cholmod_common per thread.  Contexts live in a process-wide table; a context
// is used by one thread at a time (two threads that factorize matrices with the same symbolic structure get
// two contexts; concurrent solves with one factor take turns), and a context in use is never evicted.

#define _GNU_SOURCE
#include "cholmod.h"
#include "../../include/cholmod_b200.h"
#include <dlfcn.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

// ---------------------------------------------------------------------------------------------
// per-factor contexts, keyed by the content of the symbolic factor
// ---------------------------------------------------------------------------------------------

#define B200_SHIM_SLOTS 16

typedef struct
{
    b200_factor *ctx ;
    b200_multi *multi ;         // CHOLMOD_B200_NGPUS > 1: the same factorization spread over several devices
    uint64_t key ;
    int64_t n, nsuper, ssize, xsize ;
    int dtype ;
    uint64_t stamp ;
    int refs ;                  // threads that hold this slot (never evicted while > 0)
    int temp ;                  // not in the table: destroyed when released (all table slots were in use)
    pthread_mutex_t use ;       // serialises the holders
    // the numeric factor whose values the device arena holds now
    const cholmod_factor *last_L ;
    const void *last_Lx ;
    int values_valid ;          // the last factorization through this context ended with CHOLMOD_OK
    int host_stale ;            // lazy mode: host last_Lx has not been filled yet
} shim_slot ;

static shim_slot g_slots [B200_SHIM_SLOTS] ;
static int g_slots_init = 0 ;
static uint64_t g_clock = 0 ;
static pthread_mutex_t g_lock = PTHREAD_MUTEX_INITIALIZER ;

static int env_int (const char *name, int dflt)
{
    const char *e = getenv (name) ;
    return e ? atoi (e) : dflt ;
}

static uint64_t hash_i64 (uint64_t h, const int64_t *p, size_t n, size_t stride)
{
    for (size_t k = 0 ; k < n ; k += stride)
    {
        h ^= (uint64_t) p [k] ;
        h *= 0x100000001b3ULL ;
    }
    return h ;
}

// Content key of a symbolic factor: L->super and L->pi in full, L->s sampled (every entry of a short list,
// ~64k evenly spaced entries of a long one -- hashing all of a 50 MB list on every call would cost more than
// a small factorization).  Two analyses of different matrices that agree in all of that and in n, nsuper,
// ssize, xsize are taken to be the same structure.
static uint64_t factor_key (const cholmod_factor *L)
{
    uint64_t h = 0xcbf29ce484222325ULL ;
    h = hash_i64 (h, (const int64_t *) L->super, L->nsuper + 1, 1) ;
    h = hash_i64 (h, (const int64_t *) L->pi, L->nsuper + 1, 1) ;
    size_t stride = L->ssize / 65536 ;
    h = hash_i64 (h, (const int64_t *) L->s, L->ssize, stride ? stride : 1) ;
    return h ;
}

static int shim_device (void) { return env_int ("CHOLMOD_B200_DEVICE", 0) ; }

int cholmod_l_b200_probe (cholmod_common *Common)
{
    (void) Common ;
    return b200_probe (shim_device ()) ;
}

// cholmod_l_gpu_allocate (cholmod_gpu.c:426-543) sizes one device pool from Common->maxGpuMemBytes and
// maxGpuMemFraction (the environment variables CHOLMOD_GPU_MEM_BYTES / _FRACTION feed those fields,
// cholmod_super_symbolic.c:264-281).  Here the same numbers, by the same rules, become the cap on what the
// module's contexts may hold on the device; a context that does not fit is refused (CHOLMOD_OUT_OF_MEMORY).
// Returns 0 like the reference; Common->status = CHOLMOD_GPU_PROBLEM if the device cannot be queried.
int cholmod_l_b200_allocate (cholmod_common *Common)
{
    if (!Common) return 0 ;
    size_t max_bytes = Common->maxGpuMemBytes ;
    double max_fraction = Common->maxGpuMemFraction ;
    const char *eb = getenv ("CHOLMOD_GPU_MEM_BYTES"), *ef = getenv ("CHOLMOD_GPU_MEM_FRACTION") ;
    if (max_bytes == 0 && eb) { long long v = atoll (eb) ; max_bytes = (v > 0) ? (size_t) v : 0 ; Common->maxGpuMemBytes = max_bytes ; }
    if (max_fraction == 0 && ef) { max_fraction = atof (ef) ; Common->maxGpuMemFraction = max_fraction ; }
    if (max_fraction < 0) max_fraction = 0 ;
    if (max_fraction > 1) max_fraction = 1 ;
    const int dev = shim_device (), ngpus = env_int ("CHOLMOD_B200_NGPUS", 1) ;
    for (int d = dev ; d < dev + (ngpus > 1 ? ngpus : 1) ; d++)
    {
        size_t fdm = 0, tdm = 0, limit = 0 ;
        if (b200_memorysize (d, &fdm, &tdm) != B200_OK)
        {
            cholmod_l_error (CHOLMOD_GPU_PROBLEM, __FILE__, __LINE__, "gpu memorysize failure", Common) ;
            return 0 ;
        }
        if (max_bytes == 0 && max_fraction == 0) limit = 0 ;                         // no request: whatever is free
        else if (max_bytes > 0 && max_fraction > 0)
        {
            limit = max_bytes ;
            if ((double) limit > (double) tdm * max_fraction) limit = (size_t) ((double) tdm * max_fraction) ;
        }
        else if (max_fraction > 0) limit = (size_t) (max_fraction * (double) tdm) ;
        else limit = max_bytes ;
        b200_set_memory_limit (d, limit) ;
        if (d == dev) Common->dev_mempool_size = limit ;        // informational: what cholmod_l_gpu_allocate would hold
    }
    return 0 ;
}

static void slots_init (void)
{
    if (g_slots_init) return ;
    for (int k = 0 ; k < B200_SHIM_SLOTS ; k++) pthread_mutex_init (&g_slots [k].use, NULL) ;
    g_slots_init = 1 ;
}

static void slot_destroy_ctx (shim_slot *s)
{
    if (s->multi) b200_multi_destroy (s->multi) ;
    else if (s->ctx) b200_destroy (s->ctx) ;
    s->ctx = NULL ; s->multi = NULL ;
}

// lazy mode: bring host L->x up to date from the device copy the slot still holds (slot held by the caller)
static int slot_materialize (shim_slot *s, cholmod_common *Common)
{
    if (!s->host_stale) return 1 ;
    int st = s->multi ? b200_multi_download_factor (s->multi, (void *) s->last_Lx)
                      : b200_download_factor (s->ctx, (void *) s->last_Lx) ;
    if (st < 0)
    {
        if (Common) cholmod_l_error (st, __FILE__, __LINE__, "CHOLMOD/B200: copying L->x to the host failed", Common) ;
        return 0 ;
    }
    s->host_stale = 0 ;
    return 1 ;
}

static void slot_release (shim_slot *s)
{
    pthread_mutex_unlock (&s->use) ;
    if (s->temp)
    {
        slot_destroy_ctx (s) ;
        pthread_mutex_destroy (&s->use) ;
        free (s) ;
        return ;
    }
    pthread_mutex_lock (&g_lock) ;
    s->refs-- ;
    pthread_mutex_unlock (&g_lock) ;
}

// Find (or build) the context of a supernodal symbolic factor and take it for this thread; the caller releases
// it with slot_release.  NULL + Common->status on failure.
static shim_slot *slot_acquire (cholmod_factor *L, int dtype, cholmod_common *Common, int want_multi)
{
    // several GPUs: the real double symmetric case (north_star's); everything else runs on one device
    want_multi = want_multi && dtype == B200_DOUBLE && env_int ("CHOLMOD_B200_NGPUS", 1) > 1 ;
    uint64_t key = factor_key (L) ;
    shim_slot *got = NULL ;
    pthread_mutex_lock (&g_lock) ;
    slots_init () ;
    int victim = -1 ;
    for (int k = 0 ; k < B200_SHIM_SLOTS ; k++)
    {
        shim_slot *s = &g_slots [k] ;
        if ((s->ctx || s->multi) && s->refs == 0 && s->key == key && s->n == (int64_t) L->n
            && s->nsuper == (int64_t) L->nsuper && s->ssize == (int64_t) L->ssize && s->xsize == (int64_t) L->xsize
            && s->dtype == dtype && (s->multi != NULL) == (want_multi != 0))
        {
            got = s ;
            break ;
        }
        if (s->refs == 0 && (victim < 0 || s->stamp < g_slots [victim].stamp)) victim = k ;
    }
    if (got)
    {
        got->refs++ ; got->stamp = ++g_clock ;
        pthread_mutex_unlock (&g_lock) ;
        pthread_mutex_lock (&got->use) ;
        return got ;
    }
    // a new context: into the least recently used idle slot, or a temporary one when every slot is in use
    shim_slot *s ;
    if (victim >= 0)
    {
        s = &g_slots [victim] ;
        s->refs = 1 ;                       // reserved: nobody else picks it while it is being rebuilt
    }
    else
    {
        s = (shim_slot *) calloc (1, sizeof (shim_slot)) ;
        if (!s) { pthread_mutex_unlock (&g_lock) ; Common->status = CHOLMOD_OUT_OF_MEMORY ; return NULL ; }
        pthread_mutex_init (&s->use, NULL) ;
        s->temp = 1 ; s->refs = 1 ;
    }
    pthread_mutex_unlock (&g_lock) ;
    pthread_mutex_lock (&s->use) ;
    if (s->ctx || s->multi)
    {
        // the evicted context may still owe its factor a host copy (lazy mode)
        if (s->host_stale) slot_materialize (s, Common) ;
        slot_destroy_ctx (s) ;
    }
    s->last_L = NULL ; s->last_Lx = NULL ; s->values_valid = 0 ; s->host_stale = 0 ; s->key = 0 ;
    cholmod_l_b200_allocate (Common) ;      // the reference sizes its pool during analysis; we cap ours here
    int status = 0 ;
    const int dev = shim_device (), ngpus = env_int ("CHOLMOD_B200_NGPUS", 1) ;
    if (want_multi)
        s->multi = b200_multi_create ((int64_t) L->n, (int64_t) L->nsuper, (const int64_t *) L->super,
            (const int64_t *) L->pi, (const int64_t *) L->px, (const int64_t *) L->s, dtype, dev, ngpus, &status) ;
    else
        s->ctx = b200_create ((int64_t) L->n, (int64_t) L->nsuper, (const int64_t *) L->super,
            (const int64_t *) L->pi, (const int64_t *) L->px, (const int64_t *) L->s, dtype, dev, &status) ;
    if (!s->ctx && !s->multi)
    {
        Common->status = status ;           // the B200_* codes are CHOLMOD's
        slot_release (s) ;
        return NULL ;
    }
    s->key = key ; s->n = (int64_t) L->n ; s->nsuper = (int64_t) L->nsuper ;
    s->ssize = (int64_t) L->ssize ; s->xsize = (int64_t) L->xsize ; s->dtype = dtype ;
    pthread_mutex_lock (&g_lock) ;
    s->stamp = ++g_clock ;
    pthread_mutex_unlock (&g_lock) ;
    return s ;
}

// the slot whose device arena holds the values of numeric factor L, taken for this thread; or NULL
static shim_slot *slot_of_factor (const cholmod_factor *L)
{
    shim_slot *got = NULL ;
    pthread_mutex_lock (&g_lock) ;
    for (int k = 0 ; k < B200_SHIM_SLOTS && g_slots_init ; k++)
    {
        shim_slot *s = &g_slots [k] ;
        if ((s->ctx || s->multi) && s->last_L == L && s->last_Lx == L->x && L->x) { got = s ; s->refs++ ; break ; }
    }
    pthread_mutex_unlock (&g_lock) ;
    if (!got) return NULL ;
    pthread_mutex_lock (&got->use) ;
    if (got->last_L != L || got->last_Lx != L->x)       // it changed hands while we waited
    {
        slot_release (got) ;
        return NULL ;
    }
    return got ;
}

// B200 entry type of a factor: CHOLMOD's xtype + dtype for complex, dtype alone for real
static int shim_dtype (int xtype, int dtype)
{
    if (xtype == CHOLMOD_COMPLEX || xtype == CHOLMOD_ZOMPLEX)
        return (dtype == CHOLMOD_SINGLE) ? B200_COMPLEX_SINGLE : B200_COMPLEX_DOUBLE ;
    return (dtype == CHOLMOD_SINGLE) ? B200_SINGLE : B200_DOUBLE ;
}

int cholmod_l_b200_symbolic (cholmod_factor *L, cholmod_common *Common)
{
    if (!L || !Common || !L->is_super) return 0 ;
    // a symbolic factor does not know yet whether it will hold real or complex values: real is assumed,
    // the numeric phase builds the context of the other kind when it turns out to be needed
    shim_slot *s = slot_acquire (L, shim_dtype (L->xtype, L->dtype), Common, 1) ;
    if (!s) return 0 ;
    slot_release (s) ;
    return 1 ;
}

int cholmod_l_b200_deallocate (cholmod_common *Common)
{
    pthread_mutex_lock (&g_lock) ;
    slots_init () ;
    for (int k = 0 ; k < B200_SHIM_SLOTS ; k++)
    {
        shim_slot *s = &g_slots [k] ;
        if (s->refs > 0 || (!s->ctx && !s->multi)) continue ;      // a context in use stays (its holder is mid-call)
        if (s->host_stale) slot_materialize (s, Common) ;
        slot_destroy_ctx (s) ;
        s->key = 0 ; s->stamp = 0 ; s->last_L = NULL ; s->last_Lx = NULL ; s->values_valid = 0 ; s->host_stale = 0 ;
    }
    pthread_mutex_unlock (&g_lock) ;
    return 1 ;
}

// ---------------------------------------------------------------------------------------------
// host L->x: page-lock bookkeeping and the lazy copy
// ---------------------------------------------------------------------------------------------

typedef int (*free_factor_fn) (cholmod_factor **, cholmod_common *) ;
typedef int (*change_factor_fn) (int, int, int, int, int, cholmod_factor *, cholmod_common *) ;
typedef cholmod_factor *(*copy_factor_fn) (cholmod_factor *, cholmod_common *) ;
typedef double (*rcond_fn) (cholmod_factor *, cholmod_common *) ;
typedef int (*super_solve_fn) (cholmod_factor *, cholmod_dense *, cholmod_dense *, cholmod_common *) ;

// Host code is about to read, free or reallocate L->x.  fill != 0: first bring it up to date (lazy mode);
// forget != 0: the memory is going away, drop the page lock and the association with the device copy.
static void shim_host_access (cholmod_factor *L, cholmod_common *Common, int fill, int forget)
{
    if (!L || !L->x) return ;
    shim_slot *s = slot_of_factor (L) ;
    if (s)
    {
        if (fill) slot_materialize (s, Common) ;
        if (forget) { s->last_L = NULL ; s->last_Lx = NULL ; s->values_valid = 0 ; s->host_stale = 0 ; }
        slot_release (s) ;
    }
    if (!forget) return ;
    pthread_mutex_lock (&g_lock) ;
    for (int k = 0 ; k < B200_SHIM_SLOTS && g_slots_init ; k++)
    {
        shim_slot *t = &g_slots [k] ;
        if (t->refs > 0) continue ;
        if (t->multi) b200_multi_forget_host_buffer (t->multi, L->x) ;
        else if (t->ctx) b200_forget_host_buffer (t->ctx, L->x) ;
    }
    pthread_mutex_unlock (&g_lock) ;
}

// cholmod_l_free_factor (CHOLMOD/Utility/t_cholmod_free_factor.c): the numeric phase page-locks host L->x
// the first time it writes into it; release that before the reference frees the memory.
int cholmod_l_free_factor (cholmod_factor **L, cholmod_common *Common)
{
    if (L && *L) shim_host_access (*L, Common, 0, 1) ;
    free_factor_fn next = (free_factor_fn) dlsym (RTLD_NEXT, "cholmod_l_free_factor") ;
    return next ? next (L, Common) : 0 ;
}

// cholmod_l_change_factor (CHOLMOD/Cholesky/cholmod_change_factor.c) reads and then frees or reallocates L->x
// when a supernodal numeric factor is converted (to simplicial, to symbolic)
int cholmod_l_change_factor (int to_xtype, int to_ll, int to_super, int to_packed, int to_monotonic,
    cholmod_factor *L, cholmod_common *Common)
{
    if (L && L->is_super && L->xtype != CHOLMOD_PATTERN) shim_host_access (L, Common, 1, 1) ;
    change_factor_fn next = (change_factor_fn) dlsym (RTLD_NEXT, "cholmod_l_change_factor") ;
    return next ? next (to_xtype, to_ll, to_super, to_packed, to_monotonic, L, Common) : 0 ;
}

// readers of L->x inside CHOLMOD that do not go through change_factor
cholmod_factor *cholmod_l_copy_factor (cholmod_factor *L, cholmod_common *Common)
{
    if (L && L->is_super && L->xtype != CHOLMOD_PATTERN) shim_host_access (L, Common, 1, 0) ;
    copy_factor_fn next = (copy_factor_fn) dlsym (RTLD_NEXT, "cholmod_l_copy_factor") ;
    return next ? next (L, Common) : NULL ;
}

double cholmod_l_rcond (cholmod_factor *L, cholmod_common *Common)
{
    if (L && L->is_super && L->xtype != CHOLMOD_PATTERN) shim_host_access (L, Common, 1, 0) ;
    rcond_fn next = (rcond_fn) dlsym (RTLD_NEXT, "cholmod_l_rcond") ;
    return next ? next (L, Common) : -1 ;
}

// ---------------------------------------------------------------------------------------------
// cholmod_l_super_lsolve / cholmod_l_super_ltsolve (CHOLMOD/Supernodal/cholmod_super_solve.c:40-128, :140-230)
// ---------------------------------------------------------------------------------------------

static int shim_sweep (int which, const char *name, cholmod_factor *L, cholmod_dense *X, cholmod_dense *E,
    cholmod_common *Common)
{
    const int lazy = env_int ("CHOLMOD_B200_LAZY_LX", 0), dev_solve = lazy || env_int ("CHOLMOD_B200_DEVICE_SOLVE", 0) ;
    // what the device sweeps cover: a supernodal LL' factor in double precision (real or complex) whose values
    // are in the arena, X with the layout cholmod_solve2 gives Y (leading dimension = n)
    int ours = dev_solve && Common && L && X && E && L->is_super && L->is_ll && L->dtype == CHOLMOD_DOUBLE
        && (L->xtype == CHOLMOD_REAL || L->xtype == CHOLMOD_COMPLEX) && X->xtype == L->xtype && X->dtype == L->dtype
        && X->nrow == L->n && X->d == X->nrow && L->minor == L->n ;
    shim_slot *s = ours ? slot_of_factor (L) : NULL ;
    if (s && (!s->values_valid || s->multi)) { slot_release (s) ; s = NULL ; }
    if (s)
    {
        int st = B200_OK ;
        Common->status = CHOLMOD_OK ;
        if (L->n > 0 && X->ncol > 0) st = b200_solve_sweep (s->ctx, which, (const double *) X->x, (double *) X->x, (int64_t) X->ncol) ;
        slot_release (s) ;
        if (st < 0)
        {
            cholmod_l_error (st == B200_OUT_OF_MEMORY ? CHOLMOD_OUT_OF_MEMORY : CHOLMOD_GPU_PROBLEM, __FILE__, __LINE__,
                "CHOLMOD/B200 device solve failed", Common) ;
            return 0 ;
        }
        return 1 ;
    }
    // the host sweep of the reference reads L->x
    if (L && L->is_super && L->xtype != CHOLMOD_PATTERN) shim_host_access (L, Common, 1, 0) ;
    super_solve_fn next = (super_solve_fn) dlsym (RTLD_NEXT, name) ;
    return next ? next (L, X, E, Common) : 0 ;
}

int cholmod_l_super_lsolve (cholmod_factor *L, cholmod_dense *X, cholmod_dense *E, cholmod_common *Common)
{
    return shim_sweep (1, "cholmod_l_super_lsolve", L, X, E, Common) ;
}

int cholmod_l_super_ltsolve (cholmod_factor *L, cholmod_dense *X, cholmod_dense *E, cholmod_common *Common)
{
    return shim_sweep (2, "cholmod_l_super_ltsolve", L, X, E, Common) ;
}

// ---------------------------------------------------------------------------------------------
// the worker entry: same contract as rd_/rs_cholmod_super_numeric_worker
// ---------------------------------------------------------------------------------------------

// what cholmod_gpu_stats prints (CHOLMOD/Check/cholmod_check.c:615-674): BLAS call counts and times on the GPU.
// Counts are the reference's (one dsyrk per update pair, one dgemm per pair with rows below the ancestor's
// columns, one dpotrf per supernode, one dtrsm per supernode with rows below its diagonal block).  The kernels
// here are fused (syrk + gemm + assembly in one launch), so the device time of the factorization is
// apportioned to the four kinds by their flop counts.
static void shim_gpu_stats (shim_slot *s, cholmod_common *Common)
{
    b200_factor *f = s->multi ? b200_multi_context (s->multi, 0) : s->ctx ;
    if (!f) return ;
    Common->cholmod_gpu_syrk_calls += (size_t) b200_get_int (f, "npairs") ;
    Common->cholmod_gpu_gemm_calls += (size_t) b200_get_int (f, "npairs_gemm") ;
    Common->cholmod_gpu_potrf_calls += (size_t) b200_get_int (f, "nsuper") ;
    Common->cholmod_gpu_trsm_calls += (size_t) b200_get_int (f, "nsuper_trsm") ;
    const double t = 1e-3 * b200_get_double (f, "ms_factor") ;
    const double fs = b200_get_double (f, "flops_syrk"), fg = b200_get_double (f, "flops_gemm"),
        fp = b200_get_double (f, "flops_potrf"), ft = b200_get_double (f, "flops_trsm"), all = fs + fg + fp + ft ;
    if (t > 0 && all > 0)
    {
        Common->cholmod_gpu_syrk_time += t * fs / all ;
        Common->cholmod_gpu_gemm_time += t * fg / all ;
        Common->cholmod_gpu_potrf_time += t * fp / all ;
        Common->cholmod_gpu_trsm_time += t * ft / all ;
    }
    Common->gpuNumKernelLaunches += (int) b200_get_int (f, "kernel_launches") ;
    Common->gpuKernelTime += t ;
}

static int shim_worker (cholmod_sparse *A, cholmod_sparse *F, double beta [2], cholmod_factor *L,
    cholmod_common *Common, int dtype)
{
    shim_slot *s = slot_acquire (L, dtype, Common, A->stype != 0) ;
    if (!s)
    {
        cholmod_l_error (Common->status < 0 ? Common->status : CHOLMOD_GPU_PROBLEM, __FILE__, __LINE__,
            "CHOLMOD/B200: no device context for this factor", Common) ;
        return 0 ;
    }
    // lazy mode: the arena may still hold another factor's values that the host has not seen
    if (s->host_stale && (s->last_L != L || s->last_Lx != L->x)) slot_materialize (s, Common) ;
    const int lazy = env_int ("CHOLMOD_B200_LAZY_LX", 0) && !s->multi ;
    void *Lx = lazy ? NULL : L->x ;
    int64_t minor = (int64_t) L->n ;
    const int64_t *Anz = A->packed ? NULL : (const int64_t *) A->nz ;
    const int quick = Common->quick_return_if_not_posdef ;
    int st ;
    // zomplex A: real parts in A->x, imaginary parts in A->z (L is complex either way)
    const void *Az = (A->xtype == CHOLMOD_ZOMPLEX) ? A->z : NULL ;
    if (s->multi)
    {
        st = b200_multi_super_numeric_d (s->multi, (const int64_t *) A->p, (const int64_t *) A->i, Anz,
            (const double *) A->x, beta [0], quick, (double *) L->x, &minor) ;
    }
    else if (A->stype == 0)
    {
        // the unsymmetric case: A*F is factorized (t_cholmod_super_numeric_worker.c:471-506)
        const int64_t *Fnz = F->packed ? NULL : (const int64_t *) F->nz ;
        st = b200_super_numeric_aat (s->ctx, (int64_t) A->ncol, (const int64_t *) A->p, (const int64_t *) A->i, Anz,
            A->x, Az, (const int64_t *) F->p, (const int64_t *) F->i, Fnz, F->x,
            (F->xtype == CHOLMOD_ZOMPLEX) ? F->z : NULL, beta [0], quick, Lx, &minor) ;
    }
    else if (dtype == B200_DOUBLE)
        st = b200_super_numeric_d (s->ctx, (const int64_t *) A->p, (const int64_t *) A->i, Anz,
            (const double *) A->x, beta [0], quick, (double *) Lx, &minor) ;
    else if (dtype == B200_SINGLE)
        st = b200_super_numeric_s (s->ctx, (const int64_t *) A->p, (const int64_t *) A->i, Anz,
            (const float *) A->x, beta [0], quick, (float *) Lx, &minor) ;
    else if (dtype == B200_COMPLEX_DOUBLE)
        st = b200_super_numeric_z (s->ctx, (const int64_t *) A->p, (const int64_t *) A->i, Anz,
            (const double *) A->x, (const double *) Az, beta [0], quick, (double *) Lx, &minor) ;
    else
        st = b200_super_numeric_c (s->ctx, (const int64_t *) A->p, (const int64_t *) A->i, Anz,
            (const float *) A->x, (const float *) Az, beta [0], quick, (float *) Lx, &minor) ;
    if (env_int ("CHOLMOD_B200_TRACE", 0))
        fprintf (stderr, "cholmod_b200: numeric factorization on device%s %d (n = %lld, status %d)\n",
            s->multi ? "s from" : "", shim_device (), (long long) L->n, st) ;
    s->last_L = L ; s->last_Lx = L->x ;
    s->values_valid = (st == B200_OK) ;
    s->host_stale = (lazy && st >= 0) ;
    if (st >= 0) shim_gpu_stats (s, Common) ;
    slot_release (s) ;
    if (st < 0)
    {
        cholmod_l_error (st, __FILE__, __LINE__, "CHOLMOD/B200 numeric factorization failed", Common) ;
        return 0 ;
    }
    L->minor = (size_t) minor ;
    if (st == B200_NOT_POSDEF)
        cholmod_l_error (CHOLMOD_NOT_POSDEF, __FILE__, __LINE__, "matrix not positive definite", Common) ;
    return 1 ;      // TRUE whether or not the matrix is positive definite (SNW:130-131)
}

int rd_cholmod_l_b200_super_numeric (cholmod_sparse *A, cholmod_sparse *F, double beta [2],
    cholmod_factor *L, cholmod_common *Common)
{
    return shim_worker (A, F, beta, L, Common, B200_DOUBLE) ;
}

int rs_cholmod_l_b200_super_numeric (cholmod_sparse *A, cholmod_sparse *F, double beta [2],
    cholmod_factor *L, cholmod_common *Common)
{
    return shim_worker (A, F, beta, L, Common, B200_SINGLE) ;
}

// cd_/zd_ (complex double L; A complex or zomplex) and cs_/zs_ (complex single)
int cd_cholmod_l_b200_super_numeric (cholmod_sparse *A, cholmod_sparse *F, double beta [2],
    cholmod_factor *L, cholmod_common *Common)
{
    return shim_worker (A, F, beta, L, Common, B200_COMPLEX_DOUBLE) ;
}

int zd_cholmod_l_b200_super_numeric (cholmod_sparse *A, cholmod_sparse *F, double beta [2],
    cholmod_factor *L, cholmod_common *Common)
{
    return shim_worker (A, F, beta, L, Common, B200_COMPLEX_DOUBLE) ;
}

int cs_cholmod_l_b200_super_numeric (cholmod_sparse *A, cholmod_sparse *F, double beta [2],
    cholmod_factor *L, cholmod_common *Common)
{
    return shim_worker (A, F, beta, L, Common, B200_COMPLEX_SINGLE) ;
}

int zs_cholmod_l_b200_super_numeric (cholmod_sparse *A, cholmod_sparse *F, double beta [2],
    cholmod_factor *L, cholmod_common *Common)
{
    return shim_worker (A, F, beta, L, Common, B200_COMPLEX_SINGLE) ;
}

// ---------------------------------------------------------------------------------------------
// cholmod_l_super_numeric: the public seam (cholmod_super_numeric.c:100-335)
// ---------------------------------------------------------------------------------------------

typedef int (*super_numeric_fn) (cholmod_sparse *, cholmod_sparse *, double *, cholmod_factor *, cholmod_common *) ;

int cholmod_l_super_numeric (cholmod_sparse *A, cholmod_sparse *F, double beta [2], cholmod_factor *L,
    cholmod_common *Common)
{
    // Activation: Common->useGPU == 1, or the reference's own environment switch CHOLMOD_USE_GPU=1
    // (cholmod_super_symbolic.c:250-289).  A libcholmod built without CHOLMOD_HAS_CUDA resets
    // Common->useGPU to 0 during analysis (cholmod_super_symbolic.c:320), so with an unpatched CPU-only
    // libcholmod the environment variable is the switch.
    const char *env = getenv ("CHOLMOD_USE_GPU") ;
    int want = Common && (Common->useGPU == 1 || (env && atoi (env) == 1)) ;
    // scope of the module: int64; symmetric / Hermitian with the lower triangle stored, or unsymmetric with
    // F = A^H given; real, complex or zomplex; double or single (the argument checks are the reference's,
    // cholmod_super_numeric.c:127-180; anything they would reject is left for the reference to reject)
    int cplx = A && (A->xtype == CHOLMOD_COMPLEX || A->xtype == CHOLMOD_ZOMPLEX) ;
    int shape = A && L && ((A->stype < 0 && A->nrow == A->ncol && A->nrow == L->n)
        || (A->stype == 0 && A->nrow == L->n && F && F->stype == 0 && F->nrow == A->ncol && F->ncol == A->nrow
            && F->xtype == A->xtype && F->dtype == A->dtype && F->itype == CHOLMOD_LONG)) ;
    int ours = want && shape && L->is_super && (A->xtype == CHOLMOD_REAL || cplx)
        && A->itype == CHOLMOD_LONG && L->itype == CHOLMOD_LONG
        && (L->xtype == CHOLMOD_PATTERN || L->xtype == (cplx ? CHOLMOD_COMPLEX : CHOLMOD_REAL)) ;
    // several GPUs: the real double case (north_star's); complex and single precision run on one device
    if (!ours)
    {
        if (L && L->is_super && L->xtype != CHOLMOD_PATTERN) shim_host_access (L, Common, 0, 0) ;
        super_numeric_fn next = (super_numeric_fn) dlsym (RTLD_NEXT, "cholmod_l_super_numeric") ;
        if (next) return next (A, F, beta, L, Common) ;
        if (Common) cholmod_l_error (CHOLMOD_INVALID, __FILE__, __LINE__, "input outside the B200 module", Common) ;
        return 0 ;
    }
    if (L->xtype != CHOLMOD_PATTERN && A->dtype != L->dtype)
    {
        cholmod_l_error (CHOLMOD_INVALID, __FILE__, __LINE__, "A and L must have the same dtype", Common) ;
        return 0 ;
    }
    Common->status = CHOLMOD_OK ;
    if (L->xtype == CHOLMOD_PATTERN)
    {
        // allocate L->x on the host, as cholmod_super_numeric.c:219-231 does
        L->dtype = A->dtype ;
        cholmod_l_change_factor (cplx ? CHOLMOD_COMPLEX : CHOLMOD_REAL, 1, 1, 1, 1, L, Common) ;
        if (Common->status < CHOLMOD_OK) return 0 ;
    }
    L->is_ll = 1 ;
    switch ((A->xtype + A->dtype) % 8)      // the dispatch of cholmod_super_numeric.c:293-322
    {
        case CHOLMOD_REAL    + CHOLMOD_SINGLE: return rs_cholmod_l_b200_super_numeric (A, F, beta, L, Common) ;
        case CHOLMOD_COMPLEX + CHOLMOD_SINGLE: return cs_cholmod_l_b200_super_numeric (A, F, beta, L, Common) ;
        case CHOLMOD_ZOMPLEX + CHOLMOD_SINGLE: return zs_cholmod_l_b200_super_numeric (A, F, beta, L, Common) ;
        case CHOLMOD_COMPLEX + CHOLMOD_DOUBLE: return cd_cholmod_l_b200_super_numeric (A, F, beta, L, Common) ;
        case CHOLMOD_ZOMPLEX + CHOLMOD_DOUBLE: return zd_cholmod_l_b200_super_numeric (A, F, beta, L, Common) ;
        default: break ;
    }
    return rd_cholmod_l_b200_super_numeric (A, F, beta, L, Common) ;
}
