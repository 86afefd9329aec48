// cholmod_b200_shim.c -- the cholmod_common / cholmod_factor level glue of the CHOLMOD/B200 module.
//
// This is the thin C file a CHOLMOD maintainer adds next to CHOLMOD/GPU.  It is
// compiled against the reference's own cholmod.h and exports, with CHOLMOD's naming conventions
// (SURVEY.md §8b):
//
//   cholmod_l_b200_probe            <->  cholmod_l_gpu_probe            (CHOLMOD/GPU/cholmod_gpu.c:148)
//   cholmod_l_b200_symbolic         <->  gpu_init + cholmod_l_gpu_allocate (t_cholmod_gpu.c:84, cholmod_gpu.c:426)
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
//
// A failure of the device path is reported as CHOLMOD_GPU_PROBLEM; it is NOT retried on the CPU.

#define _GNU_SOURCE
#include "cholmod.h"
#include "../../include/cholmod_b200.h"
#include <dlfcn.h>
#include <stdint.h>
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
    uint64_t key ;
    int64_t n, nsuper, ssize, xsize ;
    int dtype ;
    uint64_t stamp ;
} shim_slot ;

static shim_slot g_slots [B200_SHIM_SLOTS] ;
static uint64_t g_clock = 0 ;
static pthread_mutex_t g_lock = PTHREAD_MUTEX_INITIALIZER ;

static uint64_t hash_i64 (uint64_t h, const int64_t *p, size_t n)
{
    for (size_t k = 0 ; k < n ; k++)
    {
        h ^= (uint64_t) p [k] ;
        h *= 0x100000001b3ULL ;
    }
    return h ;
}

static uint64_t factor_key (const cholmod_factor *L)
{
    uint64_t h = 0xcbf29ce484222325ULL ;
    h = hash_i64 (h, (const int64_t *) L->super, L->nsuper + 1) ;
    h = hash_i64 (h, (const int64_t *) L->pi, L->nsuper + 1) ;
    h = hash_i64 (h, (const int64_t *) L->s, L->ssize) ;
    return h ;
}

int cholmod_l_b200_probe (cholmod_common *Common)
{
    (void) Common ;
    int dev = 0 ;
    const char *e = getenv ("CHOLMOD_B200_DEVICE") ;
    if (e) dev = atoi (e) ;
    return b200_probe (dev) ;
}

// find or build the device context of a supernodal symbolic factor; NULL + Common->status on failure
static b200_factor *shim_context (cholmod_factor *L, int dtype, cholmod_common *Common)
{
    uint64_t key = factor_key (L) ;
    b200_factor *ctx = NULL ;
    pthread_mutex_lock (&g_lock) ;
    int victim = 0 ;
    for (int k = 0 ; k < B200_SHIM_SLOTS ; k++)
    {
        shim_slot *s = &g_slots [k] ;
        if (s->ctx && s->key == key && s->n == (int64_t) L->n && s->nsuper == (int64_t) L->nsuper
            && s->ssize == (int64_t) L->ssize && s->xsize == (int64_t) L->xsize && s->dtype == dtype)
        {
            s->stamp = ++g_clock ;
            ctx = s->ctx ;
            break ;
        }
        if (g_slots [k].stamp < g_slots [victim].stamp) victim = k ;
    }
    if (!ctx)
    {
        int dev = 0, status = 0 ;
        const char *e = getenv ("CHOLMOD_B200_DEVICE") ;
        if (e) dev = atoi (e) ;
        ctx = b200_create ((int64_t) L->n, (int64_t) L->nsuper, (const int64_t *) L->super,
            (const int64_t *) L->pi, (const int64_t *) L->px, (const int64_t *) L->s, dtype, dev, &status) ;
        if (!ctx)
        {
            Common->status = status ;       // the B200_* codes are CHOLMOD's
        }
        else
        {
            shim_slot *s = &g_slots [victim] ;
            if (s->ctx) b200_destroy (s->ctx) ;
            s->ctx = ctx ; s->key = key ; s->n = (int64_t) L->n ; s->nsuper = (int64_t) L->nsuper ;
            s->ssize = (int64_t) L->ssize ; s->xsize = (int64_t) L->xsize ; s->dtype = dtype ;
            s->stamp = ++g_clock ;
        }
    }
    pthread_mutex_unlock (&g_lock) ;
    return ctx ;
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
    return shim_context (L, shim_dtype (L->xtype, L->dtype), Common) != NULL ;
}

int cholmod_l_b200_deallocate (cholmod_common *Common)
{
    (void) Common ;
    pthread_mutex_lock (&g_lock) ;
    for (int k = 0 ; k < B200_SHIM_SLOTS ; k++)
    {
        if (g_slots [k].ctx) b200_destroy (g_slots [k].ctx) ;
        memset (&g_slots [k], 0, sizeof (shim_slot)) ;
    }
    pthread_mutex_unlock (&g_lock) ;
    return 1 ;
}

// cholmod_l_free_factor (CHOLMOD/Utility/t_cholmod_free_factor.c): the numeric phase page-locks host L->x
// the first time it writes into it; release that before the reference frees the memory.
typedef int (*free_factor_fn) (cholmod_factor **, cholmod_common *) ;
typedef int (*change_factor_fn) (int, int, int, int, int, cholmod_factor *, cholmod_common *) ;

static void shim_forget (const void *Lx)
{
    if (!Lx) return ;
    pthread_mutex_lock (&g_lock) ;
    for (int k = 0 ; k < B200_SHIM_SLOTS ; k++)
        if (g_slots [k].ctx) b200_forget_host_buffer (g_slots [k].ctx, Lx) ;
    pthread_mutex_unlock (&g_lock) ;
}

int cholmod_l_free_factor (cholmod_factor **L, cholmod_common *Common)
{
    if (L && *L) shim_forget ((*L)->x) ;
    free_factor_fn next = (free_factor_fn) dlsym (RTLD_NEXT, "cholmod_l_free_factor") ;
    return next ? next (L, Common) : 0 ;
}

// cholmod_l_change_factor (CHOLMOD/Cholesky/cholmod_change_factor.c) frees or reallocates L->x when a
// supernodal numeric factor is converted (to simplicial, to symbolic): same release first.
int cholmod_l_change_factor (int to_xtype, int to_ll, int to_super, int to_packed, int to_monotonic,
    cholmod_factor *L, cholmod_common *Common)
{
    if (L && L->is_super && L->xtype != CHOLMOD_PATTERN) shim_forget (L->x) ;
    change_factor_fn next = (change_factor_fn) dlsym (RTLD_NEXT, "cholmod_l_change_factor") ;
    return next ? next (to_xtype, to_ll, to_super, to_packed, to_monotonic, L, Common) : 0 ;
}

// ---------------------------------------------------------------------------------------------
// the worker entry: same contract as rd_/rs_cholmod_super_numeric_worker for stype < 0
// ---------------------------------------------------------------------------------------------

static int shim_worker (cholmod_sparse *A, cholmod_sparse *F, double beta [2], cholmod_factor *L,
    cholmod_common *Common, int dtype)
{
    b200_factor *ctx = shim_context (L, dtype, Common) ;
    if (!ctx) return 0 ;
    int64_t minor = (int64_t) L->n ;
    const int64_t *Anz = A->packed ? NULL : (const int64_t *) A->nz ;
    int st ;
    // zomplex A: real parts in A->x, imaginary parts in A->z (L is complex either way)
    const void *Az = (A->xtype == CHOLMOD_ZOMPLEX) ? A->z : NULL ;
    if (A->stype == 0)
    {
        // the unsymmetric case: A*F is factorized (t_cholmod_super_numeric_worker.c:471-506)
        const int64_t *Fnz = F->packed ? NULL : (const int64_t *) F->nz ;
        st = b200_super_numeric_aat (ctx, (int64_t) A->ncol, (const int64_t *) A->p, (const int64_t *) A->i, Anz,
            A->x, Az, (const int64_t *) F->p, (const int64_t *) F->i, Fnz, F->x,
            (F->xtype == CHOLMOD_ZOMPLEX) ? F->z : NULL, beta [0], Common->quick_return_if_not_posdef, L->x, &minor) ;
    }
    else if (dtype == B200_DOUBLE)
        st = b200_super_numeric_d (ctx, (const int64_t *) A->p, (const int64_t *) A->i, Anz,
            (const double *) A->x, beta [0], Common->quick_return_if_not_posdef, (double *) L->x, &minor) ;
    else if (dtype == B200_SINGLE)
        st = b200_super_numeric_s (ctx, (const int64_t *) A->p, (const int64_t *) A->i, Anz,
            (const float *) A->x, beta [0], Common->quick_return_if_not_posdef, (float *) L->x, &minor) ;
    else if (dtype == B200_COMPLEX_DOUBLE)
        st = b200_super_numeric_z (ctx, (const int64_t *) A->p, (const int64_t *) A->i, Anz,
            (const double *) A->x, (const double *) Az, beta [0], Common->quick_return_if_not_posdef,
            (double *) L->x, &minor) ;
    else
        st = b200_super_numeric_c (ctx, (const int64_t *) A->p, (const int64_t *) A->i, Anz,
            (const float *) A->x, (const float *) Az, beta [0], Common->quick_return_if_not_posdef,
            (float *) L->x, &minor) ;
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
    if (!ours)
    {
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
