// oracle/ref_driver.c -- TEST INFRASTRUCTURE (oracle side), not product code.
//
// A plain-pointer C driver over the *unmodified* reference CHOLMOD built into
// oracle/_ref/libcholmod_ref.so.  It gives tests/ and bench.py's CPU-baseline legs:
//   * the host symbolic analysis (ordering, etree, supernode partition) -- which by
//     BASELINE.json's north_star stays the reference's own host code;
//   * the matrix exactly as it arrives at the numeric seam, S = tril(P A P') in CSC,
//     built the way cholmod_factorize_p does it (CHOLMOD/Cholesky/cholmod_factorize.c:172-250);
//   * the reference CPU supernodal numeric phase at that seam
//     (cholmod_l_super_numeric, CHOLMOD/Supernodal/cholmod_super_numeric.c:100) and through the
//     whole driver (cholmod_l_factorize_p, cholmod_factorize.c:93);
//   * the reference solve and the demo's scaled residual
//     (CHOLMOD/Demo/cholmod_dl_demo.c:609-617) applied to ANY factor values, so the factor the
//     CUDA path produced can be pushed through the reference's own cholmod_l_solve.
//
// Everything is int64 ("cholmod_l_"); real, complex or zomplex; dtype is double or single.

#include "cholmod.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

typedef struct
{
    cholmod_common cm ;
    cholmod_sparse *A ;     // user matrix (symmetric, one triangle stored)
    cholmod_sparse *S ;     // seam matrix: tril (P A P'), stype -1 (or A(p,:), stype 0, when A is unsymmetric)
    cholmod_sparse *F ;     // unsymmetric case only: F = A(p,:)^H, so that S*F is what gets factorized
    cholmod_factor *L ;
    double t_analyze, t_factor, t_solve ;
} refd ;

static void quiet_handler (int status, const char *file, int line, const char *msg)
{
    (void) status ; (void) file ; (void) line ; (void) msg ;
}

void *refd_open (int nthreads_max)
{
    refd *h = (refd *) calloc (1, sizeof (refd)) ;
    if (!h) return NULL ;
    cholmod_l_start (&h->cm) ;
    h->cm.useGPU = 0 ;
    h->cm.error_handler = quiet_handler ;
    h->cm.print = 0 ;
    if (nthreads_max > 0) h->cm.nthreads_max = nthreads_max ;
    return h ;
}

void refd_close (void *hv)
{
    refd *h = (refd *) hv ;
    if (!h) return ;
    cholmod_l_free_factor (&h->L, &h->cm) ;
    cholmod_l_free_sparse (&h->S, &h->cm) ;
    cholmod_l_free_sparse (&h->F, &h->cm) ;
    cholmod_l_free_sparse (&h->A, &h->cm) ;
    cholmod_l_finish (&h->cm) ;
    free (h) ;
}

// A: n-by-n CSC, one triangle stored (stype -1 lower / +1 upper), int64 indices.
// single != 0 stores the values as float (CHOLMOD_SINGLE).
int refd_set_matrix (void *hv, int64_t n, const int64_t *Ap, const int64_t *Ai,
    const double *Ax, int stype, int single)
{
    refd *h = (refd *) hv ;
    cholmod_l_free_sparse (&h->A, &h->cm) ;
    cholmod_l_free_sparse (&h->S, &h->cm) ;
    cholmod_l_free_sparse (&h->F, &h->cm) ;
    cholmod_l_free_factor (&h->L, &h->cm) ;
    int64_t nz = Ap [n] ;
    int xdtype = CHOLMOD_REAL + (single ? CHOLMOD_SINGLE : CHOLMOD_DOUBLE) ;
    h->A = cholmod_l_allocate_sparse ((size_t) n, (size_t) n, (size_t) nz, 1, 1, stype, xdtype, &h->cm) ;
    if (!h->A) return -1 ;
    memcpy (h->A->p, Ap, (size_t) (n+1) * sizeof (int64_t)) ;
    memcpy (h->A->i, Ai, (size_t) nz * sizeof (int64_t)) ;
    if (single)
    {
        float *x = (float *) h->A->x ;
        for (int64_t p = 0 ; p < nz ; p++) x [p] = (float) Ax [p] ;
    }
    else
    {
        memcpy (h->A->x, Ax, (size_t) nz * sizeof (double)) ;
    }
    return 0 ;
}

// The same with an xtype: CHOLMOD_REAL (1; Ax real), CHOLMOD_COMPLEX (2; Ax = nz (re,im) pairs) or
// CHOLMOD_ZOMPLEX (3; Ax real parts, Az imaginary parts).  A is Hermitian, one triangle stored.
int refd_set_matrix_x (void *hv, int64_t n, const int64_t *Ap, const int64_t *Ai,
    const double *Ax, const double *Az, int stype, int single, int xtype)
{
    refd *h = (refd *) hv ;
    if (xtype == CHOLMOD_REAL) return refd_set_matrix (hv, n, Ap, Ai, Ax, stype, single) ;
    cholmod_l_free_sparse (&h->A, &h->cm) ;
    cholmod_l_free_sparse (&h->S, &h->cm) ;
    cholmod_l_free_sparse (&h->F, &h->cm) ;
    cholmod_l_free_factor (&h->L, &h->cm) ;
    int64_t nz = Ap [n] ;
    int xdtype = xtype + (single ? CHOLMOD_SINGLE : CHOLMOD_DOUBLE) ;
    h->A = cholmod_l_allocate_sparse ((size_t) n, (size_t) n, (size_t) nz, 1, 1, stype, xdtype, &h->cm) ;
    if (!h->A) return -1 ;
    memcpy (h->A->p, Ap, (size_t) (n+1) * sizeof (int64_t)) ;
    memcpy (h->A->i, Ai, (size_t) nz * sizeof (int64_t)) ;
    int64_t nx = (xtype == CHOLMOD_COMPLEX) ? 2*nz : nz ;
    if (single)
    {
        float *x = (float *) h->A->x ;
        for (int64_t p = 0 ; p < nx ; p++) x [p] = (float) Ax [p] ;
        if (xtype == CHOLMOD_ZOMPLEX) { float *z = (float *) h->A->z ; for (int64_t p = 0 ; p < nz ; p++) z [p] = (float) Az [p] ; }
    }
    else
    {
        memcpy (h->A->x, Ax, (size_t) nx * sizeof (double)) ;
        if (xtype == CHOLMOD_ZOMPLEX) memcpy (h->A->z, Az, (size_t) nz * sizeof (double)) ;
    }
    return 0 ;
}

// An unsymmetric (stype 0) nrow-by-ncol matrix: cholmod_l_analyze / factorize then work on A*A^H
// (cholmod_factorize.c:211-247).  xtype as for refd_set_matrix_x.
int refd_set_matrix_rect (void *hv, int64_t nrow, int64_t ncol, const int64_t *Ap, const int64_t *Ai,
    const double *Ax, const double *Az, int single, int xtype)
{
    refd *h = (refd *) hv ;
    cholmod_l_free_sparse (&h->A, &h->cm) ;
    cholmod_l_free_sparse (&h->S, &h->cm) ;
    cholmod_l_free_sparse (&h->F, &h->cm) ;
    cholmod_l_free_factor (&h->L, &h->cm) ;
    int64_t nz = Ap [ncol] ;
    int xdtype = xtype + (single ? CHOLMOD_SINGLE : CHOLMOD_DOUBLE) ;
    h->A = cholmod_l_allocate_sparse ((size_t) nrow, (size_t) ncol, (size_t) nz, 1, 1, 0, xdtype, &h->cm) ;
    if (!h->A) return -1 ;
    memcpy (h->A->p, Ap, (size_t) (ncol+1) * sizeof (int64_t)) ;
    memcpy (h->A->i, Ai, (size_t) nz * sizeof (int64_t)) ;
    int64_t nx = (xtype == CHOLMOD_COMPLEX) ? 2*nz : nz ;
    if (single)
    {
        float *x = (float *) h->A->x ;
        for (int64_t p = 0 ; p < nx ; p++) x [p] = (float) Ax [p] ;
        if (xtype == CHOLMOD_ZOMPLEX) { float *z = (float *) h->A->z ; for (int64_t p = 0 ; p < nz ; p++) z [p] = (float) Az [p] ; }
    }
    else
    {
        memcpy (h->A->x, Ax, (size_t) nx * sizeof (double)) ;
        if (xtype == CHOLMOD_ZOMPLEX) memcpy (h->A->z, Az, (size_t) nz * sizeof (double)) ;
    }
    return 0 ;
}

// read a Matrix Market / triplet file with the reference's own reader (golden generation only)
int refd_read_matrix (void *hv, const char *path)
{
    refd *h = (refd *) hv ;
    FILE *f = fopen (path, "r") ;
    if (!f) return -1 ;
    cholmod_l_free_sparse (&h->A, &h->cm) ;
    cholmod_l_free_sparse (&h->S, &h->cm) ;
    cholmod_l_free_sparse (&h->F, &h->cm) ;
    cholmod_l_free_factor (&h->L, &h->cm) ;
    h->A = cholmod_l_read_sparse2 (f, CHOLMOD_DOUBLE, &h->cm) ;
    fclose (f) ;
    return h->A ? 0 : -2 ;
}

// ordering: CHOLMOD_NATURAL 0, GIVEN 1 (unused), AMD 2, METIS 3, NESDIS 4; -1 = default strategy.
// supernodal: CHOLMOD_SIMPLICIAL 0, AUTO 1, SUPERNODAL 2.
int refd_analyze (void *hv, int ordering, int supernodal)
{
    refd *h = (refd *) hv ;
    if (!h->A) return -1 ;
    cholmod_l_free_factor (&h->L, &h->cm) ;
    cholmod_l_free_sparse (&h->S, &h->cm) ;
    cholmod_l_free_sparse (&h->F, &h->cm) ;
    h->cm.supernodal = supernodal ;
    if (ordering >= 0)
    {
        h->cm.nmethods = 1 ;
        h->cm.method [0].ordering = ordering ;
        h->cm.postorder = 1 ;
    }
    double t = SUITESPARSE_TIME ;
    h->L = cholmod_l_analyze (h->A, &h->cm) ;
    h->t_analyze = SUITESPARSE_TIME - t ;
    if (!h->L) return -2 ;
    return h->cm.status ;
}

// build the seam matrix S = tril (P A P') exactly as cholmod_factorize_p does
int refd_seam (void *hv)
{
    refd *h = (refd *) hv ;
    if (!h->A || !h->L) return -1 ;
    cholmod_l_free_sparse (&h->S, &h->cm) ;
    cholmod_l_free_sparse (&h->F, &h->cm) ;
    cholmod_sparse *A = h->A, *A1 = NULL, *A2 = NULL ;
    int64_t *Perm = (h->L->ordering == CHOLMOD_NATURAL) ? NULL : (int64_t *) h->L->Perm ;
    if (A->stype == 0)
    {
        // F = A(p,:)^H, S = F^H = A(p,:)   (cholmod_factorize.c:201-209, :236-247)
        h->F = cholmod_l_ptranspose (A, 2, Perm, NULL, 0, &h->cm) ;
        if (!h->F) return -2 ;
        h->S = cholmod_l_ptranspose (h->F, 2, NULL, NULL, 0, &h->cm) ;
        return h->S ? 0 : -2 ;
    }
    if (A->stype > 0)
    {
        A1 = cholmod_l_ptranspose (A, 2, Perm, NULL, 0, &h->cm) ;
    }
    else if (Perm == NULL)
    {
        A1 = cholmod_l_copy_sparse (A, &h->cm) ;
    }
    else
    {
        A2 = cholmod_l_ptranspose (A, 2, Perm, NULL, 0, &h->cm) ;
        A1 = cholmod_l_ptranspose (A2, 2, NULL, NULL, 0, &h->cm) ;
        cholmod_l_free_sparse (&A2, &h->cm) ;
    }
    h->S = A1 ;
    return h->S ? 0 : -2 ;
}

// reference CPU numeric phase at the seam: cholmod_l_super_numeric (S, NULL, beta, L)
int refd_super_numeric (void *hv, double beta)
{
    refd *h = (refd *) hv ;
    if (!h->S || !h->L) return -1 ;
    double b [2] = { beta, 0 } ;
    double t = SUITESPARSE_TIME ;
    int ok = cholmod_l_super_numeric (h->S, h->F, b, h->L, &h->cm) ;
    h->t_factor = SUITESPARSE_TIME - t ;
    return ok ? h->cm.status : -100 + h->cm.status ;
}

// the whole reference driver: cholmod_l_factorize_p (A, beta, NULL, 0, L)
int refd_factorize (void *hv, double beta)
{
    refd *h = (refd *) hv ;
    if (!h->A || !h->L) return -1 ;
    double b [2] = { beta, 0 } ;
    double t = SUITESPARSE_TIME ;
    int ok = cholmod_l_factorize_p (h->A, b, NULL, 0, h->L, &h->cm) ;
    h->t_factor = SUITESPARSE_TIME - t ;
    return ok ? h->cm.status : -100 + h->cm.status ;
}

// make sure L has a numeric part (what cholmod_super_numeric.c:219-231 does) and return it,
// so that externally computed factor values can be written into host L->x.
void *refd_numeric_Lx (void *hv, int single)
{
    refd *h = (refd *) hv ;
    if (!h->L || !h->L->is_super) return NULL ;
    if (h->L->xtype == CHOLMOD_PATTERN)
    {
        h->L->dtype = single ? CHOLMOD_SINGLE : CHOLMOD_DOUBLE ;
        // a complex or zomplex A gives a complex supernodal L (cholmod_super_numeric.c:219-231)
        int xt = (h->A && h->A->xtype != CHOLMOD_REAL) ? CHOLMOD_COMPLEX : CHOLMOD_REAL ;
        cholmod_l_change_factor (xt, 1, 1, 1, 1, h->L, &h->cm) ;
        if (h->cm.status < CHOLMOD_OK) return NULL ;
    }
    h->L->is_ll = 1 ;
    h->L->minor = h->L->n ;
    return h->L->x ;
}

void refd_set_minor (void *hv, int64_t minor) { ((refd *) hv)->L->minor = (size_t) minor ; }

// x = A\b with the reference's cholmod_l_solve (CHOLMOD_A); b, x are n-by-nrhs column-major doubles
// ((re,im) pairs of doubles when the factor is complex)
int refd_solve (void *hv, const double *b, double *x, int64_t nrhs)
{
    refd *h = (refd *) hv ;
    if (!h->L || !h->A) return -1 ;
    int64_t n = (int64_t) h->L->n ;
    int single = (h->L->dtype == CHOLMOD_SINGLE) ;
    int cplx = (h->L->xtype == CHOLMOD_COMPLEX) ;
    cholmod_dense *B = cholmod_l_allocate_dense ((size_t) n, (size_t) nrhs, (size_t) n,
        (cplx ? CHOLMOD_COMPLEX : CHOLMOD_REAL) + h->L->dtype, &h->cm) ;
    if (!B) return -2 ;
    if (cplx) n *= 2 ;      // from here on n counts doubles per column
    if (single) { float *bx = (float *) B->x ; for (int64_t i = 0 ; i < n*nrhs ; i++) bx [i] = (float) b [i] ; }
    else memcpy (B->x, b, (size_t) (n*nrhs) * sizeof (double)) ;
    double t = SUITESPARSE_TIME ;
    cholmod_dense *X = cholmod_l_solve (CHOLMOD_A, h->L, B, &h->cm) ;
    h->t_solve = SUITESPARSE_TIME - t ;
    cholmod_l_free_dense (&B, &h->cm) ;
    if (!X) return -3 ;
    if (single) { float *xx = (float *) X->x ; for (int64_t i = 0 ; i < n*nrhs ; i++) x [i] = xx [i] ; }
    else memcpy (x, X->x, (size_t) (n*nrhs) * sizeof (double)) ;
    cholmod_l_free_dense (&X, &h->cm) ;
    return 0 ;
}

// scaled residual ||A x - b||inf / (||A||inf ||x||inf + ||b||inf), one RHS, computed in double
double refd_resid (void *hv, const double *x, const double *b)
{
    refd *h = (refd *) hv ;
    cholmod_sparse *A = h->A ;
    int64_t n = (int64_t) A->nrow ;
    const int64_t *Ap = (const int64_t *) A->p, *Ai = (const int64_t *) A->i ;
    int single = (A->dtype == CHOLMOD_SINGLE) ;
    const double *Ad = (const double *) A->x ; const float *Af = (const float *) A->x ;
    double *r = (double *) malloc ((size_t) n * sizeof (double)) ;
    double *rowsum = (double *) calloc ((size_t) n, sizeof (double)) ;
    for (int64_t i = 0 ; i < n ; i++) r [i] = -b [i] ;
    for (int64_t j = 0 ; j < n ; j++)
    {
        for (int64_t p = Ap [j] ; p < Ap [j+1] ; p++)
        {
            int64_t i = Ai [p] ;
            double a = single ? (double) Af [p] : Ad [p] ;
            if (A->stype < 0 && i < j) continue ;
            if (A->stype > 0 && i > j) continue ;
            r [i] += a * x [j] ; rowsum [i] += (a < 0 ? -a : a) ;
            if (i != j) { r [j] += a * x [i] ; rowsum [j] += (a < 0 ? -a : a) ; }
        }
    }
    double rn = 0, xn = 0, bn = 0, an = 0 ;
    for (int64_t i = 0 ; i < n ; i++)
    {
        double v = r [i] < 0 ? -r [i] : r [i] ; if (v > rn) rn = v ;
        v = x [i] < 0 ? -x [i] : x [i] ; if (v > xn) xn = v ;
        v = b [i] < 0 ? -b [i] : b [i] ; if (v > bn) bn = v ;
        if (rowsum [i] > an) an = rowsum [i] ;
    }
    free (r) ; free (rowsum) ;
    return rn / (an * xn + bn + (n == 0 ? 1 : 0)) ;
}

// cholmod_l_rcond of the current factor (reads the diagonal of L->x on the host)
double refd_rcond (void *hv)
{
    refd *h = (refd *) hv ;
    if (!h->L) return -1 ;
    return cholmod_l_rcond (h->L, &h->cm) ;
}

int refd_set_int (void *hv, const char *name, int64_t v)
{
    refd *h = (refd *) hv ;
    if (!strcmp (name, "quick_return_if_not_posdef")) { h->cm.quick_return_if_not_posdef = (int) v ; return 0 ; }
    if (!strcmp (name, "nthreads_max")) { h->cm.nthreads_max = (int) v ; return 0 ; }
    if (!strcmp (name, "print")) { h->cm.print = (int) v ; return 0 ; }
    if (!strcmp (name, "try_catch")) { h->cm.try_catch = (int) v ; return 0 ; }
    if (!strcmp (name, "useGPU")) { h->cm.useGPU = (int) v ; return 0 ; }
    if (!strcmp (name, "maxGpuMemBytes")) { h->cm.maxGpuMemBytes = (size_t) v ; return 0 ; }
    return -1 ;
}

int64_t refd_get_int (void *hv, const char *name)
{
    refd *h = (refd *) hv ;
    cholmod_factor *L = h->L ;
    if (!strcmp (name, "status")) return h->cm.status ;
    if (!strcmp (name, "nthreads_max")) return h->cm.nthreads_max ;
    if (!strcmp (name, "useGPU")) return h->cm.useGPU ;
    if (!strcmp (name, "gpu_syrk_calls")) return (int64_t) h->cm.cholmod_gpu_syrk_calls ;
    if (!strcmp (name, "gpu_gemm_calls")) return (int64_t) h->cm.cholmod_gpu_gemm_calls ;
    if (!strcmp (name, "gpu_potrf_calls")) return (int64_t) h->cm.cholmod_gpu_potrf_calls ;
    if (!strcmp (name, "gpu_trsm_calls")) return (int64_t) h->cm.cholmod_gpu_trsm_calls ;
    if (!strcmp (name, "cpu_syrk_calls")) return (int64_t) h->cm.cholmod_cpu_syrk_calls ;
    if (!strcmp (name, "cpu_gemm_calls")) return (int64_t) h->cm.cholmod_cpu_gemm_calls ;
    if (!strcmp (name, "cpu_potrf_calls")) return (int64_t) h->cm.cholmod_cpu_potrf_calls ;
    if (!strcmp (name, "cpu_trsm_calls")) return (int64_t) h->cm.cholmod_cpu_trsm_calls ;
    if (!strcmp (name, "maxGpuMemBytes")) return (int64_t) h->cm.maxGpuMemBytes ;
    if (h->A && !strcmp (name, "n")) return (int64_t) h->A->nrow ;
    if (h->A && !strcmp (name, "ncol")) return (int64_t) h->A->ncol ;
    if (h->F && !strcmp (name, "nnzF")) return ((int64_t *) h->F->p) [h->F->ncol] ;
    if (h->A && !strcmp (name, "nnzA")) return ((int64_t *) h->A->p) [h->A->ncol] ;
    if (h->A && !strcmp (name, "stypeA")) return h->A->stype ;
    if (h->A && !strcmp (name, "xtypeA")) return h->A->xtype ;
    if (h->A && !strcmp (name, "dtypeA")) return h->A->dtype ;
    if (h->S && !strcmp (name, "nnzS")) return ((int64_t *) h->S->p) [h->S->ncol] ;
    if (!L) return -1 ;
    if (!strcmp (name, "is_super")) return L->is_super ;
    if (!strcmp (name, "nsuper")) return (int64_t) L->nsuper ;
    if (!strcmp (name, "ssize")) return (int64_t) L->ssize ;
    if (!strcmp (name, "xsize")) return (int64_t) L->xsize ;
    if (!strcmp (name, "maxcsize")) return (int64_t) L->maxcsize ;
    if (!strcmp (name, "maxesize")) return (int64_t) L->maxesize ;
    if (!strcmp (name, "minor")) return (int64_t) L->minor ;
    if (!strcmp (name, "ordering")) return L->ordering ;
    if (!strcmp (name, "xtype")) return L->xtype ;
    if (!strcmp (name, "dtype")) return L->dtype ;
    if (!strcmp (name, "is_ll")) return L->is_ll ;
    return -1 ;
}

double refd_get_double (void *hv, const char *name)
{
    refd *h = (refd *) hv ;
    if (!strcmp (name, "fl")) return h->cm.fl ;
    if (!strcmp (name, "lnz")) return h->cm.lnz ;
    if (!strcmp (name, "anz")) return h->cm.anz ;
    if (!strcmp (name, "t_analyze")) return h->t_analyze ;
    if (!strcmp (name, "t_factor")) return h->t_factor ;
    if (!strcmp (name, "t_solve")) return h->t_solve ;
    if (!strcmp (name, "gpu_syrk_time")) return h->cm.cholmod_gpu_syrk_time ;
    if (!strcmp (name, "gpu_gemm_time")) return h->cm.cholmod_gpu_gemm_time ;
    if (!strcmp (name, "gpu_potrf_time")) return h->cm.cholmod_gpu_potrf_time ;
    if (!strcmp (name, "gpu_trsm_time")) return h->cm.cholmod_gpu_trsm_time ;
    if (!strcmp (name, "maxGpuMemFraction")) return h->cm.maxGpuMemFraction ;
    return -1 ;
}

const void *refd_get_ptr (void *hv, const char *name)
{
    refd *h = (refd *) hv ;
    cholmod_factor *L = h->L ;
    if (h->A && !strcmp (name, "Ap")) return h->A->p ;
    if (h->A && !strcmp (name, "Ai")) return h->A->i ;
    if (h->A && !strcmp (name, "Ax")) return h->A->x ;
    if (h->A && !strcmp (name, "Az")) return h->A->z ;
    if (h->S && !strcmp (name, "Sp")) return h->S->p ;
    if (h->S && !strcmp (name, "Si")) return h->S->i ;
    if (h->S && !strcmp (name, "Sx")) return h->S->x ;
    if (h->S && !strcmp (name, "Sz")) return h->S->z ;
    if (h->F && !strcmp (name, "Fp")) return h->F->p ;
    if (h->F && !strcmp (name, "Fi")) return h->F->i ;
    if (h->F && !strcmp (name, "Fx")) return h->F->x ;
    if (h->F && !strcmp (name, "Fz")) return h->F->z ;
    if (!L) return NULL ;
    if (!strcmp (name, "Perm")) return L->Perm ;
    if (!strcmp (name, "ColCount")) return L->ColCount ;
    if (!strcmp (name, "super")) return L->super ;
    if (!strcmp (name, "pi")) return L->pi ;
    if (!strcmp (name, "px")) return L->px ;
    if (!strcmp (name, "s")) return L->s ;
    if (!strcmp (name, "Lx")) return L->x ;
    return NULL ;
}
