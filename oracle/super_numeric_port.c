// oracle/super_numeric_port.c -- TEST INFRASTRUCTURE (the oracle), not product code.
//
// A plain-C restatement of the reference's left-looking supernodal numeric Cholesky,
// CHOLMOD/Supernodal/t_cholmod_super_numeric_worker.c ("SNW" below), int64 indices, for
// double (sn_port_d), float (sn_port_s) and the Hermitian case in complex double (sn_port_z)
// and complex float (sn_port_c; entries are (re,im) pairs, the CHOLMOD_COMPLEX layout).  It takes the arrays of the
// supernodal symbolic factor (L->super, pi, px, s) and the seam matrix S = tril(P A P')
// (or, for the unsymmetric case, S = A(p,:) and F = S^H: then S*F is what gets factorized)
// as plain pointers; it does not include cholmod.h and calls no BLAS: the four dense
// kernels are written out as loops, so it is slow and meant for small cases.
//
// Parity status: PINNED.  tests/test_oracle.py checks this port against the reference
// itself (oracle/_ref/libcholmod_ref.so, built from /root/reference by oracle/Makefile)
// and against the golden fixtures in tests/golden/ that were generated from it.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use this file.

#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <complex.h>

#define SNP_OK 0
#define SNP_NOT_POSDEF 1
#define SNP_OUT_OF_MEMORY (-2)

#define SNP_TEMPLATE(NAME, REAL, SCALAR, CONJ, REPART)                                                                \
int NAME (int64_t n, int64_t nsuper, const int64_t *Super, const int64_t *Lpi,                    \
    const int64_t *Lpx, const int64_t *Ls, int64_t xsize,                                         \
    const int64_t *Ap, const int64_t *Ai, const int64_t *Anz, const SCALAR *Ax,                   \
    const int64_t *Fp, const int64_t *Fi, const int64_t *Fnz, const SCALAR *Fx,                   \
    double beta, int quick_return_if_not_posdef, SCALAR *Lx, int64_t *minor_out)                  \
{                                                                                                 \
    /* workspace, SNW:211-221 */                                                                  \
    int64_t *SuperMap = malloc ((size_t) (n+1) * sizeof (int64_t)) ;                              \
    int64_t *Map      = malloc ((size_t) (n+1) * sizeof (int64_t)) ;                              \
    int64_t *Head = malloc ((size_t) (nsuper+1) * sizeof (int64_t)) ;                             \
    int64_t *Next = malloc ((size_t) (nsuper+1) * sizeof (int64_t)) ;                             \
    int64_t *Lpos = malloc ((size_t) (nsuper+1) * sizeof (int64_t)) ;                             \
    int64_t *Next_save = malloc ((size_t) (nsuper+1) * sizeof (int64_t)) ;                        \
    int64_t *Lpos_save = malloc ((size_t) (nsuper+1) * sizeof (int64_t)) ;                        \
    int64_t maxc = 1 ;                                                                            \
    SCALAR *C = NULL ;                                                                            \
    int status = SNP_OK ;                                                                         \
    if (!SuperMap || !Map || !Head || !Next || !Lpos || !Next_save || !Lpos_save)                 \
    { status = SNP_OUT_OF_MEMORY ; goto done ; }                                                  \
    /* SuperMap [k] = supernode holding column k, cholmod_super_numeric.c:276-285 */              \
    for (int64_t s = 0 ; s < nsuper ; s++)                                                        \
    {                                                                                             \
        for (int64_t k = Super [s] ; k < Super [s+1] ; k++) SuperMap [k] = s ;                    \
        Head [s] = -1 ; Next [s] = -1 ; Lpos [s] = 0 ;                                            \
    }                                                                                             \
    for (int64_t i = 0 ; i < n ; i++) Map [i] = -1 ;                                              \
    *minor_out = n ;                                                                              \
    int repeat = 0 ; int64_t nscol_new = 0 ;                                                      \
    for (int64_t s = 0 ; s < nsuper ; s++)                                                        \
    {                                                                                             \
        int64_t k1 = Super [s], k2 = Super [s+1], nscol = k2 - k1 ;                               \
        int64_t psi = Lpi [s], psx = Lpx [s], nsrow = Lpi [s+1] - psi ;                           \
        /* zero the supernode, SNW:361-379 */                                                     \
        for (int64_t p = psx ; p < psx + nsrow*nscol ; p++) Lx [p] = 0 ;                          \
        /* Map: global row -> local row of s, SNW:388-399 */                                      \
        for (int64_t k = 0 ; k < nsrow ; k++) Map [Ls [psi+k]] = k ;                              \
        /* unsymmetric case (Fp != NULL): columns k1..k2-1 of A*F, SNW:471-506 */                 \
        for (int64_t k = k1 ; Fp && k < k2 ; k++)                                                 \
        {                                                                                         \
            int64_t pf = Fp [k], pfend = Fnz ? pf + Fnz [k] : Fp [k+1] ;                          \
            for ( ; pf < pfend ; pf++)                                                            \
            {                                                                                     \
                int64_t j = Fi [pf] ;                                                             \
                SCALAR fjk = Fx [pf] ;                                                            \
                int64_t p = Ap [j], pend = Anz ? p + Anz [j] : Ap [j+1] ;                         \
                for ( ; p < pend ; p++)                                                           \
                {                                                                                 \
                    int64_t i = Ai [p] ;                                                          \
                    if (i >= k)                                                                   \
                    {                                                                             \
                        int64_t im = Map [i] ;                                                    \
                        if (im >= 0 && im < nsrow) Lx [psx + im + (k-k1)*nsrow] += Ax [p] * fjk ; \
                    }                                                                             \
                }                                                                                 \
            }                                                                                     \
        }                                                                                         \
        /* scatter the lower part of columns k1..k2-1 of A, SNW:437-468 */                        \
        for (int64_t k = k1 ; !Fp && k < k2 ; k++)                                                \
        {                                                                                         \
            int64_t p = Ap [k], pend = Anz ? p + Anz [k] : Ap [k+1] ;                             \
            for ( ; p < pend ; p++)                                                               \
            {                                                                                     \
                int64_t i = Ai [p] ;                                                              \
                if (i >= k)                                                                       \
                {                                                                                 \
                    int64_t im = Map [i] ;                                                        \
                    if (im >= 0 && im < nsrow) Lx [psx + im + (k-k1)*nsrow] = Ax [p] ;            \
                }                                                                                 \
            }                                                                                     \
        }                                                                                         \
        /* beta on the diagonal, SNW:510-520 */                                                   \
        if (beta != 0)                                                                            \
            for (int64_t k = 0 ; k < nscol ; k++) Lx [psx + k*(nsrow+1)] += (REAL) beta ;         \
        /* save / restore the descendant list, SNW:531-551 */                                     \
        for (int64_t d = Head [s] ; d != -1 ; d = Next [d])                                       \
        {                                                                                         \
            if (!repeat) { Lpos_save [d] = Lpos [d] ; Next_save [d] = Next [d] ; }                \
            else         { Lpos [d] = Lpos_save [d] ; Next [d] = Next_save [d] ; }                \
        }                                                                                         \
        /* descendant updates, SNW:588-969 */                                                     \
        int64_t dnext = Head [s] ;                                                                \
        while (dnext != -1)                                                                       \
        {                                                                                         \
            int64_t d = dnext ;                                                                   \
            int64_t ndcol = Super [d+1] - Super [d] ;                                             \
            int64_t pdi = Lpi [d], pdx = Lpx [d], pdend = Lpi [d+1], ndrow = pdend - pdi ;        \
            int64_t pdi1 = pdi + Lpos [d], pdx1 = pdx + Lpos [d], pdi2 ;                          \
            for (pdi2 = pdi1 ; pdi2 < pdend && Ls [pdi2] < k2 ; pdi2++) ;   /* SNW:686 */         \
            int64_t ndrow1 = pdi2 - pdi1, ndrow2 = pdend - pdi1 ;                                 \
            if (ndrow1*ndrow2 > maxc || !C)                                                       \
            {                                                                                     \
                if (ndrow1*ndrow2 > maxc) maxc = ndrow1*ndrow2 ;                                  \
                free (C) ; C = malloc ((size_t) maxc * sizeof (SCALAR)) ;                         \
                if (!C) { status = SNP_OUT_OF_MEMORY ; goto done ; }                              \
            }                                                                                     \
            /* C = [L1;L2] * L1', lower part of the top block only: the dsyrk of SNW:769 */       \
            /* and the dgemm of SNW:824 written as one loop nest */                               \
            for (int64_t j = 0 ; j < ndrow1 ; j++)                                                \
                for (int64_t i = j ; i < ndrow2 ; i++)                                            \
                {                                                                                 \
                    SCALAR acc = 0 ;                                                              \
                    for (int64_t k = 0 ; k < ndcol ; k++)                                         \
                        acc += Lx [pdx1 + i + k*ndrow] * CONJ (Lx [pdx1 + j + k*ndrow]) ;         \
                    C [i + j*ndrow2] = acc ;                                                      \
                }                                                                                 \
            /* relative map + assembly, SNW:897-930 */                                            \
            for (int64_t j = 0 ; j < ndrow1 ; j++)                                                \
            {                                                                                     \
                int64_t px = psx + Map [Ls [pdi1+j]] * nsrow ;                                    \
                for (int64_t i = j ; i < ndrow2 ; i++)                                            \
                    Lx [px + Map [Ls [pdi1+i]]] -= C [i + j*ndrow2] ;                             \
            }                                                                                     \
            /* hand d to its next ancestor, SNW:950-967 */                                        \
            dnext = Next [d] ;                                                                    \
            if (!repeat)                                                                          \
            {                                                                                     \
                Lpos [d] = pdi2 - pdi ;                                                           \
                if (Lpos [d] < ndrow)                                                             \
                {                                                                                 \
                    int64_t anc = SuperMap [Ls [pdi2]] ;                                          \
                    Next [d] = Head [anc] ; Head [anc] = d ;                                      \
                }                                                                                 \
            }                                                                                     \
        }                                                                                         \
        /* dense Cholesky of the diagonal block (the dpotrf of SNW:1023), lda = nsrow */          \
        int64_t nscol2 = repeat ? nscol_new : nscol ;                                             \
        int64_t info = 0 ;                                                                        \
        for (int64_t j = 0 ; j < nscol2 && !info ; j++)                                           \
        {                                                                                         \
            REAL ajj = REPART (Lx [psx + j + j*nsrow]) ;      /* zpotrf: the real part only */    \
            for (int64_t k = 0 ; k < j ; k++)                                                     \
                ajj -= REPART (Lx [psx + j + k*nsrow] * CONJ (Lx [psx + j + k*nsrow])) ;          \
            if (!(ajj > 0)) { Lx [psx + j + j*nsrow] = ajj ; info = j+1 ; break ; }               \
            ajj = (REAL) sqrt ((double) ajj) ;                                                    \
            Lx [psx + j + j*nsrow] = ajj ;                                                        \
            for (int64_t i = j+1 ; i < nscol2 ; i++)                                              \
            {                                                                                     \
                SCALAR v = Lx [psx + i + j*nsrow] ;                                               \
                for (int64_t k = 0 ; k < j ; k++)                                                 \
                    v -= Lx [psx + i + k*nsrow] * CONJ (Lx [psx + j + k*nsrow]) ;                 \
                Lx [psx + i + j*nsrow] = v / ajj ;                                                \
            }                                                                                     \
        }                                                                                         \
        if (repeat)                                                                               \
        {                                                                                         \
            info = 0 ;      /* SNW:1064-1077 */                                                   \
            for (int64_t p = psx + nsrow*nscol_new ; p < psx + nsrow*nscol ; p++) Lx [p] = 0 ;    \
        }                                                                                         \
        if (info != 0)                                                                            \
        {                                                                                         \
            /* not positive definite, SNW:1084-1146 */                                            \
            status = SNP_NOT_POSDEF ;                                                             \
            *minor_out = k1 + info - 1 ;                                                          \
            for (int64_t ss = s+1 ; ss < nsuper ; ss++) Head [ss] = -1 ;                          \
            for (int64_t p = psx ; p < xsize ; p++) Lx [p] = 0 ;                                  \
            if (info == 1 || quick_return_if_not_posdef) { Head [s] = -1 ; goto done ; }          \
            repeat = 1 ; nscol_new = info - 1 ; s-- ;                                             \
            continue ;                                                                            \
        }                                                                                         \
        /* L2 = S2 / L1' (the dtrsm of SNW:1175) on the rows below the factorized block */        \
        int64_t nsrow2 = nsrow - nscol2 ;                                                         \
        if (nsrow2 > 0)                                                                           \
        {                                                                                         \
            for (int64_t j = 0 ; j < nscol2 ; j++)                                                \
            {                                                                                     \
                REAL ljj = REPART (Lx [psx + j + j*nsrow]) ;                                      \
                for (int64_t i = nscol2 ; i < nsrow ; i++)                                        \
                {                                                                                 \
                    SCALAR v = Lx [psx + i + j*nsrow] ;                                           \
                    for (int64_t k = 0 ; k < j ; k++)                                             \
                        v -= Lx [psx + i + k*nsrow] * CONJ (Lx [psx + j + k*nsrow]) ;             \
                    Lx [psx + i + j*nsrow] = v / ljj ;                                            \
                }                                                                                 \
            }                                                                                     \
            if (!repeat)                                                                          \
            {                                                                                     \
                /* link s into its parent's list, SNW:1222-1235 */                                \
                Lpos [s] = nscol ;                                                                \
                int64_t sparent = SuperMap [Ls [psi + nscol]] ;                                   \
                Next [s] = Head [sparent] ; Head [sparent] = s ;                                  \
            }                                                                                     \
        }                                                                                         \
        Head [s] = -1 ;                                                                           \
        if (repeat) goto done ;     /* SNW:1253-1265 */                                           \
    }                                                                                             \
done:                                                                                             \
    free (C) ; free (SuperMap) ; free (Map) ; free (Head) ; free (Next) ; free (Lpos) ;           \
    free (Next_save) ; free (Lpos_save) ;                                                         \
    return status ;                                                                               \
}

#define SNP_ID(x) (x)
SNP_TEMPLATE (sn_port_d, double, double, SNP_ID, SNP_ID)
SNP_TEMPLATE (sn_port_s, float, float, SNP_ID, SNP_ID)
/* the COMPLEX variants: zherk/zgemm "C" (SNW:786-803, :850-876), zpotrf (:1030-1048), ztrsm "C" (:1184-1208) */
SNP_TEMPLATE (sn_port_z, double, double _Complex, conj, creal)
SNP_TEMPLATE (sn_port_c, float, float _Complex, conjf, crealf)

// Supernodal forward/back solve L L' x = b for one right-hand side, double, following
// CHOLMOD/Supernodal/t_cholmod_super_solve_worker.c:20-303 (lsolve) and :309-586 (ltsolve),
// dense kernels written as loops.  x is overwritten (b on input), in the permuted ordering.
int sn_port_solve_d (int64_t nsuper, const int64_t *Super, const int64_t *Lpi, const int64_t *Lpx,
    const int64_t *Ls, const double *Lx, double *x)
{
    for (int64_t s = 0 ; s < nsuper ; s++)
    {
        int64_t k1 = Super [s], nscol = Super [s+1] - k1, psi = Lpi [s], psx = Lpx [s] ;
        int64_t nsrow = Lpi [s+1] - psi ;
        for (int64_t j = 0 ; j < nscol ; j++)
        {
            double xj = x [k1+j] / Lx [psx + j + j*nsrow] ;
            x [k1+j] = xj ;
            for (int64_t i = j+1 ; i < nsrow ; i++) x [Ls [psi+i]] -= Lx [psx + i + j*nsrow] * xj ;
        }
    }
    for (int64_t s = nsuper-1 ; s >= 0 ; s--)
    {
        int64_t k1 = Super [s], nscol = Super [s+1] - k1, psi = Lpi [s], psx = Lpx [s] ;
        int64_t nsrow = Lpi [s+1] - psi ;
        for (int64_t j = nscol-1 ; j >= 0 ; j--)
        {
            double v = x [k1+j] ;
            for (int64_t i = j+1 ; i < nsrow ; i++) v -= Lx [psx + i + j*nsrow] * x [Ls [psi+i]] ;
            x [k1+j] = v / Lx [psx + j + j*nsrow] ;
        }
    }
    return 0 ;
}

// The same for a complex factor: L L^H x = b (the "C" variants of the same file).
int sn_port_solve_z (int64_t nsuper, const int64_t *Super, const int64_t *Lpi, const int64_t *Lpx,
    const int64_t *Ls, const double _Complex *Lx, double _Complex *x)
{
    for (int64_t s = 0 ; s < nsuper ; s++)
    {
        int64_t k1 = Super [s], nscol = Super [s+1] - k1, psi = Lpi [s], psx = Lpx [s] ;
        int64_t nsrow = Lpi [s+1] - psi ;
        for (int64_t j = 0 ; j < nscol ; j++)
        {
            double _Complex xj = x [k1+j] / creal (Lx [psx + j + j*nsrow]) ;
            x [k1+j] = xj ;
            for (int64_t i = j+1 ; i < nsrow ; i++) x [Ls [psi+i]] -= Lx [psx + i + j*nsrow] * xj ;
        }
    }
    for (int64_t s = nsuper-1 ; s >= 0 ; s--)
    {
        int64_t k1 = Super [s], nscol = Super [s+1] - k1, psi = Lpi [s], psx = Lpx [s] ;
        int64_t nsrow = Lpi [s+1] - psi ;
        for (int64_t j = nscol-1 ; j >= 0 ; j--)
        {
            double _Complex v = x [k1+j] ;
            for (int64_t i = j+1 ; i < nsrow ; i++) v -= conj (Lx [psx + i + j*nsrow]) * x [Ls [psi+i]] ;
            x [k1+j] = v / creal (Lx [psx + j + j*nsrow]) ;
        }
    }
    return 0 ;
}
