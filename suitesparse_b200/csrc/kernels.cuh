// kernels.cuh -- sm_100a kernels of the CHOLMOD/B200 numeric phase.
//
//   k_scatter_A   zero-initialised supernodes <- A, + beta on the diagonal          (SNW:361-520)
//   k_upd_simt    small descendant updates, one warp per 32x8 tile                  (SNW:769-930)
//   k_upd_dmma    descendant / trailing updates on the FP64 tensor pipe (DMMA),
//                 cp.async-staged shared-memory tiles, fused relative-map scatter    (SNW:769-930)
//   k_potrf       dense Cholesky (+ inverse) of <=64-wide diagonal blocks, row per thread (SNW:1023)
//   k_trsm        X <- X * L11^{-T} by substitution, one thread per row                (SNW:1175)
//   k_trsm_inv    X <- X * W', W = L11^{-1}, on the FP64 tensor pipe                   (SNW:1175)
//   k_solve_*     supernodal forward / backward solve                               (SSOL:20-586)
//
// "SNW" = CHOLMOD/Supernodal/t_cholmod_super_numeric_worker.c, "SSOL" = t_cholmod_super_solve_worker.c.
// FP64 has no tcgen05 kind: the FP64 tensor path on sm_100a is mma.sync (SASS DMMA), so operands are
// staged global -> shared with cp.async (panels have arbitrary 8-byte-aligned leading dimensions, which
// rules out TMA tensor maps) and fed to mma.sync.m8n8k4 from registers.
#pragma once
#include "plan.hpp"
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

// ---------------------------------------------------------------------------------------------
// scalars: real (float, double) and complex (re,im interleaved, the layout of CHOLMOD_COMPLEX)
// ---------------------------------------------------------------------------------------------

// The kernels below are written once over T.  For a real T every helper is the identity / the plain
// operator, so the real instantiations compile to what they were before the complex types existed; for
// cx<R> they spell out the Hermitian case of the reference (zherk/zgemm "C", zpotrf, ztrsm "C":
// SNW:786-803, :850-876, :1030-1048, :1184-1208): A = L*L^H, the diagonal of L real and positive.
template <typename R> struct __align__ (2 * sizeof (R)) cx { R re, im ; } ;
typedef cx<double> zdouble ;
typedef cx<float> cfloat ;

template <typename T> struct num { typedef T real ; static constexpr bool is_complex = false ; } ;
template <typename R> struct num<cx<R>> { typedef R real ; static constexpr bool is_complex = true ; } ;

template <typename R> __host__ __device__ __forceinline__ cx<R> operator+ (cx<R> a, cx<R> b) { return { a.re + b.re, a.im + b.im } ; }
template <typename R> __host__ __device__ __forceinline__ cx<R> operator- (cx<R> a, cx<R> b) { return { a.re - b.re, a.im - b.im } ; }
template <typename R> __host__ __device__ __forceinline__ cx<R> operator- (cx<R> a) { return { -a.re, -a.im } ; }
template <typename R> __host__ __device__ __forceinline__ cx<R> operator* (cx<R> a, cx<R> b)
{
    return { a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re } ;
}
template <typename R> __host__ __device__ __forceinline__ cx<R> operator* (cx<R> a, R b) { return { a.re * b, a.im * b } ; }
template <typename R> __host__ __device__ __forceinline__ cx<R> operator/ (cx<R> a, R b) { return { a.re / b, a.im / b } ; }
template <typename R> __host__ __device__ __forceinline__ cx<R> &operator+= (cx<R> &a, cx<R> b) { a.re += b.re ; a.im += b.im ; return a ; }
template <typename R> __host__ __device__ __forceinline__ cx<R> &operator-= (cx<R> &a, cx<R> b) { a.re -= b.re ; a.im -= b.im ; return a ; }

__host__ __device__ __forceinline__ float conj_ (float a) { return a ; }
__host__ __device__ __forceinline__ double conj_ (double a) { return a ; }
template <typename R> __host__ __device__ __forceinline__ cx<R> conj_ (cx<R> a) { return { a.re, -a.im } ; }
__host__ __device__ __forceinline__ float real_ (float a) { return a ; }
__host__ __device__ __forceinline__ double real_ (double a) { return a ; }
template <typename R> __host__ __device__ __forceinline__ R real_ (cx<R> a) { return a.re ; }

// a * conj (b)
__host__ __device__ __forceinline__ float mulc (float a, float b) { return a * b ; }
__host__ __device__ __forceinline__ double mulc (double a, double b) { return a * b ; }
template <typename R> __host__ __device__ __forceinline__ cx<R> mulc (cx<R> a, cx<R> b)
{
    return { a.re * b.re + a.im * b.im, a.im * b.re - a.re * b.im } ;
}

// the scalar r (a real number) as a T
template <typename T> struct from_real_t { static __host__ __device__ __forceinline__ T make (T r) { return r ; } } ;
template <typename R> struct from_real_t<cx<R>> { static __host__ __device__ __forceinline__ cx<R> make (R r) { return { r, (R) 0 } ; } } ;
template <typename T> __host__ __device__ __forceinline__ T from_real (typename num<T>::real r) { return from_real_t<T>::make (r) ; }

// conversion between precisions (the solve works in double whatever the factor's precision is)
template <typename X, typename T> struct cvt_t { static __device__ __forceinline__ X make (T a) { return (X) a ; } } ;
template <typename RX, typename R> struct cvt_t<cx<RX>, cx<R>> { static __device__ __forceinline__ cx<RX> make (cx<R> a) { return { (RX) a.re, (RX) a.im } ; } } ;
template <typename X, typename T> __device__ __forceinline__ X cvt (T a) { return cvt_t<X, T>::make (a) ; }

// *p -= v, atomically (per component for a complex entry: the components are independent sums)
__device__ __forceinline__ void atomic_sub (float *p, float v) { atomicAdd (p, -v) ; }
__device__ __forceinline__ void atomic_sub (double *p, double v) { atomicAdd (p, -v) ; }
template <typename R> __device__ __forceinline__ void atomic_sub (cx<R> *p, cx<R> v)
{
    atomicAdd (&p->re, -v.re) ;
    atomicAdd (&p->im, -v.im) ;
}

__device__ __forceinline__ float shfl_ (unsigned mask, float v, int src) { return __shfl_sync (mask, v, src) ; }
__device__ __forceinline__ double shfl_ (unsigned mask, double v, int src) { return __shfl_sync (mask, v, src) ; }
template <typename R> __device__ __forceinline__ cx<R> shfl_ (unsigned mask, cx<R> v, int src)
{
    return { __shfl_sync (mask, v.re, src), __shfl_sync (mask, v.im, src) } ;
}
__device__ __forceinline__ float shfl_xor_ (unsigned mask, float v, int o) { return __shfl_xor_sync (mask, v, o) ; }
__device__ __forceinline__ double shfl_xor_ (unsigned mask, double v, int o) { return __shfl_xor_sync (mask, v, o) ; }
template <typename R> __device__ __forceinline__ cx<R> shfl_xor_ (unsigned mask, cx<R> v, int o)
{
    return { __shfl_xor_sync (mask, v.re, o), __shfl_xor_sync (mask, v.im, o) } ;
}

// 1 / d for a pivot.  An IEEE double division is a long dependent instruction sequence, and the panel chain
// has one on its critical path per column: the hardware seed (relative error 2^-23) and two Newton steps give
// the reciprocal to within an ulp or two in a fraction of the latency.  d <= 0 or NaN (matrix not positive
// definite) gives garbage, as a division would: the column is reported through `minor` either way.
__device__ __forceinline__ float rcp_ (float d) { return 1.0f / d ; }
__device__ __forceinline__ double rcp_ (double d)
{
    double r ;
    asm ("rcp.approx.ftz.f64 %0, %1;" : "=d" (r) : "d" (d)) ;
    r = fma (fma (-d, r, 1.0), r, r) ;
    r = fma (fma (-d, r, 1.0), r, r) ;
    return r ;
}

// ---------------------------------------------------------------------------------------------
// scatter A
// ---------------------------------------------------------------------------------------------

// One thread per column k of A.  Lx must already be zero.  Row i >= k of column k goes to local row
// Map [i] of supernode s = SuperMap [k]; the reference's Map array is replaced by: i - k1 inside the
// diagonal block, a binary search of the (ascending) off-diagonal row list otherwise.  Entries above the
// diagonal or outside the pattern of L are dropped (SNW:455-468); duplicates: the last one wins, as in
// the reference's assignment.
template <typename T>
__global__ void __launch_bounds__ (256) k_scatter_A (int64_t kfirst, int64_t klast,
    const int64_t *__restrict__ Ap, const int64_t *__restrict__ Ai, const int64_t *__restrict__ Anz,
    const T *__restrict__ Ax, const int32_t *__restrict__ supermap, const int64_t *__restrict__ super,
    const int64_t *__restrict__ pi, const int64_t *__restrict__ px, const int32_t *__restrict__ Ls,
    typename num<T>::real beta, T *__restrict__ Lx, const unsigned char *__restrict__ colmine)
{
    int64_t k = kfirst + (int64_t) blockIdx.x * blockDim.x + threadIdx.x ;
    if (k >= klast) return ;
    if (colmine && !colmine [k]) return ;       // distributed: another rank holds (the real copy of) this column
    int32_t s = supermap [k] ;
    int64_t k1 = super [s], k2 = super [s+1], psi = pi [s] ;
    int64_t nsrow = pi [s+1] - psi, nscol = k2 - k1 ;
    T *col = Lx + px [s] + (k - k1) * nsrow ;
    int64_t p = Ap [k], pend = Anz ? p + Anz [k] : Ap [k+1] ;
    for ( ; p < pend ; p++)
    {
        int64_t i = Ai [p] ;
        if (i < k) continue ;
        int64_t pos ;
        if (i < k2)
        {
            pos = i - k1 ;
        }
        else
        {
            int64_t lo = nscol, hi = nsrow ;        // first position with Ls >= i
            while (lo < hi)
            {
                int64_t mid = (lo + hi) >> 1 ;
                if ((int64_t) Ls [psi + mid] < i) lo = mid + 1 ; else hi = mid ;
            }
            if (lo >= nsrow || (int64_t) Ls [psi + lo] != i) continue ;
            pos = lo ;
        }
        col [pos] = Ax [p] ;
    }
    // only the real part of beta exists (SNW:510-520)
    if (beta != (typename num<T>::real) 0) col [k - k1] += from_real<T> (beta) ;
}

// The unsymmetric case (A->stype == 0, SNW:471-506): the matrix to factorize is A*F with F = A(:,f)^H given
// explicitly (both already row-permuted by the caller, cholmod_factorize.c:236-247).  One thread per column k
// of L forms column k of A*F restricted to the pattern of L:
//     L (i,k) += A (i,j) * F (j,k)   for every j in the pattern of F (:,k), every i >= k in A (:,j)
// The thread owns its column, so the sums need no atomics.  A is nrow-by-ncol, F ncol-by-nrow.
template <typename T>
__global__ void __launch_bounds__ (256) k_scatter_AF (int64_t kfirst, int64_t klast,
    const int64_t *__restrict__ Ap, const int64_t *__restrict__ Ai, const int64_t *__restrict__ Anz,
    const T *__restrict__ Ax, const int64_t *__restrict__ Fp, const int64_t *__restrict__ Fi,
    const int64_t *__restrict__ Fnz, const T *__restrict__ Fx, const int32_t *__restrict__ supermap,
    const int64_t *__restrict__ super, const int64_t *__restrict__ pi, const int64_t *__restrict__ px,
    const int32_t *__restrict__ Ls, typename num<T>::real beta, T *__restrict__ Lx,
    const unsigned char *__restrict__ colmine)
{
    int64_t k = kfirst + (int64_t) blockIdx.x * blockDim.x + threadIdx.x ;
    if (k >= klast) return ;
    if (colmine && !colmine [k]) return ;
    int32_t s = supermap [k] ;
    int64_t k1 = super [s], k2 = super [s+1], psi = pi [s] ;
    int64_t nsrow = pi [s+1] - psi, nscol = k2 - k1 ;
    T *col = Lx + px [s] + (k - k1) * nsrow ;
    int64_t pf = Fp [k], pfend = Fnz ? pf + Fnz [k] : Fp [k+1] ;
    for ( ; pf < pfend ; pf++)
    {
        const int64_t j = Fi [pf] ;
        const T fjk = Fx [pf] ;
        int64_t p = Ap [j], pend = Anz ? p + Anz [j] : Ap [j+1] ;
        for ( ; p < pend ; p++)
        {
            int64_t i = Ai [p] ;
            if (i < k) continue ;
            int64_t pos ;
            if (i < k2)
            {
                pos = i - k1 ;
            }
            else
            {
                int64_t lo = nscol, hi = nsrow ;
                while (lo < hi)
                {
                    int64_t mid = (lo + hi) >> 1 ;
                    if ((int64_t) Ls [psi + mid] < i) lo = mid + 1 ; else hi = mid ;
                }
                if (lo >= nsrow || (int64_t) Ls [psi + lo] != i) continue ;
                pos = lo ;
            }
            col [pos] += Ax [p] * fjk ;
        }
    }
    if (beta != (typename num<T>::real) 0) col [k - k1] += from_real<T> (beta) ;
}

// ---------------------------------------------------------------------------------------------
// small updates: one warp per 32 x 8 tile, operands straight from L1/L2
// ---------------------------------------------------------------------------------------------

template <typename T>
__global__ void __launch_bounds__ (256) k_upd_simt (const UpdTile *__restrict__ tiles, int ntiles,
    const UpdItem *__restrict__ items, const int32_t *__restrict__ relmap, T *Lx)
{
    int w = (int) (((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5) ;
    int lane = threadIdx.x & 31 ;
    if (w >= ntiles) return ;
    UpdTile t = tiles [w] ;
    const UpdItem it = items [t.item] ;
    const int i = (int) t.ti * SIMT_TM + lane ;
    const int j0 = (int) t.tj * SIMT_TN ;
    const int nj = min (SIMT_TN, it.N - j0) ;
    const bool rowok = i < it.M ;
    const T *A = Lx + it.a_off ;
    const int64_t lda = it.lda ;
    const T zero = from_real<T> (0) ;
    T acc [SIMT_TN] ;
    #pragma unroll
    for (int jj = 0 ; jj < SIMT_TN ; jj++) acc [jj] = zero ;
    for (int k = t.k0 ; k < t.k1 ; k++)
    {
        const T *Ak = A + (int64_t) k * lda ;
        T a = rowok ? Ak [i] : zero ;
        #pragma unroll
        for (int jj = 0 ; jj < SIMT_TN ; jj++)
        {
            T b = (jj < nj) ? Ak [j0 + jj - it.shift] : zero ;
            acc [jj] += mulc (a, b) ;
        }
    }
    if (!rowok) return ;
    const bool mapped = it.map_off >= 0 ;
    const int32_t *map = relmap + (mapped ? it.map_off : 0) ;
    const int64_t ri = mapped ? (int64_t) map [i] : (int64_t) i ;
    T *C = Lx + it.c_off + ri ;
    #pragma unroll
    for (int jj = 0 ; jj < SIMT_TN ; jj++)
    {
        int j = j0 + jj ;
        if (jj < nj && i + it.shift >= j)
        {
            int64_t cj = mapped ? (int64_t) map [j] : (int64_t) j ;
            atomic_sub (C + cj * it.ldc, acc [jj]) ;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// tiny updates and tiny supernodes: one THREAD per object
// ---------------------------------------------------------------------------------------------

// The bulk of the update pairs near the leaves (K = 1..TINY_K descendant columns, a few dozen target entries):
// thread t does item tiles [t].item whole.  No shared memory, no idle lanes; the panel entries a thread re-reads
// for every target column come from L1.
template <typename T>
__global__ void __launch_bounds__ (128) k_upd_tiny (const UpdTile *__restrict__ tiles, int ntiles,
    const UpdItem *__restrict__ items, const int32_t *__restrict__ relmap, T *Lx)
{
    const int t = (int) (blockIdx.x * blockDim.x + threadIdx.x) ;
    if (t >= ntiles) return ;
    const UpdItem it = items [tiles [t].item] ;
    const T *A = Lx + it.a_off ;
    const int64_t lda = it.lda, ldc = it.ldc ;
    const bool mapped = it.map_off >= 0 ;
    const int32_t *map = relmap + (mapped ? it.map_off : 0) ;
    const T zero = from_real<T> (0) ;
    for (int j = 0 ; j < it.N ; j++)
    {
        T b [TINY_K] ;
        #pragma unroll
        for (int k = 0 ; k < TINY_K ; k++) b [k] = (k < it.K) ? A [j + k * lda] : zero ;
        T *Cj = Lx + it.c_off + (mapped ? (int64_t) map [j] : (int64_t) j) * ldc ;
        for (int i = j ; i < it.M ; i++)
        {
            T acc = zero ;
            #pragma unroll
            for (int k = 0 ; k < TINY_K ; k++) if (k < it.K) acc += mulc (A [i + k * lda], b [k]) ;
            atomic_sub (Cj + (mapped ? (int64_t) map [i] : (int64_t) i), acc) ;
        }
    }
}

// Whole supernodes of at most TINY_COLS columns with at most TINY_ROWS rows below the diagonal block (most of the
// leaves of the etree): thread t factorizes block t -- dpotrf of a <= 4 x 4 block in registers, then the rows
// below it one after the other by substitution (SNW:1023, :1175).  Same failure rule as the other potrf kernels.
template <typename T>
__global__ void __launch_bounds__ (128) k_factor_tiny (const PotrfBlk *__restrict__ blks, int nblks, T *Lx,
    unsigned long long *minor)
{
    const int t = (int) (blockIdx.x * blockDim.x + threadIdx.x) ;
    if (t >= nblks) return ;
    const PotrfBlk b = blks [t] ;
    const int jb = b.jb ;
    const int64_t lda = b.lda ;
    T *D = Lx + b.off ;
    typedef typename num<T>::real R ;
    const T one = from_real<T> (1), zero = from_real<T> (0) ;
    T a [TINY_COLS][TINY_COLS] ;        // a [r][c], lower part; identity beyond jb
    #pragma unroll
    for (int c = 0 ; c < TINY_COLS ; c++)
        #pragma unroll
        for (int r = 0 ; r < TINY_COLS ; r++)
            a [r][c] = (r < jb && c <= r) ? D [r + c * lda] : ((r == c) ? one : zero) ;
    R rdiag [TINY_COLS] ;
    #pragma unroll
    for (int j = 0 ; j < TINY_COLS ; j++)
    {
        const R d = real_ (a [j][j]) ;
        if (j < jb && !(d > (R) 0)) atomicMin (minor, (unsigned long long) (b.col + j)) ;
        const R ljj = sqrt (d), rl = (R) 1 / ljj ;
        rdiag [j] = rl ;
        a [j][j] = from_real<T> (ljj) ;
        #pragma unroll
        for (int r = j + 1 ; r < TINY_COLS ; r++) a [r][j] = a [r][j] * rl ;
        #pragma unroll
        for (int c = j + 1 ; c < TINY_COLS ; c++)
            #pragma unroll
            for (int r = c ; r < TINY_COLS ; r++) a [r][c] -= mulc (a [r][j], a [c][j]) ;
    }
    #pragma unroll
    for (int c = 0 ; c < TINY_COLS ; c++)
        #pragma unroll
        for (int r = 0 ; r < TINY_COLS ; r++)
            if (r < jb && c <= r) D [r + c * lda] = a [r][c] ;
    // rows below: x <- x * L11^{-H}
    for (int64_t row = jb ; row < lda ; row++)
    {
        T x [TINY_COLS] ;
        #pragma unroll
        for (int j = 0 ; j < TINY_COLS ; j++) x [j] = (j < jb) ? D [row + j * lda] : zero ;
        #pragma unroll
        for (int j = 0 ; j < TINY_COLS ; j++)
        {
            const T xj = x [j] * rdiag [j] ;
            x [j] = xj ;
            #pragma unroll
            for (int k = j + 1 ; k < TINY_COLS ; k++) x [k] -= mulc (xj, a [k][j]) ;
        }
        #pragma unroll
        for (int j = 0 ; j < TINY_COLS ; j++) if (j < jb) D [row + j * lda] = x [j] ;
    }
}

// ---------------------------------------------------------------------------------------------
// DMMA update kernel
// ---------------------------------------------------------------------------------------------

__device__ __forceinline__ void cp_async8 (void *smem_dst, const void *gsrc, bool valid)
{
    unsigned d = (unsigned) __cvta_generic_to_shared (smem_dst) ;
    int sz = valid ? 8 : 0 ;
    asm volatile ("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" :: "r" (d), "l" (gsrc), "r" (sz)) ;
}
__device__ __forceinline__ void cp_async_commit () { asm volatile ("cp.async.commit_group;\n" ::) ; }
template <int N> __device__ __forceinline__ void cp_async_wait () { asm volatile ("cp.async.wait_group %0;\n" :: "n" (N)) ; }

__device__ __forceinline__ void dmma884 (double &d0, double &d1, double a, double b)
{
    asm volatile ("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
        : "+d" (d0), "+d" (d1) : "d" (a), "d" (b)) ;
}

constexpr int DM_KC = 16 ;          // K depth of one pipeline stage
constexpr int DM_STAGES = 3 ;

template <int TM, int TN> constexpr size_t dmma_smem_bytes ()
{
    return (size_t) DM_STAGES * DM_KC * ((TM + 4) + (TN + 4)) * sizeof (double) + (TM + TN) * sizeof (int32_t) ;
}

// C tile (TM x TN) of one item, inner range [k0,k1): CTA of (TM/WM)*(TN/WN) warps, each warp owns a
// WM x WN sub-tile as (WM/8) x (WN/8) m8n8k4 accumulators.  Shared tiles are stored [k][m] with the row
// pitch padded to TM+4 doubles, so that the 16 lanes of a half-warp (k = lane%4, m = lane/4) hit 16
// distinct 8-byte banks.  Epilogue: red.global.add.f64 of -C through the relative map.
template <int TM, int TN, int WM, int WN>
__global__ void __launch_bounds__ ((TM / WM) * (TN / WN) * 32) k_upd_dmma (const UpdTile *__restrict__ tiles,
    const UpdItem *__restrict__ items, const int32_t *__restrict__ relmap, double *Lx)
{
    constexpr int NW_M = TM / WM, NW_N = TN / WN ;
    constexpr int NT = NW_M * NW_N * 32 ;
    constexpr int LDA_S = TM + 4, LDB_S = TN + 4 ;
    constexpr int MT = WM / 8, NTL = WN / 8 ;
    extern __shared__ __align__ (16) unsigned char smem_raw [] ;
    double *sA = reinterpret_cast<double *> (smem_raw) ;                 // [STAGES][KC][LDA_S]
    double *sB = sA + DM_STAGES * DM_KC * LDA_S ;                        // [STAGES][KC][LDB_S]
    int32_t *smap_r = reinterpret_cast<int32_t *> (sB + DM_STAGES * DM_KC * LDB_S) ;    // [TM]
    int32_t *smap_c = smap_r + TM ;                                      // [TN]

    const UpdTile t = tiles [blockIdx.x] ;
    const UpdItem it = items [t.item] ;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5 ;
    const int wm = warp % NW_M, wn = warp / NW_M ;
    const int i0 = (int) t.ti * TM, j0 = (int) t.tj * TN ;
    const int mA = min (TM, it.M - i0), nB = min (TN, it.N - j0) ;
    const int64_t lda = it.lda ;
    const double *A = Lx + it.a_off ;
    const double *Arow = A + i0, *Brow = A - it.shift + j0 ;     // B (0,0) lies `shift` rows above A (0,0)
    const int kbeg = t.k0, kend = t.k1 ;
    const int nk = (kend - kbeg + DM_KC - 1) / DM_KC ;

    // relative maps of this tile's rows and columns
    {
        const bool mapped = it.map_off >= 0 ;
        const int32_t *map = relmap + (mapped ? it.map_off : 0) ;
        for (int m = tid ; m < TM ; m += NT) smap_r [m] = (m < mA) ? (mapped ? map [i0 + m] : i0 + m) : 0 ;
        for (int m = tid ; m < TN ; m += NT) smap_c [m] = (m < nB) ? (mapped ? map [j0 + m] : j0 + m) : 0 ;
    }

    // Global -> shared staging.  Thread `tid` always copies the same tile row lm and the k columns
    // lk, lk + KPP, ... of a stage, so its source pointer just advances by a constant number of
    // columns; stages are requested in increasing k order.  A predicated-off copy has src-size 0
    // (zero fill, nothing is read).
    static_assert (NT % TM == 0 && NT % TN == 0, "thread count must be a multiple of the tile edges") ;
    constexpr int KPP_A = NT / TM, KPP_B = NT / TN ;
    const int lmA = tid % TM, lkA = tid / TM, lmB = tid % TN, lkB = tid / TN ;
    const bool rokA = lmA < mA, rokB = lmB < nB ;
    const double *gA = Arow + lmA + (int64_t) (kbeg + lkA) * lda ;
    const double *gB = Brow + lmB + (int64_t) (kbeg + lkB) * lda ;
    const int64_t stepA = (int64_t) KPP_A * lda, stepB = (int64_t) KPP_B * lda ;
    const unsigned sA0 = (unsigned) __cvta_generic_to_shared (sA) + (unsigned) ((lkA * LDA_S + lmA) * 8) ;
    const unsigned sB0 = (unsigned) __cvta_generic_to_shared (sB) + (unsigned) ((lkB * LDB_S + lmB) * 8) ;
    int kleft = kend - kbeg ;           // columns not yet requested
    auto load_stage = [&] (int buf)
    {
        const unsigned dA = sA0 + (unsigned) (buf * DM_KC * LDA_S * 8) ;
        const unsigned dB = sB0 + (unsigned) (buf * DM_KC * LDB_S * 8) ;
        const double *pa = gA, *pb = gB ;
        #pragma unroll
        for (int i = 0 ; i < DM_KC / KPP_A ; i++)
        {
            int sz = (rokA && (lkA + i * KPP_A < kleft)) ? 8 : 0 ;
            asm volatile ("cp.async.ca.shared.global [%0], [%1], 8, %2;\n"
                :: "r" (dA + (unsigned) (i * KPP_A * LDA_S * 8)), "l" (pa), "r" (sz)) ;
            pa += stepA ;
        }
        #pragma unroll
        for (int i = 0 ; i < DM_KC / KPP_B ; i++)
        {
            int sz = (rokB && (lkB + i * KPP_B < kleft)) ? 8 : 0 ;
            asm volatile ("cp.async.ca.shared.global [%0], [%1], 8, %2;\n"
                :: "r" (dB + (unsigned) (i * KPP_B * LDB_S * 8)), "l" (pb), "r" (sz)) ;
            pb += stepB ;
        }
        gA = pa ; gB = pb ; kleft -= DM_KC ;
    } ;

    double acc [MT][NTL][2] ;
    #pragma unroll
    for (int a = 0 ; a < MT ; a++)
        #pragma unroll
        for (int b = 0 ; b < NTL ; b++) { acc [a][b][0] = 0.0 ; acc [a][b][1] = 0.0 ; }

    // a warp whose whole sub-tile lies strictly above the diagonal has nothing to compute
    const bool warp_live = (i0 + wm * WM + WM - 1 + it.shift) >= (j0 + wn * WN) ;

    #pragma unroll
    for (int s = 0 ; s < DM_STAGES - 1 ; s++)
    {
        if (s < nk) load_stage (s) ;
        cp_async_commit () ;
    }
    const int fk = lane & 3, fm = lane >> 2 ;
    int cbuf = 0, lbuf = DM_STAGES - 1 ;        // buffer being computed on / being filled
    for (int kt = 0 ; kt < nk ; kt++)
    {
        cp_async_wait<DM_STAGES - 2> () ;
        __syncthreads () ;
        if (kt + DM_STAGES - 1 < nk) load_stage (lbuf) ;
        cp_async_commit () ;
        lbuf = (lbuf + 1 == DM_STAGES) ? 0 : lbuf + 1 ;
        const int cur = cbuf ;
        cbuf = (cbuf + 1 == DM_STAGES) ? 0 : cbuf + 1 ;
        if (warp_live)
        {
            const double *cA = sA + cur * DM_KC * LDA_S + wm * WM + fm ;
            const double *cB = sB + cur * DM_KC * LDB_S + wn * WN + fm ;
            #pragma unroll
            for (int kk = 0 ; kk < DM_KC ; kk += 4)
            {
                double fa [MT], fb [NTL] ;
                #pragma unroll
                for (int a = 0 ; a < MT ; a++) fa [a] = cA [(kk + fk) * LDA_S + a * 8] ;
                #pragma unroll
                for (int b = 0 ; b < NTL ; b++) fb [b] = cB [(kk + fk) * LDB_S + b * 8] ;
                #pragma unroll
                for (int a = 0 ; a < MT ; a++)
                    #pragma unroll
                    for (int b = 0 ; b < NTL ; b++)
                        dmma884 (acc [a][b][0], acc [a][b][1], fa [a], fb [b]) ;
            }
        }
    }
    cp_async_wait<0> () ;
    __syncthreads () ;
    if (!warp_live) return ;

    // epilogue: C (i,j) held by lane: i = 8a + lane/4, j = 8b + 2*(lane%4) + {0,1} within the warp tile
    double *C = Lx + it.c_off ;
    const int64_t ldc = it.ldc ;
    #pragma unroll
    for (int b = 0 ; b < NTL ; b++)
    {
        #pragma unroll
        for (int e = 0 ; e < 2 ; e++)
        {
            const int jl = wn * WN + b * 8 + 2 * fk + e ;
            if (jl >= nB) continue ;
            const int j = j0 + jl ;
            double *Cj = C + (int64_t) smap_c [jl] * ldc ;
            #pragma unroll
            for (int a = 0 ; a < MT ; a++)
            {
                const int il = wm * WM + a * 8 + fm ;
                if (il < mA && i0 + il + it.shift >= j) atomicAdd (Cj + smap_r [il], -acc [a][b][e]) ;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// complex double: the same tiles, four real DMMAs per complex product
// ---------------------------------------------------------------------------------------------

// C = A * B^H on (re,im)-interleaved operands (zherk / zgemm "N","C" of SNW:786-803, :850-876):
//     Cre += Are Bre' + Aim Bim'          Cim += Aim Bre' - Are Bim'
// One m8n8k4 fragment element is a 16-byte (re,im) pair fetched with one LDS.128; the shared tiles are
// [k][m] with the row pitch padded to TM+2 pairs, so the 8 lanes of a quarter-warp (k = lane%4, m = lane/4)
// hit 8 distinct 16-byte bank groups.  Twice the bytes but four times the flops of the real kernel per
// element staged, so a stage is 8 columns deep instead of 16.
constexpr int ZM_KC = 8 ;

template <int TM, int TN> constexpr size_t zmma_smem_bytes ()
{
    return (size_t) DM_STAGES * ZM_KC * ((TM + 2) + (TN + 2)) * sizeof (double2) + (TM + TN) * sizeof (int32_t) ;
}

__device__ __forceinline__ double neg_ (double x)
{
    // sign flip on the integer pipe: the FP64 pipe is the one this kernel saturates
    return __hiloint2double (__double2hiint (x) ^ (int) 0x80000000, __double2loint (x)) ;
}

// acc (re,im) [MT][NTL][2] += A-fragments * B-fragments^H over one shared-memory stage of KC columns
template <int MT, int NTL, int KC, int LDA_S, int LDB_S>
__device__ __forceinline__ void zmma_stage (const double2 *cA, const double2 *cB, int fk,
    double (&cre) [MT][NTL][2], double (&cim) [MT][NTL][2])
{
    #pragma unroll
    for (int kk = 0 ; kk < KC ; kk += 4)
    {
        double2 fa [MT], fb [NTL] ;
        double nre [MT] ;
        #pragma unroll
        for (int a = 0 ; a < MT ; a++) { fa [a] = cA [(kk + fk) * LDA_S + a * 8] ; nre [a] = neg_ (fa [a].x) ; }
        #pragma unroll
        for (int b = 0 ; b < NTL ; b++) fb [b] = cB [(kk + fk) * LDB_S + b * 8] ;
        #pragma unroll
        for (int a = 0 ; a < MT ; a++)
            #pragma unroll
            for (int b = 0 ; b < NTL ; b++)
            {
                dmma884 (cre [a][b][0], cre [a][b][1], fa [a].x, fb [b].x) ;
                dmma884 (cre [a][b][0], cre [a][b][1], fa [a].y, fb [b].y) ;
                dmma884 (cim [a][b][0], cim [a][b][1], fa [a].y, fb [b].x) ;
                dmma884 (cim [a][b][0], cim [a][b][1], nre [a], fb [b].y) ;
            }
    }
}

template <int TM, int TN, int WM, int WN>
__global__ void __launch_bounds__ ((TM / WM) * (TN / WN) * 32) k_upd_zmma (const UpdTile *__restrict__ tiles,
    const UpdItem *__restrict__ items, const int32_t *__restrict__ relmap, zdouble *Lx)
{
    constexpr int NW_M = TM / WM, NW_N = TN / WN ;
    constexpr int NT = NW_M * NW_N * 32 ;
    constexpr int LDA_S = TM + 2, LDB_S = TN + 2 ;
    constexpr int MT = WM / 8, NTL = WN / 8 ;
    extern __shared__ __align__ (16) unsigned char smem_raw [] ;
    double2 *sA = reinterpret_cast<double2 *> (smem_raw) ;              // [STAGES][KC][LDA_S]
    double2 *sB = sA + DM_STAGES * ZM_KC * LDA_S ;                      // [STAGES][KC][LDB_S]
    int32_t *smap_r = reinterpret_cast<int32_t *> (sB + DM_STAGES * ZM_KC * LDB_S) ;
    int32_t *smap_c = smap_r + TM ;

    const UpdTile t = tiles [blockIdx.x] ;
    const UpdItem it = items [t.item] ;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5 ;
    const int wm = warp % NW_M, wn = warp / NW_M ;
    const int i0 = (int) t.ti * TM, j0 = (int) t.tj * TN ;
    const int mA = min (TM, it.M - i0), nB = min (TN, it.N - j0) ;
    const int64_t lda = it.lda ;
    const zdouble *A = Lx + it.a_off ;
    const int kbeg = t.k0, kend = t.k1 ;
    const int nk = (kend - kbeg + ZM_KC - 1) / ZM_KC ;
    {
        const bool mapped = it.map_off >= 0 ;
        const int32_t *map = relmap + (mapped ? it.map_off : 0) ;
        for (int m = tid ; m < TM ; m += NT) smap_r [m] = (m < mA) ? (mapped ? map [i0 + m] : i0 + m) : 0 ;
        for (int m = tid ; m < TN ; m += NT) smap_c [m] = (m < nB) ? (mapped ? map [j0 + m] : j0 + m) : 0 ;
    }
    static_assert (NT % TM == 0 && NT % TN == 0, "thread count must be a multiple of the tile edges") ;
    constexpr int KPP_A = NT / TM, KPP_B = NT / TN ;
    static_assert (ZM_KC % KPP_A == 0 && ZM_KC % KPP_B == 0, "a stage must be a whole number of passes") ;
    const int lmA = tid % TM, lkA = tid / TM, lmB = tid % TN, lkB = tid / TN ;
    const bool rokA = lmA < mA, rokB = lmB < nB ;
    const zdouble *gA = A + i0 + lmA + (int64_t) (kbeg + lkA) * lda ;
    const zdouble *gB = A - it.shift + j0 + lmB + (int64_t) (kbeg + lkB) * lda ;
    const int64_t stepA = (int64_t) KPP_A * lda, stepB = (int64_t) KPP_B * lda ;
    const unsigned sA0 = (unsigned) __cvta_generic_to_shared (sA) + (unsigned) ((lkA * LDA_S + lmA) * 16) ;
    const unsigned sB0 = (unsigned) __cvta_generic_to_shared (sB) + (unsigned) ((lkB * LDB_S + lmB) * 16) ;
    int kleft = kend - kbeg ;
    auto load_stage = [&] (int buf)
    {
        const unsigned dA = sA0 + (unsigned) (buf * ZM_KC * LDA_S * 16) ;
        const unsigned dB = sB0 + (unsigned) (buf * ZM_KC * LDB_S * 16) ;
        const zdouble *pa = gA, *pb = gB ;
        #pragma unroll
        for (int i = 0 ; i < ZM_KC / KPP_A ; i++)
        {
            int sz = (rokA && (lkA + i * KPP_A < kleft)) ? 16 : 0 ;
            asm volatile ("cp.async.cg.shared.global [%0], [%1], 16, %2;\n"
                :: "r" (dA + (unsigned) (i * KPP_A * LDA_S * 16)), "l" (pa), "r" (sz)) ;
            pa += stepA ;
        }
        #pragma unroll
        for (int i = 0 ; i < ZM_KC / KPP_B ; i++)
        {
            int sz = (rokB && (lkB + i * KPP_B < kleft)) ? 16 : 0 ;
            asm volatile ("cp.async.cg.shared.global [%0], [%1], 16, %2;\n"
                :: "r" (dB + (unsigned) (i * KPP_B * LDB_S * 16)), "l" (pb), "r" (sz)) ;
            pb += stepB ;
        }
        gA = pa ; gB = pb ; kleft -= ZM_KC ;
    } ;

    double cre [MT][NTL][2], cim [MT][NTL][2] ;
    #pragma unroll
    for (int a = 0 ; a < MT ; a++)
        #pragma unroll
        for (int b = 0 ; b < NTL ; b++) { cre [a][b][0] = 0.0 ; cre [a][b][1] = 0.0 ; cim [a][b][0] = 0.0 ; cim [a][b][1] = 0.0 ; }
    const bool warp_live = (i0 + wm * WM + WM - 1 + it.shift) >= (j0 + wn * WN) ;

    #pragma unroll
    for (int s = 0 ; s < DM_STAGES - 1 ; s++)
    {
        if (s < nk) load_stage (s) ;
        cp_async_commit () ;
    }
    const int fk = lane & 3, fm = lane >> 2 ;
    int cbuf = 0, lbuf = DM_STAGES - 1 ;
    for (int kt = 0 ; kt < nk ; kt++)
    {
        cp_async_wait<DM_STAGES - 2> () ;
        __syncthreads () ;
        if (kt + DM_STAGES - 1 < nk) load_stage (lbuf) ;
        cp_async_commit () ;
        lbuf = (lbuf + 1 == DM_STAGES) ? 0 : lbuf + 1 ;
        const int cur = cbuf ;
        cbuf = (cbuf + 1 == DM_STAGES) ? 0 : cbuf + 1 ;
        if (warp_live)
            zmma_stage<MT, NTL, ZM_KC, LDA_S, LDB_S> (sA + cur * ZM_KC * LDA_S + wm * WM + fm,
                sB + cur * ZM_KC * LDB_S + wn * WN + fm, fk, cre, cim) ;
    }
    cp_async_wait<0> () ;
    __syncthreads () ;
    if (!warp_live) return ;

    zdouble *C = Lx + it.c_off ;
    const int64_t ldc = it.ldc ;
    #pragma unroll
    for (int b = 0 ; b < NTL ; b++)
    {
        #pragma unroll
        for (int e = 0 ; e < 2 ; e++)
        {
            const int jl = wn * WN + b * 8 + 2 * fk + e ;
            if (jl >= nB) continue ;
            const int j = j0 + jl ;
            zdouble *Cj = C + (int64_t) smap_c [jl] * ldc ;
            #pragma unroll
            for (int a = 0 ; a < MT ; a++)
            {
                const int il = wm * WM + a * 8 + fm ;
                if (il < mA && i0 + il + it.shift >= j)
                {
                    zdouble *dst = Cj + smap_r [il] ;
                    atomicAdd (&dst->re, -cre [a][b][e]) ;
                    // the diagonal of L1*L1^H is real (zherk leaves its imaginary part exactly zero)
                    if (i0 + il + it.shift > j) atomicAdd (&dst->im, -cim [a][b][e]) ;
                }
            }
        }
    }
}

// complex single (and any T without a tensor path): register-blocked tiles on the FMA pipe, a 4 x 4
// micro-tile per thread, the same 3-stage cp.async ring and relative-map epilogue
template <typename T, int TM, int TN> constexpr size_t cfma_smem_bytes ()
{
    return (size_t) DM_STAGES * DM_KC * (TM + TN) * sizeof (T) + (TM + TN) * sizeof (int32_t) ;
}

template <typename T, int TM, int TN>
__global__ void __launch_bounds__ ((TM / 4) * (TN / 4)) k_upd_cfma (const UpdTile *__restrict__ tiles,
    const UpdItem *__restrict__ items, const int32_t *__restrict__ relmap, T *Lx)
{
    constexpr int NTX = TM / 4, NTY = TN / 4, NT = NTX * NTY ;
    constexpr int ES = (int) sizeof (T) ;
    static_assert (ES == 4 || ES == 8 || ES == 16, "cp.async moves 4, 8 or 16 bytes") ;
    extern __shared__ __align__ (16) unsigned char smem_raw [] ;
    T *sA = reinterpret_cast<T *> (smem_raw) ;                          // [STAGES][KC][TM]
    T *sB = sA + DM_STAGES * DM_KC * TM ;                               // [STAGES][KC][TN]
    int32_t *smap_r = reinterpret_cast<int32_t *> (sB + DM_STAGES * DM_KC * TN) ;
    int32_t *smap_c = smap_r + TM ;
    const UpdTile t = tiles [blockIdx.x] ;
    const UpdItem it = items [t.item] ;
    const int tid = threadIdx.x, tx = tid % NTX, ty = tid / NTX ;
    const int i0 = (int) t.ti * TM, j0 = (int) t.tj * TN ;
    const int mA = min (TM, it.M - i0), nB = min (TN, it.N - j0) ;
    const int64_t lda = it.lda ;
    const T *A = Lx + it.a_off ;
    const int kbeg = t.k0, kend = t.k1 ;
    const int nk = (kend - kbeg + DM_KC - 1) / DM_KC ;
    {
        const bool mapped = it.map_off >= 0 ;
        const int32_t *map = relmap + (mapped ? it.map_off : 0) ;
        for (int m = tid ; m < TM ; m += NT) smap_r [m] = (m < mA) ? (mapped ? map [i0 + m] : i0 + m) : 0 ;
        for (int m = tid ; m < TN ; m += NT) smap_c [m] = (m < nB) ? (mapped ? map [j0 + m] : j0 + m) : 0 ;
    }
    auto load_stage = [&] (int kt, int buf)
    {
        const int kb = kbeg + kt * DM_KC ;
        #pragma unroll
        for (int e = tid ; e < DM_KC * TM ; e += NT)
        {
            const int m = e % TM, k = e / TM ;
            const int sz = (m < mA && kb + k < kend) ? ES : 0 ;
            const unsigned d = (unsigned) __cvta_generic_to_shared (sA + (buf * DM_KC + k) * TM + m) ;
            asm volatile ("cp.async.ca.shared.global [%0], [%1], %2, %3;\n"
                :: "r" (d), "l" (A + i0 + m + (int64_t) (kb + k) * lda), "n" (ES), "r" (sz)) ;
        }
        #pragma unroll
        for (int e = tid ; e < DM_KC * TN ; e += NT)
        {
            const int m = e % TN, k = e / TN ;
            const int sz = (m < nB && kb + k < kend) ? ES : 0 ;
            const unsigned d = (unsigned) __cvta_generic_to_shared (sB + (buf * DM_KC + k) * TN + m) ;
            asm volatile ("cp.async.ca.shared.global [%0], [%1], %2, %3;\n"
                :: "r" (d), "l" (A - it.shift + j0 + m + (int64_t) (kb + k) * lda), "n" (ES), "r" (sz)) ;
        }
    } ;
    const T zero = from_real<T> (0) ;
    T acc [4][4] ;
    #pragma unroll
    for (int a = 0 ; a < 4 ; a++)
        #pragma unroll
        for (int b = 0 ; b < 4 ; b++) acc [a][b] = zero ;
    const bool live = (i0 + tx * 4 + 3 + it.shift) >= (j0 + ty * 4) ;
    #pragma unroll
    for (int s = 0 ; s < DM_STAGES - 1 ; s++)
    {
        if (s < nk) load_stage (s, s) ;
        cp_async_commit () ;
    }
    for (int kt = 0 ; kt < nk ; kt++)
    {
        cp_async_wait<DM_STAGES - 2> () ;
        __syncthreads () ;
        if (kt + DM_STAGES - 1 < nk) load_stage (kt + DM_STAGES - 1, (kt + DM_STAGES - 1) % DM_STAGES) ;
        cp_async_commit () ;
        if (live)
        {
            const T *cA = sA + (kt % DM_STAGES) * DM_KC * TM + tx * 4 ;
            const T *cB = sB + (kt % DM_STAGES) * DM_KC * TN + ty * 4 ;
            #pragma unroll
            for (int k = 0 ; k < DM_KC ; k++)
            {
                T fa [4], fb [4] ;
                #pragma unroll
                for (int a = 0 ; a < 4 ; a++) fa [a] = cA [k * TM + a] ;
                #pragma unroll
                for (int b = 0 ; b < 4 ; b++) fb [b] = cB [k * TN + b] ;
                #pragma unroll
                for (int a = 0 ; a < 4 ; a++)
                    #pragma unroll
                    for (int b = 0 ; b < 4 ; b++) acc [a][b] += mulc (fa [a], fb [b]) ;
            }
        }
    }
    cp_async_wait<0> () ;
    __syncthreads () ;
    if (!live) return ;
    T *C = Lx + it.c_off ;
    const int64_t ldc = it.ldc ;
    #pragma unroll
    for (int b = 0 ; b < 4 ; b++)
    {
        const int jl = ty * 4 + b ;
        if (jl >= nB) continue ;
        const int j = j0 + jl ;
        T *Cj = C + (int64_t) smap_c [jl] * ldc ;
        #pragma unroll
        for (int a = 0 ; a < 4 ; a++)
        {
            const int il = tx * 4 + a ;
            if (il < mA && i0 + il + it.shift >= j) atomic_sub (Cj + smap_r [il], acc [a][b]) ;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// single precision: the same tiles on the FP32 FMA pipe
// ---------------------------------------------------------------------------------------------

// FP32 has no tensor path of FP32 accuracy (tf32 keeps 10 mantissa bits), so the single-precision
// factorization (config 5b) runs the update tiles as a classic register-blocked SGEMM: TM x TN tile,
// (TM/8) x (TN/8) threads with an 8 x 8 micro-tile each, operands staged [k][m] by a 3-stage cp.async
// ring exactly like k_upd_dmma, epilogue = red.global.add.f32 of -C through the relative map.
template <int TM, int TN> constexpr size_t ffma_smem_bytes ()
{
    return (size_t) DM_STAGES * DM_KC * (TM + TN) * sizeof (float) + (TM + TN) * sizeof (int32_t) ;
}

template <int TM, int TN>
__global__ void __launch_bounds__ ((TM / 8) * (TN / 8)) k_upd_ffma (const UpdTile *__restrict__ tiles,
    const UpdItem *__restrict__ items, const int32_t *__restrict__ relmap, float *Lx)
{
    constexpr int NTX = TM / 8, NTY = TN / 8, NT = NTX * NTY ;
    extern __shared__ __align__ (16) unsigned char smem_raw [] ;
    float *sA = reinterpret_cast<float *> (smem_raw) ;                   // [STAGES][KC][TM]
    float *sB = sA + DM_STAGES * DM_KC * TM ;                            // [STAGES][KC][TN]
    int32_t *smap_r = reinterpret_cast<int32_t *> (sB + DM_STAGES * DM_KC * TN) ;
    int32_t *smap_c = smap_r + TM ;
    const UpdTile t = tiles [blockIdx.x] ;
    const UpdItem it = items [t.item] ;
    const int tid = threadIdx.x, tx = tid % NTX, ty = tid / NTX ;
    const int i0 = (int) t.ti * TM, j0 = (int) t.tj * TN ;
    const int mA = min (TM, it.M - i0), nB = min (TN, it.N - j0) ;
    const int64_t lda = it.lda ;
    const float *A = Lx + it.a_off ;
    const int kbeg = t.k0, kend = t.k1 ;
    const int nk = (kend - kbeg + DM_KC - 1) / DM_KC ;
    {
        const bool mapped = it.map_off >= 0 ;
        const int32_t *map = relmap + (mapped ? it.map_off : 0) ;
        for (int m = tid ; m < TM ; m += NT) smap_r [m] = (m < mA) ? (mapped ? map [i0 + m] : i0 + m) : 0 ;
        for (int m = tid ; m < TN ; m += NT) smap_c [m] = (m < nB) ? (mapped ? map [j0 + m] : j0 + m) : 0 ;
    }
    auto load_stage = [&] (int kt, int buf)
    {
        const int kb = kbeg + kt * DM_KC ;
        #pragma unroll
        for (int e = tid ; e < DM_KC * TM ; e += NT)
        {
            const int m = e % TM, k = e / TM ;
            const int sz = (m < mA && kb + k < kend) ? 4 : 0 ;
            const unsigned d = (unsigned) __cvta_generic_to_shared (sA + (buf * DM_KC + k) * TM + m) ;
            asm volatile ("cp.async.ca.shared.global [%0], [%1], 4, %2;\n"
                :: "r" (d), "l" (A + i0 + m + (int64_t) (kb + k) * lda), "r" (sz)) ;
        }
        #pragma unroll
        for (int e = tid ; e < DM_KC * TN ; e += NT)
        {
            const int m = e % TN, k = e / TN ;
            const int sz = (m < nB && kb + k < kend) ? 4 : 0 ;
            const unsigned d = (unsigned) __cvta_generic_to_shared (sB + (buf * DM_KC + k) * TN + m) ;
            asm volatile ("cp.async.ca.shared.global [%0], [%1], 4, %2;\n"
                :: "r" (d), "l" (A - it.shift + j0 + m + (int64_t) (kb + k) * lda), "r" (sz)) ;
        }
    } ;
    float acc [8][8] ;
    #pragma unroll
    for (int a = 0 ; a < 8 ; a++)
        #pragma unroll
        for (int b = 0 ; b < 8 ; b++) acc [a][b] = 0.f ;
    // a thread whose whole micro-tile lies strictly above the diagonal has nothing to compute
    const bool live = (i0 + tx * 8 + 7 + it.shift) >= (j0 + ty * 8) ;
    #pragma unroll
    for (int s = 0 ; s < DM_STAGES - 1 ; s++)
    {
        if (s < nk) load_stage (s, s) ;
        cp_async_commit () ;
    }
    for (int kt = 0 ; kt < nk ; kt++)
    {
        cp_async_wait<DM_STAGES - 2> () ;
        __syncthreads () ;
        if (kt + DM_STAGES - 1 < nk) load_stage (kt + DM_STAGES - 1, (kt + DM_STAGES - 1) % DM_STAGES) ;
        cp_async_commit () ;
        if (live)
        {
            const float *cA = sA + (kt % DM_STAGES) * DM_KC * TM + tx * 8 ;
            const float *cB = sB + (kt % DM_STAGES) * DM_KC * TN + ty * 8 ;
            #pragma unroll
            for (int k = 0 ; k < DM_KC ; k++)
            {
                const float4 a0 = *reinterpret_cast<const float4 *> (cA + k * TM) ;
                const float4 a1 = *reinterpret_cast<const float4 *> (cA + k * TM + 4) ;
                const float4 b0 = *reinterpret_cast<const float4 *> (cB + k * TN) ;
                const float4 b1 = *reinterpret_cast<const float4 *> (cB + k * TN + 4) ;
                const float fa [8] = { a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w } ;
                const float fb [8] = { b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w } ;
                #pragma unroll
                for (int a = 0 ; a < 8 ; a++)
                    #pragma unroll
                    for (int b = 0 ; b < 8 ; b++) acc [a][b] += fa [a] * fb [b] ;
            }
        }
    }
    cp_async_wait<0> () ;
    __syncthreads () ;
    if (!live) return ;
    float *C = Lx + it.c_off ;
    const int64_t ldc = it.ldc ;
    #pragma unroll
    for (int b = 0 ; b < 8 ; b++)
    {
        const int jl = ty * 8 + b ;
        if (jl >= nB) continue ;
        const int j = j0 + jl ;
        float *Cj = C + (int64_t) smap_c [jl] * ldc ;
        #pragma unroll
        for (int a = 0 ; a < 8 ; a++)
        {
            const int il = tx * 8 + a ;
            if (il < mA && i0 + il + it.shift >= j) atomicAdd (Cj + smap_r [il], -acc [a][b]) ;
        }
    }
}

// Register-resident DMMA loop: the FP64 tensor peak this chip sustains (MEASURED_PEAKS.json has no FP64
// figure, so bench.py measures the roofline denominator with this kernel on the box it runs on).
static __global__ void __launch_bounds__ (256) k_dmma_peak (double *out, int iters)
{
    double acc [16][2] ;
    #pragma unroll
    for (int i = 0 ; i < 16 ; i++) { acc [i][0] = 0.0 ; acc [i][1] = 0.0 ; }
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9 ;
    for (int it = 0 ; it < iters ; it++)
    {
        #pragma unroll
        for (int i = 0 ; i < 16 ; i++) dmma884 (acc [i][0], acc [i][1], a, b) ;
    }
    double sum = 0 ;
    #pragma unroll
    for (int i = 0 ; i < 16 ; i++) sum += acc [i][0] + acc [i][1] ;
    out [(size_t) blockIdx.x * blockDim.x + threadIdx.x] = sum ;
}

// ---------------------------------------------------------------------------------------------
// dense Cholesky of diagonal blocks (jb <= NB <= 64), one thread per row, the row in registers
// ---------------------------------------------------------------------------------------------

constexpr int PB_LD = POTRF_NB + 1 ;
constexpr int WINV_LD = POTRF_NB ;              // an inverse slot is a 64 x 64 column-major array

// Narrow blocks (jb <= NB = 16, the leaves: 98 % of all supernodes): a CTA of 64 threads factorizes 4
// blocks, one 16-thread group each.  Thread r keeps row r of its block in registers.  Right-looking: at
// step j every thread publishes its (still unscaled) entry of column j through shared memory, then
// subtracts (a_rj / d) * a_cj from its row (d = the pivot) and finally scales its own entry by
// 1/sqrt(d); one barrier per column.  On a pivot that is not > 0 (or NaN) the global column index is
// recorded in *minor (atomicMin): the smallest failing column over the whole factorization is the
// reference's L->minor (SNW:1097).
template <typename T, int NB>
__global__ void __launch_bounds__ (64) k_potrf_reg (const PotrfBlk *__restrict__ blks, int nblks, T *Lx,
    unsigned long long *minor)
{
    constexpr int G = 64 / NB ;
    __shared__ T colbuf [2][64] ;
    const int g = threadIdx.x / NB, r = threadIdx.x % NB ;
    const int bi = blockIdx.x * G + g ;
    const bool active = bi < nblks ;
    PotrfBlk b ;
    if (active) b = blks [bi] ; else { b.off = 0 ; b.col = 0 ; b.lda = 0 ; b.jb = 0 ; b.slot = -1 ; }
    const int jb = b.jb ;
    const int64_t lda = b.lda ;
    T *D = Lx + b.off ;
    typedef typename num<T>::real R ;
    const T one = from_real<T> (1), zero = from_real<T> (0) ;
    T a [NB] ;
    #pragma unroll
    for (int c = 0 ; c < NB ; c++)
        a [c] = (r < jb && c <= r) ? D [r + c * lda] : ((r == c) ? one : zero) ;
    #pragma unroll
    for (int j = 0 ; j < NB ; j++)
    {
        T *cb = colbuf [j & 1] + g * NB ;
        if (r >= j) cb [r] = a [j] ;
        __syncthreads () ;
        const R d = real_ (cb [j]) ;            // the pivot: only its real part exists (zpotrf)
        if (r == j && j < jb && !(d > (R) 0)) atomicMin (minor, (unsigned long long) (b.col + j)) ;
        const T t = a [j] * rcp_ (d) ;
        #pragma unroll
        for (int c = j + 1 ; c < NB ; c++) if (c <= r) a [c] -= mulc (t, cb [c]) ;
        const R ljj = sqrt (d) ;
        if (r == j) a [j] = from_real<T> (ljj) ;
        else if (r > j) a [j] = a [j] * rcp_ (ljj) ;
    }
    #pragma unroll
    for (int c = 0 ; c < NB ; c++) if (r < jb && c <= r) D [r + c * lda] = a [c] ;
}

// Whole small supernodes (nscol <= 16, at most SMALL_ROWS rows below the diagonal block; the bulk of the
// leaves): the same register Cholesky as k_potrf_reg, and then the same 16-thread group solves the rows
// below, 16 at a time, by substitution against the factor it has just put in shared memory -- dpotrf and
// dtrsm of SNW:1023 and :1175 in one launch, no second kernel, no 128-thread tile for a handful of rows.
template <typename T>
__global__ void __launch_bounds__ (64) k_factor_small (const PotrfBlk *__restrict__ blks, int nblks, T *Lx,
    unsigned long long *minor)
{
    constexpr int NB = 16, G = 4 ;
    __shared__ T colbuf [2][64] ;
    __shared__ T sL [G][NB][NB + 1] ;
    const int g = threadIdx.x / NB, r = threadIdx.x % NB ;
    const int bi = blockIdx.x * G + g ;
    const bool active = bi < nblks ;
    PotrfBlk b ;
    if (active) b = blks [bi] ; else { b.off = 0 ; b.col = 0 ; b.lda = 0 ; b.jb = 0 ; b.slot = -1 ; }
    const int jb = b.jb ;
    const int64_t lda = b.lda ;
    T *D = Lx + b.off ;
    typedef typename num<T>::real R ;
    const T one = from_real<T> (1), zero = from_real<T> (0) ;
    T a [NB] ;
    #pragma unroll
    for (int c = 0 ; c < NB ; c++)
        a [c] = (r < jb && c <= r) ? D [r + c * lda] : ((r == c) ? one : zero) ;
    #pragma unroll
    for (int j = 0 ; j < NB ; j++)
    {
        T *cb = colbuf [j & 1] + g * NB ;
        if (r >= j) cb [r] = a [j] ;
        __syncthreads () ;
        const R d = real_ (cb [j]) ;            // the pivot: only its real part exists (zpotrf)
        if (r == j && j < jb && !(d > (R) 0)) atomicMin (minor, (unsigned long long) (b.col + j)) ;
        const T t = a [j] * rcp_ (d) ;
        #pragma unroll
        for (int c = j + 1 ; c < NB ; c++) if (c <= r) a [c] -= mulc (t, cb [c]) ;
        const R ljj = sqrt (d) ;
        if (r == j) a [j] = from_real<T> (ljj) ;
        else if (r > j) a [j] = a [j] * rcp_ (ljj) ;
    }
    #pragma unroll
    for (int c = 0 ; c < NB ; c++)
    {
        if (r < jb && c <= r) D [r + c * lda] = a [c] ;
        sL [g][r][c] = (c <= r) ? a [c] : zero ;        // identity padded: rows / columns >= jb are e_r
    }
    __syncthreads () ;
    // rows below the diagonal block: x <- x * L11^{-T}, one row per thread
    const int below = (int) lda - jb ;
    for (int r0 = 0 ; r0 < below ; r0 += NB)
    {
        const int row = r0 + r ;
        if (row < below)
        {
            T *X = D + jb + row ;
            T x [NB] ;
            #pragma unroll
            for (int j = 0 ; j < NB ; j++) x [j] = (j < jb) ? X [j * lda] : zero ;
            #pragma unroll
            for (int j = 0 ; j < NB ; j++)
            {
                const T xj = x [j] / real_ (sL [g][j][j]) ;
                x [j] = xj ;
                #pragma unroll
                for (int k = j + 1 ; k < NB ; k++) x [k] -= mulc (xj, sL [g][k][j]) ;
            }
            #pragma unroll
            for (int j = 0 ; j < NB ; j++) if (j < jb) X [j * lda] = x [j] ;
        }
    }
}

// W = L^{-1} of a (padded) 64 x 64 lower triangular block held in shared memory, by a CTA of 256 threads:
// s [c * PB_LD + r] = L (r,c), rsq [r] = 1 / L (r,r), result w [r * PB_LD + c] = W (r,c), tmp = 32 x 33 scratch.
// Block-recursive, so that the serial chain is 16 rows instead of 64:
//  (A) the four 16 x 16 diagonal blocks: thread c < 64 solves L16 * w = e_c for ITS column by forward
//      substitution entirely in registers (no shuffle, no barrier inside the chain; the entries of L are
//      broadcast reads);
//  (B), (C) the off-diagonal blocks follow from W21 = -W22 * (L21 * W11), first for the two 16-blocks inside each
//      32-block, then for the 32-block.  In both products a thread owns nb/8 entries of one column j (rows
//      tt%8, tt%8+8, ...): the operand that depends on j alone is loaded once for all of them, the other one
//      is conflict-free across the lanes.
template <typename T>
__device__ __forceinline__ void inv64_lower (const T *s, T *w, const T *rsq, T *tmp)
{
    const T one = from_real<T> (1), zero = from_real<T> (0) ;
    if (threadIdx.x < POTRF_NB)
    {
        const int c = threadIdx.x, b0 = c & ~15, cl = c & 15 ;
        T wc [16] ;
        #pragma unroll
        for (int rl = 0 ; rl < 16 ; rl++)
        {
            T acc0 = zero, acc1 = zero ;
            #pragma unroll
            for (int kl = 0 ; kl + 1 < rl ; kl += 2)
            {
                acc0 += s [(b0 + kl) * PB_LD + b0 + rl] * wc [kl] ;
                acc1 += s [(b0 + kl + 1) * PB_LD + b0 + rl] * wc [kl + 1] ;
            }
            if (rl & 1) acc0 += s [(b0 + rl - 1) * PB_LD + b0 + rl] * wc [rl - 1] ;
            // rows above the diagonal of column c come out as exact zeros: (0 - 0) * rsq
            wc [rl] = (((rl == cl) ? one : zero) - (acc0 + acc1)) * real_ (rsq [b0 + rl]) ;
        }
        #pragma unroll
        for (int rl = 0 ; rl < 16 ; rl++) if (rl >= cl) w [(b0 + rl) * PB_LD + c] = wc [rl] ;
    }
    __syncthreads () ;
    #pragma unroll 1
    for (int pass = 0 ; pass < 2 ; pass++)
    {
        // pass 0: blocks (rows 16..31, cols 0..15) and (rows 48..63, cols 32..47); pass 1: (rows 32..63, cols 0..31)
        const int nb = pass ? 32 : 16, nu = nb >> 3 ;
        const int half = (int) threadIdx.x / (8 * nb), tt = (int) threadIdx.x % (8 * nb) ;
        const int j = tt >> 3, i0 = tt & 7 ;               // entries (i0 + 8u, j), u < nu
        const int c0 = half * 32, r0 = c0 + nb, t0 = half * 16 ;
        // step 1: tmp (i,j) = sum_{k >= j} L (r0+i, c0+k) W (c0+k, c0+j)
        T acc [4] = { zero, zero, zero, zero } ;
        for (int k = j ; k < nb ; k++)
        {
            const T wk = w [(c0 + k) * PB_LD + c0 + j] ;
            const T *sk = s + (c0 + k) * PB_LD + r0 + i0 ;
            #pragma unroll
            for (int u = 0 ; u < 4 ; u++) if (u < nu) acc [u] += sk [8 * u] * wk ;
        }
        #pragma unroll
        for (int u = 0 ; u < 4 ; u++) if (u < nu) tmp [(t0 + i0 + 8 * u) * 33 + j] = acc [u] ;
        __syncthreads () ;
        // step 2: W (r0+i, c0+j) = - sum_{k <= i} W (r0+i, r0+k) tmp (k,j)
        #pragma unroll
        for (int u = 0 ; u < 4 ; u++) acc [u] = zero ;
        const int imax = i0 + 8 * (nu - 1) ;
        for (int k = 0 ; k <= imax ; k++)
        {
            const T tk = tmp [(t0 + k) * 33 + j] ;
            #pragma unroll
            for (int u = 0 ; u < 4 ; u++)
                if (u < nu && k <= i0 + 8 * u) acc [u] += w [(r0 + i0 + 8 * u) * PB_LD + r0 + k] * tk ;
        }
        #pragma unroll
        for (int u = 0 ; u < 4 ; u++) if (u < nu) w [(r0 + i0 + 8 * u) * PB_LD + c0 + j] = -acc [u] ;
        __syncthreads () ;
    }
}

template <typename T> constexpr size_t potrf_smem_bytes ()
{
    return (size_t) (2 * POTRF_NB * PB_LD + 2 * (4 * POTRF_NB + 8) + 3 * POTRF_NB + 32 * 33) * sizeof (T) ;
}

// Wider blocks (16 < jb <= 64): one CTA of 256 threads per block.  The block lives in registers as a
// 16 x 16 grid of 4 x 4 sub-tiles (thread (ti,tj) holds rows 4ti.., columns 4tj..); the loop over tile
// columns is rolled, so the code stays small (a fully unrolled one-row-per-thread version is bound by
// cold instruction fetch, a shared-memory version by read-modify-write latency).  Right-looking on the
// UNSCALED columns: at step j the owners of column j publish it through shared memory, then every
// thread subtracts a_rj * a_cj / d_j from its registers -- one barrier per column; afterwards every
// column is scaled by 1/sqrt(d_j).
// If the block has an inverse slot (slot >= 0), W = L^{-1} is also formed (4 lanes per column, row by
// row, in shared memory) and written to winv [slot], identity-padded to nbk x nbk with leading
// dimension 64: the dtrsm on the rows below then becomes a product with W' on the tensor pipe.
template <typename T>
__global__ void __launch_bounds__ (256) k_potrf_tile (const PotrfBlk *__restrict__ blks, T *Lx,
    T *__restrict__ winv, int nbk, unsigned long long *minor)
{
    extern __shared__ __align__ (16) unsigned char smem_raw [] ;
    T *s = reinterpret_cast<T *> (smem_raw) ;           // s [c * PB_LD + r] = L (r,c), lower part
    T *w = s + POTRF_NB * PB_LD ;                       // w [r * PB_LD + c] = W (r,c), lower part
    constexpr int CB = 4 * POTRF_NB + 8 ;               // one buffer: 4 columns of 64, their 4 pivots, the 4 reciprocals
    T *colbuf = w + POTRF_NB * PB_LD ;                  // [2][CB] the 4 columns of a panel, double buffered
    T *dvec = colbuf + 2 * CB ;                         // [64] pivots d_j
    T *sqv = dvec + POTRF_NB ;                          // [64] sqrt (d_j)
    T *rsq = sqv + POTRF_NB ;                           // [64] 1 / sqrt (d_j)
    const PotrfBlk b = blks [blockIdx.x] ;
    const int jb = b.jb ;
    const int64_t lda = b.lda ;
    T *D = Lx + b.off ;
    const int ti = threadIdx.x & 15, tj = threadIdx.x >> 4 ;
    typedef typename num<T>::real R ;
    const T one = from_real<T> (1), zero = from_real<T> (0) ;
    T a [4][4] ;
    #pragma unroll
    for (int jc = 0 ; jc < 4 ; jc++)
        #pragma unroll
        for (int i = 0 ; i < 4 ; i++)
        {
            const int r = 4 * ti + i, c = 4 * tj + jc ;
            a [i][jc] = (r < jb && c < jb && r >= c) ? D [r + c * lda] : ((r == c) ? one : zero) ;
        }
    const int njt = (jb + 3) >> 2 ;
    // One barrier per PANEL of 4 columns.  The 16 owners of tile column jt (one half-warp: threadIdx = 16 tj + ti)
    // factorize their 4 columns among themselves, getting the pivot and the pivot-row entries from the
    // diagonal-tile thread by shuffles, and publish the panel; every later tile column then takes a rank-4
    // update.  The owners of panel jt+1 take theirs first and factorize at once, while the other tile columns
    // are still being updated: the serial part of the next panel hides behind the update of this one.
    const unsigned lane = threadIdx.x & 31 ;
    auto factor_panel = [&] (int jt)
    {
        T *cb = colbuf + (jt & 1) * CB ;
        const unsigned mask = 0xFFFFu << (lane & 16) ;
        const int src = (int) (lane & 16) | jt ;               // the lane with ti == jt in this half-warp
        R rdv [4] ;
        #pragma unroll
        for (int jj = 0 ; jj < 4 ; jj++)
        {
            const R d = __shfl_sync (mask, real_ (a [jj][jj]), src) ;
            const R rd = rcp_ (d) ;
            rdv [jj] = rd ;
            #pragma unroll
            for (int cc = jj + 1 ; cc < 4 ; cc++)
            {
                const T f = conj_ (shfl_ (mask, a [cc][jj], src)) * rd ;       // conj (A (4jt+cc, j)) / d_j
                #pragma unroll
                for (int i = 0 ; i < 4 ; i++) a [i][cc] -= a [i][jj] * f ;
            }
        }
        #pragma unroll
        for (int q = 0 ; q < 4 ; q++)
        {
            #pragma unroll
            for (int i = 0 ; i < 4 ; i++) cb [q * POTRF_NB + 4 * ti + i] = a [i][q] ;
            if (ti == jt) { cb [4 * POTRF_NB + q] = a [q][q] ; cb [4 * POTRF_NB + 4 + q] = from_real<T> (rdv [q]) ; }
        }
    } ;
    if (tj == 0) factor_panel (0) ;
    for (int jt = 0 ; jt < njt ; jt++)
    {
        __syncthreads () ;                                      // panel jt is published
        const T *cb = colbuf + (jt & 1) * CB ;
        R rd [4] ;
        #pragma unroll
        for (int q = 0 ; q < 4 ; q++) rd [q] = real_ (cb [4 * POTRF_NB + 4 + q]) ;   // 1 / d_q, from the owners
        if (threadIdx.x == 0)
        {
            #pragma unroll
            for (int q = 0 ; q < 4 ; q++)
            {
                const R d = real_ (cb [4 * POTRF_NB + q]) ;
                dvec [4 * jt + q] = from_real<T> (d) ;
                if (4 * jt + q < jb && !(d > (R) 0)) atomicMin (minor, (unsigned long long) (b.col + 4 * jt + q)) ;
            }
        }
        if (ti >= tj && tj > jt)
        {
            #pragma unroll
            for (int q = 0 ; q < 4 ; q++)
            {
                T ar [4], ac [4] ;
                #pragma unroll
                for (int i = 0 ; i < 4 ; i++)
                {
                    ar [i] = cb [q * POTRF_NB + 4 * ti + i] ;
                    ac [i] = conj_ (cb [q * POTRF_NB + 4 * tj + i]) * rd [q] ;
                }
                #pragma unroll
                for (int jc = 0 ; jc < 4 ; jc++)
                    #pragma unroll
                    for (int i = 0 ; i < 4 ; i++) a [i][jc] -= ar [i] * ac [jc] ;
            }
        }
        // the other buffer held panel jt-1, which everybody has finished reading before the barrier above
        if (tj == jt + 1 && jt + 1 < njt) factor_panel (jt + 1) ;
    }
    __syncthreads () ;
    if (threadIdx.x < POTRF_NB)
    {
        const R q = (threadIdx.x < 4 * njt) ? sqrt (real_ (dvec [threadIdx.x])) : (R) 1 ;
        sqv [threadIdx.x] = from_real<T> (q) ;
        rsq [threadIdx.x] = from_real<T> (rcp_ (q)) ;
    }
    __syncthreads () ;
    #pragma unroll
    for (int jc = 0 ; jc < 4 ; jc++)
    {
        const int c = 4 * tj + jc ;
        const R ljj = real_ (sqv [c]), rl = real_ (rsq [c]) ;
        #pragma unroll
        for (int i = 0 ; i < 4 ; i++)
        {
            const int r = 4 * ti + i ;
            if (r < c) continue ;
            if (r < jb && c < jb)
            {
                const T v = (r == c) ? from_real<T> (ljj) : a [i][jc] * rl ;
                s [c * PB_LD + r] = v ;
                D [r + c * lda] = v ;
            }
            else s [c * PB_LD + r] = (r == c) ? one : zero ;        // identity padding up to 64 x 64
        }
    }
    if (b.slot < 0) return ;
    __syncthreads () ;
    inv64_lower<T> (s, w, rsq, rsq + POTRF_NB) ;      // tmp: [32][33] scratch after the three 64-vectors
    T *W = winv + (int64_t) b.slot * (WINV_LD * WINV_LD) ;
    for (int e = threadIdx.x ; e < nbk * nbk ; e += 256)
    {
        const int r = e % nbk, c = e / nbk ;
        T v ;
        if (r < jb && c < jb) v = (r >= c) ? w [r * PB_LD + c] : zero ;
        else v = (r == c) ? one : zero ;
        W [r + c * WINV_LD] = v ;
    }
}

// ---------------------------------------------------------------------------------------------
// rows below a factorized diagonal block through the explicit inverse:  X <- X * W',  W = L11^{-1}
// ---------------------------------------------------------------------------------------------

template <int NBK> constexpr size_t trsm_inv_smem_bytes ()
{
    return (size_t) NBK * ((TRSM_ROWS + 4) + (NBK + 4)) * sizeof (double) ;
}

// One CTA = 128 rows x NBK columns, 8 warps as 4 x 2, DMMA m8n8k4 with the same shared-memory layout
// as k_upd_dmma.  The whole X tile is in shared memory before the first store, so the product is done
// in place.  (W is lower triangular; its zero half is multiplied too -- K is only NBK.)
template <int NBK>
__global__ void __launch_bounds__ (256) k_trsm_inv (const TrsmTile *__restrict__ tiles, double *Lx,
    const double *__restrict__ winv)
{
    constexpr int LDX = TRSM_ROWS + 4, LDW = NBK + 4 ;
    constexpr int WN = NBK / 2, MT = 4, NTL = WN / 8 ;
    extern __shared__ __align__ (16) unsigned char smem_raw [] ;
    double *sX = reinterpret_cast<double *> (smem_raw) ;        // [NBK][LDX]
    double *sW = sX + NBK * LDX ;                               // [NBK][LDW]
    const TrsmTile t = tiles [blockIdx.x] ;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5 ;
    const int wm = warp & 3, wn = warp >> 2 ;
    const int jb = t.jb, nrows = t.nrows ;
    const int64_t lda = t.lda ;
    double *X = Lx + t.row_off ;
    const double *W = winv + (int64_t) t.slot * (WINV_LD * WINV_LD) ;
    #pragma unroll
    for (int e = tid ; e < NBK * TRSM_ROWS ; e += 256)
    {
        int m = e % TRSM_ROWS, k = e / TRSM_ROWS ;
        bool ok = (m < nrows) && (k < jb) ;
        cp_async8 (sX + k * LDX + m, ok ? (const void *) (X + m + (int64_t) k * lda) : (const void *) W, ok) ;
    }
    #pragma unroll
    for (int e = tid ; e < NBK * NBK ; e += 256)
    {
        int n = e % NBK, k = e / NBK ;
        cp_async8 (sW + k * LDW + n, W + n + k * WINV_LD, true) ;
    }
    cp_async_commit () ;
    cp_async_wait<0> () ;
    __syncthreads () ;
    double acc [MT][NTL][2] ;
    #pragma unroll
    for (int a = 0 ; a < MT ; a++)
        #pragma unroll
        for (int b = 0 ; b < NTL ; b++) { acc [a][b][0] = 0.0 ; acc [a][b][1] = 0.0 ; }
    const int fk = lane & 3, fm = lane >> 2 ;
    const double *cA = sX + wm * 32 + fm ;
    const double *cB = sW + wn * WN + fm ;
    #pragma unroll
    for (int kk = 0 ; kk < NBK ; kk += 4)
    {
        double fa [MT], fb [NTL] ;
        #pragma unroll
        for (int a = 0 ; a < MT ; a++) fa [a] = cA [(kk + fk) * LDX + a * 8] ;
        #pragma unroll
        for (int b = 0 ; b < NTL ; b++) fb [b] = cB [(kk + fk) * LDW + b * 8] ;
        #pragma unroll
        for (int a = 0 ; a < MT ; a++)
            #pragma unroll
            for (int b = 0 ; b < NTL ; b++)
                dmma884 (acc [a][b][0], acc [a][b][1], fa [a], fb [b]) ;
    }
    #pragma unroll
    for (int b = 0 ; b < NTL ; b++)
        #pragma unroll
        for (int e = 0 ; e < 2 ; e++)
        {
            const int jl = wn * WN + b * 8 + 2 * fk + e ;
            if (jl >= jb) continue ;
            #pragma unroll
            for (int a = 0 ; a < MT ; a++)
            {
                const int il = wm * 32 + a * 8 + fm ;
                if (il < nrows) X [il + (int64_t) jl * lda] = acc [a][b][e] ;
            }
        }
}

// complex double:  X <- X * W^H  (ztrsm "R","L","C","N" of SNW:1198 through the inverse), the same
// A * B^H product as k_upd_zmma with A = the X tile and B = W
template <int NBK> constexpr size_t trsm_zinv_smem_bytes ()
{
    return (size_t) NBK * ((TRSM_ROWS + 2) + (NBK + 2)) * sizeof (double2) ;
}

template <int NBK>
__global__ void __launch_bounds__ (256) k_trsm_zinv (const TrsmTile *__restrict__ tiles, zdouble *Lx,
    const zdouble *__restrict__ winv)
{
    constexpr int LDX = TRSM_ROWS + 2, LDW = NBK + 2 ;
    constexpr int WN = NBK / 2, MT = 4, NTL = WN / 8 ;
    extern __shared__ __align__ (16) unsigned char smem_raw [] ;
    double2 *sX = reinterpret_cast<double2 *> (smem_raw) ;      // [NBK][LDX]
    double2 *sW = sX + NBK * LDX ;                              // [NBK][LDW]
    const TrsmTile t = tiles [blockIdx.x] ;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5 ;
    const int wm = warp & 3, wn = warp >> 2 ;
    const int jb = t.jb, nrows = t.nrows ;
    const int64_t lda = t.lda ;
    zdouble *X = Lx + t.row_off ;
    const zdouble *W = winv + (int64_t) t.slot * (WINV_LD * WINV_LD) ;
    #pragma unroll
    for (int e = tid ; e < NBK * TRSM_ROWS ; e += 256)
    {
        int m = e % TRSM_ROWS, k = e / TRSM_ROWS ;
        bool ok = (m < nrows) && (k < jb) ;
        const void *src = ok ? (const void *) (X + m + (int64_t) k * lda) : (const void *) W ;
        unsigned d = (unsigned) __cvta_generic_to_shared (sX + k * LDX + m) ;
        asm volatile ("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" :: "r" (d), "l" (src), "r" (ok ? 16 : 0)) ;
    }
    #pragma unroll
    for (int e = tid ; e < NBK * NBK ; e += 256)
    {
        int n = e % NBK, k = e / NBK ;
        unsigned d = (unsigned) __cvta_generic_to_shared (sW + k * LDW + n) ;
        asm volatile ("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" :: "r" (d), "l" (W + n + k * WINV_LD), "r" (16)) ;
    }
    cp_async_commit () ;
    cp_async_wait<0> () ;
    __syncthreads () ;
    double cre [MT][NTL][2], cim [MT][NTL][2] ;
    #pragma unroll
    for (int a = 0 ; a < MT ; a++)
        #pragma unroll
        for (int b = 0 ; b < NTL ; b++) { cre [a][b][0] = 0.0 ; cre [a][b][1] = 0.0 ; cim [a][b][0] = 0.0 ; cim [a][b][1] = 0.0 ; }
    const int fk = lane & 3, fm = lane >> 2 ;
    zmma_stage<MT, NTL, NBK, LDX, LDW> (sX + wm * 32 + fm, sW + wn * WN + fm, fk, cre, cim) ;
    #pragma unroll
    for (int b = 0 ; b < NTL ; b++)
        #pragma unroll
        for (int e = 0 ; e < 2 ; e++)
        {
            const int jl = wn * WN + b * 8 + 2 * fk + e ;
            if (jl >= jb) continue ;
            #pragma unroll
            for (int a = 0 ; a < MT ; a++)
            {
                const int il = wm * 32 + a * 8 + fm ;
                if (il < nrows) X [il + (int64_t) jl * lda] = { cre [a][b][e], cim [a][b][e] } ;
            }
        }
}

// ---------------------------------------------------------------------------------------------
// triangular solve of the rows below a factorized diagonal block: one thread per row
// ---------------------------------------------------------------------------------------------

template <typename T, int NB>
__device__ __forceinline__ void trsm_row (const T *sL, T *X, int64_t lda, int jb)
{
    T x [NB] ;
    #pragma unroll
    for (int j = 0 ; j < NB ; j++) x [j] = (j < jb) ? X [j * lda] : from_real<T> (0) ;
    #pragma unroll
    for (int j = 0 ; j < NB ; j++)
    {
        // columns >= jb of sL hold the identity, so the padded part is a no-op
        T xj = x [j] / real_ (sL [j * PB_LD + j]) ;
        x [j] = xj ;
        #pragma unroll
        for (int k = j + 1 ; k < NB ; k++) x [k] -= mulc (xj, sL [j * PB_LD + k]) ;
    }
    #pragma unroll
    for (int j = 0 ; j < NB ; j++) if (j < jb) X [j * lda] = x [j] ;
}

template <typename T>
__global__ void __launch_bounds__ (TRSM_ROWS) k_trsm (const TrsmTile *__restrict__ tiles, T *Lx)
{
    // 16-byte entries: blocks wider than 16 columns always take the inverse path (k_trsm_zinv), and a
    // 64 x 65 tile of them would not fit the 48 KB of static shared memory
    constexpr int MAXNB = (sizeof (T) == 16) ? 32 : POTRF_NB ;
    __shared__ T sL [(MAXNB - 1) * PB_LD + MAXNB] ;
    const TrsmTile t = tiles [blockIdx.x] ;
    const int jb = t.jb ;
    const int64_t lda = t.lda ;
    const int nbp = (jb <= 8) ? 8 : (jb <= 16) ? 16 : (jb <= 32) ? 32 : 64 ;
    if (nbp > MAXNB) { __trap () ; }
    const T *D = Lx + t.diag_off ;
    for (int e = threadIdx.x ; e < nbp * nbp ; e += TRSM_ROWS)
    {
        int r = e % nbp, c = e / nbp ;
        T v ;
        if (r < jb && c < jb) v = (r >= c) ? D [r + c * lda] : from_real<T> (0) ;
        else v = from_real<T> ((r == c) ? 1 : 0) ;
        sL [c * PB_LD + r] = v ;
    }
    __syncthreads () ;
    if ((int) threadIdx.x >= t.nrows) return ;
    T *X = Lx + t.row_off + threadIdx.x ;
    if (nbp == 8) trsm_row<T, 8> (sL, X, lda, jb) ;
    else if (nbp == 16) trsm_row<T, 16> (sL, X, lda, jb) ;
    else if (nbp == 32) trsm_row<T, 32> (sL, X, lda, jb) ;
    else if constexpr (MAXNB >= 64) trsm_row<T, 64> (sL, X, lda, jb) ;
}

// ---------------------------------------------------------------------------------------------
// supernodal solve, one CTA per supernode of a level (forward: leaves first; backward: root first)
// ---------------------------------------------------------------------------------------------

// The right-hand sides are double (real factor) or complex double (complex factor) whatever the factor's
// own precision; the backward sweep uses L^H (SSOL:309-586, the "C" variants for complex).
template <typename T> struct solve_x { typedef double type ; } ;
template <typename R> struct solve_x<cx<R>> { typedef cx<double> type ; } ;

// Supernodes of at most SOLVE_SMALL columns: one CTA of 128 threads each.  The diagonal block goes to shared
// memory once (pitch mc+1, mc = the widest supernode of the launch), the triangular solve is done by ONE warp
// with the unknowns in registers (two per lane, the pivot broadcast by shuffle: no barrier, no memory latency
// inside the serial chain), the rows below by all threads with independent loads.
constexpr int SOLVE_SMALL = 64 ;

template <typename T> inline size_t solve_small_smem_bytes (int64_t mc)
{
    return (size_t) (mc * (mc + 1) + mc + 128) * sizeof (typename solve_x<T>::type) ;
}

// forward:  x1 <- L11 \ x1 ;  x [rows below] -= L21 * x1      (SSOL:20-303; atomics because siblings of
// one level update common ancestor rows concurrently)
template <typename T>
__global__ void __launch_bounds__ (128) k_solve_fwd (const int32_t *__restrict__ list,
    const int64_t *__restrict__ super, const int64_t *__restrict__ pi, const int64_t *__restrict__ px,
    const int32_t *__restrict__ Ls, const T *__restrict__ Lx, typename solve_x<T>::type *X, int64_t n, int nrhs,
    int mc)
{
    typedef typename solve_x<T>::type XT ;
    extern __shared__ __align__ (16) unsigned char solve_smem [] ;
    const int ld = mc + 1 ;
    XT *sL = reinterpret_cast<XT *> (solve_smem) ;          // sL [c * ld + r] = L (r,c), r >= c
    XT *sx = sL + mc * ld ;                                 // [mc]
    const int s = list [blockIdx.x] ;
    const int64_t k1 = super [s], psi = pi [s], nsrow = pi [s+1] - psi ;
    const int nscol = (int) (super [s+1] - k1) ;
    const T *L = Lx + px [s] ;
    const int tid = threadIdx.x, lane = tid & 31 ;
    const XT zero = from_real<XT> (0) ;
    for (int e = tid ; e < nscol * nscol ; e += 128)
    {
        const int r = e % nscol, c = e / nscol ;
        if (r >= c) sL [c * ld + r] = cvt<XT> (L [r + (int64_t) c * nsrow]) ;
    }
    for (int rhs = 0 ; rhs < nrhs ; rhs++)
    {
        XT *x = X + (int64_t) rhs * n ;
        if (tid < nscol) sx [tid] = x [k1 + tid] ;
        __syncthreads () ;
        if (tid < 32)
        {
            XT a0 = (lane < nscol) ? sx [lane] : zero, a1 = (lane + 32 < nscol) ? sx [lane + 32] : zero ;
            for (int j = 0 ; j < nscol ; j++)
            {
                const XT xj = shfl_ (0xffffffffu, (j < 32) ? a0 : a1, j & 31) / real_ (sL [j * ld + j]) ;
                if (lane == (j & 31)) { if (j < 32) a0 = xj ; else a1 = xj ; }
                if (lane > j && lane < nscol) a0 -= sL [j * ld + lane] * xj ;
                if (lane + 32 > j && lane + 32 < nscol) a1 -= sL [j * ld + lane + 32] * xj ;
            }
            if (lane < nscol) sx [lane] = a0 ;
            if (lane + 32 < nscol) sx [lane + 32] = a1 ;
        }
        __syncthreads () ;
        if (tid < nscol) x [k1 + tid] = sx [tid] ;
        for (int64_t i = nscol + tid ; i < nsrow ; i += 128)
        {
            XT acc = zero ;
            #pragma unroll 8
            for (int j = 0 ; j < nscol ; j++) acc += cvt<XT> (L [i + (int64_t) j * nsrow]) * sx [j] ;
            atomic_sub (x + Ls [psi + i], acc) ;
        }
        __syncthreads () ;
    }
}

// backward:  x1 <- L11^H \ (x1 - L21^H * x [rows below])        (SSOL:309-586)
template <typename T>
__global__ void __launch_bounds__ (128) k_solve_bwd (const int32_t *__restrict__ list,
    const int64_t *__restrict__ super, const int64_t *__restrict__ pi, const int64_t *__restrict__ px,
    const int32_t *__restrict__ Ls, const T *__restrict__ Lx, typename solve_x<T>::type *X, int64_t n, int nrhs,
    int mc)
{
    typedef typename solve_x<T>::type XT ;
    extern __shared__ __align__ (16) unsigned char solve_smem [] ;
    const int ld = mc + 1 ;
    XT *sL = reinterpret_cast<XT *> (solve_smem) ;
    XT *sx = sL + mc * ld ;                                 // [mc]
    XT *xr = sx + mc ;                                      // [128] the unknowns of a chunk of rows below
    const int s = list [blockIdx.x] ;
    const int64_t k1 = super [s], psi = pi [s], nsrow = pi [s+1] - psi ;
    const int nscol = (int) (super [s+1] - k1) ;
    const T *L = Lx + px [s] ;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5 ;
    const XT zero = from_real<XT> (0) ;
    for (int e = tid ; e < nscol * nscol ; e += 128)
    {
        const int r = e % nscol, c = e / nscol ;
        if (r >= c) sL [c * ld + r] = cvt<XT> (L [r + (int64_t) c * nsrow]) ;
    }
    for (int rhs = 0 ; rhs < nrhs ; rhs++)
    {
        XT *x = X + (int64_t) rhs * n ;
        if (tid < nscol) sx [tid] = x [k1 + tid] ;
        // sx [j] -= sum_{i >= nscol} conj (L (i,j)) * x [Ls [i]], 128 rows at a time, a warp per column
        for (int64_t i0 = nscol ; i0 < nsrow ; i0 += 128)
        {
            __syncthreads () ;
            xr [tid] = (i0 + tid < nsrow) ? x [Ls [psi + i0 + tid]] : zero ;
            __syncthreads () ;
            for (int j = warp ; j < nscol ; j += 4)
            {
                const T *Lj = L + i0 + (int64_t) j * nsrow ;
                XT acc = zero ;
                #pragma unroll
                for (int k = 0 ; k < 4 ; k++)
                {
                    const int i = lane + 32 * k ;
                    if (i0 + i < nsrow) acc += mulc (xr [i], cvt<XT> (Lj [i])) ;
                }
                for (int o = 16 ; o > 0 ; o >>= 1) acc += shfl_xor_ (0xffffffffu, acc, o) ;
                if (lane == 0) sx [j] -= acc ;              // column j belongs to this warp alone
            }
        }
        __syncthreads () ;
        if (tid < 32)
        {
            // L11^H x = sx: after x_j is known, row j of L takes it out of the earlier unknowns
            XT a0 = (lane < nscol) ? sx [lane] : zero, a1 = (lane + 32 < nscol) ? sx [lane + 32] : zero ;
            for (int j = nscol - 1 ; j >= 0 ; j--)
            {
                const XT xj = shfl_ (0xffffffffu, (j < 32) ? a0 : a1, j & 31) / real_ (sL [j * ld + j]) ;
                if (lane == (j & 31)) { if (j < 32) a0 = xj ; else a1 = xj ; }
                if (lane < j) a0 -= mulc (xj, sL [lane * ld + j]) ;
                if (lane + 32 < j) a1 -= mulc (xj, sL [(lane + 32) * ld + j]) ;
            }
            if (lane < nscol) x [k1 + lane] = a0 ;
            if (lane + 32 < nscol) x [k1 + lane + 32] = a1 ;
        }
        __syncthreads () ;
    }
}

// ---------------------------------------------------------------------------------------------
// wide supernodes (nscol > 256) in the solve: 64 columns at a time, the rows below them spread over
// many CTAs, instead of one CTA walking thousands of columns one by one
// ---------------------------------------------------------------------------------------------

// Diagonal block q of a wide supernode and one 256-row tile below it.  The 64 x 64 diagonal blocks are
// inverted once per factorization (k_solve_inv), so that a block step is ONE launch: forward, every tile
// CTA forms y = W * x_blk itself (4096 multiply-adds, W comes from L2) and subtracts L (rows, blk) * y from
// the rows it owns; backward, the tiles subtract L (rows, blk)^H * x [rows] from the block's entries and the
// CTA that finishes last multiplies by W^H.
struct SolveBlk  { int32_t s, c0, jb, pad ; } ;
struct SolveTile { int32_t s, c0, jb, r0, nrows, writer, ntiles, blk ; } ;   // tile `writer` stores y; ntiles per block

constexpr int SOLVE_WLD = 64 ;      // an inverse is a 64 x 64 column-major array, lower triangular, identity padded

template <typename T> constexpr size_t solve_inv_smem_bytes ()
{
    return (size_t) (2 * POTRF_NB * PB_LD + POTRF_NB + 32 * 33) * sizeof (typename solve_x<T>::type) ;
}

// W = (diagonal block)^{-1}, in the precision of the solve
template <typename T>
__global__ void __launch_bounds__ (256) k_solve_inv (const SolveBlk *__restrict__ blks,
    const int64_t *__restrict__ pi, const int64_t *__restrict__ px, const T *__restrict__ Lx,
    typename solve_x<T>::type *__restrict__ Winv)
{
    typedef typename solve_x<T>::type XT ;
    extern __shared__ __align__ (16) unsigned char solve_smem [] ;
    XT *s = reinterpret_cast<XT *> (solve_smem) ;       // s [c * PB_LD + r] = L (r,c)
    XT *w = s + POTRF_NB * PB_LD ;
    XT *rsq = w + POTRF_NB * PB_LD ;
    XT *tmp = rsq + POTRF_NB ;
    const SolveBlk b = blks [blockIdx.x] ;
    const int64_t nsrow = pi [b.s+1] - pi [b.s] ;
    const T *L = Lx + px [b.s] + b.c0 + (int64_t) b.c0 * nsrow ;
    const int jb = b.jb ;
    const XT one = from_real<XT> (1), zero = from_real<XT> (0) ;
    for (int e = threadIdx.x ; e < POTRF_NB * POTRF_NB ; e += 256)
    {
        const int r = e % POTRF_NB, c = e / POTRF_NB ;
        XT v ;
        if (r < jb && c < jb) v = (r >= c) ? cvt<XT> (L [r + (int64_t) c * nsrow]) : zero ;
        else v = (r == c) ? one : zero ;
        s [c * PB_LD + r] = v ;
    }
    __syncthreads () ;
    if (threadIdx.x < POTRF_NB) rsq [threadIdx.x] = from_real<XT> (1.0 / real_ (s [threadIdx.x * PB_LD + threadIdx.x])) ;
    __syncthreads () ;
    inv64_lower<XT> (s, w, rsq, tmp) ;
    XT *W = Winv + (int64_t) blockIdx.x * (SOLVE_WLD * SOLVE_WLD) ;
    for (int e = threadIdx.x ; e < POTRF_NB * POTRF_NB ; e += 256)
    {
        const int r = e % POTRF_NB, c = e / POTRF_NB ;
        W [r + c * SOLVE_WLD] = (r >= c) ? w [r * PB_LD + c] : zero ;
    }
}

__device__ __forceinline__ double ldcg_ (const double *p) { return __ldcg (p) ; }
__device__ __forceinline__ cx<double> ldcg_ (const cx<double> *p)
{
    double2 v = __ldcg (reinterpret_cast<const double2 *> (p)) ;
    return { v.x, v.y } ;
}

// forward block step:  y = L_blk \ x_blk  (dtrsv / dtrsm of SSOL:93, :204);  x [rows] -= L (rows, blk) * y  (the dgemv /
// dgemm of SSOL:100, :212).  Rows inside the supernode are its own later unknowns, rows below it belong to
// ancestors and are shared with the siblings of the level, hence the atomics.  y goes to Y (not back into X: the
// other tiles of the launch are still reading x_blk).
template <typename T>
__global__ void __launch_bounds__ (256) k_solve_fwd_blk (const SolveTile *__restrict__ tiles,
    const int64_t *__restrict__ super, const int64_t *__restrict__ pi, const int64_t *__restrict__ px,
    const int32_t *__restrict__ Ls, const T *__restrict__ Lx, const typename solve_x<T>::type *__restrict__ Winv,
    typename solve_x<T>::type *X, typename solve_x<T>::type *Y, int64_t n, int nrhs)
{
    typedef typename solve_x<T>::type XT ;
    __shared__ XT xb [64], yb [64] ;
    const SolveTile t = tiles [blockIdx.x] ;
    const int64_t k1 = super [t.s], nscol = super [t.s+1] - k1, psi = pi [t.s], nsrow = pi [t.s+1] - psi ;
    const T *L = Lx + px [t.s] + (int64_t) t.c0 * nsrow ;
    const XT *W = Winv + (int64_t) t.blk * (SOLVE_WLD * SOLVE_WLD) ;
    const int tid = threadIdx.x, row = t.r0 + tid ;
    const XT zero = from_real<XT> (0) ;
    for (int rhs = 0 ; rhs < nrhs ; rhs++)
    {
        XT *x = X + (int64_t) rhs * n, *y = Y + (int64_t) rhs * n ;
        if (tid < 64) xb [tid] = (tid < t.jb) ? x [k1 + t.c0 + tid] : zero ;
        __syncthreads () ;
        {
            // all 16 loads of a thread are issued before the first is used
            const int r = tid >> 2, part = tid & 3 ;
            XT wv [16] ;
            #pragma unroll
            for (int k = 0 ; k < 16 ; k++) wv [k] = (part + 4 * k <= r) ? W [r + (part + 4 * k) * SOLVE_WLD] : zero ;
            XT acc = zero ;
            #pragma unroll
            for (int k = 0 ; k < 16 ; k++) acc += wv [k] * xb [part + 4 * k] ;
            acc += shfl_xor_ (0xffffffffu, acc, 1) ;
            acc += shfl_xor_ (0xffffffffu, acc, 2) ;
            if (part == 0) yb [r] = acc ;
        }
        __syncthreads () ;
        if (t.writer && tid < t.jb) y [k1 + t.c0 + tid] = yb [tid] ;
        if (tid < t.nrows)
        {
            XT acc = zero ;
            if (t.jb == 64)
            {
                #pragma unroll 16
                for (int j = 0 ; j < 64 ; j++) acc += cvt<XT> (L [row + (int64_t) j * nsrow]) * yb [j] ;
            }
            else for (int j = 0 ; j < t.jb ; j++) acc += cvt<XT> (L [row + (int64_t) j * nsrow]) * yb [j] ;
            const int64_t dst = (row < nscol) ? k1 + row : (int64_t) Ls [psi + row] ;
            atomic_sub (x + dst, acc) ;
        }
        __syncthreads () ;
    }
}

// backward block step:  Y_blk -= L (rows, blk)^H * x [rows]  (SSOL:386, dgemv "C") by every tile, then the tile that
// finishes last does  x_blk = L_blk^H \ Y_blk = W^H * Y_blk  (SSOL:386-402).  counters [blk * nrhs + rhs] counts the
// finished tiles and is left at zero for the next solve.
template <typename T>
__global__ void __launch_bounds__ (256) k_solve_bwd_blk (const SolveTile *__restrict__ tiles,
    const int64_t *__restrict__ super, const int64_t *__restrict__ pi, const int64_t *__restrict__ px,
    const int32_t *__restrict__ Ls, const T *__restrict__ Lx, const typename solve_x<T>::type *__restrict__ Winv,
    typename solve_x<T>::type *X, typename solve_x<T>::type *Y, int *counters, int64_t n, int nrhs)
{
    typedef typename solve_x<T>::type XT ;
    __shared__ XT xr [256], yb [64] ;
    __shared__ int is_last ;
    const SolveTile t = tiles [blockIdx.x] ;
    const int64_t k1 = super [t.s], nscol = super [t.s+1] - k1, psi = pi [t.s], nsrow = pi [t.s+1] - psi ;
    const int tid = threadIdx.x, j = tid >> 2, part = tid & 3 ;
    const T *Lbase = Lx + px [t.s] + (int64_t) t.c0 * nsrow + t.r0 ;     // first row of the tile, first column of the block
    const XT *W = Winv + (int64_t) t.blk * (SOLVE_WLD * SOLVE_WLD) ;
    const XT zero = from_real<XT> (0) ;
    for (int rhs = 0 ; rhs < nrhs ; rhs++)
    {
        XT *x = X + (int64_t) rhs * n, *y = Y + (int64_t) rhs * n ;
        if (tid < t.nrows)
        {
            const int row = t.r0 + tid ;
            xr [tid] = x [(row < nscol) ? k1 + row : (int64_t) Ls [psi + row]] ;
        }
        __syncthreads () ;
        // warp w takes columns 8w .. 8w+7; its lanes run down the 256 rows of the tile, 32 at a time (coalesced,
        // 8 independent loads per column), then the warp adds up
        if (t.nrows > 0)
        {
            const int lane = tid & 31, warp = tid >> 5 ;
            for (int jj = 0 ; jj < 8 ; jj++)
            {
                const int jc = warp * 8 + jj ;
                if (jc >= t.jb) break ;
                const T *Lj = Lbase + (int64_t) jc * nsrow ;
                XT acc = zero ;
                #pragma unroll
                for (int k = 0 ; k < 8 ; k++)
                {
                    const int r = lane + 32 * k ;
                    if (r < t.nrows) acc += mulc (xr [r], cvt<XT> (Lj [r])) ;
                }
                for (int o = 16 ; o > 0 ; o >>= 1) acc += shfl_xor_ (0xffffffffu, acc, o) ;
                if (lane == 0) atomic_sub (y + k1 + t.c0 + jc, acc) ;
            }
        }
        __threadfence () ;
        __syncthreads () ;
        if (tid == 0)
        {
            int *cnt = counters + (int64_t) t.blk * nrhs + rhs ;
            const int done = atomicAdd (cnt, 1) ;
            is_last = (done == t.ntiles - 1) ;
            if (is_last) *cnt = 0 ;
        }
        __syncthreads () ;
        if (is_last)
        {
            __threadfence () ;
            if (tid < 64) yb [tid] = (tid < t.jb) ? ldcg_ (y + k1 + t.c0 + tid) : zero ;
            __syncthreads () ;
            // x_c = sum_{r >= c} conj (W (r,c)) * y_r : thread (c, part) walks down column c of W
            XT wv [16] ;
            #pragma unroll
            for (int k = 0 ; k < 16 ; k++) wv [k] = (part + 4 * k >= j) ? W [part + 4 * k + j * SOLVE_WLD] : zero ;
            XT a2 = zero ;
            #pragma unroll
            for (int k = 0 ; k < 16 ; k++) a2 += mulc (yb [part + 4 * k], wv [k]) ;
            a2 += shfl_xor_ (0xffffffffu, a2, 1) ;
            a2 += shfl_xor_ (0xffffffffu, a2, 2) ;
            if (part == 0 && j < t.jb) x [k1 + t.c0 + j] = a2 ;
        }
        __syncthreads () ;
    }
}

// ---------------------------------------------------------------------------------------------
// fan-in of the contribution blocks of the top separators (several GPUs)
// ---------------------------------------------------------------------------------------------

// arena base address of every rank, as seen from this device: peer-mapped (cudaDeviceEnablePeerAccess within
// one process, cudaIpcOpenMemHandle across processes), so a plain load travels over NVLink
struct PeerPtrs { void *p [32] ; } ;

__device__ __forceinline__ float ldcg_ (const float *p) { return __ldcg (p) ; }
__device__ __forceinline__ cx<float> ldcg_ (const cx<float> *p)
{
    float2 v = __ldcg (reinterpret_cast<const float2 *> (p)) ;
    return { v.x, v.y } ;
}

// One CTA per (block, column): dst (i, col) += sum over the ranks in mask of their copy of the same entry, for
// the rows i >= col of the block (its lower part).  The top separators sit at the same arena offset on every
// rank.  Loads bypass L1 (the peers rewrite this memory in every factorization); a thread's loads from the
// different peers are independent, consecutive threads read consecutive addresses.
template <typename T>
__global__ void __launch_bounds__ (256) k_pull_add (const PullSeg *__restrict__ segs, T *Lx, PeerPtrs peers)
{
    const PullSeg g = segs [blockIdx.x] ;
    const int col = blockIdx.y ;
    if (col >= g.ncols) return ;
    const int64_t base = g.off + (int64_t) col * g.ld ;
    T *dst = Lx + base ;
    for (int i = col + (int) threadIdx.x ; i < g.nrows ; i += 256)
    {
        T acc = dst [i] ;
        for (uint32_t m = g.mask ; m ; m &= m - 1)
        {
            const int r = __ffs ((int) m) - 1 ;
            acc += ldcg_ (reinterpret_cast<const T *> (peers.p [r]) + base + i) ;
        }
        dst [i] = acc ;
    }
}

// ---------------------------------------------------------------------------------------------
// wide supernodes in the solve, 256 columns (4 diagonal blocks) per launch  (CHOLMOD_B200_SOLVE256)
// ---------------------------------------------------------------------------------------------

// One tile of the rows below a 256-column "superstep" [c0, c0+nb) of a wide supernode (64 rows in the forward
// sweep, 256 in the backward sweep).  blk0 = index of the superstep's first 64 x 64 inverse in Winv (the blocks of a
// supernode are consecutive); step = the superstep's number (its counters); ntiles = tiles of the superstep;
// next_nb > 0 (forward) marks a tile that holds rows of the NEXT superstep's diagonal block: the last of those
// tiles to finish solves that block at once, so a superstep costs one launch.
struct SolveTile256 { int32_t s, c0, nb, r0, nrows, ntiles, blk0, step, next_nb, pad ; } ;

// Programmatic dependent launch: a superstep kernel launched with the programmatic-stream-serialization attribute is
// scheduled while its predecessor drains; it may read only what no kernel of the solve writes (the schedule, the
// factor, the inverses) before solve_pdl_wait (), which returns once the predecessor has completed and its writes
// are visible, and then lets its own successor in.  Without the attribute both instructions do nothing.
__device__ __forceinline__ void solve_pdl_wait ()
{
    asm volatile ("griddepcontrol.wait;" ::: "memory") ;
    asm volatile ("griddepcontrol.launch_dependents;") ;
}

// Diagnostic build (-DB200_SOLVE_TRACE): SM-clock stamps at the phase boundaries of the CTA that ends up doing the
// diagonal solve, summed over the launches ([0..15] forward, [16..31] backward; slot 15 / 31 counts the launches).
#ifdef B200_SOLVE_TRACE
__device__ unsigned long long g_solve_trace [32] ;
#define SOLVE_TR(k) do { if (threadIdx.x == 0) tr [k] = clock64 () ; } while (0)
#define SOLVE_TR_DECL long long tr [10] = { 0, 0, 0, 0, 0, 0, 0, 0, 0, 0 }
#define SOLVE_TR_FLUSH(base, cnt) do { if (threadIdx.x == 0) { for (int k_ = 1 ; k_ < (cnt) ; k_++) \
    atomicAdd (&g_solve_trace [(base) + k_ - 1], (unsigned long long) (tr [k_] - tr [k_-1])) ; \
    atomicAdd (&g_solve_trace [(base) + 15], 1ull) ; } } while (0)
#else
#define SOLVE_TR(k) do { } while (0)
#define SOLVE_TR_DECL do { } while (0)
#define SOLVE_TR_FLUSH(base, cnt) do { } while (0)
#endif

// Ask L2 for what the diagonal solve of the superstep [c0, c0+nb) will read -- its 64 x 64 inverses and the lower
// triangle of its diagonal block -- long before the CTA that does the solve is known: one CTA of 256 threads, a few
// lines per thread.  A hint only; every address lies inside Winv or inside the supernode's columns.
template <typename T, typename XT>
__device__ __forceinline__ void solve_diag256_prefetch (const T *__restrict__ Lsn, int64_t nsrow, int c0, int nb,
    const XT *__restrict__ W0)
{
    const int tid = threadIdx.x ;
    const char *wp = reinterpret_cast<const char *> (W0) ;
    const int wbytes = ((nb + 63) >> 6) * SOLVE_WLD * SOLVE_WLD * (int) sizeof (XT) ;
    for (int off = tid * 128 ; off < wbytes ; off += 256 * 128)
        asm volatile ("prefetch.global.L2 [%0];" :: "l" (wp + off)) ;
    constexpr int PER = 128 / (int) sizeof (T) ;            // entries per 128 bytes
    const int chunks = (nb + PER - 1) / PER ;
    for (int e = tid ; e < nb * chunks ; e += 256)
    {
        const int j = e / chunks, r = (e % chunks) * PER ;
        if (r + PER > j)                                    // the chunk reaches below the diagonal
        {
            const T *q = Lsn + (c0 + max (r, j)) + (int64_t) (c0 + j) * nsrow ;
            asm volatile ("prefetch.global.L2 [%0];" :: "l" (q)) ;
        }
    }
}

// the same for a whole-block inverse: the lower triangle of W256 below nb
template <typename XT>
__device__ __forceinline__ void solve_w256_prefetch (const XT *__restrict__ W, int nb)
{
    constexpr int PER = 128 / (int) sizeof (XT) ;
    const int chunks = (nb + PER - 1) / PER ;
    for (int e = threadIdx.x ; e < nb * chunks ; e += 256)
    {
        const int c = e / chunks, r = (e % chunks) * PER ;
        if (r + PER > c) asm volatile ("prefetch.global.L2 [%0];" :: "l" (W + max (r, c) + (int64_t) c * 256)) ;
    }
}

// Forward diagonal solve of a superstep by a CTA of 256 threads: xs [0..nb) (shared) holds the right-hand side and
// is destroyed; ys [0..nb) receives y = L_blk \ xs.  Four sub-steps of 64: y_q = W_q * x_q with the pre-inverted
// diagonal block (4 threads per row), then the later unknowns take x -= L (rows, block q) * y_q, 4 threads per row
// with 16 independent loads each.
template <typename T, typename XT, int PRE_MAX>
__device__ __forceinline__ void solve_diag256_fwd (const T *__restrict__ Lsn, int64_t nsrow, int c0, int nb,
    const XT *__restrict__ W0, XT *xs, XT *ys)
{
    const int tid = threadIdx.x, r4 = tid >> 2, part = tid & 3 ;
    const XT zero = from_real<XT> (0) ;
    for (int q0 = 0 ; q0 < nb ; q0 += 64)
    {
        const int jb = min (64, nb - q0) ;
        const XT *W = W0 + (int64_t) (q0 >> 6) * (SOLVE_WLD * SOLVE_WLD) ;
        // the entries of L that the update below needs do not depend on y_q: ask for them before the product with
        // W_q, all of them at once (up to three later 64-row groups, 16 columns each per thread)
        constexpr int PRE = (sizeof (T) <= 8) ? PRE_MAX : 1 ;   // 16-byte entries: one group ahead, for the registers' sake
        T lv [PRE][16] ;
        #pragma unroll
        for (int g = 0 ; g < PRE ; g++)
        {
            const int i = q0 + 64 * (g + 1) + r4 ;
            const T *Lr = Lsn + (c0 + i) + (int64_t) (c0 + q0 + 16 * part) * nsrow ;
            #pragma unroll
            for (int j = 0 ; j < 16 ; j++) lv [g][j] = (i < nb) ? Lr [(int64_t) j * nsrow] : T () ;
        }
        {
            XT wv [16] ;
            #pragma unroll
            for (int k = 0 ; k < 16 ; k++) wv [k] = (part + 4 * k <= r4) ? W [r4 + (part + 4 * k) * SOLVE_WLD] : zero ;
            XT acc = zero ;
            #pragma unroll
            for (int k = 0 ; k < 16 ; k++) acc += wv [k] * xs [q0 + part + 4 * k] ;
            acc += shfl_xor_ (0xffffffffu, acc, 1) ;
            acc += shfl_xor_ (0xffffffffu, acc, 2) ;
            if (part == 0 && r4 < jb) ys [q0 + r4] = acc ;
        }
        __syncthreads () ;
        // rows below block q exist only when the block is a full one (jb == 64)
        #pragma unroll
        for (int g = 0 ; g < 3 ; g++)
        {
            const int i = q0 + 64 * (g + 1) + r4 ;
            XT acc = zero ;
            if (g < PRE)
            {
                #pragma unroll
                for (int j = 0 ; j < 16 ; j++) acc += cvt<XT> (lv [g < PRE ? g : 0][j]) * ys [q0 + 16 * part + j] ;
            }
            else if (i < nb)
            {
                const T *Lr = Lsn + (c0 + i) + (int64_t) (c0 + q0 + 16 * part) * nsrow ;
                #pragma unroll
                for (int j = 0 ; j < 16 ; j++) acc += cvt<XT> (Lr [(int64_t) j * nsrow]) * ys [q0 + 16 * part + j] ;
            }
            acc += shfl_xor_ (0xffffffffu, acc, 1) ;
            acc += shfl_xor_ (0xffffffffu, acc, 2) ;
            if (part == 0 && i < nb) xs [i] -= acc ;
        }
        __syncthreads () ;
    }
}

// v [u] (u < 16) summed over the 32 lanes of the warp, 31 shuffles: on return lane l holds the sum of v [l & 15] in
// v [0].  The first step adds the two half-warps; every later step keeps one half of the remaining values and hands
// the other half to the partner lane, which keeps those.
template <typename XT, int H>
__device__ __forceinline__ void warp_sum_step (XT (&v) [16], int lane)
{
    const bool up = (lane & H) != 0 ;
    #pragma unroll
    for (int u = 0 ; u < H ; u++)
    {
        const XT keep = up ? v [u + H] : v [u], send = up ? v [u] : v [u + H] ;
        v [u] = keep + shfl_xor_ (0xffffffffu, send, H) ;
    }
}

template <typename XT>
__device__ __forceinline__ void warp_sum16 (XT (&v) [16], int lane)
{
    #pragma unroll
    for (int u = 0 ; u < 16 ; u++) v [u] += shfl_xor_ (0xffffffffu, v [u], 16) ;
    warp_sum_step<XT, 8> (v, lane) ;
    warp_sum_step<XT, 4> (v, lane) ;
    warp_sum_step<XT, 2> (v, lane) ;
    warp_sum_step<XT, 1> (v, lane) ;
}

// ---- the whole 256 x 256 diagonal block of a superstep inverted (once per factorization) -----------------------
// The serial part of a superstep is its diagonal solve; with the four 64 x 64 inverses it is four dependent
// sub-steps of two memory round trips each, all through one SM (measured: 27k of the 50k cycles of a forward
// launch).  With W256 = (diagonal block)^{-1} it is one triangular matrix-vector product with every load
// independent.  W256 of superstep `step` is a 256 x 256 column-major array at W256 + step * 65536; only its lower
// triangle below nb is written and read.

constexpr int SOLVE_W256_LD = 256 ;
template <typename T> constexpr size_t solve_inv256_smem_bytes ()
{
    return (size_t) (2 * 64 * 64) * sizeof (typename solve_x<T>::type) ;
}

// C (64 x 64) += A * B with A in As [k * 64 + row], B in Bt [k * 64 + col]; thread (ty, tx) holds C (4 ty + a, 4 tx + b)
template <typename XT>
__device__ __forceinline__ void mm64_acc (const XT *As, const XT *Bt, int ty, int tx, XT (&acc) [4][4])
{
    #pragma unroll 4
    for (int k = 0 ; k < 64 ; k++)
    {
        XT a [4], b [4] ;
        #pragma unroll
        for (int u = 0 ; u < 4 ; u++) { a [u] = As [k * 64 + 4 * ty + u] ; b [u] = Bt [k * 64 + 4 * tx + u] ; }
        #pragma unroll
        for (int u = 0 ; u < 4 ; u++)
            #pragma unroll
            for (int v = 0 ; v < 4 ; v++) acc [u][v] += a [u] * b [v] ;
    }
}

// One CTA per superstep (steps [blockIdx.x] = { s, c0, nb, first 64 x 64 inverse }): block (i,j) of the inverse, i > j,
// is  - W_ii * sum_{k = j .. i-1} L_ik * W_kj  (blocks of 64; W_ii are the inverses k_solve_inv left in Winv).
template <typename T>
__global__ void __launch_bounds__ (256) k_solve_inv256 (const SolveBlk *__restrict__ steps,
    const int64_t *__restrict__ pi, const int64_t *__restrict__ px, const T *__restrict__ Lx,
    const typename solve_x<T>::type *__restrict__ Winv, typename solve_x<T>::type *W256all)
{
    typedef typename solve_x<T>::type XT ;
    extern __shared__ __align__ (16) unsigned char solve_smem [] ;
    XT *As = reinterpret_cast<XT *> (solve_smem) ;
    XT *Bt = As + 64 * 64 ;
    const SolveBlk b = steps [blockIdx.x] ;
    const int64_t nsrow = pi [b.s+1] - pi [b.s] ;
    const T *Lsn = Lx + px [b.s] ;
    const int c0 = b.c0, nb = b.jb, nq = (nb + 63) >> 6 ;
    const XT *W64 = Winv + (int64_t) b.pad * (SOLVE_WLD * SOLVE_WLD) ;
    XT *W = W256all + (int64_t) blockIdx.x * (SOLVE_W256_LD * SOLVE_W256_LD) ;
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15 ;
    const XT zero = from_real<XT> (0) ;
    // the diagonal blocks: copies of the 64 x 64 inverses (lower triangles)
    for (int q = 0 ; q < nq ; q++)
        for (int e = tid ; e < 64 * 64 ; e += 256)
        {
            const int r = e & 63, c = e >> 6 ;
            if (r >= c && 64 * q + r < nb) W [(64 * q + r) + (int64_t) (64 * q + c) * SOLVE_W256_LD] = W64 [(int64_t) q * (SOLVE_WLD * SOLVE_WLD) + r + c * SOLVE_WLD] ;
        }
    __syncthreads () ;
    for (int j = 0 ; j + 1 < nq ; j++)
        for (int i = j + 1 ; i < nq ; i++)
        {
            const int ri = min (64, nb - 64 * i) ;              // rows of block row i (block column j < i is a full one)
            XT acc [4][4] ;
            #pragma unroll
            for (int u = 0 ; u < 4 ; u++)
                #pragma unroll
                for (int v = 0 ; v < 4 ; v++) acc [u][v] = zero ;
            for (int k = j ; k < i ; k++)
            {
                // A = L (block i, block k), B = W (block k, block j): lower triangular when k == j
                for (int e = tid ; e < 64 * 64 ; e += 256)
                {
                    const int r = e & 63, c = e >> 6 ;
                    As [c * 64 + r] = (r < ri) ? cvt<XT> (Lsn [(c0 + 64 * i + r) + (int64_t) (c0 + 64 * k + c) * nsrow]) : zero ;
                    Bt [r * 64 + c] = (k > j || r >= c) ? W [(64 * k + r) + (int64_t) (64 * j + c) * SOLVE_W256_LD] : zero ;
                }
                __syncthreads () ;
                mm64_acc<XT> (As, Bt, ty, tx, acc) ;
                __syncthreads () ;
            }
            // W (i,j) = - W_ii * S
            #pragma unroll
            for (int u = 0 ; u < 4 ; u++)
                #pragma unroll
                for (int v = 0 ; v < 4 ; v++) { Bt [(4 * ty + u) * 64 + 4 * tx + v] = acc [u][v] ; acc [u][v] = zero ; }
            for (int e = tid ; e < 64 * 64 ; e += 256)
            {
                const int r = e & 63, c = e >> 6 ;
                As [c * 64 + r] = W64 [(int64_t) i * (SOLVE_WLD * SOLVE_WLD) + r + c * SOLVE_WLD] ;     // zero above the diagonal
            }
            __syncthreads () ;
            mm64_acc<XT> (As, Bt, ty, tx, acc) ;
            #pragma unroll
            for (int u = 0 ; u < 4 ; u++)
                #pragma unroll
                for (int v = 0 ; v < 4 ; v++)
                    if (4 * ty + u < ri) { XT neg = zero ; neg -= acc [u][v] ; W [(64 * i + 4 * ty + u) + (int64_t) (64 * j + 4 * tx + v) * SOLVE_W256_LD] = neg ; }
            __syncthreads () ;
        }
}

// ys [0..nb) = W256 * xs [0..nb) (lower triangular), by a CTA of 256 threads.  Thread (h, p) takes the rows p and
// 255 - p (together 257 entries, so the threads are balanced) and the columns of parity h; every load is
// independent, consecutive threads read consecutive rows.  red: 512 entries of scratch.
template <typename XT>
__device__ __forceinline__ void solve_diag256w_fwd (const XT *__restrict__ W, int nb, const XT *xs, XT *ys, XT *red)
{
    const int tid = threadIdx.x, h = tid >> 7, p = tid & 127, ra = p, rb = 255 - p ;
    const XT zero = from_real<XT> (0) ;
    XT acca = zero, accb = zero ;
    const XT *Wa = W + ra, *Wb = W + rb ;
    const int la = (ra < nb) ? ra : -1, lb = (rb < nb) ? rb : -1 ;     // last column of each row (none if the row is not there)
    #pragma unroll 2
    for (int c0 = h ; c0 <= lb || c0 <= la ; c0 += 16)
    {
        XT va [8], vb [8] ;
        #pragma unroll
        for (int u = 0 ; u < 8 ; u++)
        {
            const int c = c0 + 2 * u ;
            va [u] = (c <= la) ? Wa [(int64_t) c * SOLVE_W256_LD] : zero ;
            vb [u] = (c <= lb) ? Wb [(int64_t) c * SOLVE_W256_LD] : zero ;
        }
        #pragma unroll
        for (int u = 0 ; u < 8 ; u++)
        {
            const int c = c0 + 2 * u ;
            const XT xv = (c < 256) ? xs [c] : zero ;
            acca += va [u] * xv ; accb += vb [u] * xv ;
        }
    }
    red [h * 256 + ra] = acca ; red [h * 256 + rb] = accb ;
    __syncthreads () ;
    ys [tid] = red [tid] + red [256 + tid] ;
    __syncthreads () ;
}

// xo [c] = sum_{r >= c} conj (W256 (r,c)) * yb [r] for c < nb, by a CTA of 256 threads: warp w takes the columns
// 16 w + u and 255 - (16 w + u), u < 16 (balanced), its lanes run down the rows (coalesced), warp_sum16 adds up.
// Lane l < 16 of warp w returns column 16 w + l in x0 and column 255 - (16 w + l) in x1.
template <typename XT>
__device__ __forceinline__ void solve_diag256w_bwd (const XT *__restrict__ W, int nb, const XT *yb, XT &x0, XT &x1)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5 ;
    const XT zero = from_real<XT> (0) ;
    #pragma unroll
    for (int pass = 0 ; pass < 2 ; pass++)
    {
        XT v [16] ;
        #pragma unroll
        for (int u = 0 ; u < 16 ; u++) v [u] = zero ;
        const int cfirst = pass ? 255 - (16 * warp + 15) : 16 * warp ;      // the smallest column of the pass
        if (cfirst < nb)
            #pragma unroll 2
            for (int k = cfirst >> 5 ; 32 * k < nb ; k++)
            {
                const int r = lane + 32 * k ;
                const XT yv = (r < nb) ? yb [r] : zero ;
                XT w [16] ;
                #pragma unroll
                for (int u = 0 ; u < 16 ; u++)
                {
                    const int c = pass ? 255 - (16 * warp + u) : 16 * warp + u ;
                    w [u] = (r >= c && r < nb) ? W [r + (int64_t) c * SOLVE_W256_LD] : zero ;
                }
                #pragma unroll
                for (int u = 0 ; u < 16 ; u++) v [u] += mulc (yv, w [u]) ;
            }
        warp_sum16<XT> (v, lane) ;
        if (pass) x1 = v [0] ; else x0 = v [0] ;
    }
}

// First superstep of every wide supernode of a level: y = L_blk \ x_blk, into Y.
template <typename T>
__global__ void __launch_bounds__ (256) k_solve_diag_fwd256 (const SolveTile256 *__restrict__ firsts,
    const int64_t *__restrict__ super, const int64_t *__restrict__ pi, const int64_t *__restrict__ px,
    const T *__restrict__ Lx, const typename solve_x<T>::type *__restrict__ Winv,
    const typename solve_x<T>::type *__restrict__ W256,
    const typename solve_x<T>::type *X, typename solve_x<T>::type *Y, int64_t n, int nrhs)
{
    typedef typename solve_x<T>::type XT ;
    __shared__ XT xs [256], ys [256], red [512] ;
    const SolveTile256 t = firsts [blockIdx.x] ;
    const int64_t k1 = super [t.s], nsrow = pi [t.s+1] - pi [t.s] ;
    const T *Lsn = Lx + px [t.s] ;
    const int tid = threadIdx.x ;
    if (W256) solve_w256_prefetch<XT> (W256 + (int64_t) t.step * (SOLVE_W256_LD * SOLVE_W256_LD), t.nb) ;
    solve_pdl_wait () ;
    for (int rhs = 0 ; rhs < nrhs ; rhs++)
    {
        xs [tid] = (tid < t.nb) ? X [(int64_t) rhs * n + k1 + t.c0 + tid] : from_real<XT> (0) ;
        __syncthreads () ;
        if (W256) solve_diag256w_fwd<XT> (W256 + (int64_t) t.step * (SOLVE_W256_LD * SOLVE_W256_LD), t.nb, xs, ys, red) ;
        else solve_diag256_fwd<T, XT, 3> (Lsn, nsrow, t.c0, t.nb, Winv + (int64_t) t.blk0 * (SOLVE_WLD * SOLVE_WLD), xs, ys) ;
        if (tid < t.nb) Y [(int64_t) rhs * n + k1 + t.c0 + tid] = ys [tid] ;
        __syncthreads () ;
    }
}

// Forward superstep, one CTA per 64 rows below it: x [rows] -= L (rows, superstep) * y, four threads per row (one per
// 64-column block of the superstep, 16 independent loads per batch).  Rows inside the supernode's own columns are
// touched by this CTA alone; rows below it belong to ancestors and are shared with the siblings of the level:
// atomics.  The (up to four) tiles that hold the rows of the next superstep's diagonal block count themselves off;
// the last one solves that block, so the next launch finds its y in Y.
template <typename T>
__global__ void __launch_bounds__ (256, 2) k_solve_fwd256 (const SolveTile256 *__restrict__ tiles,
    const int64_t *__restrict__ super, const int64_t *__restrict__ pi, const int64_t *__restrict__ px,
    const int32_t *__restrict__ Ls, const T *__restrict__ Lx, const typename solve_x<T>::type *__restrict__ Winv,
    const typename solve_x<T>::type *__restrict__ W256,
    typename solve_x<T>::type *X, typename solve_x<T>::type *Y, int *counters, int64_t n, int nrhs)
{
    typedef typename solve_x<T>::type XT ;
    __shared__ XT yb [256], xs [256], ys [256], red [512] ;
    __shared__ int is_last ;
    const SolveTile256 t = tiles [blockIdx.x] ;
    const int64_t k1 = super [t.s], nscol = super [t.s+1] - k1, psi = pi [t.s], nsrow = pi [t.s+1] - psi ;
    const T *Lsn = Lx + px [t.s] ;
    const int tid = threadIdx.x, r4 = tid >> 2, part = tid & 3, row = t.r0 + r4 ;
    const XT zero = from_real<XT> (0) ;
    const int jq = min (64, t.nb - 64 * part) ;         // columns of this thread's block of the superstep (may be <= 0)
    SOLVE_TR_DECL ;
    SOLVE_TR (0) ;
    if (t.next_nb > 0 && t.r0 == t.c0 + t.nb)
    {
        if (W256) solve_w256_prefetch<XT> (W256 + (int64_t) (t.step + 1) * (SOLVE_W256_LD * SOLVE_W256_LD), t.next_nb) ;
        else solve_diag256_prefetch<T, XT> (Lsn, nsrow, t.c0 + t.nb, t.next_nb,
            Winv + (int64_t) (t.blk0 + ((t.nb + 63) >> 6)) * (SOLVE_WLD * SOLVE_WLD)) ;
    }
    solve_pdl_wait () ;
    for (int rhs = 0 ; rhs < nrhs ; rhs++)
    {
        XT *x = X + (int64_t) rhs * n, *y = Y + (int64_t) rhs * n ;
        yb [tid] = (tid < t.nb) ? y [k1 + t.c0 + tid] : zero ;
        __syncthreads () ;
        SOLVE_TR (1) ;
        XT acc = zero ;
        if (r4 < t.nrows && jq > 0)
        {
            const T *Lr = Lsn + row + (int64_t) (t.c0 + 64 * part) * nsrow ;
            const XT *yq = yb + 64 * part ;
            if (jq == 64)
            {
                #pragma unroll 16
                for (int j = 0 ; j < 64 ; j++) acc += cvt<XT> (Lr [(int64_t) j * nsrow]) * yq [j] ;
            }
            else for (int j = 0 ; j < jq ; j++) acc += cvt<XT> (Lr [(int64_t) j * nsrow]) * yq [j] ;
        }
        acc += shfl_xor_ (0xffffffffu, acc, 1) ;
        acc += shfl_xor_ (0xffffffffu, acc, 2) ;
        SOLVE_TR (2) ;
        if (part == 0 && r4 < t.nrows)
        {
            if (row < nscol) x [k1 + row] -= acc ;       // a later unknown of this supernode: ours alone
            else atomic_sub (x + Ls [psi + row], acc) ;
        }
        SOLVE_TR (3) ;
        if (t.next_nb > 0)
        {
            __threadfence () ;
            __syncthreads () ;
            SOLVE_TR (4) ;
            if (tid == 0)
            {
                int *cnt = counters + (int64_t) t.step * nrhs + rhs ;
                const int ndiag = (t.next_nb + 63) >> 6 ;
                const int done = atomicAdd (cnt, 1) ;
                is_last = (done == ndiag - 1) ;
                if (is_last) *cnt = 0 ;
            }
            __syncthreads () ;
            SOLVE_TR (5) ;
            if (is_last)
            {
                __threadfence () ;
                xs [tid] = (tid < t.next_nb) ? ldcg_ (x + k1 + t.c0 + t.nb + tid) : zero ;
                __syncthreads () ;
                SOLVE_TR (6) ;
                if (W256) solve_diag256w_fwd<XT> (W256 + (int64_t) (t.step + 1) * (SOLVE_W256_LD * SOLVE_W256_LD), t.next_nb, xs, ys, red) ;
                else solve_diag256_fwd<T, XT, 2> (Lsn, nsrow, t.c0 + t.nb, t.next_nb,
                    Winv + (int64_t) (t.blk0 + ((t.nb + 63) >> 6)) * (SOLVE_WLD * SOLVE_WLD), xs, ys) ;
                SOLVE_TR (7) ;
                if (tid < t.next_nb) y [k1 + t.c0 + t.nb + tid] = ys [tid] ;
                SOLVE_TR (8) ;
                SOLVE_TR_FLUSH (0, 9) ;
            }
        }
        __syncthreads () ;
    }
}

// Backward superstep, one CTA per 64 rows below it: Y_blk -= L (rows, superstep)^H * x [rows] (atomics).  A warp
// takes 32 columns, 16 at a time; a lane holds two rows of each (32 independent, coalesced loads per pass) and the
// warp adds up with warp_sum16, so that 16 lanes subtract 16 consecutive entries of Y.  The CTA that finishes
// last then solves the diagonal block backwards, four sub-steps of 64 from the bottom up:
// x_q = W_q^H * y_q ;  y_c -= L (block q, column c)^H * x_q for the columns c of the blocks above (4 threads per column).
template <typename T>
__global__ void __launch_bounds__ (256, (sizeof (typename solve_x<T>::type) <= 8) ? 2 : 1) k_solve_bwd256 (const SolveTile256 *__restrict__ tiles,
    const int64_t *__restrict__ super, const int64_t *__restrict__ pi, const int64_t *__restrict__ px,
    const int32_t *__restrict__ Ls, const T *__restrict__ Lx, const typename solve_x<T>::type *__restrict__ Winv,
    const typename solve_x<T>::type *__restrict__ W256,
    typename solve_x<T>::type *X, typename solve_x<T>::type *Y, int *counters, int64_t n, int nrhs)
{
    typedef typename solve_x<T>::type XT ;
    __shared__ XT xr [64], yb [256], xq [64] ;
    __shared__ int is_last ;
    const SolveTile256 t = tiles [blockIdx.x] ;
    const int64_t k1 = super [t.s], nscol = super [t.s+1] - k1, psi = pi [t.s], nsrow = pi [t.s+1] - psi ;
    const T *Lsn = Lx + px [t.s] ;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, r4 = tid >> 2, part = tid & 3 ;
    const XT zero = from_real<XT> (0) ;
    SOLVE_TR_DECL ;
    SOLVE_TR (0) ;
    if (t.r0 == t.c0 + t.nb)
    {
        if (W256) solve_w256_prefetch<XT> (W256 + (int64_t) t.step * (SOLVE_W256_LD * SOLVE_W256_LD), t.nb) ;
        else solve_diag256_prefetch<T, XT> (Lsn, nsrow, t.c0, t.nb, Winv + (int64_t) t.blk0 * (SOLVE_WLD * SOLVE_WLD)) ;
    }
    solve_pdl_wait () ;
    for (int rhs = 0 ; rhs < nrhs ; rhs++)
    {
        XT *x = X + (int64_t) rhs * n, *y = Y + (int64_t) rhs * n ;
        if (tid < 64)
        {
            const int row = t.r0 + tid ;
            xr [tid] = (tid < t.nrows) ? x [(row < nscol) ? k1 + row : (int64_t) Ls [psi + row]] : zero ;
        }
        __syncthreads () ;
        SOLVE_TR (1) ;
        if (t.nrows > 0)
        {
            const T *Lbase = Lsn + t.r0 + (int64_t) t.c0 * nsrow ;
            const XT x0 = xr [lane], x1 = xr [lane + 32] ;
            const bool ok0 = lane < t.nrows, ok1 = lane + 32 < t.nrows ;
            for (int h = 0 ; h < 2 ; h++)
            {
                const int cb = warp * 32 + 16 * h ;
                if (cb >= t.nb) break ;
                T a [16], b [16] ;
                if (t.nrows == 64 && cb + 16 <= t.nb)
                {
                    #pragma unroll
                    for (int u = 0 ; u < 16 ; u++)
                    {
                        const T *Lc = Lbase + (int64_t) (cb + u) * nsrow ;
                        a [u] = Lc [lane] ; b [u] = Lc [lane + 32] ;
                    }
                }
                else
                {
                    #pragma unroll
                    for (int u = 0 ; u < 16 ; u++)
                    {
                        const T *Lc = Lbase + (int64_t) (cb + u) * nsrow ;
                        const bool cok = cb + u < t.nb ;
                        a [u] = (cok && ok0) ? Lc [lane] : T () ;
                        b [u] = (cok && ok1) ? Lc [lane + 32] : T () ;
                    }
                }
                XT v [16] ;
                #pragma unroll
                for (int u = 0 ; u < 16 ; u++) v [u] = mulc (x0, cvt<XT> (a [u])) + mulc (x1, cvt<XT> (b [u])) ;
                warp_sum16<XT> (v, lane) ;
                if (lane < 16 && cb + lane < t.nb) atomic_sub (y + k1 + t.c0 + cb + lane, v [0]) ;
            }
        }
        SOLVE_TR (2) ;
        __threadfence () ;
        __syncthreads () ;
        SOLVE_TR (3) ;
        if (tid == 0)
        {
            int *cnt = counters + (int64_t) t.step * nrhs + rhs ;
            const int done = atomicAdd (cnt, 1) ;
            is_last = (done == t.ntiles - 1) ;
            if (is_last) *cnt = 0 ;
        }
        __syncthreads () ;
        SOLVE_TR (4) ;
        if (is_last)
        {
            __threadfence () ;
            yb [tid] = (tid < t.nb) ? ldcg_ (y + k1 + t.c0 + tid) : zero ;
            __syncthreads () ;
            SOLVE_TR (5) ;
            if (W256)
            {
                XT x0, x1 ;
                solve_diag256w_bwd<XT> (W256 + (int64_t) t.step * (SOLVE_W256_LD * SOLVE_W256_LD), t.nb, yb, x0, x1) ;
                if (lane < 16)
                {
                    const int ca = 16 * warp + lane, cb = 255 - ca ;
                    if (ca < t.nb) x [k1 + t.c0 + ca] = x0 ;
                    if (cb < t.nb) x [k1 + t.c0 + cb] = x1 ;
                }
            }
            const int nq = W256 ? 0 : (t.nb + 63) >> 6 ;
            for (int q = nq - 1 ; q >= 0 ; q--)
            {
                const int q0 = q * 64, jb = min (64, t.nb - q0) ;
                const XT *W = Winv + (int64_t) (t.blk0 + q) * (SOLVE_WLD * SOLVE_WLD) ;
                // rows of block q in the columns of the blocks above it: independent of x_q, asked for first
                constexpr int PRE = (sizeof (T) <= 8) ? 2 : 1 ;
                T lv [PRE][16] ;
                #pragma unroll
                for (int g = 0 ; g < PRE ; g++)
                {
                    const T *Lc = Lsn + (t.c0 + q0 + 16 * part) + (int64_t) (t.c0 + 64 * g + r4) * nsrow ;
                    #pragma unroll
                    for (int r = 0 ; r < 16 ; r++) lv [g][r] = (64 * g < q0 && 16 * part + r < jb) ? Lc [r] : T () ;
                }
                {
                    // x_c = sum_{r >= c} conj (W (r,c)) * y_r : thread (c, part) walks down column c of W
                    XT wv [16] ;
                    #pragma unroll
                    for (int k = 0 ; k < 16 ; k++) wv [k] = (part + 4 * k >= r4) ? W [part + 4 * k + r4 * SOLVE_WLD] : zero ;
                    XT acc = zero ;
                    #pragma unroll
                    for (int k = 0 ; k < 16 ; k++)
                        if (part + 4 * k < jb) acc += mulc (yb [q0 + part + 4 * k], wv [k]) ;
                    acc += shfl_xor_ (0xffffffffu, acc, 1) ;
                    acc += shfl_xor_ (0xffffffffu, acc, 2) ;
                    if (part == 0 && r4 < jb) { xq [r4] = acc ; x [k1 + t.c0 + q0 + r4] = acc ; }
                }
                __syncthreads () ;
                #pragma unroll
                for (int g = 0 ; g < 3 ; g++)
                {
                    const int c = 64 * g + r4 ;         // a column of a block above (c < q0 when 64 g < q0: q0 is a multiple of 64)
                    XT acc = zero ;
                    if (g < PRE)
                    {
                        #pragma unroll
                        for (int r = 0 ; r < 16 ; r++)
                            if (16 * part + r < jb) acc += mulc (xq [16 * part + r], cvt<XT> (lv [g < PRE ? g : 0][r])) ;
                    }
                    else if (64 * g < q0)
                    {
                        const T *Lc = Lsn + (t.c0 + q0 + 16 * part) + (int64_t) (t.c0 + c) * nsrow ;
                        #pragma unroll
                        for (int r = 0 ; r < 16 ; r++)
                            if (16 * part + r < jb) acc += mulc (xq [16 * part + r], cvt<XT> (Lc [r])) ;
                    }
                    acc += shfl_xor_ (0xffffffffu, acc, 1) ;
                    acc += shfl_xor_ (0xffffffffu, acc, 2) ;
                    if (part == 0 && 64 * g < q0) yb [c] -= acc ;
                }
                __syncthreads () ;
            }
            SOLVE_TR (6) ;
            SOLVE_TR_FLUSH (16, 7) ;
        }
        __syncthreads () ;
    }
}

// After a forward sweep on its own (cholmod_super_lsolve without the ltsolve that normally follows): the
// results of the wide supernodes' columns are in Y, everything else in X; gather them into X.
template <typename XT>
__global__ void __launch_bounds__ (64) k_solve_merge (const SolveBlk *__restrict__ blks, const int64_t *__restrict__ super,
    XT *X, const XT *__restrict__ Y, int64_t n, int nrhs)
{
    const SolveBlk b = blks [blockIdx.x] ;
    const int64_t at = super [b.s] + b.c0 + threadIdx.x ;
    if ((int) threadIdx.x < b.jb)
        for (int rhs = 0 ; rhs < nrhs ; rhs++) X [(int64_t) rhs * n + at] = Y [(int64_t) rhs * n + at] ;
}

}   // namespace b200
