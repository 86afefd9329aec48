// symbolic.cuh -- device-side symbolic helpers: the update-pair list and the relative maps
// of a supernodal symbolic factor, built on the GPU from L->super, L->pi, L->s.
//
// The host planner (plan.cpp) enumerates the (descendant, ancestor) pairs with the walk the reference uses to size
// its update workspace (CHOLMOD/Supernodal/cholmod_super_symbolic.c:896-921) and fills every pair's relative map
// RelativeMap [i] = Map [Ls [pdi1 + i]] (t_cholmod_super_numeric_worker.c:388-399, :897-903).  The relative maps are
// the big table (config 2: 17.9 M entries, 72 MB, a third of the planner's time plus the upload); here they never
// exist on the host: the pairs are enumerated per descendant, ordered by target with a stable radix sort, and one
// warp per pair looks its rows up in the target's row list.  The tables are bit-identical to the host's
// (tests/test_gpu_symbolic.py).
#pragma once
#include "plan.hpp"
#include <cuda_runtime.h>
#include <cub/cub.cuh>
#include <stdint.h>

namespace b200 {

// Walk the rows of descendant d below its diagonal block: every maximal run of rows that fall into the columns of
// one supernode s is a pair (d, s).  emit == nullptr: count only.
__device__ __forceinline__ int sym_walk (int64_t d, const int64_t *__restrict__ super, const int64_t *__restrict__ pi,
    const int32_t *__restrict__ Ls, const int32_t *__restrict__ supermap, Pair *emit)
{
    const int64_t ndcol = super [d+1] - super [d] ;
    int64_t p = pi [d] + ndcol ;
    const int64_t pend = pi [d+1] ;
    int n = 0 ;
    while (p < pend)
    {
        const int32_t s = supermap [Ls [p]] ;
        const int64_t k2 = super [s+1] ;
        int64_t q = p ;
        while (q < pend && (int64_t) Ls [q] < k2) q++ ;
        if (emit)
        {
            Pair pr ;
            pr.d = (int32_t) d ; pr.s = s ; pr.lpos = (int32_t) (p - pi [d]) ;
            pr.ndrow1 = (int32_t) (q - p) ; pr.ndrow2 = (int32_t) (pend - p) ;
            emit [n] = pr ;
        }
        n++ ;
        p = q ;
    }
    return n ;
}

static __global__ void __launch_bounds__ (128) k_sym_pair_count (int64_t nsuper, const int64_t *__restrict__ super,
    const int64_t *__restrict__ pi, const int32_t *__restrict__ Ls, const int32_t *__restrict__ supermap, int64_t *count)
{
    const int64_t d = (int64_t) blockIdx.x * blockDim.x + threadIdx.x ;
    if (d < nsuper) count [d] = sym_walk (d, super, pi, Ls, supermap, nullptr) ;
}

static __global__ void __launch_bounds__ (128) k_sym_pair_emit (int64_t nsuper, const int64_t *__restrict__ super,
    const int64_t *__restrict__ pi, const int32_t *__restrict__ Ls, const int32_t *__restrict__ supermap,
    const int64_t *__restrict__ first, Pair *pairs, int32_t *key_s, int32_t *idx)
{
    const int64_t d = (int64_t) blockIdx.x * blockDim.x + threadIdx.x ;
    if (d >= nsuper) return ;
    const int64_t at = first [d] ;
    const int n = sym_walk (d, super, pi, Ls, supermap, pairs + at) ;
    for (int k = 0 ; k < n ; k++) { key_s [at + k] = pairs [at + k].s ; idx [at + k] = (int32_t) (at + k) ; }
}

// ndrow2 of the pairs in target order (the sizes of their relative maps, in the order the maps are laid out)
static __global__ void __launch_bounds__ (256) k_sym_gather_len (int64_t npairs, const Pair *__restrict__ pairs,
    const int32_t *__restrict__ order, int64_t *len)
{
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x ;
    if (t < npairs) len [t] = pairs [order [t]].ndrow2 ;
}

// One warp per pair (in target order): relmap [off + i] = position of row i of the pair inside the row list of
// its target s -- r - k1 inside the diagonal block, a binary search of the ascending off-diagonal rows otherwise
// (the reference's Map array).  Also stores the pair's offset: pair_map [pair] = off.
static __global__ void __launch_bounds__ (256) k_sym_relmap (int64_t npairs, const Pair *__restrict__ pairs,
    const int32_t *__restrict__ order, const int64_t *__restrict__ off, const int64_t *__restrict__ super,
    const int64_t *__restrict__ pi, const int32_t *__restrict__ Ls, int64_t *pair_map, int32_t *relmap, int *bad)
{
    const int64_t t = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5 ;
    const int lane = threadIdx.x & 31 ;
    if (t >= npairs) return ;
    const int32_t pid = order [t] ;
    const Pair pr = pairs [pid] ;
    const int64_t o = off [t] ;
    if (lane == 0) pair_map [pid] = o ;
    const int64_t k1 = super [pr.s], k2 = super [pr.s+1], psi = pi [pr.s] ;
    const int64_t nscol = k2 - k1, nsrow = pi [pr.s+1] - psi ;
    const int32_t *rows = Ls + pi [pr.d] + pr.lpos ;
    for (int i = lane ; i < pr.ndrow2 ; i += 32)
    {
        const int64_t r = rows [i] ;
        int64_t pos ;
        if (r < k2) pos = r - k1 ;
        else
        {
            int64_t lo = nscol, hi = nsrow ;
            while (lo < hi)
            {
                const int64_t mid = (lo + hi) >> 1 ;
                if ((int64_t) Ls [psi + mid] < r) lo = mid + 1 ; else hi = mid ;
            }
            pos = lo ;
            // the supernodal inclusion property: every remaining row of d is a row of s
            if (lo >= nsrow || (int64_t) Ls [psi + lo] != r) *bad = 1 ;
        }
        relmap [o + i] = (int32_t) pos ;
    }
}

}   // namespace b200
