// plan.hpp -- host-side launch plan of the CHOLMOD/B200 numeric phase.
//
// Everything the reference recomputes per factorization that depends only on the symbolic factor
// (CHOLMOD/Supernodal/t_cholmod_super_numeric_worker.c "SNW": descendant lists :531-551, ndrow1/ndrow2
// :686-698, relative maps :897-903, re-linking :950-967, parent links :1222-1235) is computed here once
// per analysis and turned into flat tables the kernels index directly.
#pragma once
#include <cstdint>
#include <vector>

namespace b200 {

// One dense product  C (M x N) = A (M x K) * B (N x K)',  A and B in the same K columns of one panel, B starting
// `shift` rows above A (shift = 0: B is the first N rows of A).  Only the entries with  i + shift >= j  are wanted
// -- the lower trapezoid when shift = 0, everything when shift >= N (a block of rows far below the diagonal).
// It is subtracted from the factor:
//   map_off >= 0 :  Lx [c_off + map [i] + map [j] * ldc] -= C (i,j)     (descendant -> ancestor update,
//                   the dsyrk + dgemm + assembly of SNW:769-930; map = the pair's relative map; shift = 0)
//   map_off <  0 :  Lx [c_off + i + j * ldc]             -= C (i,j)     (trailing update inside a supernode
//                   during its blocked dpotrf/dtrsm, SNW:1023-1181; c_off = the target's (0,0))
struct UpdItem
{
    int64_t a_off ;     // arena offset of A (0,0); A is column-major with leading dimension lda
    int64_t c_off ;     // arena offset of the target (supernode origin, or C (0,0) when direct)
    int64_t map_off ;   // offset into the relative-map table, or -1
    int32_t lda, ldc ;
    int32_t M, N, K ;
    int32_t shift ;     // B (0,0) = A (0,0) - shift rows
} ;

// One tile of one item: rows [ti*TM, ..), columns [tj*TN, ..), inner range [k0, k1).
struct UpdTile
{
    int32_t item ;
    int32_t ti, tj ;    // 32 bits: a supernode may have up to 2^30 rows (build_plan), i.e. 2^25 SIMT tile rows
    int32_t k0, k1 ;
} ;

// Diagonal block of a supernode to factorize in place (jb <= POTRF_NB), lower Cholesky.
struct PotrfBlk
{
    int64_t off ;       // Lx offset of the block's (0,0)
    int64_t col ;       // global column index of the block's first column (for L->minor)
    int32_t lda, jb ;
    int32_t slot ;      // inverse slot in the winv scratch (k_trsm_inv follows), or -1
    int32_t pad ;
} ;

// Rows below a factorized diagonal block:  X <- X * L11^{-T}   (the dtrsm of SNW:1175, R,L,C,N)
struct TrsmTile
{
    int64_t diag_off ;  // Lx offset of L11 (0,0)
    int64_t row_off ;   // Lx offset of X (0,0)
    int32_t lda, jb, nrows ;
    int32_t slot ;      // inverse slot of the diagonal block (K_TRSM_INV*), or -1
} ;

enum LaunchKind : int32_t
{
    K_UPD_SIMT = 0,     // UpdTile list, 32 x 8 tiles, one warp each
    K_UPD_D64  = 1,     // UpdTile list, 64 x 64 DMMA tiles
    K_UPD_D128 = 2,     // UpdTile list, BIG_TM x BIG_TN DMMA tiles
    K_POTRF16  = 3,     // PotrfBlk list, jb <= 16
    K_POTRF32  = 4,     // PotrfBlk list, jb <= 32
    K_POTRF64  = 5,     // PotrfBlk list, jb <= 64
    K_TRSM     = 6,     // TrsmTile list, substitution
    K_TRSM_INV32 = 7,   // TrsmTile list, product with the inverse of a <= 32-wide block
    K_TRSM_INV64 = 8,   // TrsmTile list, product with the inverse of a <= 64-wide block
    K_FACTOR_SMALL = 9, // PotrfBlk list: whole supernodes of <= 16 columns and few rows, potrf + trsm fused
    K_FACTOR_TINY = 10, // PotrfBlk list: whole supernodes of <= TINY_COLS columns and <= TINY_ROWS rows below, one THREAD each
    K_UPD_TINY = 11,    // UpdTile list, one tile = one whole item of a few dozen entries and K <= TINY_K, one THREAD each
    K_NKINDS   = 12,
    K_JOIN     = 12,    // no kernel: only the launch's wait_ev / rec_ev are honoured
    K_COPYOUT  = 13,    // no kernel: CopySeg list, finished parts of L->x go to the host (copy stream)
    K_BCAST    = 14,    // no kernel: BcastSeg list, ncclBroadcast of finished panels (communication stream)
    K_BARRIER  = 15,    // no kernel: all ranks meet (a one-word all-reduce on the communication stream)
    K_PULL     = 16,    // PullSeg list: the owner of a block of a top separator adds the other ranks' contribution
                        // blocks to its own, reading them from their arenas over NVLink (k_pull_add)
} ;

struct Launch
{
    int32_t kind ;
    int32_t count ;     // number of list entries
    int64_t first ;     // first entry in the kind's list
    int32_t level ;     // etree level this launch belongs to
    int32_t stream ;    // 0 = main (panel) stream, 1 = look-ahead stream of the wide trailing updates,
                        // 2 = copy stream (K_COPYOUT), 3 = communication stream (K_BCAST, K_BARRIER),
                        // 4 = far stream (the rows of a panel far below the diagonal block, see plan_factor)
    int32_t wait_ev ;   // event the stream waits for before this launch, or -1
    int32_t rec_ev ;    // event recorded on the stream after this launch, or -1
} ;

constexpr int POTRF_NB   = 64 ;     // width of a block column factorized by one POTRF block
constexpr int OUTER_NB   = 256 ;    // width of an outer block column (trailing updates with K = OUTER_NB)
constexpr int TRSM_ROWS  = 128 ;    // rows per TRSM tile
constexpr int SMALL_ROWS = 256 ;    // a supernode of <= 16 columns with at most this many rows below its
                                    // diagonal block is factorized by one 16-thread group (K_FACTOR_SMALL)
constexpr int SIMT_TM = 32, SIMT_TN = 8 ;
// The leaves of a nested-dissection etree: most supernodes have 1-4 columns and a handful of rows, most update
// pairs a few dozen entries with K = 1..4.  A warp (or a 16-thread group) per such object idles 30 of its lanes;
// they get one thread each instead.
constexpr int TINY_COLS = 4, TINY_ROWS = 64 ;      // K_FACTOR_TINY: nscol <= TINY_COLS, nsrow - nscol <= TINY_ROWS
constexpr int TINY_K = 4, TINY_MN = 128 ;          // K_UPD_TINY: K <= TINY_K, M * N <= TINY_MN
// the large DMMA tile: 128 x 64 with 4 warps of 32 x 64, two CTAs per SM so that one CTA's prologue,
// barriers and epilogue overlap the other's tensor work
constexpr int BIG_TM = 128, BIG_TN = 64 ;
// multi-GPU: consecutive outer blocks of a distributed supernode that stay on one rank (block-cyclic with
// blocks of DIST_RUN * OUTER_NB columns): inside a run the panel chain never waits for a broadcast
constexpr int DIST_RUN = 4 ;

// A finished, contiguous piece of L->x that can go to the host early: `len` values at offset `off` of the
// device arena belong at offset `hoff` of the caller's L->x (the two differ only in a distributed plan, whose
// arena holds the top separators and this rank's subtrees, nothing else).
struct CopySeg { int64_t off, len, hoff ; } ;

// A finished, contiguous piece of the arena that its owner rank broadcasts to all other ranks.
struct BcastSeg { int64_t off, len ; int32_t root, pad ; } ;

// Contribution blocks: rows [0, nrows) x columns [0, ncols) with leading dimension ld at arena
// offset `off` -- the lower part of one 256-column block of a top separator, at the same offset on every rank.
// Every rank has accumulated there the updates of ITS subtrees; the block's owner adds up the copies of the
// ranks in `mask` (bit r = rank r has something to contribute).
struct PullSeg { int64_t off ; int32_t nrows, ncols, ld ; uint32_t mask ; } ;

// One (descendant d, ancestor s) update: rows lpos.. of d; the first ndrow1 fall in the columns of s.
struct Pair { int32_t d, s, lpos, ndrow1, ndrow2 ; } ;

struct Plan
{
    int64_t n = 0, nsuper = 0, ssize = 0, xsize = 0 ;
    int32_t nlevels = 0 ;
    int64_t npairs = 0 ;
    double flops_update = 0, flops_potrf = 0, flops_trsm = 0 ;
    // the reference's BLAS call statistics (cholmod_gpu_stats): one dsyrk per pair, one dgemm per pair with rows
    // below the ancestor's columns, one dpotrf per supernode, one dtrsm per supernode with rows below its block
    double flops_syrk = 0, flops_gemm = 0 ;
    int64_t npairs_gemm = 0, nsuper_trsm = 0 ;
    double flops_kind [K_NKINDS] = {} ;     // algorithmic flops executed by each launch kind
    int64_t count_kind [K_NKINDS] = {} ;    // list entries (tiles / blocks) per launch kind
    int64_t winv_slots = 0 ;                // inverse slots (64 x 64 values each) the largest launch needs
    int32_t nevents = 0 ;                   // events the look-ahead schedule uses
    // single-GPU plans: the levels [fork_lo, fork_hi) are factorized subtree group by subtree group, each
    // group on its own pair of streams (5 + 2g: panel chain, 6 + 2g: look-ahead), see build_plan
    int32_t nstreams = 5 ;                  // stream ids a launch may carry: 0 main, 1 look-ahead, 2 copy, 3 comm, 4 far, 5.. groups
    int32_t ngroups = 1, fork_lo = -1, fork_hi = -1 ;
    std::vector<int32_t> group ;            // [nsuper] group of a supernode below fork_hi (else -1)

    std::vector<int32_t> level ;        // [nsuper] height above the leaves in the supernodal etree
    std::vector<int64_t> parent ;       // [nsuper] etree parent or -1
    std::vector<int32_t> supermap ;     // [n] column -> supernode (cholmod_super_numeric.c:276-285)
    std::vector<int32_t> relmap ;       // concatenated relative maps of all pairs (empty when they are built on the
                                        // device instead, symbolic.cuh: only their offsets pair_map exist here then)
    int64_t relmap_len = 0 ;            // total length of the relative maps
    std::vector<UpdItem> items ;
    std::vector<UpdTile> tiles ;
    std::vector<PotrfBlk> potrf ;
    std::vector<TrsmTile> trsm ;
    std::vector<Launch> launches ;
    std::vector<CopySeg> copies ;       // K_COPYOUT segments; together they cover what this rank owns exactly once
    std::vector<BcastSeg> bcasts ;      // K_BCAST segments (identical on every rank)
    std::vector<PullSeg> pulls ;        // K_PULL segments of the blocks this rank owns
    std::vector<CopySeg> owned ;        // everything this rank owns, as maximal runs (one run [0, xsize) on one GPU)

    // device layout: supernode s lives at dpx [s] of the arena (-1: not on this rank), dxsize values in all.
    // One GPU: dpx == px.  Distributed: the top separators first (same offsets on every rank), then this
    // rank's subtrees in postorder.
    std::vector<int64_t> dpx ;
    int64_t dxsize = 0, top_size = 0 ;
    std::vector<unsigned char> colmine ;    // [n] distributed: 1 if this rank scatters column k of A (it owns the
                                            // column's subtree, or the column's block of a top separator)
    int64_t nvlink_in_values = 0 ;      // what one factorization brings in over NVLink: panels received + blocks pulled
    int64_t trunc_s = -1, trunc_ncol = 0 ;  // >= 0: supernodes after trunc_s are left zero and trunc_s itself is
                                            // factorized for its first trunc_ncol columns only (SNW:1084-1146)

    // Distribution over the GPUs of one box.  nranks == 1: everything belongs to rank 0.
    // "top" supernodes (the separators whose subtree carries more than 1/(4 nranks) of the work) are split
    // by outer block columns, block T of s belonging to rank (owner [s] + T / dist_run) % nranks (1-D block
    // cyclic); every other supernode belongs, with its whole subtree, to one rank, is stored on that rank only,
    // and reaches the top separators as that rank's contribution blocks (build_dist in plan.cpp).
    int32_t nranks = 1, rank = 0 ;
    int32_t dist_run = 0 ;              // outer blocks per block-cyclic unit; 0 = chosen per supernode (env CHOLMOD_B200_DIST_RUN overrides)
    std::vector<int32_t> run ;          // [nsuper] the unit of top supernode s: about a third of an even share of its
                                        // blocks, at most DIST_RUN (long runs keep the chain off the broadcasts, short
                                        // runs balance the update work of a supernode with few blocks)
    std::vector<char> top ;             // [nsuper]
    std::vector<int32_t> owner ;        // [nsuper] owning rank (top: the rank of outer block 0)
    std::vector<int32_t> sroot ;        // [nsuper] root of the local subtree of a non-top supernode

    std::vector<Pair> pairs ;           // all update pairs, in descendant order
    std::vector<int64_t> pair_map ;     // [npairs] offset of the pair's relative map in relmap
    std::vector<int64_t> target_ptr ;   // [nsuper+1] pairs grouped by target supernode ...
    std::vector<int32_t> target_list ;  // ... as indices into pairs

    // solve schedule: supernodes grouped by level (forward: ascending levels; backward: descending)
    std::vector<int64_t> level_ptr ;    // [nlevels+1]
    std::vector<int32_t> level_list ;   // [nsuper] supernodes sorted by level
} ;

// Builds the plan.  Returns 0, or a negative B200_* status (invalid input / too large).
// use_dmma = false (single precision): no inverse-based panel solves, no near / far split.
// nranks > 1 builds rank `rank`'s share of the distributed plan (build_dist in plan.cpp): its own subtrees, its
// contribution blocks, the pulls of the blocks it owns, and for the top separators the same launch structure on
// every rank, each launch holding only the blocks / tiles that rank owns, plus the K_BARRIER and K_BCAST launches
// all ranks share.
// host_relmap = false: only the offsets of the relative maps are computed; the maps themselves are built on the
// device (symbolic.cuh).
// trunc_s >= 0 (distributed plans only): the plan of the not-positive-definite case -- everything up to supernode
// trunc_s as usual, trunc_s itself for its first trunc_ncol columns, nothing after it.
int build_plan (int64_t n, int64_t nsuper, const int64_t *super, const int64_t *pi, const int64_t *px,
    const int64_t *Ls, bool use_dmma, Plan &plan, int nranks = 1, int rank = 0, int64_t trunc_s = -1,
    int64_t trunc_ncol = 0, bool host_relmap = true) ;

// SNW:1084-1146: a one-supernode plan that re-applies every descendant update to supernode s and then
// factorizes only its first nscol_new columns (dtrsm on all rows below them).  Reuses P.relmap.
int build_repair_plan (const Plan &P, const int64_t *super, const int64_t *pi, const int64_t *px,
    int64_t s, int64_t nscol_new, bool use_dmma, Plan &R) ;

}   // namespace b200
