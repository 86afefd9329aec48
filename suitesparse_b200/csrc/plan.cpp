// plan.cpp -- builds the level-ordered launch plan from the supernodal symbolic factor (host, once
// per analysis).  See plan.hpp.  Reference behaviour restated here (not copied):
//   * pair enumeration = the walk cholmod_super_symbolic.c:896-921 uses to size C, which visits exactly
//     the (d, s) pairs the numeric loop later visits through its Head/Next/Lpos lists (SNW:588-969);
//   * relative map of a pair = Map [Ls [pdi1 + i]] with Map built from the rows of s (SNW:388-399, :897-903);
//   * etree parent of s = SuperMap [Ls [pi [s] + nscol]] (SNW:1226).
#include "plan.hpp"
#include "../../include/cholmod_b200.h"
#include <algorithm>
#include <cstring>
#include <numeric>
#include <cstdlib>
#include <chrono>
#include <cstdio>

namespace b200 {

namespace {

struct TileSink
{
    std::vector<UpdTile> cls [4] ;      // per update kernel class: SIMT, DMMA 64, DMMA 128 x 64, tiny
    void clear () { for (auto &v : cls) v.clear () ; }
} ;

inline int64_t cdiv (int64_t a, int64_t b) { return (a + b - 1) / b ; }

// Who does what.  `D` carries the distribution (Plan::nranks/rank/top/owner); a repair plan is built
// with local == true: every rank then does everything itself on its own complete copy of the factor.
struct Own
{
    const Plan *D ;
    bool local ;
    bool dist () const { return !local && D->nranks > 1 ; }
    bool is_top (int64_t s) const { return dist () && D->top [s] ; }
    int block_owner (int64_t s, int64_t T) const
    {
        if (!dist ()) return D->rank ;
        return D->top [s] ? (int) ((D->owner [s] + T / D->run [s]) % D->nranks) : D->owner [s] ;
    }
    bool mine (int64_t s, int64_t T) const { return block_owner (s, T) == D->rank ; }
} ;

// choose the kernel class of an item and emit its tiles
void emit_tiles (Plan &P, int32_t item_id, const UpdItem &it, TileSink &sink, bool use_dmma)
{
    const int64_t M = it.M, N = it.N, K = it.K ;
    int cls ;
    (void) use_dmma ;       // both precisions use the tile classes (double: DMMA, single: FFMA kernels)
    if (it.shift == 0 && K <= TINY_K && M * N <= TINY_MN)
    {
        // one thread does the whole item: a single "tile"
        P.flops_kind [K_UPD_TINY] += (double) K * (double) N * (2.0 * (double) M - (double) N) ;
        UpdTile t ;
        t.item = item_id ; t.ti = 0 ; t.tj = 0 ; t.k0 = 0 ; t.k1 = (int32_t) K ;
        sink.cls [3].push_back (t) ;
        return ;
    }
    if (N < 16 || K < 8 || M * N < 2048)
    {
        cls = K_UPD_SIMT ;
    }
    else
    {
        // padded work of the two DMMA tilings; the 64-tile kernel re-reads operands twice as often
        double c64  = 1.25 * (double) (cdiv (M, 64) * 64) * (double) (cdiv (N, 64) * 64) ;
        double c128 = (double) (cdiv (M, BIG_TM) * BIG_TM) * (double) (cdiv (N, BIG_TN) * BIG_TN) ;
        cls = (c128 <= c64) ? K_UPD_D128 : K_UPD_D64 ;
    }
    const int TM = (cls == K_UPD_SIMT) ? SIMT_TM : (cls == K_UPD_D64 ? 64 : BIG_TM) ;
    const int TN = (cls == K_UPD_SIMT) ? SIMT_TN : (cls == K_UPD_D64 ? 64 : BIG_TN) ;
    const int64_t nti = cdiv (M, TM), ntj = cdiv (N, TN) ;
    // the tiles that hold a wanted entry (i + shift >= j): tile row ti reaches down to row ti*TM + TM-1
    auto first_ti = [&] (int64_t tj) -> int64_t { return std::max<int64_t> (0, tj * TN - it.shift) / TM ; } ;
    int64_t ntiles = 0 ;
    for (int64_t tj = 0 ; tj < ntj ; tj++) ntiles += nti - first_ti (tj) ;
    // split K when a long inner dimension meets few tiles (the output is accumulated with atomics,
    // so partial sums need no workspace)
    int64_t nsplit = 1 ;
    if (cls != K_UPD_SIMT && K >= 1024 && ntiles < 296)
    {
        nsplit = std::min<int64_t> (K / 256, cdiv (592, ntiles)) ;
        if (nsplit < 1) nsplit = 1 ;
    }
    int64_t kchunk = cdiv (cdiv (K, nsplit), 16) * 16 ;
    auto &out = sink.cls [cls] ;
    P.flops_kind [cls] += (it.shift >= N) ? 2.0 * (double) K * (double) M * (double) N
        : (double) K * (double) N * (2.0 * (double) M - (double) N) ;
    for (int64_t tj = 0 ; tj < ntj ; tj++)
    {
        for (int64_t ti = first_ti (tj) ; ti < nti ; ti++)
        {
            for (int64_t k0 = 0 ; k0 < K ; k0 += kchunk)
            {
                UpdTile t ;
                t.item = item_id ;
                t.ti = (int32_t) ti ; t.tj = (int32_t) tj ;
                t.k0 = (int32_t) k0 ; t.k1 = (int32_t) std::min<int64_t> (K, k0 + kchunk) ;
                out.push_back (t) ;
            }
        }
    }
}

void add_item (Plan &P, const UpdItem &it, TileSink &sink, bool use_dmma)
{
    P.items.push_back (it) ;
    emit_tiles (P, (int32_t) (P.items.size () - 1), it, sink, use_dmma) ;
}

Launch make_launch (int kind, int64_t first, int64_t count, int level, int stream)
{
    Launch L ;
    L.kind = kind ; L.count = (int32_t) count ; L.first = first ; L.level = level ;
    L.stream = stream ; L.wait_ev = -1 ; L.rec_ev = -1 ;
    return L ;
}

// Append the gathered tiles as one launch per kernel class on `stream`.  The first launch waits for
// wait_ev and the last records rec_ev; with nothing to launch the events are carried by a K_JOIN.
void flush_tiles (Plan &P, TileSink &sink, int level, int stream = 0, int wait_ev = -1, int rec_ev = -1)
{
    const size_t before = P.launches.size () ;
    for (int c = 0 ; c < 4 ; c++)
    {
        auto &v = sink.cls [c] ;
        if (v.empty ()) continue ;
        const int kind = (c == 3) ? (int) K_UPD_TINY : c ;
        // longest tiles first: the CTAs still running when the launch drains are then the short ones
        // (the one-warp tiles of the SIMT class and the one-thread items are all short: their order does not matter)
        if (c == K_UPD_D64 || c == K_UPD_D128)
            std::stable_sort (v.begin (), v.end (), [] (const UpdTile &a, const UpdTile &b) { return (a.k1 - a.k0) > (b.k1 - b.k0) ; }) ;
        P.launches.push_back (make_launch (kind, (int64_t) P.tiles.size (), (int64_t) v.size (), level, stream)) ;
        P.count_kind [kind] += (int64_t) v.size () ;
        P.tiles.insert (P.tiles.end (), v.begin (), v.end ()) ;
    }
    sink.clear () ;
    if (wait_ev < 0 && rec_ev < 0) return ;
    if (P.launches.size () == before) P.launches.push_back (make_launch (K_JOIN, 0, 0, level, stream)) ;
    P.launches [before].wait_ev = wait_ev ;
    P.launches.back ().rec_ev = rec_ev ;
}

// The descendant updates into supernode s; `px` holds the ARENA offsets of the supernodes (Plan::dpx).
//
// One GPU (and repair plans): every pair, whole.
// Distributed, `which` selects the part of the work that is planned:
//   UPD_ALL            s is an ordinary supernode of one of this rank's subtrees: all its descendants are in the
//                      same subtree, every pair whole;
//   UPD_FROM_SUBTREES  s is a top separator: the pairs whose descendant belongs to one of THIS rank's subtrees,
//                      whole, into this rank's own copy of s -- its contribution block, which the owners of the
//                      blocks of s later pull (K_PULL);
//   UPD_FROM_TOP       s is a top separator: the pairs whose descendant is a top separator too (its panels have
//                      been broadcast).  Each pair is cut into slabs of target columns, one slab per outer block,
//                      and this rank keeps the slabs of the blocks it owns: a slab [ja, jb) of an item is the item
//                      of the panel rows ja.. (B = its first jb-ja rows), so the same kernels apply.
enum { UPD_ALL = 0, UPD_FROM_SUBTREES = 1, UPD_FROM_TOP = 2 } ;

void plan_updates_into (Plan &out, const Plan &P, const Own &own, const int64_t *super, const int64_t *pi,
    const int64_t *px, int64_t s, TileSink &sink, bool use_dmma, int which = UPD_ALL)
{
    const bool dist = own.dist () ;
    const bool split = own.is_top (s) && which == UPD_FROM_TOP ;
    if (dist && which == UPD_ALL && (own.is_top (s) || !own.mine (s, 0))) return ;
    for (int64_t t = P.target_ptr [s] ; t < P.target_ptr [s+1] ; t++)
    {
        const int32_t pid = P.target_list [t] ;
        const Pair &pr = P.pairs [pid] ;
        const int64_t d = pr.d ;
        if (dist && which == UPD_FROM_SUBTREES && (P.top [d] || P.owner [d] != P.rank)) continue ;
        if (dist && which == UPD_FROM_TOP && !P.top [d]) continue ;
        UpdItem it ;
        it.lda = (int32_t) (pi [d+1] - pi [d]) ;
        it.a_off = px [d] + pr.lpos ;
        it.c_off = px [s] ;
        it.ldc = (int32_t) (pi [s+1] - pi [s]) ;
        it.map_off = P.pair_map [pid] ;
        it.M = pr.ndrow2 ; it.N = pr.ndrow1 ; it.K = (int32_t) (super [d+1] - super [d]) ;
        it.shift = 0 ;
        if (!split) { add_item (out, it, sink, use_dmma) ; continue ; }
        const int32_t *map = P.relmap.data () + it.map_off ;    // map [j] < nscol (s) for j < N
        for (int64_t ja = 0 ; ja < it.N ; )
        {
            const int64_t T = map [ja] / OUTER_NB ;
            int64_t jb = ja + 1 ;
            while (jb < it.N && map [jb] / OUTER_NB == T) jb++ ;
            if (own.mine (s, T))
            {
                UpdItem sl = it ;
                sl.a_off = it.a_off + ja ; sl.map_off = it.map_off + ja ;
                sl.M = (int32_t) (it.M - ja) ; sl.N = (int32_t) (jb - ja) ;
                add_item (out, sl, sink, use_dmma) ;
            }
            ja = jb ;
        }
    }
}

// Blocked dpotrf + dtrsm of a set of independent supernodes (SNW:1023-1181), batched step by step.
// ncol_eff [q] = number of leading columns of supernode S [q] to factorize (its nscol normally).
//
// Two-level blocking: an outer block is OUTER_NB columns, factorized in POTRF_NB-wide steps
// (potrf of the diagonal block, trsm of the rows below, "inner" update of the rest of the outer block,
// K = POTRF_NB); when an outer block T is done its "outer" update (K = OUTER_NB) goes to all later
// columns.  With use_dmma, blocks wider than 16 columns also produce their inverse and the rows below
// them are multiplied by it on the tensor pipe (K_TRSM_INV*); narrower blocks use substitution.
//
// Look-ahead: the outer update of block T is split into Ua (columns of outer block T+1, main stream) and
// Ub (all later columns, look-ahead stream).  The panel chain F(T+1) = {potrf, trsm, inner updates} then
// runs on the main stream while Ub(T) fills the machine:
//     main:        F(T)  [bcast T]  rec eF(T)   Ua(T)   [wait eB(T-1)]  F(T+1)  rec eF(T+1)  Ua(T+1) ...
//     look-ahead:                   [wait eF(T)]  Ub(T)  rec eB(T)       [wait eF(T+1)]  Ub(T+1) ...
// Ub(T) and Ua(T+1) may touch the same columns concurrently; both only subtract with atomics.
//
// Near / far (sf >= 0, supernodes with at least FAR_MIN_ROWS rows beyond the next two diagonal blocks): what
// the NEXT diagonal block waits for is only the work on the rows close to the diagonal, so only that stays on
// the main stream.  With D = the diagonal block of T, N1 = the 256 rows below it, far = all rows below N1:
//     main:   per step  potrf, trsm of the rows of D u N1, inner update of D u N1, rec eS(step) ;
//             after the last step  rec eN(T) ;  Ua_near(T) = next diagonal block only ;  F(T+1) starts, its first
//             trsm waits for eM(T)
//     far:    per step  [wait eS(step)]  trsm of the far rows, inner update of the far rows ;  rec eFar(T) ;
//             [wait eN(T)]  Ua_mid(T) = rows N1 of block T+1, rec eM(T) ;  Ua_far(T) = its far rows
//     look-ahead:  [wait eN(T), eFar(T)]  Ub(T)
// The far stream works through the tall, throughput-bound launches while the main stream is already on the next
// diagonal block.  An update of a block of rows far below the diagonal is a rectangular product: an UpdItem with
// shift > 0.  Inverses of the diagonal blocks of such supernodes live in slots that rotate over two outer blocks,
// because a far trsm may still read the inverse of step j while the main stream factorizes the steps after it.
//
// Distributed (top) supernodes: block T is factorized by its owner and then broadcast (K_BCAST, issued
// by every rank in the same order); Ua(T) is done by the owner of block T+1 and Ub(T) as one slab per
// later block by that block's owner -- a 1-D block-cyclic right-looking Cholesky.
constexpr int64_t FAR_MIN_ROWS = 1024 ;     // (CHOLMOD_B200_FAR_MIN overrides it: the CPU tests use small matrices)

void plan_factor (Plan &out, const Own &own, const int64_t *super, const int64_t *pi, const int64_t *px,
    const int32_t *S, const int64_t *ncol_eff, int64_t ns, int level, TileSink &sink, bool use_dmma,
    char *copied = nullptr, int sm = 0, int sl = 1, int32_t slot_base = 0, const int64_t *hpx = nullptr, int sf = -1)
{
    // sm / sl: the stream of the panel chain and the look-ahead stream of this call (0 / 1 unless the level
    // is being factorized group by group, see build_plan); sf: the far stream, or -1 (no near / far split);
    // slot_base: first inverse slot it may use
    // px: arena offsets (Plan::dpx); hpx: offsets in the caller's L->x (the copy-out needs both; NULL: the same)
    if (!hpx) hpx = px ;
    if (!use_dmma) sf = -1 ;
    int64_t far_min = FAR_MIN_ROWS ;
    if (const char *e = getenv ("CHOLMOD_B200_FAR_MIN")) far_min = std::max (1, atoi (e)) ;
    int64_t maxcol = 0 ;
    for (int64_t q = 0 ; q < ns ; q++) maxcol = std::max (maxcol, ncol_eff [q]) ;
    // far_from (q, J0): first far row of supernode q while outer block J0 is factorized, or nsrow (no far part)
    auto far_from = [&] (int64_t q, int64_t J0) -> int64_t
    {
        const int64_t s = S [q], nsrow = pi [s+1] - pi [s] ;
        const int64_t F = J0 + 2 * OUTER_NB ;
        if (sf < 0 || ncol_eff [q] <= J0 || nsrow - F < far_min || !own.mine (s, J0 / OUTER_NB)) return nsrow ;
        return F ;
    } ;
    // supernodes that ever have a far part get rotating inverse slots: 2 outer blocks x (OUTER_NB / POTRF_NB) steps
    constexpr int ROT = 2 * (OUTER_NB / POTRF_NB) ;
    std::vector<int32_t> farq ((size_t) ns, -1) ;
    int32_t nfarq = 0 ;
    for (int64_t q = 0 ; q < ns ; q++)
        for (int64_t J0 = 0 ; J0 < ncol_eff [q] ; J0 += OUTER_NB)
            if (far_from (q, J0) < pi [S [q] + 1] - pi [S [q]]) { farq [q] = nfarq++ ; break ; }
    const int32_t step_slot_base = slot_base + ROT * nfarq ;
    int64_t stepctr = 0 ;

    std::vector<PotrfBlk> pb [3] ;      // jb <= 16, <= 32, <= 64
    std::vector<TrsmTile> tt [3], ttf [3] ;     // substitution, inverse 32, inverse 64; near and far rows
    TileSink sinkF ;                    // far updates of the current step / far parts of Ua
    int evB_prev = -1, evB_prev2 = -1 ; // eB(T-1), eB(T-2)
    int last_bcast_ev = -1, last_far_ev = -1 ;
    int evM_prev = -1 ;                 // eM(T-1): Ua_mid (T-1) is done (the first trsm of block T waits for it)
    auto push_trsm = [&] (std::vector<TrsmTile> (&lists) [3], int stream, int wait_ev) -> bool
    {
        bool any = false ;
        for (int c = 0 ; c < 3 ; c++)
        {
            if (lists [c].empty ()) continue ;
            Launch L = make_launch (K_TRSM + c, (int64_t) out.trsm.size (), (int64_t) lists [c].size (), level, stream) ;
            if (!any) L.wait_ev = wait_ev ;
            any = true ;
            out.count_kind [L.kind] += L.count ;
            out.trsm.insert (out.trsm.end (), lists [c].begin (), lists [c].end ()) ;
            out.launches.push_back (L) ;
        }
        return any ;
    } ;
    auto join = [&] (int stream, int wait_ev, int rec_ev)
    {
        Launch J = make_launch (K_JOIN, 0, 0, level, stream) ; J.wait_ev = wait_ev ; J.rec_ev = rec_ev ;
        out.launches.push_back (J) ;
    } ;
    for (int64_t J0 = 0 ; J0 < maxcol ; J0 += OUTER_NB)
    {
        const int64_t T = J0 / OUTER_NB ;
        bool block_has_far = false ;
        for (int64_t q = 0 ; q < ns ; q++) if (far_from (q, J0) < pi [S [q] + 1] - pi [S [q]]) block_has_far = true ;
        // ---- F(T): the panel chain of outer block T -------------------------------------------------
        // it reads columns that Ub(T-2) and earlier wrote; Ub(T-1) is only needed by F(T+1)
        int pending_wait = evB_prev2 ;
        evB_prev2 = -1 ;
        int trsm_wait = evM_prev ;      // rows N1 of this block got their Ua on the far stream
        evM_prev = -1 ;
        for (int64_t j0 = J0 ; j0 < std::min (maxcol, J0 + OUTER_NB) ; j0 += POTRF_NB, stepctr++)
        {
            for (auto &v : pb) v.clear () ;
            for (auto &v : tt) v.clear () ;
            for (auto &v : ttf) v.clear () ;
            std::vector<PotrfBlk> small, tiny ;
            int32_t nslots = step_slot_base ;
            bool far_step = false ;
            for (int64_t q = 0 ; q < ns ; q++)
            {
                const int64_t s = S [q] ;
                const int64_t nscol = ncol_eff [q], nsrow = pi [s+1] - pi [s] ;
                if (nscol <= j0 || !own.mine (s, T)) continue ;
                const int64_t jb = std::min<int64_t> (POTRF_NB, nscol - j0) ;
                const int64_t r0 = j0 + jb ;
                const int64_t F = far_from (q, J0) ;        // rows [r0, F) near, [F, nsrow) far
                const int pcls = (jb <= 16) ? 0 : (jb <= 32) ? 1 : 2 ;
                const bool inv = use_dmma && pcls > 0 && r0 < nsrow ;
                PotrfBlk b ;
                b.off = px [s] + j0 + j0 * nsrow ; b.col = super [s] + j0 ;
                b.lda = (int32_t) nsrow ; b.jb = (int32_t) jb ;
                if (j0 == 0 && nscol <= TINY_COLS && nsrow - nscol <= TINY_ROWS)
                {
                    // a whole tiny supernode (most of the leaves): one thread
                    b.slot = -1 ; b.pad = 0 ;
                    tiny.push_back (b) ;
                    out.flops_kind [K_FACTOR_TINY] += (double) jb * jb * jb / 3.0 + (double) (nsrow - nscol) * jb * jb ;
                    continue ;
                }
                if (j0 == 0 && nscol <= 16 && nsrow - nscol <= SMALL_ROWS)
                {
                    // a whole small supernode: potrf and the trsm of its few rows in one kernel
                    b.slot = -1 ; b.pad = 0 ;
                    small.push_back (b) ;
                    out.flops_kind [K_FACTOR_SMALL] += (double) jb * jb * jb / 3.0 + (double) (nsrow - nscol) * jb * jb ;
                    continue ;
                }
                if (!inv) b.slot = -1 ;
                else if (farq [q] >= 0) b.slot = slot_base + (int32_t) (stepctr % ROT) * nfarq + farq [q] ;
                else b.slot = nslots++ ;
                b.pad = 0 ;
                pb [pcls].push_back (b) ;
                out.flops_kind [K_POTRF16 + pcls] += (double) jb * jb * jb / 3.0 ;
                const int tcls = inv ? pcls : 0 ;
                for (int64_t r = r0 ; r < nsrow ; r += TRSM_ROWS)
                {
                    // tiles are cut at F so that no tile straddles the two zones
                    const int64_t rend = (r < F) ? std::min (F, nsrow) : nsrow ;
                    TrsmTile t ;
                    t.diag_off = b.off ; t.row_off = px [s] + r + j0 * nsrow ;
                    t.lda = (int32_t) nsrow ; t.jb = (int32_t) jb ;
                    t.nrows = (int32_t) std::min<int64_t> (TRSM_ROWS, rend - r) ; t.slot = b.slot ;
                    ((r < F) ? tt : ttf) [tcls].push_back (t) ;
                    out.flops_kind [K_TRSM + tcls] += (double) t.nrows * jb * jb ;
                    if (r < F && r + TRSM_ROWS > F) r = F - TRSM_ROWS ;     // the next tile starts at F
                }
                // inner update: columns [r0, Jend) of the current outer block, K = jb
                const int64_t Jend = std::min<int64_t> (nscol, J0 + OUTER_NB) ;
                if (r0 < Jend)
                {
                    UpdItem it ;
                    it.lda = it.ldc = (int32_t) nsrow ; it.map_off = -1 ; it.shift = 0 ;
                    it.a_off = px [s] + r0 + j0 * nsrow ;
                    it.c_off = px [s] + r0 + r0 * nsrow ;
                    it.M = (int32_t) (std::min (F, nsrow) - r0) ; it.N = (int32_t) (Jend - r0) ; it.K = (int32_t) jb ;
                    add_item (out, it, sink, use_dmma) ;
                    if (F < nsrow)
                    {
                        // the far rows: a rectangular block, B = the rows [r0, Jend) of the same columns
                        it.shift = (int32_t) (F - r0) ;
                        it.a_off = px [s] + F + j0 * nsrow ;
                        it.c_off = px [s] + F + r0 * nsrow ;
                        it.M = (int32_t) (nsrow - F) ;
                        add_item (out, it, sinkF, use_dmma) ;
                    }
                }
                if (F < nsrow) far_step = true ;
            }
            out.winv_slots = std::max<int64_t> (out.winv_slots, nslots) ;
            for (int w = 0 ; w < 2 ; w++)
            {
                std::vector<PotrfBlk> &v = w ? small : tiny ;
                if (v.empty ()) continue ;
                Launch L = make_launch (w ? K_FACTOR_SMALL : K_FACTOR_TINY, (int64_t) out.potrf.size (), (int64_t) v.size (), level, sm) ;
                L.wait_ev = pending_wait ; pending_wait = -1 ;
                out.count_kind [L.kind] += L.count ;
                out.potrf.insert (out.potrf.end (), v.begin (), v.end ()) ;
                out.launches.push_back (L) ;
            }
            for (int c = 0 ; c < 3 ; c++)
            {
                if (pb [c].empty ()) continue ;
                Launch L = make_launch (K_POTRF16 + c, (int64_t) out.potrf.size (), (int64_t) pb [c].size (), level, sm) ;
                L.wait_ev = pending_wait ; pending_wait = -1 ;
                out.count_kind [L.kind] += L.count ;
                out.potrf.insert (out.potrf.end (), pb [c].begin (), pb [c].end ()) ;
                out.launches.push_back (L) ;
            }
            if (trsm_wait >= 0) { join (sm, trsm_wait, -1) ; trsm_wait = -1 ; }
            push_trsm (tt, sm, -1) ;
            if (far_step)
            {
                // eS: the inverse of this step and the near rows of its columns are there
                const int evS = out.nevents++ ;
                join (sm, -1, evS) ;
                if (!push_trsm (ttf, sf, evS)) join (sf, evS, -1) ;
                flush_tiles (out, sinkF, level, sf) ;
            }
            flush_tiles (out, sink, level, sm) ;
        }
        if (pending_wait >= 0)
        {
            // this rank had no panel work in block T: the wait still has to be honoured before the
            // broadcast / updates below may overwrite or read what Ub(T-2) touched
            join (sm, pending_wait, -1) ;
            pending_wait = -1 ;
        }
        if (trsm_wait >= 0) { join (sm, trsm_wait, -1) ; trsm_wait = -1 ; }
        int evFar = -1 ;                // far (T) is through on the far stream
        if (block_has_far) { evFar = out.nevents++ ; join (sf, -1, evFar) ; last_far_ev = evFar ; }

        // ---- distributed supernodes: the owner of block T sends the finished panel to everyone -----------
        size_t bcast_first = 0, bcast_count = 0 ;
        bool foreign_ua = false ;       // this rank applies Ua(T) from a panel it has to receive first
        {
            const size_t first = out.bcasts.size () ;
            for (int64_t q = 0 ; q < ns ; q++)
            {
                const int64_t s = S [q] ;
                const int64_t nscol = ncol_eff [q], nsrow = pi [s+1] - pi [s] ;
                if (nscol <= J0 || !own.is_top (s)) continue ;
                const int64_t Jend = std::min<int64_t> (nscol, J0 + OUTER_NB) ;
                BcastSeg b ;
                b.off = px [s] + J0 * nsrow ; b.len = (Jend - J0) * nsrow ;
                b.root = own.block_owner (s, T) ; b.pad = 0 ;
                out.bcasts.push_back (b) ;
            }
            bcast_first = first ; bcast_count = out.bcasts.size () - first ;
        }

        // ---- outer update of block T -------------------------------------------------------------------
        const int64_t Jn = J0 + OUTER_NB ;      // first column of the next outer block
        TileSink later ;
        bool any_b = false, any_ua_far = false ;
        bool ua_reads_far = false ;     // an undivided Ua on the main stream reads far rows of panel T
        for (int64_t q = 0 ; q < ns ; q++)
        {
            const int64_t s = S [q] ;
            const int64_t nscol = ncol_eff [q], nsrow = pi [s+1] - pi [s] ;
            if (nscol <= Jn) continue ;
            UpdItem it ;
            it.lda = it.ldc = (int32_t) nsrow ; it.map_off = -1 ; it.shift = 0 ; it.K = (int32_t) OUTER_NB ;
            // Ua: columns [Jn, Jn2) -- the next panel
            const int64_t Jn2 = std::min<int64_t> (nscol, Jn + OUTER_NB) ;
            if (own.mine (s, T + 1))
            {
                if (own.is_top (s) && !own.mine (s, T)) foreign_ua = true ;
                const int64_t Fn = far_from (q, Jn) ;       // zones of block T+1: near rows end at Fn
                it.a_off = px [s] + Jn + J0 * nsrow ;
                it.c_off = px [s] + Jn + Jn * nsrow ;
                it.N = (int32_t) (Jn2 - Jn) ;
                if (Fn >= nsrow)
                {
                    it.M = (int32_t) (nsrow - Jn) ;
                    add_item (out, it, sink, use_dmma) ;
                    if (far_from (q, J0) < nsrow) ua_reads_far = true ;
                }
                else
                {
                    // Ua_near: the next diagonal block only (main stream); Ua_mid: the rows N1 of block T+1, which
                    // its first trsm needs (far stream, first); Ua_far: its far rows (far stream)
                    it.M = (int32_t) OUTER_NB ;
                    add_item (out, it, sink, use_dmma) ;
                    UpdItem r = it ;
                    r.shift = (int32_t) OUTER_NB ;
                    r.a_off = px [s] + Jn + OUTER_NB + J0 * nsrow ;
                    r.c_off = px [s] + Jn + OUTER_NB + Jn * nsrow ;
                    r.M = (int32_t) (Fn - (Jn + OUTER_NB)) ;
                    add_item (out, r, sinkF, use_dmma) ;        // flushed first: Ua_mid
                    r.shift = (int32_t) (Fn - Jn) ;
                    r.a_off = px [s] + Fn + J0 * nsrow ;
                    r.c_off = px [s] + Fn + Jn * nsrow ;
                    r.M = (int32_t) (nsrow - Fn) ;
                    add_item (out, r, later, use_dmma) ;        // parked in `later` until Ua_mid is flushed: Ua_far
                    any_ua_far = true ;
                }
            }
        }
        // Ua_mid and Ua_far go to the far stream now, after far (T) (in order) and the near rows of panel T
        // (eN, or the broadcast for a panel of another rank); Ub is collected afterwards
        TileSink ua_far ;
        if (any_ua_far) { ua_far = later ; later.clear () ; }
        for (int64_t q = 0 ; q < ns ; q++)
        {
            const int64_t s = S [q] ;
            const int64_t nscol = ncol_eff [q], nsrow = pi [s+1] - pi [s] ;
            if (nscol <= Jn) continue ;
            UpdItem it ;
            it.lda = it.ldc = (int32_t) nsrow ; it.map_off = -1 ; it.shift = 0 ; it.K = (int32_t) OUTER_NB ;
            const int64_t Jn2 = std::min<int64_t> (nscol, Jn + OUTER_NB) ;
            if (Jn2 >= nscol) continue ;
            // Ub: columns [Jn2, nscol) -- everything after the next panel
            if (!own.is_top (s))
            {
                if (!own.mine (s, 0)) continue ;
                it.a_off = px [s] + Jn2 + J0 * nsrow ;
                it.c_off = px [s] + Jn2 + Jn2 * nsrow ;
                it.M = (int32_t) (nsrow - Jn2) ; it.N = (int32_t) (nscol - Jn2) ;
                add_item (out, it, later, use_dmma) ;
                any_b = true ;
            }
            else
            {
                // one slab per later outer block, done by the rank that owns the block
                for (int64_t Jc = Jn2 ; Jc < nscol ; Jc += OUTER_NB)
                {
                    if (!own.mine (s, Jc / OUTER_NB)) continue ;
                    const int64_t Je = std::min<int64_t> (nscol, Jc + OUTER_NB) ;
                    it.a_off = px [s] + Jc + J0 * nsrow ;
                    it.c_off = px [s] + Jc + Jc * nsrow ;
                    it.M = (int32_t) (nsrow - Jc) ; it.N = (int32_t) (Je - Jc) ;
                    add_item (out, it, later, use_dmma) ;
                    any_b = true ;
                }
            }
        }
        // copy-out: the finished panel [J0, Jend) of every wide supernode this rank holds can go to the host
        const size_t copy_first = out.copies.size () ;
        if (copied)
        {
            for (int64_t q = 0 ; q < ns ; q++)
            {
                const int64_t s = S [q] ;
                const int64_t nscol = ncol_eff [q], nsrow = pi [s+1] - pi [s] ;
                if (own.is_top (s))
                {
                    // a block of a top separator goes to the host from the rank that owns (and has just
                    // factorized) it -- every rank's PCIe link carries its share of L->x
                    if (nscol <= J0) continue ;
                    copied [s] = 1 ;
                    if (!own.mine (s, T)) continue ;
                }
                else if (nscol <= OUTER_NB || nscol <= J0 || !own.mine (s, 0)) continue ;
                const int64_t Jend = std::min<int64_t> (nscol, J0 + OUTER_NB) ;
                CopySeg c ; c.off = px [s] + J0 * nsrow ; c.len = (Jend - J0) * nsrow ; c.hoff = hpx [s] + J0 * nsrow ;
                out.copies.push_back (c) ;
                copied [s] = 1 ;
            }
        }
        const bool any_copy = out.copies.size () > copy_first ;
        // eN / evF: the near part of F(T) ends here on the main stream (for a broadcast panel: the panel is here)
        int evF = -1, evB = -1 ;
        if (bcast_count > 0)
        {
            // the panels travel on the communication stream, off the owner's critical path, once both the near
            // part (evP, main stream) and the far part (evFar) are finished; evF says "panel T is on this rank"
            const int evP = out.nevents++ ;
            evF = out.nevents++ ;
            join (sm, -1, evP) ;
            if (evFar >= 0) join (3, evFar, -1) ;
            Launch Bc = make_launch (K_BCAST, (int64_t) bcast_first, (int64_t) bcast_count, level, 3) ;
            Bc.wait_ev = evP ; Bc.rec_ev = evF ;
            out.launches.push_back (Bc) ;
            last_bcast_ev = evF ;
            if (any_b) evB = out.nevents++ ;
        }
        else if (any_b || any_copy || any_ua_far)
        {
            evF = out.nevents++ ; if (any_b) evB = out.nevents++ ;
            if (!out.launches.empty () && out.launches.back ().stream == sm && out.launches.back ().rec_ev < 0)
                out.launches.back ().rec_ev = evF ;
            else join (sm, -1, evF) ;
        }
        if (any_copy)
        {
            // (a broadcast panel is complete at evF; otherwise the far rows are finished at evFar)
            if (evFar >= 0 && bcast_count == 0) join (2, evFar, -1) ;
            Launch C = make_launch (K_COPYOUT, (int64_t) copy_first, (int64_t) (out.copies.size () - copy_first), level, 2) ;
            C.wait_ev = evF ;
            out.launches.push_back (C) ;
        }
        if (ua_reads_far && evFar >= 0) join (sm, evFar, -1) ;
        flush_tiles (out, sink, level, sm, foreign_ua ? evF : -1, -1) ; // Ua(T) / Ua_near(T), main stream
        if (any_ua_far)
        {
            const int evM = out.nevents++ ;
            flush_tiles (out, sinkF, level, sf, evF, evM) ;            // Ua_mid(T)
            flush_tiles (out, ua_far, level, sf) ;                      // Ua_far(T)
            evM_prev = evM ;
            last_far_ev = out.nevents++ ;
            join (sf, -1, last_far_ev) ;
        }
        if (any_b)
        {
            if (evFar >= 0 && bcast_count == 0) join (sl, evFar, -1) ;
            flush_tiles (out, later, level, sl, evF, evB) ;            // Ub(T), look-ahead stream
        }
        evB_prev2 = evB_prev ;
        evB_prev = evB ;
    }
    // the main stream joins the look-ahead, the far and the communication stream before anything else uses these
    // supernodes
    for (int ev : { evB_prev2, evB_prev, last_bcast_ev, last_far_ev, evM_prev })
    {
        if (ev < 0) continue ;
        join (sm, ev, -1) ;
    }
}

// Split the supernodal etree among the ranks: "top" = supernodes whose subtree holds more than
// 1/(4 nranks) of all work (closed under taking ancestors); every maximal subtree below them goes, whole,
// to the currently least loaded rank (largest first).
int partition (Plan &P, const int64_t *super, const int64_t *pi, const int64_t *px)
{
    const int64_t nsuper = P.nsuper ;
    const int R = P.nranks ;
    P.top.assign ((size_t) nsuper, 0) ;
    P.owner.assign ((size_t) nsuper, 0) ;
    P.sroot.resize ((size_t) nsuper) ;
    std::iota (P.sroot.begin (), P.sroot.end (), 0) ;
    if (R <= 1) return B200_OK ;
    std::vector<double> work ((size_t) nsuper, 0.0) ;
    for (int64_t s = 0 ; s < nsuper ; s++)
    {
        double nscol = (double) (super [s+1] - super [s]), nsrow = (double) (pi [s+1] - pi [s]) ;
        work [s] = nscol * nscol * nscol / 3.0 + nscol * nscol * (nsrow - nscol) + nsrow * nscol ;
    }
    for (const Pair &pr : P.pairs)
    {
        double K = (double) (super [pr.d+1] - super [pr.d]) ;
        work [pr.s] += K * pr.ndrow1 * (2.0 * pr.ndrow2 - pr.ndrow1) ;
    }
    std::vector<double> sub (work) ;                // subtree work; children precede parents
    std::vector<int64_t> first ((size_t) nsuper) ;
    std::iota (first.begin (), first.end (), 0) ;
    double total = 0 ;
    for (int64_t s = 0 ; s < nsuper ; s++)
    {
        int64_t p = P.parent [s] ;
        if (p >= 0) { sub [p] += sub [s] ; first [p] = std::min (first [p], first [s]) ; }
        else total += sub [s] ;
    }
    // (CHOLMOD_B200_TOP_DIV: the 4 of "1/(4 nranks) of all work", for experiments)
    double top_div = 4.0 ;
    if (const char *e = getenv ("CHOLMOD_B200_TOP_DIV")) top_div = std::max (0.5, atof (e)) ;
    const double thresh = total / (top_div * R) ;
    for (int64_t s = 0 ; s < nsuper ; s++) P.top [s] = (sub [s] > thresh) ? 1 : 0 ;
    // a subtree's owner also computes the updates of its supernodes into the top separators (its
    // contribution blocks): that work counts towards the subtree's load
    {
        std::vector<double> out ((size_t) nsuper, 0.0) ;
        for (const Pair &pr : P.pairs)
            if (!P.top [pr.d] && P.top [pr.s])
                out [pr.d] += (double) (super [pr.d+1] - super [pr.d]) * pr.ndrow1 * (2.0 * pr.ndrow2 - pr.ndrow1) ;
        for (int64_t s = 0 ; s < nsuper ; s++)
        {
            if (P.top [s]) continue ;
            sub [s] += out [s] ;
            const int64_t p = P.parent [s] ;
            if (p >= 0 && !P.top [p]) out [p] += out [s] ;
        }
    }
    // local subtree roots, largest first, to the least loaded rank
    std::vector<int64_t> roots ;
    for (int64_t s = 0 ; s < nsuper ; s++)
        if (!P.top [s] && (P.parent [s] < 0 || P.top [P.parent [s]])) roots.push_back (s) ;
    std::sort (roots.begin (), roots.end (), [&] (int64_t a, int64_t b) { return sub [a] != sub [b] ? sub [a] > sub [b] : a < b ; }) ;
    std::vector<double> load ((size_t) R, 0.0) ;
    for (int64_t r : roots)
    {
        int g = (int) (std::min_element (load.begin (), load.end ()) - load.begin ()) ;
        load [g] += sub [r] ;
        P.owner [r] = g ;
    }
    int ntop = 0 ;
    P.run.assign ((size_t) nsuper, 1) ;
    for (int64_t s = nsuper - 1 ; s >= 0 ; s--)
    {
        if (P.top [s])
        {
            P.owner [s] = (ntop++) % R ;
            const int64_t nblocks = cdiv (super [s+1] - super [s], OUTER_NB) ;
            P.run [s] = (P.dist_run > 0) ? P.dist_run : (int32_t) std::max<int64_t> (1, std::min<int64_t> (DIST_RUN, nblocks / (3 * R))) ;
            continue ;
        }
        int64_t p = P.parent [s] ;
        if (p >= 0 && !P.top [p]) { P.sroot [s] = P.sroot [p] ; P.owner [s] = P.owner [p] ; }
    }
    // a local subtree must be one contiguous range of supernodes (the factor is postordered)
    for (int64_t r : roots)
        for (int64_t s = first [r] ; s <= r ; s++)
            if (P.top [s] || P.sroot [s] != r) return B200_INVALID ;
    (void) px ;
    return B200_OK ;
}

// Decide whether and where a single-GPU plan forks into subtree groups (see build_plan).
void plan_groups (Plan &P, const int64_t *super, const int64_t *pi)
{
    const int64_t nsuper = P.nsuper ;
    P.ngroups = 1 ; P.fork_lo = P.fork_hi = -1 ;
    P.group.assign ((size_t) nsuper, -1) ;
    int want = 8 ;
    if (const char *e = getenv ("CHOLMOD_B200_GROUPS")) want = atoi (e) ;
    if (P.nranks > 1 || want < 2 || P.nlevels < 4) return ;
    want = std::min (want, 16) ;
    // fork_lo: the first level with fewer than 2048 supernodes; fork_hi: the first level after it with fewer
    // than 8 supernodes or a supernode of 2048 columns or more (from there on a level fills the machine)
    int lo = -1, hi = -1 ;
    for (int lev = 0 ; lev < P.nlevels ; lev++)
    {
        const int64_t cnt = P.level_ptr [lev+1] - P.level_ptr [lev] ;
        int64_t maxcol = 0 ;
        for (int64_t q = P.level_ptr [lev] ; q < P.level_ptr [lev+1] ; q++)
            maxcol = std::max (maxcol, super [P.level_list [q] + 1] - super [P.level_list [q]]) ;
        if (lo < 0) { if (cnt < 2048) lo = lev ; else continue ; }
        if (cnt < 8 || maxcol >= 2048) { hi = lev ; break ; }
    }
    if (lo < 0) return ;
    if (hi < 0) hi = P.nlevels ;
    if (hi - lo < 2) return ;
    // subtree work below fork_hi
    std::vector<double> sub ((size_t) nsuper, 0.0) ;
    for (int64_t s = 0 ; s < nsuper ; s++)
    {
        double nscol = (double) (super [s+1] - super [s]), nsrow = (double) (pi [s+1] - pi [s]) ;
        sub [s] = nscol * nscol * nscol / 3.0 + nscol * nscol * (nsrow - nscol) + nsrow * nscol ;
    }
    for (const Pair &pr : P.pairs)
    {
        double K = (double) (super [pr.d+1] - super [pr.d]) ;
        sub [pr.s] += K * pr.ndrow1 * (2.0 * pr.ndrow2 - pr.ndrow1) ;
    }
    std::vector<int64_t> roots ;
    for (int64_t s = 0 ; s < nsuper ; s++)      // children precede parents
    {
        if (P.level [s] >= hi) continue ;
        const int64_t p = P.parent [s] ;
        if (p >= 0 && P.level [p] < hi) sub [p] += sub [s] ; else roots.push_back (s) ;
    }
    std::sort (roots.begin (), roots.end (), [&] (int64_t a, int64_t b) { return sub [a] != sub [b] ? sub [a] > sub [b] : a < b ; }) ;
    // only subtrees that reach into the forked levels matter for the balance; count them
    int64_t tall = 0 ;
    for (int64_t r : roots) if (P.level [r] >= lo) tall++ ;
    const int G = (int) std::min<int64_t> (want, tall) ;
    if (G < 2) return ;
    std::vector<double> load ((size_t) G, 0.0) ;
    for (int64_t r : roots)
    {
        int g = (int) (std::min_element (load.begin (), load.end ()) - load.begin ()) ;
        load [g] += sub [r] ;
        P.group [r] = g ;
    }
    for (int64_t s = nsuper - 1 ; s >= 0 ; s--)
    {
        if (P.level [s] >= hi || P.group [s] >= 0) continue ;
        P.group [s] = P.group [P.parent [s]] ;      // the parent is below fork_hi too, and already has its group
    }
    P.ngroups = G ; P.fork_lo = lo ; P.fork_hi = hi ;
}

// The distributed plan of rank P.rank.
//
//   arena     the top separators first, at the same offsets on every rank, then this rank's subtrees; other
//             ranks' subtrees are not stored here at all
//   phase A   this rank's subtrees, level by level -- no other rank is involved.  Then, into this rank's own
//             (zero-initialised) copies of the top separators, every update that a supernode of its subtrees
//             sends there: the rank's CONTRIBUTION BLOCKS
//   fan-in    all ranks meet (K_BARRIER); the owner of each 256-column block of a top separator adds the
//             other ranks' contribution blocks to its own, reading them straight out of their arenas (K_PULL)
//   phase B   the top separators, level by level, 1-D block cyclic: updates between top separators by the
//             owners of the target blocks, panel chain by the block's owner, finished panels broadcast
//   copy-out  every rank sends what it owns -- its subtrees after phase A, its blocks of the top separators as
//             they are finished -- to the same host L->x, each over its own PCIe link
int build_dist (Plan &P, const int64_t *super, const int64_t *pi, const int64_t *px, bool use_dmma)
{
    const int64_t nsuper = P.nsuper ;
    const int rank = P.rank, R = P.nranks ;
    const Own own { &P, false } ;
    const int nlevels = P.nlevels ;
    const int64_t last = (P.trunc_s >= 0) ? P.trunc_s : nsuper - 1 ;       // nothing after `last` is computed
    if (R > 32) return B200_TOO_LARGE ;                                      // PullSeg::mask
    auto ncol_of = [&] (int64_t s) -> int64_t
    {
        return (s == P.trunc_s) ? P.trunc_ncol : super [s+1] - super [s] ;
    } ;

    // ---- arena layout -------------------------------------------------------------------------------------
    P.dpx.assign ((size_t) nsuper + 1, -1) ;
    int64_t at = 0 ;
    for (int64_t s = 0 ; s < nsuper ; s++) if (P.top [s]) { P.dpx [s] = at ; at += px [s+1] - px [s] ; }
    P.top_size = at ;
    for (int64_t s = 0 ; s < nsuper ; s++)
        if (!P.top [s] && P.owner [s] == rank) { P.dpx [s] = at ; at += px [s+1] - px [s] ; }
    P.dxsize = at ;
    P.dpx [nsuper] = at ;
    const int64_t *dpx = P.dpx.data () ;

    // ---- who scatters which column of A, and what this rank owns ---------------------------------------------
    P.colmine.assign ((size_t) P.n, 0) ;
    for (int64_t s = 0 ; s < nsuper ; s++)
    {
        const int64_t nscol = super [s+1] - super [s], nsrow = pi [s+1] - pi [s] ;
        if (!P.top [s])
        {
            if (P.owner [s] != rank) continue ;
            for (int64_t k = super [s] ; k < super [s+1] ; k++) P.colmine [k] = 1 ;
            if (!P.owned.empty () && P.owned.back ().hoff + P.owned.back ().len == px [s]
                && P.owned.back ().off + P.owned.back ().len == dpx [s]) P.owned.back ().len += nscol * nsrow ;
            else P.owned.push_back (CopySeg { dpx [s], nscol * nsrow, px [s] }) ;
            continue ;
        }
        for (int64_t J0 = 0 ; J0 < nscol ; J0 += OUTER_NB)
        {
            if (!own.mine (s, J0 / OUTER_NB)) continue ;
            const int64_t Jend = std::min (nscol, J0 + OUTER_NB) ;
            for (int64_t k = super [s] + J0 ; k < super [s] + Jend ; k++) P.colmine [k] = 1 ;
            P.owned.push_back (CopySeg { dpx [s] + J0 * nsrow, (Jend - J0) * nsrow, px [s] + J0 * nsrow }) ;
        }
    }

    TileSink sink ;
    std::vector<int64_t> ncols ;
    std::vector<int32_t> Sg ;
    std::vector<char> copied ((size_t) nsuper, 0) ;
    const bool copy_early = (P.trunc_s < 0) ;       // a truncated (not positive definite) run is copied at its end

    // ---- phase A: this rank's subtrees -----------------------------------------------------------------------
    for (int lev = 0 ; lev < nlevels ; lev++)
    {
        Sg.clear () ;
        for (int64_t q = P.level_ptr [lev] ; q < P.level_ptr [lev+1] ; q++)
        {
            const int32_t s = P.level_list [q] ;
            if (!P.top [s] && P.owner [s] == rank && s <= last && ncol_of (s) > 0) Sg.push_back (s) ;
        }
        if (Sg.empty ()) continue ;
        for (int32_t s : Sg) plan_updates_into (P, P, own, super, pi, dpx, s, sink, use_dmma, UPD_ALL) ;
        flush_tiles (P, sink, lev) ;
        ncols.resize (Sg.size ()) ;
        for (size_t q = 0 ; q < Sg.size () ; q++) ncols [q] = ncol_of (Sg [q]) ;
        plan_factor (P, own, super, pi, dpx, Sg.data (), ncols.data (), (int64_t) Sg.size (), lev, sink, use_dmma,
            nullptr, 0, 1, 0, nullptr, 4) ;
    }
    // the subtrees are final: they can leave for the host while the rest is computed
    if (copy_early)
    {
        const size_t first = P.copies.size () ;
        for (const CopySeg &c : P.owned) if (c.off >= P.top_size) P.copies.push_back (c) ;
        if (P.copies.size () > first)
        {
            const int ev = P.nevents++ ;
            Launch J = make_launch (K_JOIN, 0, 0, 0, 0) ; J.rec_ev = ev ;
            P.launches.push_back (J) ;
            Launch C = make_launch (K_COPYOUT, (int64_t) first, (int64_t) (P.copies.size () - first), 0, 2) ;
            C.wait_ev = ev ;
            P.launches.push_back (C) ;
        }
    }
    // contribution blocks: what this rank's subtrees send into the top separators, into its own copies of them
    std::vector<uint32_t> contrib ((size_t) nsuper, 0) ;        // ranks that have something for top separator s
    int first_top_level = nlevels ;
    for (int64_t s = 0 ; s < nsuper ; s++)
    {
        if (!P.top [s]) continue ;
        first_top_level = std::min<int> (first_top_level, P.level [s]) ;
        if (s > last || ncol_of (s) == 0) continue ;
        for (int64_t t = P.target_ptr [s] ; t < P.target_ptr [s+1] ; t++)
        {
            const int64_t d = P.pairs [P.target_list [t]].d ;
            if (!P.top [d]) contrib [s] |= 1u << P.owner [d] ;
        }
        plan_updates_into (P, P, own, super, pi, dpx, s, sink, use_dmma, UPD_FROM_SUBTREES) ;
    }
    flush_tiles (P, sink, std::max (0, first_top_level - 1)) ;

    // ---- fan-in --------------------------------------------------------------------------------------------------
    // One barrier, then the pulls level by level on a stream of their own (5): the main stream only waits for the
    // pulls of the level it is about to factorize, so the big separators' contribution blocks cross NVLink while
    // the smaller ones below them are already being factorized.
    std::vector<int> ev_pull ((size_t) nlevels, -1) ;
    {
        const int lev0 = std::max (0, std::min (first_top_level, nlevels - 1)) ;
        const int evA = P.nevents++, evM = P.nevents++ ;
        Launch J = make_launch (K_JOIN, 0, 0, lev0, 0) ; J.rec_ev = evA ;
        P.launches.push_back (J) ;
        Launch B = make_launch (K_BARRIER, 0, 0, lev0, 3) ; B.wait_ev = evA ; B.rec_ev = evM ;
        P.launches.push_back (B) ;
        P.nstreams = 6 ;
        bool first_pull = true ;
        for (int lev = 0 ; lev < nlevels ; lev++)
        {
            const size_t first = P.pulls.size () ;
            for (int64_t q = P.level_ptr [lev] ; q < P.level_ptr [lev+1] ; q++)
            {
                const int64_t s = P.level_list [q] ;
                if (!P.top [s] || s > last || ncol_of (s) == 0) continue ;
                const uint32_t others = contrib [s] & ~(1u << rank) ;
                if (!others) continue ;
                int po = 0 ; for (uint32_t m = others ; m ; m &= m - 1) po++ ;
                const int64_t nscol = super [s+1] - super [s], nsrow = pi [s+1] - pi [s] ;
                for (int64_t J0 = 0 ; J0 < nscol ; J0 += OUTER_NB)
                {
                    if (!own.mine (s, J0 / OUTER_NB)) continue ;
                    const int64_t Jend = std::min (nscol, J0 + OUTER_NB) ;
                    PullSeg g ;
                    g.off = dpx [s] + J0 + J0 * nsrow ;        // rows J0.. of columns J0..Jend-1: the lower part
                    g.nrows = (int32_t) (nsrow - J0) ; g.ncols = (int32_t) (Jend - J0) ; g.ld = (int32_t) nsrow ;
                    g.mask = others ;
                    P.pulls.push_back (g) ;
                    P.nvlink_in_values += (int64_t) po * g.nrows * g.ncols ;
                }
            }
            bool has_top = false ;
            for (int64_t q = P.level_ptr [lev] ; q < P.level_ptr [lev+1] ; q++) has_top = has_top || P.top [P.level_list [q]] ;
            if (!has_top) continue ;
            // (a level without anything to pull still takes its turn: its event orders it after the barrier)
            ev_pull [lev] = P.nevents++ ;
            Launch W = make_launch (P.pulls.size () > first ? K_PULL : K_JOIN, (int64_t) first,
                (int64_t) (P.pulls.size () - first), lev, 5) ;
            W.wait_ev = first_pull ? evM : -1 ;
            W.rec_ev = ev_pull [lev] ;
            first_pull = false ;
            P.launches.push_back (W) ;
        }
        if (first_pull)
        {
            // no top separator at all (cannot happen with nranks > 1 and a non-empty factor, but keep the barrier joined)
            Launch W = make_launch (K_JOIN, 0, 0, lev0, 0) ; W.wait_ev = evM ;
            P.launches.push_back (W) ;
        }
    }

    // ---- phase B: the top separators ---------------------------------------------------------------------------
    for (int lev = 0 ; lev < nlevels ; lev++)
    {
        Sg.clear () ;
        for (int64_t q = P.level_ptr [lev] ; q < P.level_ptr [lev+1] ; q++)
        {
            const int32_t s = P.level_list [q] ;
            if (P.top [s] && s <= last && ncol_of (s) > 0) Sg.push_back (s) ;
        }
        if (Sg.empty ()) continue ;
        if (ev_pull [lev] >= 0)
        {
            Launch W = make_launch (K_JOIN, 0, 0, lev, 0) ; W.wait_ev = ev_pull [lev] ;
            P.launches.push_back (W) ;
        }
        for (int32_t s : Sg) plan_updates_into (P, P, own, super, pi, dpx, s, sink, use_dmma, UPD_FROM_TOP) ;
        flush_tiles (P, sink, lev) ;
        ncols.resize (Sg.size ()) ;
        for (size_t q = 0 ; q < Sg.size () ; q++) ncols [q] = ncol_of (Sg [q]) ;
        plan_factor (P, own, super, pi, dpx, Sg.data (), ncols.data (), (int64_t) Sg.size (), lev, sink, use_dmma,
            copy_early ? copied.data () : nullptr, 0, 1, 0, px, 4) ;
    }
    for (const BcastSeg &b : P.bcasts) if (b.root != rank) P.nvlink_in_values += b.len ;
    if (copy_early)
    {
        // together the segments are exactly what this rank owns
        int64_t a = 0, b = 0 ;
        for (const CopySeg &c : P.copies) a += c.len ;
        for (const CopySeg &c : P.owned) b += c.len ;
        if (a != b) return B200_INVALID ;
    }
    return B200_OK ;
}

}   // namespace

int build_plan (int64_t n, int64_t nsuper, const int64_t *super, const int64_t *pi, const int64_t *px,
    const int64_t *Ls, bool use_dmma, Plan &P, int nranks, int rank, int64_t trunc_s, int64_t trunc_ncol,
    bool host_relmap)
{
    if (nranks > 1) host_relmap = true ;        // the distributed planner cuts updates by the maps' contents
    if (n < 0 || nsuper < 0 || !super || !pi || !px || (!Ls && nsuper > 0)) return B200_INVALID ;
    if (nranks < 1 || rank < 0 || rank >= nranks) return B200_INVALID ;
    if (n >= (int64_t) 1 << 31 || nsuper >= (int64_t) 1 << 31) return B200_TOO_LARGE ;
    P = Plan () ;
    P.n = n ; P.nsuper = nsuper ; P.nranks = nranks ; P.rank = rank ;
    if (trunc_s >= 0 && (nranks == 1 || trunc_s >= nsuper || trunc_ncol < 0)) return B200_INVALID ;
    P.trunc_s = trunc_s ; P.trunc_ncol = trunc_ncol ;
    // CHOLMOD_B200_PLAN_TIMING=1: where the planner's own time goes (stderr)
    const bool timing = getenv ("CHOLMOD_B200_PLAN_TIMING") != nullptr ;
    auto t_last = std::chrono::steady_clock::now () ;
    auto lap = [&] (const char *what)
    {
        if (!timing) return ;
        auto now = std::chrono::steady_clock::now () ;
        fprintf (stderr, "cholmod_b200 plan: %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli> (now - t_last).count ()) ;
        t_last = now ;
    } ;
    if (const char *e = getenv ("CHOLMOD_B200_DIST_RUN")) P.dist_run = std::max (1, atoi (e)) ;
    if (nsuper == 0)
    {
        if (n != 0) return B200_INVALID ;
        P.level_ptr.assign (1, 0) ;
        P.target_ptr.assign (1, 0) ;
        P.dpx.assign (1, 0) ;
        return B200_OK ;
    }
    if (super [0] != 0 || super [nsuper] != n || pi [0] != 0) return B200_INVALID ;
    P.ssize = pi [nsuper] ; P.xsize = px [nsuper] ;
    for (int64_t s = 0 ; s < nsuper ; s++)
    {
        int64_t nscol = super [s+1] - super [s], nsrow = pi [s+1] - pi [s] ;
        if (nscol <= 0 || nsrow < nscol || px [s+1] - px [s] != nscol * nsrow) return B200_INVALID ;
        if (nsrow >= (int64_t) 1 << 30) return B200_TOO_LARGE ;
        // the leading rows of a supernode are its own columns, in order (cholmod_super_symbolic.c:779-782)
        if (Ls [pi [s]] != super [s] || Ls [pi [s] + nscol - 1] != super [s+1] - 1) return B200_INVALID ;
    }

    // column -> supernode
    P.supermap.resize ((size_t) n) ;
    for (int64_t s = 0 ; s < nsuper ; s++)
        for (int64_t k = super [s] ; k < super [s+1] ; k++) P.supermap [k] = (int32_t) s ;

    // supernodal etree and levels (children precede parents: the factor is postordered)
    P.parent.assign ((size_t) nsuper, -1) ;
    P.level.assign ((size_t) nsuper, 0) ;
    for (int64_t s = 0 ; s < nsuper ; s++)
    {
        int64_t nscol = super [s+1] - super [s], nsrow = pi [s+1] - pi [s] ;
        if (nsrow > nscol)
        {
            int64_t r = Ls [pi [s] + nscol] ;
            if (r < super [s+1] || r >= n) return B200_INVALID ;
            int64_t p = P.supermap [r] ;
            P.parent [s] = p ;
            P.level [p] = std::max (P.level [p], P.level [s] + 1) ;
        }
    }
    int32_t nlevels = 0 ;
    for (int64_t s = 0 ; s < nsuper ; s++) nlevels = std::max (nlevels, P.level [s] + 1) ;
    P.nlevels = nlevels ;
    P.level_ptr.assign ((size_t) nlevels + 1, 0) ;
    for (int64_t s = 0 ; s < nsuper ; s++) P.level_ptr [P.level [s] + 1]++ ;
    for (int l = 0 ; l < nlevels ; l++) P.level_ptr [l+1] += P.level_ptr [l] ;
    P.level_list.resize ((size_t) nsuper) ;
    {
        std::vector<int64_t> w (P.level_ptr.begin (), P.level_ptr.end () - 1) ;
        for (int64_t s = 0 ; s < nsuper ; s++) P.level_list [w [P.level [s]]++] = (int32_t) s ;
    }

    lap ("checks, etree, levels") ;
    // enumerate the update pairs (d -> s) by walking the rows of every descendant
    std::vector<Pair> &pairs = P.pairs ;
    std::vector<int64_t> &tcount = P.target_ptr ;
    tcount.assign ((size_t) nsuper + 1, 0) ;
    for (int64_t d = 0 ; d < nsuper ; d++)
    {
        int64_t ndcol = super [d+1] - super [d] ;
        int64_t p = pi [d] + ndcol, pend = pi [d+1] ;
        while (p < pend)
        {
            int64_t r = Ls [p] ;
            if (r < 0 || r >= n) return B200_INVALID ;
            int64_t s = P.supermap [r] ;
            if (s <= d) return B200_INVALID ;
            int64_t k2 = super [s+1], q = p ;
            while (q < pend && Ls [q] < k2) q++ ;
            Pair pr ;
            pr.d = (int32_t) d ; pr.s = (int32_t) s ; pr.lpos = (int32_t) (p - pi [d]) ;
            pr.ndrow1 = (int32_t) (q - p) ; pr.ndrow2 = (int32_t) (pend - p) ;
            pairs.push_back (pr) ;
            tcount [s+1]++ ;
            P.flops_update += (double) ndcol * (double) pr.ndrow1 * (2.0 * pr.ndrow2 - pr.ndrow1) ;
            P.flops_syrk += (double) ndcol * (double) pr.ndrow1 * (double) pr.ndrow1 ;
            P.flops_gemm += 2.0 * (double) ndcol * (double) pr.ndrow1 * (double) (pr.ndrow2 - pr.ndrow1) ;
            if (pr.ndrow2 > pr.ndrow1) P.npairs_gemm++ ;
            p = q ;
        }
    }
    if (pairs.size () >= ((size_t) 1 << 31)) return B200_TOO_LARGE ;
    P.npairs = (int64_t) pairs.size () ;
    for (int64_t s = 0 ; s < nsuper ; s++) tcount [s+1] += tcount [s] ;
    P.target_list.resize (pairs.size ()) ;
    {
        std::vector<int64_t> w (tcount.begin (), tcount.end () - 1) ;
        for (size_t k = 0 ; k < pairs.size () ; k++) P.target_list [w [pairs [k].s]++] = (int32_t) k ;
    }

    lap ("update pairs") ;
    // relative maps: position of every remaining row of d inside the row list of s
    P.pair_map.resize (pairs.size ()) ;
    {
        int64_t total = 0 ;
        for (size_t k = 0 ; k < pairs.size () ; k++) total += pairs [k].ndrow2 ;
        P.relmap_len = total ;
        if (!host_relmap)
        {
            // only where each pair's map will lie (the same layout: by target, pairs of a target in order); the maps
            // themselves are built on the device (symbolic.cuh), which also checks that every row is found
            int64_t off = 0 ;
            for (int64_t s = 0 ; s < nsuper ; s++)
                for (int64_t t = tcount [s] ; t < tcount [s+1] ; t++)
                {
                    P.pair_map [P.target_list [t]] = off ;
                    off += pairs [P.target_list [t]].ndrow2 ;
                }
        }
        else
        {
        P.relmap.resize ((size_t) total) ;
        std::vector<int32_t> Map ((size_t) n, -1), stamp ((size_t) n, -1) ;
        int64_t off = 0 ;
        for (int64_t s = 0 ; s < nsuper ; s++)
        {
            if (tcount [s] == tcount [s+1]) continue ;
            int64_t nsrow = pi [s+1] - pi [s] ;
            for (int64_t k = 0 ; k < nsrow ; k++)
            {
                int64_t r = Ls [pi [s] + k] ;
                if (r < 0 || r >= n) return B200_INVALID ;
                Map [r] = (int32_t) k ; stamp [r] = (int32_t) s ;
            }
            for (int64_t t = tcount [s] ; t < tcount [s+1] ; t++)
            {
                const int32_t pid = P.target_list [t] ;
                const Pair &pr = pairs [pid] ;
                P.pair_map [pid] = off ;
                const int64_t *rows = Ls + pi [pr.d] + pr.lpos ;
                for (int32_t i = 0 ; i < pr.ndrow2 ; i++)
                {
                    int64_t r = rows [i] ;
                    if (stamp [r] != (int32_t) s) return B200_INVALID ;  // pattern of d not inside pattern of s
                    P.relmap [off + i] = Map [r] ;
                }
                off += pr.ndrow2 ;
            }
        }
        }
    }

    lap ("relative maps") ;
    // flop counts of the dense factorizations
    for (int64_t s = 0 ; s < nsuper ; s++)
    {
        double nscol = (double) (super [s+1] - super [s]), nsrow = (double) (pi [s+1] - pi [s]) ;
        P.flops_potrf += nscol * nscol * nscol / 3.0 ;
        P.flops_trsm += nscol * nscol * (nsrow - nscol) ;
        if (nsrow > nscol) P.nsuper_trsm++ ;
    }

    // who owns what
    int st = partition (P, super, pi, px) ;
    if (st != B200_OK) return st ;
    const Own own { &P, false } ;
    if (nranks > 1)
    {
        st = build_dist (P, super, pi, px, use_dmma) ;
        lap ("distributed plan") ;
        if (st != B200_OK) return st ;
        if (P.items.size () >= ((size_t) 1 << 31)) return B200_TOO_LARGE ;
        return B200_OK ;
    }
    // one GPU: the arena is L->x itself
    P.dpx.assign (px, px + nsuper + 1) ;
    P.dxsize = P.xsize ;
    P.owned.push_back (CopySeg { 0, P.xsize, 0 }) ;
    auto avail_level = [&] (int64_t s) -> int { return P.level [s] ; } ;

    // Copy-out schedule: from the first level above which at most 1024 supernodes remain, every level end
    // is a checkpoint at which all maximal runs of finished, not yet copied supernodes (contiguous in L->x:
    // the factor is postordered) are sent to the host; wide supernodes go panel by panel (plan_factor).
    int first_ckpt = nlevels - 1 ;
    for (int lev = 0 ; lev < nlevels ; lev++)
        if (nsuper - P.level_ptr [lev+1] <= 1024) { first_ckpt = lev ; break ; }
    std::vector<char> copied ((size_t) nsuper, 0) ;

    // Single GPU: subtree groups.  A level with thousands of supernodes fills the machine when all of them
    // are advanced in lock-step, and the few huge supernodes at the top are bound by their updates; the
    // levels in between -- tens of supernodes a few hundred to a few thousand columns wide -- are bound by
    // the latency of their panel chains, one level after the other.  Those levels [fork_lo, fork_hi) are
    // therefore factorized by subtree: the subtrees hanging below level fork_hi are dealt to `ngroups` groups
    // (largest first, to the least loaded), every group walks ITS supernodes level by level on its own
    // stream pair, and the chains of different groups overlap.  Nothing crosses between groups (an update
    // only ever goes from a supernode to one of its ancestors), so the only events are a fork and a join.
    lap ("partition, copy-out setup") ;
    plan_groups (P, super, pi) ;
    lap ("subtree groups") ;
    std::vector<int32_t> group_slot_base ((size_t) P.ngroups + 1, 0) ;
    if (P.ngroups > 1)
    {
        // inverse slots: a group never needs more than its supernode count on one level
        std::vector<int32_t> need ((size_t) P.ngroups, 0), cnt ((size_t) P.ngroups, 0) ;
        for (int lev = P.fork_lo ; lev < P.fork_hi ; lev++)
        {
            std::fill (cnt.begin (), cnt.end (), 0) ;
            for (int64_t q = P.level_ptr [lev] ; q < P.level_ptr [lev+1] ; q++) cnt [P.group [P.level_list [q]]]++ ;
            for (int g = 0 ; g < P.ngroups ; g++) need [g] = std::max (need [g], cnt [g]) ;
        }
        for (int g = 0 ; g < P.ngroups ; g++) group_slot_base [g+1] = group_slot_base [g] + need [g] ;
        P.nstreams = 5 + 2 * P.ngroups ;
    }

    // checkpoint after level `lev`: all maximal runs of finished, not yet copied supernodes go to the host
    auto checkpoint = [&] (int lev)
    {
        const size_t copy_first = P.copies.size () ;
        for (int64_t s = 0 ; s < nsuper ; )
        {
            if (copied [s] || avail_level (s) > lev) { s++ ; continue ; }
            int64_t e = s ;
            while (e < nsuper && !copied [e] && avail_level (e) <= lev) copied [e++] = 1 ;
            CopySeg c ; c.off = px [s] ; c.len = px [e] - px [s] ; c.hoff = c.off ;
            P.copies.push_back (c) ;
            s = e ;
        }
        if (P.copies.size () == copy_first) return ;
        const int evL = P.nevents++ ;
        Launch J = make_launch (K_JOIN, 0, 0, lev, 0) ; J.rec_ev = evL ;
        P.launches.push_back (J) ;
        Launch C = make_launch (K_COPYOUT, (int64_t) copy_first, (int64_t) (P.copies.size () - copy_first), lev, 2) ;
        C.wait_ev = evL ;
        P.launches.push_back (C) ;
    } ;

    // the launch plan, level by level: (1) every descendant update whose target lives on the level,
    // (2) the blocked factorization of the level's supernodes, (3) broadcasts of finished local subtrees
    TileSink sink ;
    std::vector<int64_t> ncols ;
    std::vector<int32_t> Sg ;
    P.items.reserve (pairs.size () + (size_t) nsuper / 4 + 1024) ;
    P.tiles.reserve (2 * pairs.size () + 1024) ;
    P.potrf.reserve ((size_t) nsuper + 1024) ;
    for (int lev = 0 ; lev < nlevels ; lev++)
    {
        const int32_t *S = P.level_list.data () + P.level_ptr [lev] ;
        const int64_t ns = P.level_ptr [lev+1] - P.level_ptr [lev] ;
        if (P.ngroups > 1 && lev >= P.fork_lo && lev < P.fork_hi)
        {
            if (lev == P.fork_lo)
            {
                // fork: every group's chain stream starts after what the main stream has done so far
                const int evFork = P.nevents++ ;
                Launch J = make_launch (K_JOIN, 0, 0, lev, 0) ; J.rec_ev = evFork ;
                P.launches.push_back (J) ;
                for (int g = 0 ; g < P.ngroups ; g++)
                {
                    Launch W = make_launch (K_JOIN, 0, 0, lev, 5 + 2 * g) ; W.wait_ev = evFork ;
                    P.launches.push_back (W) ;
                }
            }
            for (int g = 0 ; g < P.ngroups ; g++)
            {
                Sg.clear () ;
                for (int64_t q = 0 ; q < ns ; q++) if (P.group [S [q]] == g) Sg.push_back (S [q]) ;
                if (Sg.empty ()) continue ;
                const int sm = 5 + 2 * g, sl = 6 + 2 * g ;
                for (int32_t sn : Sg) plan_updates_into (P, P, own, super, pi, px, sn, sink, use_dmma) ;
                flush_tiles (P, sink, lev, sm) ;
                ncols.resize (Sg.size ()) ;
                for (size_t q = 0 ; q < Sg.size () ; q++) ncols [q] = super [Sg [q] + 1] - super [Sg [q]] ;
                plan_factor (P, own, super, pi, px, Sg.data (), ncols.data (), (int64_t) Sg.size (), lev, sink, use_dmma,
                    nullptr, sm, sl, group_slot_base [g]) ;
            }
            if (lev == P.fork_hi - 1)
            {
                // join: the main stream goes on when every group is through
                for (int g = 0 ; g < P.ngroups ; g++)
                {
                    const int evDone = P.nevents++ ;
                    Launch J = make_launch (K_JOIN, 0, 0, lev, 5 + 2 * g) ; J.rec_ev = evDone ;
                    P.launches.push_back (J) ;
                    Launch W = make_launch (K_JOIN, 0, 0, lev, 0) ; W.wait_ev = evDone ;
                    P.launches.push_back (W) ;
                }
            }
            continue ;      // what these levels finished goes to the host at the first checkpoint after the join
        }
        for (int64_t q = 0 ; q < ns ; q++) plan_updates_into (P, P, own, super, pi, px, S [q], sink, use_dmma) ;
        flush_tiles (P, sink, lev) ;
        ncols.resize ((size_t) ns) ;
        for (int64_t q = 0 ; q < ns ; q++) ncols [q] = super [S [q] + 1] - super [S [q]] ;
        const bool ckpt = lev >= first_ckpt ;
        plan_factor (P, own, super, pi, px, S, ncols.data (), ns, lev, sink, use_dmma, ckpt ? copied.data () : nullptr,
            0, 1, 0, nullptr, 4) ;
        if (ckpt) checkpoint (lev) ;
    }
    // Whatever no checkpoint has sent yet goes now.  This is the only copy-out of a plan whose subtree-group
    // fork runs to the last level (a forest of many equal trees: the forked levels skip the checkpoints).
    checkpoint (nlevels - 1) ;
    {
        // the segments must cover [0, xsize) exactly once: the drop-in calls return the host copy they made
        std::vector<CopySeg> segs (P.copies) ;
        std::sort (segs.begin (), segs.end (), [] (const CopySeg &a, const CopySeg &b) { return a.off < b.off ; }) ;
        int64_t at = 0 ;
        for (const CopySeg &c : segs) { if (c.off != at || c.len <= 0) return B200_INVALID ; at += c.len ; }
        if (at != P.xsize) return B200_INVALID ;
    }
    lap ("items, tiles, launches") ;
    if (P.items.size () >= ((size_t) 1 << 31)) return B200_TOO_LARGE ;
    return B200_OK ;
}

int build_repair_plan (const Plan &P, const int64_t *super, const int64_t *pi, const int64_t *px,
    int64_t s, int64_t nscol_new, bool use_dmma, Plan &R)
{
    R = Plan () ;
    R.n = P.n ; R.nsuper = P.nsuper ; R.xsize = P.xsize ; R.ssize = P.ssize ;
    if (s < 0 || s >= P.nsuper || nscol_new < 0 || nscol_new > super [s+1] - super [s]) return B200_INVALID ;
    // every rank repairs its own (complete) copy of the factor by itself
    const Own own { &P, true } ;
    TileSink sink ;
    plan_updates_into (R, P, own, super, pi, px, s, sink, use_dmma) ;
    flush_tiles (R, sink, P.level [s]) ;
    const int32_t S [1] = { (int32_t) s } ;
    const int64_t nc [1] = { nscol_new } ;
    if (nscol_new > 0) plan_factor (R, own, super, pi, px, S, nc, 1, P.level [s], sink, use_dmma) ;
    return B200_OK ;
}

}   // namespace b200
