// b200_multi.cpp -- one factorization over several GPUs driven from ONE process: the form of the multi-GPU
// path that a cholmod_l_factorize caller can reach (the shim's CHOLMOD_B200_NGPUS).  It is a thin driver over
// the per-rank contexts of b200_create_dist: one context per device, one host thread per context for the
// duration of a call, the same distributed plan and the same kernels as the one-process-per-GPU form.
// Written against the C ABI only.
#include "../../include/cholmod_b200.h"
#include <cstdio>
#include <cstring>
#include <new>
#include <thread>
#include <vector>

struct b200_multi
{
    std::vector<b200_factor *> ctx ;
    int dtype = B200_DOUBLE ;
} ;

namespace {

// run fn (rank) on one host thread per rank and wait; returns the ranks' statuses
template <typename Fn> std::vector<int> on_all_ranks (int n, Fn fn)
{
    std::vector<int> st ((size_t) n, B200_OK) ;
    std::vector<std::thread> th ;
    th.reserve ((size_t) n) ;
    for (int r = 0 ; r < n ; r++) th.emplace_back ([&, r] { st [(size_t) r] = fn (r) ; }) ;
    for (auto &t : th) t.join () ;
    return st ;
}

// the most serious status: any error beats NOT_POSDEF beats OK
int worst (const std::vector<int> &st)
{
    int w = B200_OK ;
    for (int s : st) if (s < 0 && (w >= 0 || s < w)) w = s ;
    if (w < 0) return w ;
    for (int s : st) if (s > w) w = s ;
    return w ;
}

}   // namespace

extern "C" {

b200_multi *b200_multi_create (int64_t n, int64_t nsuper, const int64_t *super, const int64_t *pi, const int64_t *px,
    const int64_t *s, int dtype, int first_device, int ngpus, int *status)
{
    int dummy ; if (!status) status = &dummy ;
    *status = B200_OK ;
    if (ngpus < 1 || first_device < 0 || first_device + ngpus > b200_device_count ()) { *status = B200_GPU_PROBLEM ; return nullptr ; }
    b200_multi *m = new (std::nothrow) b200_multi ;
    if (!m) { *status = B200_OUT_OF_MEMORY ; return nullptr ; }
    m->dtype = dtype ;
    m->ctx.assign ((size_t) ngpus, nullptr) ;
    // every rank builds its share of the common plan and uploads its tables, all at the same time
    std::vector<int> st = on_all_ranks (ngpus, [&] (int r)
    {
        int s1 = B200_OK ;
        m->ctx [(size_t) r] = b200_create_dist (n, nsuper, super, pi, px, s, dtype, first_device + r, ngpus, r, &s1) ;
        return s1 ;
    }) ;
    *status = worst (st) ;
    if (*status == B200_OK && ngpus > 1)
    {
        unsigned char id [128] ;
        *status = b200_comm_unique_id (id) ;
        // ncclCommInitRank of the ranks of one process must be entered concurrently
        if (*status == B200_OK) *status = worst (on_all_ranks (ngpus, [&] (int r) { return b200_comm_init (m->ctx [(size_t) r], id) ; })) ;
        // the ranks reach each other's arenas directly (one address space): exchange the arena addresses
        if (*status == B200_OK)
        {
            std::vector<void *> arena ((size_t) ngpus) ;
            for (int r = 0 ; r < ngpus ; r++) arena [(size_t) r] = b200_peer_arena (m->ctx [(size_t) r]) ;
            *status = worst (on_all_ranks (ngpus, [&] (int r)
            {
                for (int q = 0 ; q < ngpus ; q++)
                {
                    if (q == r) continue ;
                    int s1 = b200_peer_attach (m->ctx [(size_t) r], q, arena [(size_t) q], nullptr) ;
                    if (s1 < 0) return s1 ;
                }
                return (int) B200_OK ;
            })) ;
        }
    }
    if (*status != B200_OK) { b200_multi_destroy (m) ; return nullptr ; }
    return m ;
}

void b200_multi_destroy (b200_multi *m)
{
    if (!m) return ;
    // NCCL communicators of one process are torn down concurrently, like they were set up
    on_all_ranks ((int) m->ctx.size (), [&] (int r) { b200_destroy (m->ctx [(size_t) r]) ; return (int) B200_OK ; }) ;
    delete m ;
}

int b200_multi_ngpus (const b200_multi *m) { return m ? (int) m->ctx.size () : 0 ; }

b200_factor *b200_multi_context (b200_multi *m, int rank)
{
    return (m && rank >= 0 && rank < (int) m->ctx.size ()) ? m->ctx [(size_t) rank] : nullptr ;
}

int b200_multi_super_numeric_d (b200_multi *m, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const double *Ax, double beta, int quick_return_if_not_posdef, double *Lx, int64_t *minor)
{
    if (!m || m->dtype != B200_DOUBLE) return m ? B200_INVALID : B200_GPU_PROBLEM ;
    const int n = (int) m->ctx.size () ;
    std::vector<int64_t> mn ((size_t) n, -1) ;
    // every rank takes A from the caller's arrays over its own PCIe link and writes the parts of L->x it owns
    // into the caller's one host buffer over the same link
    std::vector<int> st = on_all_ranks (n, [&] (int r)
    {
        return b200_super_numeric_d (m->ctx [(size_t) r], Ap, Ai, Anz, Ax, beta, quick_return_if_not_posdef, Lx, &mn [(size_t) r]) ;
    }) ;
    if (minor) *minor = mn [0] ;
    return worst (st) ;
}

int b200_multi_download_factor (b200_multi *m, void *Lx)
{
    if (!m) return B200_GPU_PROBLEM ;
    return worst (on_all_ranks ((int) m->ctx.size (), [&] (int r) { return b200_download_factor (m->ctx [(size_t) r], Lx) ; })) ;
}

int b200_multi_forget_host_buffer (b200_multi *m, const void *Lx)
{
    if (!m) return B200_OK ;
    for (b200_factor *f : m->ctx) b200_forget_host_buffer (f, Lx) ;
    return B200_OK ;
}

}   // extern "C"
