// b200_noalloc.cu -- b200_panel_factor_noalloc_d: the one entry point of the library that allocates nothing.
//
// Who needs it.  Two kinds of callers of this library cannot go through cudaMalloc:
//   * solvers that factorize inside a time step or a control loop with a hard latency budget.  cudaMalloc
//     and cudaFree synchronize the whole device and take an unbounded time, so such codes grab their
//     device memory once at start-up and forbid any allocation afterwards;
//   * applications that have already handed all of HBM to their own pool (the reference's GPU module works
//     this way too: one pool sized by CHOLMOD_GPU_MEM_BYTES at cholmod_l_gpu_allocate time and nothing
//     else, CHOLMOD/GPU/cholmod_gpu.c:426-543).  For them a library-side cudaMalloc either fails or
//     fragments a memory budget they have accounted for to the byte.
// Both can still pass in device buffers they own.  What they cannot supply is the scratch the panel
// kernels need behind the scenes (the inverse of the diagonal block, the pivot flag, the tile
// descriptors); in the normal path b200_create cudaMallocs it.  Here it lives in a file-scope __device__
// array instead, so it is part of the loaded module: every GPU that loads this library holds these
// ~49 KB from load until exit, whether or not the entry point is ever called.
//
// The work is one panel step of the blocked supernodal factorization -- dpotrf of the leading jb x jb block,
// then the dtrsm of every row below it (CHOLMOD/Supernodal/t_cholmod_super_numeric_worker.c:1023, :1175) --
// done by the SAME kernels as the normal path (k_potrf_tile, k_trsm_inv), which read their descriptors and
// write the inverse through pointers into the static array.
#include "../../include/cholmod_b200.h"
#include "plan.hpp"
#include "kernels.cuh"
#include <cstdio>
#include <cstdint>
#include <mutex>

using namespace b200 ;

namespace {

constexpr int     NA_MAX_COLS  = B200_NOALLOC_MAX_COLS ;
constexpr int64_t NA_MAX_ROWS  = B200_NOALLOC_MAX_ROWS ;
constexpr int     NA_MAX_TILES = (int) (NA_MAX_ROWS / TRSM_ROWS) ;
static_assert (NA_MAX_COLS == POTRF_NB, "one diagonal block per call") ;

struct NoAllocScratch
{
    double winv [WINV_LD * WINV_LD] ;       // inverse of the diagonal block (32 KB)
    unsigned long long minor ;              // smallest failing column, ~0 when none
    PotrfBlk blk ;                          // descriptor of the diagonal block
    TrsmTile tiles [NA_MAX_TILES] ;         // descriptors of the 128-row tiles below it (16 KB)
} ;

// The scratch itself.  In host code this name is only a stand-in for the symbol: the device address is
// obtained once per device with cudaGetSymbolAddress (below) and that pointer is what the kernels get.
static __device__ __align__ (256) unsigned char g_noalloc_scratch [sizeof (NoAllocScratch)] ;

// Fills the descriptors for this call and re-arms the pivot flag.  Everything in the scratch that a call
// reads is written by that same call first: the flag and the descriptors here, the inverse by k_potrf_tile
// (all nbk x nbk entries).  Nothing is carried over from the previous call.
__global__ void k_noalloc_setup (NoAllocScratch *s, int64_t lda, int64_t nrows, int jb)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x ;
    if (t == 0)
    {
        s->minor = ~0ull ;
        s->blk.off = 0 ; s->blk.col = 0 ; s->blk.lda = (int32_t) lda ; s->blk.jb = jb ;
        s->blk.slot = 0 ; s->blk.pad = 0 ;
    }
    const int64_t below = nrows - jb ;
    const int64_t r = (int64_t) t * TRSM_ROWS ;
    if (r < below)
    {
        TrsmTile &tl = s->tiles [t] ;
        tl.diag_off = 0 ; tl.row_off = jb + r ;
        tl.lda = (int32_t) lda ; tl.jb = jb ;
        tl.nrows = (int32_t) ((below - r < TRSM_ROWS) ? below - r : TRSM_ROWS) ; tl.slot = 0 ;
    }
}

struct DeviceState { NoAllocScratch *ptr = nullptr ; bool ready = false ; } ;
DeviceState g_state [64] ;
std::mutex g_mutex ;

bool ok (cudaError_t e, const char *what)
{
    if (e == cudaSuccess) return true ;
    fprintf (stderr, "cholmod_b200: b200_panel_factor_noalloc_d: %s: %s\n", what, cudaGetErrorString (e)) ;
    return false ;
}

// First call on a device: look up where the static array lives on THAT device (the address differs from
// device to device, and the host-side symbol is not an address at all) and raise the kernels' dynamic
// shared-memory limits.  Neither call allocates.
int device_setup (int device, NoAllocScratch **out)
{
    std::lock_guard<std::mutex> lock (g_mutex) ;
    DeviceState &st = g_state [device] ;
    if (!st.ready)
    {
        void *p = nullptr ;
        if (!ok (cudaGetSymbolAddress (&p, g_noalloc_scratch), "cudaGetSymbolAddress")) return B200_GPU_PROBLEM ;
        if (!ok (cudaFuncSetAttribute (k_potrf_tile<double>, cudaFuncAttributeMaxDynamicSharedMemorySize,
            (int) potrf_smem_bytes<double> ()), "attr potrf")) return B200_GPU_PROBLEM ;
        if (!ok (cudaFuncSetAttribute (k_trsm_inv<32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
            (int) trsm_inv_smem_bytes<32> ()), "attr trsm32")) return B200_GPU_PROBLEM ;
        if (!ok (cudaFuncSetAttribute (k_trsm_inv<64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
            (int) trsm_inv_smem_bytes<64> ()), "attr trsm64")) return B200_GPU_PROBLEM ;
        st.ptr = reinterpret_cast<NoAllocScratch *> (p) ;
        st.ready = true ;
    }
    *out = st.ptr ;
    return B200_OK ;
}

}   // namespace

extern "C" int b200_panel_factor_noalloc_d (double *d_panel, int64_t lda, int64_t nrows, int64_t jb,
    int64_t *info, void *stream)
{
    if (info) *info = 0 ;
    if (!d_panel || lda < nrows || nrows < jb || jb < 1)
    {
        fprintf (stderr, "cholmod_b200: b200_panel_factor_noalloc_d: invalid argument "
            "(need d_panel != NULL, 1 <= jb <= nrows <= lda)\n") ;
        return B200_INVALID ;
    }
    if (lda > (int64_t) INT32_MAX)
    {
        // the tile descriptors carry the leading dimension as a 32-bit integer
        fprintf (stderr, "cholmod_b200: b200_panel_factor_noalloc_d: lda = %lld exceeds 2^31-1\n", (long long) lda) ;
        return B200_TOO_LARGE ;
    }
    if (jb > NA_MAX_COLS || nrows > NA_MAX_ROWS)
    {
        fprintf (stderr, "cholmod_b200: b200_panel_factor_noalloc_d: panel of %lld x %lld exceeds the fixed "
            "scratch of this entry point (at most %lld rows x %d columns); use b200_create / "
            "b200_super_numeric_d, which size their scratch with cudaMalloc\n",
            (long long) nrows, (long long) jb, (long long) NA_MAX_ROWS, NA_MAX_COLS) ;
        return B200_TOO_LARGE ;
    }
    int device = -1 ;
    if (!ok (cudaGetDevice (&device), "cudaGetDevice") || device < 0 || device >= 64) return B200_GPU_PROBLEM ;
    NoAllocScratch *s = nullptr ;
    int st = device_setup (device, &s) ;
    if (st != B200_OK) return st ;

    cudaStream_t cs = reinterpret_cast<cudaStream_t> (stream) ;
    const int ijb = (int) jb ;
    const int64_t below = nrows - jb ;
    const int ntiles = (int) ((below + TRSM_ROWS - 1) / TRSM_ROWS) ;
    const int nbk = (ijb <= 32) ? 32 : 64 ;
    k_noalloc_setup<<<(ntiles + 255) / 256 + 1, 256, 0, cs>>> (s, lda, nrows, ijb) ;
    k_potrf_tile<double><<<1, 256, potrf_smem_bytes<double> (), cs>>> (&s->blk, d_panel, s->winv, nbk, &s->minor) ;
    if (ntiles > 0)
    {
        if (nbk == 32) k_trsm_inv<32><<<ntiles, 256, trsm_inv_smem_bytes<32> (), cs>>> (s->tiles, d_panel, s->winv) ;
        else           k_trsm_inv<64><<<ntiles, 256, trsm_inv_smem_bytes<64> (), cs>>> (s->tiles, d_panel, s->winv) ;
    }
    if (!ok (cudaGetLastError (), "kernel launch")) return B200_GPU_PROBLEM ;
    unsigned long long minor = ~0ull ;
    if (!ok (cudaMemcpyAsync (&minor, &s->minor, sizeof (minor), cudaMemcpyDeviceToHost, cs), "read pivot flag"))
        return B200_GPU_PROBLEM ;
    if (!ok (cudaStreamSynchronize (cs), "synchronize")) return B200_GPU_PROBLEM ;
    if (minor == ~0ull) return B200_OK ;
    if (info) *info = (int64_t) minor + 1 ;
    return B200_NOT_POSDEF ;
}
