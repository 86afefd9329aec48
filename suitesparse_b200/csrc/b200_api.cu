// b200_api.cu -- the C ABI of include/cholmod_b200.h: context, device memory, and the executor that
// walks the launch plan on one CUDA stream.  No CPU fallback anywhere: without a usable device every
// numeric entry point returns B200_GPU_PROBLEM.
#include "../../include/cholmod_b200.h"
#include "plan.hpp"
#include "kernels.cuh"
#include "symbolic.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>
#include <type_traits>
#include <mutex>
#include <thread>
#include <algorithm>
#include <utility>
#include <map>
#include <dlfcn.h>
#include <nccl.h>       // types only: libnccl is dlopen'ed, so a single-GPU user needs no NCCL at all
#include <cuda.h>       // types only: the few driver entry points of the green-context API are dlsym'ed

using namespace b200 ;

namespace {

// NCCL entry points, resolved at run time from the libnccl.so.2 already in the process (torch's) or the
// system one.  Only the multi-GPU path (b200_comm_*, K_BCAST) touches them.
struct NcclApi
{
    void *handle = nullptr ;
    ncclResult_t (*GetUniqueId) (ncclUniqueId *) = nullptr ;
    ncclResult_t (*CommInitRank) (ncclComm_t *, int, ncclUniqueId, int) = nullptr ;
    ncclResult_t (*CommDestroy) (ncclComm_t) = nullptr ;
    ncclResult_t (*Broadcast) (const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr ;
    ncclResult_t (*AllReduce) (const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr ;
    ncclResult_t (*GroupStart) () = nullptr ;
    ncclResult_t (*GroupEnd) () = nullptr ;
    const char *(*GetErrorString) (ncclResult_t) = nullptr ;
    bool ok = false ;
} ;

NcclApi *nccl_api ()
{
    static NcclApi api ;
    static bool tried = false ;
    static std::mutex lock ;
    std::lock_guard<std::mutex> guard (lock) ;
    if (tried) return api.ok ? &api : nullptr ;
    tried = true ;
    api.handle = dlopen ("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL) ;
    if (!api.handle) api.handle = dlopen ("libnccl.so", RTLD_NOW | RTLD_GLOBAL) ;
    if (!api.handle) { fprintf (stderr, "cholmod_b200: cannot load libnccl.so.2: %s\n", dlerror ()) ; return nullptr ; }
    #define B200_NCCL_SYM(field, name) *(void **) (&api.field) = dlsym (api.handle, name)
    B200_NCCL_SYM (GetUniqueId, "ncclGetUniqueId") ;
    B200_NCCL_SYM (CommInitRank, "ncclCommInitRank") ;
    B200_NCCL_SYM (CommDestroy, "ncclCommDestroy") ;
    B200_NCCL_SYM (Broadcast, "ncclBroadcast") ;
    B200_NCCL_SYM (AllReduce, "ncclAllReduce") ;
    B200_NCCL_SYM (GroupStart, "ncclGroupStart") ;
    B200_NCCL_SYM (GroupEnd, "ncclGroupEnd") ;
    B200_NCCL_SYM (GetErrorString, "ncclGetErrorString") ;
    #undef B200_NCCL_SYM
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.Broadcast && api.AllReduce
        && api.GroupStart && api.GroupEnd && api.GetErrorString ;
    return api.ok ? &api : nullptr ;
}

}   // namespace

namespace {

// Green contexts (CUDA 12.4+): a spatial partition of the SMs with streams of its own.  Used to keep a few SMs free
// of the throughput kernels of a distributed plan, so that the latency-bound kernels of the panel chain never wait
// for a tile of a look-ahead update to drain before they get an SM (reserve_sms_for_chain).  Resolved from the
// libcuda.so.1 the runtime has already loaded; absent or failing, everything runs on ordinary streams.
struct GreenApi
{
    CUresult (*DeviceGetDevResource) (CUdevice, CUdevResource *, CUdevResourceType) = nullptr ;
    CUresult (*DevSmResourceSplitByCount) (CUdevResource *, unsigned int *, const CUdevResource *, CUdevResource *, unsigned int, unsigned int) = nullptr ;
    CUresult (*DevResourceGenerateDesc) (CUdevResourceDesc *, CUdevResource *, unsigned int) = nullptr ;
    CUresult (*GreenCtxCreate) (CUgreenCtx *, CUdevResourceDesc, CUdevice, unsigned int) = nullptr ;
    CUresult (*GreenCtxDestroy) (CUgreenCtx) = nullptr ;
    CUresult (*GreenCtxStreamCreate) (CUstream *, CUgreenCtx, unsigned int, int) = nullptr ;
    bool ok = false ;
} ;

GreenApi *green_api ()
{
    static GreenApi api ;
    static bool tried = false ;
    static std::mutex lock ;
    std::lock_guard<std::mutex> guard (lock) ;
    if (tried) return api.ok ? &api : nullptr ;
    tried = true ;
    void *h = dlopen ("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL) ;
    if (!h) return nullptr ;
    #define B200_CU_SYM(field, name) *(void **) (&api.field) = dlsym (h, name)
    B200_CU_SYM (DeviceGetDevResource, "cuDeviceGetDevResource") ;
    B200_CU_SYM (DevSmResourceSplitByCount, "cuDevSmResourceSplitByCount") ;
    B200_CU_SYM (DevResourceGenerateDesc, "cuDevResourceGenerateDesc") ;
    B200_CU_SYM (GreenCtxCreate, "cuGreenCtxCreate") ;
    B200_CU_SYM (GreenCtxDestroy, "cuGreenCtxDestroy") ;
    B200_CU_SYM (GreenCtxStreamCreate, "cuGreenCtxStreamCreate") ;
    #undef B200_CU_SYM
    api.ok = api.DeviceGetDevResource && api.DevSmResourceSplitByCount && api.DevResourceGenerateDesc
        && api.GreenCtxCreate && api.GreenCtxDestroy && api.GreenCtxStreamCreate ;
    return api.ok ? &api : nullptr ;
}

template <typename T> struct DevBuf
{
    T *p = nullptr ; size_t n = 0 ;
    cudaError_t alloc (size_t count)
    {
        release () ;
        n = count ;
        if (count == 0) return cudaSuccess ;
        cudaError_t e = cudaMalloc ((void **) &p, count * sizeof (T)) ;
        if (e != cudaSuccess) { p = nullptr ; n = 0 ; }
        return e ;
    }
    cudaError_t upload (const std::vector<T> &v, cudaStream_t st)
    {
        cudaError_t e = alloc (v.size ()) ;
        if (e != cudaSuccess || v.empty ()) return e ;
        return cudaMemcpyAsync (p, v.data (), v.size () * sizeof (T), cudaMemcpyHostToDevice, st) ;
    }
    void release () { if (p) cudaFree (p) ; p = nullptr ; n = 0 ; }
    size_t bytes () const { return n * sizeof (T) ; }
} ;

}   // namespace

struct b200_factor
{
    Plan plan ;
    int dtype = B200_DOUBLE ;
    int device = -1 ;
    size_t esize = 8 ;
    std::vector<int64_t> h_super, h_pi, h_px ;
    std::vector<int64_t> h_Ls ;             // distributed contexts keep L->s: the not-positive-definite case re-plans
    // distributed: arena base of every rank as this device sees it (own entry = d_Lx.p); the peers' arenas are
    // peer-mapped (same process) or opened from their IPC handles (other processes)
    PeerPtrs peers = {} ;
    bool peer_ipc [32] = {} ;               // peers.p [r] came from cudaIpcOpenMemHandle and has to be closed
    bool one_process = false ;              // peers attached by address: all ranks are threads of this process
    int peers_attached = 0 ;
    DevBuf<unsigned char> d_colmine ;
    DevBuf<PullSeg> d_pulls ;
    DevBuf<unsigned long long> d_barrier ;  // the word the K_BARRIER all-reduce works on
    double ms_pull = 0 ;                    // profile mode: the fan-in kernel (K_PULL)

    cudaStream_t stream = nullptr ;         // main stream: everything except the look-ahead updates
    cudaStream_t stream2 = nullptr ;        // look-ahead stream (lower priority) of the wide trailing updates
    cudaStream_t stream3 = nullptr ;        // copy stream: early D2H of finished parts of L->x
    cudaStream_t stream4 = nullptr ;        // communication stream: every NCCL call of this context
    cudaStream_t stream5 = nullptr ;        // far stream: panel rows far below the diagonal block (plan_factor)
    CUgreenCtx green = nullptr ;            // the SM partition stream2 and stream5 live in (distributed contexts), or NULL
    int reserved_sms = 0 ;                  // SMs outside that partition: only the other streams' kernels reach them
    std::vector<cudaStream_t> gstreams ;    // subtree-group streams of a single-GPU plan: [2g] chain, [2g+1] look-ahead
    cudaEvent_t ev_comm [2] = {} ;
    void *copy_dst = nullptr ;              // host L->x of the factorization in flight (copy-out), or NULL
    ncclComm_t comm = nullptr ;             // multi-GPU: communicator over the ranks of the plan (b200_comm_init)
    std::vector<cudaEvent_t> plan_ev ;      // events of the plan's look-ahead schedule
    cudaEvent_t ev [8] = {} ;
    // symbolic tables
    DevBuf<int64_t> d_super, d_pi, d_px ;
    DevBuf<int32_t> d_Ls, d_supermap, d_relmap, d_level_list ;
    DevBuf<Pair> d_pairs ;                  // device-built symbolic tables (symbolic.cuh): the pair list ...
    DevBuf<int64_t> d_pair_map ;            // ... and where each pair's relative map lies in d_relmap
    bool device_symbolic = false ;          // d_relmap was built on the device, not uploaded
    float ms_symbolic = -1 ;
    DevBuf<UpdItem> d_items ;
    DevBuf<UpdTile> d_tiles ;
    DevBuf<PotrfBlk> d_potrf ;
    DevBuf<TrsmTile> d_trsm ;
    // numeric
    DevBuf<unsigned char> d_Lx ;
    DevBuf<int64_t> d_Ap, d_Ai, d_Anz ;
    DevBuf<unsigned char> d_Ax ;
    DevBuf<int64_t> d_Fp, d_Fi, d_Fnz ;     // F = A(:,f)^H of the unsymmetric case (A*F is factorized)
    DevBuf<unsigned char> d_Fx ;
    DevBuf<unsigned long long> d_minor ;
    DevBuf<unsigned char> d_winv ;          // scratch: inverses of the diagonal blocks of one launch
    DevBuf<double> d_X, d_Y ;               // solve: right-hand sides / forward results of the wide supernodes
    DevBuf<unsigned char> d_solveW ;        // solve: inverses of the 64 x 64 diagonal blocks of the wide supernodes
    DevBuf<int> d_solve_cnt ;               // solve: finished-tile counters (block x rhs), zero between solves
    bool solve_w_valid = false ;
    double ms_solve_part [4] = {} ;         // profile mode: small fwd, wide fwd, small bwd, wide bwd
    int64_t solve_launch_count [4] = {} ;            // d_solveW belongs to the factor values now in d_Lx
    // solve schedule (built at the first solve)
    // kind 0: one CTA per supernode; 2: 64-column block step of the wide ones (both sweeps); 3, 4: first diagonal solves
    // and forward tiles of a 256-column superstep (forward sweep only); 5: its backward tiles (backward sweep only)
    struct SolveLaunch { int kind ; int64_t first, count, maxcol ; } ;
    std::vector<SolveLaunch> solve_launches ;
    bool solve_ready = false ;
    DevBuf<int32_t> d_ssmall ;
    DevBuf<SolveBlk> d_sblk ;
    DevBuf<SolveTile> d_stile ;
    DevBuf<SolveTile256> d_stile256 ;
    DevBuf<SolveBlk> d_sstep ;              // solve: the 256-column supersteps { s, c0, nb, first 64 x 64 inverse }
    DevBuf<unsigned char> d_solveW256 ;     // solve: inverses of the supersteps' whole diagonal blocks (or empty)
    int solve_cols = 64 ;                   // columns of a wide supernode per launch: 64 or 256 (CHOLMOD_B200_SOLVE_COLS)
    int solve_pdl = 1 ;                     // superstep kernels launched programmatically dependent (CHOLMOD_B200_SOLVE_PDL)
    std::vector<unsigned char> h_Ax ;       // zomplex input (Ax, Az) interleaved before the upload
    std::vector<unsigned char> h_Fx ;
    bool have_A = false, A_unpacked = false ;
    bool have_F = false, F_unpacked = false ;
    int64_t A_nz = 0 ;
    // host-registered output buffer (pinned once, reused by later factorizations)
    void *reg_ptr = nullptr ; size_t reg_bytes = 0 ;     // the buffer last prepared by ensure_registered
    std::vector<std::pair<void *, size_t>> reg_pieces ; // what this context page-locked of it (cudaHostRegister)

    int64_t kernel_launches = 0 ;
    int64_t nccl_calls = 0 ;
    int profile = 0 ;                       // 1: bracket every launch with events (diagnostic runs only)
    // The launch list of a one-GPU plan can be captured once as a CUDA graph and replayed by every later
    // factorization (the list is static).  One graph per destination of the copy-out (NULL = the factor stays on
    // the device); use_graph: -1 = not decided yet, 0 = plain launches (the default), 1 = graphs.
    // A graph's copy nodes are bound to the page-locked mapping the destination had at capture: reg_gen (bumped by
    // every change of this context's registrations) and `pinned` are compared before a replay.
    struct GraphSlot { void *dst = nullptr ; cudaGraphExec_t exec = nullptr ; int64_t launches = 0 ; uint64_t used = 0 ;
        uint64_t reg_gen = 0 ; bool pinned = false ; } ;
    uint64_t reg_gen = 0 ;
    GraphSlot graphs [2] ;
    int use_graph = -1 ;
    uint64_t graph_clock = 0 ;
    int64_t graph_replays = 0, graph_captures = 0 ;
    cudaEvent_t ev_fork = nullptr ;
    std::vector<cudaEvent_t> ev_join ;
    std::vector<cudaEvent_t> prof_ev ;
    double ms_kind [K_NKINDS] = {} ;
    int64_t launches_kind [K_NKINDS] = {} ;
    float ms_level_end [64] = {} ;          // profile mode: end time of each etree level on the device timeline
    std::vector<float> ms_launch, ms_launch_start ;     // profile mode: duration / start time of every plan launch
    float ms_factor = -1, ms_h2d = -1, ms_d2h = -1, ms_scatter = -1, ms_solve = -1 ;
    // H2D staging ring for the caller's pageable A arrays (h2d_staged): STAGE_SLOTS chunks of page-locked memory
    unsigned char *stage = nullptr ;
    cudaEvent_t stage_ev [8] = {} ;
    size_t device_bytes = 0 ;
    bool mem_counted = false ;              // device_bytes is part of g_mem_held [device]
} ;

namespace {

bool cuda_ok (cudaError_t e, const char *what)
{
    if (e == cudaSuccess) return true ;
    fprintf (stderr, "cholmod_b200: %s: %s\n", what, cudaGetErrorString (e)) ;
    return false ;
}

#define CK(call, what) do { if (!cuda_ok ((call), what)) return B200_GPU_PROBLEM ; } while (0)

bool dtype_ok (int dtype)
{
    return dtype == B200_DOUBLE || dtype == B200_SINGLE || dtype == B200_COMPLEX_DOUBLE || dtype == B200_COMPLEX_SINGLE ;
}
bool dtype_complex (int dtype) { return dtype == B200_COMPLEX_DOUBLE || dtype == B200_COMPLEX_SINGLE ; }
// double and complex double have the FP64 tensor pipe (and with it the inverse-based panel solves)
bool dtype_dmma (int dtype) { return dtype == B200_DOUBLE || dtype == B200_COMPLEX_DOUBLE ; }
size_t dtype_esize (int dtype)
{
    return (dtype == B200_DOUBLE) ? 8 : (dtype == B200_SINGLE) ? 4 : (dtype == B200_COMPLEX_DOUBLE) ? 16 : 8 ;
}

// run `call` with T = the entry type of the context
#define B200_DISPATCH(dtype, call) \
    ((dtype) == B200_DOUBLE ? call<double> : (dtype) == B200_SINGLE ? call<float> : \
     (dtype) == B200_COMPLEX_DOUBLE ? call<zdouble> : call<cfloat>)

int set_device (const b200_factor *f)
{
    if (!f || f->device < 0) return B200_GPU_PROBLEM ;
    CK (cudaSetDevice (f->device), "cudaSetDevice") ;
    return B200_OK ;
}

bool g_attr_done [64] = {} ;
std::mutex g_attr_lock ;

// per-device cap on the device memory the contexts of this library may hold together (0 = no cap), and what
// they hold now: the counterpart of the pool cholmod_l_gpu_allocate sizes from Common->maxGpuMemBytes /
// maxGpuMemFraction (CHOLMOD/GPU/cholmod_gpu.c:426-543)
std::mutex g_mem_lock ;
size_t g_mem_limit [64] = {}, g_mem_held [64] = {} ;

int kernel_attrs (int device)
{
    std::lock_guard<std::mutex> guard (g_attr_lock) ;
    if (device < 64 && g_attr_done [device]) return B200_OK ;
    CK (cudaFuncSetAttribute (k_upd_dmma<BIG_TM, BIG_TN, 32, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) dmma_smem_bytes<BIG_TM, BIG_TN> ()), "attr d128") ;
    CK (cudaFuncSetAttribute (k_upd_dmma<64, 64, 32, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) dmma_smem_bytes<64, 64> ()), "attr d64") ;
    CK (cudaFuncSetAttribute (k_trsm_inv<32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) trsm_inv_smem_bytes<32> ()), "attr trsm_inv32") ;
    CK (cudaFuncSetAttribute (k_trsm_inv<64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) trsm_inv_smem_bytes<64> ()), "attr trsm_inv64") ;
    CK (cudaFuncSetAttribute (k_potrf_tile<double>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) potrf_smem_bytes<double> ()), "attr potrf") ;
    CK (cudaFuncSetAttribute (k_potrf_tile<float>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) potrf_smem_bytes<float> ()), "attr potrf") ;
    CK (cudaFuncSetAttribute (k_potrf_tile<zdouble>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) potrf_smem_bytes<zdouble> ()), "attr potrf") ;
    CK (cudaFuncSetAttribute (k_potrf_tile<cfloat>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) potrf_smem_bytes<cfloat> ()), "attr potrf") ;
    CK (cudaFuncSetAttribute (k_upd_zmma<BIG_TM, BIG_TN, 32, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) zmma_smem_bytes<BIG_TM, BIG_TN> ()), "attr z128") ;
    CK (cudaFuncSetAttribute (k_upd_zmma<64, 64, 32, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) zmma_smem_bytes<64, 64> ()), "attr z64") ;
    CK (cudaFuncSetAttribute (k_trsm_zinv<32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) trsm_zinv_smem_bytes<32> ()), "attr trsm_zinv32") ;
    CK (cudaFuncSetAttribute (k_trsm_zinv<64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) trsm_zinv_smem_bytes<64> ()), "attr trsm_zinv64") ;
    CK (cudaFuncSetAttribute (k_upd_cfma<cfloat, BIG_TM, BIG_TN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) cfma_smem_bytes<cfloat, BIG_TM, BIG_TN> ()), "attr c128") ;
    CK (cudaFuncSetAttribute (k_upd_cfma<cfloat, 64, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) cfma_smem_bytes<cfloat, 64, 64> ()), "attr c64") ;
    CK (cudaFuncSetAttribute (k_solve_fwd<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024), "attr fwd") ;
    CK (cudaFuncSetAttribute (k_solve_bwd<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024), "attr bwd") ;
    CK (cudaFuncSetAttribute (k_solve_fwd<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024), "attr fwd") ;
    CK (cudaFuncSetAttribute (k_solve_bwd<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024), "attr bwd") ;
    CK (cudaFuncSetAttribute (k_solve_fwd<zdouble>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024), "attr fwd") ;
    CK (cudaFuncSetAttribute (k_solve_bwd<zdouble>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024), "attr bwd") ;
    CK (cudaFuncSetAttribute (k_solve_fwd<cfloat>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024), "attr fwd") ;
    CK (cudaFuncSetAttribute (k_solve_bwd<cfloat>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024), "attr bwd") ;
    CK (cudaFuncSetAttribute (k_solve_inv<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) solve_inv_smem_bytes<double> ()), "attr inv") ;
    CK (cudaFuncSetAttribute (k_solve_inv<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) solve_inv_smem_bytes<float> ()), "attr inv") ;
    CK (cudaFuncSetAttribute (k_solve_inv<zdouble>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) solve_inv_smem_bytes<zdouble> ()), "attr inv") ;
    CK (cudaFuncSetAttribute (k_solve_inv<cfloat>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) solve_inv_smem_bytes<cfloat> ()), "attr inv") ;
    CK (cudaFuncSetAttribute (k_solve_inv256<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) solve_inv256_smem_bytes<double> ()), "attr inv256") ;
    CK (cudaFuncSetAttribute (k_solve_inv256<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) solve_inv256_smem_bytes<float> ()), "attr inv256") ;
    CK (cudaFuncSetAttribute (k_solve_inv256<zdouble>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) solve_inv256_smem_bytes<zdouble> ()), "attr inv256") ;
    CK (cudaFuncSetAttribute (k_solve_inv256<cfloat>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) solve_inv256_smem_bytes<cfloat> ()), "attr inv256") ;
    if (device < 64) g_attr_done [device] = true ;
    return B200_OK ;
}

// run the launch list [first, last) of a plan on the context's stream
template <typename T>
int run_launches (b200_factor *f, const Plan &P, const UpdItem *items, const UpdTile *tiles,
    const PotrfBlk *potrf, const TrsmTile *trsm, const int32_t *relmap, const PullSeg *pulls = nullptr)
{
    T *Lx = reinterpret_cast<T *> (f->d_Lx.p) ;
    T *winv = reinterpret_cast<T *> (f->d_winv.p) ;
    size_t nev = 0 ;
    while ((int) f->gstreams.size () < P.nstreams - 5)
    {
        // same priorities as the main pair: chains high, look-ahead updates low
        int lo = 0, hi = 0 ;
        CK (cudaDeviceGetStreamPriorityRange (&lo, &hi), "priority range") ;
        cudaStream_t g ;
        CK (cudaStreamCreateWithPriority (&g, cudaStreamNonBlocking, (f->gstreams.size () & 1) ? lo : hi), "group stream") ;
        f->gstreams.push_back (g) ;
    }
    auto stream_of = [&] (int id) -> cudaStream_t
    {
        return (id == 0) ? f->stream : (id == 1) ? f->stream2 : (id == 2) ? f->stream3 : (id == 3) ? f->stream4
            : (id == 4) ? f->stream5 : f->gstreams [(size_t) (id - 5)] ;
    } ;
    while ((int64_t) f->plan_ev.size () < (int64_t) P.nevents)
    {
        cudaEvent_t e ;
        CK (cudaEventCreateWithFlags (&e, cudaEventDisableTiming), "plan event") ;
        f->plan_ev.push_back (e) ;
    }
    if ((size_t) P.winv_slots * WINV_LD * WINV_LD * sizeof (T) > f->d_winv.n) return B200_INVALID ;
    if (f->profile)
    {
        while (f->prof_ev.size () < 2 * P.launches.size ())
        {
            cudaEvent_t e ;
            CK (cudaEventCreate (&e), "profile event") ;
            f->prof_ev.push_back (e) ;
        }
    }
    for (const Launch &L : P.launches)
    {
        if (L.kind == K_COPYOUT)
        {
            // finished pieces of L->x leave for the caller's (pinned) buffer while later levels compute
            if (!f->copy_dst) continue ;
            if (L.wait_ev >= 0) CK (cudaStreamWaitEvent (f->stream3, f->plan_ev [L.wait_ev], 0), "wait event") ;
            for (int64_t c = L.first ; c < L.first + L.count ; c++)
            {
                const CopySeg &seg = P.copies [c] ;
                CK (cudaMemcpyAsync ((char *) f->copy_dst + (size_t) seg.hoff * sizeof (T), Lx + seg.off,
                    (size_t) seg.len * sizeof (T), cudaMemcpyDeviceToHost, f->stream3), "copy-out") ;
            }
            continue ;
        }
        if (L.stream < 0 || L.stream >= std::max (5, P.nstreams)) return B200_INVALID ;
        const cudaStream_t st = stream_of (L.stream) ;
        if (L.wait_ev >= 0) CK (cudaStreamWaitEvent (st, f->plan_ev [L.wait_ev], 0), "wait event") ;
        if (L.kind == K_BCAST)
        {
            // finished panels / subtrees travel from their owner to every other rank over NVLink
            NcclApi *nc = nccl_api () ;
            if (!nc || !f->comm) { fprintf (stderr, "cholmod_b200: distributed plan without b200_comm_init\n") ; return B200_INVALID ; }
            typedef typename num<T>::real R ;
            const ncclDataType_t dt = (sizeof (R) == 8) ? ncclDouble : ncclFloat ;
            constexpr size_t per = sizeof (T) / sizeof (R) ;       // 2 for (re,im) entries
            nc->GroupStart () ;
            for (int64_t c = L.first ; c < L.first + L.count ; c++)
            {
                const BcastSeg &seg = P.bcasts [c] ;
                ncclResult_t r = nc->Broadcast (Lx + seg.off, Lx + seg.off, (size_t) seg.len * per, dt, seg.root, f->comm, st) ;
                if (r != ncclSuccess) { nc->GroupEnd () ; fprintf (stderr, "cholmod_b200: ncclBroadcast: %s\n", nc->GetErrorString (r)) ; return B200_GPU_PROBLEM ; }
            }
            if (nc->GroupEnd () != ncclSuccess) return B200_GPU_PROBLEM ;
            f->nccl_calls += L.count ;
            if (L.rec_ev >= 0) CK (cudaEventRecord (f->plan_ev [L.rec_ev], st), "record event") ;
            continue ;
        }
        if (L.kind == K_BARRIER)
        {
            // all ranks meet: nobody pulls a contribution block before every rank has finished writing its own
            NcclApi *nc = nccl_api () ;
            if (!nc || !f->comm) { fprintf (stderr, "cholmod_b200: distributed plan without b200_comm_init\n") ; return B200_INVALID ; }
            if (nc->AllReduce (f->d_barrier.p, f->d_barrier.p, 1, ncclUint64, ncclSum, f->comm, st) != ncclSuccess) return B200_GPU_PROBLEM ;
            f->nccl_calls++ ;
            if (L.rec_ev >= 0) CK (cudaEventRecord (f->plan_ev [L.rec_ev], st), "record event") ;
            continue ;
        }
        if (L.kind == K_JOIN || L.count <= 0)
        {
            if (L.rec_ev >= 0) CK (cudaEventRecord (f->plan_ev [L.rec_ev], st), "record event") ;
            continue ;
        }
        if (L.kind == K_PULL && (!pulls || f->peers_attached < P.nranks))
        {
            fprintf (stderr, "cholmod_b200: distributed plan without b200_peer_attach for every rank\n") ;
            return B200_INVALID ;
        }
        if (f->profile) cudaEventRecord (f->prof_ev [nev++], st) ;
        switch (L.kind)
        {
            case K_UPD_SIMT:
            {
                int64_t nthreads = (int64_t) L.count * 32 ;
                unsigned grid = (unsigned) ((nthreads + 255) / 256) ;
                k_upd_simt<T><<<grid, 256, 0, st>>> (tiles + L.first, L.count, items, relmap, Lx) ;
                break ;
            }
            case K_UPD_D64:
                if constexpr (std::is_same<T, double>::value)
                    k_upd_dmma<64, 64, 32, 32><<<(unsigned) L.count, 128, dmma_smem_bytes<64, 64> (), st>>>
                        (tiles + L.first, items, relmap, Lx) ;
                else if constexpr (std::is_same<T, float>::value)
                    k_upd_ffma<64, 64><<<(unsigned) L.count, 64, ffma_smem_bytes<64, 64> (), st>>>
                        (tiles + L.first, items, relmap, Lx) ;
                else if constexpr (std::is_same<T, zdouble>::value)
                    k_upd_zmma<64, 64, 32, 16><<<(unsigned) L.count, 256, zmma_smem_bytes<64, 64> (), st>>>
                        (tiles + L.first, items, relmap, Lx) ;
                else
                    k_upd_cfma<T, 64, 64><<<(unsigned) L.count, 256, cfma_smem_bytes<T, 64, 64> (), st>>>
                        (tiles + L.first, items, relmap, Lx) ;
                break ;
            case K_UPD_D128:
                if constexpr (std::is_same<T, double>::value)
                    k_upd_dmma<BIG_TM, BIG_TN, 32, 64><<<(unsigned) L.count, (BIG_TM / 32) * (BIG_TN / 64) * 32,
                        dmma_smem_bytes<BIG_TM, BIG_TN> (), st>>>
                        (tiles + L.first, items, relmap, Lx) ;
                else if constexpr (std::is_same<T, float>::value)
                    k_upd_ffma<BIG_TM, BIG_TN><<<(unsigned) L.count, (BIG_TM / 8) * (BIG_TN / 8),
                        ffma_smem_bytes<BIG_TM, BIG_TN> (), st>>>
                        (tiles + L.first, items, relmap, Lx) ;
                else if constexpr (std::is_same<T, zdouble>::value)
                    k_upd_zmma<BIG_TM, BIG_TN, 32, 32><<<(unsigned) L.count, (BIG_TM / 32) * (BIG_TN / 32) * 32,
                        zmma_smem_bytes<BIG_TM, BIG_TN> (), st>>>
                        (tiles + L.first, items, relmap, Lx) ;
                else
                    k_upd_cfma<T, BIG_TM, BIG_TN><<<(unsigned) L.count, (BIG_TM / 4) * (BIG_TN / 4),
                        cfma_smem_bytes<T, BIG_TM, BIG_TN> (), st>>>
                        (tiles + L.first, items, relmap, Lx) ;
                break ;
            case K_POTRF16:
                k_potrf_reg<T, 16><<<(unsigned) ((L.count + 3) / 4), 64, 0, st>>> (potrf + L.first, L.count, Lx, f->d_minor.p) ;
                break ;
            case K_FACTOR_TINY:
                k_factor_tiny<T><<<(unsigned) ((L.count + 127) / 128), 128, 0, st>>> (potrf + L.first, L.count, Lx, f->d_minor.p) ;
                break ;
            case K_UPD_TINY:
                k_upd_tiny<T><<<(unsigned) ((L.count + 127) / 128), 128, 0, st>>> (tiles + L.first, L.count, items, relmap, Lx) ;
                break ;
            case K_FACTOR_SMALL:
                k_factor_small<T><<<(unsigned) ((L.count + 3) / 4), 64, 0, st>>> (potrf + L.first, L.count, Lx, f->d_minor.p) ;
                break ;
            case K_POTRF32:
                k_potrf_tile<T><<<(unsigned) L.count, 256, potrf_smem_bytes<T> (), st>>> (potrf + L.first, Lx, winv, 32, f->d_minor.p) ;
                break ;
            case K_POTRF64:
                k_potrf_tile<T><<<(unsigned) L.count, 256, potrf_smem_bytes<T> (), st>>> (potrf + L.first, Lx, winv, 64, f->d_minor.p) ;
                break ;
            case K_TRSM:
                k_trsm<T><<<(unsigned) L.count, TRSM_ROWS, 0, st>>> (trsm + L.first, Lx) ;
                break ;
            case K_TRSM_INV32:
                if constexpr (std::is_same<T, double>::value)
                    k_trsm_inv<32><<<(unsigned) L.count, 256, trsm_inv_smem_bytes<32> (), st>>> (trsm + L.first, Lx, winv) ;
                else if constexpr (std::is_same<T, zdouble>::value)
                    k_trsm_zinv<32><<<(unsigned) L.count, 256, trsm_zinv_smem_bytes<32> (), st>>> (trsm + L.first, Lx, winv) ;
                else return B200_INVALID ;
                break ;
            case K_TRSM_INV64:
                if constexpr (std::is_same<T, double>::value)
                    k_trsm_inv<64><<<(unsigned) L.count, 256, trsm_inv_smem_bytes<64> (), st>>> (trsm + L.first, Lx, winv) ;
                else if constexpr (std::is_same<T, zdouble>::value)
                    k_trsm_zinv<64><<<(unsigned) L.count, 256, trsm_zinv_smem_bytes<64> (), st>>> (trsm + L.first, Lx, winv) ;
                else return B200_INVALID ;
                break ;
            case K_PULL:
                k_pull_add<T><<<dim3 ((unsigned) L.count, OUTER_NB), 256, 0, st>>> (pulls + L.first, Lx, f->peers) ;
                break ;
            default:
                return B200_INVALID ;
        }
        f->kernel_launches++ ;
        if (f->profile) cudaEventRecord (f->prof_ev [nev++], st) ;
        if (L.rec_ev >= 0) CK (cudaEventRecord (f->plan_ev [L.rec_ev], st), "record event") ;
    }
    CK (cudaGetLastError (), "kernel launch") ;
    if (f->profile)
    {
        CK (cudaStreamSynchronize (f->stream), "profile sync") ;
        CK (cudaStreamSynchronize (f->stream2), "profile sync") ;
        CK (cudaStreamSynchronize (f->stream4), "profile sync") ;
        CK (cudaStreamSynchronize (f->stream5), "profile sync") ;
        for (cudaStream_t g : f->gstreams) CK (cudaStreamSynchronize (g), "profile sync") ;
        size_t e = 0, li = 0 ;
        f->ms_launch.assign (P.launches.size (), 0.f) ;
        f->ms_launch_start.assign (P.launches.size (), 0.f) ;
        for (const Launch &L : P.launches)
        {
            const size_t idx = li++ ;
            if (L.kind == K_JOIN || L.kind == K_COPYOUT || L.kind == K_BCAST || L.kind == K_BARRIER || L.count <= 0) continue ;
            float ms = 0 ;
            cudaEventElapsedTime (&ms, f->prof_ev [e], f->prof_ev [e+1]) ;
            f->ms_launch [idx] = ms ;
            cudaEventElapsedTime (&f->ms_launch_start [idx], f->prof_ev [0], f->prof_ev [e]) ;
            e += 2 ;
            if (L.kind < K_NKINDS) { f->ms_kind [L.kind] += ms ; f->launches_kind [L.kind]++ ; }
            else if (L.kind == K_PULL) f->ms_pull += ms ;
            // timeline: when (after the first launch of the plan) the last launch of each level ended
            if (L.level >= 0 && L.level < 64)
            {
                float t_end = 0 ;
                cudaEventElapsedTime (&t_end, f->prof_ev [0], f->prof_ev [e-1]) ;
                if (t_end > f->ms_level_end [L.level]) f->ms_level_end [L.level] = t_end ;
            }
        }
    }
    return B200_OK ;
}

template <typename T>
int scatter_A (b200_factor *f, int64_t kfirst, int64_t klast, double beta)
{
    if (klast <= kfirst) return B200_OK ;
    unsigned grid = (unsigned) ((klast - kfirst + 255) / 256) ;
    if (f->have_F)
        k_scatter_AF<T><<<grid, 256, 0, f->stream>>> (kfirst, klast, f->d_Ap.p, f->d_Ai.p,
            f->A_unpacked ? f->d_Anz.p : nullptr, reinterpret_cast<const T *> (f->d_Ax.p), f->d_Fp.p, f->d_Fi.p,
            f->F_unpacked ? f->d_Fnz.p : nullptr, reinterpret_cast<const T *> (f->d_Fx.p), f->d_supermap.p,
            f->d_super.p, f->d_pi.p, f->d_px.p, f->d_Ls.p, (typename num<T>::real) beta, reinterpret_cast<T *> (f->d_Lx.p),
            f->d_colmine.p) ;
    else
    k_scatter_A<T><<<grid, 256, 0, f->stream>>> (kfirst, klast, f->d_Ap.p, f->d_Ai.p,
        f->A_unpacked ? f->d_Anz.p : nullptr, reinterpret_cast<const T *> (f->d_Ax.p), f->d_supermap.p,
        f->d_super.p, f->d_pi.p, f->d_px.p, f->d_Ls.p, (typename num<T>::real) beta, reinterpret_cast<T *> (f->d_Lx.p),
        f->d_colmine.p) ;
    f->kernel_launches++ ;
    CK (cudaGetLastError (), "k_scatter_A") ;
    return B200_OK ;
}

// SNW:1084-1146 on the device: the matrix is not positive definite at global column `minor`.  Zero
// supernode s and everything after it; unless info == 1 or quick return, rebuild supernode s from A and
// its descendants and factorize only its first info-1 columns, leaving its other columns zero.
template <typename T>
int repair_not_posdef (b200_factor *f, int64_t minor, double beta, int quick)
{
    const Plan &P = f->plan ;
    const int64_t s = P.supermap [minor] ;
    const int64_t k1 = f->h_super [s], nscol = f->h_super [s+1] - k1 ;
    const int64_t psx = f->h_px [s], nsrow = f->h_pi [s+1] - f->h_pi [s] ;
    const int64_t info = minor - k1 + 1 ;
    T *Lx = reinterpret_cast<T *> (f->d_Lx.p) ;
    CK (cudaMemsetAsync (Lx + psx, 0, (size_t) (P.xsize - psx) * sizeof (T), f->stream), "memset tail") ;
    if (info == 1 || quick) return B200_OK ;
    const int64_t nscol_new = info - 1 ;
    Plan R ;
    int st = build_repair_plan (P, f->h_super.data (), f->h_pi.data (), f->h_px.data (), s, nscol_new,
        dtype_dmma (f->dtype), R) ;
    if (st) return st ;
    DevBuf<UpdItem> items ; DevBuf<UpdTile> tiles ; DevBuf<PotrfBlk> potrf ; DevBuf<TrsmTile> trsm ;
    bool ok = items.upload (R.items, f->stream) == cudaSuccess && tiles.upload (R.tiles, f->stream) == cudaSuccess
        && potrf.upload (R.potrf, f->stream) == cudaSuccess && trsm.upload (R.trsm, f->stream) == cudaSuccess ;
    if (ok)
    {
        st = scatter_A<T> (f, k1, k1 + nscol, beta) ;
        if (!st) st = run_launches<T> (f, R, items.p, tiles.p, potrf.p, trsm.p, f->d_relmap.p) ;
        if (!st && cudaMemsetAsync (Lx + psx + nsrow * nscol_new, 0, (size_t) (nsrow * (nscol - nscol_new)) * sizeof (T),
            f->stream) != cudaSuccess) st = B200_GPU_PROBLEM ;
        if (cudaStreamSynchronize (f->stream) != cudaSuccess) st = B200_GPU_PROBLEM ;
    }
    else st = B200_OUT_OF_MEMORY ;
    items.release () ; tiles.release () ; potrf.release () ; trsm.release () ;
    return st ;
}

}   // namespace

// -------------------------------------------------------------------------------------------------
// C ABI
// -------------------------------------------------------------------------------------------------

extern "C" {

double b200_dmma_peak_tflops (int device, int iters)
{
    if (!b200_probe (device) || iters <= 0) return -1 ;
    if (cudaSetDevice (device) != cudaSuccess) return -1 ;
    cudaDeviceProp prop ;
    if (cudaGetDeviceProperties (&prop, device) != cudaSuccess) return -1 ;
    double *out = nullptr ;
    const int grid = prop.multiProcessorCount * 4, block = 256 ;
    if (cudaMalloc ((void **) &out, (size_t) grid * block * sizeof (double)) != cudaSuccess) return -1 ;
    cudaEvent_t e0, e1 ;
    cudaEventCreate (&e0) ; cudaEventCreate (&e1) ;
    k_dmma_peak<<<grid, block>>> (out, iters / 4 + 1) ;          // warm-up
    cudaEventRecord (e0) ;
    k_dmma_peak<<<grid, block>>> (out, iters) ;
    cudaEventRecord (e1) ;
    double tf = -1 ;
    if (cudaEventSynchronize (e1) == cudaSuccess)
    {
        float ms = 0 ;
        cudaEventElapsedTime (&ms, e0, e1) ;
        // 16 m8n8k4 DMMAs (512 flops each) per warp and iteration
        double flops = (double) grid * (block / 32) * (double) iters * 16.0 * 512.0 ;
        tf = flops / (ms * 1e-3) / 1e12 ;
    }
    cudaEventDestroy (e0) ; cudaEventDestroy (e1) ; cudaFree (out) ;
    cudaGetLastError () ;
    return tf ;
}

const char *b200_version (void) { return "cholmod_b200 0.1 sm_100a" ; }

int b200_device_count (void)
{
    int n = 0 ;
    if (cudaGetDeviceCount (&n) != cudaSuccess) { cudaGetLastError () ; return 0 ; }
    return n ;
}

int b200_probe (int device)
{
    int n = b200_device_count () ;
    if (device < 0 || device >= n) return 0 ;
    cudaDeviceProp prop ;
    if (cudaGetDeviceProperties (&prop, device) != cudaSuccess) { cudaGetLastError () ; return 0 ; }
    return (prop.major == 10) ? 1 : 0 ;
}

int b200_memorysize (int device, size_t *free_bytes, size_t *total_bytes)
{
    if (!b200_probe (device)) return B200_GPU_PROBLEM ;
    CK (cudaSetDevice (device), "cudaSetDevice") ;
    size_t fr = 0, tot = 0 ;
    CK (cudaMemGetInfo (&fr, &tot), "cudaMemGetInfo") ;
    if (free_bytes) *free_bytes = fr ;
    if (total_bytes) *total_bytes = tot ;
    return B200_OK ;
}

// The pair list and the relative maps, on the device, from the symbolic tables already there (symbolic.cuh).
// Fails with B200_INVALID if the counts disagree with the host planner's or a row of a descendant is missing from
// its ancestor's pattern.
static int build_device_symbolic (b200_factor *f)
{
    const Plan &P = f->plan ;
    cudaStream_t s = f->stream ;
    const int64_t ns = P.nsuper ;
    if (ns == 0 || P.npairs == 0) { f->device_symbolic = true ; return B200_OK ; }
    DevBuf<int64_t> cnt, first, len, off ;
    DevBuf<int32_t> key, key2, idx, order ;
    DevBuf<unsigned char> tmp ;
    DevBuf<int> bad ;
    auto fail = [&] (int st) { cudaGetLastError () ; return st ; } ;
    if (cnt.alloc ((size_t) ns + 1) != cudaSuccess || first.alloc ((size_t) ns + 1) != cudaSuccess
        || bad.alloc (1) != cudaSuccess) return fail (B200_OUT_OF_MEMORY) ;
    CK (cudaEventRecord (f->ev [0], s), "record") ;
    CK (cudaMemsetAsync (cnt.p, 0, ((size_t) ns + 1) * sizeof (int64_t), s), "memset") ;
    CK (cudaMemsetAsync (bad.p, 0, sizeof (int), s), "memset") ;
    const unsigned gd = (unsigned) ((ns + 127) / 128) ;
    k_sym_pair_count<<<gd, 128, 0, s>>> (ns, f->d_super.p, f->d_pi.p, f->d_Ls.p, f->d_supermap.p, cnt.p) ;
    size_t tb = 0 ;
    CK (cub::DeviceScan::ExclusiveSum (nullptr, tb, cnt.p, first.p, (int) (ns + 1), s), "scan size") ;
    if (tmp.alloc (tb + 16) != cudaSuccess) return fail (B200_OUT_OF_MEMORY) ;
    CK (cub::DeviceScan::ExclusiveSum (tmp.p, tb, cnt.p, first.p, (int) (ns + 1), s), "scan") ;
    int64_t np = -1 ;
    CK (cudaMemcpyAsync (&np, first.p + ns, sizeof (np), cudaMemcpyDeviceToHost, s), "npairs") ;
    CK (cudaStreamSynchronize (s), "pair count") ;
    if (np != P.npairs) { fprintf (stderr, "cholmod_b200: device pair count %lld, host %lld\n", (long long) np, (long long) P.npairs) ; return B200_INVALID ; }
    if (f->d_pairs.alloc ((size_t) np) != cudaSuccess || f->d_pair_map.alloc ((size_t) np) != cudaSuccess
        || key.alloc ((size_t) np) != cudaSuccess || key2.alloc ((size_t) np) != cudaSuccess
        || idx.alloc ((size_t) np) != cudaSuccess || order.alloc ((size_t) np) != cudaSuccess
        || len.alloc ((size_t) np + 1) != cudaSuccess || off.alloc ((size_t) np + 1) != cudaSuccess
        || f->d_relmap.alloc ((size_t) std::max<int64_t> (P.relmap_len, 1)) != cudaSuccess) return fail (B200_OUT_OF_MEMORY) ;
    k_sym_pair_emit<<<gd, 128, 0, s>>> (ns, f->d_super.p, f->d_pi.p, f->d_Ls.p, f->d_supermap.p, first.p, f->d_pairs.p, key.p, idx.p) ;
    // by target, pairs of one target in descendant order (stable): the order the maps are laid out in
    int bits = 1 ; while (((int64_t) 1 << bits) < ns) bits++ ;
    size_t sb = 0 ;
    CK (cub::DeviceRadixSort::SortPairs (nullptr, sb, key.p, key2.p, idx.p, order.p, (int) np, 0, bits, s), "sort size") ;
    if (sb + 16 > tmp.n && tmp.alloc (sb + 16) != cudaSuccess) return fail (B200_OUT_OF_MEMORY) ;
    CK (cub::DeviceRadixSort::SortPairs (tmp.p, sb, key.p, key2.p, idx.p, order.p, (int) np, 0, bits, s), "sort") ;
    CK (cudaMemsetAsync (len.p + np, 0, sizeof (int64_t), s), "memset") ;
    k_sym_gather_len<<<(unsigned) ((np + 255) / 256), 256, 0, s>>> (np, f->d_pairs.p, order.p, len.p) ;
    size_t tb2 = 0 ;
    CK (cub::DeviceScan::ExclusiveSum (nullptr, tb2, len.p, off.p, (int) (np + 1), s), "scan size") ;
    if (tb2 + 16 > tmp.n && tmp.alloc (tb2 + 16) != cudaSuccess) return fail (B200_OUT_OF_MEMORY) ;
    CK (cub::DeviceScan::ExclusiveSum (tmp.p, tb2, len.p, off.p, (int) (np + 1), s), "scan") ;
    k_sym_relmap<<<(unsigned) ((np * 32 + 255) / 256), 256, 0, s>>> (np, f->d_pairs.p, order.p, off.p, f->d_super.p, f->d_pi.p,
        f->d_Ls.p, f->d_pair_map.p, f->d_relmap.p, bad.p) ;
    int64_t total = -1 ; int hbad = 0 ;
    CK (cudaMemcpyAsync (&total, off.p + np, sizeof (total), cudaMemcpyDeviceToHost, s), "relmap length") ;
    CK (cudaMemcpyAsync (&hbad, bad.p, sizeof (hbad), cudaMemcpyDeviceToHost, s), "flag") ;
    CK (cudaEventRecord (f->ev [1], s), "record") ;
    CK (cudaStreamSynchronize (s), "device symbolic") ;
    CK (cudaGetLastError (), "device symbolic kernels") ;
    cudaEventElapsedTime (&f->ms_symbolic, f->ev [0], f->ev [1]) ;
    cnt.release () ; first.release () ; len.release () ; off.release () ; key.release () ; key2.release () ;
    idx.release () ; order.release () ; tmp.release () ; bad.release () ;
    if (total != P.relmap_len || hbad) return B200_INVALID ;
    f->device_symbolic = true ;
    return B200_OK ;
}

static void unregister_pieces (b200_factor *f) ;

static int create_device_side (b200_factor *f, const int64_t *Ls)
{
    const Plan &P = f->plan ;
    if (!b200_probe (f->device)) return B200_GPU_PROBLEM ;
    CK (cudaSetDevice (f->device), "cudaSetDevice") ;
    int st = kernel_attrs (f->device) ;
    if (st != B200_OK) return st ;
    {
        // the panel chain (main stream) gets the higher priority: its small kernels must be scheduled
        // as soon as a slot frees up while a look-ahead update fills the machine
        int lo = 0, hi = 0 ;
        CK (cudaDeviceGetStreamPriorityRange (&lo, &hi), "priority range") ;
        CK (cudaStreamCreateWithPriority (&f->stream, cudaStreamNonBlocking, hi), "stream") ;
        // distributed plans: the throughput streams (look-ahead updates, far rows) get all SMs but a few, which stay
        // free for the panel chain, the broadcasts and the fan-in (CHOLMOD_B200_RESERVE_SMS, default 8 when nranks > 1,
        // 0 otherwise: on one GPU the updates are the bottleneck, not the chain)
        {
            int want = (P.nranks > 1) ? 8 : 0 ;
            if (const char *e = getenv ("CHOLMOD_B200_RESERVE_SMS")) want = atoi (e) ;
            GreenApi *ga = (want > 0) ? green_api () : nullptr ;
            cudaDeviceProp prop ;
            if (ga && cudaGetDeviceProperties (&prop, f->device) == cudaSuccess && want < prop.multiProcessorCount - 16)
            {
                CUdevResource sm, grp, rem ;
                unsigned int n = 1 ;
                CUdevResourceDesc desc = nullptr ;
                CUstream s2 = nullptr, s5 = nullptr ;
                const unsigned keep = (unsigned) (prop.multiProcessorCount - want) ;
                bool ok = ga->DeviceGetDevResource ((CUdevice) f->device, &sm, CU_DEV_RESOURCE_TYPE_SM) == CUDA_SUCCESS
                    && ga->DevSmResourceSplitByCount (&grp, &n, &sm, &rem, 0, keep & ~7u) == CUDA_SUCCESS && n == 1
                    && ga->DevResourceGenerateDesc (&desc, &grp, 1) == CUDA_SUCCESS
                    && ga->GreenCtxCreate (&f->green, desc, (CUdevice) f->device, CU_GREEN_CTX_DEFAULT_STREAM) == CUDA_SUCCESS ;
                if (ok) ok = ga->GreenCtxStreamCreate (&s2, f->green, CU_STREAM_NON_BLOCKING, lo) == CUDA_SUCCESS
                    && ga->GreenCtxStreamCreate (&s5, f->green, CU_STREAM_NON_BLOCKING, (hi < lo - 1) ? hi + 1 : hi) == CUDA_SUCCESS ;
                if (ok)
                {
                    f->stream2 = (cudaStream_t) s2 ; f->stream5 = (cudaStream_t) s5 ;
                    f->reserved_sms = prop.multiProcessorCount - (int) grp.sm.smCount ;
                }
                else
                {
                    if (s2) cudaStreamDestroy ((cudaStream_t) s2) ;
                    if (f->green) { ga->GreenCtxDestroy (f->green) ; f->green = nullptr ; }
                    cudaGetLastError () ;
                }
            }
        }
        if (!f->stream2) CK (cudaStreamCreateWithPriority (&f->stream2, cudaStreamNonBlocking, lo), "stream2") ;
        CK (cudaStreamCreateWithFlags (&f->stream3, cudaStreamNonBlocking), "stream3") ;
        CK (cudaStreamCreateWithPriority (&f->stream4, cudaStreamNonBlocking, hi), "stream4") ;
        // between the panel chain and the look-ahead updates when the device has more than two priority levels
        if (!f->stream5) CK (cudaStreamCreateWithPriority (&f->stream5, cudaStreamNonBlocking, (hi < lo - 1) ? hi + 1 : hi), "stream5") ;
        for (auto &e : f->ev_comm) CK (cudaEventCreateWithFlags (&e, cudaEventDisableTiming), "event") ;
    }
    for (auto &e : f->ev) CK (cudaEventCreate (&e), "event") ;
    {
        // refuse before allocating anything if the context cannot fit under the cap set for this device
        const size_t need = (size_t) (3 * (P.nsuper + 1)) * 8 + (size_t) P.ssize * 4 + (size_t) P.n * 4 + (size_t) P.relmap_len * 4
            + (size_t) P.nsuper * 4 + P.items.size () * sizeof (UpdItem) + P.tiles.size () * sizeof (UpdTile)
            + P.potrf.size () * sizeof (PotrfBlk) + P.trsm.size () * sizeof (TrsmTile) + (size_t) P.dxsize * f->esize
            + (size_t) std::max<int64_t> (P.winv_slots, 1) * WINV_LD * WINV_LD * f->esize ;
        std::lock_guard<std::mutex> guard (g_mem_lock) ;
        const int d = f->device ;
        if (d < 64 && g_mem_limit [d] > 0 && g_mem_held [d] + need > g_mem_limit [d])
        {
            fprintf (stderr, "cholmod_b200: context needs %.1f MB on device %d, over the limit of %.1f MB (%.1f MB held)\n",
                need / 1048576.0, d, g_mem_limit [d] / 1048576.0, g_mem_held [d] / 1048576.0) ;
            return B200_OUT_OF_MEMORY ;
        }
    }
    std::vector<int32_t> Ls32 ((size_t) P.ssize) ;
    for (int64_t k = 0 ; k < P.ssize ; k++) Ls32 [k] = (int32_t) Ls [k] ;
    cudaStream_t s = f->stream ;
    bool ok = true ;
    ok = ok && f->d_super.upload (f->h_super, s) == cudaSuccess ;
    ok = ok && f->d_pi.upload (f->h_pi, s) == cudaSuccess ;
    ok = ok && f->d_px.upload (P.dpx, s) == cudaSuccess ;        // ARENA offsets (== L->px on one GPU)
    ok = ok && f->d_Ls.upload (Ls32, s) == cudaSuccess ;
    ok = ok && f->d_supermap.upload (P.supermap, s) == cudaSuccess ;
    if (!P.relmap.empty () || P.relmap_len == 0) ok = ok && f->d_relmap.upload (P.relmap, s) == cudaSuccess ;
    ok = ok && f->d_level_list.upload (P.level_list, s) == cudaSuccess ;
    ok = ok && f->d_items.upload (P.items, s) == cudaSuccess ;
    ok = ok && f->d_tiles.upload (P.tiles, s) == cudaSuccess ;
    ok = ok && f->d_potrf.upload (P.potrf, s) == cudaSuccess ;
    ok = ok && f->d_trsm.upload (P.trsm, s) == cudaSuccess ;
    // (a distributed arena is never empty: its address is what the other ranks attach to)
    ok = ok && f->d_Lx.alloc ((size_t) (P.nranks > 1 ? std::max<int64_t> (P.dxsize, 1) : P.dxsize) * f->esize) == cudaSuccess ;
    ok = ok && f->d_minor.alloc (1) == cudaSuccess ;
    if (P.nranks > 1)
    {
        ok = ok && f->d_colmine.upload (P.colmine, s) == cudaSuccess ;
        ok = ok && f->d_pulls.upload (P.pulls, s) == cudaSuccess ;
        ok = ok && f->d_barrier.alloc (1) == cudaSuccess ;
        if (ok) ok = cudaMemsetAsync (f->d_barrier.p, 0, sizeof (unsigned long long), s) == cudaSuccess ;
        f->peers.p [P.rank] = f->d_Lx.p ;
        f->peers_attached = 1 ;
    }
    ok = ok && f->d_winv.alloc ((size_t) std::max<int64_t> (P.winv_slots, 1) * WINV_LD * WINV_LD * f->esize) == cudaSuccess ;
    if (!ok) { cudaGetLastError () ; return B200_OUT_OF_MEMORY ; }
    CK (cudaStreamSynchronize (s), "upload symbolic") ;
    if (P.relmap.empty () && P.relmap_len > 0)
    {
        // the relative maps were left out of the host plan: build them here, from the tables just uploaded
        st = build_device_symbolic (f) ;
        if (st != B200_OK) return st ;
    }
    f->device_bytes = f->d_super.bytes () + f->d_pi.bytes () + f->d_px.bytes () + f->d_Ls.bytes ()
        + f->d_supermap.bytes () + f->d_relmap.bytes () + f->d_level_list.bytes () + f->d_items.bytes ()
        + f->d_tiles.bytes () + f->d_potrf.bytes () + f->d_trsm.bytes () + f->d_Lx.bytes () + f->d_winv.bytes ()
        + f->d_colmine.bytes () + f->d_pulls.bytes () + f->d_pairs.bytes () + f->d_pair_map.bytes () ;
    if (f->device < 64)
    {
        std::lock_guard<std::mutex> guard (g_mem_lock) ;
        g_mem_held [f->device] += f->device_bytes ;
        f->mem_counted = true ;
    }
    return B200_OK ;
}

b200_factor *b200_create (int64_t n, int64_t nsuper, const int64_t *super, const int64_t *pi,
    const int64_t *px, const int64_t *s, int dtype, int device, int *status)
{
    return b200_create_dist (n, nsuper, super, pi, px, s, dtype, device, 1, 0, status) ;
}

b200_factor *b200_create_dist (int64_t n, int64_t nsuper, const int64_t *super, const int64_t *pi,
    const int64_t *px, const int64_t *s, int dtype, int device, int nranks, int rank, int *status)
{
    int dummy ; if (!status) status = &dummy ;
    *status = B200_OK ;
    if (!dtype_ok (dtype)) { *status = B200_INVALID ; return nullptr ; }
    b200_factor *f = new (std::nothrow) b200_factor ;
    if (!f) { *status = B200_OUT_OF_MEMORY ; return nullptr ; }
    f->dtype = dtype ; f->esize = dtype_esize (dtype) ; f->device = device ;
    try
    {
        // plan-only contexts (device < 0, the CPU tests of the planner) can ask for the truncated plan the
        // executor builds after a failed pivot (repair_dist): CHOLMOD_B200_TRUNC="<supernode>,<columns>"
        int64_t trunc_s = -1, trunc_ncol = 0 ;
        if (device < 0 && nranks > 1)
            if (const char *e = getenv ("CHOLMOD_B200_TRUNC"))
            {
                long long a = -1, b = 0 ;
                if (sscanf (e, "%lld,%lld", &a, &b) == 2) { trunc_s = a ; trunc_ncol = b ; }
            }
        // the relative maps -- the planner's biggest table -- are built on the device (symbolic.cuh) for a one-GPU
        // context; plan-only contexts, distributed plans (which read the maps) and CHOLMOD_B200_HOST_RELMAP=1 keep
        // the host version
        const bool host_relmap = device < 0 || nranks > 1 || getenv ("CHOLMOD_B200_HOST_RELMAP") != nullptr ;
        *status = build_plan (n, nsuper, super, pi, px, s, dtype_dmma (dtype), f->plan, nranks, rank, trunc_s, trunc_ncol,
            host_relmap) ;
        if (*status == B200_OK)
        {
            f->h_super.assign (super, super + nsuper + 1) ;
            f->h_pi.assign (pi, pi + nsuper + 1) ;
            f->h_px.assign (px, px + nsuper + 1) ;
            if (nranks > 1) f->h_Ls.assign (s, s + pi [nsuper]) ;
        }
    }
    catch (const std::bad_alloc &) { *status = B200_OUT_OF_MEMORY ; }
    if (*status == B200_OK && device >= 0)
    {
        try { *status = create_device_side (f, s) ; }
        catch (const std::bad_alloc &) { *status = B200_OUT_OF_MEMORY ; }
    }
    if (*status != B200_OK) { b200_destroy (f) ; return nullptr ; }
    return f ;
}

int b200_set_memory_limit (int device, size_t bytes)
{
    if (device < 0 || device >= 64) return B200_INVALID ;
    std::lock_guard<std::mutex> guard (g_mem_lock) ;
    g_mem_limit [device] = bytes ;
    return B200_OK ;
}

void b200_destroy (b200_factor *f)
{
    if (!f) return ;
    if (f->mem_counted && f->device >= 0 && f->device < 64)
    {
        std::lock_guard<std::mutex> guard (g_mem_lock) ;
        g_mem_held [f->device] -= std::min (g_mem_held [f->device], f->device_bytes) ;
    }
    if (f->device >= 0 && f->stream)
    {
        cudaSetDevice (f->device) ;
        cudaStreamSynchronize (f->stream) ;
        unregister_pieces (f) ;
        f->d_super.release () ; f->d_pi.release () ; f->d_px.release () ; f->d_Ls.release () ;
        f->d_supermap.release () ; f->d_relmap.release () ; f->d_level_list.release () ;
        f->d_pairs.release () ; f->d_pair_map.release () ;
        f->d_items.release () ; f->d_tiles.release () ; f->d_potrf.release () ; f->d_trsm.release () ;
        f->d_Lx.release () ; f->d_Ap.release () ; f->d_Ai.release () ; f->d_Anz.release () ;
        f->d_Ax.release () ; f->d_Fp.release () ; f->d_Fi.release () ; f->d_Fnz.release () ; f->d_Fx.release () ;
        f->d_minor.release () ; f->d_X.release () ; f->d_Y.release () ; f->d_solveW.release () ; f->d_solve_cnt.release () ;
        f->d_winv.release () ; f->d_colmine.release () ; f->d_pulls.release () ; f->d_barrier.release () ;
        for (int r = 0 ; r < 32 ; r++)
            if (f->peer_ipc [r] && f->peers.p [r]) { cudaIpcCloseMemHandle (f->peers.p [r]) ; f->peers.p [r] = nullptr ; }
        f->d_ssmall.release () ; f->d_sblk.release () ; f->d_stile.release () ; f->d_stile256.release () ; f->d_sstep.release () ; f->d_solveW256.release () ;
        for (auto &e : f->ev) if (e) cudaEventDestroy (e) ;
        if (f->stage) { cudaFreeHost (f->stage) ; for (auto &e : f->stage_ev) if (e) cudaEventDestroy (e) ; }
        for (auto &e : f->prof_ev) cudaEventDestroy (e) ;
        for (auto &e : f->plan_ev) cudaEventDestroy (e) ;
        for (auto &g : f->graphs) if (g.exec) { cudaGraphExecDestroy (g.exec) ; g.exec = nullptr ; }
        if (f->ev_fork) cudaEventDestroy (f->ev_fork) ;
        for (auto &e : f->ev_join) cudaEventDestroy (e) ;
        if (f->comm) { NcclApi *nc = nccl_api () ; if (nc) nc->CommDestroy (f->comm) ; f->comm = nullptr ; }
        cudaStreamSynchronize (f->stream2) ;
        cudaStreamDestroy (f->stream2) ;
        for (cudaStream_t g : f->gstreams) { cudaStreamSynchronize (g) ; cudaStreamDestroy (g) ; }
        if (f->stream3) { cudaStreamSynchronize (f->stream3) ; cudaStreamDestroy (f->stream3) ; }
        if (f->stream4) { cudaStreamSynchronize (f->stream4) ; cudaStreamDestroy (f->stream4) ; }
        if (f->stream5) { cudaStreamSynchronize (f->stream5) ; cudaStreamDestroy (f->stream5) ; }
        for (auto &e : f->ev_comm) if (e) cudaEventDestroy (e) ;
        cudaStreamDestroy (f->stream) ;
        if (f->green) { GreenApi *ga = green_api () ; if (ga) ga->GreenCtxDestroy (f->green) ; f->green = nullptr ; }
        cudaGetLastError () ;
    }
    delete f ;
}

int64_t b200_get_int (const b200_factor *f, const char *name)
{
    if (!f || !name) return -1 ;
    const Plan &P = f->plan ;
    if (!strcmp (name, "n")) return P.n ;
    if (!strcmp (name, "nsuper")) return P.nsuper ;
    if (!strcmp (name, "xsize")) return P.xsize ;
    if (!strcmp (name, "ssize")) return P.ssize ;
    if (!strcmp (name, "nlevels")) return P.nlevels ;
    if (!strcmp (name, "npairs")) return P.npairs ;
    if (!strcmp (name, "npairs_gemm")) return P.npairs_gemm ;
    if (!strcmp (name, "nsuper_trsm")) return P.nsuper_trsm ;
    if (!strcmp (name, "relmap_len")) return P.relmap_len ;
    if (!strcmp (name, "device_symbolic")) return f->device_symbolic ? 1 : 0 ;
    if (!strcmp (name, "nitems")) return (int64_t) P.items.size () ;
    if (!strcmp (name, "ntiles")) return (int64_t) P.tiles.size () ;
    if (!strcmp (name, "npotrf")) return (int64_t) P.potrf.size () ;
    if (!strcmp (name, "ntrsm")) return (int64_t) P.trsm.size () ;
    if (!strcmp (name, "nlaunches")) return (int64_t) P.launches.size () ;
    if (!strcmp (name, "nevents")) return P.nevents ;
    if (!strcmp (name, "winv_slots")) return P.winv_slots ;
    if (!strcmp (name, "device_bytes")) return (int64_t) f->device_bytes ;
    if (!strcmp (name, "device")) return f->device ;
    if (!strcmp (name, "dtype")) return f->dtype ;
    if (!strcmp (name, "kernel_launches")) return f->kernel_launches ;
    if (!strcmp (name, "ngpus")) return P.nranks ;
    if (!strcmp (name, "rank")) return P.rank ;
    if (!strcmp (name, "nccl_calls")) return f->nccl_calls ;
    if (!strcmp (name, "nbcasts")) return (int64_t) P.bcasts.size () ;
    if (!strcmp (name, "npulls")) return (int64_t) P.pulls.size () ;
    if (!strcmp (name, "dxsize")) return P.dxsize ;
    if (!strcmp (name, "top_size")) return P.top_size ;
    if (!strcmp (name, "nvlink_in_bytes")) return P.nvlink_in_values * (int64_t) f->esize ;
    if (!strcmp (name, "peers_attached")) return f->peers_attached ;
    if (!strcmp (name, "reserved_sms")) return f->reserved_sms ;
    if (!strcmp (name, "host_registered")) { int64_t b = 0 ; for (auto &pc : f->reg_pieces) b += (int64_t) pc.second ; return b ; }
    if (!strcmp (name, "ngroups")) return P.ngroups ;
    if (!strcmp (name, "fork_lo")) return P.fork_lo ;
    if (!strcmp (name, "fork_hi")) return P.fork_hi ;
    if (!strcmp (name, "nstreams")) return P.nstreams ;
    if (!strcmp (name, "ntop")) { int64_t c = 0 ; for (char t : P.top) c += t ; return c ; }
    if (!strcmp (name, "profile")) return f->profile ;
    if (!strncmp (name, "launches_kind", 13) && name [13] >= '0' && name [13] <= '9' && atoi (name + 13) < K_NKINDS) return f->launches_kind [atoi (name + 13)] ;
    if (!strncmp (name, "tiles_kind", 10) && name [10] >= '0' && name [10] <= '9' && atoi (name + 10) < K_NKINDS) return P.count_kind [atoi (name + 10)] ;
    return -1 ;
}

int b200_set_int (b200_factor *f, const char *name, int64_t value)
{
    if (!f || !name) return B200_INVALID ;
    if (!strcmp (name, "profile")) { f->profile = value ? 1 : 0 ; return B200_OK ; }
    if (!strcmp (name, "use_graph")) { f->use_graph = value ? 1 : 0 ; return B200_OK ; }
    return B200_INVALID ;
}

double b200_get_double (const b200_factor *f, const char *name)
{
    if (!f || !name) return -1 ;
    const Plan &P = f->plan ;
    if (!strcmp (name, "flops_executed")) return P.flops_update + P.flops_potrf + P.flops_trsm ;
    if (!strcmp (name, "flops_update")) return P.flops_update ;
    if (!strcmp (name, "flops_potrf")) return P.flops_potrf ;
    if (!strcmp (name, "flops_trsm")) return P.flops_trsm ;
    if (!strcmp (name, "flops_syrk")) return P.flops_syrk ;
    if (!strcmp (name, "flops_gemm")) return P.flops_gemm ;
    if (!strcmp (name, "ms_factor")) return f->ms_factor ;
    if (!strcmp (name, "ms_h2d")) return f->ms_h2d ;
    if (!strcmp (name, "ms_d2h")) return f->ms_d2h ;
    if (!strcmp (name, "ms_scatter")) return f->ms_scatter ;
    if (!strcmp (name, "ms_solve")) return f->ms_solve ;
    if (!strcmp (name, "ms_pull")) return f->ms_pull ;
    if (!strcmp (name, "ms_symbolic")) return f->ms_symbolic ;
    if (!strncmp (name, "ms_kind", 7) && name [7] >= '0' && name [7] <= '9' && atoi (name + 7) < K_NKINDS) return f->ms_kind [atoi (name + 7)] ;
    if (!strncmp (name, "ms_solve_part", 13) && name [13] >= '0' && name [13] < '4') return f->ms_solve_part [name [13] - '0'] ;
    if (!strcmp (name, "solve_cols")) return f->solve_cols ;
    if (!strcmp (name, "solve_w256_bytes")) return (double) f->d_solveW256.n ;
#ifdef B200_SOLVE_TRACE
    if (!strncmp (name, "solve_trace", 11))
    {
        unsigned long long h [32] ;
        const int k = atoi (name + 11) ;
        if (k < 0 || k >= 32 || cudaMemcpyFromSymbol (h, g_solve_trace, sizeof (h)) != cudaSuccess) return -1 ;
        return (double) h [k] ;
    }
#endif
    if (!strcmp (name, "graph_replays")) return (double) f->graph_replays ;
    if (!strcmp (name, "graph_captures")) return (double) f->graph_captures ;
    if (!strcmp (name, "use_graph")) return f->use_graph ;
    if (!strncmp (name, "solve_launches", 14) && name [14] >= '0' && name [14] < '4') return (double) f->solve_launch_count [name [14] - '0'] ;
    if (!strncmp (name, "ms_launch_start", 15)) { size_t k = (size_t) atoll (name + 15) ; return k < f->ms_launch_start.size () ? f->ms_launch_start [k] : -1 ; }
    if (!strncmp (name, "ms_launch", 9)) { size_t k = (size_t) atoll (name + 9) ; return k < f->ms_launch.size () ? f->ms_launch [k] : -1 ; }
    if (!strncmp (name, "ms_level_end", 12)) { int k = atoi (name + 12) ; return (k >= 0 && k < 64) ? f->ms_level_end [k] : -1 ; }
    if (!strncmp (name, "flops_kind", 10) && name [10] >= '0' && name [10] <= '9' && atoi (name + 10) < K_NKINDS) return P.flops_kind [atoi (name + 10)] ;
    return -1 ;
}

// Copy plan tables to caller buffers (tests of the host planner; works on a device < 0 context).
// which: "level" int32[nsuper], "relmap" int32[relmap_len], "items" (nitems x 48 bytes),
// "tiles" (ntiles x 20), "potrf" (npotrf x 24), "trsm" (ntrsm x 32), "launches" (nlaunches x 24),
// "parent" int64[nsuper].  Returns the number of bytes written, or -1.
}   // extern "C"

// The solve schedule (host only; b200_plan_export "solve_*" hands it to the CPU tests, which replay it).  Per level:
// supernodes of at most SOLVE_SMALL columns get one CTA each (k_solve_fwd / k_solve_bwd, kind 0); wider ones advance
// 256 columns ("superstep") per launch, batched over the level -- or 64 columns per launch with
// CHOLMOD_B200_SOLVE_COLS=64 (kind 2).  The backward sweep reads the same list in reverse.
struct SolveSchedule
{
    int cols = 256 ;
    std::vector<int32_t> small ;
    std::vector<SolveBlk> blks, steps ;         // 64 x 64 diagonal blocks; supersteps { s, c0, nb, first block }
    std::vector<SolveTile> tiles ;
    std::vector<SolveTile256> tiles256 ;
    std::vector<b200_factor::SolveLaunch> launches ;
} ;

static void build_solve_schedule (const Plan &P, const std::vector<int64_t> &h_super, const std::vector<int64_t> &h_pi,
    SolveSchedule &S)
{
    constexpr int64_t SOLVE_BIG = SOLVE_SMALL, SOLVE_NB = 64, SOLVE_ROWS = 256, SOLVE_SUPER = 256, SOLVE_FROWS = 64 ;
    const char *ecols = getenv ("CHOLMOD_B200_SOLVE_COLS") ;
    S.cols = (ecols && atoi (ecols) == 64) ? 64 : 256 ;
    std::vector<int32_t> &small = S.small ;
    std::vector<SolveBlk> &blks = S.blks, &steps = S.steps ;
    std::vector<SolveTile> &tiles = S.tiles ;
    std::vector<SolveTile256> &tiles256 = S.tiles256 ;
    std::vector<int32_t> blk_first ((size_t) P.nsuper, -1), step_first ((size_t) P.nsuper, -1) ;
    small.clear () ; blks.clear () ; steps.clear () ; tiles.clear () ; tiles256.clear () ; S.launches.clear () ;
    for (int lev = 0 ; lev < P.nlevels ; lev++)
    {
        const int64_t first = P.level_ptr [lev], cnt = P.level_ptr [lev+1] - first ;
        const size_t s0 = small.size () ;
        std::vector<int32_t> big ;
        int64_t maxcol = 0, maxbig = 0 ;
        for (int64_t q = 0 ; q < cnt ; q++)
        {
            const int32_t sn = P.level_list [first + q] ;
            const int64_t nscol = h_super [sn+1] - h_super [sn] ;
            if (nscol <= SOLVE_BIG) { small.push_back (sn) ; maxcol = std::max (maxcol, nscol) ; }
            else
            {
                // the 64 x 64 diagonal blocks of one supernode are consecutive in blks (and in Winv)
                big.push_back (sn) ; maxbig = std::max (maxbig, nscol) ;
                blk_first [sn] = (int32_t) blks.size () ;
                for (int64_t c0 = 0 ; c0 < nscol ; c0 += SOLVE_NB)
                    blks.push_back ({ sn, (int32_t) c0, (int32_t) std::min<int64_t> (SOLVE_NB, nscol - c0), 0 }) ;
                // and so are its supersteps
                step_first [sn] = (int32_t) steps.size () ;
                if (S.cols != 64)
                    for (int64_t c0 = 0 ; c0 < nscol ; c0 += SOLVE_SUPER)
                        steps.push_back ({ sn, (int32_t) c0, (int32_t) std::min<int64_t> (SOLVE_SUPER, nscol - c0),
                            blk_first [sn] + (int32_t) (c0 / SOLVE_NB) }) ;
            }
        }
        if (small.size () > s0) S.launches.push_back ({ 0, (int64_t) s0, (int64_t) (small.size () - s0), maxcol }) ;
        if (S.cols == 64)
        {
            for (int64_t c0 = 0 ; c0 < maxbig ; c0 += SOLVE_NB)
            {
                const size_t t0 = tiles.size () ;
                for (int32_t sn : big)
                {
                    const int64_t nscol = h_super [sn+1] - h_super [sn], nsrow = h_pi [sn+1] - h_pi [sn] ;
                    if (nscol <= c0) continue ;
                    const int32_t jb = (int32_t) std::min<int64_t> (SOLVE_NB, nscol - c0) ;
                    const int32_t blk = blk_first [sn] + (int32_t) (c0 / SOLVE_NB) ;
                    const size_t tb = tiles.size () ;
                    for (int64_t r = c0 + jb ; r < nsrow ; r += SOLVE_ROWS)
                        tiles.push_back ({ sn, (int32_t) c0, jb, (int32_t) r, (int32_t) std::min<int64_t> (SOLVE_ROWS, nsrow - r), 0, 0, blk }) ;
                    // a block with nothing below it still needs its diagonal solve
                    if (tiles.size () == tb) tiles.push_back ({ sn, (int32_t) c0, jb, (int32_t) (c0 + jb), 0, 0, 0, blk }) ;
                    tiles [tb].writer = 1 ;
                    for (size_t k = tb ; k < tiles.size () ; k++) tiles [k].ntiles = (int32_t) (tiles.size () - tb) ;
                }
                S.launches.push_back ({ 2, (int64_t) t0, (int64_t) (tiles.size () - t0), 0 }) ;
            }
            continue ;
        }
        // 256 columns per launch.  Forward: the first superstep of every wide supernode of the level is solved by
        // a launch of its own (kind 3); after that the tiles of superstep k (kind 4, 64 rows each) also solve the
        // diagonal block of superstep k+1.  Backward: tiles of 64 rows too (kind 5), the last one to finish solves
        // the superstep's own diagonal block.
        if (!big.empty ())
        {
            const size_t t0 = tiles256.size () ;
            for (int32_t sn : big)
            {
                const int64_t nscol = h_super [sn+1] - h_super [sn] ;
                tiles256.push_back ({ sn, 0, (int32_t) std::min<int64_t> (SOLVE_SUPER, nscol), 0, 0, 1, blk_first [sn], step_first [sn], 0, 0 }) ;
            }
            S.launches.push_back ({ 3, (int64_t) t0, (int64_t) (tiles256.size () - t0), 0 }) ;
        }
        for (int64_t c0 = 0 ; c0 < maxbig ; c0 += SOLVE_SUPER)
        {
            const size_t tf = tiles256.size () ;
            std::vector<SolveTile256> back ;
            for (int32_t sn : big)
            {
                const int64_t nscol = h_super [sn+1] - h_super [sn], nsrow = h_pi [sn+1] - h_pi [sn] ;
                if (nscol <= c0) continue ;
                const int32_t nb = (int32_t) std::min<int64_t> (SOLVE_SUPER, nscol - c0) ;
                const int32_t next_nb = (int32_t) std::min<int64_t> (SOLVE_SUPER, nscol - (c0 + nb)) ;      // 0 at the last one
                const int32_t blk0 = blk_first [sn] + (int32_t) (c0 / SOLVE_NB), step = step_first [sn] + (int32_t) (c0 / SOLVE_SUPER) ;
                for (int64_t r = c0 + nb ; r < nsrow ; r += SOLVE_FROWS)
                    tiles256.push_back ({ sn, (int32_t) c0, nb, (int32_t) r, (int32_t) std::min<int64_t> (SOLVE_FROWS, nsrow - r), 0,
                        blk0, step, (r < c0 + nb + next_nb) ? next_nb : 0, 0 }) ;
                const size_t tb = back.size () ;
                for (int64_t r = c0 + nb ; r < nsrow ; r += SOLVE_FROWS)
                    back.push_back ({ sn, (int32_t) c0, nb, (int32_t) r, (int32_t) std::min<int64_t> (SOLVE_FROWS, nsrow - r), 0,
                        blk0, step, 0, 0 }) ;
                if (back.size () == tb) back.push_back ({ sn, (int32_t) c0, nb, (int32_t) (c0 + nb), 0, 0, blk0, step, 0, 0 }) ;
                for (size_t k = tb ; k < back.size () ; k++) back [k].ntiles = (int32_t) (back.size () - tb) ;
            }
            S.launches.push_back ({ 4, (int64_t) tf, (int64_t) (tiles256.size () - tf), 0 }) ;
            S.launches.push_back ({ 5, (int64_t) tiles256.size (), (int64_t) back.size (), 0 }) ;
            tiles256.insert (tiles256.end (), back.begin (), back.end ()) ;
        }
    }
}

extern "C" {

int64_t b200_plan_export (const b200_factor *f, const char *which, void *out, int64_t out_bytes)
{
    if (!f || !which) return -1 ;
    const Plan &P = f->plan ;
    const void *src = nullptr ; size_t bytes = 0 ;
    if (!strcmp (which, "level")) { src = P.level.data () ; bytes = P.level.size () * sizeof (int32_t) ; }
    else if (!strcmp (which, "parent")) { src = P.parent.data () ; bytes = P.parent.size () * sizeof (int64_t) ; }
    else if (!strcmp (which, "relmap")) { src = P.relmap.data () ; bytes = P.relmap.size () * sizeof (int32_t) ; }
    else if (!strcmp (which, "items")) { src = P.items.data () ; bytes = P.items.size () * sizeof (UpdItem) ; }
    else if (!strcmp (which, "tiles")) { src = P.tiles.data () ; bytes = P.tiles.size () * sizeof (UpdTile) ; }
    else if (!strcmp (which, "potrf")) { src = P.potrf.data () ; bytes = P.potrf.size () * sizeof (PotrfBlk) ; }
    else if (!strcmp (which, "trsm")) { src = P.trsm.data () ; bytes = P.trsm.size () * sizeof (TrsmTile) ; }
    else if (!strcmp (which, "launches")) { src = P.launches.data () ; bytes = P.launches.size () * sizeof (Launch) ; }
    else if (!strcmp (which, "copies")) { src = P.copies.data () ; bytes = P.copies.size () * sizeof (CopySeg) ; }
    else if (!strcmp (which, "bcasts")) { src = P.bcasts.data () ; bytes = P.bcasts.size () * sizeof (BcastSeg) ; }
    else if (!strcmp (which, "owner")) { src = P.owner.data () ; bytes = P.owner.size () * sizeof (int32_t) ; }
    else if (!strcmp (which, "top")) { src = P.top.data () ; bytes = P.top.size () ; }
    else if (!strcmp (which, "group")) { src = P.group.data () ; bytes = P.group.size () * sizeof (int32_t) ; }
    else if (!strcmp (which, "dpx")) { src = P.dpx.data () ; bytes = P.dpx.size () * sizeof (int64_t) ; }
    else if (!strcmp (which, "colmine")) { src = P.colmine.data () ; bytes = P.colmine.size () ; }
    else if (!strcmp (which, "pulls")) { src = P.pulls.data () ; bytes = P.pulls.size () * sizeof (PullSeg) ; }
    else if (!strcmp (which, "owned")) { src = P.owned.data () ; bytes = P.owned.size () * sizeof (CopySeg) ; }
    else if (!strncmp (which, "dev_", 4) && f->device >= 0 && f->device_symbolic)
    {
        // the device-built symbolic tables, copied back (tests compare them with the host planner's)
        const void *dsrc = nullptr ;
        if (!strcmp (which, "dev_pairs")) { dsrc = f->d_pairs.p ; bytes = f->d_pairs.bytes () ; }
        else if (!strcmp (which, "dev_pair_map")) { dsrc = f->d_pair_map.p ; bytes = f->d_pair_map.bytes () ; }
        else if (!strcmp (which, "dev_relmap")) { dsrc = f->d_relmap.p ; bytes = (size_t) P.relmap_len * sizeof (int32_t) ; }
        else return -1 ;
        if (!out) return (int64_t) bytes ;
        if ((int64_t) bytes > out_bytes) return -1 ;
        if (cudaSetDevice (f->device) != cudaSuccess) return -1 ;
        if (bytes && cudaMemcpy (out, dsrc, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError () ; return -1 ; }
        return (int64_t) bytes ;
    }
    else if (!strncmp (which, "solve_", 6))
    {
        // the solve schedule, built on request (host only): "solve_small" int32, "solve_blks" / "solve_steps" SolveBlk,
        // "solve_tiles" SolveTile, "solve_tiles256" SolveTile256, "solve_launches" { kind, first, count, maxcol } int64
        static thread_local SolveSchedule S ;
        static thread_local std::vector<int64_t> flat ;
        build_solve_schedule (P, f->h_super, f->h_pi, S) ;
        if (!strcmp (which, "solve_small")) { src = S.small.data () ; bytes = S.small.size () * sizeof (int32_t) ; }
        else if (!strcmp (which, "solve_blks")) { src = S.blks.data () ; bytes = S.blks.size () * sizeof (SolveBlk) ; }
        else if (!strcmp (which, "solve_steps")) { src = S.steps.data () ; bytes = S.steps.size () * sizeof (SolveBlk) ; }
        else if (!strcmp (which, "solve_tiles")) { src = S.tiles.data () ; bytes = S.tiles.size () * sizeof (SolveTile) ; }
        else if (!strcmp (which, "solve_tiles256")) { src = S.tiles256.data () ; bytes = S.tiles256.size () * sizeof (SolveTile256) ; }
        else if (!strcmp (which, "solve_launches"))
        {
            flat.clear () ;
            for (const auto &L : S.launches) { flat.push_back (L.kind) ; flat.push_back (L.first) ; flat.push_back (L.count) ; flat.push_back (L.maxcol) ; }
            src = flat.data () ; bytes = flat.size () * sizeof (int64_t) ;
        }
        else return -1 ;
    }
    else if (!strcmp (which, "pairs")) { src = P.pairs.data () ; bytes = P.pairs.size () * sizeof (Pair) ; }
    else if (!strcmp (which, "pair_map")) { src = P.pair_map.data () ; bytes = P.pair_map.size () * sizeof (int64_t) ; }
    else return -1 ;
    if (!out) return (int64_t) bytes ;
    if ((int64_t) bytes > out_bytes) return -1 ;
    if (bytes) memcpy (out, src, bytes) ;
    return (int64_t) bytes ;
}

int b200_comm_unique_id (void *id128)
{
    NcclApi *nc = nccl_api () ;
    if (!nc || !id128) return B200_GPU_PROBLEM ;
    static_assert (sizeof (ncclUniqueId) == 128, "ncclUniqueId is 128 bytes") ;
    ncclUniqueId id ;
    if (nc->GetUniqueId (&id) != ncclSuccess) return B200_GPU_PROBLEM ;
    memcpy (id128, &id, sizeof (id)) ;
    return B200_OK ;
}

int b200_comm_init (b200_factor *f, const void *id128)
{
    int st = set_device (f) ; if (st) return st ;
    if (!id128) return B200_INVALID ;
    if (f->plan.nranks <= 1) return B200_OK ;
    NcclApi *nc = nccl_api () ;
    if (!nc) return B200_GPU_PROBLEM ;
    if (f->comm) { nc->CommDestroy (f->comm) ; f->comm = nullptr ; }
    ncclUniqueId id ;
    memcpy (&id, id128, sizeof (id)) ;
    ncclResult_t r = nc->CommInitRank (&f->comm, f->plan.nranks, id, f->plan.rank) ;
    if (r != ncclSuccess)
    {
        fprintf (stderr, "cholmod_b200: ncclCommInitRank: %s\n", nc->GetErrorString (r)) ;
        f->comm = nullptr ;
        return B200_GPU_PROBLEM ;
    }
    return B200_OK ;
}

void *b200_peer_arena (b200_factor *f) { return (f && f->device >= 0) ? (void *) f->d_Lx.p : nullptr ; }

int b200_peer_handle (b200_factor *f, void *handle64)
{
    int st = set_device (f) ; if (st) return st ;
    if (!handle64 || !f->d_Lx.p) return B200_INVALID ;
    static_assert (sizeof (cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes") ;
    cudaIpcMemHandle_t h ;
    CK (cudaIpcGetMemHandle (&h, f->d_Lx.p), "cudaIpcGetMemHandle") ;
    memcpy (handle64, &h, sizeof (h)) ;
    return B200_OK ;
}

int b200_peer_attach (b200_factor *f, int rank, void *arena, const void *handle64)
{
    int st = set_device (f) ; if (st) return st ;
    const int R = f->plan.nranks ;
    if (rank < 0 || rank >= R || R > 32) return B200_INVALID ;
    if (rank != f->plan.rank)
    {
        if (f->peer_ipc [rank] && f->peers.p [rank]) { cudaIpcCloseMemHandle (f->peers.p [rank]) ; f->peer_ipc [rank] = false ; }
        f->peers.p [rank] = nullptr ;
        if (arena)
        {
            // same process: make the peer's device reachable from ours, then its pointer is good as it is
            cudaPointerAttributes at ;
            CK (cudaPointerGetAttributes (&at, arena), "cudaPointerGetAttributes") ;
            if (at.type != cudaMemoryTypeDevice) return B200_INVALID ;
            if (at.device != f->device)
            {
                int can = 0 ;
                CK (cudaDeviceCanAccessPeer (&can, f->device, at.device), "cudaDeviceCanAccessPeer") ;
                if (!can) { fprintf (stderr, "cholmod_b200: device %d cannot reach device %d\n", f->device, at.device) ; return B200_GPU_PROBLEM ; }
                cudaError_t e = cudaDeviceEnablePeerAccess (at.device, 0) ;
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { cuda_ok (e, "cudaDeviceEnablePeerAccess") ; return B200_GPU_PROBLEM ; }
                cudaGetLastError () ;
            }
            f->peers.p [rank] = arena ;
            f->one_process = true ;
        }
        else if (handle64)
        {
            cudaIpcMemHandle_t h ;
            memcpy (&h, handle64, sizeof (h)) ;
            void *p = nullptr ;
            CK (cudaIpcOpenMemHandle (&p, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle") ;
            f->peers.p [rank] = p ; f->peer_ipc [rank] = true ;
        }
        else return B200_INVALID ;
    }
    int cnt = 0 ;
    for (int r = 0 ; r < R ; r++) if (f->peers.p [r]) cnt++ ;
    f->peers_attached = cnt ;
    return B200_OK ;
}

// Host -> device copy of a large array in the caller's pageable memory.  cudaMemcpyAsync from pageable memory is
// one thread of the driver copying into its own bounce buffers (10.8 GB/s measured on the box, 6.6 ms for config
// 2's A and 30 ms for config 5's, a third of its end-to-end time); here STAGE_THREADS host threads copy 4 MB chunks
// into a page-locked ring owned by the context and queue the DMA of each chunk behind it, so the memcpy of one chunk
// overlaps the DMA of the others.  Small arrays take the plain path.
constexpr size_t STAGE_CHUNK = (size_t) 4 << 20 ;
constexpr int STAGE_SLOTS = 8, STAGE_THREADS = 4 ;

static int h2d_staged (b200_factor *f, void *dst, const void *src, size_t bytes, cudaStream_t s)
{
    if (bytes < 4 * STAGE_CHUNK)
    {
        CK (cudaMemcpyAsync (dst, src, bytes, cudaMemcpyHostToDevice, s), "H2D") ;
        return B200_OK ;
    }
    if (!f->stage)
    {
        if (cudaHostAlloc ((void **) &f->stage, STAGE_SLOTS * STAGE_CHUNK, cudaHostAllocDefault) != cudaSuccess)
        {
            cudaGetLastError () ; f->stage = nullptr ;
            CK (cudaMemcpyAsync (dst, src, bytes, cudaMemcpyHostToDevice, s), "H2D") ;
            return B200_OK ;
        }
        for (auto &e : f->stage_ev) CK (cudaEventCreateWithFlags (&e, cudaEventDisableTiming), "stage event") ;
    }
    const size_t nchunks = (bytes + STAGE_CHUNK - 1) / STAGE_CHUNK ;
    int status [STAGE_THREADS] = {} ;
    std::thread th [STAGE_THREADS] ;
    for (int t = 0 ; t < STAGE_THREADS ; t++)
        th [t] = std::thread ([=, &status]
        {
            if (cudaSetDevice (f->device) != cudaSuccess) { status [t] = 1 ; return ; }
            size_t turn = 0 ;
            for (size_t c = (size_t) t ; c < nchunks ; c += STAGE_THREADS, turn++)
            {
                // thread t owns the slots t and t + STAGE_THREADS and alternates between them
                const int slot = t + (int) (turn & 1) * STAGE_THREADS ;
                const size_t off = c * STAGE_CHUNK, len = std::min (STAGE_CHUNK, bytes - off) ;
                unsigned char *buf = f->stage + (size_t) slot * STAGE_CHUNK ;
                if (cudaEventSynchronize (f->stage_ev [slot]) != cudaSuccess) { status [t] = 1 ; return ; }   // its last DMA is done
                memcpy (buf, (const unsigned char *) src + off, len) ;
                if (cudaMemcpyAsync ((unsigned char *) dst + off, buf, len, cudaMemcpyHostToDevice, s) != cudaSuccess
                    || cudaEventRecord (f->stage_ev [slot], s) != cudaSuccess) { status [t] = 1 ; return ; }
            }
        }) ;
    for (auto &t : th) t.join () ;
    for (int st : status) if (st) { cudaGetLastError () ; return B200_GPU_PROBLEM ; }
    return B200_OK ;
}

// one CSC matrix (ncols columns) to the device; returns its entry count through *nz_out
static int upload_csc (b200_factor *f, int64_t ncols, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const void *Ax, DevBuf<int64_t> &dp, DevBuf<int64_t> &di, DevBuf<int64_t> &dnz, DevBuf<unsigned char> &dx,
    int64_t *nz_out)
{
    if (!Ap || ncols < 0) return B200_INVALID ;
    // unpacked matrices: entries of column k live at Ap [k] .. Ap [k] + Anz [k]; copy up to max end
    int64_t nz = Ap [ncols] ;
    if (Anz) for (int64_t k = 0 ; k < ncols ; k++) nz = (Ap [k] + Anz [k] > nz) ? Ap [k] + Anz [k] : nz ;
    if (nz < 0 || (nz > 0 && (!Ai || !Ax))) return B200_INVALID ;
    cudaStream_t s = f->stream ;
    if (dp.n < (size_t) ncols + 1 && dp.alloc ((size_t) ncols + 1) != cudaSuccess) return B200_OUT_OF_MEMORY ;
    if (di.n < (size_t) nz && di.alloc ((size_t) nz) != cudaSuccess) return B200_OUT_OF_MEMORY ;
    if (dx.n < (size_t) nz * f->esize && dx.alloc ((size_t) nz * f->esize) != cudaSuccess) return B200_OUT_OF_MEMORY ;
    if (Anz && dnz.n < (size_t) ncols && dnz.alloc ((size_t) ncols) != cudaSuccess) return B200_OUT_OF_MEMORY ;
    // several ranks with a communicator: every rank was given the same matrix, so only rank 0 sends it over PCIe
    // and the others receive it over NVLink (one host feeding N GPUs the same 70-350 MB at once is slower than that).
    // Not when the ranks are threads of one process: a rank that is already inside its broadcast kernel waits for
    // rank 0, and rank 0's allocations and pageable copies above can wait for that rank's device (peer access makes
    // an allocation a matter of every device): every rank uploads its own copy then.
    NcclApi *nc = (f->plan.nranks > 1 && f->comm && !f->one_process) ? nccl_api () : nullptr ;
    if (!nc || f->plan.rank == 0)
    {
        CK (cudaMemcpyAsync (dp.p, Ap, (size_t) (ncols + 1) * sizeof (int64_t), cudaMemcpyHostToDevice, s), "H2D Ap") ;
        if (nz > 0)
        {
            int st = h2d_staged (f, di.p, Ai, (size_t) nz * sizeof (int64_t), s) ;
            if (!st) st = h2d_staged (f, dx.p, Ax, (size_t) nz * f->esize, s) ;
            if (st) return st ;
        }
        if (Anz && ncols > 0) CK (cudaMemcpyAsync (dnz.p, Anz, (size_t) ncols * sizeof (int64_t), cudaMemcpyHostToDevice, s), "H2D Anz") ;
    }
    if (nc)
    {
        bool ok = nc->GroupStart () == ncclSuccess ;
        ok = ok && nc->Broadcast (dp.p, dp.p, (size_t) (ncols + 1) * 8, ncclChar, 0, f->comm, s) == ncclSuccess ;
        if (nz > 0)
        {
            ok = ok && nc->Broadcast (di.p, di.p, (size_t) nz * 8, ncclChar, 0, f->comm, s) == ncclSuccess ;
            ok = ok && nc->Broadcast (dx.p, dx.p, (size_t) nz * f->esize, ncclChar, 0, f->comm, s) == ncclSuccess ;
        }
        if (Anz && ncols > 0) ok = ok && nc->Broadcast (dnz.p, dnz.p, (size_t) ncols * 8, ncclChar, 0, f->comm, s) == ncclSuccess ;
        ok = (nc->GroupEnd () == ncclSuccess) && ok ;
        if (!ok) return B200_GPU_PROBLEM ;
    }
    *nz_out = nz ;
    return B200_OK ;
}

int b200_upload_matrix (b200_factor *f, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const void *Ax)
{
    int st = set_device (f) ; if (st) return st ;
    cudaStream_t s = f->stream ;
    int64_t nz = 0 ;
    CK (cudaEventRecord (f->ev [0], s), "record") ;
    st = upload_csc (f, f->plan.n, Ap, Ai, Anz, Ax, f->d_Ap, f->d_Ai, f->d_Anz, f->d_Ax, &nz) ;
    if (st) return st ;
    CK (cudaEventRecord (f->ev [1], s), "record") ;
    CK (cudaStreamSynchronize (s), "H2D A") ;
    cudaEventElapsedTime (&f->ms_h2d, f->ev [0], f->ev [1]) ;
    f->have_A = true ; f->A_unpacked = (Anz != nullptr) ; f->A_nz = nz ;
    f->have_F = false ;
    return B200_OK ;
}

int b200_upload_matrix_aat (b200_factor *f, int64_t ncol, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const void *Ax, const int64_t *Fp, const int64_t *Fi, const int64_t *Fnz, const void *Fx)
{
    int st = set_device (f) ; if (st) return st ;
    cudaStream_t s = f->stream ;
    int64_t nz = 0, fnz = 0 ;
    f->have_A = false ; f->have_F = false ;
    CK (cudaEventRecord (f->ev [0], s), "record") ;
    st = upload_csc (f, ncol, Ap, Ai, Anz, Ax, f->d_Ap, f->d_Ai, f->d_Anz, f->d_Ax, &nz) ;
    if (st) return st ;
    st = upload_csc (f, f->plan.n, Fp, Fi, Fnz, Fx, f->d_Fp, f->d_Fi, f->d_Fnz, f->d_Fx, &fnz) ;
    if (st) return st ;
    CK (cudaEventRecord (f->ev [1], s), "record") ;
    CK (cudaStreamSynchronize (s), "H2D A, F") ;
    cudaEventElapsedTime (&f->ms_h2d, f->ev [0], f->ev [1]) ;
    f->have_A = true ; f->A_unpacked = (Anz != nullptr) ; f->A_nz = nz ;
    f->have_F = true ; f->F_unpacked = (Fnz != nullptr) ;
    return B200_OK ;
}

int b200_upload_values (b200_factor *f, const void *Ax)
{
    int st = set_device (f) ; if (st) return st ;
    if (!f->have_A || f->have_F) return B200_INVALID ;      // the pattern of a symmetric A must be on the device
    if (f->A_nz == 0) return B200_OK ;
    if (!Ax) return B200_INVALID ;
    CK (cudaEventRecord (f->ev [0], f->stream), "record") ;
    NcclApi *nc = (f->plan.nranks > 1 && f->comm && !f->one_process) ? nccl_api () : nullptr ;      // as in upload_csc
    if (!nc || f->plan.rank == 0)
    {
        st = h2d_staged (f, f->d_Ax.p, Ax, (size_t) f->A_nz * f->esize, f->stream) ;
        if (st) return st ;
    }
    if (nc && nc->Broadcast (f->d_Ax.p, f->d_Ax.p, (size_t) f->A_nz * f->esize, ncclChar, 0, f->comm, f->stream) != ncclSuccess)
        return B200_GPU_PROBLEM ;
    CK (cudaEventRecord (f->ev [1], f->stream), "record") ;
    CK (cudaStreamSynchronize (f->stream), "H2D Ax") ;
    cudaEventElapsedTime (&f->ms_h2d, f->ev [0], f->ev [1]) ;
    return B200_OK ;
}

}   // extern "C"

// Host ranges this library has page-locked, process wide, with the number of contexts that rely on each: several
// contexts may write into one buffer (the devices of b200_multi; the shim's contexts of different threads that meet
// the same L->x one after the other), and a range must stay locked until the last of them lets go -- another
// context's copy-out may be in flight into it.
struct HostReg { size_t bytes ; int refs ; } ;
static std::mutex g_reg_lock ;
static std::map<void *, HostReg> g_reg ;

static void unregister_pieces (b200_factor *f)
{
    std::lock_guard<std::mutex> guard (g_reg_lock) ;
    for (auto &pc : f->reg_pieces)
    {
        auto it = g_reg.find (pc.first) ;
        if (it == g_reg.end ()) continue ;
        if (--it->second.refs <= 0)
        {
            cudaHostUnregister (pc.first) ;
            cudaGetLastError () ;
            g_reg.erase (it) ;
        }
    }
    f->reg_pieces.clear () ;
    f->reg_ptr = nullptr ; f->reg_bytes = 0 ;
    f->reg_gen++ ;
}

static bool is_page_locked (const void *p, size_t bytes)
{
    unsigned fl = 0 ;
    if (bytes == 0) return true ;
    if (cudaHostGetFlags (&fl, const_cast<void *> (p)) == cudaSuccess
        && cudaHostGetFlags (&fl, (char *) const_cast<void *> (p) + (bytes - 1)) == cudaSuccess) return true ;
    cudaGetLastError () ;
    return false ;
}

// take a reference on [p, p + len) if this library has it locked already, else lock it (unless somebody else has)
static void lock_piece (b200_factor *f, void *p, size_t len)
{
    std::lock_guard<std::mutex> guard (g_reg_lock) ;
    f->reg_gen++ ;
    auto it = g_reg.find (p) ;
    if (it != g_reg.end () && it->second.bytes == len)
    {
        it->second.refs++ ;
        f->reg_pieces.push_back ({ p, len }) ;
        return ;
    }
    if (is_page_locked (p, len)) return ;       // the caller's own pinned memory (or an overlapping range of ours)
    if (cudaHostRegister (p, len, cudaHostRegisterPortable) == cudaSuccess)
    {
        g_reg [p] = HostReg { len, 1 } ;
        f->reg_pieces.push_back ({ p, len }) ;
    }
    else cudaGetLastError () ;
}

// Page-lock the caller's L->x once: later factorizations into the same buffer then copy at PCIe speed and
// overlap with the computation.  One GPU, or several GPUs driven from one process (b200_multi): the whole buffer
// (portable; the first context to get here registers it, the others find it registered).  One process per GPU:
// every process page-locks only the pieces of the (shared) buffer that its rank writes -- its subtrees and its
// blocks of the top separators -- not N times the whole factor.
static void ensure_registered (b200_factor *f, void *Lx, size_t bytes)
{
    if (f->reg_ptr == Lx && f->reg_bytes == bytes)
    {
        // same address and size as last time -- but the caller may have freed and re-obtained the memory in
        // between without telling us (b200_forget_host_buffer): ask the driver whether what we registered is
        // still a page-locked mapping before trusting it
        bool ok = true ;
        for (auto &pc : f->reg_pieces) ok = ok && is_page_locked (pc.first, pc.second) ;
        if (ok) return ;
    }
    unregister_pieces (f) ;
    f->reg_ptr = Lx ; f->reg_bytes = bytes ;
    bool other_processes = false ;
    for (int r = 0 ; r < 32 ; r++) other_processes = other_processes || f->peer_ipc [r] ;
    if (f->plan.nranks > 1 && other_processes)
    {
        // page-aligned hulls of the owned runs, merged where they touch or lie within 2 MB of each other
        const size_t page = 4096, gap = (size_t) 2 << 20 ;
        std::vector<std::pair<size_t, size_t>> rg ;
        for (const CopySeg &c : f->plan.owned)
        {
            if (c.len <= 0) continue ;
            size_t a = (size_t) c.hoff * f->esize, b = a + (size_t) c.len * f->esize ;
            const size_t base = (size_t) (uintptr_t) Lx ;
            a = ((base + a) / page) * page ; b = ((base + b + page - 1) / page) * page ;
            a = std::max (a, (base / page) * page) ; b = std::min (b, ((base + bytes + page - 1) / page) * page) ;
            rg.push_back ({ a, b }) ;
        }
        std::sort (rg.begin (), rg.end ()) ;
        std::vector<std::pair<size_t, size_t>> merged ;
        for (auto &r : rg)
        {
            if (!merged.empty () && r.first <= merged.back ().second + gap) merged.back ().second = std::max (merged.back ().second, r.second) ;
            else merged.push_back (r) ;
        }
        for (auto &r : merged)
        {
            void *p = (void *) (uintptr_t) r.first ;
            const size_t len = r.second - r.first ;
            if (len >= ((size_t) 1 << 16)) lock_piece (f, p, len) ;
        }
        return ;
    }
    // page-locked already -- by the caller, or by another context of this process that writes into the same
    // buffer (b200_multi)?  Then there is nothing to do, and nothing of ours to undo later.
    if (bytes >= ((size_t) 1 << 20)) lock_piece (f, Lx, bytes) ;
}

// everything this rank owns (all of L->x on one GPU) from the arena to its place in the caller's L->x
template <typename T>
static int copy_owned_to_host (b200_factor *f, void *host_dst, cudaStream_t st)
{
    const T *Lx = reinterpret_cast<const T *> (f->d_Lx.p) ;
    for (const CopySeg &seg : f->plan.owned)
        if (seg.len > 0)
            CK (cudaMemcpyAsync ((char *) host_dst + (size_t) seg.hoff * sizeof (T), Lx + seg.off, (size_t) seg.len * sizeof (T),
                cudaMemcpyDeviceToHost, st), "D2H Lx") ;
    return B200_OK ;
}

// SNW:1084-1146 for a distributed factorization.  No rank holds the whole factor, so instead of repairing one
// supernode every rank re-plans the factorization TRUNCATED at the failing supernode s (everything before s as
// usual, s itself for its first info-1 columns -- or not at all when info == 1 or quick return is set --,
// nothing after s, which stays zero) and runs it again.  `minor` is the same on all ranks (all-reduced), so all
// ranks take this path together.  Rare, and twice the work of a factorization; the values before s come out
// as in the first run.
template <typename T>
static int repair_dist (b200_factor *f, int64_t minor, double beta, int quick)
{
    const Plan &P = f->plan ;
    cudaStream_t st0 = f->stream ;
    const int64_t s = P.supermap [minor] ;
    const int64_t k1 = f->h_super [s], nscol = f->h_super [s+1] - k1, nsrow = f->h_pi [s+1] - f->h_pi [s] ;
    const int64_t info = minor - k1 + 1 ;
    const int64_t ncol_new = (info == 1 || quick) ? 0 : info - 1 ;
    Plan R ;
    int st = build_plan (P.n, P.nsuper, f->h_super.data (), f->h_pi.data (), f->h_px.data (), f->h_Ls.data (),
        dtype_dmma (f->dtype), R, P.nranks, P.rank, s, ncol_new) ;
    if (st) return st ;
    DevBuf<UpdItem> items ; DevBuf<UpdTile> tiles ; DevBuf<PotrfBlk> potrf ; DevBuf<TrsmTile> trsm ; DevBuf<PullSeg> pulls ;
    bool ok = items.upload (R.items, st0) == cudaSuccess && tiles.upload (R.tiles, st0) == cudaSuccess
        && potrf.upload (R.potrf, st0) == cudaSuccess && trsm.upload (R.trsm, st0) == cudaSuccess
        && pulls.upload (R.pulls, st0) == cudaSuccess ;
    if (!ok) st = B200_OUT_OF_MEMORY ;
    const unsigned long long big = ~0ull ;
    T *Lx = reinterpret_cast<T *> (f->d_Lx.p) ;
    if (!st && cudaMemcpyAsync (f->d_minor.p, &big, sizeof (big), cudaMemcpyHostToDevice, st0) != cudaSuccess) st = B200_GPU_PROBLEM ;
    if (!st && P.dxsize > 0 && cudaMemsetAsync (Lx, 0, (size_t) P.dxsize * sizeof (T), st0) != cudaSuccess) st = B200_GPU_PROBLEM ;
    if (!st) st = scatter_A<T> (f, 0, ncol_new > 0 ? k1 + nscol : k1, beta) ;
    if (!st) st = run_launches<T> (f, R, items.p, tiles.p, potrf.p, trsm.p, f->d_relmap.p, pulls.p) ;
    if (!st && ncol_new > 0 && P.dpx [s] >= 0
        && cudaMemsetAsync (Lx + P.dpx [s] + nsrow * ncol_new, 0, (size_t) (nsrow * (nscol - ncol_new)) * sizeof (T), st0) != cudaSuccess)
        st = B200_GPU_PROBLEM ;
    if (cudaStreamSynchronize (st0) != cudaSuccess && !st) st = B200_GPU_PROBLEM ;
    if (cudaStreamSynchronize (f->stream2) != cudaSuccess && !st) st = B200_GPU_PROBLEM ;
    if (cudaStreamSynchronize (f->stream4) != cudaSuccess && !st) st = B200_GPU_PROBLEM ;
    if (cudaStreamSynchronize (f->stream5) != cudaSuccess && !st) st = B200_GPU_PROBLEM ;
    items.release () ; tiles.release () ; potrf.release () ; trsm.release () ; pulls.release () ;
    return st ;
}

// The plan's launch list through a CUDA graph: replay the graph captured for this copy-out destination, or capture
// it now (the launches are recorded, not run), instantiate and launch it.  Returns 1 when the graph ran, 0 when the
// caller has to issue the launches itself (capture not possible: nothing has run), < 0 on an error.
template <typename T>
static int run_graph (b200_factor *f)
{
    const Plan &P = f->plan ;
    cudaStream_t s = f->stream ;
    b200_factor::GraphSlot *slot = nullptr ;
    const bool pinned = f->copy_dst && is_page_locked (f->copy_dst, (size_t) P.xsize * sizeof (T)) ;
    for (auto &g : f->graphs)
        if (g.exec && g.dst == f->copy_dst)
        {
            if (!g.dst || (g.reg_gen == f->reg_gen && g.pinned == pinned)) slot = &g ;
            else { cudaGraphExecDestroy (g.exec) ; g.exec = nullptr ; }     // the destination's mapping has changed
        }
    if (!slot)
    {
        // every stream of the plan joins the capture behind the main one, and returns to it at the end
        std::vector<cudaStream_t> side = { f->stream2, f->stream3, f->stream5 } ;
        while ((int) f->gstreams.size () < P.nstreams - 5)
        {
            int lo = 0, hi = 0 ;
            CK (cudaDeviceGetStreamPriorityRange (&lo, &hi), "priority range") ;
            cudaStream_t g ;
            CK (cudaStreamCreateWithPriority (&g, cudaStreamNonBlocking, (f->gstreams.size () & 1) ? lo : hi), "group stream") ;
            f->gstreams.push_back (g) ;
        }
        for (cudaStream_t g : f->gstreams) side.push_back (g) ;
        while ((int64_t) f->plan_ev.size () < (int64_t) P.nevents)
        {
            cudaEvent_t e ;
            CK (cudaEventCreateWithFlags (&e, cudaEventDisableTiming), "plan event") ;
            f->plan_ev.push_back (e) ;
        }
        if (!f->ev_fork) CK (cudaEventCreateWithFlags (&f->ev_fork, cudaEventDisableTiming), "fork event") ;
        while (f->ev_join.size () < side.size ())
        {
            cudaEvent_t e ;
            CK (cudaEventCreateWithFlags (&e, cudaEventDisableTiming), "join event") ;
            f->ev_join.push_back (e) ;
        }
        // nothing of an earlier call may still be running on the side streams when they enter the capture
        for (cudaStream_t t : side) if (t) CK (cudaStreamSynchronize (t), "side stream") ;
        if (cudaStreamBeginCapture (s, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError () ; return 0 ; }
        bool ok = cudaEventRecord (f->ev_fork, s) == cudaSuccess ;
        for (cudaStream_t t : side) if (t) ok = ok && cudaStreamWaitEvent (t, f->ev_fork, 0) == cudaSuccess ;
        const int64_t before = f->kernel_launches ;
        int st = ok ? run_launches<T> (f, P, f->d_items.p, f->d_tiles.p, f->d_potrf.p, f->d_trsm.p, f->d_relmap.p, f->d_pulls.p) : B200_GPU_PROBLEM ;
        const int64_t launches = f->kernel_launches - before ;
        f->kernel_launches = before ;
        for (size_t k = 0 ; k < side.size () ; k++)
            if (side [k])
                ok = ok && cudaEventRecord (f->ev_join [k], side [k]) == cudaSuccess && cudaStreamWaitEvent (s, f->ev_join [k], 0) == cudaSuccess ;
        cudaGraph_t graph = nullptr ;
        cudaError_t e = cudaStreamEndCapture (s, &graph) ;
        cudaGraphExec_t exec = nullptr ;
        if (st == B200_OK && ok && e == cudaSuccess && graph) e = cudaGraphInstantiate (&exec, graph, 0) ;
        if (graph) cudaGraphDestroy (graph) ;
        if (st != B200_OK || !ok || e != cudaSuccess || !exec)
        {
            cudaGetLastError () ;
            if (exec) cudaGraphExecDestroy (exec) ;
            f->use_graph = 0 ;          // this context issues its launches one by one from now on
            return (st < 0 && st != B200_GPU_PROBLEM) ? st : 0 ;
        }
        slot = &f->graphs [0] ;
        if (f->graphs [0].exec && (!f->graphs [1].exec || f->graphs [1].used < f->graphs [0].used)) slot = &f->graphs [1] ;
        if (slot->exec) cudaGraphExecDestroy (slot->exec) ;
        slot->exec = exec ; slot->dst = f->copy_dst ; slot->launches = launches ;
        slot->reg_gen = f->reg_gen ; slot->pinned = pinned ;
        f->graph_captures++ ;
    }
    slot->used = ++f->graph_clock ;
    CK (cudaGraphLaunch (slot->exec, s), "graph launch") ;
    f->kernel_launches += slot->launches ;
    f->graph_replays++ ;
    return 1 ;
}

// host_dst != NULL: finished parts of the factor are copied to host_dst [0..xsize) while the rest is still
// being computed (the plan's K_COPYOUT schedule); on return the host copy is complete.
template <typename T>
static int factor_device_T (b200_factor *f, double beta, int quick, int64_t *minor_out, void *host_dst = nullptr)
{
    const Plan &P = f->plan ;
    cudaStream_t s = f->stream ;
    f->copy_dst = host_dst ;
    struct Reset { b200_factor *f ; ~Reset () { f->copy_dst = nullptr ; } } reset { f } ;
    const unsigned long long big = ~0ull ;
    f->kernel_launches = 0 ; f->nccl_calls = 0 ;
    f->solve_w_valid = false ;              // the solve's block inverses belong to the previous factor
    for (int k = 0 ; k < K_NKINDS ; k++) { f->ms_kind [k] = 0 ; f->launches_kind [k] = 0 ; }
    for (int k = 0 ; k < 64 ; k++) f->ms_level_end [k] = 0 ;
    f->ms_pull = 0 ;
    CK (cudaEventRecord (f->ev [2], s), "record") ;
    CK (cudaMemcpyAsync (f->d_minor.p, &big, sizeof (big), cudaMemcpyHostToDevice, s), "minor init") ;
    if (P.dxsize > 0) CK (cudaMemsetAsync (f->d_Lx.p, 0, (size_t) P.dxsize * sizeof (T), s), "memset Lx") ;
    int st = scatter_A<T> (f, 0, P.n, beta) ;
    if (st) return st ;
    CK (cudaEventRecord (f->ev [3], s), "record") ;
    if (f->use_graph < 0)
    {
        // opt-in (CHOLMOD_B200_GRAPH=1): measured on config 5 (9725 launches) a replay takes 23.1 ms against 24.4 ms
        // of plain launches, but capture and instantiation add 0.5 s to the first call (1 s on config 2's 4035)
        const char *e = getenv ("CHOLMOD_B200_GRAPH") ;
        f->use_graph = (e && atoi (e) != 0) ? 1 : 0 ;
    }
    int ran = 0 ;
    if (f->use_graph == 1 && P.nranks == 1 && !f->profile)
    {
        ran = run_graph<T> (f) ;
        if (ran < 0) return ran ;
    }
    if (!ran) st = run_launches<T> (f, P, f->d_items.p, f->d_tiles.p, f->d_potrf.p, f->d_trsm.p, f->d_relmap.p, f->d_pulls.p) ;
    if (st) return st ;
    if (P.nranks > 1)
    {
        // the smallest failing column over all ranks is L->minor
        NcclApi *nc = nccl_api () ;
        if (!nc || !f->comm) return B200_INVALID ;
        CK (cudaEventRecord (f->ev_comm [0], s), "record") ;
        CK (cudaStreamWaitEvent (f->stream4, f->ev_comm [0], 0), "wait") ;
        if (nc->AllReduce (f->d_minor.p, f->d_minor.p, 1, ncclUint64, ncclMin, f->comm, f->stream4) != ncclSuccess) return B200_GPU_PROBLEM ;
        CK (cudaEventRecord (f->ev_comm [1], f->stream4), "record") ;
        CK (cudaStreamWaitEvent (s, f->ev_comm [1], 0), "wait") ;
    }
    unsigned long long minor = big ;
    CK (cudaMemcpyAsync (&minor, f->d_minor.p, sizeof (minor), cudaMemcpyDeviceToHost, s), "minor read") ;
    CK (cudaEventRecord (f->ev [4], s), "record") ;
    CK (cudaStreamSynchronize (s), "factorize") ;
    if (host_dst) CK (cudaStreamSynchronize (f->stream3), "copy-out") ;
    cudaEventElapsedTime (&f->ms_scatter, f->ev [2], f->ev [3]) ;
    cudaEventElapsedTime (&f->ms_factor, f->ev [2], f->ev [4]) ;
    if (minor == big)
    {
        if (minor_out) *minor_out = P.n ;
        return B200_OK ;
    }
    // not positive definite: reproduce SNW:1084-1146
    if (minor_out) *minor_out = (int64_t) minor ;
    f->copy_dst = nullptr ;
    st = (P.nranks > 1) ? repair_dist<T> (f, (int64_t) minor, beta, quick) : repair_not_posdef<T> (f, (int64_t) minor, beta, quick) ;
    if (st < 0) return st ;
    if (host_dst && P.xsize > 0)     // rare path: send the repaired factor again, as a whole
    {
        st = copy_owned_to_host<T> (f, host_dst, s) ;
        if (st < 0) return st ;
    }
    CK (cudaStreamSynchronize (s), "repair") ;
    return B200_NOT_POSDEF ;
}

extern "C" {

int b200_factor_device (b200_factor *f, double beta, int quick_return_if_not_posdef, int64_t *minor)
{
    int st = set_device (f) ; if (st) return st ;
    if (!f->have_A) return B200_INVALID ;
    return B200_DISPATCH (f->dtype, factor_device_T) (f, beta, quick_return_if_not_posdef, minor, nullptr) ;
}

int b200_download_factor (b200_factor *f, void *Lx)
{
    int st = set_device (f) ; if (st) return st ;
    if (!Lx && f->plan.xsize > 0) return B200_INVALID ;
    size_t bytes = (size_t) f->plan.xsize * f->esize ;
    if (bytes == 0) return B200_OK ;
    ensure_registered (f, Lx, bytes) ;
    CK (cudaEventRecord (f->ev [5], f->stream), "record") ;
    st = B200_DISPATCH (f->dtype, copy_owned_to_host) (f, Lx, f->stream) ;
    if (st < 0) return st ;
    CK (cudaEventRecord (f->ev [6], f->stream), "record") ;
    CK (cudaStreamSynchronize (f->stream), "D2H Lx") ;
    cudaEventElapsedTime (&f->ms_d2h, f->ev [5], f->ev [6]) ;
    return B200_OK ;
}

int b200_upload_factor (b200_factor *f, const void *Lx)
{
    int st = set_device (f) ; if (st) return st ;
    size_t bytes = (size_t) f->plan.xsize * f->esize ;
    if (bytes == 0) return B200_OK ;
    if (!Lx || f->plan.nranks > 1) return B200_INVALID ;    // a distributed arena holds only this rank's share
    f->solve_w_valid = false ;
    CK (cudaMemcpyAsync (f->d_Lx.p, Lx, bytes, cudaMemcpyHostToDevice, f->stream), "H2D Lx") ;
    CK (cudaStreamSynchronize (f->stream), "H2D Lx") ;
    return B200_OK ;
}

void *b200_device_Lx (b200_factor *f) { return (f && f->device >= 0) ? (void *) f->d_Lx.p : nullptr ; }

int b200_forget_host_buffer (b200_factor *f, const void *Lx)
{
    if (!f || f->device < 0) return B200_OK ;
    if (f->reg_ptr && (Lx == nullptr || Lx == f->reg_ptr))
    {
        if (set_device (f)) return B200_GPU_PROBLEM ;
        cudaStreamSynchronize (f->stream3) ;
        unregister_pieces (f) ;
    }
    return B200_OK ;
}

int b200_super_numeric_d (b200_factor *f, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const double *Ax, double beta, int quick_return_if_not_posdef, double *Lx, int64_t *minor)
{
    if (!f || f->dtype != B200_DOUBLE) return f ? B200_INVALID : B200_GPU_PROBLEM ;
    int st = b200_upload_matrix (f, Ap, Ai, Anz, Ax) ;
    if (st < 0) return st ;
    if (Lx) ensure_registered (f, Lx, (size_t) f->plan.xsize * sizeof (double)) ;
    return factor_device_T<double> (f, beta, quick_return_if_not_posdef, minor, Lx) ;
}

int b200_super_numeric_s (b200_factor *f, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const float *Ax, double beta, int quick_return_if_not_posdef, float *Lx, int64_t *minor)
{
    if (!f || f->dtype != B200_SINGLE) return f ? B200_INVALID : B200_GPU_PROBLEM ;
    int st = b200_upload_matrix (f, Ap, Ai, Anz, Ax) ;
    if (st < 0) return st ;
    if (Lx) ensure_registered (f, Lx, (size_t) f->plan.xsize * sizeof (float)) ;
    return factor_device_T<float> (f, beta, quick_return_if_not_posdef, minor, Lx) ;
}

}   // extern "C"

// complex: Ax holds (re,im) pairs when Az == NULL (CHOLMOD_COMPLEX), else Ax = real parts and Az =
// imaginary parts (CHOLMOD_ZOMPLEX; interleaved here on the host, O(nnz(A))).  L is always (re,im) pairs,
// as in the reference (a zomplex A gives a complex L, SNW:42-84).
template <typename R>
static int super_numeric_complex (b200_factor *f, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const R *Ax, const R *Az, double beta, int quick, R *Lx, int64_t *minor)
{
    if (!f) return B200_GPU_PROBLEM ;
    if (f->dtype != ((sizeof (R) == 8) ? B200_COMPLEX_DOUBLE : B200_COMPLEX_SINGLE)) return B200_INVALID ;
    const void *vals = Ax ;
    if (Az)
    {
        if (!Ap) return B200_INVALID ;
        const int64_t n = f->plan.n ;
        int64_t nz = Ap [n] ;
        if (Anz) for (int64_t k = 0 ; k < n ; k++) nz = (Ap [k] + Anz [k] > nz) ? Ap [k] + Anz [k] : nz ;
        if (nz < 0 || (nz > 0 && !Ax)) return B200_INVALID ;
        try { f->h_Ax.resize ((size_t) nz * 2 * sizeof (R)) ; }
        catch (const std::bad_alloc &) { return B200_OUT_OF_MEMORY ; }
        R *dst = reinterpret_cast<R *> (f->h_Ax.data ()) ;
        for (int64_t p = 0 ; p < nz ; p++) { dst [2*p] = Ax [p] ; dst [2*p+1] = Az [p] ; }
        vals = dst ;
    }
    int st = b200_upload_matrix (f, Ap, Ai, Anz, vals) ;
    if (st < 0) return st ;
    if (Lx) ensure_registered (f, Lx, (size_t) f->plan.xsize * 2 * sizeof (R)) ;
    return factor_device_T<cx<R>> (f, beta, quick, minor, Lx) ;
}

extern "C" {

int b200_super_numeric_z (b200_factor *f, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const double *Ax, const double *Az, double beta, int quick_return_if_not_posdef, double *Lx, int64_t *minor)
{
    return super_numeric_complex<double> (f, Ap, Ai, Anz, Ax, Az, beta, quick_return_if_not_posdef, Lx, minor) ;
}

int b200_super_numeric_c (b200_factor *f, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const float *Ax, const float *Az, double beta, int quick_return_if_not_posdef, float *Lx, int64_t *minor)
{
    return super_numeric_complex<float> (f, Ap, Ai, Anz, Ax, Az, beta, quick_return_if_not_posdef, Lx, minor) ;
}

// (re,im) pairs from split real / imaginary arrays of nz entries of the context's precision
static const void *interleave_zomplex (b200_factor *f, std::vector<unsigned char> &buf, int64_t ncols,
    const int64_t *Ap, const int64_t *Anz, const void *Ax, const void *Az)
{
    if (!Az) return Ax ;
    if (!Ap || !dtype_complex (f->dtype)) return nullptr ;
    int64_t nz = Ap [ncols] ;
    if (Anz) for (int64_t k = 0 ; k < ncols ; k++) nz = (Ap [k] + Anz [k] > nz) ? Ap [k] + Anz [k] : nz ;
    if (nz < 0 || (nz > 0 && !Ax)) return nullptr ;
    const size_t rs = f->esize / 2 ;
    buf.resize ((size_t) nz * f->esize + 1) ;
    unsigned char *dst = buf.data () ;
    const unsigned char *re = (const unsigned char *) Ax, *im = (const unsigned char *) Az ;
    for (int64_t p = 0 ; p < nz ; p++)
    {
        memcpy (dst + (size_t) p * 2 * rs, re + (size_t) p * rs, rs) ;
        memcpy (dst + (size_t) p * 2 * rs + rs, im + (size_t) p * rs, rs) ;
    }
    return dst ;
}

int b200_super_numeric_aat (b200_factor *f, int64_t ncol, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const void *Ax, const void *Az, const int64_t *Fp, const int64_t *Fi, const int64_t *Fnz, const void *Fx,
    const void *Fz, double beta, int quick_return_if_not_posdef, void *Lx, int64_t *minor)
{
    if (!f) return B200_GPU_PROBLEM ;
    if ((Az != nullptr) != (Fz != nullptr)) return B200_INVALID ;      // A and F have the same xtype (SNUM:151)
    const void *av, *fv ;
    try
    {
        av = interleave_zomplex (f, f->h_Ax, ncol, Ap, Anz, Ax, Az) ;
        fv = interleave_zomplex (f, f->h_Fx, f->plan.n, Fp, Fnz, Fx, Fz) ;
    }
    catch (const std::bad_alloc &) { return B200_OUT_OF_MEMORY ; }
    if ((Az && !av) || (Fz && !fv)) return B200_INVALID ;
    int st = b200_upload_matrix_aat (f, ncol, Ap, Ai, Anz, av, Fp, Fi, Fnz, fv) ;
    if (st < 0) return st ;
    if (Lx) ensure_registered (f, Lx, (size_t) f->plan.xsize * f->esize) ;
    return B200_DISPATCH (f->dtype, factor_device_T) (f, beta, quick_return_if_not_posdef, minor, Lx) ;
}

}   // extern "C"

// which: bit 0 = the forward sweep (L y = b), bit 1 = the backward sweep (L^H x = y)
template <typename T>
static int solve_T (b200_factor *f, int nrhs, int which)
{
    const Plan &P = f->plan ;
    cudaStream_t s = f->stream ;
    typedef typename solve_x<T>::type XT ;
    const T *Lx = reinterpret_cast<const T *> (f->d_Lx.p) ;
    XT *X = reinterpret_cast<XT *> (f->d_X.p) ;
    XT *Y = reinterpret_cast<XT *> (f->d_Y.p) ;
    if (!f->solve_ready)
    {
        // the schedule is built (and uploaded) at the first solve
        if (const char *e = getenv ("CHOLMOD_B200_SOLVE_PDL")) f->solve_pdl = atoi (e) != 0 ;
        SolveSchedule S ;
        build_solve_schedule (P, f->h_super, f->h_pi, S) ;
        f->solve_cols = S.cols ;
        f->solve_launches = S.launches ;
        const std::vector<int32_t> &small = S.small ;
        const std::vector<SolveBlk> &blks = S.blks, &steps = S.steps ;
        const std::vector<SolveTile> &tiles = S.tiles ;
        const std::vector<SolveTile256> &tiles256 = S.tiles256 ;
        bool ok = f->d_ssmall.upload (small, s) == cudaSuccess ;
        ok = ok && f->d_sblk.upload (blks, s) == cudaSuccess ;
        ok = ok && f->d_stile.upload (tiles, s) == cudaSuccess ;
        ok = ok && f->d_stile256.upload (tiles256, s) == cudaSuccess ;
        ok = ok && f->d_sstep.upload (steps, s) == cudaSuccess ;
        // the whole-block inverses are an accelerator, not a requirement: without the memory for them (or with
        // CHOLMOD_B200_SOLVE_W256=0) the diagonal solves go through the 64 x 64 inverses
        f->d_solveW256.release () ;
        const char *ew = getenv ("CHOLMOD_B200_SOLVE_W256") ;
        if (ok && !steps.empty () && !(ew && atoi (ew) == 0))
        {
            size_t free_b = 0, total_b = 0 ;
            const size_t need = steps.size () * (size_t) (SOLVE_W256_LD * SOLVE_W256_LD) * sizeof (XT) ;
            if (cudaMemGetInfo (&free_b, &total_b) == cudaSuccess && need < free_b / 2
                && f->d_solveW256.alloc (need) != cudaSuccess) cudaGetLastError () ;
        }
        ok = ok && f->d_solveW.alloc (blks.size () * (size_t) (SOLVE_WLD * SOLVE_WLD) * sizeof (XT)) == cudaSuccess ;
        if (!ok) { cudaGetLastError () ; return B200_OUT_OF_MEMORY ; }
        CK (cudaStreamSynchronize (s), "solve plan upload") ;
        f->solve_ready = true ;
        f->solve_w_valid = false ;
    }
    const size_t nblk = f->d_sblk.n ;
    XT *Winv = reinterpret_cast<XT *> (f->d_solveW.p) ;
    const XT *W256 = reinterpret_cast<const XT *> (f->d_solveW256.p) ;
    if (nblk > 0)
    {
        if (!f->solve_w_valid)
        {
            k_solve_inv<T><<<(unsigned) nblk, 256, solve_inv_smem_bytes<T> (), s>>> (f->d_sblk.p, f->d_pi.p, f->d_px.p, Lx, Winv) ;
            if (f->d_solveW256.p && f->d_sstep.n > 0)
                k_solve_inv256<T><<<(unsigned) f->d_sstep.n, 256, solve_inv256_smem_bytes<T> (), s>>> (f->d_sstep.p, f->d_pi.p, f->d_px.p,
                    Lx, Winv, reinterpret_cast<XT *> (f->d_solveW256.p)) ;
            f->solve_w_valid = true ;
        }
        if (f->d_solve_cnt.n < nblk * (size_t) nrhs)
        {
            if (f->d_solve_cnt.alloc (nblk * (size_t) nrhs) != cudaSuccess) { cudaGetLastError () ; return B200_OUT_OF_MEMORY ; }
            CK (cudaMemsetAsync (f->d_solve_cnt.p, 0, f->d_solve_cnt.bytes (), s), "solve counters") ;
        }
    }
    const size_t nl = f->solve_launches.size () ;
    size_t nev = 0 ;
    if (f->profile)
        while (f->prof_ev.size () < 4 * nl)
        {
            cudaEvent_t e ;
            CK (cudaEventCreate (&e), "profile event") ;
            f->prof_ev.push_back (e) ;
        }
    for (size_t pass = 0 ; pass < 2 ; pass++)
    {
        if (!(which & (1 << pass))) continue ;
        // the wide supernodes hand their forward results to the backward sweep through Y: a backward sweep on its
        // own starts from the right-hand side there too
        if (pass == 1 && which == 2 && nblk > 0)
            CK (cudaMemcpyAsync (Y, X, (size_t) P.n * (size_t) nrhs * sizeof (XT), cudaMemcpyDeviceToDevice, s), "solve Y <- X") ;
        for (size_t q = 0 ; q < nl ; q++)
        {
            const b200_factor::SolveLaunch &L = f->solve_launches [pass == 0 ? q : nl - 1 - q] ;
            if (L.count <= 0 || (pass == 0 && L.kind == 5) || (pass == 1 && (L.kind == 3 || L.kind == 4))) continue ;
            const unsigned grid = (unsigned) L.count ;
            if (f->profile) cudaEventRecord (f->prof_ev [nev++], s) ;
            if (L.kind == 0)
            {
                const size_t sm = solve_small_smem_bytes<T> (L.maxcol) ;
                if (pass == 0)
                    k_solve_fwd<T><<<grid, 128, sm, s>>> (f->d_ssmall.p + L.first, f->d_super.p, f->d_pi.p, f->d_px.p,
                        f->d_Ls.p, Lx, X, P.n, nrhs, (int) L.maxcol) ;
                else
                    k_solve_bwd<T><<<grid, 128, sm, s>>> (f->d_ssmall.p + L.first, f->d_super.p, f->d_pi.p, f->d_px.p,
                        f->d_Ls.p, Lx, X, P.n, nrhs, (int) L.maxcol) ;
            }
            else if (L.kind >= 3 && L.kind <= 5)
            {
                cudaLaunchConfig_t cfg = {} ;
                cfg.gridDim = dim3 (grid) ; cfg.blockDim = dim3 (256) ; cfg.dynamicSmemBytes = 0 ; cfg.stream = s ;
                cudaLaunchAttribute at [1] ;
                at [0].id = cudaLaunchAttributeProgrammaticStreamSerialization ;
                at [0].val.programmaticStreamSerializationAllowed = 1 ;
                cfg.attrs = at ; cfg.numAttrs = f->solve_pdl ? 1 : 0 ;
                const SolveTile256 *tl = f->d_stile256.p + L.first ;
                const int64_t *dsuper = f->d_super.p, *dpi = f->d_pi.p, *dpx = f->d_px.p ;
                const int32_t *dLs = f->d_Ls.p ;
                const XT *cW = Winv, *cX = X ;
                int *cnt = f->d_solve_cnt.p ;
                const int64_t nn = P.n ;
                cudaError_t e ;
                if (L.kind == 3) e = cudaLaunchKernelEx (&cfg, k_solve_diag_fwd256<T>, tl, dsuper, dpi, dpx, Lx, cW, W256, cX, Y, nn, nrhs) ;
                else if (L.kind == 4) e = cudaLaunchKernelEx (&cfg, k_solve_fwd256<T>, tl, dsuper, dpi, dpx, dLs, Lx, cW, W256, X, Y, cnt, nn, nrhs) ;
                else e = cudaLaunchKernelEx (&cfg, k_solve_bwd256<T>, tl, dsuper, dpi, dpx, dLs, Lx, cW, W256, X, Y, cnt, nn, nrhs) ;
                CK (e, "superstep launch") ;
            }
            else
            {
                if (pass == 0)
                    k_solve_fwd_blk<T><<<grid, 256, 0, s>>> (f->d_stile.p + L.first, f->d_super.p, f->d_pi.p, f->d_px.p,
                        f->d_Ls.p, Lx, Winv, X, Y, P.n, nrhs) ;
                else
                    k_solve_bwd_blk<T><<<grid, 256, 0, s>>> (f->d_stile.p + L.first, f->d_super.p, f->d_pi.p, f->d_px.p,
                        f->d_Ls.p, Lx, Winv, X, Y, f->d_solve_cnt.p, P.n, nrhs) ;
            }
            if (f->profile) cudaEventRecord (f->prof_ev [nev++], s) ;
        }
        // a forward sweep on its own: gather the wide supernodes' results (in Y) into X
        if (pass == 0 && which == 1 && nblk > 0)
            k_solve_merge<XT><<<(unsigned) nblk, 64, 0, s>>> (f->d_sblk.p, f->d_super.p, X, Y, P.n, nrhs) ;
    }
    CK (cudaGetLastError (), "solve kernels") ;
    if (f->profile)
    {
        CK (cudaStreamSynchronize (s), "profile sync") ;
        for (int k = 0 ; k < 4 ; k++) { f->ms_solve_part [k] = 0 ; f->solve_launch_count [k] = 0 ; }
        size_t e = 0 ;
        for (size_t pass = 0 ; pass < 2 ; pass++)
            for (size_t q = 0 ; q < nl ; q++)
            {
                const b200_factor::SolveLaunch &L = f->solve_launches [pass == 0 ? q : nl - 1 - q] ;
                if (L.count <= 0 || !(which & (1 << pass))) continue ;
                if ((pass == 0 && L.kind == 5) || (pass == 1 && (L.kind == 3 || L.kind == 4))) continue ;
                float ms = 0 ;
                cudaEventElapsedTime (&ms, f->prof_ev [e], f->prof_ev [e+1]) ;
                e += 2 ;
                const int k = (int) pass * 2 + (L.kind == 0 ? 0 : 1) ;
                f->ms_solve_part [k] += ms ; f->solve_launch_count [k]++ ;
            }
    }
    return B200_OK ;
}

extern "C" {

static int solve_host (b200_factor *f, const double *B, double *X, int64_t nrhs, bool want_complex, int which = 3)
{
    int st = set_device (f) ; if (st) return st ;
    if (dtype_complex (f->dtype) != want_complex || which < 1 || which > 3) return B200_INVALID ;
    if (f->plan.nranks > 1)
    {
        fprintf (stderr, "cholmod_b200: a distributed factor is not solved on the device: no rank holds all of it "
            "(solve on the host, or upload the assembled L->x into a one-GPU context)\n") ;
        return B200_INVALID ;
    }
    const int64_t n = f->plan.n ;
    if (nrhs < 0 || (!B && n * nrhs > 0) || (!X && n * nrhs > 0)) return B200_INVALID ;
    if (n == 0 || nrhs == 0) return B200_OK ;
    size_t cnt = (size_t) n * (size_t) nrhs * (want_complex ? 2 : 1) ;      // doubles
    if (f->d_X.n < cnt && f->d_X.alloc (cnt) != cudaSuccess) return B200_OUT_OF_MEMORY ;
    if (f->d_Y.n < cnt && f->d_Y.alloc (cnt) != cudaSuccess) return B200_OUT_OF_MEMORY ;
    cudaStream_t s = f->stream ;
    CK (cudaMemcpyAsync (f->d_X.p, B, cnt * 8, cudaMemcpyHostToDevice, s), "H2D B") ;
    CK (cudaEventRecord (f->ev [0], s), "record") ;
    st = B200_DISPATCH (f->dtype, solve_T) (f, (int) nrhs, which) ;
    if (st) return st ;
    CK (cudaEventRecord (f->ev [1], s), "record") ;
    CK (cudaMemcpyAsync (X, f->d_X.p, cnt * 8, cudaMemcpyDeviceToHost, s), "D2H X") ;
    CK (cudaStreamSynchronize (s), "solve") ;
    cudaEventElapsedTime (&f->ms_solve, f->ev [0], f->ev [1]) ;
    return B200_OK ;
}

int b200_solve_d (b200_factor *f, const double *B, double *X, int64_t nrhs)
{
    return solve_host (f, B, X, nrhs, false) ;
}

int b200_solve_z (b200_factor *f, const double *B, double *X, int64_t nrhs)
{
    return solve_host (f, B, X, nrhs, true) ;
}

int b200_solve_sweep (b200_factor *f, int which, const double *B, double *X, int64_t nrhs)
{
    if (!f) return B200_GPU_PROBLEM ;
    return solve_host (f, B, X, nrhs, dtype_complex (f->dtype), which) ;
}

}   // extern "C"
