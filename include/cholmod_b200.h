/* cholmod_b200.h -- C ABI of the CHOLMOD/B200 module: a B200-native (sm_100a) supernodal numeric
 * Cholesky for CHOLMOD.  It replaces CHOLMOD/GPU/t_cholmod_gpu.c + cholmod_gpu.c +
 * cholmod_gpu_kernels.cu and the useGPU branch of CHOLMOD/Supernodal/t_cholmod_super_numeric_worker.c
 * of SuiteSparse 7.8.3 / CHOLMOD 5.3.0.  Paths cited below are relative to the SuiteSparse tree.
 *
 * Everything here is plain pointers and sizes: no cholmod.h, no torch, no C++ types.  Indices are
 * int64 (the reference GPU path is "cholmod_l_" only, t_cholmod_super_numeric_worker.c:174),
 * values are double (…_d) or float (…_s).  The cholmod_common / cholmod_factor level glue
 * (cholmod_l_b200_* and the rd_/rs_ worker entry) is the thin C file csrc/cholmod_b200_shim.c;
 * INTEGRATION.md shows where a maintainer hooks it into the reference tree.
 *
 * Status codes are CHOLMOD's (CHOLMOD/Include/cholmod.h:829-840).  There is NO CPU fallback: if the
 * device or the kernels are unavailable every entry point fails with B200_GPU_PROBLEM.
 */
#ifndef CHOLMOD_B200_H
#define CHOLMOD_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK              0      /* CHOLMOD_OK */
#define B200_NOT_POSDEF      1      /* CHOLMOD_NOT_POSDEF (a warning: the call itself succeeded) */
#define B200_OUT_OF_MEMORY (-2)     /* CHOLMOD_OUT_OF_MEMORY */
#define B200_TOO_LARGE     (-3)     /* CHOLMOD_TOO_LARGE */
#define B200_INVALID       (-4)     /* CHOLMOD_INVALID */
#define B200_GPU_PROBLEM   (-5)     /* CHOLMOD_GPU_PROBLEM */

#define B200_DOUBLE 0               /* CHOLMOD_DOUBLE */
#define B200_SINGLE 4               /* CHOLMOD_SINGLE */
#define B200_COMPLEX_DOUBLE 2       /* CHOLMOD_COMPLEX + CHOLMOD_DOUBLE: entries are (re,im) pairs of doubles */
#define B200_COMPLEX_SINGLE 6       /* CHOLMOD_COMPLEX + CHOLMOD_SINGLE: (re,im) pairs of floats */

/* Opaque factorization context: device-resident symbolic tables, the launch plan derived from them,
 * the L->x arena in HBM, streams and events.  One per (analysis, device); it is reused by every
 * numeric factorization of the same symbolic factor (cholmod_factorize.c:67-70). */
typedef struct b200_factor b200_factor ;

/* 1 if `device` exists and is an sm_100 part this library can run on, else 0.
 * Replaces cholmod_l_gpu_probe (CHOLMOD/GPU/cholmod_gpu.c:148-177). */
int b200_probe (int device) ;

/* Number of visible CUDA devices (0 when there is no driver). */
int b200_device_count (void) ;

/* Free and total bytes of HBM on `device`.  Replaces cholmod_l_gpu_memorysize (cholmod_gpu.c:63-136).
 * Returns B200_OK or B200_GPU_PROBLEM. */
int b200_memorysize (int device, size_t *free_bytes, size_t *total_bytes) ;

/* Cap, in bytes, on the device memory all contexts of this library may hold together on `device` (0 = no cap,
 * the default).  b200_create / b200_create_dist fail with B200_OUT_OF_MEMORY, before allocating anything,
 * when the new context (symbolic tables + plan + the L->x arena) would not fit under it.  Replaces the pool
 * sizing of cholmod_l_gpu_allocate from Common->maxGpuMemBytes / maxGpuMemFraction and CHOLMOD_GPU_MEM_BYTES /
 * CHOLMOD_GPU_MEM_FRACTION (CHOLMOD/GPU/cholmod_gpu.c:426-543, cholmod_super_symbolic.c:264-281). */
int b200_set_memory_limit (int device, size_t bytes) ;

/* Build the context for one supernodal symbolic factor: L->super [nsuper+1], L->pi [nsuper+1],
 * L->px [nsuper+1], L->s [pi[nsuper]] exactly as cholmod_super_symbolic2 produced them
 * (CHOLMOD/Supernodal/cholmod_super_symbolic.c:147-938).  Computes -- on the host, once -- everything
 * the reference recomputes per factorization and that is purely symbolic: the supernodal etree and its
 * levels, the list of (descendant, ancestor) update pairs with ndrow1/ndrow2
 * (t_cholmod_super_numeric_worker.c:686-698, :950-967), their relative maps (:897-903), and the
 * level-ordered launch plan; then uploads the tables and allocates the L->x arena.
 * Replaces gpu_init (CHOLMOD/GPU/t_cholmod_gpu.c:84-160) and cholmod_l_gpu_allocate
 * (cholmod_gpu.c:426-543).  device < 0 builds the host-side plan only (no CUDA call is made; used by
 * the CPU tests of the planner) -- such a context cannot factorize.
 * dtype is B200_DOUBLE, B200_SINGLE, B200_COMPLEX_DOUBLE or B200_COMPLEX_SINGLE (the factor is then
 * Hermitian: A = L*L^H).  On failure returns NULL and sets *status. */
b200_factor *b200_create (int64_t n, int64_t nsuper, const int64_t *super, const int64_t *pi,
    const int64_t *px, const int64_t *s, int dtype, int device, int *status) ;

/* --- one factorization spread over the GPUs of one box (the reference has no counterpart: "this version
 * of CHOLMOD can only utilize a single GPU", CHOLMOD/GPU/cholmod_gpu.c:138-143) ------------------------
 * Every rank (one per GPU; one process each, or all in one process -- see b200_multi_create) calls
 * b200_create_dist with the same symbolic factor, its own device, the number of ranks and its rank, and gets its
 * share of one common plan:
 *   - the supernodal etree is cut into independent subtrees, each owned -- stored, factorized, copied to the
 *     host -- by one rank, and a top part (the big separators) whose supernodes are split over all ranks by
 *     256-column blocks (1-D block cyclic);
 *   - the updates a rank's subtrees send into the top separators are accumulated on that rank, in its own
 *     zero-initialised copies of them: its contribution blocks.  After a barrier the owner of each block adds
 *     the other ranks' contribution blocks to its own, reading them from their arenas over NVLink;
 *   - the top separators are then factorized cooperatively; only their finished panels travel
 *     (ncclBroadcast).
 * A rank's arena holds the top separators and its own subtrees, nothing else ("device_bytes").  Setting up:
 * rank 0 makes a NCCL unique id (b200_comm_unique_id, 128 bytes) and every rank an arena handle
 * (b200_peer_handle, 64 bytes); the host program carries them around by any means (torch.distributed, MPI, a
 * file); every rank then calls b200_comm_init and, for every other rank, b200_peer_attach.  After that all
 * ranks make the same sequence of b200_upload_* / b200_factor_device / b200_super_numeric_* calls, with the same
 * matrix (rank 0's copy is the one that is used: it goes over PCIe once and to the other GPUs over NVLink).  The host Lx of those calls
 * receives only what the rank owns, at its place in L->x: hand every rank a pointer to the SAME memory (one
 * buffer in one process, a shared mapping across processes) and it comes out complete, each rank's share over
 * its own PCIe link.  b200_download_factor likewise.  A distributed factor is not solved on the device
 * (b200_solve_* return B200_INVALID): no rank holds all of it.  nranks == 1 is b200_create. */
b200_factor *b200_create_dist (int64_t n, int64_t nsuper, const int64_t *super, const int64_t *pi,
    const int64_t *px, const int64_t *s, int dtype, int device, int nranks, int rank, int *status) ;
int b200_comm_unique_id (void *id128) ;
int b200_comm_init (b200_factor *f, const void *id128) ;

/* The arena of a distributed context, for the other ranks to attach to: its device address (ranks in the same
 * process) or a 64-byte inter-process handle (cudaIpcMemHandle_t; ranks in other processes). */
void *b200_peer_arena (b200_factor *f) ;
int b200_peer_handle (b200_factor *f, void *handle64) ;
/* Make rank `rank`'s arena readable from this context's device: pass its address (same process; peer access
 * between the two devices is enabled) or its handle (other process; opened with cudaIpcOpenMemHandle), the
 * other argument NULL.  Every other rank has to be attached before the first factorization. */
int b200_peer_attach (b200_factor *f, int rank, void *arena, const void *handle64) ;

/* The same factorization driven from ONE process: `ngpus` contexts on devices first_device .. first_device +
 * ngpus - 1, communicator and peer attachments set up inside, one host thread per device for the duration of
 * a call.  This is the form an unmodified cholmod_l_factorize caller reaches (the shim's CHOLMOD_B200_NGPUS).
 * b200_multi_super_numeric_d has the contract of b200_super_numeric_d: one host A in, one complete host L->x
 * out, every device moving its share. */
typedef struct b200_multi b200_multi ;
b200_multi *b200_multi_create (int64_t n, int64_t nsuper, const int64_t *super, const int64_t *pi,
    const int64_t *px, const int64_t *s, int dtype, int first_device, int ngpus, int *status) ;
void b200_multi_destroy (b200_multi *m) ;
int b200_multi_ngpus (const b200_multi *m) ;
b200_factor *b200_multi_context (b200_multi *m, int rank) ;
int b200_multi_super_numeric_d (b200_multi *m, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const double *Ax, double beta, int quick_return_if_not_posdef, double *Lx, int64_t *minor) ;
int b200_multi_download_factor (b200_multi *m, void *Lx) ;
int b200_multi_forget_host_buffer (b200_multi *m, const void *Lx) ;

/* Release everything.  Replaces cholmod_l_gpu_deallocate / gpu_end (cholmod_gpu.c:187-228, :319-420). */
void b200_destroy (b200_factor *f) ;

/* Integer property of the context by name: "n", "nsuper", "xsize", "nlevels", "npairs", "npairs_gemm" (pairs
 * with rows below the ancestor's columns: the reference's dgemm calls), "nsuper_trsm" (its dtrsm calls),
 * "relmap_len", "ntiles", "nlaunches", "device_bytes", "device", "kernel_launches" (kernels launched
 * by the last factorization), "ngpus", "rank"; for a distributed context "dxsize" (values in this rank's arena),
 * "top_size" (of which top separators), "npulls", "nvlink_in_bytes" (what one factorization brings in over NVLink),
 * "peers_attached", "reserved_sms" (SMs the throughput streams stay off), "host_registered" (bytes of the caller's
 * L->x this context has page-locked); "device_symbolic" (1: pairs and relative maps were built on the device).
 * Returns -1 for an unknown name. */
int64_t b200_get_int (const b200_factor *f, const char *name) ;

/* Floating-point property by name: "flops_syrk", "flops_gemm", "flops_potrf", "flops_trsm" (the flops of the
 * reference's four BLAS/LAPACK call kinds), "flops_executed" (supernodal flop count of the plan:
 * sum over pairs of ndcol*ndrow1*(2*ndrow2-ndrow1) plus nscol^3/3 + nscol^2*(nsrow-nscol) per
 * supernode), "ms_factor" (device time of the last b200_factor_device, CUDA events), "ms_h2d",
 * "ms_d2h", "ms_scatter", "ms_solve".  Returns -1 for an unknown name. */
double b200_get_double (const b200_factor *f, const char *name) ;

/* --- the numeric phase, host buffers in and out (the drop-in call) --------------------------------
 * Same contract as rd_/rs_cholmod_super_numeric_worker (t_cholmod_super_numeric_worker.c:133-145) for
 * the symmetric case (A->stype < 0, F unused): A = tril (P A P') in CSC (Ap [n+1], Ai, Ax; Anz [n]
 * non-NULL iff A is unpacked; columns may be unsorted; entries outside the pattern of L are dropped,
 * :455-468), beta [0] is added to the diagonal (:510-520).  Fills host Lx [xsize] with the supernodal
 * factor in the reference layout (column-major per supernode, lda = nsrow, :342-348).  Lx == NULL (any of the
 * b200_super_numeric_* calls): the factor stays on the device only -- solve with b200_solve_*, fetch it
 * later with b200_download_factor.
 * Returns B200_OK, or B200_NOT_POSDEF with *minor = the failing column and L zeroed / partially
 * factorized exactly as :1084-1146 prescribe (quick_return_if_not_posdef as in cholmod_common), or a
 * negative status on a fatal error.  *minor = n on success (:1269). */
int b200_super_numeric_d (b200_factor *f, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const double *Ax, double beta, int quick_return_if_not_posdef, double *Lx, int64_t *minor) ;
int b200_super_numeric_s (b200_factor *f, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const float *Ax, double beta, int quick_return_if_not_posdef, float *Lx, int64_t *minor) ;

/* The Hermitian case: cd_/zd_ and cs_/zs_cholmod_super_numeric_worker (the COMPLEX and ZOMPLEX variants of
 * t_cholmod_super_numeric_worker.c, selected at cholmod_super_numeric.c:262-322; zherk/zgemm/zpotrf/ztrsm at
 * :786-803, :850-876, :1030-1048, :1184-1208).  A = tril of a Hermitian matrix.  Az == NULL: Ax holds nnz
 * (re,im) pairs (CHOLMOD_COMPLEX); Az != NULL: Ax holds the real and Az the imaginary parts
 * (CHOLMOD_ZOMPLEX).  Lx [2*xsize] receives (re,im) pairs whatever the input was, as in the reference
 * (a supernodal L is never zomplex).  Only the real part of beta exists (:510-520); the imaginary part of
 * the diagonal of A is ignored and the diagonal of L is real, as zpotrf does.  The context must have been
 * created with B200_COMPLEX_DOUBLE (_z) or B200_COMPLEX_SINGLE (_c). */
int b200_super_numeric_z (b200_factor *f, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const double *Ax, const double *Az, double beta, int quick_return_if_not_posdef, double *Lx, int64_t *minor) ;
int b200_super_numeric_c (b200_factor *f, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const float *Ax, const float *Az, double beta, int quick_return_if_not_posdef, float *Lx, int64_t *minor) ;

/* The unsymmetric case, A->stype == 0 (cholmod_super_numeric.c:136-154, t_cholmod_super_numeric_worker.c:471-506):
 * the matrix that is factorized is A*F + beta*I, where A is n-by-ncol (rows already permuted:
 * A(p,f) of cholmod_factorize.c:236-247) and F = A(p,f)^H is ncol-by-n, both CSC, packed or not (Anz / Fnz),
 * values of the context's type (any of the four); A*F is never formed: each column of L gathers its part of
 * it.  Az and Fz are both NULL, or both non-NULL for zomplex input to a complex context (real parts in Ax /
 * Fx, imaginary parts in Az / Fz).  Lx [xsize] of the context's type.  Status and *minor as above. */
int b200_super_numeric_aat (b200_factor *f, int64_t ncol, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const void *Ax, const void *Az, const int64_t *Fp, const int64_t *Fi, const int64_t *Fnz, const void *Fx,
    const void *Fz, double beta, int quick_return_if_not_posdef, void *Lx, int64_t *minor) ;

/* --- the same phase in three steps, for callers that keep A and L resident in HBM ---------------- */

/* Copy A (host CSC as above, values of the context's dtype) to the device.  The pattern may change from
 * call to call.  Replaces the per-supernode host scatter + H2D of t_cholmod_gpu.c:669-700. */
int b200_upload_matrix (b200_factor *f, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const void *Ax) ;

/* The same for the unsymmetric case: A (n-by-ncol) and F (ncol-by-n) as for b200_super_numeric_aat, values
 * of the context's type ((re,im) pairs for a complex context).  The next b200_factor_device factorizes
 * A*F + beta*I. */
int b200_upload_matrix_aat (b200_factor *f, int64_t ncol, const int64_t *Ap, const int64_t *Ai, const int64_t *Anz,
    const void *Ax, const int64_t *Fp, const int64_t *Fi, const int64_t *Fnz, const void *Fx) ;

/* Re-factorization fast path (cholmod_factorize.c:67-70: one analysis, many numeric factorizations): the
 * pattern uploaded by the last b200_upload_matrix is kept on the device and only the values Ax (same
 * order, same count) are replaced -- 8 bytes per entry instead of 24. */
int b200_upload_values (b200_factor *f, const void *Ax) ;

/* Factorize the device-resident A into the device-resident L: zero + scatter A (:361-520), then level by
 * level the batched descendant updates (dsyrk/dgemm/assemble, :769-930) and the per-supernode
 * dpotrf/dtrsm (:1023-1181).  Blocks until the device is done.  Returns status and *minor as above. */
int b200_factor_device (b200_factor *f, double beta, int quick_return_if_not_posdef, int64_t *minor) ;

/* Copy the device-resident factor (a distributed context: the part this rank owns) to its place in host
 * Lx [xsize] (values of the context's dtype).
 * Replaces the per-supernode D2H of t_cholmod_gpu.c:1263-1273, :1403-1443. */
int b200_download_factor (b200_factor *f, void *Lx) ;

/* The host L->x passed to b200_super_numeric_* / b200_download_factor is page-locked (cudaHostRegister) the
 * first time it is seen and stays so for later factorizations.  Call this before that memory is freed
 * (cholmod_free_factor): Lx = the buffer, or NULL for "whatever is registered".  Always B200_OK unless the
 * device cannot be selected. */
int b200_forget_host_buffer (b200_factor *f, const void *Lx) ;

/* Device address of the L->x arena (xsize values of the context's dtype), or NULL. */
void *b200_device_Lx (b200_factor *f) ;

/* Overwrite the device-resident factor with host values (used to solve with a factor computed
 * elsewhere, and by tests). */
int b200_upload_factor (b200_factor *f, const void *Lx) ;

/* --- solve with the device-resident factor -------------------------------------------------------
 * X = L'\(L\B) in the permuted ordering, B and X n-by-nrhs column-major doubles on the host (X may
 * alias B).  Device counterpart of cholmod_super_lsolve / cholmod_super_ltsolve
 * (CHOLMOD/Supernodal/t_cholmod_super_solve_worker.c:20-303, :309-586). */
int b200_solve_d (b200_factor *f, const double *B, double *X, int64_t nrhs) ;

/* The same with a complex factor: X = L^H\(L\B), B and X n-by-nrhs column-major (re,im) pairs of doubles
 * (the "C" variants of the same file). */
int b200_solve_z (b200_factor *f, const double *B, double *X, int64_t nrhs) ;

/* One sweep only, for callers that drive the two halves themselves as cholmod_solve2 does
 * (CHOLMOD/Cholesky/cholmod_solve.c:696-708): which = 1 is cholmod_super_lsolve (X = L\B,
 * CHOLMOD/Supernodal/cholmod_super_solve.c:40-128), which = 2 is cholmod_super_ltsolve (X = L^H\B, :140-230),
 * which = 3 both (= b200_solve_d / _z).  B and X are doubles, (re,im) pairs for a complex context. */
int b200_solve_sweep (b200_factor *f, int which, const double *B, double *X, int64_t nrhs) ;

/* --- one panel step without any allocation ---------------------------------------------------------
 * This entry point has no counterpart in the reference.
 *
 * It exists for callers that are not allowed to reach cudaMalloc: solvers that run inside a time step or
 * control loop and have banned allocation after start-up (cudaMalloc/cudaFree synchronize the device and
 * have no latency bound), and applications that have given all of HBM to a pool of their own, so that an
 * allocation made by a library fails or upsets a budget they track themselves.  Such a caller passes a
 * device buffer it owns; the scratch the kernels need (inverse of the diagonal block, pivot flag, tile
 * descriptors: about 49 KB) is a static __device__ array inside the library, resident on every GPU that
 * loads the library from load time until the process exits.
 *
 * Work: one panel step of the blocked factorization, in place on the device.  d_panel is a column-major
 * nrows-by-jb panel with leading dimension lda (device pointer, 8-byte aligned).  On entry its leading
 * jb-by-jb block holds the lower triangle of a symmetric matrix A11 and the rows below hold A21; on return
 * the leading block holds L11 with A11 = L11*L11' (strict upper triangle untouched) and the rows below hold
 * A21 * L11^{-T} -- the dpotrf and dtrsm of t_cholmod_super_numeric_worker.c:1023 and :1175 for a supernode
 * of jb columns.  `stream` is a cudaStream_t passed as a pointer (NULL = the default stream).  The call
 * returns after the stream has finished the work.
 *
 * Limits: jb <= B200_NOALLOC_MAX_COLS, nrows <= B200_NOALLOC_MAX_ROWS and lda <= 2^31-1; anything larger is refused with
 * B200_TOO_LARGE and a message on stderr (the fixed scratch cannot hold it; use b200_create +
 * b200_super_numeric_d).  B200_INVALID for a NULL panel or inconsistent sizes.
 * Returns B200_OK, or B200_NOT_POSDEF with *info = the 1-based column of the first pivot that is not
 * positive (LAPACK's convention; *info = 0 otherwise), or B200_GPU_PROBLEM.
 *
 * The device that owns d_panel must be current.  The first call on each device looks up the device address
 * of the scratch and raises two kernel attributes; make that first call before starting any stream capture.
 * No cudaMalloc, cudaFree or other allocating call is made, on the first call or later.
 * All calls on a device share the one scratch array: each call rewrites every part of it that it reads
 * (nothing is inherited from the previous call), but two calls that overlap on the GPU -- different streams,
 * or different host threads -- race on it and corrupt each other.  Serialise them. */
#define B200_NOALLOC_MAX_COLS 64
#define B200_NOALLOC_MAX_ROWS 65536
int b200_panel_factor_noalloc_d (double *d_panel, int64_t lda, int64_t nrows, int64_t jb, int64_t *info,
    void *stream) ;

/* --- diagnostics (the reference's counterparts are the BLAS_TIMER counters and cholmod_gpu_stats,
 * CHOLMOD/Check/cholmod_check.c:615-674) -------------------------------------------------------- */

/* Set an integer option by name.  "profile" = 1 brackets every kernel launch of the next factorizations (and
 * solves) with CUDA events and accumulates device time per kernel kind: "ms_kind<k>", "launches_kind<k>" and the
 * algorithmic "flops_kind<k>" in b200_get_double / b200_get_int, k = 0 SIMT update, 1 tensor update 64x64,
 * 2 tensor update 128x64, 3..5 POTRF of <= 16 / 32 / 64 columns, 6 TRSM by substitution, 7..8 TRSM through the
 * 32 / 64-wide inverse, 9 whole small supernodes, 10 whole tiny supernodes (one thread each), 11 tiny updates (one
 * thread each); "ms_launch<i>" / "ms_launch_start<i>" are the duration and
 * start time of launch i of the plan, in the order of the "launches" table of b200_plan_export; "ms_level_end<l>"
 * when etree level l was finished; "ms_solve_part0..3" the solve's small-forward, wide-forward, small-backward
 * and wide-backward parts.  Profiling serialises nothing but adds event overhead: never time a benchmark step
 * with it on.
 * "use_graph" = 1 (or CHOLMOD_B200_GRAPH=1 in the environment; one-GPU contexts): the launch list is captured once
 * per copy-out destination as a CUDA graph and replayed by later factorizations ("graph_captures",
 * "graph_replays" in b200_get_double).  A graph is captured again when the page-locked registration of its
 * destination has changed.  Off by default: on config 5 a replay takes 23.1 ms against 24.4 ms of plain launches,
 * but the first call pays 0.5-1 s for capture and instantiation.
 * The solve of supernodes wider than 64 columns advances 256 columns per launch, with the whole 256 x 256 diagonal
 * block of every such superstep inverted once per factorization (512 KB each; "solve_w256_bytes" in
 * b200_get_double) and the superstep kernels launched programmatically dependent.  Environment, read at a context's
 * first solve: CHOLMOD_B200_SOLVE_COLS=64 selects the older 64-column schedule ("solve_cols"),
 * CHOLMOD_B200_SOLVE_W256=0 the diagonal solve through the 64 x 64 inverses (also used when the memory for the
 * 256 x 256 ones is not there), CHOLMOD_B200_SOLVE_PDL=0 plain launches. */
int b200_set_int (b200_factor *f, const char *name, int64_t value) ;

/* Copy one table of the host-side launch plan to `out` (or return its size in bytes when out == NULL):
 * "level", "parent", "relmap", "pairs", "pair_map", "items", "tiles", "potrf", "trsm", "launches", "copies",
 * "owned", "bcasts", "pulls", "dpx", "colmine", "owner", "top", "group" (layouts in csrc/plan.hpp).  Works on a
 * device < 0 context; used by the CPU tests of the planner.  "solve_small", "solve_blks", "solve_steps",
 * "solve_tiles", "solve_tiles256", "solve_launches": the schedule of the device solve (layouts in csrc/kernels.cuh
 * and tests/solve_interp.py, which replays it on the host).  "dev_pairs", "dev_pair_map", "dev_relmap": the
 * symbolic tables a one-GPU context builds on the device (csrc/symbolic.cuh: the pair walk of
 * cholmod_super_symbolic.c:896-921 and the relative maps of t_cholmod_super_numeric_worker.c:897-903), copied
 * back for comparison with the host planner's. */
int64_t b200_plan_export (const b200_factor *f, const char *which, void *out, int64_t out_bytes) ;

/* Measured FP64 tensor-pipe (DMMA, mma.sync.m8n8k4.f64) peak of `device` in TFLOP/s from a
 * register-resident loop of `iters` iterations per warp; < 0 on failure.  This is the roofline
 * denominator bench.py reports against (MEASURED_PEAKS.json carries no FP64 figure). */
double b200_dmma_peak_tflops (int device, int iters) ;

/* Library/kernels self-description, e.g. "cholmod_b200 0.1 sm_100a".  Never NULL. */
const char *b200_version (void) ;

#ifdef __cplusplus
}
#endif
#endif
