/*
 * rmb_mobility.h -- C ABI of the B200 blob-mobility product (librmb_mobility.so).
 *
 * Drop-in boundary: the module-level functions of the reference's
 * mobility/mobility.py,
 *     f(r_vectors, vec, eta, a, *args, periodic_length=L) -> flat 3N array
 * (mobility.py:222-548 PyCUDA wrappers, :1119-1341 numba wrappers).  Where the
 * reference's PyCUDA path binds one __global__ kernel per block
 * (mobility_pycuda.py:2235-2746), a maintainer binds ONE entry point,
 * rmb_mobility_product(), selected by (block, wall).  See INTEGRATION.md for
 * the ctypes stub.
 *
 * Conventions
 *  - every pointer marked "device" is a CUDA device pointer (e.g. a PyTorch
 *    tensor's data_ptr()); "host" pointers are ordinary host memory;
 *  - all device arrays are float64, C-contiguous, interleaved xyz, exactly the
 *    reference's layout; `precision` selects only the
 *    arithmetic of the O(N^2) pair kernel;
 *  - calls enqueue work on `stream` (a cudaStream_t, NULL = default stream)
 *    and never synchronise; inputs are not modified;
 *  - the return value is 0 on success, otherwise a negative RMB_E_* code or a
 *    positive cudaError_t; rmb_last_error() describes the last failure of the
 *    calling thread.  Like the reference, bad geometry (coincident blobs) is
 *    not an error: it shows up as NaN/Inf in the output.
 */
#ifndef RMB_MOBILITY_H
#define RMB_MOBILITY_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RMB_VERSION 105 /* 0.1.2: pair-symmetric kernels also for the RMB_NO_WALL rotation blocks */

/* block: which mobility block is applied (mobility.py naming in comments) */
#define RMB_UF 0    /* *_mobility_trans_times_force        (mobility.py:222,288,1119,1132) */
#define RMB_UT 1    /* *_mobility_trans_times_torque       (:427,482,1200,1213)            */
#define RMB_WF 2    /* *_mobility_rot_times_force          (:300,328,1258,1273)            */
#define RMB_WT 3    /* *_mobility_rot_times_torque         (:342,370,1301,1316)            */
#define RMB_UF_UT 4 /* *_mobility_trans_times_force_torque (:384,413), vec_b = torque      */

/* wall */
#define RMB_NO_WALL 0     /* no_wall_* : unbounded RPY                                   */
#define RMB_SINGLE_WALL 1 /* single_wall_* : Swan-Brady wall, heights clamped to z >= a,
                             damping B applied on both sides (mobility.py:52-84)          */

#define RMB_FREE_SURFACE 2 /* free_surface_mobility_trans_times_force_* (mobility.py:533,1390): RPY plus its
                             image across a stress-free plane z = 0; RMB_UF only, no clamp, no damping    */

/* precision of the pair arithmetic */
#define RMB_FP64 0
#define RMB_FP32 1

/* flags */
#define RMB_IN_PLANE 1    /* in_plane_* (mobility.py:255,454,1166,1235): z column and row zero */
#define RMB_NO_TMA 2      /* stage source tiles with plain loads instead of cp.async.bulk (debug) */
#define RMB_SYMMETRIC 4   /* evaluate every unordered pair once and update both blobs.  Exists for fp64 and fp32 (fp32 pair
                             arithmetic; the sums are fp64: 32 partners at a time in fp32, then fp64 registers / atomics), n_targets ==
                             n_sources, L[2] <= 0 (open or pseudo-periodic in x / y), blocks RMB_UF / RMB_UT / RMB_WF /
                             RMB_WT / RMB_UF_UT with RMB_SINGLE_WALL or RMB_NO_WALL; any other call (free
                             surface, target shards, periodic z) ignores the flag and takes the row kernel.  Sums are accumulated with atomics: results agree with
                             the row kernel to rounding, but are not bit-reproducible run to run.  Without this
                             flag the symmetric kernel is still chosen for eligible calls with
                             n_sources >= RMB_SYMMETRIC_MIN_N */
#define RMB_DETERMINISTIC 8 /* never use the pair-symmetric kernel: one thread per target, fixed summation order */
#define RMB_SYMMETRIC_MIN_N 8000    /* measured r02o (one-tile work items): row = symmetric at 6,000 blobs, symmetric +15 % at 8,000, +31 % at 12,000 */

/* error codes */
#define RMB_E_ARG (-1)        /* invalid argument                          */
#define RMB_E_WORKSPACE (-2)  /* workspace too small                       */
#define RMB_E_NODEVICE (-3)   /* no CUDA device / wrong architecture       */

int rmb_version(void);
const char *rmb_last_error(void);

/* Number of SMs of the current device (>0), or a negative RMB_E_* code. */
int rmb_device_sm_count(void);

/*
 * Bytes of device scratch rmb_mobility_product() needs for this problem
 * (packed source records + damping factors + per-split partial sums).
 * L is a host pointer to 3 doubles (NULL = not periodic).
 */
size_t rmb_workspace_bytes(int block, int precision, long n_targets, long n_sources, const double *L);

/* Kernels one rmb_mobility_product() call launches for this problem (2 or 3). */
int rmb_launches_per_product(int block, int wall, int precision, int flags, long n_targets, long n_sources,
                             const double *L);

/*
 * out[0:3*n_targets] = rows [target_offset, target_offset + n_targets) of
 *     B * M(r_eff) * (B * vec)            (single wall)
 *     M(r) * vec                          (no wall)
 * where M is the (block, wall) blob mobility over ALL n_sources blobs and the
 * pair sum runs over the 3^d nearest image boxes of the d periodic directions
 * (mobility_numba.py:166-197).  n_targets = n_sources, target_offset = 0 is
 * the reference's call; a rank that owns a shard of the targets passes its
 * range, and the self term stays correct because it is matched on the global
 * index.
 *
 *   r_sources   device, 3*n_sources   blob positions (unclamped)
 *   vec_a       device, 3*n_sources   force (UF, WF, UF_UT) or torque (UT, WT)
 *   vec_b       device, 3*n_sources   torque for RMB_UF_UT, else NULL
 *   L           host, 3 doubles or NULL; component <= 0: not periodic
 *   out         device, 3*n_targets
 *   workspace   device, >= rmb_workspace_bytes(...) bytes, 16-byte aligned
 */
int rmb_mobility_product(int block, int wall, int precision, int flags,
                         long n_targets, long n_sources, long target_offset,
                         const double *r_sources, const double *vec_a, const double *vec_b,
                         double eta, double a, const double *L,
                         double *out, void *workspace, size_t workspace_bytes, void *stream);

/*
 * Multi-rank form of the pair-symmetric product (fp64; block = RMB_UF / RMB_UT / RMB_WF / RMB_WT / RMB_UF_UT
 * (vec_b = torque for RMB_UF_UT, else NULL), wall = RMB_SINGLE_WALL or RMB_NO_WALL;
 * all n blobs targets and sources; L = NULL or
 * pseudo-periodic in x and/or y only: L[2] must be 0).  Every rank calls
 *   rmb_symmetric_partial()  raw pair sums, for ALL blobs, of this rank's share of the unordered pairs (row groups
 *                            rank, rank + world, ... of the triangular schedule) -> acc, overwritten;
 *                            acc: device, rmb_symmetric_acc_len(n) doubles
 *   [sum acc over the ranks, e.g. ncclAllReduce]
 *   rmb_symmetric_finish()   out = B (acc + self term) / (8 pi eta a^k) for targets [target_offset, +n_targets);
 *                            same block, wall, a, r, vec_a, vec_b as the partial call.  It reads ONLY its arguments
 *                            (the height clamp, the damping B and B * v are recomputed per target): nothing is
 *                            carried over from the partial call, whose workspace is pure scratch and may be reused
 *                            by any other product in between.
 * rank = 0, world = 1 is what rmb_mobility_product(block, wall, RMB_FP64, RMB_SYMMETRIC, ...) does.
 * workspace (partial only): >= rmb_workspace_bytes(block, RMB_FP64, n, n, L).
 */
size_t rmb_symmetric_acc_len(long n);
int rmb_symmetric_partial(int block, int wall, int flags, long n, int rank, int world, const double *r,
                          const double *vec_a, const double *vec_b, double eta, double a, const double *L, double *acc, void *workspace,
                          size_t workspace_bytes, void *stream);
int rmb_symmetric_finish(int block, int wall, int flags, long n, long target_offset, long n_targets, const double *r,
                         const double *vec_a, const double *vec_b, const double *acc, double eta, double a, double *out,
                         void *stream);

/*
 * Source-target UF product with per-blob radii: the flow at n_targets blobs
 * created by forces on n_sources other (or the same) blobs,
 *     out = B_t * M(x_eff, y_eff; radius_t, radius_s) * (B_s * force)      (wall = RMB_SINGLE_WALL)
 *     out = M(x, y; radius_t, radius_s) * force                            (RMB_NO_WALL, RMB_FREE_SURFACE)
 * in units of 1/(8 pi eta).  Replaces
 *   mobility.{single_wall,no_wall,free_surface}_mobility_trans_times_force_source_target_{numba,pycuda}
 *   (mobility/mobility.py:494, :551, :588, :1407 -> mobility_numba.py:1480-1658, :1941-2092;
 *   mobility_pycuda.py:1394) and mobility.mobility_radii_trans_times_force (:1369, sources == targets).
 * RPY for unequal radii (far / partial overlap / complete overlap, the latter also covering a target that
 * coincides with a source: there is no self branch); the wall form clamps every height to the blob's own radius
 * and damps with B_ii = z_i / radius_i where z_i < radius_i (mobility.py:87-119); pseudo-periodic image boxes
 * as in rmb_mobility_product.
 *
 *   r_targets       device, 3*n_targets     radius_targets  device, n_targets
 *   r_sources       device, 3*n_sources     radius_sources  device, n_sources
 *   force           device, 3*n_sources     out             device, 3*n_targets
 *   L               host, 3 doubles or NULL
 *   workspace       device, >= rmb_source_target_workspace_bytes(...) bytes, 16-byte aligned
 */
size_t rmb_source_target_workspace_bytes(int precision, long n_targets, long n_sources, const double *L);
int rmb_mobility_source_target(int wall, int precision, int flags, long n_targets, long n_sources,
                               const double *r_targets, const double *r_sources, const double *force,
                               const double *radius_targets, const double *radius_sources, double eta,
                               const double *L, double *out, void *workspace, size_t workspace_bytes,
                               void *stream);

/*
 * Allocation-free variant of rmb_mobility_product.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments and result as rmb_mobility_product minus the workspace:
 * the scratch (packed records, damping factors, partial sums / pair-symmetric accumulator) lives in a file-scope
 * `static __device__` array of rmb_static_workspace_max() = 64 MiB that every GPU which loads the library holds until
 * the process exits.  The call performs no cudaMalloc / cudaFree and calls no allocating library routine; inputs and
 * the output are the caller's.
 *   limit    products whose rmb_workspace_bytes(...) exceeds 64 MiB (about 450,000 blobs for an fp64 single-vector
 *            block, 330,000 for RMB_UF_UT; 80 / 112 bytes per blob plus up to 27 MB of partial sums) are rejected with RMB_E_WORKSPACE and a message in rmb_last_error()
 *   sharing  every call on a device uses the SAME array, which keeps the last call's contents; whatever must start
 *            at zero is cleared by the call itself on `stream`.  Calls that overlap on the GPU (different streams,
 *            or two host threads) race on it: serialise them, or give the concurrent ones their own workspace
 *            through rmb_mobility_product.
 *   set-up   the first call on a device resolves the array's device address (cudaGetSymbolAddress, that device
 *            current); make that first call outside stream capture.
 */
size_t rmb_static_workspace_max(void);
int rmb_mobility_product_static(int block, int wall, int precision, int flags,
                                long n_targets, long n_sources, long target_offset,
                                const double *r_sources, const double *vec_a, const double *vec_b,
                                double eta, double a, const double *L, double *out, void *stream);

/*
 * FP64 (precision = RMB_FP64) or FP32 FMA-pipe peak probe: runs `iters`
 * dependent-chain-free FMAs per thread on a full grid and returns the
 * measured rate in FLOP/s through *flops (host).  Synchronises.  Used by
 * bench.py for the roofline denominator (MEASURED_PEAKS.json has no FP64 figure).
 */
int rmb_fma_peak(int precision, int iters, double *flops, double *elapsed_ms);

#ifdef __cplusplus
}
#endif
#endif /* RMB_MOBILITY_H */
