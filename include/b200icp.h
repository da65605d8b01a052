/* b200icp.h -- C ABI of libb200icp.so, the B200 (sm_100a) batched ICP engine.
 *
 * This is the drop-in boundary for the reference's ICP hot path
 * (fadli0029/LiDAR-SLAM).  Each entry point names the reference interface it
 * replaces.  Plain pointers and sizes only; no torch / C++ types.
 *
 * Ownership: the caller allocates and frees EVERY buffer (device memory unless
 * stated otherwise), including the workspace; the entries never allocate device
 * memory and are re-entrant.  All launches are asynchronous on the given CUDA
 * stream (pass a cudaStream_t cast to void*; NULL = the legacy default stream).
 * Documented exceptions: the HOST entries (b200icp_run_host*, b200icp_run_pair_host,
 * b200icp_submit_host_ranges / b200icp_wait_host) take host pointers and own grow-only
 * per-device arenas and streams (freed by b200icp_host_release; calls on one device are
 * serialised, two submitted calls may be in flight); b200icp_host_alloc / _free hand out
 * page-locked host memory; b200icp_run_static uses a fixed scratch array inside the
 * library's device image instead of a caller workspace; b200icp_ogm_build keeps one
 * auxiliary stream per device.  Functions return 0 or a negative
 * B200ICP_ERR_* code; per-pair anomalies are reported in `status_out`, never
 * as a call failure, because the reference does not raise on them either.
 */
#ifndef B200ICP_H_
#define B200ICP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ICP_VERSION 100 /* 0.1.0 */

/* ---- call-level error codes ------------------------------------------------ */
#define B200ICP_OK 0
#define B200ICP_ERR_BAD_ARG (-1)   /* null pointer, non-positive size, bad dim        */
#define B200ICP_ERR_TOO_LARGE (-2) /* n_max beyond what this build supports           */
#define B200ICP_ERR_WORKSPACE (-3) /* workspace smaller than b200icp_workspace_bytes  */
#define B200ICP_ERR_CUDA (-4)      /* launch failed; see cudaGetLastError             */
#define B200ICP_ERR_NO_DEVICE (-5) /* no sm_100 device / wrong architecture           */
#define B200ICP_ERR_BUSY (-7)      /* b200icp_submit_host_ranges: two calls are already in flight on this device */
#define B200ICP_ERR_STATIC_SCRATCH (-6) /* b200icp_run_static: the problem needs more scratch than
                                           B200ICP_STATIC_SCRATCH_BYTES; use b200icp_run       */

/* ---- per-pair status (status_out) --------------------------------------------
 * Which of the reference's three stop tests fired (modules/icp.py:172-178), or
 * why the pair was not run. */
#define B200ICP_STATUS_EPSILON 0   /* error < epsilon                      icp.py:172 */
#define B200ICP_STATUS_MAX_ITERS 1 /* iteration >= max_iters               icp.py:174 */
#define B200ICP_STATUS_STALLED 2   /* |last_err - error| < stopping_thresh icp.py:177 */
#define B200ICP_STATUS_NONFINITE 3 /* error became NaN/Inf; stopped early             */
#define B200ICP_STATUS_EMPTY 4     /* source or target scan has no points; not run    */
#define B200ICP_STATUS_NONPLANAR 5 /* 2-D entry given an init that moves points out of
                                      the plane; not run (use the 3-D entry)           */
#define B200ICP_STATUS_SKIPPED 6   /* pair index < 0 (padding entry); not run         */
#define B200ICP_STATUS_NOT_RESIDENT 7 /* gated launch: the pair's scans had not arrived
                                      within the wait limit; not run (T = init, NaN error) */

/* ---- launch flags (the `flags` argument of b200icp_run_ex and of the host entries) ---------
 * Behaviour switches are arguments, not environment look-ups on the launch path.  (The
 * developer overrides B200ICP_SWEEP=exhaustive, B200ICP_HOST_GATED=0, B200ICP_HOST_SLICES=n,
 * B200ICP_PIPELINE_CHUNKS=n are read ONCE, when the library is loaded, and OR-ed into every
 * call's flags.) */
#define B200ICP_FLAG_EXHAUSTIVE 0x1u   /* evaluate every (source, target) candidate: no block
                                          pruning.  Same results bit for bit; for A/B timing and as
                                          the FP32-issue-rate figure of the inner loop.          */
#define B200ICP_FLAG_NO_PRUNE 0x4u     /* the shipped row kernel with the block bounds switched off
                                          (every block swept, same summation order: bit-identical
                                          results); what the pruned == unpruned tests compare      */
#define B200ICP_FLAG_LATENCY 0x8u      /* few planar pairs (<= one per SM): 16 warps on each pair instead
                                          of 4, i.e. shortest time per pair instead of most pairs per
                                          second.  Another summation order: results equal to rounding
                                          (1e-13), not bit for bit; b200icp_run_pair_host sets it.    */
#define B200ICP_FLAG_HOST_CHUNKED 0x2u /* host entries: pair chunks alternating on two streams
                                          instead of ONE gated launch (A/B)                     */
#define B200ICP_FLAG_HOST_NO_DIRECT 0x10u /* host entries: always copy the results back; by default a result
                                             array in page-locked memory is written by the kernel itself */
#define B200ICP_FLAG_SLICES(n) (((unsigned)(n)&0xffu) << 8)  /* host entries: pool slices (0 = 24) */
#define B200ICP_FLAG_CHUNKS(n) (((unsigned)(n)&0xffu) << 16) /* chunked host path: chunks (0 = 6)  */

/* Library version (B200ICP_VERSION of the build). */
int b200icp_version(void);

/* Human-readable text for a B200ICP_ERR_* code (static storage). */
const char* b200icp_error_string(int code);

/* Bytes of device workspace b200icp_run / b200icp_nearest need for this problem size: the work
 * queue (256 B) and, for the general kernel (dim 3, or dim 2 with n_max > 2048), one fp32 copy of
 * a target per resident cluster.  Concurrent launches need separate workspaces.
 * After a run of the planar kernel, bytes 8..15 of the workspace hold (uint64) the number of
 * (32 source points x 32 target points) tiles whose candidate distances were really evaluated --
 * the block-pruned sweep skips the tiles that provably hold no candidate (same results as the
 * exhaustive sweep; B200ICP_FLAG_EXHAUSTIVE turns pruning off). */
int b200icp_workspace_bytes(int n_pairs, int n_max, int dim, size_t* bytes_out);

/* Batched point-to-point ICP.  Replaces, for every pair b independently,
 *
 *   run_icp(pc1, pc2, init_transform, epsilon, max_iters, stopping_thresh,
 *           return_error=True, normalize_error, return_correspondences=True)
 *                                                 -- modules/icp.py:123-189
 *   (== code/icp_warm_up/icp.py:130-196), i.e. the loop of icp_iteration
 *   (icp.py:100-121): transform, exact nearest neighbour (icp.py:29-41,161),
 *   Kabsch solve (icp.py:43-74), compose, SSE / normalised error (icp.py:76-98),
 *   and the three stop tests in the reference's order (icp.py:172-181).
 *
 * scans      [n_scans, n_max, dim] float64, row-major; scan i has counts[i] valid
 *            leading points, the rest is ignored.  dim is 2 or 3.  16-byte aligned.
 * counts     [n_scans] int32.
 * pairs      [n_pairs, 2] int32: (source scan, target scan).  The source is moved
 *            onto the target (pc1 -> pc2).  A negative source index skips the pair.
 * init       [n_pairs, 4, 4] float64 row-major homogeneous initial transforms
 *            (init_transform).  With dim == 2 it must keep z = 0 planar (what
 *            TSE3_from_TSE2, modules/utils.py:227-240, produces).
 * epsilon    [n_pairs] float64 (epsilon of each call; icp.py:127).
 * normalize  [n_pairs] uint8 (normalize_error of each call; icp.py:131).
 * max_iters, stopping_thresh: as icp.py:128-129 (same for the whole batch).
 *
 * T_out      [n_pairs, 4, 4] float64: transforms[-1] (icp.py:183-189).
 * err_out    [n_pairs] float64: error of the last executed pass (describes the
 *            transform BEFORE the final update, as in the reference).
 * iters_out  [n_pairs] int32: number of executed passes (>= 1 when run).
 * status_out [n_pairs] int32: B200ICP_STATUS_*.
 * corr_out   [n_pairs, n_max] int32 or NULL: correspondences of the last pass
 *            (index into the target scan; -1 beyond the source count).  Exact
 *            float64 nearest neighbour; the lowest index wins exact ties.
 * workspace  device scratch of at least b200icp_workspace_bytes(...) bytes.
 */
int b200icp_run(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                const int32_t* pairs, const double* init, const double* epsilon,
                const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                int32_t* corr_out, void* workspace, size_t workspace_bytes, void* stream);

/* b200icp_run for a scan pool that is still arriving (planar kernel only: dim 2, n_max <= 2048).
 * ready_scans points to a device uint32 holding the number of leading pool rows that are resident;
 * the caller raises it from another stream (e.g. a 4-byte copy after each slice of the pool) while
 * the kernel runs, and the kernel waits on it (system-scope acquire) before it stages a pair.
 * Everything else (counts, pairs, init, epsilon, normalize) must be resident at launch.
 * peer_pose_out / n_peers / peer_row0 as in b200icp_run_exchange (NULL, 0, 0: no exchange).  This is
 * how b200icp_run_host overlaps the upload of the pool with ONE launch over the whole pair list.
 * Wait limit: a pair whose scans have not become resident within 5 s of its turn is NOT run: it
 * gets status B200ICP_STATUS_NOT_RESIDENT, T = its init and a NaN error (a stalled copy stream must
 * neither hang the device nor produce poses from missing data), workspace word 5 (uint32 at byte
 * 20) is set to 1, and every later pair whose scans are still missing is given up at once.  A
 * caller that cannot tolerate missing pairs checks word 5 or the statuses (b200icp_run_host and
 * lidar_slam_b200.engine do, and fail the call). */
int b200icp_run_gated(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                      const int32_t* pairs, const double* init, const double* epsilon,
                      const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                      double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                      int32_t* corr_out, void* workspace, size_t workspace_bytes, void* stream,
                      const unsigned int* ready_scans, double* const* peer_pose_out, int n_peers,
                      long long peer_row0);

/* b200icp_run with the multi-GPU pose exchange fused into the kernel epilogue: the thread that
 * finishes pair b also stores its packed row (x, y, theta, error) -- the b200icp_pack_poses
 * layout -- into row peer_row0 + b of each of the n_peers gather buffers.  The buffers are
 * peer-mapped device memory of every rank (this rank's own included), e.g. the buffer_ptrs of a
 * torch symmetric-memory allocation, so the rows travel as plain stores over NVLink / NVSwitch
 * while the rest of the batch is still being aligned; no separate gather kernel or collective
 * runs afterwards, only a cross-rank barrier before the rows are read.  This is what returns the
 * relative poses to the pose-graph builder on every rank (main.py:148-172).
 * peer_pose_out: HOST array of n_peers (<= 16) device pointers, each 32-byte aligned and holding
 * at least peer_row0 + n_pairs rows of 4 float64.  Pairs that are not run (status 4/5/6) store
 * the pose of their init with a NaN error. */
int b200icp_run_exchange(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                         const int32_t* pairs, const double* init, const double* epsilon,
                         const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                         double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                         int32_t* corr_out, void* workspace, size_t workspace_bytes, void* stream,
                         double* const* peer_pose_out, int n_peers, long long peer_row0);

/* The fused multi-GPU pose exchange of a launch (see b200icp_run_exchange), with optional ARRIVAL
 * FLAGS that replace the cross-rank barrier after the launch: the thread that retires the launch's
 * last pair stores flag_value -- after all pose rows, release at system scope -- into the uint32 at
 * byte offset flag_offset_bytes + 4 * flag_index of EVERY peer buffer.  A consumer on any rank waits
 * until its own buffer holds flag_value in the words of all ranks (b200icp_exchange_wait) and may
 * then read every row.  flag_offset_bytes = 0: no flags (the caller synchronises the ranks). */
typedef struct b200icp_exchange {
  double* const* peer_pose_out; /* HOST array of n_peers device pointers (32-byte aligned)        */
  int n_peers;                  /* <= 16                                                          */
  long long peer_row0;          /* row of this launch's pair 0 in every gather buffer             */
  long long flag_offset_bytes;  /* multiple of 4, beyond the rows; 0 = no arrival flags           */
  int flag_index;               /* this rank's flag word                                          */
  unsigned int flag_value;      /* value that announces this launch (e.g. step number + 1)        */
} b200icp_exchange;

/* Stream-ordered wait for the arrival flags: returns (on the stream) once the n_flags uint32 words at
 * `flags` (this rank's own buffer + flag_offset_bytes) all hold a value >= `value`, or after
 * timeout_ms milliseconds, in which case *timed_out (device uint32, may be NULL) is set to 1. */
int b200icp_exchange_wait(const unsigned int* flags, int n_flags, unsigned int value, int timeout_ms,
                          unsigned int* timed_out, void* stream);

/* Stream-ordered raise of this rank's arrival flag on every peer buffer (what the kernels do when
 * their last pair retires), for callers that split one logical launch over several streams. */
int b200icp_exchange_signal(const b200icp_exchange* exchange, void* stream);

/* The general device entry: b200icp_run with every optional extra in one signature.
 * ready_scans     NULL, or the gate of b200icp_run_gated;
 * exchange        NULL, or the fused pose exchange (above);
 * flags           B200ICP_FLAG_* (0 = the shipped defaults).
 * b200icp_run, b200icp_run_gated and b200icp_run_exchange are this entry with flags = 0. */
int b200icp_run_ex(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                   const int32_t* pairs, const double* init, const double* epsilon,
                   const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                   double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                   int32_t* corr_out, void* workspace, size_t workspace_bytes, void* stream,
                   const unsigned int* ready_scans, const b200icp_exchange* exchange, unsigned flags);

/* b200icp_run_ex over a pool of RANGES instead of points (planar kernel only: n_beams <= 2048).
 * Replaces get_lidar_data (modules/localization.py:252-307) AND run_icp (modules/icp.py:123-189)
 * in one launch: the kernel builds the two scans of every pair itself -- beams with
 * range_min <= r <= range_max kept in beam order (localization.py:288-290), x = r*cos(a) + off_x,
 * y = r*sin(a) + off_y in the reference's operation order (localization.py:293-301; off = (0.13323, 0)),
 * bit-identical to b200icp_scans_from_ranges -- so the pool costs 8 bytes per beam (4 with float32
 * ranges) instead of 16 per point, in HBM and on the way there.
 * ranges      [n_scans, n_beams] float64 (ranges_dtype 0) or float32 (ranges_dtype 1; exact when
 *             every range is float32-representable, as Hokuyo millimetre data is: the caller checks).
 * cos_tab / sin_tab [n_beams] float64: cos / sin of the beam angles evaluated by the caller (numpy).
 * counts_out  [n_scans] int32 or NULL: valid beams of every scan that a pair referenced.
 * corr_out    [n_pairs, n_max] int32 or NULL with n_max = n_beams rounded up to even.
 * ready_scans NULL or the gate of b200icp_run_gated, counting resident RANGE rows.
 * Everything else as b200icp_run_ex. */
int b200icp_run_ranges(const void* ranges, int ranges_dtype, int n_scans, int n_beams, const double* cos_tab,
                       const double* sin_tab, double range_min, double range_max, double off_x, double off_y,
                       const int32_t* pairs, const double* init, const double* epsilon,
                       const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                       double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                       int32_t* corr_out, int32_t* counts_out, void* workspace, size_t workspace_bytes,
                       void* stream, const unsigned int* ready_scans, const b200icp_exchange* exchange,
                       unsigned flags);

/* b200icp_run for callers that live in HOST memory, like the reference (numpy arrays): every
 * pointer below is a host pointer (pinned or pageable), results included; the arguments mean
 * what they mean for b200icp_run.  The library owns the device side: a grow-only arena on the
 * current device, a copy stream and two compute streams.  The scan pool is streamed to the
 * device in slices while the pair list runs in chunks (consecutive pairs (i, i-1), the shape of
 * modules/localization.py:112-130, overlap copy and compute; any other pair list is correct,
 * it just waits for more of the pool).  With dim == 2, initial transforms that leave the z = 0
 * plane send the whole batch through the general kernel on z-padded scans, as the reference's
 * lifting does (icp.py:153-156).  Synchronous: outputs are filled on return.  err_out, iters_out,
 * status_out and corr_out may be NULL.  Calls on one device are serialised.
 * b200icp_host_release frees the arena of the current device. */
int b200icp_run_host(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                     const int32_t* pairs, const double* init, const double* epsilon,
                     const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                     double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                     int32_t* corr_out);
/* b200icp_run_host with the optional extras: the fused multi-GPU pose exchange (`exchange`, NULL =
 * none; its peer pointers are peer-mapped DEVICE pointers) and the launch flags (B200ICP_FLAG_*).
 * With an exchange the same entry serves one and several GPUs: every rank calls it on its shard. */
int b200icp_run_host_ex(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                        const int32_t* pairs, const double* init, const double* epsilon,
                        const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                        double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                        int32_t* corr_out, const b200icp_exchange* exchange, unsigned flags);

/* b200icp_run_ranges for callers that live in HOST memory: the reference's pipeline
 * get_lidar_data -> poses_from_scan_matching (modules/localization.py:252-307, then :95-134) with
 * the synced lidar ranges [n_scans, n_beams] handed over as they are (float64, ranges_dtype 0; or a
 * float32 copy, ranges_dtype 1).  Half (a quarter) of the bytes of the point pool cross PCIe; the
 * range rows are streamed in slices under ONE gated launch whose kernel builds the scans
 * (bit-identical to get_lidar_data) as it stages each pair.  All pointers are HOST pointers except
 * the peer pointers inside `exchange`.  counts_out [n_scans] int32 or NULL: valid beams per scan (-1 for a
 * scan no pair referenced).  Initial transforms must keep the plane (status 5 otherwise).
 * Synchronous; calls on one device are serialised. */
int b200icp_run_host_ranges(const void* ranges, int ranges_dtype, int n_scans, int n_beams, const double* cos_tab,
                            const double* sin_tab, double range_min, double range_max, double off_x, double off_y,
                            const int32_t* pairs, const double* init, const double* epsilon,
                            const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                            double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                            int32_t* corr_out, int32_t* counts_out, const b200icp_exchange* exchange,
                            unsigned flags);
/* run_icp for ONE pair with HOST arrays, the direct C form of the reference call
 *   run_icp(pc1, pc2, init_transform, epsilon, max_iters, stopping_thresh,
 *           return_error=True, normalize_error, return_correspondences)   -- modules/icp.py:123-189
 * as modules/localization.py:120 and main.py:162 make it pair after pair (the un-batched drop-in).
 * src [n_src, dim], tgt [n_tgt, dim] float64 (dim 2 or 3), init [4,4]; T_out [4,4]; err_out,
 * iters_out, status_out single values (may be NULL); corr_out [n_src] int32 or NULL.
 * One pinned staging block, one H2D copy, one launch (a latency-oriented CTA shape: 16 warps on the
 * pair), one D2H copy.  Synchronous.  A planar pair whose init leaves the plane is lifted to 3-D. */
int b200icp_run_pair_host(const double* src, int n_src, const double* tgt, int n_tgt, int dim,
                          const double* init, double epsilon, int max_iters, double stopping_thresh,
                          int normalize, double* T_out, double* err_out, int32_t* iters_out,
                          int32_t* status_out, int32_t* corr_out);
int b200icp_host_release(void);

/* b200icp_run_host_ranges in two halves, for a caller with several batches (several logs, a serving loop):
 * b200icp_submit_host_ranges enqueues the uploads, the launch and the copies back and returns a ticket
 * without waiting; b200icp_wait_host(ticket) returns once the outputs of that call are filled.  Up to TWO
 * calls may be in flight per device (B200ICP_ERR_BUSY otherwise); each runs on an arena and streams of its
 * own, so the upload and the first pairs of call k + 1 overlap the queue tail and the result copies of call
 * k (a launch of 5 000 pairs ends with the SMs 13 % idle).  Every array passed in (inputs and
 * outputs, and the exchange struct's pointers) must stay valid and untouched until the wait returns; use
 * page-locked memory (below), or the copies are not asynchronous.  Results are those of the synchronous
 * entry, bit for bit.  No counterpart in the reference (its loop is synchronous). */
int b200icp_submit_host_ranges(const void* ranges, int ranges_dtype, int n_scans, int n_beams, const double* cos_tab,
                               const double* sin_tab, double range_min, double range_max, double off_x, double off_y,
                               const int32_t* pairs, const double* init, const double* epsilon,
                               const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                               double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                               int32_t* corr_out, int32_t* counts_out, const b200icp_exchange* exchange,
                               unsigned flags, int* ticket_out);
int b200icp_wait_host(int ticket);

/* Page-locked host memory for the arrays handed to the host entries (cudaHostAlloc / cudaFreeHost).
 * Optional: the host entries accept any host pointer, but copies from and to pageable memory are
 * staged by the driver and block the calling thread, which costs the overlap the entries are built
 * for.  lidar_slam_b200.hostapi keeps its reusable result buffers in memory from here.  No counterpart
 * in the reference (numpy arrays are pageable).  b200icp_host_alloc returns NULL on failure. */
void* b200icp_host_alloc(size_t bytes);
int b200icp_host_free(void* p);

/* b200icp_run without a workspace argument, for callers that cannot allocate device memory at all
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * NO COUNTERPART IN THE REFERENCE: this entry was added.
 * Same arguments, outputs and results as b200icp_run (bit for bit: the same kernels run), but the
 * scratch memory (work queue; for the general kernel also the per-pair state, the work-item ring, the
 * partial sums and the Morton-sorted copies of the referenced scans) lives in ONE fixed array of
 * B200ICP_STATIC_SCRATCH_BYTES inside the library's device image: every GPU that loads the library
 * holds that much memory until the process exits.  The entry calls no cudaMalloc / cudaFree / any
 * allocating routine.  A problem whose b200icp_workspace_bytes(n_pairs, n_max, dim) exceeds
 * B200ICP_STATIC_SCRATCH_BYTES is rejected with B200ICP_ERR_STATIC_SCRATCH before anything is launched
 * (the planar kernel, dim 2 and n_max <= 2048, needs 256 bytes for any number of pairs; the general
 * kernel needs about 100 bytes per point of every referenced scan, e.g. 24 pairs of 3000-point
 * clouds = 7 MiB).  The array is SHARED by every call on a device and keeps the last call's contents
 * (each call re-initialises the words it needs at zero, on the given stream): two calls that overlap
 * on the GPU -- different streams, or different host threads -- RACE on it; serialise them on one
 * stream or use b200icp_run with separate workspaces. */
#define B200ICP_STATIC_SCRATCH_BYTES (16u << 20)
int b200icp_run_static(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                       const int32_t* pairs, const double* init, const double* epsilon,
                       const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                       double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                       int32_t* corr_out, void* stream);

/* One exact nearest-neighbour pass, exposed on its own for tests and for the
 * roofline measurement.  Replaces get_correspondences / KDTree.query
 * (modules/icp.py:29-41 with :161) for a batch of (query set, target set) pairs
 * taken from the same scan pool; queries are used as they are (no transform).
 * corr_out [n_pairs, n_max] int32; dist2_out [n_pairs, n_max] float64 or NULL. */
int b200icp_nearest(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                    const int32_t* pairs, int n_pairs, int32_t* corr_out, double* dist2_out,
                    void* workspace, size_t workspace_bytes, void* stream);

/* Pack results for the multi-GPU exchange: out[b] = (x, y, theta, error) with
 * x, y = T[0][3], T[1][3] and theta = atan2(T[1][0], T[0][0]), i.e. pose_from_T of
 * modules/utils.py:196-210 applied to TSE2_from_TSE3(T) (utils.py:212-225), which is what
 * the pose-graph builder consumes (main.py:152-155,163-171).  T [n,4,4], err [n],
 * out [n,4] float64. */
int b200icp_pack_poses(const double* T, const double* err, int n, double* out, void* stream);

/* FP32 FFMA issue-rate probe used as the roofline denominator: launches a
 * register-only FFMA loop on every SM.  Performs `iters` * 64 dependent-free
 * FFMAs per thread on a grid of n_blocks x 256 threads; the caller times it with
 * CUDA events.  sink [n_blocks*256] float32 receives one value per thread so the
 * loop cannot be removed. */
int b200icp_ffma_probe(float* sink, int n_blocks, int iters, void* stream);

/* Voxel-grid down-sampling on the device.  Replaces voxel_downsample
 * (modules/icp.py:4-27): voxel index floor((p - min)/voxel_size), mean of the
 * points in every occupied voxel, voxels in lexicographic (x, y, z) index order.
 * points [n, 3] float64; out [n, 3] float64 (first *n_out rows written);
 * n_out  device int32.  Workspace from b200icp_voxel_workspace_bytes. */
int b200icp_voxel_workspace_bytes(int n_points, size_t* bytes_out);
int b200icp_voxel_downsample(const double* points, int n_points, double voxel_size, double* out,
                             int32_t* n_out, void* workspace, size_t workspace_bytes,
                             void* stream);

/* ---- scan construction (the producer of the scan pool) ---------------------------------
 * Replaces get_lidar_data (modules/localization.py:252-307): for every scan, beams with
 * range_min <= r <= range_max are kept in beam order (localization.py:288-290), converted
 * x = r*cos(a) + off_x, y = r*sin(a) + off_y (localization.py:293-301; the reference's lidar
 * offset is off_x = 0.13323, off_y = 0) and written as the leading points of row `scan` of the
 * padded pool b200icp_run reads.  cos_tab / sin_tab [n_beams] are cos / sin of
 * linspace(-135 deg, +135 deg, n_beams) (localization.py:273-278) evaluated by the caller, so
 * that the result is bit-identical to numpy's.  ranges [n_scans, n_beams] float64;
 * scans_out [n_scans, n_max, 2] float64 (n_max even, rows zero-filled past the count);
 * counts_out [n_scans] int32. */
int b200icp_scans_from_ranges(const double* ranges, int n_scans, int n_beams, const double* cos_tab,
                              const double* sin_tab, double range_min, double range_max,
                              double off_x, double off_y, double* scans_out, int32_t* counts_out,
                              int n_max, void* stream);

/* Replaces statistical_outliers_removal (modules/localization.py:223-250) on the pool, in
 * place: d = sqrt(x^2 + y^2) of every valid point of every scan, threshold = mean(d) +
 * k_std * std(d) (population std, as np.std), then every scan keeps, in order, its points
 * with d < threshold; counts are updated and freed slots zeroed.  stats_out [4] device
 * float64 receives {mean, std, threshold, n_points}.  The two global sums are reduced in a
 * fixed (deterministic) order that differs from numpy's pairwise order, so mean/std agree
 * with the reference to a few ulp; a point within that of the threshold may differ. */
int b200icp_range_filter_workspace_bytes(size_t* bytes_out);
int b200icp_range_filter(double* scans, int32_t* counts, int n_scans, int n_max, double k_std,
                         double* stats_out, void* workspace, size_t workspace_bytes, void* stream);

/* ---- occupancy-grid mapping (the consumer of the poses) ----------------------------------
 * Replaces OccupancyGridMap.build_map / update_map / bresenham2D / world2grid
 * (modules/ogm.py:46-60, :149-195, :197-231, :104-126) for n_maps maps built side by side from
 * the same scans and n_maps pose sequences.  Per scan s and ray i: end point to the world
 * frame with the pose's rotation (ogm.py:160-162), start cell = world2grid(pose + (0.13323, 0))
 * (ogm.py:165-169; the reference does not rotate the offset), Bresenham line, every in-bounds
 * cell but the last gets  cell = f32(f64(cell) - log 4),  the last in-bounds cell
 * cell = f32(f64(cell) + log 4)  (ogm.py:172-188), in ray order; after each scan the grid is
 * clipped to [-20, 20] (ogm.py:191).  Results are bit-identical to the reference's float32 grid
 * (the per-cell update order is reproduced), except that the end-point rotation uses separate
 * rounded multiply/adds where numpy's BLAS may fuse them: an end point within 1 ulp of a cell
 * border can land in the neighbouring cell.
 *
 * log_odds [n_maps, width, height] float32, row-major, updated in place (zero it for a new map;
 *          values must lie in [-20, 20] on entry, which the clip guarantees for any map built by
 *          this entry or by the reference).
 * poses    [n_maps, n_scans, 3] float64 (x, y, yaw);  cos_sin [n_maps, n_scans, 2] float64 =
 *          (cos yaw, sin yaw) evaluated by the caller (bit-identical to numpy's).
 * scans / counts: the scan pool of b200icp_run with dim == 2 (robot-frame end points).
 * n_max <= 32767, n_maps <= 65535.  Workspace from b200icp_ogm_workspace_bytes (two sets of up to
 * 64 int32 count planes per map, <= 48 MB per set and map, plus the event lists).  The scans are
 * processed 64 at a time by four kernels; the per-cell replay of batch b runs on an internal
 * auxiliary stream while the ray passes of batch b + 1 run on `stream`, and `stream` is joined
 * with the auxiliary stream before the call returns (stream-ordered for the caller as usual).
 * B200ICP_OGM_EVENTS_PER_RAY (default 16) sizes the event lists; lists that do not fit are
 * replayed by an exact closed-form ray test instead. */
int b200icp_ogm_workspace_bytes(int width, int height, int n_max, int n_maps, size_t* bytes_out);
int b200icp_ogm_build(float* log_odds, int width, int height, double res, double min_x, double min_y,
                      const double* poses, const double* cos_sin, const double* scans,
                      const int32_t* counts, int n_scans, int n_max, int n_maps, void* workspace,
                      size_t workspace_bytes, void* stream);

/* The thresholding at the end of build_map (ogm.py:62-65): grid = 1 where
 * 1/(1+exp(log_odds)) > 0.5, else 0.  The float32 comparison is monotone in log_odds, so the
 * caller passes the largest float32 log-odds value that the reference maps to 1
 * (`one_at_or_below`, found once on the host with the reference's own float32 expression). */
int b200icp_ogm_binarize(const float* log_odds, uint8_t* grid, long long n_cells, float one_at_or_below,
                         void* stream);

/* ---- texture mapping (RGB-D splat onto the grid) -----------------------------------------
 * Replaces the frame loop of generate_texture_map (modules/texture_mapping.py:49-96) with its
 * helpers get_depth_image (:132-143), get_rgbi_rgbj (:145-160), vectorized_generate_point_cloud
 * (:162-224) and transform_point_cloud (:226-240): every pixel of every frame is turned into a
 * depth (float32 chain), an RGB look-up position and an optical-frame point (float64), carried
 * optical -> camera -> robot -> world, mapped to a cell with world2grid (ogm.py:118-119) and,
 * if inside the grid, painted.  The reference paints frames in order and pixels in row-major
 * order with last-writer-wins; the kernel folds (frame, pixel, colour) into one 64-bit key per
 * cell with atomicMax, which yields the same winner without ordering the work.
 *
 * keys        [width*height] uint64 device state, zero = unpainted; successive calls with
 *             increasing frame_base accumulate (frame_base + n_frames < 2^20 - 1).
 * disparity   [n_frames, img_h, img_w] float32;  rgb [n_frames, img_h, img_w, 3] uint8 (RGB).
 * T_wr        [n_frames, 16] float64 (device) row-major world <- robot (texture_mapping.py:71-81, built
 *             by the caller with numpy's cos/sin); K_inv [9] = inv(K); T_cam_from_optical [16] =
 *             inv(T_oc) (:213-222); T_robot_from_cam [16] = T_rc -- these three are HOST pointers (small
 *             constants copied into the launch).  img_h*img_w < 2^20.
 * Products and sums are rounded one by one; numpy's BLAS may fuse them, so a point within an
 * ulp of a cell border can land in the neighbouring cell. */
int b200icp_texture_paint(unsigned long long* keys, int width, int height, double res, double min_x,
                          double min_y, const float* disparity, const uint8_t* rgb, int n_frames,
                          int img_h, int img_w, int frame_base, const double* T_wr,
                          const double* K_inv, const double* T_cam_from_optical,
                          const double* T_robot_from_cam, void* stream);

/* texture_out [width, height, 3] float32 = (painted colour, else the grid_map value on all three
 * channels (texture_mapping.py:46-47; grid may be NULL = 0)) / 255 (:98). */
int b200icp_texture_resolve(const unsigned long long* keys, const uint8_t* grid, float* texture_out,
                            int width, int height, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200ICP_H_ */
