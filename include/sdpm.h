/* include/sdpm.h -- C ABI of the B200-native semi-dense probability-mapping path.
 *
 * Drop-in boundary for the hot path of atlas-jj/ORB-SLAM-free-space-carving:
 *   reference class   include/Modeler/ProbabilityMapping.h:61-105   ("PM.h")
 *   reference code    src/Modeler/ProbabilityMapping.cc              ("PM.cc")
 *
 * The reference has no FFI for this path (it is a C++ class that nothing calls); these
 * entry points are what a maintainer's binding would call from the C++ `ProbabilityMapping`
 * methods (see INTEGRATION.md and orb-slam-free-space-carving_b200/host/ProbabilityMapping.h).
 * Plain pointers and sizes only; no CUDA, torch or OpenCV types.  All functions return an
 * `int` status (0 = SD_OK, negative = SdStatus); the reference's `void` methods silently
 * `return`/`continue` on bad input (PM.cc:141-142, 160, 393, 408) -- here bad arguments are errors.
 *
 * One sd_ctx per GPU.  A context is not thread-safe; several contexts may be driven from
 * several host threads.  Work is enqueued asynchronously on the context's CUDA streams;
 * sd_download / sd_sync wait for it.  There is NO CPU fallback: sd_create fails if no CUDA
 * device is usable.
 */
#ifndef SDPM_H
#define SDPM_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SD_MAX_NEIGHBOURS 40 /* BASELINE.json config 5 sweeps 5..40 covisible neighbours; PM.h:38 covisN = 7 */

typedef enum SdStatus {
    SD_OK = 0,
    SD_ERR_INVALID_ARG = -1,
    SD_ERR_UNKNOWN_KEYFRAME = -2,
    SD_ERR_NOT_READY = -3, /* stage prerequisites missing (e.g. neighbour without semi-dense maps) */
    SD_ERR_CUDA = -4,
    SD_ERR_NO_DEVICE = -5,
    SD_ERR_SHAPE = -6 /* key frame size differs from the context's */
} SdStatus;

/* Run-time copy of the compile-time tunables PM.h:38-49 and of the literals in PM.cc. */
typedef struct SdParams {
    float lambdaG;     /* PM.h:40 = 8    gradient-magnitude gate (PM.cc:201, 411, 562)            */
    float lambdaL;     /* PM.h:41 = 80   gradient vs epipolar-line angle, degrees (PM.cc:421)     */
    float lambdaTheta; /* PM.h:42 = 45   gradient-orientation difference, degrees (PM.cc:431)     */
    int lambdaN;       /* PM.h:43 = 3    min compatible hypotheses / neighbours (PM.cc:221,623,762) */
    double theta;      /* PM.h:47 = 0.23 THETA (double literal, PM.cc:436, 455-456)               */
    double chi2_fuse;  /* 5.99, PM.cc:917, 923                                                    */
    double chi2_inter; /* 3.84, PM.cc:710                                                         */
    float slope_limit; /* 4,    PM.cc:393                                                         */
    int border;        /* 2,    PM.cc:198-199, 492, 555, 659: must be 2 (hard-coded in every reference loop) */
    int interkf_inplace; /* 0 (default): InterKeyFrameDepthChecking reads its neighbours' maps as they were after the
                          * intra-key-frame stage (snapshot, order-independent, what multi-GPU runs need).
                          * 1: the reference's in-place behaviour (PM.cc:764, 793 overwrite depth_map_ inside the loop over
                          * key frames, PM.cc:262-307): a neighbour that has already been checked is read as checked, so
                          * the result depends on the order of the sd_interkf_check calls, as the reference's depends on
                          * its visiting order.  Single GPU only. */
} SdParams;

/* Device-resident planes of one key frame (for a GPU consumer such as a device-side CARV feed).
 * Float planes have `pitch_f32` elements per row, the u8 image `pitch_u8` bytes per row. */
typedef struct SdDevicePlanes {
    int width, height, pitch_f32, pitch_u8;
    const uint8_t* gray;        /* KeyFrame::im_                                  */
    const float* grad;          /* KeyFrame::GradImg                              */
    const float* theta;         /* KeyFrame::GradTheta (degrees)                  */
    const float* depth;         /* KeyFrame::depth_map_ after the intra-KF stage  */
    const float* sigma;         /* KeyFrame::depth_sigma_                         */
    const float* depth_checked; /* depth_map_ after InterKeyFrameDepthChecking    */
    const float* points;        /* KeyFrame::SemiDensePointSets_, H x 3W floats   */
} SdDevicePlanes;

typedef struct sd_ctx sd_ctx;

int sd_default_params(SdParams* p);

/* ProbabilityMapping::ProbabilityMapping(Map*), PM.cc:61: one context per GPU. */
int sd_create(int device, const SdParams* params /* NULL = defaults */, sd_ctx** out);
int sd_destroy(sd_ctx* ctx);

/* Key-frame image upload + gradient front end on device: the fields PM.cc reads as kf->im_,
 * kf->GradImg, kf->GradTheta, kf->I_stddev (PM.cc:186, 201, 214, 457).  K = {fx, fy, cx, cy};
 * Tcw = 3x4 row-major world-to-camera pose (KeyFrame::GetRotation/GetTranslation, KeyFrame.cc:109-119).
 * `gray` is a HOST buffer with `stride` bytes per row, borrowed for the duration of the call only when
 * it is pageable; pinned buffers must stay valid until the next sd_sync. */
int sd_upload_keyframe(sd_ctx* ctx, int kf_id, const uint8_t* gray, int w, int h, int stride, const float K[4],
                       const float Tcw[12]);
/* Same, `gray_dev` already resident in device memory of this context's GPU. */
int sd_upload_keyframe_device(sd_ctx* ctx, int kf_id, const uint8_t* gray_dev, int w, int h, int stride,
                              const float K[4], const float Tcw[12]);
/* KeyFrame::SetPose (KeyFrame.cc:70-84) after BA / loop closing; marks poseChanged (PM.cc:329). */
int sd_set_pose(sd_ctx* ctx, int kf_id, const float Tcw[12]);
int sd_release_keyframe(sd_ctx* ctx, int kf_id);

/* SemiDenseLoop per-key-frame half, PM.cc:137-257: ComputeFundamental per neighbour (189-195),
 * the pixel loop with EpipolarSearch + InverseDepthHypothesisFusion (197-231),
 * IntraKeyFrameDepthChecking + IntraKeyFrameDepthGrowing (237-238).  The caller supplies what the
 * reference computes on the host before the loop: the neighbour pick (151-160), the median in-plane
 * rotation per neighbour in degrees (170-179) and the StereoSearchConstraints interval (184;
 * NB min_invdepth is the LARGER value, PM.cc:381-382). */
int sd_semidense(sd_ctx* ctx, int kf_id, const int* nbr_ids, const float* nbr_rot_deg, int n, float min_invdepth,
                 float max_invdepth);

/* InterKeyFrameDepthChecking, PM.cc:628-799, snapshot semantics (reads the neighbours' maps as
 * they were after sd_semidense, writes this key frame's `depth_checked`). */
int sd_interkf_check(sd_ctx* ctx, int kf_id, const int* nbr_ids, int n);

/* UpdateSemiDensePointSet, PM.cc:337-367 (uses depth_checked when present, else depth). */
int sd_update_pointset(sd_ctx* ctx, int kf_id);
/* InterKeyFrameDepthChecking followed by UpdateSemiDensePointSet, as PM.cc:302-304 calls them, in one kernel launch:
 * same results as sd_interkf_check + sd_update_pointset. */
int sd_interkf_check_points(sd_ctx* ctx, int kf_id, const int* nbr_ids, int n);

/* The loop of ProbabilityMapping::SemiDenseLoop over many key frames (PM.cc:137-257 per key frame, then PM.cc:262-307)
 * issued by one call: sd_semidense for every item in order, and -- with SD_LOOP_INTERKF -- sd_interkf_check_points of an
 * item as soon as the maps of all its neighbours have been queued (neighbours outside the list must already have maps;
 * an item whose neighbours lack them is skipped like PM.cc:289-297 and reported in inter_done).  Downloads (dense or
 * compact, see below) of an item's results are queued right after its last stage when the item asks for them.
 * A skipped item has no point set: its SD_DL_DENSE / SD_DL_COMPACT download still delivers the intra-key-frame maps,
 * SD_DL_COMPACT_POINTS sets *out_count = 0, SD_DL_DENSE_POINTS leaves the buffers untouched -- pass inter_done to tell.
 * With SdParams.interkf_inplace the checks are order-dependent, so they are issued strictly in list order after all the
 * searches (PM.cc:262's loop order) instead of as early as possible. */
enum { SD_LOOP_INTERKF = 1 };
enum { SD_DL_NONE = 0, SD_DL_DENSE = 1, SD_DL_DENSE_POINTS = 2, SD_DL_COMPACT = 3, SD_DL_COMPACT_POINTS = 4 };
typedef struct SdLoopItem {
    int kf_id;
    const int* nbr_ids;       /* n covisible key frames, best first (PM.cc:151-160) */
    const float* nbr_rot_deg; /* median in-plane rotation per neighbour (PM.cc:170-179); NULL = zeros */
    int n;
    float min_invdepth, max_invdepth; /* StereoSearchConstraints (PM.cc:184) */
    int finish_only;          /* 1: the key frame already has its maps (an earlier call queued sd_semidense): only the
                               * inter-key-frame stage and the download are issued for it */
    int download;             /* SD_DL_*: host buffers below, pinned, valid until the next sd_sync */
    float* out_depth;         /* SD_DL_DENSE*: W*H floats each */
    float* out_sigma;
    float* out_points;        /* SD_DL_DENSE_POINTS: H*3W floats */
    void* out_records;        /* SD_DL_COMPACT*: `capacity` SdCompactDepth / SdCompactPoint records */
    int capacity;
    int* out_count;           /* SD_DL_COMPACT*: records available (copy stops at capacity); written by sd_sync at the latest */
} SdLoopItem;
int sd_semidense_loop(sd_ctx* ctx, const SdLoopItem* items, int n_items, int flags, int* inter_done /* nullable, n_items */);

/* Confident semi-dense points of one key frame, compacted on the device: the feed for the CARV transcript
 * (extra 3-D points per key frame enter the Delaunay carving through
 * SFMTranscriptInterface_ORBSLAM::addKeyFrameInsertionWithLinesEntry, src/Modeler/SFMTranscriptInterface_ORBSLAM.cpp:319-374,
 * called from src/Modeler/Modeler.cc:118-124) and for the .obj dump of PM.cc:100-132.
 * Selection is the reference's own (PM.cc:120-121): skip if `sigma > max_sigma` (the float against a double, 0.01 there),
 * keep if `depth > 0.000001`; whole image, raster order; depth = depth_checked when the inter-KF stage ran.
 * If max_points > 0 and more pixels qualify, the max_points with the smallest sigma are kept (ties: first in raster order),
 * still emitted in raster order.  xyz are world points with the arithmetic and the current pose of UpdateSemiDensePointSet
 * (PM.cc:337-367); sd_update_pointset need not have run.  Host outputs hold `capacity` entries each (xyz 3 floats per
 * entry; sigma and pixel = y << 16 | x may be NULL); *count is the number selected (copy stops at capacity).  Synchronises. */
int sd_confident_points(sd_ctx* ctx, int kf_id, double max_sigma, int max_points, int capacity, float* xyz, float* sigma,
                        uint32_t* pixel, int* count);
/* Same, results left in device scratch owned by the context (valid until the next sd_confident_points* call on it);
 * *confident (nullable) = pixels passing the predicate before the max_points cut. */
int sd_confident_points_device(sd_ctx* ctx, int kf_id, double max_sigma, int max_points, const float** xyz_dev,
                               const float** sigma_dev, const uint32_t** pixel_dev, int* count, int* confident);
/* Copy (up to capacity entries of) the result the last sd_confident_points_device call left on the device. */
int sd_confident_points_fetch(sd_ctx* ctx, int capacity, float* xyz, float* sigma, uint32_t* pixel, int* count);

/* Copy results to host buffers (each may be NULL): depth = depth_checked when the inter-KF stage
 * ran, else the intra-KF map; W*H floats each, points H*3W floats.  Synchronises. */
int sd_download(sd_ctx* ctx, int kf_id, float* depth, float* sigma, float* points);
/* Asynchronous variant for pinned host buffers; completes at the next sd_sync. */
int sd_download_async(sd_ctx* ctx, int kf_id, float* depth, float* sigma, float* points);
/* Compact form of the same results: one record per pixel that carries a depth or a sigma (about a quarter of the
 * pixels), in no particular order; every other pixel of depth_map_ / depth_sigma_ / SemiDensePointSets_ is zero.
 * `depth` is depth_checked when the inter-key-frame stage ran, else the intra-key-frame map; x, y, z is the pixel's
 * entry of SemiDensePointSets_ (needs sd_update_pointset / sd_interkf_check_points).  The host expands the records into
 * the dense planes (ProbabilityMapping::FetchResults does); a 720p key frame moves ~3 MB (~6 MB with points) instead of
 * 7.4 MB (18.4 MB).  The async variant completes at the next sd_sync: the record count is only known on the device, so
 * the copy itself is issued by a later sd_* call on this context once the count has arrived.  *count = records available
 * (the copy stops at capacity).  Key frames whose maps came from sd_import_maps cannot be downloaded this way. */
typedef struct SdCompactDepth { uint32_t pixel /* y << 16 | x */; float depth, sigma; } SdCompactDepth;
typedef struct SdCompactPoint { uint32_t pixel; float depth, sigma, x, y, z; } SdCompactPoint;
int sd_download_compact(sd_ctx* ctx, int kf_id, int with_points, int capacity, void* records, int* count);
int sd_download_compact_async(sd_ctx* ctx, int kf_id, int with_points, int capacity, void* records, int* count);
/* Any single plane, for tests: 0 grad, 1 theta, 2 depth (intra), 3 sigma, 4 depth_checked, 5 points. */
int sd_download_plane(sd_ctx* ctx, int kf_id, int which, float* out);
int sd_get_istddev(sd_ctx* ctx, int kf_id, float* out);        /* KeyFrame::I_stddev */
int sd_get_active_count(sd_ctx* ctx, int kf_id, int* out);     /* pixels passing the PM.cc:201 gate */
int sd_device_ptrs(sd_ctx* ctx, int kf_id, SdDevicePlanes* out);
/* Replace the (depth, sigma) maps of a key frame from device buffers (tightly packed W*H floats):
 * how a rank installs a neighbour's maps received over NCCL before sd_interkf_check. */
int sd_import_maps(sd_ctx* ctx, int kf_id, const float* depth_dev, const float* sigma_dev);
/* Pack the (depth, sigma) maps of a key frame into device buffers (tightly packed W*H floats). */
int sd_export_maps(sd_ctx* ctx, int kf_id, float* depth_dev, float* sigma_dev);
/* Same as sd_import_maps from HOST buffers (W*H floats each): how the host class re-installs the maps of a key frame
 * whose device planes it had released (sd_release_keyframe) when its pose changes later (PM.cc:321-334). */
int sd_import_maps_host(sd_ctx* ctx, int kf_id, const float* depth, const float* sigma);

/* ---- multi-GPU: the one exchange step of the path ----
 * Key frames are sharded by index range, one context (one process) per GPU; InterKeyFrameDepthChecking (PM.cc:705-752)
 * reads the (depth, sigma) maps of neighbours owned by other ranks.  sd_halo_exchange sends the maps of this rank's
 * boundary key frames and receives the halo key frames' maps with grouped ncclSend / ncclRecv straight out of / into the
 * key-frame planes on the context's compute stream (no staging copies).  NCCL is loaded with dlopen("libnccl.so.2") at
 * sd_comm_init time; libsdpm.so itself does not link it. */
int sd_nccl_unique_id(void* id128 /* 128 bytes out; rank 0 calls this and distributes the bytes */);
int sd_comm_init(sd_ctx* ctx, int world, int rank, const void* id128);
int sd_comm_destroy(sd_ctx* ctx);
typedef struct SdHaloPeer {
    int peer;            /* rank */
    int n_send;
    const int* send_ids; /* key frames of ours the peer needs, ascending; their maps must exist (sd_semidense) */
    int n_recv;
    const int* recv_ids; /* key frames of the peer we need, ascending; must be resident here (sd_upload_keyframe) */
} SdHaloPeer;
int sd_halo_exchange(sd_ctx* ctx, const SdHaloPeer* peers, int n_peers);

/* Page-locked host memory for the asynchronous downloads above (SdLoopItem::out_records, sd_download_*_async): what a
 * host without its own CUDA runtime calls uses -- host/ProbabilityMapping.cc sees the GPU through this header only.  (No
 * reference counterpart.)  Allocate at start-up and reuse: like any page-locking call these are slow. */
int sd_alloc_host(sd_ctx* ctx, uint64_t bytes, void** out);
int sd_free_host(sd_ctx* ctx, void* ptr);

/* Wait for downloads without waiting for the rest of the GPU's queue (sd_sync waits for everything).  sd_download_fence
 * returns a ticket that stands for every asynchronous download queued on the context so far (SdLoopItem downloads,
 * sd_download_async, sd_download_compact_async); sd_download_wait(ticket) returns once those have landed in their host
 * buffers and their counts are written, while later work -- the next window of key frames -- keeps running.  This is how
 * ProbabilityMapping::SemiDenseLoop keeps two windows in flight.  (No reference counterpart.) */
int sd_download_fence(sd_ctx* ctx, uint64_t* ticket);
int sd_download_wait(sd_ctx* ctx, uint64_t ticket);

int sd_sync(sd_ctx* ctx);
const char* sd_last_error(sd_ctx* ctx);
/* cudaStream_t of the compute stream, for callers that order their own device work after it */
int sd_compute_stream(sd_ctx* ctx, void** stream);
/* number of kernels this context has launched, and device time of the last timed span */
int sd_launch_count(sd_ctx* ctx, uint64_t* out);
/* Device-time span over ALL of the context's streams: begin = wait until idle, record an event;
 * end = join the streams, record an event, wait, return the elapsed milliseconds (CUDA events). */
int sd_span_begin(sd_ctx* ctx);
int sd_span_end(sd_ctx* ctx, double* ms);
/* Per-kernel device timing (CUDA events on the launching stream). enable=1 starts collecting; enable=2 also drains every
 * stream before each timed kernel, so that the durations are those of the kernel running alone (diagnostic: it
 * serialises the pipeline). */
int sd_timing_enable(sd_ctx* ctx, int enable);
/* names: "grad", "search", "fuse", "intra" (checking + growing, one kernel), "inter", "points"; returns accumulated ms and launches */
int sd_timing_get(sd_ctx* ctx, const char* name, double* ms, uint64_t* launches);
int sd_timing_reset(sd_ctx* ctx);

/* ---- host-side helpers that stay on the CPU in the reference too ---- */
/* StereoSearchConstraints, PM.cc:370-383, on the vector GetAllPointDepths() would return. */
int sd_stereo_search_constraints(const float* orb_depths, int n, float* min_invdepth, float* max_invdepth);
/* GetRotInPlane + median, PM.cc:467-484 and 172-177.  Map points by id, id < 0 = none. */
float sd_rot_in_plane_median(const long long* mp1, const float* ang1, int n1, const long long* mp2,
                             const float* ang2, int n2);
/* ComputeFundamental PM.cc:972-987 and the R21/t21 of PM.cc:890-891, OpenCV-exact arithmetic. */
int sd_pair_geometry(const float K1[4], const float Tcw1[12], const float K2[4], const float Tcw2[12], float F12[9],
                     float R21[9], float t21[3]);

/* GetSearchRange, PM.cc:877-910: the column interval [umin, umax] of key frame 2 (clamped to
 * [0, cols] inclusive, as the reference does) that the inverse-depth interval maps to for pixel (px, py). */
int sd_search_range(const float K1[4], const float Tcw1[12], const float Tcw2[12], int cols, int px, int py, float mind,
                    float maxd, float* umin, float* umax);

/* IntraKeyFrameDepthChecking (PM.cc:486-547) / IntraKeyFrameDepthGrowing (PM.cc:549-596) on caller-owned
 * HOST planes of w*h floats each, updated in place like the reference's cv::Mat& arguments: the
 * planes are copied to the device, one stencil kernel runs (the same one sd_semidense launches),
 * and the result is copied back.  Synchronous. */
int sd_intra_depth_check_planes(sd_ctx* ctx, float* depth, float* sigma, const float* grad, int w, int h);
int sd_intra_depth_grow_planes(sd_ctx* ctx, float* depth, float* sigma, const float* grad, int w, int h);

/* sd_intra_depth_check_planes for callers that must not allocate device memory.  (No counterpart in the reference: it has
 * no GPU code, so the question never arises there.)
 * Who needs it: sd_intra_depth_check_planes obtains five working planes with cudaMalloc on every call and returns them
 * with cudaFree.  Both go to the driver's allocator, take an unpredictable time, and cudaFree in addition waits for all
 * work on the GPU -- including a key-frame pipeline this or another context has in flight.  A tracking or local-mapping
 * thread with a frame deadline cannot afford that stall, and a process that handed all of its device memory to its own
 * pool at start-up (or runs under a hard memory quota) may simply get cudaErrorMemoryAllocation.  This variant never calls
 * cudaMalloc, cudaFree or any other allocating routine: the five working planes live in an array that is part of
 * libsdpm.so's device image, 5 * SD_STATIC_MAX_PIXELS floats (17.6 MiB) on every GPU that loads the library, until the
 * process exits.
 * Same arguments, same arithmetic and the same results, bit for bit, as sd_intra_depth_check_planes; w * h above
 * SD_STATIC_MAX_PIXELS is refused with SD_ERR_INVALID_ARG (sd_last_error names the limit).  Synchronous.
 * First call on each GPU: the library looks up the array's device address (cudaGetSymbolAddress, with the context's
 * device current) and remembers it; this makes the driver load the module, so make that first call during start-up,
 * before any CUDA stream capture begins.
 * Sharing: all calls on one GPU, from every context in the process, use the same array, and it keeps what the last call
 * left in it.  Nothing in it has to be cleared between calls (each call overwrites the input planes and the kernel writes
 * every output pixel), but two calls that are on the GPU at the same time overwrite each other's planes: the caller
 * serialises them per GPU. */
#define SD_STATIC_MAX_PIXELS (1280 * 720)
int sd_intra_depth_check_planes_static(sd_ctx* ctx, float* depth, float* sigma, const float* grad, int w, int h);

/* ---- single-pixel probes (run the same device code as the batch kernels; for unit tests) ---- */
/* EpipolarSearch PM.cc:385-465 for n pixels xy[2*i], xy[2*i+1] of kf1 against kf2.
 * out[6*i..] = {inverse depth, sigma, best integer uj (or -1), supported, best_u (the refined u*, PM.cc:459),
 *               best_v (PM.cc:460)} */
int sd_probe_epipolar(sd_ctx* ctx, int kf1, int kf2, const int* xy, int n, float min_invdepth, float max_invdepth,
                      float rot_deg, float* out);
/* InverseDepthHypothesisFusion PM.cc:598-626 on one hypothesis list (n <= SD_MAX_NEIGHBOURS). */
int sd_probe_fusion(sd_ctx* ctx, const float* depth, const float* sigma, int n, float* out_depth, float* out_sigma,
                    int* supported);

#ifdef __cplusplus
}
#endif
#endif /* SDPM_H */
