/* orbx.h — C ABI of the B200-native ORB frontend (sm_100a CUDA kernels behind plain pointers).
 *
 * This is the drop-in boundary for the ONE hot path of ORB-SLAM3 that the ROS 2 wrapper drives:
 * ORB_SLAM3::ORBextractor::operator() and brute-force / banded top-2 Hamming matching with
 * ORBmatcher::DescriptorDistance semantics.  Citations are file:line under the reference tree
 * (ORB_SLAM3/...).  The C++ adapter `compat/ORBextractor.h` forwards the reference's class surface
 * (ORB_SLAM3/include/ORBextractor.h:43-109) to these entry points; INTEGRATION.md shows the
 * reference-side binding.
 *
 * Conventions: every function returns ORBX_OK (0) or a negative orbx_status; no exceptions cross
 * the boundary; orbx_last_error() returns a thread-local message for the last failure.  There is
 * NO CPU fallback: without a CUDA device every compute entry point fails with ORBX_ERR_CUDA.
 * Pointers named d_* are device pointers, all others are host pointers.
 */
#ifndef ORBX_H_
#define ORBX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum orbx_status {
    ORBX_OK = 0,
    ORBX_ERR_ARG = -1,      /* bad argument (null pointer, non-positive size, image too small/large) */
    ORBX_ERR_CUDA = -2,     /* CUDA runtime/driver failure, or no device */
    ORBX_ERR_CAPACITY = -3, /* caller's output buffer too small; *n_out holds the needed count */
    ORBX_ERR_NCCL = -4,     /* NCCL failure */
    ORBX_ERR_STATE = -5     /* call order violated (e.g. results requested before an extract) */
} orbx_status;

/* Same 28-byte layout as cv::KeyPoint (pt.x, pt.y, size, angle, response, octave, class_id), so a
 * std::vector<cv::KeyPoint>::data() can be passed directly (ORBextractor.cc:1112). */
typedef struct orbx_keypoint {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orbx_keypoint;

/* The five constructor arguments of ORB_SLAM3::ORBextractor (ORBextractor.cc:409-411, fed from
 * YAML by Settings.cc:443-450 / Tracking.cc:1223-1289) plus device-side knobs that never reach the
 * reference class surface. */
typedef struct orbx_params {
    int nfeatures;
    float scale_factor;
    int nlevels;
    int ini_th_fast;
    int min_th_fast;
    int device;  /* CUDA device ordinal; -1 = current device */
    int fp_mode; /* 0 = STRICT (no FMA contraction in computeOrbDescriptor's rotation, -ffp-contract=off build
                    of the reference); 1 = NATIVE (the reference's -O3 -march=native build: fmaf(x,b,y*a)) */
} orbx_params;

typedef struct orbx_extractor* orbx_handle;

const char* orbx_last_error(void);
int orbx_version(void);
/* number of kernel launches issued by this library in the calling process so far (bench bookkeeping) */
unsigned long long orbx_launch_count(void);

/* ---- extractor life cycle: ORBextractor::ORBextractor (ORBextractor.cc:409-469) -------------------- */
int orbx_create(const orbx_params* params, orbx_handle* out);
void orbx_destroy(orbx_handle h);
/* Debugging aid (the reference has no sanitizer set-up; compute-sanitizer is closed on the GPU pool this was
 * built on): with ORBX_GUARD=1 in the environment every device buffer of a handle carries a 256-byte guard band; this call
 * synchronises the device and returns ORBX_ERR_CAPACITY naming the buffer if a kernel wrote behind one, ORBX_OK otherwise. */
int orbx_debug_check_guards(orbx_handle h);

/* Per-level tables the reference exposes through GetScaleFactors(), GetInverseScaleFactors(),
 * GetScaleSigmaSquares(), GetInverseScaleSigmaSquares() (ORBextractor.h:61-83) and mnFeaturesPerLevel
 * (ORBextractor.cc:434-445).  Each output may be NULL; arrays hold nlevels entries. */
int orbx_get_tables(orbx_handle h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                    int* features_per_level);

/* Upper bound of the keypoints operator() returns for a w x hgt frame = the room kps / desc / cap_per_frame must offer:
 * the sum over the levels of max(N_l + 4, 4 * nIni_l + 1).  DistributeOctTree (ORBextractor.cc:555-779) stops splitting at
 * the first node count >= N_l (each split adds at most 3), but its first breadth pass splits every one of the
 * nIni_l = round(width / height) root nodes whatever N_l is (:610-681), so small nfeatures return more than N_l + 3. */
int orbx_capacity(orbx_handle h, int w, int hgt, int* cap);

/* ---- single frame: ORBextractor::operator() (ORBextractor.cc:1086-1168) ---------------------------
 * img: host, 8-bit single channel, `stride` bytes per row.  lap0/lap1 = vLappingArea[0..1].
 * kps/desc: host outputs with room for `cap` keypoints (desc is cap x 32 bytes).  *n_out = number of
 * keypoints, *mono_index = the value operator() returns.  An empty image (img NULL or w/h <= 0) is not
 * an error: *mono_index = -1, *n_out = 0 (ORBextractor.cc:1090-1091). */
int orbx_extract(orbx_handle h, const uint8_t* img, int w, int hgt, int stride, int lap0, int lap1,
                 orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out, int* mono_index);

/* The drop-in path of one frame with ONE host synchronisation (what compat/ORBextractor.cc calls): the frame is copied into
 * a page-locked staging buffer owned by the handle, and the keypoints (28-byte cv::KeyPoint records), the descriptors, the
 * counts and - if want_pyramid - the frame's whole padded pyramid block come back into the same staging buffer behind one
 * cudaStreamSynchronize.  The returned pointers stay valid until the next call on this handle.  out->pyramid + level[l].offset
 * is the padded plane of level l: (h + 38) rows of `pitch` bytes, (w + 38) of them meaningful; mvImagePyramid[l] is the ROI at
 * (19, 19) (ORBextractor.cc:1176-1178).  An empty image is not an error: mono_index = -1, n = 0 (:1090-1091). */
typedef struct orbx_frame_level {
    int w, h;            /* level size (ORBextractor.cc:1175) */
    int pitch;           /* bytes per row of the padded plane */
    unsigned offset;     /* byte offset of the padded plane inside `pyramid` */
} orbx_frame_level;
typedef struct orbx_frame_result {
    const orbx_keypoint* kps;  /* n records, operator()'s output order */
    const uint8_t* desc;       /* n x 32 bytes */
    int n;
    int mono_index;
    const uint8_t* pyramid;    /* NULL unless want_pyramid */
    int nlevels;
    orbx_frame_level level[16];
} orbx_frame_result;
int orbx_extract_frame(orbx_handle h, const uint8_t* img, int w, int hgt, int stride, int lap0, int lap1, int want_pyramid,
                       orbx_frame_result* out);

/* Input staging fused ahead of the pyramid: System::TrackMonocular's cv::resize(im, resized, newImSize) (System.cc:420-425)
 * followed by Tracking::GrabImageMonocular's cvtColor(... -> GRAY) (Tracking.cc:1569-1582), in the reference's order
 * (resize on the interleaved colour image, then grey), then operator().  img: host, `channels` (1, 3 or 4) interleaved
 * bytes per pixel, `stride` bytes per row; rgb = the reference's mbRGB (1: RGB/RGBA order, 0: BGR/BGRA);
 * new_w/new_h = Settings::newImSize() or 0/0 for no resize.  The other arguments are orbx_extract's. */
int orbx_extract_color(orbx_handle h, const uint8_t* img, int w, int hgt, int stride, int channels, int rgb, int new_w,
                       int new_h, int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out,
                       int* mono_index);
/* The grey frame the last orbx_extract_color handed to the extractor (w x h compact; `out` may be NULL to query sizes). */
int orbx_get_staged_gray(orbx_handle h, uint8_t* out, int* w, int* hgt);

/* mvImagePyramid[level] of the last orbx_extract / batch frame `frame` (ORBextractor.cc:1170-1195): the
 * level is a ROI at (19,19) of a padded (w+38) x (h+38) plane.  `out` (host, may be NULL to query the
 * size) receives the padded plane compactly, (h+38) rows of (w+38) bytes. */
int orbx_get_pyramid(orbx_handle h, int frame, int level, uint8_t* out, int* w, int* hgt);

/* ---- batch of equally sized frames (the throughput path; frames are independent) ------------------
 * orbx_extract_batch: host frames (frame f at imgs + f*frame_stride, rows `stride` bytes apart) are
 * copied to the device, extracted, and the results copied back: kps (nframes x cap_per_frame),
 * desc (nframes x cap_per_frame x 32), n_out[nframes], mono_index[nframes].  Any of the four outputs may
 * be NULL to leave that result on the device only.
 * orbx_extract_batch_device: same with frames already resident on the device; results stay there and
 * are read with orbx_batch_results / orbx_batch_download. */
int orbx_extract_batch(orbx_handle h, const uint8_t* imgs, int nframes, int w, int hgt, int stride,
                       size_t frame_stride, int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc,
                       int cap_per_frame, int* n_out, int* mono_index);
int orbx_extract_batch_device(orbx_handle h, const uint8_t* d_imgs, int nframes, int w, int hgt, int stride,
                              size_t frame_stride, int lap0, int lap1);
/* Asynchronous form of orbx_extract_batch: the H2D copy of the frames, the kernels and the D2H copies of the results
 * are only ENQUEUED on the handle's stream (host buffers should be page-locked, and must stay untouched until
 * orbx_sync / orbx_check returns).  cap_per_frame must be at least the extractor's capacity (sum over levels of
 * N_l + 4).  Two handles driven this way overlap one handle's copies with the other's kernels, the way the reference
 * runs its left / right extractors on two threads (Frame.cc:122-125).  orbx_check = orbx_sync + the device-side
 * invariant flags (ORBX_ERR_CAPACITY if one tripped). */
int orbx_extract_batch_async(orbx_handle h, const uint8_t* imgs, int nframes, int w, int hgt, int stride,
                             size_t frame_stride, int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc,
                             int cap_per_frame, int* n_out, int* mono_index);
int orbx_check(orbx_handle h);
/* Device-resident results of the last batch: keypoints/descriptors of frame f start at
 * d_kps + f*cap, d_desc + f*cap*32; d_n[f] keypoints are valid. */
int orbx_batch_results(orbx_handle h, const orbx_keypoint** d_kps, const uint8_t** d_desc, const int** d_n,
                       const int** d_mono, int* cap_per_frame);
int orbx_batch_download(orbx_handle h, int frame, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out,
                        int* mono_index);
/* Keypoint counts (and monoIndex values) of ALL frames of the last batch in one copy; either pointer may be NULL. */
int orbx_batch_counts(orbx_handle h, int32_t* n_out, int32_t* mono_out);
/* Batch scheduling knob (default 1): with n > 1 a batch of at least 8n frames is cut into n contiguous parts that run
 * the whole pipeline concurrently on n internal stream pairs (frames are independent), forked from and joined back into
 * the handle's stream, so one part's latency-bound stages (quadtree, describe) share the SMs with another part's
 * issue-bound ones.  Results are identical; orbx_last_timing then reports only the total. */
int orbx_set_subbatches(orbx_handle h, int n);
/* Blocks until the handle's stream is idle (batch_device is asynchronous). */
int orbx_sync(orbx_handle h);
/* The handle's cudaStream_t (for callers that record their own events around a run of batch calls). */
void* orbx_get_stream(orbx_handle h);
/* Elapsed device time in milliseconds of the last batch (CUDA events on the handle's stream); the
 * per-stage array (pyramid, fast, quadtree, blur, describe) may be NULL. */
int orbx_last_timing(orbx_handle h, float* total_ms, float stage_ms[5]);
/* Debug/parity taps of the last extraction: FAST candidates of (frame, level) in the reference's
 * order (cell row-major, then FAST's row-major inside the cell; ORBextractor.cc:863-868) as
 * (x, y, response) int32 triplets in cell-offset coordinates; returns the count through *n_out. */
int orbx_get_candidates(orbx_handle h, int frame, int level, int32_t* out, int cap, int* n_out);
/* Blurred level (GaussianBlur 7x7 sigma 2 of the un-padded level, ORBextractor.cc:1132-1133), w x h compact. */
int orbx_get_blurred(orbx_handle h, int frame, int level, uint8_t* out, int* w, int* hgt);

/* Fills d_out (device, nframes x hgt x w, compact) with frames k0..k0+nframes-1 of the deterministic
 * synthetic stream (bit-identical to orb_slam3_ros2_b200/synth.py). */
int orbx_synth_frames_device(uint8_t* d_out, int k0, int nframes, int w, int hgt, uint64_t seed, int right,
                             int device);

/* ---- Hamming matching ----------------------------------------------------------------------------
 * ORBmatcher::DescriptorDistance (ORBmatcher.cc:2058-2074): 256-bit Hamming distance (host helper). */
int orbx_hamming(const uint8_t* a, const uint8_t* b);

/* Brute-force k=2: for every query the two database rows with the smallest distance, ordered by
 * (distance asc, index asc) == cv::BFMatcher(NORM_HAMMING).knnMatch(q, db, m, 2) (Frame.cc:1144) and the
 * reference's strict-< best/second loops visiting candidates in ascending index (ORBmatcher.cc:83-120).
 * Missing entries (nd < 2) get idx = -1, dist = 256.  idx outputs are int64 row numbers. */
int orbx_match_top2(const uint8_t* q, int nq, const uint8_t* db, int64_t nd, int64_t* idx1, int32_t* d1,
                    int64_t* idx2, int32_t* d2, int device);
int orbx_match_top2_device(const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t nd, int64_t db_index_base,
                           int64_t* d_idx1, int32_t* d_d1, int64_t* d_idx2, int32_t* d_d2, void* stream);
/* The matching entry points that take no handle (orbx_match_top2, _device, orbx_match_stereo_banded, orbx_accept_top2,
 * orbx_search_*, orbx_compute_stereo_matches, ...) keep their device scratch and one CUDA stream PER CALLING THREAD.  Consequences:
 * calls of one thread are serialised through that scratch - orbx_match_top2_device may only be given streams whose work is ordered
 * with the thread's previous call (the same stream, or after a synchronisation) -, and a thread that is about to exit should call
 * orbx_thread_cleanup() to release the scratch (it is not released automatically). */
void orbx_thread_cleanup(void);

/* orbx_match_top2_device for callers that may not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.  Same arguments, same
 * kernels and bit-identical results as orbx_match_top2_device; it calls no cudaMalloc / cudaFree / allocating routine and creates
 * no stream - the per-chunk partial keys go to a file-scope device array of 2 x ORBX_NOALLOC_MAX_KEYS x 8 B = 16 MiB that every
 * GPU loading the library holds until exit.  Limit: nq x (database chunks) <= ORBX_NOALLOC_MAX_KEYS, else ORBX_ERR_CAPACITY with a
 * message naming the sizes (on a 148-SM B200 every nq <= 32768 with nd <= 100 M rows fits: the chunk count is about 592 / ceil(nq /
 * 1024)).  All calls on one GPU share the array, from every thread and stream: calls that overlap on the GPU race - order them on
 * one stream or with events.  The array keeps the last call's keys; nothing in it has to start cleared (every slot read is
 * written first by the same call).  The work is only enqueued on `stream` (a cudaStream_t, may be 0); the current device is used. */
#define ORBX_NOALLOC_MAX_KEYS (1u << 20)
int orbx_match_top2_device_noalloc(const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t nd, int64_t db_index_base,
                                   int64_t* d_idx1, int32_t* d_d1, int64_t* d_idx2, int32_t* d_d2, void* stream);

/* Frame-to-frame top-2 over the last batch held by `h` (config 2): frame f's descriptors against frame
 * f+1's, f = 0..nframes-2.  Outputs are (nframes-1) x cap entries; to_host = 0: device pointers, results left on the
 * stream; 1: host pointers, synchronous; 2: host pointers, copies only enqueued (see orbx_extract_batch_async). */
int orbx_match_batch_consecutive(orbx_handle h, int32_t* idx1, int32_t* d1, int32_t* idx2, int32_t* d2,
                                 int to_host);

/* Frame::ComputeStereoMatches' Hamming stage (Frame.cc:811-894): for each left keypoint the best right
 * keypoint among those whose row band [floor(vR - 2*sf[octR]), ceil(vR + 2*sf[octR])] contains (int)vL, with
 * octave within +-1 and uR in [uL - max_d, uL] (minD = 0, Frame.cc:841); initial best = TH_HIGH (100), strict <,
 * right keypoints in ascending index.  best_idx = 0 / best_dist = 100 when nothing beats TH_HIGH (:867-868). */
int orbx_match_stereo_banded(const orbx_keypoint* kps_l, const uint8_t* desc_l, int nl, const orbx_keypoint* kps_r,
                             const uint8_t* desc_r, int nr, const float* scale_factors, int nlevels, int n_rows,
                             float max_d, int32_t* best_idx, int32_t* best_dist, int device);

/* ---- accept rules on top-2 results (ORBmatcher constants TH_LOW = 50, TH_HIGH = 100: ORBmatcher.cc:35-36) ------
 * One flag per query from (d1, d2), the float comparisons written exactly as the reference writes them.  A missing
 * second candidate is d2 = 256, the reference's initial bestDist2 (ORBmatcher.cc:83, :271, :669). */
typedef enum orbx_accept_rule {
    ORBX_ACCEPT_PROJ_MAP = 0,   /* SearchByProjection(F, map points), ORBmatcher.cc:123-128: d1 <= TH_HIGH and
                                   (level1 != level2 or (float)d1 <= nnratio * d2); levels NULL = always equal */
    ORBX_ACCEPT_BOW_KF_F = 1,   /* SearchByBoW(KF, F), :327-329: d1 <= TH_LOW and (float)d1 < nnratio * (float)d2 */
    ORBX_ACCEPT_INIT = 2,       /* SearchForInitialization, :702-705: d1 <= TH_LOW and d1 < (float)d2 * nnratio */
    ORBX_ACCEPT_BOW_KF_KF = 3,  /* SearchByBoW(KF, KF), :848-850: d1 < TH_LOW (strict) and (float)d1 < nnratio * (float)d2 */
    ORBX_ACCEPT_BEST_HIGH = 4,  /* SearchByProjection(frame, frame) :1770, SearchBySim3 :1569, :1649: d1 <= TH_HIGH, best only */
    ORBX_ACCEPT_BEST_LOW = 5,   /* Fuse / SearchBySim3 style, :1312, :1437: d1 <= TH_LOW, best only */
    ORBX_ACCEPT_KNN_RATIO = 6,  /* fisheye stereo on knnMatch results, Frame.cc:1151: two results and
                                   (double)d1 < (double)d2 * 0.7 (the literal is a double; nnratio is ignored) */
    ORBX_ACCEPT_LOW_RATIO = 7   /* SearchByProjection(KF, Sim3, points, ..., th, ratioHamming), :523, :636:
                                   (float)d1 <= TH_LOW * ratioHamming, best only; nnratio carries ratioHamming */
} orbx_accept_rule;
/* d1, d2: n distances; idx2: second indices (only ORBX_ACCEPT_KNN_RATIO reads it: < 0 = fewer than two results; may be
 * NULL = always two); level1/level2: octaves of the best / second candidate (PROJ_MAP only, may be NULL).
 * accept: n bytes, 1 = accepted.  Host pointers; the _device form takes device pointers and a cudaStream_t. */
int orbx_accept_top2(int rule, float nnratio, const int32_t* d1, const int32_t* d2, const int64_t* idx2,
                     const int32_t* level1, const int32_t* level2, int n, uint8_t* accept, int device);
int orbx_accept_top2_device(int rule, float nnratio, const int32_t* d_d1, const int32_t* d_d2, const int64_t* d_idx2,
                            const int32_t* d_level1, const int32_t* d_level2, int n, uint8_t* d_accept, void* stream);

/* ---- ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono), whole function ------------------
 * (ORBmatcher.cc:1676-1887; Tracking::TrackWithMotionModel, Tracking.cc:2876, :2906, :2914), frames with Nleft == -1 (monocular,
 * rectified stereo, RGB-D).  `pairs` independent (last, current) frame pairs in one call; arrays are [pair][cap_last] /
 * [pair][cap_cur], host pointers.
 *   kps_last   LastFrame.mvKeysUn (octave and angle are read)          last_flags  bit 0: mvpMapPoints[i] != NULL and !mvbOutlier[i],
 *   world      MapPoint::GetWorldPos(), 3 floats per feature                        bit 1: that map point's Observations() > 0
 *   mp_desc    MapPoint::GetDescriptor(), 32 bytes per feature         n_last / n_cur  features per pair
 *   kps_cur    CurrentFrame.mvKeysUn, desc_cur = mDescriptors          cur_u_right mvuRight or NULL (monocular)
 *   cur_taken  != 0: CurrentFrame.mvpMapPoints[i2] already holds a map point with observations (NULL = none; TrackWithMotionModel
 *              clears the vector first, Tracking.cc:2866)
 *   poses      14 floats per pair: Tcw then Tlw, each (qx, qy, qz, qw, tx, ty, tz) - Sophus::SE3f's own storage
 * Results: match_cur[pair][i2] = index of the LAST-frame feature whose map point CurrentFrame.mvpMapPoints[i2] holds after the
 * call (-1 = none); nmatches[pair] = the function's return value (it counts overwritten matches too, as the reference does).
 * Float arithmetic: the operations of Sophus' SE3f * point (so3.hpp:358-368, se3.hpp:321-325) and Pinhole::project
 * (CameraModels/Pinhole.cpp:43-49), each rounded once, in source order (the reference built with -ffp-contract=off). */
typedef struct orbx_proj_params {
    float fx, fy, cx, cy;                  /* CurrentFrame.mpCamera (pinhole) */
    float min_x, max_x, min_y, max_y;      /* Frame::mnMinX .. mnMaxY */
    float scale_factors[16];               /* CurrentFrame.mvScaleFactors */
    int nlevels;
    float mbf, mb;                         /* CurrentFrame.mbf, mb */
    float th;
    int mono;                              /* bMono */
    int check_orientation;                 /* ORBmatcher::mbCheckOrientation */
} orbx_proj_params;
int orbx_search_by_projection(const orbx_keypoint* kps_last, const uint8_t* last_flags, const float* world, const uint8_t* mp_desc,
                              const int32_t* n_last, const orbx_keypoint* kps_cur, const uint8_t* desc_cur, const float* cur_u_right,
                              const uint8_t* cur_taken, const int32_t* n_cur, const float* poses, int pairs, int cap_last, int cap_cur,
                              const orbx_proj_params* params, int32_t* match_cur, int32_t* nmatches, int device);

/* ---- ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, th, bFarPoints, thFarPoints), whole function ----
 * (ORBmatcher.cc:43-210 + RadiusByViewingCos :215-221; Tracking::SearchLocalPoints, Tracking.cc:3410, :3433), frames with Nleft == -1.
 * The map points arrive projected, as Frame::isInFrustum leaves them: track = 5 floats per point (mTrackProjX, mTrackProjY,
 * mTrackProjXR, mTrackViewCos, mTrackDepth), level = mnTrackScaleLevel, flags bit 0 = mbTrackInView, bit 1 = Observations() > 0,
 * bit 2 = isBad(), mp_desc = GetDescriptor().  The frame side, `pairs`, the [pair][cap] layout, cur_taken and the results are those
 * of orbx_search_by_projection; of `params` the image bounds, scale factors and th are read; nnratio = ORBmatcher::mfNNratio (0.8 in
 * SearchLocalPoints).  match_cur[pair][i] = index of the map point F.mvpMapPoints[i] holds after the call, -1 = none. */
int orbx_search_by_projection_map(const float* track, const int32_t* level, const uint8_t* flags, const uint8_t* mp_desc, const int32_t* n_mp,
                                  const orbx_keypoint* kps_cur, const uint8_t* desc_cur, const float* cur_u_right, const uint8_t* cur_taken,
                                  const int32_t* n_cur, int pairs, int cap_mp, int cap_cur, const orbx_proj_params* params, float nnratio,
                                  int far_points, float th_far_points, int32_t* match_cur, int32_t* nmatches, int device);

/* ---- guided search: Frame::AssignFeaturesToGrid + Frame::GetFeaturesInArea + the best/second loop ---------------
 * The inner loop of ORBmatcher::SearchByProjection (frame to frame, ORBmatcher.cc:1676-1887), SearchForInitialization
 * (:648-763) and SearchByProjection(F, map points) (:45-138): for every query (u, v, radius, [min_level, max_level],
 * descriptor) the candidates are the target keypoints GetFeaturesInArea returns (Frame.cc:657-723: grid cells
 * ix outer / iy inner, insertion order inside a cell, octave filter, |dx| < r and |dy| < r), visited in that order with
 * strict-< best/second updates.  The grid is the reference's 64 x 48 (Frame.h:44-45) over [min_x, max_x) x
 * [min_y, max_y) (Frame::ComputeImageBounds), cell of a keypoint = round((x - min) * inv) (Frame::PosInGrid,
 * Frame.cc:725-736).  Outputs per query: idx1/idx2 = target keypoint indices (-1 = none), d1/d2 (256 = none),
 * lvl1/lvl2 = octaves of the two (may be NULL).  max_level < 0 = no upper bound, as in the reference.
 * Two host-state filters of the reference loops are optional inputs: skip_t[i] != 0 drops target i (a map point with
 * observations is already attached, ORBmatcher.cc:87-89, :1746-1748), and with u_right_t != NULL a target whose
 * u_right_t[i] > 0 is dropped when fabs(query.ur - u_right_t[i]) > query.r (:91-96, :1750-1756). */
typedef struct orbx_grid_query {
    float u, v, r;
    int32_t min_level, max_level;
    float ur; /* projected right-image coordinate; only read when u_right_t != NULL */
} orbx_grid_query;
int orbx_search_grid_top2(const orbx_keypoint* kps_t, const uint8_t* desc_t, int nt, float min_x, float max_x, float min_y,
                          float max_y, const uint8_t* skip_t, const float* u_right_t, const orbx_grid_query* queries,
                          const uint8_t* desc_q, int nq, int32_t* idx1, int32_t* d1, int32_t* idx2, int32_t* d2,
                          int32_t* lvl1, int32_t* lvl2, int device);

/* ---- Frame::ComputeStereoMatches, whole function (Frame.cc:811-980) -------------------------------------------
 * After the banded Hamming stage (orbx_match_stereo_banded) the reference refines every match with best < 75
 * ((TH_HIGH + TH_LOW) / 2) by an 11 x 11 L1 patch correlation over +-5 px on pyramid level kpL.octave of the two
 * extractors (:896-935), fits a parabola (:937-949), converts to disparity / depth (:951-966) and finally removes
 * matches whose correlation distance is >= 1.5 * 1.4 * median (:969-980).  `left` / `right` are the two extractor
 * handles whose LAST single-frame extraction produced kps_l / kps_r (their device-resident pyramids are read in place:
 * no mvImagePyramid download).  bf = mbf, b = mb.  u_right / depth: nl floats, -1 where unmatched (:813-814). */
int orbx_compute_stereo_matches(orbx_handle left, orbx_handle right, int frame, const orbx_keypoint* kps_l,
                                const uint8_t* desc_l, int nl, const orbx_keypoint* kps_r, const uint8_t* desc_r, int nr,
                                float bf, float b, float* u_right, float* depth);
/* The same for EVERY pair of the two handles' last batches (pair f = frame f of `left` with frame f of `right`), keypoints,
 * descriptors and pyramids all read in place on the device: config 3 of BASELINE.json.  u_right / depth: host,
 * [nframes][cap] floats with cap = cap_per_frame of orbx_batch_results; entries past a frame's keypoint count are unspecified. */
int orbx_compute_stereo_matches_batch(orbx_handle left, orbx_handle right, float bf, float b, float* u_right, float* depth);

/* ---- Frame::UndistortKeyPoints / Frame::ComputeImageBounds (row N1, Frame.cc:747-809), pinhole camera -----------------
 * Replaces cv::undistortPoints(mat, mat, Pinhole::toK(), mDistCoef, cv::Mat(), mK) (Frame.cc:766, :795): K and P are the two
 * 3 x 3 float matrices row-major, dist = mDistCoef (4, 5, 8 or 12 floats: k1 k2 p1 p2 [k3 [k4 k5 k6 [s1..s4]]]).  Arithmetic
 * in doubles, 5 fixed iterations, results rounded to float: equal to OpenCV's scalar path bit for bit.  With dist[0] == 0 the
 * keypoints are copied unchanged (:749-753) and the bounds are (0, cols, 0, rows) (:804-809).
 * orbx_undistort_keypoints: host arrays (kps_un may alias kps).  orbx_image_bounds: bounds = (mnMinX, mnMaxX, mnMinY,
 * mnMaxY), what orbx_search_grid_top2 takes.  orbx_undistort_batch: mvKeysUn of EVERY frame of the handle's last batch, read
 * and written on the device ([frame][cap_per_frame] as orbx_batch_results; entries past a frame's count are unspecified);
 * kps_un (host, may be NULL) receives a copy, *d_kps_un (may be NULL) the device array, valid until the next call. */
int orbx_undistort_keypoints(const orbx_keypoint* kps, int n, const float* K, const float* dist, int ndist, const float* P,
                             orbx_keypoint* kps_un, int device);
int orbx_image_bounds(int cols, int rows, const float* K, const float* dist, int ndist, const float* P, float bounds[4], int device);
int orbx_undistort_batch(orbx_handle h, const float* K, const float* dist, int ndist, const float* P, orbx_keypoint* kps_un,
                         const orbx_keypoint** d_kps_un);

/* ---- ORBmatcher::SearchForInitialization (ORBmatcher.cc:648-763), whole function, monocular frames ---------------------
 * F1 = (kps1 = mvKeysUn, desc1), F2 = (kps2, desc2) with its 64 x 48 grid over [min_x, max_x) x [min_y, max_y)
 * (Frame::AssignFeaturesToGrid).  For every level-0 keypoint i1 in index order: candidates = F2.GetFeaturesInArea(
 * vbPrevMatched[i1], windowSize, 0, 0), best / second Hamming distance over the candidates NOT already held by an earlier
 * i1 at a smaller or equal distance (vMatchedDistance, :687-688), accept best <= TH_LOW && best < second * nnratio, take the
 * point over from an earlier worse match (:701-709), rotation histogram + ComputeThreeMaxima (:2012-2053) when
 * check_orientation.  prev_matched: n1 x 2 floats = vbPrevMatched, updated for matched keypoints (:756-759).
 * matches12: n1 ints = vnMatches12 (-1 = none).  *nmatches = the return value.  The window queries and distances (order
 * independent) run one warp per keypoint; the i1 loop itself is order dependent and runs in order on one warp per frame pair.
 * _batch: pair p = frame p -> frame p + 1 of the handle's last batch, read in place on the device (use_undistorted != 0: the
 * mvKeysUn of orbx_undistort_batch), vbPrevMatched = frame p's own keypoints as Tracking::MonocularInitialization sets it
 * (Tracking.cc:2479-2481, called at :2512).  matches12: [frames-1][cap_per_frame], nmatches: [frames-1], prev_matched (may be NULL):
 * [frames-1][cap_per_frame][2]. */
int orbx_search_for_initialization(const orbx_keypoint* kps1, const uint8_t* desc1, int n1, const orbx_keypoint* kps2,
                                   const uint8_t* desc2, int n2, float min_x, float max_x, float min_y, float max_y, float* prev_matched,
                                   int window_size, float nnratio, int check_orientation, int32_t* matches12, int32_t* nmatches,
                                   int device);
int orbx_search_for_initialization_batch(orbx_handle h, int use_undistorted, float min_x, float max_x, float min_y, float max_y,
                                         int window_size, float nnratio, int check_orientation, int32_t* matches12, int32_t* nmatches,
                                         float* prev_matched);

/* ---- Frame::ComputeBoW (row N4): DBoW2 vocabulary-tree transform --------------------------------------------------
 * Replaces mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4) (ORB_SLAM3/src/Frame.cc:738-745), i.e.
 * DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>::transform(features, BowVector&, FeatureVector&, levelsup)
 * (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1126-1196) with the per-feature descent of :1217-1263 and
 * FORB::distance (FORB.cpp:81-100), for the ORB vocabulary's settings: TF_IDF weighting, L1_NORM scoring.
 *
 * The vocabulary is handed over as flat arrays (what loadFromTextFile builds, TemplatedVocabulary.h:1340-1420):
 * node 0 = root, parent[i] < i for i >= 1, children of a node are visited in increasing node id (the loader appends
 * them in file order), a node without children is a word and words are numbered in node order, weight[i] = the
 * node's idf weight (only read for words; <= 0 = stopped word), L = the tree depth of the file header.
 *
 * Outputs per frame f (strides: `cap` entries, node_begin cap + 1), in the iteration order of the reference's std::maps:
 *   BowVector      word_ids[f][0 .. n_words[f])   ascending, word_vals = L1-normalised tf-idf (doubles, bit-exact)
 *   FeatureVector  node_ids[f][0 .. n_nodes[f])   ascending node ids at level L - levelsup (0 = root if that is <= 0),
 *                  node_begin[f][j] .. node_begin[f][j+1] = the slice of feat_idx[f] holding node j's feature indices
 *                  in insertion (= ascending) order.
 * Where the reference reads an uninitialised NodeId (a leaf above level L - levelsup, :1225-1258) this returns 0. */
typedef struct orbx_vocabulary* orbx_voc;
int orbx_voc_create(int n_nodes, const int32_t* parent, const uint8_t* node_desc, const double* weight, int L, int device,
                    orbx_voc* out);
void orbx_voc_destroy(orbx_voc v);
int orbx_voc_info(orbx_voc v, int* n_nodes, int* n_words, int* max_children, int* L);
/* desc: host, [nframes][cap][32]; n_per_frame: host, descriptors used of each frame (<= cap <= 16384). */
int orbx_bow_transform(orbx_voc v, const uint8_t* desc, const int32_t* n_per_frame, int nframes, int cap, int levelsup,
                       int32_t* word_ids, double* word_vals, int32_t* n_words, int32_t* node_ids, int32_t* node_begin,
                       int32_t* feat_idx, int32_t* n_nodes);
/* The same on the descriptors of the handle's LAST extraction, read in place on the device (no descriptor download);
 * cap = the cap_per_frame of orbx_batch_results, nframes = the frames of that extraction. */
int orbx_bow_transform_extracted(orbx_voc v, orbx_handle h, int levelsup, int32_t* word_ids, double* word_vals,
                                 int32_t* n_words, int32_t* node_ids, int32_t* node_begin, int32_t* feat_idx,
                                 int32_t* n_nodes);

/* ---- ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)  (src/ORBmatcher.cc:223-425), monocular case ------
 * The consumer of the FeatureVectors: keyframe features are matched to frame features of the SAME vocabulary node only.
 * Inputs are the two descriptor sets, the two FeatureVectors in the CSR form orbx_bow_transform returns them in, the
 * keypoint angles (kp.angle of pKF->mvKeysUn / F.mvKeys) and valid_kf[i]: 1 = keyframe feature i has a good MapPoint,
 * 0 = none, 2 = MapPoint::isBad().  Reproduced: node-by-node walk in ascending node id, keyframe features in node order,
 * frame features that already received a MapPoint are skipped (:279), strict-< best / second, accept
 * best1 <= TH_LOW && (float)best1 < nnratio * (float)best2 (:323-327), rotation histogram + ComputeThreeMaxima
 * (:345-352, :403-420, :2012-2053) when check_orientation != 0.
 * match_f[j] = index of the keyframe feature whose MapPoint goes to frame feature j (vpMapPointMatches[j] =
 * vpMapPointsKF[match_f[j]]), or -1; *n_matches = the function's return value.  n_f < 65536. */
int orbx_search_by_bow(const uint8_t* desc_kf, int n_kf, const uint8_t* valid_kf, const float* angle_kf, const int32_t* kf_node_ids,
                       const int32_t* kf_node_begin, const int32_t* kf_feat_idx, int kf_nodes, const uint8_t* desc_f, int n_f,
                       const float* angle_f, const int32_t* f_node_ids, const int32_t* f_node_begin, const int32_t* f_feat_idx,
                       int f_nodes, float nnratio, int check_orientation, int32_t* match_f, int32_t* n_matches, int device);
/* ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12) (ORBmatcher.cc:765-905, monocular
 * keyframes; loop closing and map merging): the same node-by-node walk, but a candidate of keyframe 2 needs a good MapPoint
 * and must not be matched yet (vbMatched2, :826-830), the accept rule is bestDist1 < TH_LOW (strict, :848) and the result is
 * indexed by keyframe 1: match12[i] = feature of keyframe 2 whose MapPoint goes to feature i, or -1.  valid1 / valid2 as
 * valid_kf above. */
int orbx_search_by_bow_kf(const uint8_t* desc1, int n1, const uint8_t* valid1, const float* angle1, const int32_t* node_ids1,
                          const int32_t* node_begin1, const int32_t* feat_idx1, int nodes1, const uint8_t* desc2, int n2,
                          const uint8_t* valid2, const float* angle2, const int32_t* node_ids2, const int32_t* node_begin2,
                          const int32_t* feat_idx2, int nodes2, float nnratio, int check_orientation, int32_t* match12,
                          int32_t* n_matches, int device);

/* Measured sustained 32-bit POPC rate of the device in giga-POPC/s (8 independent XOR+POPC chains per thread):
 * the denominator of the matcher's roofline, pairs/s = POPC/s / 8. */
int orbx_bench_popc(int device, double* gpopc_per_s);

/* ---- descriptor database sharded over ranks (config 5): one process per GPU ----------------------
 * Rank 0 calls orbx_comm_unique_id and distributes the 128 bytes out of band (the host launcher's
 * rendezvous); every rank then calls orbx_comm_init.  orbx_match_top2_sharded computes the local top-2 of
 * the replicated queries against this rank's shard (rows [db_index_base, db_index_base + nd_local) of the
 * global database), ncclAllGather's the per-rank candidates and merges them on every rank by
 * (distance asc, global index asc). */
typedef struct orbx_comm_s* orbx_comm;
int orbx_comm_unique_id(uint8_t id[128]);
int orbx_comm_init(int rank, int nranks, const uint8_t id[128], int device, orbx_comm* out);
void orbx_comm_destroy(orbx_comm c);
int orbx_match_top2_sharded(orbx_comm c, const uint8_t* d_q, int nq, const uint8_t* d_db_local, int64_t nd_local,
                            int64_t db_index_base, int64_t* d_idx1, int32_t* d_d1, int64_t* d_idx2, int32_t* d_d2);
/* The same, only ENQUEUED on the communicator's stream (no host synchronisation): consecutive calls on one communicator run back to
 * back on the device; orbx_comm_sync waits for all of them.  nq must not grow between un-synchronised calls. */
int orbx_match_top2_sharded_async(orbx_comm c, const uint8_t* d_q, int nq, const uint8_t* d_db_local, int64_t nd_local,
                                  int64_t db_index_base, int64_t* d_idx1, int32_t* d_d1, int64_t* d_idx2, int32_t* d_d2);
int orbx_comm_sync(orbx_comm c);
/* The merge alone, on already gathered candidates laid out [rank][query] (used by the CPU-side tests of
 * the exchange step and by the single-process emulation of G shards on one GPU). */
int orbx_merge_top2_device(const uint64_t* d_keys1, const uint64_t* d_keys2, int nranks, int nq, int64_t* d_idx1,
                           int32_t* d_d1, int64_t* d_idx2, int32_t* d_d2, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H_ */
