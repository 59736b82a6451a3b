/*
 * ibl.h -- C ABI of libibl.so: the B200 (sm_100a) implementation of the per-frame
 * object-localisation hot path of instance-based-loc.
 *
 * The reference has no FFI on this path: its "operator API" is the Python signatures of
 * utils/depth_utils.py, utils/similarity_volume.py, utils/fpfh_register.py and the Open3D 0.17
 * calls they make.  Each entry point below names the reference call (file:line) it replaces.
 * The Python mirror in instance_based_loc_b200/ binds these through ctypes.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name starts with h_ ; plain C types only
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), except where a
 *     count must be returned to the host (documented per function: "syncs")
 *   - return value: >= 0 success (sometimes a count), < 0 an IBL_E_* code;
 *     ibl_last_error_string() gives detail for the calling thread
 *   - points / colours / normals are row-major double[n][3] (Open3D stores f64; exact f64
 *     arithmetic in nanoflann accumulation order is what makes counts and index sets bit-exact)
 *   - all floating-point arithmetic is compiled with FMA contraction disabled
 */
#ifndef IBL_H_
#define IBL_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IBL_OK 0
#define IBL_E_CUDA (-1)
#define IBL_E_ARG (-2)
#define IBL_E_NOMEM (-3)
#define IBL_E_RANGE (-4)

const char *ibl_last_error_string(void);
int ibl_version(void);

/* ---------------------------------------------------------------- depth -> clouds ----------- */
/* utils/depth_utils.py:46-90 + :176-206 (get_mask_coloured_pointclouds_from_depth): ONE pass over
 * the frame for all masks.  depth is float32[H*W] (depth_is_f64==0) or float64[H*W]; rgb uint8
 * [H*W*3] or NULL; masks uint8[n_masks][H*W] (0/1); hor float32[W], ver float32[H] are the
 * reference's np.linspace vectors.  Points of mask i are written, in row-major pixel order, to
 * out_pts/out_cols rows [h_offsets[i], h_offsets[i+1]).  out_* must hold n_masks*H*W rows
 * (or pass out_pts==NULL to only count).  h_offsets: host int64[n_masks+1].  Syncs. */
int ibl_backproject_masks(const void *depth, int depth_is_f64, const uint8_t *rgb, const uint8_t *masks,
                          int n_masks, int H, int W, const float *hor, const float *ver, double fx, double fy,
                          double *out_pts, double *out_cols, int64_t *h_offsets, void *stream);

/* utils/depth_utils.py:211-265 voxel_down_sample_with_colors.  vox_id (optional) receives, per
 * input point, the rank of its voxel in first-occurrence order; h_voxel_keys (optional, host
 * int64[n][3]) is not written by the device path.  cols / nrm may be NULL.  Returns the number of
 * voxels (syncs).  Outputs must hold n rows. */
int64_t ibl_voxel_downsample(const double *pts, const double *cols, const double *nrm, int64_t n, double voxel,
                             double *out_pts, double *out_cols, double *out_nrm, int32_t *vox_id,
                             int64_t *out_keys /* int64[n][3] floor(p/voxel), may be NULL */, void *stream);

/* order-preserving row selection (Open3D SelectByIndex used by remove_radius_outlier).
 * keep: uint8[n]; a/b/c are optional double[n][3] arrays compacted identically; out_index
 * (optional int32) receives the kept indices.  Returns kept count (syncs). */
int64_t ibl_select_rows(const uint8_t *keep, int64_t n, const double *a, const double *b, const double *c,
                        double *out_a, double *out_b, double *out_c, int32_t *out_index, void *stream);

/* Open3D PointCloud::Transform / translate (object_memory.py:1033-1034, fpfh_register ICP):
 * out = ((T00*x + T01*y) + T02*z) + T03 per row; T is HOST row-major double[16]. rotate_only=1
 * applies R only (normals). in==out allowed. */
int ibl_transform_points(const double *h_T16, const double *in, int64_t n, int rotate_only, double *out,
                         void *stream);
/* column means in numpy's exact order (np.mean(points, axis=0), object_memory.py:1030-1031, adds the rows
 * one by one): bit-identical to the reference's detected_mean / memory_mean.  h_mean host double[3]. Syncs. */
int ibl_mean3(const double *pts, int64_t n, double *h_mean, void *stream);
/* ibl_mean3 of k clouds in one launch and one synchronisation (object_memory.py:1030-1031: one mean per assignment union);
 * ptrs / ns host arrays of device pointers and row counts, h_means host double[k][3]. */
int ibl_mean3_many(const double *const *ptrs, const int64_t *ns, int32_t k, double *h_means, void *stream);

/* ---------------------------------------------------------------- uniform grid -------------- */
typedef struct ibl_grid ibl_grid_t;
/* Cell-sorted uniform grid over pts (replaces Open3D KDTreeFlann construction at every call
 * site).  cell is a request: it is enlarged until nx*ny*nz <= max_cells (0 -> 2^24).  Syncs. */
int ibl_grid_create(const double *pts, int64_t n, double cell, int64_t max_cells, ibl_grid_t **out, void *stream);
/* One grid over the concatenation of several clouds, seg int32[n] (device) = the cloud of every point.  Radius queries on
 * it only count neighbours of the query's own cloud: one build + one query replace a build + a query per cloud (the
 * per-mask outlier filters of a frame, utils/depth_utils.py:85-88, object_memory.py:992-998).  Only ibl_radius_count and
 * ibl_radius_outlier_mask are defined on a segmented grid.  Syncs. */
int ibl_grid_create_segmented(const double *pts, int64_t n, const int32_t *seg, double cell, int64_t max_cells,
                              ibl_grid_t **out, void *stream);
void ibl_grid_destroy(ibl_grid_t *g);
double ibl_grid_cell(const ibl_grid_t *g);
int64_t ibl_grid_occupied(const ibl_grid_t *g);
/* Chebyshev ring-distance field (uint8 per cell, capped) used to skip empty space in nn queries */
int ibl_grid_build_ringfield(ibl_grid_t *g, int max_ring, void *stream);

/* Exact nearest-neighbour candidate field for "nearest target point within max_d" queries (replaces the KD-tree of
 * GetRegistrationResultAndCorrespondences, Registration.cpp, for RANSAC validation / ICP / evaluate_registration):
 * sparse voxel table (side `voxel`, 0 -> max_d/4) whose entries list every point that can be the answer for a query
 * inside the voxel.  Once built, ibl_nn_score / ibl_ransac_score / ibl_icp use it whenever their max_d <= this
 * max_d; results are identical to the ring search.  Syncs (two size read-backs). */
int ibl_grid_build_nnfield(ibl_grid_t *g, double max_d, double voxel, void *stream);
int64_t ibl_grid_nnfield_bytes(const ibl_grid_t *g);

/* KDTreeFlann::SearchRadius count for every grid point against the grid (A.2):
 * count[i] = #{j : d2(i,j) < r*r}.  depth_utils.py:41,87, object_memory.py:994. */
int ibl_radius_count(const ibl_grid_t *g, double radius, int32_t *count, void *stream);
/* keep[i] = count[i] > nb_points fused variant */
int ibl_radius_outlier_mask(const ibl_grid_t *g, double radius, int32_t nb_points, uint8_t *keep, void *stream);

/* Open3D PointCloud::cluster_dbscan(eps, min_points) (memory build: object_memory/object_memory.py:315,511,632) for the
 * grid's points: labels int32[n] device, -1 = noise, clusters numbered in the order of their smallest core index (what
 * Open3D's index-order scan produces); core iff #{j : d2(i,j) < eps*eps} (self included) >= min_points; a border point
 * takes the smallest cluster number among the core points within eps.  The grid's cell should be >= eps (a hair larger is
 * cheapest).  h_nclusters (host, may be NULL) receives the number of clusters; syncs when it is given. */
int ibl_dbscan(const ibl_grid_t *g, double eps, int32_t min_points, int32_t *labels, int32_t *h_nclusters, void *stream);
/* Synthetic RGB-D frames (SURVEY.md 8(d) inputs / 8(f) row S1; stands in for the reference's habitat-sim rendering,
 * hm3d_data_generation.py): one ray per pixel of an H x W pin-hole frame (focal f, pixel grid = the float32 linspace
 * values utils/depth_utils.py:57-63 back-projects with) against the room (floor z = 0, ceiling `height`, walls at
 * +-room) and m oriented boxes (kind 0) / ellipsoids (kind 1): half double[m][3], R double[m][9] object->world
 * row-major, t double[m][3], all device.  h_Twc16 = HOST row-major camera->world 4x4.  Outputs (device): out_t
 * double[H*W] camera depth of the first hit (+inf: none), out_inst int32[H*W] object id (-1: room or nothing),
 * out_nrm double[H*W][3] world-frame surface normal.  Produces inputs only; never timed. */
int ibl_raycast_scene(const double *h_Twc16, int32_t H, int32_t W, double f, const int32_t *kind, const double *half,
                      const double *R, const double *t, int32_t m, double room, double height, double *out_t,
                      int32_t *out_inst, double *out_nrm, void *stream);
/* axis-aligned bounds of a cloud, bit-identical to numpy's pcd.min(axis=-1) / pcd.max(axis=-1) (utils/IoU_ops.py:24-28,
 * 71-75): h_minmax6 = host double[6] (min xyz, max xyz).  n >= 1.  Syncs. */
int ibl_bounds3(const double *pts, int64_t n, double *h_minmax6, void *stream);

/* KDTreeFlann::SearchHybrid(q, radius, max_nn) (A.1) for nq queries; queries==NULL means the
 * grid's own points (row i of the output = point i).  idx int32[nq][max_nn] (-1 padded), d2
 * double[nq][max_nn] (may be NULL), cnt int32[nq].  Ascending (d2, index). */
int ibl_knn_hybrid(const ibl_grid_t *g, const double *queries, int64_t nq, double radius, int32_t max_nn,
                   int32_t *idx, double *d2, int32_t *cnt, void *stream);

/* SearchHybrid(q, max_d, 1) for every row of T*src (h_T16 may be NULL = identity): the body of
 * GetRegistrationResultAndCorrespondences (A.6), used by evaluate_registration
 * (fpfh_register.py:145-150) and each ICP iteration.  nn_idx int32[n] (-1 = none) and nn_d2
 * optional.  h_result: host {int64 n_inliers; double err2}.  Syncs. */
typedef struct { int64_t n_inliers; double err2; } ibl_score_t;
int ibl_nn_score(const ibl_grid_t *g, const double *src, int64_t n, const double *h_T16, double max_d,
                 int32_t *nn_idx, double *nn_d2, ibl_score_t *h_result, void *stream);

/* ibl_nn_score without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same kernels and reduction order as ibl_nn_score, so n_inliers / err2 / nn_idx / nn_d2 are bit-identical; the
 * per-CTA partials live in a file-scope static device array (64 KiB, held by every GPU that loads the library until
 * exit) instead of stream-ordered scratch.  Limits and rules:
 *   - n <= IBL_NN_SCORE_STATIC_MAX_POINTS, else IBL_E_RANGE with a message (use ibl_nn_score);
 *   - d_result is a DEVICE ibl_score_t owned by the caller (required); nothing is copied to the host and the call does
 *     not synchronise -- everything is ordered on `stream`;
 *   - calls no cudaMalloc / cudaFree / allocating library routine.  The first call on each device looks the array's
 *     address up (cudaGetSymbolAddress, that device current): make one call before starting a graph capture;
 *   - all calls share the one array and it keeps the last call's contents (no part of it needs to start at zero: a call
 *     overwrites every partial it later reads).  Calls that overlap on the GPU RACE on it: serialise them. */
#define IBL_NN_SCORE_STATIC_MAX_POINTS (1 << 20)
int ibl_nn_score_static(const ibl_grid_t *g, const double *src, int64_t n, const double *h_T16, double max_d,
                        int32_t *nn_idx, double *nn_d2, ibl_score_t *d_result, void *stream);

/* ---------------------------------------------------------------- normals / FPFH ------------ */
/* PointCloud::EstimateNormals(Hybrid) covariance + FastEigen3x3 (A.3), fpfh_register.py:90-92 */
int ibl_normals(const double *pts, int64_t n, const int32_t *nbr_idx, const int32_t *nbr_cnt, int32_t max_nn,
                double *normals, void *stream);
/* ComputeSPFHFeature / ComputeFPFHFeature (A.4), fpfh_register.py:94-97.  Device layout is
 * POINT-MAJOR: spfh, fpfh double[n][33] (the Python mirror exposes Open3D's [33][n] view). */
int ibl_spfh(const double *pts, const double *nrm, int64_t n, const int32_t *nbr_idx, const int32_t *nbr_cnt,
             int32_t max_nn, double *spfh, void *stream);
int ibl_fpfh(int64_t n, const int32_t *nbr_idx, const double *nbr_d2, const int32_t *nbr_cnt, int32_t max_nn,
             const double *spfh, double *fpfh, void *stream);
/* exact 33-d nearest neighbour of every column of fa among the columns of fb (A.5 KDTreeFlann on
 * features, SearchKNN k=1); lowest index wins ties.  fa double[na][33], fb double[nb][33]. */
int ibl_feature_nn(const double *fa, int64_t na, const double *fb, int64_t nb, int32_t *nn, void *stream);
/* Cached matcher operands of ONE feature set (a memory-side union's FPFH, matched against a new source cloud every frame):
 * the filter's centre and power-of-two scale, packed query-role and target-role tile images, norms, fp32 rows.  `f` must
 * stay alive as long as the pack.  No reference counterpart (Open3D rebuilds its feature KD-trees on every call). */
typedef struct ibl_feature_pack ibl_feature_pack_t;
int ibl_feature_pack_create(const double *f, int64_t n, ibl_feature_pack_t **out, void *stream);
void ibl_feature_pack_destroy(ibl_feature_pack_t *pack);
/* Both directions of the A.5 search for a fresh set fa double[na][33] against a pack: nn_ab int32[na] (nearest pack row per
 * row of fa) and nn_ba int32[pack rows] (nearest row of fa per pack row); identical to two ibl_feature_nn calls.  fa is
 * read once.  d_flag (device uint32, may be NULL) is OR-ed with 1 if fa does not fit the pack's fp16 range: the outputs
 * are then invalid and the caller must fall back to ibl_feature_nn. */
int ibl_feature_match_packed(const double *fa, int64_t na, const ibl_feature_pack_t *pack, int32_t *nn_ab, int32_t *nn_ba,
                             uint32_t *d_flag, void *stream);
/* The same for callers that apply the mutual filter (CorrespondencesFromFeatures keeps (i, j = nn_ab[i]) iff nn_ba[j] == i,
 * fpfh_register.py:110-112 with mutual_filter = True): nn_ab as above; nn_ba[j] is computed for the pack rows j that occur
 * in nn_ab -- the only ones the filter reads -- and is -1 for every other row.  The reverse search then runs over a few
 * thousand rows instead of the whole memory side. */
int ibl_feature_match_packed_mutual(const double *fa, int64_t na, const ibl_feature_pack_t *pack, int32_t *nn_ab,
                                    int32_t *nn_ba, uint32_t *d_flag, void *stream);
/* Debug / test hook of the tensor-core filter inside ibl_feature_nn (csrc/match_tc.cu): the raw tcgen05 (kind::f16, 3-way fp16 split) scores
 * s'(a_i, b_j) = a32_i . b32_j - |b32_j|^2 / 2 as float32 [na][nb] (row-major), a32 / b32 being the rows minus the filter's
 * centre, times its power-of-two scale; centre (33) and scale (1) are written to center_out (device double[34]) when
 * not NULL.  swap_lbo_sbo must be 0 (1 is a
 * descriptor-layout probe).  Not on the product path. */
int ibl_feature_scores_debug(const double *fa, int64_t na, const double *fb, int64_t nb, float *scores,
                             double *center_out, int swap_lbo_sbo, void *stream);
/* Test hook: cap ibl_feature_nn's candidate pool at `chunks` chunks of 64 candidates (0 = default sizing) to force the
 * pool-exhausted path (candidates evaluated on the spot).  Process-wide; results are identical either way. */
int ibl_feature_nn_set_pool_chunks(int chunks);
/* Instrumentation of ibl_feature_nn: h_out2 = {band candidates, exact f64 evaluations} since the last reset.  Syncs. */
int ibl_feature_nn_stats(int64_t *h_out2, int reset);
/* instrumentation: who waits for whom inside the tcgen05 matcher kernel.  mode 1 zeroes the counters and switches
 * collection on, mode 0 reads them and switches it off.  h_out8 = {CTAs, target tiles streamed, MMA thread's loop clocks,
 * of those waiting for a free accumulator (epilogue-bound), waiting for a target tile (copy-bound), epilogue warp 0's loop
 * clocks, of those waiting for an accumulator (MMA-bound), producer's clocks waiting for a free stage}.  No reference
 * counterpart.  Syncs. */
int ibl_feature_nn_wait_stats(int64_t *h_out8, int mode);
/* mutual filter + fallback (A.5). ij int32[ns], ji int32[nt] -> corres int32[<=ns][2]; returns C. Syncs. */
int64_t ibl_mutual_corres(const int32_t *ij, int64_t ns, const int32_t *ji, int64_t nt, int32_t ransac_n,
                          int32_t *corres, void *stream);

/* ---------------------------------------------------------------- RANSAC -------------------- */
typedef struct {
    double T[16];          /* row-major 4x4 */
    double fitness;        /* n_inliers / n_src */
    double inlier_rmse;    /* sqrt(err2 / n_inliers) */
    int64_t n_inliers;
    double err2;
    int64_t best_hyp;      /* index of the winning hypothesis in the sample stream (-1 none) */
    int64_t drawn;         /* hypotheses consumed before the adaptive stop */
    int64_t validated;     /* hypotheses that passed the cheap checkers and were scored */
    int64_t n_corres;
    int32_t iterations;    /* ICP iterations performed */
    int32_t status;        /* 0 ok; 1 = empty result (too few correspondences etc.) */
} ibl_reg_result_t;

/* Hypothesis samples: either an explicit host-seeded array samples int32[.][3] (device), or
 * samples==NULL and the counter-based stream ibl_sample(seed, h, j, C) (rigid.cuh; restated in
 * oracle/sampler.py so the oracle consumes identical draws). */

/* A.7 hypothesis generation + EdgeLength/Distance checkers for hypotheses [h0, h0+H).
 * pass uint8[H]; T_out double[H][16] optional (written only where pass, or everywhere when
 * use_checkers==0).  fpfh_register.py:110-119. */
int ibl_ransac_check(const double *src, const double *tgt, const int32_t *corres, int64_t n_corres,
                     const int32_t *samples, uint64_t seed, int64_t h0, int64_t H, double edge_thr, double dist_thr,
                     int use_checkers, uint8_t *pass, double *T_out, void *stream);
/* Score hypotheses hyp_ids int64[nh] over the whole source (A.6): inliers int32[nh], err2
 * double[nh], corres_inliers int32[nh] (EvaluateInlierCorrespondenceRatio numerator), T_out
 * double[nh][16].  Source is tiled over CTAs; partials are combined in a fixed order. */
int ibl_ransac_score(const ibl_grid_t *tgt_grid, const double *src, int64_t n_src, const double *tgt,
                     const int32_t *corres, int64_t n_corres, const int32_t *samples, uint64_t seed,
                     const int64_t *hyp_ids, int64_t nh, double max_d, int32_t *inliers, double *err2,
                     int32_t *corres_inliers, double *T_out, void *stream);
/* Inlier counts only, for hypotheses hyp_ids int64[nh]: inliers[k] equals ibl_ransac_score's inliers[k] (tested), at a
 * fraction of the cost -- a point's scan stops at the first target within max_d, and voxels of the NN field that lie
 * wholly within max_d of a target need no scan at all.  T_out double[nh][16]. */
int ibl_ransac_count(const ibl_grid_t *tgt_grid, const double *src, int64_t n_src, const double *tgt,
                     const int32_t *corres, int64_t n_corres, const int32_t *samples, uint64_t seed,
                     const int64_t *hyp_ids, int64_t nh, double max_d, int32_t *inliers, double *T_out, void *stream);
/* Whole RegistrationRANSACBasedOnCorrespondence loop with sequential semantics (A.7): blocks of
 * `block` hypotheses are checked, the survivors scored, and the ordered update / adaptive stop
 * applied on the host.  Syncs. */
int ibl_ransac_run(const ibl_grid_t *tgt_grid, const double *src, int64_t n_src, const double *tgt,
                   const int32_t *corres, int64_t n_corres, const int32_t *samples, uint64_t seed,
                   int64_t max_iteration, double confidence, double max_d, double edge_thr, int use_checkers,
                   int early_exit, int64_t block, ibl_reg_result_t *h_result, void *stream);
/* The same loop over hypotheses [h_begin, max_iteration) of the sample stream only: the block one rank owns when a
 * single registration is sharded by hypothesis over GPUs (secondary axis); `drawn` counts from h_begin. */
int ibl_ransac_run_range(const ibl_grid_t *tgt_grid, const double *src, int64_t n_src, const double *tgt,
                   const int32_t *corres, int64_t n_corres, const int32_t *samples, uint64_t seed,
                   int64_t h_begin, int64_t max_iteration, double confidence, double max_d, double edge_thr, int use_checkers,
                   int early_exit, int64_t block, ibl_reg_result_t *h_result, void *stream);

/* ---- batched registration: every (frame, assignment) problem of a step advances together ---- */
/* One register_point_clouds problem (utils/fpfh_register.py:100-143) after feature matching.  All pointers are device
 * memory; src / src_col is the centred detected union, tgt* the cached memory-side union with its normals, colours and
 * colour gradients; the two grids carry the exact-NN candidate fields for the RANSAC and the ICP distances. */
typedef struct {
    const double *src;               /* [n_src][3] */
    const double *src_col;           /* [n_src][3] (coloured ICP) */
    int64_t n_src;
    const double *tgt;               /* [n_tgt][3] */
    const double *tgt_nrm, *tgt_col, *tgt_grad;   /* [n_tgt][3] (coloured ICP) */
    int64_t n_tgt;
    const ibl_grid_t *grid_ransac;   /* target grid, NN field built for the RANSAC max_d */
    const ibl_grid_t *grid_icp;      /* target grid, NN field built for the ICP max_d */
    const int32_t *corres;           /* [n_corres][2] (source index, target index) */
    int64_t n_corres;
    uint64_t seed;                   /* counter-based sample stream ibl_sample(seed, h, j, n_corres) */
} ibl_reg_problem_t;
/* ibl_ransac_run for n_problems independent problems at once (object_memory.py:1020-1108 registers up to 6 assignments
 * per frame one after the other): per round one block of hypotheses per active problem, ONE launch per stage for the whole
 * batch and two host synchronisations (inlier counts of the validated hypotheses, exact scores of the contenders).  Every field of
 * h_results[p] equals what ibl_ransac_run returns for problem p alone.  Syncs. */
int ibl_ransac_run_batch(const ibl_reg_problem_t *problems, int32_t n_problems, int64_t max_iteration, double confidence,
                         double max_d, double edge_thr, int use_checkers, int early_exit, int64_t first_block,
                         ibl_reg_result_t *h_results, void *stream);
/* Measurement switch for ibl_ransac_run_batch's count pass: 1 (default) = a hypothesis is dropped as soon as its inliers so
 * far plus the points not yet looked at cannot reach the best count of the earlier rounds (exact: such a hypothesis can
 * never be accepted by the ordered rule); 0 = every validated hypothesis is counted over the whole source.  Results are the
 * same either way. */
int ibl_ransac_set_prune(int on);
/* ibl_icp(mode 1) for n_problems problems at once, from h_T_init[p] (host double[n_problems][16]): one launch per ICP
 * iteration for the whole batch, convergence rule and 6x6 solve per problem on the device, two state read-backs (after 12 iterations
 * and at the end).  Results equal ibl_icp's.  Syncs. */
int ibl_icp_colored_batch(const ibl_reg_problem_t *problems, int32_t n_problems, const double *h_T_init, double max_d,
                          int max_iteration, double rel_fitness, double rel_rmse, double lambda_geometric,
                          ibl_reg_result_t *h_results, void *stream);
/* ibl_nn_score (evaluate_registration, fpfh_register.py:145-150) of n_queries (source, transform) pairs against ONE target
 * grid: srcs / ns / h_T16 are host arrays of device pointers, point counts and row-major 4x4 transforms.  One launch and
 * one synchronisation for all of them; h_results[q] equals ibl_nn_score's result for query q. */
int ibl_nn_score_batch(const ibl_grid_t *g, const double *const *srcs, const int64_t *ns, const double *h_T16,
                       int32_t n_queries, double max_d, ibl_score_t *h_results, void *stream);

/* ---------------------------------------------------------------- ICP ----------------------- */
/* InitializePointCloudForColoredICP (A.8): grad double[n][3] */
int ibl_color_gradient(const double *pts, const double *nrm, const double *col, int64_t n, const int32_t *nbr_idx,
                       const int32_t *nbr_cnt, int32_t max_nn, double *grad, void *stream);
/* TransformationEstimationPointToPoint::ComputeTransformation over (i, nn_idx[i]>=0) pairs
 * (A.9; Eigen::umeyama restated as Jacobi-Kabsch).  h_T16 host out.  Syncs. */
int ibl_p2p_update(const double *src, const int32_t *nn_idx, int64_t n, const double *tgt, double *h_T16,
                   void *stream);
/* TransformationEstimationForColoredICP (A.8) over (i, nn_idx[i]>=0): h_out27 = 21 upper-triangle
 * JTJ (row-major) + 6 JTr; h_T16 (optional) = the solved update.  Deterministic tree reduction. Syncs. */
int ibl_colored_update(const double *src, const double *scol, const int32_t *nn_idx, int64_t n, const double *tgt,
                       const double *tnrm, const double *tcol, const double *tgrad, double lambda_geometric,
                       double *h_out27, double *h_T16, void *stream);
/* RegistrationICP loop (A.9), fpfh_register.py:132-141.  mode 0 = point-to-point, 1 = coloured
 * (scol/tnrm/tcol/tgrad required).  src is not modified.  Syncs every iteration (27 doubles). */
int ibl_icp(const ibl_grid_t *tgt_grid, const double *src, const double *scol, int64_t n_src, const double *tgt,
            const double *tnrm, const double *tcol, const double *tgrad, int mode, double max_d,
            const double *h_T_init, int max_iteration, double rel_fitness, double rel_rmse, double lambda_geometric,
            ibl_reg_result_t *h_result, void *stream);

/* ---------------------------------------------------------------- similarity ---------------- */
/* object_memory.py:922-936: S[i][j] = max over stored embeddings e of object j of <mem_e/|mem_e|,
 * det_i/|det_i|>.  det float32[nd][D]; mem float32[E][D]; seg int32[m+1] CSR offsets. */
int ibl_cosine_segmax(const float *det, int32_t nd, const float *mem, int64_t E, int32_t D, const int32_t *seg,
                      int32_t m, float *S, void *stream);
/* utils/similarity_volume.py:102-164,213-270 without materialising (m+1)^3: for every chosen
 * combination of s=min(nd,3) detections, the k best masked fp16 products in (value desc, flat
 * index asc) order.  S float32[nd][m] -> out_flat int64[n_comb][k], out_val float32[n_comb][k]
 * (value = the fp16 product widened, -inf when exhausted with flat index 0). */
int ibl_simvolume_topk(const float *S, int32_t nd, int32_t m, int32_t k, int64_t *out_flat, float *out_val,
                       void *stream);
/* Test / measurement switch: 1 (default) = (i, j) pairs whose k-line cannot reach the value of the k-th best entry are
 * skipped in every pass (exact: fp16 products are monotone in each factor); 0 = every pass walks the whole sub-volume. */
int ibl_simvolume_set_prune(int on);

/* ---------------------------------------------------------------- host wait policy ----------- */
/* How host threads wait inside this library's (and the process's) stream synchronisations on the current device:
 * 0 = driver default (spins when cores outnumber contexts), 1 = spin, 2 = yield, 3 = blocking (sleep on an interrupt),
 * 4 = hybrid (this library's waits poll for ~60 us, then sleep in 20-50 us steps; everything else as 0).
 * The path synchronises a few hundred times per frame (data-dependent sizes, the RANSAC and ICP host rules), so with
 * several frames in flight spinning costs > 2 cores per GPU.  Measured on an 8 x B200 box with 32 cores (DESIGN.md
 * section 5): blocking waits cost 3.4x in frames/s, yield 1.15x -- spin (the default) stays the right choice there; the
 * knob exists for boxes with fewer cores per GPU.  No reference counterpart (Open3D is CPU-only). */
int ibl_set_host_wait(int mode);

/* Reserve `bytes` of stream-ordered scratch memory for `stream` (the library's temporaries come from the device's
 * default cudaMallocAsync pool, which keeps what it has held): call once per frame stream before serving, so no frame
 * waits for the driver to map memory.  No reference counterpart. */
int ibl_reserve_scratch(int64_t bytes, void *stream);
/* {reserved now, reserved high-water, used now, used high-water} bytes of that pool (host array of 4). */
int ibl_scratch_stats(int64_t *h_out4);

/* ---------------------------------------------------------------- instrumentation ------------ */
/* Number of kernels this library has launched in the calling process (bench.py's gpu_launches). */
int64_t ibl_launch_count(void);
/* Optional per-kernel timing with cudaEvents on the launching stream.  ibl_prof_read fills arrays of
 * ibl_prof_num_tags() entries (milliseconds, launches, algorithmic work units: bytes, or f64 flops for
 * feature_nn) accumulated since the previous read, and synchronises the device. */
void ibl_prof_enable(int on);
int ibl_prof_num_tags(void);
const char *ibl_prof_tag_name(int tag);
int ibl_prof_read(double *ms, int64_t *launches, double *work);

#ifdef __cplusplus
}
#endif
#endif /* IBL_H_ */
