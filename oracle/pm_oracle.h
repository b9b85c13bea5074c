/* oracle/pm_oracle.h -- TEST INFRASTRUCTURE. C interface of the CPU oracle.
 *
 * The oracle is a CPU restatement of the reference's semi-dense probability-mapping path
 * (/root/reference/src/Modeler/ProbabilityMapping.cc, "PM.cc"; constants from
 * include/Modeler/ProbabilityMapping.h:38-49, "PM.h").  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product path
 * (orb-slam-free-space-carving_b200/) never includes, links or calls anything in this directory.
 *
 * PARITY STATUS: "parity unpinned" against the reference itself -- PM.cc is orphaned, does not
 * compile against its own header and the reference ships no tests or golden
 * vectors (F5).  What IS pinned: the third-party OpenCV arithmetic (cvlike.h) against cv2 4.13
 * golden vectors, hand-derived known-answer tests, and recovery of the synthetic scenes'
 * ground-truth depth.
 */
#ifndef PM_ORACLE_H
#define PM_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Run-time copy of the compile-time tunables, PM.h:38-49, plus the magic literals in PM.cc. */
typedef struct OraParams {
    float lambdaG;      /* PM.h:40  = 8   gradient-magnitude gate                       */
    float lambdaL;      /* PM.h:41  = 80  max angle gradient vs epipolar line (degrees)   */
    float lambdaTheta;  /* PM.h:42  = 45  max gradient-orientation difference (degrees)   */
    int   lambdaN;      /* PM.h:43  = 3   min compatible hypotheses / neighbours          */
    double theta;       /* PM.h:47  = 0.23 (THETA, a double literal)                      */
    double chi2_fuse;   /* PM.cc:917,923 = 5.99                                           */
    double chi2_inter;  /* PM.cc:710      = 3.84                                           */
    float slope_limit;  /* PM.cc:393 = 4                                                  */
    int   border;       /* PM.cc:198-199 = 2                                              */
    int   interkf_inplace; /* 0: snapshot inter-KF check (H1); 1: PM.cc:764,793 in place, in call order */
} OraParams;

void ora_default_params(OraParams* p);
void* ora_create(const OraParams* p /* NULL = defaults */);
void ora_destroy(void* h);
void ora_set_threads(int n); /* OpenMP threads for the five pragma loops (PM.cc:197,339,491,554,658); 1 = as shipped */
int ora_max_threads(void);

/* Key-frame view: image + intrinsics K = {fx, fy, cx, cy} + pose Tcw (3x4 row major).  Runs the
 * build-defined gradient front end: 3x3 Sobel/8 with replicate border, magnitude,
 * fastAtan2 orientation in degrees, population std-dev of the u8 image. */
int ora_add_keyframe(void* h, int kf_id, const uint8_t* gray, int w, int hgt, int stride, const float K[4], const float Tcw[12]);
int ora_set_pose(void* h, int kf_id, const float Tcw[12]);
/* PM.cc:137-257 minus the neighbour pick / rot / prior, which the caller supplies */
int ora_semidense(void* h, int kf_id, const int* nbr_ids, const float* nbr_rot_deg, int n, float min_depth, float max_depth);
/* PM.cc:628-799, snapshot semantics */
int ora_interkf(void* h, int kf_id, const int* nbr_ids, int n);
/* PM.cc:337-367 */
int ora_pointset(void* h, int kf_id);
/* PM.cc:114-127 selection (sigma > max_sigma: skip; depth > 1e-6: keep) as a point list, optional smallest-sigma cut;
 * returns the number selected (outputs hold `capacity` entries, each nullable) */
int ora_confident_points(void* h, int kf_id, double max_sigma, int max_points, int capacity, float* xyz, float* sigma, uint32_t* pixel);

/* test hook: install (depth, sigma) maps as the intra-KF result of a key frame (counterpart of sd_import_maps) */
int ora_set_maps(void* h, int kf_id, const float* depth, const float* sigma);

enum { ORA_GRAD = 0, ORA_THETA = 1, ORA_DEPTH = 2, ORA_SIGMA = 3, ORA_DEPTH_CHECKED = 4, ORA_POINTS = 5 };
int ora_get_plane(void* h, int kf_id, int what, float* out); /* H*W floats, POINTS: H*3W */
float ora_get_istddev(void* h, int kf_id);

/* ---- single-function probes (known-answer tests) ---- */
float ora_fast_atan2(float y, float x);
float ora_bilinear_u8(const uint8_t* img, int w, int h, float y, float x);   /* PM.cc:40-59 */
float ora_bilinear_f32(const float* img, int w, int h, float y, float x);
int   ora_chi_test(float a, float b, float sa, float sb, double thr);             /* PM.cc:920-924 */
void  ora_get_fusion(const float* d, const float* s, int n, float* depth, float* sigma, float* min_sigma); /* PM.cc:947-970 */
void  ora_get_fusion_pairs(const float* d, const float* s, int n, float* depth, float* sigma);             /* PM.cc:926-945 */
int   ora_hypothesis_fusion(const OraParams* p, const float* d, const float* s, int n, float* depth, float* sigma); /* PM.cc:598-626, returns supported */
void  ora_stereo_search_constraints(const float* depths, int n, float* min_depth, float* max_depth);     /* PM.cc:370-383 */
float ora_rot_in_plane_median(const long long* mp1, const float* ang1, int n1, const long long* mp2, const float* ang2, int n2); /* PM.cc:467-484 + 172-177 */
void  ora_pair_geometry(const float K1[4], const float Tcw1[12], const float K2[4], const float Tcw2[12],
                        float F12[9], float R21[9], float t21[3]);                /* PM.cc:972-987, 890-891 */
void  ora_search_range(const float K1[4], const float R21[9], const float t21[3], int cols, int px, int py,
                       float mind, float maxd, float* umin, float* umax);         /* PM.cc:877-910 */
void  ora_intra_check_planes(const OraParams* p, float* depth, float* sigma, int w, int h);                    /* PM.cc:486-547 */
void  ora_intra_grow_planes(const OraParams* p, float* depth, float* sigma, const float* grad, int w, int h);  /* PM.cc:549-596 */
float ora_pixel_depth(const float K1[4], const float R21[9], const float t21[3], float uj, int px, int py); /* PM.cc:845-875 */
/* PM.cc:385-465 on one pixel; returns 1 if a hypothesis was produced */
int   ora_epipolar_search(void* h, int kf1, int kf2, int x, int y, float min_depth, float max_depth, float rot,
                          float* depth, float* sigma, float* best_u, float* best_v);
/* cvlike probes, replayed against tests/golden/cv_golden.npz */
void ora_cv_gemm_small_33x31(const float A[9], const float b[3], double alpha, const float* c, double beta, float d[3]);
void ora_cv_gemm_small_33x33(const float A[9], const float B[9], float D[9]);
void ora_cv_gemm_small_44x41(const float A[16], const float b[4], float d[4]);
void ora_cv_gemm_acc64_ABt(const float A[9], const float B[9], double alpha, float D[9]);
float ora_cv_gemm_acc64_dot(const float* a, const float* b, int n, double alpha);
int  ora_cv_invert33(const float A[9], float D[9]);
int  ora_cv_solve_lu33(const float A[9], const float B[9], float X[9]);

#ifdef __cplusplus
}
#endif
#endif
