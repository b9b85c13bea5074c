// Host-side mirror of the reference's ProbabilityMapping class: every method is a thin forwarder to
// the C ABI of include/sdpm.h (CUDA kernels in libsdpm.so).  See ProbabilityMapping.h.
// Reference: /root/reference/src/Modeler/ProbabilityMapping.cc ("PM.cc").
#include "ProbabilityMapping.h"

#include <algorithm>
#include <cstdio>
#include <stdexcept>

namespace sdpm_host {

ProbabilityMapping::ProbabilityMapping(Map* pMap, int device, int covis_n, const SdParams* params)
    : mpMap(pMap), ctx_(nullptr), covis_n_(covis_n) {
    if (covis_n < 1 || covis_n > SD_MAX_NEIGHBOURS) throw std::invalid_argument("ProbabilityMapping: covis_n out of range");
    int rc = sd_create(device, params, &ctx_);
    if (rc != SD_OK) throw std::runtime_error("ProbabilityMapping: sd_create failed (no usable CUDA device; there is no CPU fallback)");
}

ProbabilityMapping::~ProbabilityMapping() {
    if (ctx_) sd_destroy(ctx_);
}

void ProbabilityMapping::check(int rc, const char* what) const {
    if (rc == SD_OK) return;
    char buf[640];
    std::snprintf(buf, sizeof(buf), "%s: status %d: %s", what, rc, sd_last_error(ctx_));
    throw std::runtime_error(buf);
}

// The reference keeps im_/GradImg/GradTheta/I_stddev in the KeyFrame; here they live in the
// context, keyed by mnId, created on first use (image upload + gradient front end, kernel K0).
void ProbabilityMapping::ensure_uploaded(KeyFrameView* kf) {
    if (std::find(uploaded_.begin(), uploaded_.end(), kf->mnId) != uploaded_.end()) return;
    if (kf->im_.empty()) throw std::invalid_argument("ProbabilityMapping: key frame without image");
    check(sd_upload_keyframe(ctx_, kf->mnId, kf->im_.data, kf->im_.cols, kf->im_.rows, kf->im_.stride, kf->K, kf->Tcw),
          "sd_upload_keyframe");
    uploaded_.push_back(kf->mnId);
}

// PM.cc:151-160 (and 278-287): the first covisN covisible key frames that are neither bad nor unmapped
std::vector<KeyFrameView*> ProbabilityMapping::pick_neighbours(KeyFrameView* kf) const {
    std::vector<KeyFrameView*> closestMatches;
    for (size_t idxCov = 0; idxCov < kf->covisible.size(); idxCov++) {
        if ((int)closestMatches.size() >= covis_n_) break;
        KeyFrameView* c = kf->covisible[idxCov];
        if (c->isBad()) continue;
        if (!c->Mapped()) continue;
        closestMatches.push_back(c);
    }
    return closestMatches;
}

void ProbabilityMapping::Run() {
    SemiDenseLoop();
    UpdateAllSemiDensePointSet();
}

void ProbabilityMapping::SemiDenseLoop() {
    std::vector<KeyFrameView*> vpKFs = mpMap->GetAllKeyFrames();
    // the images of every key frame first, so that the upload stream runs ahead of the compute stream
    for (KeyFrameView* kf : vpKFs)
        if (!kf->isBad()) ensure_uploaded(kf);
    for (KeyFrameView* kf : vpKFs) SemiDenseRecon(kf);

    // PM.cc:262-313
    for (size_t i = 0; i < vpKFs.size(); i++) {
        KeyFrameView* kf = vpKFs[i];
        if (kf->isBad() || kf->interKF_depth_flag_) continue;
        if (!kf->semidense_flag_) continue;
        std::vector<KeyFrameView*> neighbors = pick_neighbours(kf);
        if ((int)neighbors.size() < covis_n_) continue;
        int num_depth_kf = 0;
        for (size_t j = 0; j < neighbors.size(); j++)
            if (neighbors[j]->semidense_flag_) num_depth_kf++;
        if (num_depth_kf < covis_n_) continue;
        InterKeyFrameDepthChecking(kf, neighbors);
        UpdateSemiDensePointSet(kf);
        kf->interKF_depth_flag_ = true;
    }
    // results back to the KeyFrame members, as the reference leaves them (PM.cc:225-226, 306, 361-363)
    for (KeyFrameView* kf : vpKFs)
        if (!kf->isBad() && kf->semidense_flag_) FetchResults(kf, kf->interKF_depth_flag_);
}

void ProbabilityMapping::SemiDenseRecon(KeyFrameView* kf) {
    if (kf->isBad() || kf->semidense_flag_) return;  // PM.cc:141-142
    std::vector<KeyFrameView*> closestMatches = pick_neighbours(kf);
    if ((int)closestMatches.size() < covis_n_) return;  // PM.cc:160

    std::vector<int> ids;
    std::vector<float> rots;
    for (KeyFrameView* kf2 : closestMatches) {  // PM.cc:170-179
        ensure_uploaded(kf2);
        ids.push_back(kf2->mnId);
        rots.push_back(MedianRotInPlane(kf, kf2));
    }
    float max_depth, min_depth;
    StereoSearchConstraints(kf, &min_depth, &max_depth);  // PM.cc:184
    ensure_uploaded(kf);
    // PM.cc:189-238: ComputeFundamental per neighbour, the pixel loop, the two intra-key-frame passes
    check(sd_semidense(ctx_, kf->mnId, ids.data(), rots.data(), (int)ids.size(), min_depth, max_depth), "sd_semidense");
    kf->semidense_flag_ = true;  // PM.cc:244
}

void ProbabilityMapping::StereoSearchConstraints(KeyFrameView* kf, float* min_depth, float* max_depth) {
    check(sd_stereo_search_constraints(kf->point_depths.data(), (int)kf->point_depths.size(), min_depth, max_depth),
          "sd_stereo_search_constraints");
}

void ProbabilityMapping::ComputeFundamental(KeyFrameView* pKF1, KeyFrameView* pKF2, float F12[9]) {
    float R21[9], t21[3];
    check(sd_pair_geometry(pKF1->K, pKF1->Tcw, pKF2->K, pKF2->Tcw, F12, R21, t21), "sd_pair_geometry");
}

float ProbabilityMapping::MedianRotInPlane(KeyFrameView* kf1, KeyFrameView* kf2) {
    const int n1 = (int)std::min(kf1->map_points.size(), kf1->keypoint_angles.size());
    const int n2 = (int)std::min(kf2->map_points.size(), kf2->keypoint_angles.size());
    if (n1 == 0 || n2 == 0) return 0.0f;  // "0 rotation for kf pair without covisibility", PM.cc:174
    return sd_rot_in_plane_median(kf1->map_points.data(), kf1->keypoint_angles.data(), n1, kf2->map_points.data(),
                                  kf2->keypoint_angles.data(), n2);
}

// Single-pixel probe.  `pixel`, `F12` and `th_pi` are derived from kf1/kf2 on the device exactly as
// SemiDenseLoop derives them (PM.cc:202, 192, 214), so the arguments are accepted for signature
// compatibility and ignored.
void ProbabilityMapping::EpipolarSearch(KeyFrameView* kf1, KeyFrameView* kf2, const int x, const int y, float /*pixel*/,
                                        float min_depth, float max_depth, depthHo* dh, const float* /*F12*/, float& best_u,
                                        float& best_v, float /*th_pi*/, float rot) {
    ensure_uploaded(kf1);
    ensure_uploaded(kf2);
    const int xy[2] = {x, y};
    float out[6] = {0, 0, -1, 0, 0, 0};
    check(sd_probe_epipolar(ctx_, kf1->mnId, kf2->mnId, xy, 1, min_depth, max_depth, rot, out), "sd_probe_epipolar");
    dh->supported = out[3] != 0.0f;
    if (dh->supported) {
        dh->depth = out[0];
        dh->sigma = out[1];
        best_u = out[4];  // PM.cc:459-460
        best_v = out[5];
    }
}

void ProbabilityMapping::GetSearchRange(float& umin, float& umax, int px, int py, float mind, float maxd, KeyFrameView* kf,
                                        KeyFrameView* kf2) {
    const int cols = kf->im_.cols;
    check(sd_search_range(kf->K, kf->Tcw, kf2->Tcw, cols, px, py, mind, maxd, &umin, &umax), "sd_search_range");
}

void ProbabilityMapping::InverseDepthHypothesisFusion(const std::vector<depthHo>& h, depthHo& dist) {
    dist.depth = 0;
    dist.sigma = 0;
    dist.supported = false;
    std::vector<float> d(h.size()), s(h.size());
    for (size_t i = 0; i < h.size(); ++i) {
        d[i] = h[i].depth;
        s[i] = h[i].sigma;
    }
    int sup = 0;
    float od = 0, os = 0;
    check(sd_probe_fusion(ctx_, d.data(), s.data(), (int)h.size(), &od, &os, &sup), "sd_probe_fusion");
    if (sup) {
        dist.depth = od;
        dist.sigma = os;
        dist.supported = true;
    }
}

void ProbabilityMapping::IntraKeyFrameDepthChecking(PlaneView<float>& depth_map, PlaneView<float>& depth_sigma,
                                                    const PlaneView<const float> gradimg) {
    if (depth_map.stride != depth_map.cols || depth_sigma.stride != depth_sigma.cols || gradimg.stride != gradimg.cols)
        throw std::invalid_argument("IntraKeyFrameDepthChecking: planes must be dense");
    check(sd_intra_depth_check_planes(ctx_, depth_map.data, depth_sigma.data, gradimg.data, depth_map.cols, depth_map.rows),
          "sd_intra_depth_check_planes");
}

void ProbabilityMapping::IntraKeyFrameDepthGrowing(PlaneView<float>& depth_map, PlaneView<float>& depth_sigma,
                                                   const PlaneView<const float> gradimg) {
    if (depth_map.stride != depth_map.cols || depth_sigma.stride != depth_sigma.cols || gradimg.stride != gradimg.cols)
        throw std::invalid_argument("IntraKeyFrameDepthGrowing: planes must be dense");
    check(sd_intra_depth_grow_planes(ctx_, depth_map.data, depth_sigma.data, gradimg.data, depth_map.cols, depth_map.rows),
          "sd_intra_depth_grow_planes");
}

void ProbabilityMapping::InterKeyFrameDepthChecking(KeyFrameView* currentKf, std::vector<KeyFrameView*> neighbors) {
    std::vector<int> ids;
    for (KeyFrameView* n : neighbors) ids.push_back(n->mnId);
    check(sd_interkf_check(ctx_, currentKf->mnId, ids.data(), (int)ids.size()), "sd_interkf_check");
}

void ProbabilityMapping::UpdateSemiDensePointSet(KeyFrameView* kf) {
    if (kf->poseChanged) check(sd_set_pose(ctx_, kf->mnId, kf->Tcw), "sd_set_pose");
    check(sd_update_pointset(ctx_, kf->mnId), "sd_update_pointset");
    kf->poseChanged = false;  // PM.cc:331
}

void ProbabilityMapping::UpdateAllSemiDensePointSet() {
    std::vector<KeyFrameView*> vpKFs = mpMap->GetAllKeyFrames();
    if (vpKFs.size() < 10) return;  // PM.cc:323
    for (size_t i = 0; i < vpKFs.size(); i++) {
        KeyFrameView* kf = vpKFs[i];
        if (kf->isBad() || !kf->interKF_depth_flag_) continue;
        if (kf->poseChanged) {
            UpdateSemiDensePointSet(kf);
            FetchResults(kf, true);
        }
    }
}

void ProbabilityMapping::FetchResults(KeyFrameView* kf, bool points) {
    const size_t px = (size_t)kf->im_.rows * kf->im_.cols;
    kf->depth_map_.resize(px);
    kf->depth_sigma_.resize(px);
    if (points) kf->SemiDensePointSets_.resize(3 * px);
    check(sd_download(ctx_, kf->mnId, kf->depth_map_.data(), kf->depth_sigma_.data(),
                      points ? kf->SemiDensePointSets_.data() : nullptr),
          "sd_download");
}

int ProbabilityMapping::ConfidentPoints(KeyFrameView* kf, std::vector<Point3f>& out, double max_sigma, int max_points) {
    if (kf->poseChanged) {  // PM.cc:329-331: points follow the current pose
        check(sd_set_pose(ctx_, kf->mnId, kf->Tcw), "sd_set_pose");
        kf->poseChanged = false;
    }
    const int npix = kf->im_.rows * kf->im_.cols;
    int cap = (max_points > 0 && max_points < npix) ? max_points : 0, n = 0;
    static_assert(sizeof(Point3f) == 3 * sizeof(float), "Point3f must be three packed floats");
    const size_t base = out.size();
    if (cap == 0) {  // size the buffer to the selection: count first (device-resident result), then copy
        check(sd_confident_points(ctx_, kf->mnId, max_sigma, max_points, 0, nullptr, nullptr, nullptr, &cap), "sd_confident_points");
        if (cap == 0) return 0;
    }
    out.resize(base + (size_t)cap);
    check(sd_confident_points(ctx_, kf->mnId, max_sigma, max_points, cap, &out[base].x, nullptr, nullptr, &n), "sd_confident_points");
    if (n < cap) out.resize(base + (size_t)n);
    return n;
}

bool ProbabilityMapping::SavePointCloudObj(const std::string& path, double max_sigma) {
    std::FILE* f = std::fopen(path.c_str(), "w");
    if (!f) {
        std::fprintf(stderr, "Failed to save points on line\n");  // PM.cc:103
        return false;
    }
    std::vector<Point3f> pts;
    for (KeyFrameView* kf : mpMap->GetAllKeyFrames()) {
        if (kf->isBad()) continue;
        if (!kf->semidense_flag_) continue;
        if (!kf->interKF_depth_flag_) continue;
        pts.clear();
        ConfidentPoints(kf, pts, max_sigma, 0);  // PM.cc:114-127 on the device
        for (const Point3f& P : pts) std::fprintf(f, "v %f %f %f\n", P.x, P.y, P.z);  // std::to_string(float) == "%f"
    }
    std::fclose(f);
    return true;
}

}  // namespace sdpm_host
