// Host-side mirror of the reference's ProbabilityMapping class: every method is a thin forwarder to
// the C ABI of include/sdpm.h (CUDA kernels in libsdpm.so).  See ProbabilityMapping.h.
// Reference: /root/reference/src/Modeler/ProbabilityMapping.cc ("PM.cc").
#include "ProbabilityMapping.h"

#include <unistd.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <unordered_set>

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

// ---- the worker thread, PM.cc:65-134 (finish handshake as in Modeler.cc:1441-1463) ----

void ProbabilityMapping::Run() {
    {
        std::unique_lock<std::mutex> lock(mMutexFinish);
        mbFinished = false;
    }
    while (1) {
        if (CheckFinish()) break;  // PM.cc:69
        SemiDenseLoop();           // PM.cc:76
        UpdateAllSemiDensePointSet();  // PM.cc:79: make point positions follow the key-frame poses
        usleep(5000);              // PM.cc:86
    }
    if (!obj_path_.empty()) SavePointCloudObj(obj_path_);  // PM.cc:99-130
    SetFinish();  // PM.cc:134
}

void ProbabilityMapping::RunOnce() {
    SemiDenseLoop();
    UpdateAllSemiDensePointSet();
}

void ProbabilityMapping::RequestFinish() {
    std::unique_lock<std::mutex> lock(mMutexFinish);
    mbFinishRequested = true;
}

bool ProbabilityMapping::CheckFinish() {
    std::unique_lock<std::mutex> lock(mMutexFinish);
    return mbFinishRequested;
}

void ProbabilityMapping::SetFinish() {
    std::unique_lock<std::mutex> lock(mMutexFinish);
    mbFinished = true;
}

bool ProbabilityMapping::isFinished() {
    std::unique_lock<std::mutex> lock(mMutexFinish);
    return mbFinished;
}

// ---- residency ----

// The reference keeps im_/GradImg/GradTheta/I_stddev in the KeyFrame; here they live in the
// context, keyed by mnId, created on first use (image upload + gradient front end, kernel K0).
// The reference reads GetRotation()/GetTranslation() at call time (PM.cc:638-644, 854-860, 973-979):
// the pose the device holds is refreshed whenever the view's has moved.
ProbabilityMapping::KfState& ProbabilityMapping::make_resident(KeyFrameView* kf) {
    float Tcw[12];
    kf->GetPose(Tcw);  // one copy under mMutexPose: SetPose on another thread cannot hand a torn pose to the device
    auto it = resident_.find(kf->mnId);
    if (it == resident_.end()) {
        if (kf->im_.empty()) throw std::invalid_argument("ProbabilityMapping: key frame without image");
        check(sd_upload_keyframe(ctx_, kf->mnId, kf->im_.data, kf->im_.cols, kf->im_.rows, kf->im_.stride, kf->K, Tcw),
              "sd_upload_keyframe");
        KfState st;
        std::memcpy(st.Tcw, Tcw, sizeof(st.Tcw));
        it = resident_.emplace(kf->mnId, st).first;
        const size_t px = (size_t)kf->im_.rows * kf->im_.cols;
        if (kf->semidense_flag_ && kf->depth_map_.size() == px && kf->depth_sigma_.size() == px) {
            // released earlier: its maps come back from the host members (checked depth once the inter-KF stage ran,
            // which is what the reference's in-place depth_map_ holds)
            check(sd_import_maps_host(ctx_, kf->mnId, kf->depth_map_.data(), kf->depth_sigma_.data()), "sd_import_maps_host");
            it->second.imported = true;
        }
    } else if (std::memcmp(it->second.Tcw, Tcw, sizeof(it->second.Tcw)) != 0) {
        check(sd_set_pose(ctx_, kf->mnId, Tcw), "sd_set_pose");
        std::memcpy(it->second.Tcw, Tcw, sizeof(it->second.Tcw));
    }
    return it->second;
}

void ProbabilityMapping::ReleaseKeyFrame(KeyFrameView* kf) {
    auto it = resident_.find(kf->mnId);
    if (it == resident_.end()) return;
    check(sd_release_keyframe(ctx_, kf->mnId), "sd_release_keyframe");
    resident_.erase(it);
}

void ProbabilityMapping::enforce_resident_limit(const std::vector<KeyFrameView*>& vpKFs) {
    for (KeyFrameView* kf : vpKFs)
        if (kf->isBad()) ReleaseKeyFrame(kf);
    if (resident_limit_ == 0 || resident_.size() <= resident_limit_) return;
    // key frames some unfinished key frame still needs: itself and its covisN picks
    std::unordered_set<int> needed;
    for (KeyFrameView* kf : vpKFs) {
        if (kf->isBad() || kf->interKF_depth_flag_) continue;
        needed.insert(kf->mnId);
        for (KeyFrameView* nb : pick_neighbours(kf)) needed.insert(nb->mnId);
        // neighbours that are not mapped yet will be picked later; keep what the full covisibility list names as well
        for (size_t i = 0; i < kf->covisible.size() && (int)i < 2 * covis_n_; ++i) needed.insert(kf->covisible[i]->mnId);
    }
    std::vector<KeyFrameView*> cand;
    for (KeyFrameView* kf : vpKFs)
        if (kf->interKF_depth_flag_ && resident_.count(kf->mnId) && !needed.count(kf->mnId)) cand.push_back(kf);
    std::sort(cand.begin(), cand.end(), [](const KeyFrameView* a, const KeyFrameView* b) { return a->mnId < b->mnId; });
    for (KeyFrameView* kf : cand) {
        if (resident_.size() <= resident_limit_) break;
        ReleaseKeyFrame(kf);
    }
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

void ProbabilityMapping::SemiDenseLoop() {
    std::vector<KeyFrameView*> vpKFs = mpMap->GetAllKeyFrames();
    changed_.clear();
    // the images of the key frames that will be reconstructed in this pass first, so that the upload stream runs ahead
    // of the compute stream
    for (KeyFrameView* kf : vpKFs)
        if (!kf->isBad() && !kf->semidense_flag_ && (int)pick_neighbours(kf).size() >= covis_n_) make_resident(kf);
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
        // InterKeyFrameDepthChecking + UpdateSemiDensePointSet (PM.cc:302-304): one fused launch
        std::vector<int> ids;
        kf->poseChanged = false;  // cleared BEFORE the pose is read: a SetPose that lands later raises it again (next pass)
        make_resident(kf);
        for (KeyFrameView* n : neighbors) {
            make_resident(n);
            ids.push_back(n->mnId);
        }
        check(sd_interkf_check_points(ctx_, kf->mnId, ids.data(), (int)ids.size()), "sd_interkf_check_points");
        kf->interKF_depth_flag_ = true;
        if (std::find(changed_.begin(), changed_.end(), kf) == changed_.end()) changed_.push_back(kf);
    }
    // results back to the KeyFrame members, as the reference leaves them (PM.cc:225-226, 306, 361-363): only what
    // this pass produced
    for (KeyFrameView* kf : changed_)
        if (!kf->isBad()) FetchResults(kf, kf->interKF_depth_flag_);
    changed_.clear();
    enforce_resident_limit(vpKFs);
    max_resident_seen_ = std::max(max_resident_seen_, resident_.size());
    passes_++;
}

void ProbabilityMapping::SemiDenseRecon(KeyFrameView* kf) {
    if (kf->isBad() || kf->semidense_flag_) return;  // PM.cc:141-142
    std::vector<KeyFrameView*> closestMatches = pick_neighbours(kf);
    if ((int)closestMatches.size() < covis_n_) return;  // PM.cc:160

    std::vector<int> ids;
    std::vector<float> rots;
    for (KeyFrameView* kf2 : closestMatches) {  // PM.cc:170-179
        make_resident(kf2);
        ids.push_back(kf2->mnId);
        rots.push_back(MedianRotInPlane(kf, kf2));
    }
    float max_depth, min_depth;
    StereoSearchConstraints(kf, &min_depth, &max_depth);  // PM.cc:184
    make_resident(kf).imported = false;
    // PM.cc:189-238: ComputeFundamental per neighbour, the pixel loop, the two intra-key-frame passes
    check(sd_semidense(ctx_, kf->mnId, ids.data(), rots.data(), (int)ids.size(), min_depth, max_depth), "sd_semidense");
    kf->semidense_flag_ = true;  // PM.cc:244
    changed_.push_back(kf);
}

void ProbabilityMapping::StereoSearchConstraints(KeyFrameView* kf, float* min_depth, float* max_depth) {
    check(sd_stereo_search_constraints(kf->point_depths.data(), (int)kf->point_depths.size(), min_depth, max_depth),
          "sd_stereo_search_constraints");
}

void ProbabilityMapping::ComputeFundamental(KeyFrameView* pKF1, KeyFrameView* pKF2, float F12[9]) {
    float R21[9], t21[3], T1[12], T2[12];
    pKF1->GetPose(T1);
    pKF2->GetPose(T2);
    check(sd_pair_geometry(pKF1->K, T1, pKF2->K, T2, F12, R21, t21), "sd_pair_geometry");
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
    make_resident(kf1);
    make_resident(kf2);
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
    float T1[12], T2[12];
    kf->GetPose(T1);
    kf2->GetPose(T2);
    check(sd_search_range(kf->K, T1, T2, cols, px, py, mind, maxd, &umin, &umax), "sd_search_range");
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
    make_resident(currentKf);
    for (KeyFrameView* n : neighbors) {
        make_resident(n);  // poses as they are now (PM.cc:638-644)
        ids.push_back(n->mnId);
    }
    check(sd_interkf_check(ctx_, currentKf->mnId, ids.data(), (int)ids.size()), "sd_interkf_check");
}

void ProbabilityMapping::UpdateSemiDensePointSet(KeyFrameView* kf) {
    kf->poseChanged = false;  // PM.cc:331; before the pose is read, so that a concurrent SetPose is not lost
    make_resident(kf);  // uploads / re-installs if needed and pushes the current pose (PM.cc:340: GetPoseInverse at call time)
    check(sd_update_pointset(ctx_, kf->mnId), "sd_update_pointset");
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
    auto st = resident_.find(kf->mnId);
    if (st == resident_.end()) throw std::runtime_error("FetchResults: key frame is not resident on the device");
    if (st->second.imported) {  // maps re-installed from the host: no compact form, dense planes
        kf->depth_map_.resize(px);
        kf->depth_sigma_.resize(px);
        if (points) kf->SemiDensePointSets_.resize(3 * px);
        check(sd_download(ctx_, kf->mnId, kf->depth_map_.data(), kf->depth_sigma_.data(),
                          points ? kf->SemiDensePointSets_.data() : nullptr),
              "sd_download");
        filled_.erase(kf->mnId);
        return;
    }
    // compact records over PCIe (about a quarter of the pixels carry anything), expanded into the dense members here
    const int cap = (kf->im_.cols - 4) * (kf->im_.rows - 4);
    if ((int)records_.size() < cap) records_.resize(cap);
    int n = 0;
    check(sd_download_compact(ctx_, kf->mnId, points ? 1 : 0, cap, records_.data(), &n), "sd_download_compact");
    std::vector<uint32_t>& filled = filled_[kf->mnId];
    const int W = kf->im_.cols;
    const bool fresh_maps = kf->depth_map_.size() != px || kf->depth_sigma_.size() != px;
    if (fresh_maps) {
        kf->depth_map_.assign(px, 0.0f);
        kf->depth_sigma_.assign(px, 0.0f);
    }
    const bool fresh_points = points && kf->SemiDensePointSets_.size() != 3 * px;
    if (fresh_points) kf->SemiDensePointSets_.assign(3 * px, 0.0f);
    const bool have_points = kf->SemiDensePointSets_.size() == 3 * px;
    if (!fresh_maps || (have_points && !fresh_points)) {
        for (uint32_t e : filled) {  // what the previous fetch wrote
            const size_t o = (size_t)(e >> 16) * W + (e & 0xffffu);
            kf->depth_map_[o] = 0.0f;
            kf->depth_sigma_[o] = 0.0f;
            if (have_points) { kf->SemiDensePointSets_[3 * o] = 0.0f; kf->SemiDensePointSets_[3 * o + 1] = 0.0f; kf->SemiDensePointSets_[3 * o + 2] = 0.0f; }
        }
    }
    filled.resize(n);
    if (points) {
        const SdCompactPoint* r = records_.data();
        for (int i = 0; i < n; ++i) {
            const size_t o = (size_t)(r[i].pixel >> 16) * W + (r[i].pixel & 0xffffu);
            kf->depth_map_[o] = r[i].depth;
            kf->depth_sigma_[o] = r[i].sigma;
            kf->SemiDensePointSets_[3 * o] = r[i].x; kf->SemiDensePointSets_[3 * o + 1] = r[i].y; kf->SemiDensePointSets_[3 * o + 2] = r[i].z;
            filled[i] = r[i].pixel;
        }
    } else {
        const SdCompactDepth* r = reinterpret_cast<const SdCompactDepth*>(records_.data());
        for (int i = 0; i < n; ++i) {
            const size_t o = (size_t)(r[i].pixel >> 16) * W + (r[i].pixel & 0xffffu);
            kf->depth_map_[o] = r[i].depth;
            kf->depth_sigma_[o] = r[i].sigma;
            filled[i] = r[i].pixel;
        }
    }
}

int ProbabilityMapping::ConfidentPoints(KeyFrameView* kf, std::vector<Point3f>& out, double max_sigma, int max_points) {
    kf->poseChanged = false;
    make_resident(kf);  // PM.cc:329-331: points follow the current pose
    static_assert(sizeof(Point3f) == 3 * sizeof(float), "Point3f must be three packed floats");
    // the selection stays on the device; its size decides the host buffer, then one copy
    const float *dx = nullptr, *ds = nullptr;
    const uint32_t* dp = nullptr;
    int n = 0;
    check(sd_confident_points_device(ctx_, kf->mnId, max_sigma, max_points, &dx, &ds, &dp, &n, nullptr), "sd_confident_points_device");
    if (n == 0) return 0;
    const size_t base = out.size();
    out.resize(base + (size_t)n);
    int got = 0;
    check(sd_confident_points_fetch(ctx_, n, &out[base].x, nullptr, nullptr, &got), "sd_confident_points_fetch");
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
