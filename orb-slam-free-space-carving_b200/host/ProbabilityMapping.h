// Host-side mirror of the reference's ProbabilityMapping class
// (/root/reference/include/Modeler/ProbabilityMapping.h:61-105, "PM.h"; code in
// /root/reference/src/Modeler/ProbabilityMapping.cc, "PM.cc") on top of the C ABI of
// include/sdpm.h.  Same class name, same method names, same argument meaning; the two types the
// reference takes from elsewhere are replaced by plain data:
//
//   ORB_SLAM2::KeyFrame*  ->  KeyFrameView*   (the reference's KeyFrame has none of the members
//                                              PM.cc reads -- SURVEY.md F3 -- so the view declares them)
//   cv::Mat               ->  PlaneView<T>    (OpenCV C++ is not part of this build)
//
// Every method forwards to libsdpm.so; nothing here computes on the CPU what the reference
// computes in its pixel loops.  The per-pixel methods (EpipolarSearch, GetSearchRange,
// InverseDepthHypothesisFusion) are single-pixel probes that run the same device code as the
// batch kernels.  Error convention: the reference's methods are `void` and return silently
// (PM.cc:141-142, 160); here a failing C-ABI call throws std::runtime_error with sd_last_error().
#ifndef SDPM_HOST_PROBABILITYMAPPING_H
#define SDPM_HOST_PROBABILITYMAPPING_H

#include <cstdint>
#include <string>
#include <vector>

#include "sdpm.h"

// PM.h:38-49 (the ones the path reads; run-time copies live in SdParams)
#define covisN 7
#define lambdaG_DEFAULT 8
#define lambdaL_DEFAULT 80
#define lambdaTheta_DEFAULT 45
#define lambdaN_DEFAULT 3
#define THETA_DEFAULT 0.23

namespace sdpm_host {

// cv::Mat stand-in: `stride` in elements
template <typename T>
struct PlaneView {
    T* data = nullptr;
    int rows = 0, cols = 0, stride = 0;
    PlaneView() = default;
    PlaneView(T* d, int r, int c, int s = 0) : data(d), rows(r), cols(c), stride(s ? s : c) {}
    T& at(int y, int x) const { return data[(size_t)y * stride + x]; }
    bool empty() const { return !data || rows <= 0 || cols <= 0; }
};

struct Point3f {  // cv::Point3f stand-in
    float x, y, z;
};

// What PM.cc reads from / writes to ORB_SLAM2::KeyFrame.
struct KeyFrameView {
    int mnId = -1;
    PlaneView<const uint8_t> im_;  // kf->im_ / GetImage(), PM.cc:186
    float K[4] = {0, 0, 0, 0};     // fx, fy, cx, cy  (kf->fx ..., PM.cc:862-865; GetCalibrationMatrix, PM.cc:675)
    float Tcw[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};  // GetRotation / GetTranslation, PM.cc:885-888
    bool bad = false;              // isBad()
    bool mapped = true;            // Mapped(), PM.cc:156
    std::vector<KeyFrameView*> covisible;  // GetVectorCovisibleKeyFrames(), best first, PM.cc:151
    std::vector<float> point_depths;       // GetAllPointDepths(), PM.cc:372
    std::vector<long long> map_points;     // GetMapPointMatches() as ids, < 0 = none, PM.cc:469-470
    std::vector<float> keypoint_angles;    // GetKeyPointsUn()[i].angle, PM.cc:477-478
    // results (host copies, W*H floats; point set H x 3W)
    std::vector<float> depth_map_, depth_sigma_, SemiDensePointSets_;
    bool semidense_flag_ = false, interKF_depth_flag_ = false, poseChanged = false;

    bool isBad() const { return bad; }
    bool Mapped() const { return mapped; }
    void SetPose(const float Tcw_new[12]) {  // KeyFrame::SetPose + poseChanged (PM.cc:329)
        for (int i = 0; i < 12; ++i) Tcw[i] = Tcw_new[i];
        poseChanged = true;
    }
};

struct Map {  // ORB_SLAM2::Map::GetAllKeyFrames(), PM.cc:106, 322
    std::vector<KeyFrameView*> keyframes;
    std::vector<KeyFrameView*> GetAllKeyFrames() const { return keyframes; }
};

class ProbabilityMapping {
public:
    struct depthHo {  // PM.h:64-70; `depth` is INVERSE depth
        depthHo() : depth(0.0f), sigma(0.0f), supported(false) { Pw[0] = Pw[1] = Pw[2] = 0.0f; }
        float depth;
        float sigma;
        bool supported;
        float Pw[3];
    };

    // PM.cc:61.  `device` = CUDA device of the context; `covis_n` = covisN (PM.h:38), a run-time
    // value here because BASELINE.json's configurations use 6..40 neighbours.
    explicit ProbabilityMapping(Map* pMap, int device = 0, int covis_n = covisN, const SdParams* params = nullptr);
    ~ProbabilityMapping();
    ProbabilityMapping(const ProbabilityMapping&) = delete;
    ProbabilityMapping& operator=(const ProbabilityMapping&) = delete;

    // One iteration of the body of Run(), PM.cc:65-87: SemiDenseLoop + UpdateAllSemiDensePointSet.
    void Run();
    // PM.cc:137-318 over every key frame of the map: first half per key frame (SemiDenseRecon),
    // then the inter-key-frame pass (PM.cc:262-313).
    void SemiDenseLoop();
    // PM.h:75 / PM.cc:137-257: one key frame, neighbour pick .. IntraKeyFrameDepthGrowing.
    void SemiDenseRecon(KeyFrameView* kf);
    void StereoSearchConstraints(KeyFrameView* kf, float* min_depth, float* max_depth);
    void EpipolarSearch(KeyFrameView* kf1, KeyFrameView* kf2, const int x, const int y, float pixel, float min_depth,
                        float max_depth, depthHo* dh, const float F12[9], float& best_u, float& best_v, float th_pi,
                        float rot);
    void GetSearchRange(float& umin, float& umax, int px, int py, float mind, float maxd, KeyFrameView* kf, KeyFrameView* kf2);
    void InverseDepthHypothesisFusion(const std::vector<depthHo>& h, depthHo& dist);
    void IntraKeyFrameDepthChecking(PlaneView<float>& depth_map, PlaneView<float>& depth_sigma, const PlaneView<const float> gradimg);
    void IntraKeyFrameDepthGrowing(PlaneView<float>& depth_map, PlaneView<float>& depth_sigma, const PlaneView<const float> gradimg);
    void UpdateSemiDensePointSet(KeyFrameView* kf);
    void UpdateAllSemiDensePointSet();
    void InterKeyFrameDepthChecking(KeyFrameView* currentKf, std::vector<KeyFrameView*> neighbors);

    // PM.cc:100-132: "v x y z" for every pixel with sigma <= 0.01 (PM.cc:120) and inverse depth > 1e-6.
    // The selection runs on the device (sd_confident_points); only the selected points cross PCIe.
    bool SavePointCloudObj(const std::string& path, double max_sigma = 0.01);
    // The same selection for one key frame as the point list the CARV transcript ingests
    // (SFMTranscriptInterface_ORBSLAM::addKeyFrameInsertionWithLinesEntry(k, kCopy, std::vector<cv::Point3f>&),
    // SFMTranscriptInterface_ORBSLAM.cpp:319-374); max_points > 0 keeps the max_points smallest-sigma points.
    // Returns the number of points appended to `out`.
    int ConfidentPoints(KeyFrameView* kf, std::vector<Point3f>& out, double max_sigma = 0.01, int max_points = 0);

    // ComputeFundamental (PM.cc:972-987) and GetRotInPlane + median (PM.cc:467-484, 172-177):
    // private in PM.h, public here so the tests can reach them.
    void ComputeFundamental(KeyFrameView* pKF1, KeyFrameView* pKF2, float F12[9]);
    float MedianRotInPlane(KeyFrameView* kf1, KeyFrameView* kf2);

    // copy device results of a key frame into its host members (depth_map_, depth_sigma_, SemiDensePointSets_)
    void FetchResults(KeyFrameView* kf, bool points);
    sd_ctx* context() const { return ctx_; }

private:
    void ensure_uploaded(KeyFrameView* kf);
    std::vector<KeyFrameView*> pick_neighbours(KeyFrameView* kf) const;  // PM.cc:151-160, 278-287
    void check(int rc, const char* what) const;

    Map* mpMap;
    sd_ctx* ctx_;
    int covis_n_;
    std::vector<int> uploaded_;  // ids resident in the context
};

}  // namespace sdpm_host
#endif
