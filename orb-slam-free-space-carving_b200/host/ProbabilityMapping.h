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
// computes in its pixel loops (the one host loop is the expansion of the compact download records
// into the dense depth_map_ / depth_sigma_ / SemiDensePointSets_ members: a scatter of values the
// GPU produced).  The per-pixel methods (EpipolarSearch, GetSearchRange,
// InverseDepthHypothesisFusion) are single-pixel probes that run the same device code as the
// batch kernels.  Error convention: the reference's methods are `void` and return silently
// (PM.cc:141-142, 160); here a failing C-ABI call throws std::runtime_error with sd_last_error().
//
// Threading: Run() is the reference's polling worker (PM.cc:65-134; the pattern of
// Modeler::Run, Modeler.cc:51-98, started like System.cc:110-111) and is meant to own the object
// from its own thread; RequestFinish() / isFinished(), Map::AddKeyFrame() and KeyFrameView::SetPose() (bundle adjustment,
// loop closing; the pose is read under the key frame's mMutexPose, one consistent copy per use) are the only calls
// another thread may make while it runs.
#ifndef SDPM_HOST_PROBABILITYMAPPING_H
#define SDPM_HOST_PROBABILITYMAPPING_H

#include <atomic>
#include <cstdint>
#include <exception>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
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
    // GetRotation / GetTranslation, PM.cc:885-888.  Guarded by mMutexPose like KeyFrame::Tcw (KeyFrame.cc:70-119): once
    // the worker runs, write it through SetPose and read it through GetPose only.
    float Tcw[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    mutable std::mutex mMutexPose;
    std::atomic<bool> bad{false};     // isBad(); other SLAM threads flip these while the worker runs
    std::atomic<bool> mapped{true};   // Mapped(), PM.cc:156
    std::vector<KeyFrameView*> covisible;  // GetVectorCovisibleKeyFrames(), best first, PM.cc:151
    std::vector<float> point_depths;       // GetAllPointDepths(), PM.cc:372
    std::vector<long long> map_points;     // GetMapPointMatches() as ids, < 0 = none, PM.cc:469-470
    std::vector<float> keypoint_angles;    // GetKeyPointsUn()[i].angle, PM.cc:477-478
    // results (host copies, W*H floats; point set H x 3W)
    std::vector<float> depth_map_, depth_sigma_, SemiDensePointSets_;
    bool semidense_flag_ = false, interKF_depth_flag_ = false;
    std::atomic<bool> poseChanged{false};

    bool isBad() const { return bad.load(); }
    bool Mapped() const { return mapped.load(); }
    void SetPose(const float Tcw_new[12]) {  // KeyFrame::SetPose (KeyFrame.cc:70-84) + poseChanged (PM.cc:329)
        {
            std::unique_lock<std::mutex> lock(mMutexPose);
            for (int i = 0; i < 12; ++i) Tcw[i] = Tcw_new[i];
        }
        poseChanged = true;
    }
    void GetPose(float out[12]) const {  // GetRotation + GetTranslation as one consistent copy (KeyFrame.cc:109-119)
        std::unique_lock<std::mutex> lock(mMutexPose);
        for (int i = 0; i < 12; ++i) out[i] = Tcw[i];
    }
};

struct Map {  // ORB_SLAM2::Map: AddKeyFrame / GetAllKeyFrames under mMutexMap (Map.cc:36-40, 97-101), PM.cc:106, 322
    std::vector<KeyFrameView*> keyframes;
    mutable std::mutex mMutexMap;
    void AddKeyFrame(KeyFrameView* kf) {
        std::unique_lock<std::mutex> lock(mMutexMap);
        keyframes.push_back(kf);
    }
    std::vector<KeyFrameView*> GetAllKeyFrames() const {
        std::unique_lock<std::mutex> lock(mMutexMap);
        return keyframes;
    }
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

    // PM.cc:65-134: the worker.  while (!CheckFinish()) { SemiDenseLoop(); UpdateAllSemiDensePointSet(); usleep(5000); }
    // then the .obj dump (PM.cc:99-130, skipped when the path is empty) and mbFinished = true.
    void Run();
    void RequestFinish();  // the pattern of Modeler.cc:1441-1463
    bool isFinished();
    // One iteration of the body of Run(): SemiDenseLoop + UpdateAllSemiDensePointSet.
    void RunOnce();
    void SetPointCloudPath(const std::string& path) { obj_path_ = path; }  // default "semi_pointcloud.obj", PM.cc:99

    // PM.cc:137-318 over every key frame of the map: first half per key frame (SemiDenseRecon),
    // then the inter-key-frame pass (PM.cc:262-313).  Only key frames whose results changed in this
    // pass are copied back to their host members.
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

    // copy device results of a key frame into its host members (depth_map_, depth_sigma_, SemiDensePointSets_):
    // compact records over PCIe (sd_download_compact), expanded here
    void FetchResults(KeyFrameView* kf, bool points);
    // Device memory of a live map: at most `n` key frames stay resident (0 = no limit).  After each pass the finished
    // key frames (interKF_depth_flag_) that no unfinished key frame still lists among its covisN neighbours are released,
    // oldest first, until the limit holds; bad key frames are released at once.  A released key frame that is needed
    // again (a later key frame lists it, or its pose changes, PM.cc:321-334) is re-installed from its host members --
    // then with the CHECKED inverse depth, which is what the reference's in-place depth_map_ holds by that time.
    void SetResidentLimit(size_t n) { resident_limit_ = n; }
    // SemiDenseLoop issues a pass in windows of at most `n` key frames to reconstruct: one sd_semidense_loop call per
    // window with the compact records of every key frame it completes downloaded asynchronously into page-locked
    // buffers (three sets of n: one being filled by the GPU, one being expanded, one in between), expanded into the KeyFrame members by `threads` host threads while the GPU works on the
    // next window.  n bounds the page-locked memory (n * W*H/2 * 24 bytes per set); 0 threads = hardware concurrency,
    // at most 16.  With SdParams.interkf_inplace a pass is one window (the checks are order-dependent).
    void SetBatchWindow(size_t n, unsigned threads = 0) { window_ = n ? n : 1; expand_threads_ = threads; }
    // Host wall-clock seconds of the last SemiDenseLoop pass by phase (diagnostics of the e2e_class bench leg).
    struct PassStats { double resident = 0, prepare = 0, issue = 0, expand = 0, sync = 0, total = 0; int windows = 0; };
    const PassStats& LastPassStats() const { return stats_; }
    void ReleaseKeyFrame(KeyFrameView* kf);
    size_t ResidentKeyFrames() const { return resident_.size(); }
    size_t MaxResidentKeyFrames() const { return max_resident_seen_; }  // high-water mark after the release step of a pass
    uint64_t Passes() const { return passes_.load(); }                  // completed SemiDenseLoop passes
    sd_ctx* context() const { return ctx_; }

private:
    struct KfState {
        float Tcw[12];                 // the pose the device holds
        bool imported = false;         // maps re-installed from the host members (no compact form)
    };
    struct Fetch {                     // one asynchronous compact download of a window
        KeyFrameView* kf = nullptr;
        bool points = false;
        int slot = 0;                  // buffer of the set
        bool sync_fallback = false;    // imported maps (no compact form): fetched the slow way after the window
    };
    struct BufferSet {
        void* base = nullptr;          // page-locked, slots * record_cap * sizeof(SdCompactPoint)
        int* counts = nullptr;         // page-locked, one per slot
        size_t slots = 0;
        int record_cap = 0;
    };
    bool CheckFinish();
    void SetFinish();
    void ensure_buffers(BufferSet& b, size_t slots, int record_cap);
    void expand_window(const BufferSet& b, const std::vector<Fetch>& fetches);
    void expand_records(KeyFrameView* kf, const void* records, int n, bool points, std::vector<uint32_t>& filled);
    static void expand_streaming(KeyFrameView* kf, const void* records, int n, bool points, std::vector<uint32_t>& filled);
    void join_expander();
    KfState& make_resident(KeyFrameView* kf);  // upload on first use, push the pose if it moved, re-install released maps
    std::vector<KeyFrameView*> pick_neighbours(KeyFrameView* kf) const;  // PM.cc:151-160, 278-287
    void enforce_resident_limit(const std::vector<KeyFrameView*>& vpKFs);
    void check(int rc, const char* what) const;

    Map* mpMap;
    sd_ctx* ctx_;
    int covis_n_;
    std::unordered_map<int, KfState> resident_;       // key frames with planes in the context, by mnId
    std::unordered_map<int, std::vector<uint32_t>> filled_;  // by mnId: pixels of the host members the last FetchResults wrote
    std::vector<KeyFrameView*> changed_;              // results to copy back at the end of the pass
    std::vector<SdCompactPoint> records_;             // download scratch
    size_t resident_limit_ = 0, max_resident_seen_ = 0;
    size_t window_ = 32;
    unsigned expand_threads_ = 0;
    bool inplace_ = false;                            // SdParams.interkf_inplace of the context
    BufferSet bufs_[3];
    PassStats stats_;
    std::thread expander_;                            // expands the previous window's records while this thread issues the next
    std::exception_ptr expander_error_;
    std::atomic<uint64_t> passes_{0};
    std::string obj_path_ = "semi_pointcloud.obj";
    std::mutex mMutexFinish;
    bool mbFinishRequested = false, mbFinished = false;
};

}  // namespace sdpm_host
#endif
