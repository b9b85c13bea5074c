// Host-side mirror of the reference's ProbabilityMapping class: every method is a thin forwarder to
// the C ABI of include/sdpm.h (CUDA kernels in libsdpm.so).  See ProbabilityMapping.h.
// Reference: /root/reference/src/Modeler/ProbabilityMapping.cc ("PM.cc").
#include "ProbabilityMapping.h"

#include <unistd.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <new>
#include <stdexcept>
#include <thread>
#include <unordered_set>

namespace sdpm_host {

ProbabilityMapping::ProbabilityMapping(Map* pMap, int device, int covis_n, const SdParams* params)
    : mpMap(pMap), ctx_(nullptr), covis_n_(covis_n) {
    if (covis_n < 1 || covis_n > SD_MAX_NEIGHBOURS) throw std::invalid_argument("ProbabilityMapping: covis_n out of range");
    int rc = sd_create(device, params, &ctx_);
    if (rc != SD_OK) throw std::runtime_error("ProbabilityMapping: sd_create failed (no usable CUDA device; there is no CPU fallback)");
    inplace_ = params && params->interkf_inplace != 0;
}

ProbabilityMapping::~ProbabilityMapping() {
    if (expander_.joinable()) expander_.join();
    if (!ctx_) return;
    sd_sync(ctx_);
    for (BufferSet& b : bufs_) {
        sd_free_host(ctx_, b.base);
        sd_free_host(ctx_, b.counts);
    }
    sd_destroy(ctx_);
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

void ProbabilityMapping::ensure_buffers(BufferSet& b, size_t slots, int record_cap) {
    if (b.base && b.slots >= slots && b.record_cap == record_cap) return;
    check(sd_sync(ctx_), "sd_sync");
    check(sd_free_host(ctx_, b.base), "sd_free_host");
    check(sd_free_host(ctx_, b.counts), "sd_free_host");
    b.base = nullptr; b.counts = nullptr;
    void* p = nullptr;
    check(sd_alloc_host(ctx_, (uint64_t)slots * record_cap * sizeof(SdCompactPoint), &p), "sd_alloc_host");
    b.base = p;
    check(sd_alloc_host(ctx_, (uint64_t)slots * sizeof(int), &p), "sd_alloc_host");
    b.counts = static_cast<int*>(p);
    b.slots = slots;
    b.record_cap = record_cap;
}

// The pass of PM.cc:137-318, issued in windows (SetBatchWindow): per window ONE sd_semidense_loop call carries the
// per-key-frame half of every key frame of the window (PM.cc:137-257), the inter-key-frame check + point set
// (PM.cc:262-307) of every key frame -- of this or an earlier window or pass -- whose neighbours all have maps by then, and
// the asynchronous download of what each of them produced.  The records of window k are expanded into the KeyFrame
// members while the GPU works on window k+1.
void ProbabilityMapping::SemiDenseLoop() {
    using clk = std::chrono::steady_clock;
    auto since = [](clk::time_point t) { return std::chrono::duration<double>(clk::now() - t).count(); };
    const clk::time_point t_pass = clk::now();
    stats_ = PassStats();
    std::vector<KeyFrameView*> vpKFs = mpMap->GetAllKeyFrames();
    changed_.clear();
    std::vector<KeyFrameView*> recon;  // PM.cc:141-142, 160
    for (KeyFrameView* kf : vpKFs)
        if (!kf->isBad() && !kf->semidense_flag_ && (int)pick_neighbours(kf).size() >= covis_n_) recon.push_back(kf);
    const size_t window = inplace_ ? std::max<size_t>(recon.size(), 1) : window_;

    struct Pending {
        std::vector<Fetch> fetches;
        int set = 0;
        bool any = false;
        uint64_t ticket = 0;  // sd_download_fence after the window was issued
    } prev;
    size_t pos = 0;
    int turn = 0;
    bool first = true;
    while (first || pos < recon.size()) {
        first = false;
        const size_t end = std::min(recon.size(), pos + window);
        std::unordered_set<int> in_window;
        for (size_t r = pos; r < end; ++r) in_window.insert(recon[r]->mnId);
        // the images of the window first, so that the upload stream runs ahead of the compute stream
        clk::time_point t0 = clk::now();
        for (size_t r = pos; r < end; ++r) make_resident(recon[r]);
        stats_.resident += since(t0);
        t0 = clk::now();
        pos = end;

        struct Entry {
            KeyFrameView* kf;
            std::vector<KeyFrameView*> nb;
            std::vector<int> ids;
            std::vector<float> rots;
            float min_depth = 0, max_depth = 0;
            bool recon = false, inter = false;
        };
        std::vector<Entry> entries;
        for (KeyFrameView* kf : vpKFs) {  // list order = the visiting order of PM.cc:137 and PM.cc:262
            if (kf->isBad() || kf->interKF_depth_flag_) continue;
            const bool in_win = in_window.count(kf->mnId) != 0;
            if (!in_win && !kf->semidense_flag_) continue;
            std::vector<KeyFrameView*> nb = pick_neighbours(kf);
            if ((int)nb.size() < covis_n_) continue;  // PM.cc:160, 287
            bool inter = true;                        // PM.cc:289-297: every neighbour has (or gets in this window) its maps
            for (KeyFrameView* n : nb)
                if (!n->semidense_flag_ && !in_window.count(n->mnId)) inter = false;
            if (!in_win && !inter) continue;
            Entry e;
            e.kf = kf; e.recon = in_win; e.inter = inter; e.nb = nb;
            for (KeyFrameView* n : nb) {
                e.ids.push_back(n->mnId);
                if (in_win) e.rots.push_back(MedianRotInPlane(kf, n));  // PM.cc:170-179
            }
            if (in_win) StereoSearchConstraints(kf, &e.min_depth, &e.max_depth);  // PM.cc:184
            entries.push_back(std::move(e));
        }
        stats_.prepare += since(t0);
        t0 = clk::now();
        // residency and poses of everything the window touches (uploads of neighbours outside the window, pose pushes,
        // re-installation of released key frames)
        for (Entry& e : entries) {
            if (e.inter) e.kf->poseChanged = false;  // cleared BEFORE the pose is read: a later SetPose raises it again
            KfState& st = make_resident(e.kf);
            if (e.recon) st.imported = false;
        }
        for (Entry& e : entries)
            for (KeyFrameView* n : e.nb) make_resident(n);

        stats_.resident += since(t0);
        t0 = clk::now();
        // three buffer sets: this window's set was last read by the expansion of three windows ago, which ended before
        // the expansion of two windows ago started (expand_window joins its predecessor)
        BufferSet& bs = bufs_[turn % 3];
        std::vector<SdLoopItem> items(entries.size());
        std::vector<int> inter_done(entries.size(), 0);
        Pending cur;
        cur.set = turn % 3;
        if (!entries.empty()) {
            const KeyFrameView* k0 = entries[0].kf;
            const int record_cap = std::max(1, (int)(((size_t)k0->im_.rows * k0->im_.cols) / 2));
            ensure_buffers(bs, std::max(entries.size(), window_), record_cap);
        }
        for (size_t i = 0; i < entries.size(); ++i) {
            Entry& e = entries[i];
            SdLoopItem& it = items[i];
            std::memset(&it, 0, sizeof(it));
            it.kf_id = e.kf->mnId;
            it.nbr_ids = e.ids.data();
            it.nbr_rot_deg = e.recon ? e.rots.data() : nullptr;
            it.n = (int)e.ids.size();
            it.min_invdepth = e.min_depth;
            it.max_invdepth = e.max_depth;
            it.finish_only = e.recon ? 0 : 1;
            Fetch f;
            f.kf = e.kf; f.points = e.inter; f.slot = (int)i;
            f.sync_fallback = resident_.at(e.kf->mnId).imported;
            if (f.sync_fallback) {
                it.download = SD_DL_NONE;
            } else {
                it.download = e.inter ? SD_DL_COMPACT_POINTS : SD_DL_COMPACT;
                it.out_records = static_cast<char*>(bs.base) + (size_t)i * bs.record_cap * sizeof(SdCompactPoint);
                it.capacity = bs.record_cap;
                it.out_count = bs.counts + i;
                bs.counts[i] = -1;
            }
            cur.fetches.push_back(f);
        }
        if (!items.empty()) {
            check(sd_semidense_loop(ctx_, items.data(), (int)items.size(), SD_LOOP_INTERKF, inter_done.data()), "sd_semidense_loop");
            check(sd_download_fence(ctx_, &cur.ticket), "sd_download_fence");
            cur.any = true;
            for (size_t i = 0; i < entries.size(); ++i) {
                Entry& e = entries[i];
                if (e.recon) e.kf->semidense_flag_ = true;  // PM.cc:244
                if (e.inter != (inter_done[i] != 0)) {
                    // the device knows better which neighbours have maps (e.g. flags reset by the caller): take its
                    // word and fetch this key frame the direct way
                    cur.fetches[i].sync_fallback = true;
                    cur.fetches[i].points = inter_done[i] != 0;
                }
                if (inter_done[i]) e.kf->interKF_depth_flag_ = true;  // PM.cc:306
            }
        }
        // the previous window's records are in host memory (sd_sync below, one iteration ago): expand them while the GPU
        // works on this window
        stats_.issue += since(t0);
        t0 = clk::now();
        // two windows in flight: this one is queued behind the previous one on the GPU.  Wait for the previous window's
        // downloads only (not for the queue), then hand its records to the expander thread and go on to prepare the next
        // window while the GPU runs this one
        if (prev.any) {
            check(sd_download_wait(ctx_, prev.ticket), "sd_download_wait");
            stats_.sync += since(t0);
            t0 = clk::now();
            expand_window(bufs_[prev.set], prev.fetches);
            stats_.expand += since(t0);
        }
        stats_.windows++;
        prev = std::move(cur);
        ++turn;
    }
    clk::time_point t1 = clk::now();
    check(sd_sync(ctx_), "sd_sync");
    stats_.sync += since(t1);
    t1 = clk::now();
    if (prev.any) expand_window(bufs_[prev.set], prev.fetches);
    join_expander();
    stats_.expand += since(t1);
    enforce_resident_limit(vpKFs);
    max_resident_seen_ = std::max(max_resident_seen_, resident_.size());
    passes_++;
    stats_.total = since(t_pass);
}

void ProbabilityMapping::join_expander() {
    if (expander_.joinable()) expander_.join();
    if (expander_error_) {
        std::exception_ptr e = expander_error_;
        expander_error_ = nullptr;
        std::rethrow_exception(e);
    }
}

// Starts the expansion of one window's records on the expander thread (which fans out over `expand_threads_` workers) and
// returns.  What needs the context -- key frames fetched the direct way -- is done here, on the calling thread, first.
void ProbabilityMapping::expand_window(const BufferSet& b, const std::vector<Fetch>& fetches) {
    join_expander();
    struct Job {
        KeyFrameView* kf;
        const void* rec;
        int n;
        bool points;
        std::vector<uint32_t>* filled;
    };
    std::vector<Job> jobs;
    for (const Fetch& f : fetches) {
        if (f.kf->isBad()) continue;
        if (f.sync_fallback || b.counts[f.slot] < 0 || b.counts[f.slot] > b.record_cap) {
            FetchResults(f.kf, f.points);  // imported maps, or more records than the window's buffers hold: the direct way
            continue;
        }
        const char* rec = static_cast<const char*>(b.base) + (size_t)f.slot * b.record_cap * sizeof(SdCompactPoint);
        jobs.push_back(Job{f.kf, rec, b.counts[f.slot], f.points, &filled_[f.kf->mnId]});  // the map is only touched here
    }
    if (jobs.empty()) return;
    unsigned nt = expand_threads_ ? expand_threads_ : std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
    nt = (unsigned)std::min<size_t>(nt, jobs.size());
    expander_ = std::thread([this, nt, jobs = std::move(jobs)] {
        try {
            std::atomic<size_t> next{0};
            auto body = [&] {
                for (size_t w = next++; w < jobs.size(); w = next++)
                    expand_streaming(jobs[w].kf, jobs[w].rec, jobs[w].n, jobs[w].points, *jobs[w].filled);
            };
            std::vector<std::thread> pool;
            for (unsigned t = 1; t < nt; ++t) pool.emplace_back(body);
            body();
            for (std::thread& t : pool) t.join();
        } catch (...) {
            expander_error_ = std::current_exception();
        }
    });
}

namespace {
// dst <- src with stores that bypass the cache (the planes are written once and not read back here)
void stream_copy(float* dst, const float* src, size_t n) {
#if defined(__SSE2__)
    if (((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15u) == 0) {
        size_t i = 0;
        for (; i + 4 <= n; i += 4) _mm_stream_ps(dst + i, _mm_load_ps(src + i));
        for (; i < n; ++i) dst[i] = src[i];
        return;
    }
#endif
    std::memcpy(dst, src, n * sizeof(float));
}
struct AlignedFloats {
    float* p = nullptr;
    size_t n = 0;
    ~AlignedFloats() { std::free(p); }
    float* get(size_t want) {
        if (want > n) {
            std::free(p);
            p = static_cast<float*>(std::aligned_alloc(64, (want * sizeof(float) + 63) / 64 * 64));
            if (!p) throw std::bad_alloc();
            n = want;
        }
        return p;
    }
};
}  // namespace

// Compact records -> the dense members (PM.cc:225-226, 361-363), written as whole planes in bands of 16 rows: the
// records are bucketed by band (they arrive in no particular order), each band is assembled in a cache-resident staging
// buffer (zeros + the band's records) and streamed to the member once.  Memory traffic is one write of every plane,
// instead of the read-for-ownership + write-back of nearly every cache line, twice (clear, then write), that a scatter
// straight into the members costs.
void ProbabilityMapping::expand_streaming(KeyFrameView* kf, const void* records, int n, bool points, std::vector<uint32_t>& filled) {
    const int W = kf->im_.cols, H = kf->im_.rows;
    const size_t px = (size_t)W * H;
    if (kf->depth_map_.size() != px) kf->depth_map_.assign(px, 0.0f);
    if (kf->depth_sigma_.size() != px) kf->depth_sigma_.assign(px, 0.0f);
    const bool had_points = kf->SemiDensePointSets_.size() == 3 * px;
    if (points && !had_points) kf->SemiDensePointSets_.assign(3 * px, 0.0f);
    // a key frame that has a point set from an earlier pass and gets maps without points now: the stale points of the
    // pixels written last time go, as in expand_records
    const bool clear_points = !points && had_points;
    constexpr int BH = 16;
    const int nb = (H + BH - 1) / BH;
    const size_t stride = points ? sizeof(SdCompactPoint) : sizeof(SdCompactDepth);
    const char* rec = static_cast<const char*>(records);
    auto pixel_of = [&](int i) { return *reinterpret_cast<const uint32_t*>(rec + (size_t)i * stride); };
    thread_local std::vector<int> start, order;
    thread_local AlignedFloats stage;
    start.assign((size_t)nb + 1, 0);
    for (int i = 0; i < n; ++i) {
        const uint32_t p = pixel_of(i);
        const int y = (int)(p >> 16), x = (int)(p & 0xffffu);
        if (y >= H || x >= W) throw std::runtime_error("expand_streaming: record outside the image");
        start[(size_t)y / BH + 1]++;
    }
    for (int b = 0; b < nb; ++b) start[(size_t)b + 1] += start[b];
    order.resize((size_t)n);
    {
        std::vector<int> at(start.begin(), start.end() - 1);
        for (int i = 0; i < n; ++i) order[(size_t)at[(pixel_of(i) >> 16) / BH]++] = i;
    }
    if (clear_points)
        for (uint32_t e : filled) {
            const size_t o = (size_t)(e >> 16) * W + (e & 0xffffu);
            kf->SemiDensePointSets_[3 * o] = 0.0f; kf->SemiDensePointSets_[3 * o + 1] = 0.0f; kf->SemiDensePointSets_[3 * o + 2] = 0.0f;
        }
    const size_t band_px = (size_t)BH * W;
    float* sd = stage.get(band_px * (points ? 5 : 2) + 64);
    float* ss = sd + (band_px + 15) / 16 * 16;
    float* sp = ss + (band_px + 15) / 16 * 16;
    for (int b = 0; b < nb; ++b) {
        const int y0 = b * BH, rows = std::min(BH, H - y0);
        const size_t cnt = (size_t)rows * W;
        std::memset(sd, 0, cnt * sizeof(float));
        std::memset(ss, 0, cnt * sizeof(float));
        if (points) std::memset(sp, 0, 3 * cnt * sizeof(float));
        for (int k = start[b]; k < start[(size_t)b + 1]; ++k) {
            const char* r = rec + (size_t)order[(size_t)k] * stride;
            const SdCompactDepth* d = reinterpret_cast<const SdCompactDepth*>(r);  // SdCompactPoint begins with the same three fields
            const size_t o = (size_t)((int)(d->pixel >> 16) - y0) * W + (d->pixel & 0xffffu);
            sd[o] = d->depth;
            ss[o] = d->sigma;
            if (points) {
                const SdCompactPoint* q = reinterpret_cast<const SdCompactPoint*>(r);
                sp[3 * o] = q->x; sp[3 * o + 1] = q->y; sp[3 * o + 2] = q->z;
            }
        }
        const size_t off = (size_t)y0 * W;
        stream_copy(kf->depth_map_.data() + off, sd, cnt);
        stream_copy(kf->depth_sigma_.data() + off, ss, cnt);
        if (points) stream_copy(kf->SemiDensePointSets_.data() + 3 * off, sp, 3 * cnt);
    }
#if defined(__SSE2__)
    _mm_sfence();
#endif
    filled.resize((size_t)n);
    for (int i = 0; i < n; ++i) filled[(size_t)i] = pixel_of(i);
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
    expand_records(kf, records_.data(), n, points, filled_[kf->mnId]);
}

// Compact records -> the dense members the reference leaves in the KeyFrame (PM.cc:225-226, 361-363).  `filled` lists the
// pixels the previous expansion of this key frame wrote: only those are cleared, instead of the whole planes.
void ProbabilityMapping::expand_records(KeyFrameView* kf, const void* records, int n, bool points, std::vector<uint32_t>& filled) {
    const size_t px = (size_t)kf->im_.rows * kf->im_.cols;
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
        const SdCompactPoint* r = static_cast<const SdCompactPoint*>(records);
        for (int i = 0; i < n; ++i) {
            const size_t o = (size_t)(r[i].pixel >> 16) * W + (r[i].pixel & 0xffffu);
            kf->depth_map_[o] = r[i].depth;
            kf->depth_sigma_[o] = r[i].sigma;
            kf->SemiDensePointSets_[3 * o] = r[i].x; kf->SemiDensePointSets_[3 * o + 1] = r[i].y; kf->SemiDensePointSets_[3 * o + 2] = r[i].z;
            filled[i] = r[i].pixel;
        }
    } else {
        const SdCompactDepth* r = static_cast<const SdCompactDepth*>(records);
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
