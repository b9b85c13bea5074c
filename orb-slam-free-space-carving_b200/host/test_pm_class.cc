// Test driver of the C++ host mirror (ProbabilityMapping.h): reads a scene file written by
// tests/test_host_class.py, runs ProbabilityMapping::Run() through the class surface and writes the
// KeyFrameView results back for the Python side to compare with the oracle.
//
//   test_pm_class <scene.bin> <out.bin> [obj path [max sigma]]   one pass (RunOnce) over the whole map on cuda:0
//   test_pm_class --stream[=limit] <scene.bin> <out.bin> [obj [max sigma]]
//        replay stub of the live system: the worker thread runs ProbabilityMapping::Run() (PM.cc:65-134)
//        while this thread inserts the recorded key frames one at a time, as LocalMapping does (LocalMapping.cc:94);
//        `limit` > 0 bounds the key frames resident on the device (SetResidentLimit)
//   test_pm_class --stale-pose <scene.bin> <out.bin>   uploads every key frame with a WRONG pose first, then SetPose(true
//        pose) and one pass: the results must be those of the true poses (the reference reads poses at call time)
//   test_pm_class --bench=steps[,warmup[,window[,threads]]] <scene.bin> <out.bin>
//        bench.py's `e2e_class` leg: `steps` timed passes of SemiDenseLoop over the whole map, every pass from scratch
//        (key frames released and flags cleared, so that images are uploaded and results expanded into the KeyFrame
//        members again, as in a first pass); prints one line `bench: {json}`; out.bin as in the plain mode (last pass)
//   test_pm_class --no-device-check                  exits 0 iff construction fails loudly without a GPU
//
// scene.bin (little endian): int32 nkf, w, h, covis_n; then per key frame:
//   int32 id; float K[4]; float Tcw[12]; int32 ncov; int32 cov[ncov] (indices into the file order);
//   int32 nd; float depths[nd]; int32 nkp; int64 mp[nkp]; float ang[nkp]; uint8 image[w*h]
// then int32 nprobe; per probe: int32 kf1_index, kf2_index, x, y; float min_depth, max_depth, rot
// out.bin: per key frame: int32 semidense_flag, interKF_flag; float depth[w*h], sigma[w*h]; int32 has_points;
//   float points[3*w*h] if has_points; then per probe: float depth, sigma, best_u, best_v, supported, umin, umax
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <thread>
#include <vector>

#include "ProbabilityMapping.h"

using namespace sdpm_host;

namespace {
template <typename T>
void rd(std::FILE* f, T* p, size_t n) {
    if (n && std::fread(p, sizeof(T), n, f) != n) throw std::runtime_error("scene file truncated");
}
template <typename T>
void wr(std::FILE* f, const T* p, size_t n) {
    if (n && std::fwrite(p, sizeof(T), n, f) != n) throw std::runtime_error("short write");
}
struct Probe {
    int kf1, kf2, x, y;
    float mind, maxd, rot;
};
}  // namespace

int main(int argc, char** argv) {
    if (argc >= 2 && !std::strcmp(argv[1], "--no-device-check")) {
        Map map;
        try {
            ProbabilityMapping pm(&map, 0);
        } catch (const std::exception& e) {
            std::printf("construction failed loudly: %s\n", e.what());
            return 0;
        }
        std::printf("a CUDA device is present: construction succeeded\n");
        return 0;
    }
    bool stream = false, stale_pose = false;
    int bench_steps = 0, bench_warmup = 1, bench_window = 32, bench_threads = 0;
    long limit = 0;
    if (argc >= 2 && !std::strncmp(argv[1], "--stream", 8)) {
        stream = true;
        if (argv[1][8] == '=') limit = std::atol(argv[1] + 9);
        --argc; ++argv;
    } else if (argc >= 2 && !std::strcmp(argv[1], "--stale-pose")) {
        stale_pose = true;
        --argc; ++argv;
    } else if (argc >= 2 && !std::strncmp(argv[1], "--bench=", 8)) {
        std::sscanf(argv[1] + 8, "%d,%d,%d,%d", &bench_steps, &bench_warmup, &bench_window, &bench_threads);
        if (bench_steps < 1) bench_steps = 1;
        --argc; ++argv;
    }
    if (argc < 3) {
        std::fprintf(stderr, "usage: %s [--stream[=limit] | --stale-pose] scene.bin out.bin [cloud.obj [max sigma]]\n", argv[0]);
        return 2;
    }
    try {
        std::FILE* f = std::fopen(argv[1], "rb");
        if (!f) throw std::runtime_error("cannot open scene file");
        int32_t hdr[4];
        rd(f, hdr, 4);
        const int nkf = hdr[0], w = hdr[1], h = hdr[2], covis_n = hdr[3];
        std::vector<std::unique_ptr<KeyFrameView>> kfs;
        std::vector<std::vector<uint8_t>> images(nkf);
        std::vector<std::vector<int32_t>> cov(nkf);
        for (int i = 0; i < nkf; ++i) {
            std::unique_ptr<KeyFrameView> kf(new KeyFrameView());
            int32_t id, n;
            rd(f, &id, 1);
            kf->mnId = id;
            rd(f, kf->K, 4);
            rd(f, kf->Tcw, 12);
            rd(f, &n, 1);
            cov[i].resize(n);
            rd(f, cov[i].data(), n);
            rd(f, &n, 1);
            kf->point_depths.resize(n);
            rd(f, kf->point_depths.data(), n);
            rd(f, &n, 1);
            kf->map_points.resize(n);
            kf->keypoint_angles.resize(n);
            rd(f, kf->map_points.data(), n);
            rd(f, kf->keypoint_angles.data(), n);
            images[i].resize((size_t)w * h);
            rd(f, images[i].data(), images[i].size());
            kf->im_ = PlaneView<const uint8_t>(images[i].data(), h, w);
            kfs.push_back(std::move(kf));
        }
        int32_t nprobe = 0;
        rd(f, &nprobe, 1);
        std::vector<Probe> probes(nprobe);
        for (auto& p : probes) {
            int32_t v[4];
            rd(f, v, 4);
            p.kf1 = v[0]; p.kf2 = v[1]; p.x = v[2]; p.y = v[3];
            float g[3];
            rd(f, g, 3);
            p.mind = g[0]; p.maxd = g[1]; p.rot = g[2];
        }
        std::fclose(f);

        Map map;
        for (int i = 0; i < nkf; ++i)
            for (int c : cov[i]) kfs[i]->covisible.push_back(kfs[c].get());

        ProbabilityMapping pm(&map, 0, covis_n);
        if (stream) {
            for (int i = 0; i < nkf; ++i) kfs[i]->mapped = false;
            pm.SetResidentLimit((size_t)limit);
            pm.SetPointCloudPath(argc >= 4 ? argv[3] : "");
            std::thread worker([&pm] { pm.Run(); });  // started like mptModeler, System.cc:110-111
            // a third thread plays bundle adjustment / loop closing: it keeps announcing poses (the true ones, so the
            // expected result does not change) through SetPose while the worker reads them under mMutexPose
            std::atomic<int> inserted{0};
            std::atomic<bool> stop_poses{false};
            std::atomic<uint64_t> pose_writes{0};
            std::thread poses([&] {
                std::vector<float> T(12);
                for (uint64_t k = 0; !stop_poses.load(); ++k) {
                    const int n = inserted.load();
                    if (n > 0) {
                        KeyFrameView* kf = kfs[(size_t)((k * 7919u) % (uint64_t)n)].get();
                        kf->GetPose(T.data());
                        kf->SetPose(T.data());
                        pose_writes++;
                    }
                    usleep(1500);
                }
            });
            for (int i = 0; i < nkf; ++i) {
                kfs[i]->mapped = true;
                map.AddKeyFrame(kfs[i].get());
                inserted = i + 1;
                usleep(i % 3 == 0 ? 20000 : 1000);     // irregular arrival, several worker passes in between
            }
            stop_poses = true;
            poses.join();
            std::printf("stream: %llu SetPose calls from a second thread while the worker ran\n", (unsigned long long)pose_writes.load());
            const uint64_t p0 = pm.Passes();
            while (pm.Passes() < p0 + 2) usleep(1000);  // two full passes after the last insertion
            pm.RequestFinish();
            worker.join();
            if (!pm.isFinished()) throw std::runtime_error("worker did not set mbFinished");
            std::printf("stream: %llu passes, resident %zu (max after release %zu, limit %ld)\n", (unsigned long long)pm.Passes(),
                        pm.ResidentKeyFrames(), pm.MaxResidentKeyFrames(), limit);
            if (limit > 0 && (long)pm.MaxResidentKeyFrames() > limit + 3 * covis_n + 1)
                throw std::runtime_error("resident limit not enforced");
        } else if (bench_steps > 0) {
            for (int i = 0; i < nkf; ++i) map.AddKeyFrame(kfs[i].get());
            pm.SetBatchWindow((size_t)bench_window, (unsigned)bench_threads);
            auto one_pass = [&] {
                for (int i = 0; i < nkf; ++i) {
                    pm.ReleaseKeyFrame(kfs[i].get());
                    kfs[i]->semidense_flag_ = false;
                    kfs[i]->interKF_depth_flag_ = false;
                }
                pm.SemiDenseLoop();
            };
            for (int i = 0; i < std::max(bench_warmup, 1); ++i) one_pass();  // first pass also allocates the KeyFrame members
            const auto t0 = std::chrono::steady_clock::now();
            for (int i = 0; i < bench_steps; ++i) one_pass();
            const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            int done = 0;
            for (int i = 0; i < nkf; ++i) done += kfs[i]->interKF_depth_flag_ ? 1 : 0;
            const ProbabilityMapping::PassStats& ps = pm.LastPassStats();
            std::printf("last pass: %d windows, resident %.1f ms, prepare %.1f ms, issue %.1f ms, expand %.1f ms, sync %.1f ms, total %.1f ms\n",
                        ps.windows, 1e3 * ps.resident, 1e3 * ps.prepare, 1e3 * ps.issue, 1e3 * ps.expand, 1e3 * ps.sync, 1e3 * ps.total);
            std::printf("bench: {\"keyframes\": %d, \"steps\": %d, \"warmup\": %d, \"window\": %d, \"seconds\": %.6f, "
                        "\"kf_per_s\": %.2f, \"ms_per_step\": %.3f, \"inter_checked\": %d}\n",
                        nkf, bench_steps, std::max(bench_warmup, 1), bench_window, sec, bench_steps * (double)nkf / sec,
                        1e3 * sec / bench_steps, done);
        } else {
            for (int i = 0; i < nkf; ++i) map.AddKeyFrame(kfs[i].get());
            if (stale_pose) {
                std::vector<std::vector<float>> truth(nkf);
                for (int i = 0; i < nkf; ++i) {
                    truth[i].assign(kfs[i]->Tcw, kfs[i]->Tcw + 12);
                    kfs[i]->Tcw[3] += 0.07f + 0.01f * i;   // a wrong translation ...
                    kfs[i]->Tcw[7] -= 0.03f;
                }
                for (int i = 0; i + 1 < nkf; ++i) {        // ... goes to the device with the upload a probe triggers
                    ProbabilityMapping::depthHo dh;
                    float bu = 0, bv = 0, F12[9] = {0};
                    pm.EpipolarSearch(kfs[i].get(), kfs[i + 1].get(), w / 2, h / 2, 0.f, 1.2f, 0.8f, &dh, F12, bu, bv, 0.f, 0.f);
                }
                for (int i = 0; i < nkf; ++i) kfs[i]->SetPose(truth[i].data());   // BA / loop closing moved them
            }
            pm.RunOnce();
        }

        std::FILE* o = std::fopen(argv[2], "wb");
        if (!o) throw std::runtime_error("cannot open output file");
        for (int i = 0; i < nkf; ++i) {
            KeyFrameView* kf = kfs[i].get();
            int32_t flags[2] = {kf->semidense_flag_, kf->interKF_depth_flag_};
            wr(o, flags, 2);
            std::vector<float> zero((size_t)w * h, 0.f);
            wr(o, kf->depth_map_.size() ? kf->depth_map_.data() : zero.data(), (size_t)w * h);
            wr(o, kf->depth_sigma_.size() ? kf->depth_sigma_.data() : zero.data(), (size_t)w * h);
            int32_t hp = kf->SemiDensePointSets_.size() == (size_t)3 * w * h;
            wr(o, &hp, 1);
            if (hp) wr(o, kf->SemiDensePointSets_.data(), (size_t)3 * w * h);
        }
        for (const auto& p : probes) {
            ProbabilityMapping::depthHo dh;
            float bu = 0, bv = 0, F12[9];
            KeyFrameView *a = kfs[p.kf1].get(), *b = kfs[p.kf2].get();
            pm.ComputeFundamental(a, b, F12);
            pm.EpipolarSearch(a, b, p.x, p.y, (float)a->im_.at(p.y, p.x), p.mind, p.maxd, &dh, F12, bu, bv, 0.f, p.rot);
            float umin = 0, umax = 0;
            pm.GetSearchRange(umin, umax, p.x, p.y, p.mind, p.maxd, a, b);
            float rec[7] = {dh.depth, dh.sigma, bu, bv, dh.supported ? 1.f : 0.f, umin, umax};
            wr(o, rec, 7);
        }
        std::fclose(o);
        if (!stream && argc >= 4 && !pm.SavePointCloudObj(argv[3], argc >= 5 ? std::atof(argv[4]) : 0.01)) return 1;
        std::printf("ok: %d key frames %dx%d, covisN %d, %d probes\n", nkf, w, h, covis_n, (int)nprobe);
        return 0;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "test_pm_class: %s\n", e.what());
        return 1;
    }
}
