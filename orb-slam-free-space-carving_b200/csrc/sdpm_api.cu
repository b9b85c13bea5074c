// C ABI (include/sdpm.h) over the CUDA kernels in sdpm_kernels.cuh: the context that owns the
// device-resident key-frame planes, the three streams (upload+K0 / K1..K5 / download) and the
// per-pair host math.  No CPU fallback exists: every stage is a kernel launch.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/sdpm.h"
#include "pair_geometry.h"
#include "sdpm_kernels.cuh"

using namespace sdpm;

namespace {

enum { T_GRAD = 0, T_SEARCH, T_FUSE, T_INTRA, T_INTER, T_POINTS, T_COUNT };
const char* kTimerNames[T_COUNT] = {"grad", "search", "fuse", "intra", "inter", "points"};

struct Slot {
    int id = -1;
    float K[4] = {0, 0, 0, 0};
    float Rcw[9], tcw[3], Twc[16];
    uint8_t* slab = nullptr;  // one allocation for everything below except points
    uint8_t* gray = nullptr;
    float *grad = nullptr, *theta = nullptr, *depth = nullptr, *sigma = nullptr, *depth_chk = nullptr;
    uint32_t* arc = nullptr;  // SD_ARC_PLANES orientation-arc bit planes (k0_gradient)
    uint4* rec = nullptr;     // search records: one 16-byte {theta, grad, grad below, im | im below << 16} per pixel (k0_gradient)
    float* points = nullptr;
    uint32_t* list = nullptr;   // active pixels (k0_gradient), input of k1_search
    uint32_t* vlist = nullptr;  // pixels carrying a depth after the intra-KF stage (k23_intra), input of k4_inter_check
    float* ray = nullptr;       // W + H floats: (x - cx) / fx per column, (y - cy) / fy per row (k0_gradient), read by k1_search
    KfMeta* meta = nullptr;
    bool vlist_ok = false;      // vlist / meta->valid_count describe `depth`
    bool vlist_from_import = false;  // the list was rebuilt from imported planes (which may carry values outside the 2-px frame)
    bool has_semidense = false, has_checked = false, has_points = false, pose_changed = false;
    cudaEvent_t ev_up = nullptr;  // upload + K0 done (s_up)
    cudaEvent_t ev_dn = nullptr;  // last download issued (s_dn)
    bool dn_pending = false;
    cudaEvent_t ev_sd = nullptr;   // (depth, sigma) maps ready: after k23_intra / sd_import_maps (s_main)
    // Record numbers (see wait_for_*): every stream is in-order, so waiting for record q of a stream covers all its earlier ones
    long long up_seq = 0;          // number of this key frame's upload on s_up
    long long sd_seq = 0;          // number of the ev_sd record on s_main; 0 = no maps yet
    long long chk_idx = -1;        // ring index of the last K4 / K5 / K6 of this key frame (s_chk); -1 = none
    long long rd_idx = -1;         // ring index of the last K4 that read this key frame's maps (own or as a neighbour)
};

struct TimedSpan {
    int which;
    cudaEvent_t a, b;
};

}  // namespace

struct sd_ctx {
    int device = 0;
    SdParams P;
    int W = 0, H = 0, pitch_f = 0, pitch_u8 = 0, arc_pitch = 0;
    // s_up: upload + K0; s_srch: k1_search (low priority, fills the machine); s_main: everything after the
    // search (k1_fuse .. K5, NCCL ordering) at high priority, so the short latency-bound kernels of key frame
    // k run in the issue slots the search of key frame k+1 leaves idle; s_dn: downloads
    // s_chk: K4 + K5 (high priority), ordered against s_main by per-key-frame events, so the inter-key-frame
    // check of key frame j overlaps the fusion / intra stage of a later one
    cudaStream_t s_up = nullptr, s_srch = nullptr, s_main = nullptr, s_chk = nullptr, s_dn = nullptr;
    static const int RING = 64;
    cudaEvent_t ev_ring[RING] = {nullptr};  // ev_ring[i % RING]: record number i of s_chk (one per K4 / K5 / K6)
    long long ring_i = 0;
    long long up_seq = 0, sd_seq = 0;       // records issued so far on s_up (uploads) and on s_main (ev_sd)
    // highest record number of the producing stream that the consuming stream has already waited for
    long long srch_seen_up = 0, main_seen_up = 0, chk_seen_sd = 0, dn_seen_sd = 0, main_seen_chk = -1, dn_seen_chk = -1;
    bool dirty = false;  // work has been queued on s_srch / s_main / s_chk / s_dn since s_up last waited for those streams
    cudaEvent_t ev_tmp = nullptr, ev_span_a = nullptr, ev_span_b = nullptr, ev_fence = nullptr;
    std::unordered_map<int, Slot*> kfs;
    std::vector<Slot*> free_slots;
    // k1_fuse output / k23_intra input, one pair per hypothesis buffer; invariant: all zero while no key frame is in
    // flight on the buffer (k23_intra is followed by the memsets that restore it, off the key frame's own dependency chain)
    float *tmp_depth[2] = {nullptr, nullptr}, *tmp_sigma[2] = {nullptr, nullptr};
    bool tmp_dirty[2] = {false, false};  // a fusion was queued into the pair and the restoring memset has not been (error path)
    float2* hyp[2] = {nullptr, nullptr};               // hypothesis buffers of k1_search, [neighbour][hyp_stride], double-buffered
    int hyp_stride = 0, hyp_rows = 0, hyp_turn = 0;
    cudaEvent_t ev_srch[2] = {nullptr, nullptr};       // k1_search into hyp[b] done (s_srch)
    cudaEvent_t ev_hypfree[2] = {nullptr, nullptr};    // k1_fuse out of hyp[b] done (s_main)
    bool hypfree_valid[2] = {false, false};
    std::string err;
    uint64_t launches = 0;
    bool timing = false;
    bool timing_serial = false;  // sd_timing_enable(ctx, 2): every timed kernel runs alone (streams drained before it)
    std::vector<TimedSpan> spans;
    std::vector<cudaEvent_t> ev_pool;
    double t_ms[T_COUNT] = {0};
    uint64_t t_n[T_COUNT] = {0};
    float chi_fuse_f = 0, chi_inter_f = 0;
    int sm_count = 0, k1_grid = 0;  // persistent k1_search grid: SMs x resident blocks
    // K6 scratch (confident-point feed), grown on demand
    K6State* conf_state = nullptr;
    uint32_t* conf_hist = nullptr;
    int* conf_seg = nullptr;      // 4 x nseg ints
    uint8_t* conf_out = nullptr;  // cap x (xyz, sigma, pixel)
    int conf_cap = 0;
    int* h_conf = nullptr;        // pinned: {total, confident}
    const float* conf_last_xyz = nullptr;   // result of the last sd_confident_points_device call (device)
    const float* conf_last_sigma = nullptr;
    const uint32_t* conf_last_pix = nullptr;
    int conf_last_count = 0;
    // compact downloads (k7_pack): ring of device staging buffers; the record count of a key frame is only known on the
    // device, so the device->host copy of a pending download is issued once its count has arrived (poll_pack)
    static const int PACK_RING = 16;
    struct PackSlot {
        uint8_t* dev = nullptr;     // [int count | pad to 256 B | cap records]
        int* h_count = nullptr;     // pinned
        cudaEvent_t ev_cnt = nullptr;
        void* host_records = nullptr;
        int capacity = 0, stride = 0;
        int* count_out = nullptr;
    };
    PackSlot pack[PACK_RING];
    long long pack_head = 0, pack_tail = 0;  // pending downloads: [head, tail)
    int pack_cap = 0;
    // NCCL (dlopen), one communicator per context
    void* nccl_lib = nullptr;
    void* nccl_comm = nullptr;
    int nccl_world = 0, nccl_rank = -1;
};

namespace {

int fail(sd_ctx* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (c) c->err = buf;
    return code;
}

#define CU(c, call)                                                                                      \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess) return fail(c, SD_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

Slot* find(sd_ctx* c, int id) {
    auto it = c->kfs.find(id);
    return it == c->kfs.end() ? nullptr : it->second;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

int alloc_slot(sd_ctx* c, Slot** out) {
    if (!c->free_slots.empty()) {
        *out = c->free_slots.back();
        c->free_slots.pop_back();
        return SD_OK;
    }
    Slot* s = new Slot();
    const size_t fplane = align_up((size_t)c->pitch_f * c->H * sizeof(float), 256);
    const size_t uplane = align_up((size_t)c->pitch_u8 * c->H, 256);
    const size_t lplane = align_up((size_t)c->W * c->H * sizeof(uint32_t), 256);
    const size_t aplane = align_up((size_t)SD_ARC_PLANES * c->H * c->arc_pitch * sizeof(uint32_t), 256);
    const size_t rplane = align_up((size_t)c->pitch_f * c->H * sizeof(uint4), 256);
    const size_t yplane = align_up((size_t)(c->W + c->H) * sizeof(float), 256);
    const size_t total = uplane + 5 * fplane + aplane + rplane + 2 * lplane + yplane + 256;
    cudaError_t e = cudaMalloc(&s->slab, total);
    if (e != cudaSuccess) {
        delete s;
        return fail(c, SD_ERR_CUDA, "cudaMalloc(%zu): %s", total, cudaGetErrorString(e));
    }
    uint8_t* p = s->slab;
    s->gray = p; p += uplane;
    s->grad = (float*)p; p += fplane;
    s->theta = (float*)p; p += fplane;
    s->arc = (uint32_t*)p; p += aplane;
    s->rec = (uint4*)p; p += rplane;
    s->depth = (float*)p; p += fplane;
    s->sigma = (float*)p; p += fplane;
    s->depth_chk = (float*)p; p += fplane;
    s->list = (uint32_t*)p; p += lplane;
    s->vlist = (uint32_t*)p; p += lplane;
    s->ray = (float*)p; p += yplane;
    s->meta = (KfMeta*)p;
    cudaEventCreateWithFlags(&s->ev_up, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&s->ev_dn, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&s->ev_sd, cudaEventDisableTiming);
    *out = s;
    return SD_OK;
}

void free_slot(Slot* s) {
    if (!s) return;
    cudaFree(s->slab);
    cudaFree(s->points);
    if (s->ev_up) cudaEventDestroy(s->ev_up);
    if (s->ev_dn) cudaEventDestroy(s->ev_dn);
    if (s->ev_sd) cudaEventDestroy(s->ev_sd);
    delete s;
}

cudaEvent_t pool_event(sd_ctx* c) {
    if (!c->ev_pool.empty()) {
        cudaEvent_t e = c->ev_pool.back();
        c->ev_pool.pop_back();
        return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}

struct ScopedTimer {  // CUDA events on the launching stream around one kernel
    sd_ctx* c;
    cudaStream_t s;
    int which;
    cudaEvent_t a = nullptr, b = nullptr;
    ScopedTimer(sd_ctx* c_, cudaStream_t s_, int w) : c(c_), s(s_), which(w) {
        if (c->timing) {
            if (c->timing_serial) {  // standalone durations: nothing else may be running while this kernel does
                cudaStreamSynchronize(c->s_up);
                cudaStreamSynchronize(c->s_srch);
                cudaStreamSynchronize(c->s_main);
                cudaStreamSynchronize(c->s_chk);
                cudaStreamSynchronize(c->s_dn);
            }
            a = pool_event(c);
            b = pool_event(c);
            cudaEventRecord(a, s);
        }
    }
    ~ScopedTimer() {
        c->launches++;
        if (a) {
            cudaEventRecord(b, s);
            c->spans.push_back({which, a, b});
        }
    }
};

void collect_spans(sd_ctx* c) {
    for (auto& sp : c->spans) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess) {
            c->t_ms[sp.which] += ms;
            c->t_n[sp.which]++;
        }
        c->ev_pool.push_back(sp.a);
        c->ev_pool.push_back(sp.b);
    }
    c->spans.clear();
}

void set_pose_fields(Slot* s, const float Tcw[12]) {
    for (int r = 0; r < 3; ++r) {
        for (int col = 0; col < 3; ++col) s->Rcw[r * 3 + col] = Tcw[r * 4 + col];
        s->tcw[r] = Tcw[r * 4 + 3];
    }
    geom::pose_inverse(s->Rcw, s->tcw, s->Twc);
}

int upload_common(sd_ctx* c, int kf_id, const uint8_t* gray, bool on_device, int w, int h, int stride, const float K[4],
                  const float Tcw[12]) {
    if (!c || !gray || !K || !Tcw || w < 8 || h < 8 || stride < w ||
        w > 65534 || h > 65535 ||                  // list entries pack (y, x) in 16 bits each; k1_search packs u_hi + 1 <= w + 1
        (long long)(w + 128) * h >= (1ll << 30))  // kernels index planes with 32-bit element offsets
        return fail(c, SD_ERR_INVALID_ARG, "sd_upload_keyframe: bad argument");
    CU(c, cudaSetDevice(c->device));
    if (c->W == 0) {
        c->W = w;
        c->H = h;
        c->pitch_f = (int)align_up(w, 32);
        c->pitch_u8 = (int)align_up(w, 128);
        c->arc_pitch = 4 * ((w + K0_TW - 1) / K0_TW);  // u32 words per bit-plane row, whole K0 tiles
        const size_t fplane = (size_t)c->pitch_f * c->H * sizeof(float);
        for (int b = 0; b < 2; ++b) {
            CU(c, cudaMalloc(&c->tmp_depth[b], 2 * fplane));  // depth | sigma: one memset restores both
            c->tmp_sigma[b] = c->tmp_depth[b] + (size_t)c->pitch_f * c->H;
            // zeroed on the stream that uses them (k1_fuse runs on s_main): the streams are non-blocking, so a memset on the
            // legacy default stream would have no formal ordering before the first fusion
            CU(c, cudaMemsetAsync(c->tmp_depth[b], 0, 2 * fplane, c->s_main));
        }
    } else if (c->W != w || c->H != h) {
        return fail(c, SD_ERR_SHAPE, "key frame %d is %dx%d, context is %dx%d", kf_id, w, h, c->W, c->H);
    }
    Slot* s = find(c, kf_id);
    bool fresh = false;
    if (!s) {
        fresh = c->free_slots.empty();
        int rc = alloc_slot(c, &s);
        if (rc != SD_OK) return rc;
        s->id = kf_id;
        c->kfs[kf_id] = s;
    }
    if (!fresh && c->dirty) {
        // the planes are being replaced (same key frame again, or a slot recycled by sd_release_keyframe):
        // wait for everything already queued that may still read them.  Once per batch of uploads: s_up is in-order,
        // so while nothing new is queued on the other streams (c->dirty) the earlier wait still covers them
        c->dirty = false;
        CU(c, cudaEventRecord(c->ev_tmp, c->s_main));
        CU(c, cudaStreamWaitEvent(c->s_up, c->ev_tmp, 0));
        CU(c, cudaEventRecord(c->ev_tmp, c->s_srch));
        CU(c, cudaStreamWaitEvent(c->s_up, c->ev_tmp, 0));
        CU(c, cudaEventRecord(c->ev_tmp, c->s_chk));
        CU(c, cudaStreamWaitEvent(c->s_up, c->ev_tmp, 0));
        CU(c, cudaEventRecord(c->ev_tmp, c->s_dn));
        CU(c, cudaStreamWaitEvent(c->s_up, c->ev_tmp, 0));
    }
    s->has_semidense = s->has_checked = s->has_points = s->pose_changed = s->vlist_ok = s->vlist_from_import = false;
    s->dn_pending = false;
    s->sd_seq = 0;
    s->chk_idx = s->rd_idx = -1;
    s->up_seq = ++c->up_seq;
    std::memcpy(s->K, K, sizeof(s->K));
    set_pose_fields(s, Tcw);

    CU(c, cudaMemcpy2DAsync(s->gray, c->pitch_u8, gray, stride, w, h, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->s_up));
    CU(c, cudaMemsetAsync(s->meta, 0, sizeof(KfMeta), c->s_up));
    {
        ScopedTimer t(c, c->s_up, T_GRAD);
        dim3 blk(K0_TX, K0_TY), grd((w + K0_TW - 1) / K0_TW, (h + K0_TY - 1) / K0_TY);
        k0_gradient<<<grd, blk, 0, c->s_up>>>(s->gray, c->pitch_u8, w, h, s->grad, s->theta, c->pitch_f, s->arc, c->arc_pitch,
                                              s->rec, s->list, s->meta, c->P.lambdaG, c->P.lambdaTheta, c->P.border,
                                              s->ray, s->K[0], s->K[1], s->K[2], s->K[3]);
    }
    CU(c, cudaGetLastError());
    CU(c, cudaEventRecord(s->ev_up, c->s_up));
    return SD_OK;
}

// Cross-stream waits.  Every stream of the context is in-order and its records are numbered (uploads on s_up: Slot::up_seq;
// ev_sd records on s_main: Slot::sd_seq; K4 / K5 / K6 on s_chk: ring indices), so a consumer stream that has waited for
// record q of a producer stream is past every earlier record of it too: one cudaStreamWaitEvent on the LATEST record
// needed, none when the consumer has already waited for that one or a later one (`seen`).  An event that has been
// re-recorded since (the key frame uploaded again, a ring slot reused) marks a later point of the same stream, which
// still implies the one asked for.
cudaError_t wait_upload(cudaStream_t stream, long long& seen, const Slot* latest) {
    if (!latest || latest->up_seq <= seen) return cudaSuccess;
    seen = latest->up_seq;
    return cudaStreamWaitEvent(stream, latest->ev_up, 0);
}
cudaError_t wait_maps(cudaStream_t stream, long long& seen, const Slot* latest) {
    if (!latest || latest->sd_seq <= seen) return cudaSuccess;  // sd_seq == 0: no maps recorded
    seen = latest->sd_seq;
    return cudaStreamWaitEvent(stream, latest->ev_sd, 0);
}
cudaError_t wait_ring(sd_ctx* c, cudaStream_t stream, long long& seen, long long idx) {
    if (idx <= seen) return cudaSuccess;  // idx == -1: none
    seen = idx;
    return cudaStreamWaitEvent(stream, c->ev_ring[idx % sd_ctx::RING], 0);
}
// record number of the next K4 / K5 / K6 on s_chk (call right after the launch)
cudaError_t record_chk(sd_ctx* c, Slot* s) {
    const long long idx = c->ring_i++;
    s->chk_idx = idx;
    return cudaEventRecord(c->ev_ring[idx % sd_ctx::RING], c->s_chk);
}
// Before s_main overwrites the maps of a key frame: wait for the K4 launches (s_chk) that read them, as the key frame's
// own or as a neighbour's, and for its own K5 / K6.
cudaError_t wait_map_readers(sd_ctx* c, Slot* s) {
    return wait_ring(c, c->s_main, c->main_seen_chk, std::max(s->rd_idx, s->chk_idx));
}
void record_maps(sd_ctx* c, Slot* s) { s->sd_seq = ++c->sd_seq; }


// ---- compact downloads: pending queue ----
// issue the device->host copy of pending download `i` (its count has arrived in pinned memory)
cudaError_t pack_issue(sd_ctx* c, long long i) {
    sd_ctx::PackSlot& ps = c->pack[i % sd_ctx::PACK_RING];
    const int avail = *ps.h_count;
    if (ps.count_out) *ps.count_out = avail;
    const int ncopy = std::min(avail, ps.capacity);
    cudaError_t e = cudaSuccess;
    if (ncopy > 0 && ps.host_records)
        e = cudaMemcpyAsync(ps.host_records, ps.dev + 256, (size_t)ncopy * ps.stride, cudaMemcpyDeviceToHost, c->s_dn);
    ps.host_records = nullptr;
    ps.count_out = nullptr;
    return e;
}

// drain == false: issue the copies of the pending downloads whose count has arrived (no waiting);
// drain == true: wait for every count and issue every copy
cudaError_t poll_pack(sd_ctx* c, bool drain) {
    while (c->pack_head < c->pack_tail) {
        sd_ctx::PackSlot& ps = c->pack[c->pack_head % sd_ctx::PACK_RING];
        cudaError_t q = drain ? cudaEventSynchronize(ps.ev_cnt) : cudaEventQuery(ps.ev_cnt);
        if (q == cudaErrorNotReady) return cudaSuccess;
        if (q != cudaSuccess) return q;
        cudaError_t e = pack_issue(c, c->pack_head);
        if (e != cudaSuccess) return e;
        c->pack_head++;
    }
    return cudaSuccess;
}

void fill_search_common(sd_ctx* c, Slot* s, SearchArgs& A) {
    A.W = c->W; A.H = c->H; A.pitch_f = c->pitch_f; A.pitch_u8 = c->pitch_u8; A.arc_pitch = c->arc_pitch;
    A.border = c->P.border; A.lambdaN = c->P.lambdaN; A.lambdaG = c->P.lambdaG;
    A.fx = s->K[0]; A.fy = s->K[1]; A.cx = s->K[2]; A.cy = s->K[3];
    A.lambdaL = c->P.lambdaL; A.lambdaTheta = c->P.lambdaTheta; A.slope = c->P.slope_limit;
    A.chi_fuse = c->chi_fuse_f;
    A.THETA = c->P.theta;
    A.inv_THETA = 1 / c->P.theta;
    A.inv_theta_f = (float)A.inv_THETA;
    A.ray = s->ray;
    A.img = s->gray; A.grad = s->grad; A.theta = s->theta; A.rec = s->rec;
    A.list = s->list; A.meta = s->meta;
    A.count_override = -1;
    A.task_counter = &s->meta->k1_task;
    A.hyp = nullptr; A.hyp_stride = 0;
    A.probe_out = nullptr;
}

void fill_nbr(Slot* s, Slot* nb, float rot, NbrConst& N) {
    geom::fundamental(s->K, s->Rcw, s->tcw, nb->K, nb->Rcw, nb->tcw, N.F);
    geom::relative_pose(s->Rcw, s->tcw, nb->Rcw, nb->tcw, N.R, N.t);
    N.rot = rot;
    for (int k = 0; k < 3; ++k) { N.Rd0[k] = (double)N.R[k]; N.Rd2[k] = (double)N.R[6 + k]; }
    N.td0 = (double)N.t[0];
    N.td2 = (double)N.t[2];
    N.img = nb->gray; N.grad = nb->grad; N.theta = nb->theta; N.arc = nb->arc; N.rec = nb->rec; N.meta = nb->meta;
}

}  // namespace

extern "C" {

int sd_default_params(SdParams* p) {
    if (!p) return SD_ERR_INVALID_ARG;
    p->lambdaG = 8;
    p->lambdaL = 80;
    p->lambdaTheta = 45;
    p->lambdaN = 3;
    p->theta = 0.23;
    p->chi2_fuse = 5.99;
    p->chi2_inter = 3.84;
    p->slope_limit = 4;
    p->border = 2;
    p->interkf_inplace = 0;
    return SD_OK;
}

int sd_create(int device, const SdParams* params, sd_ctx** out) {
    if (!out) return SD_ERR_INVALID_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || device < 0 || device >= ndev) return SD_ERR_NO_DEVICE;
    if (cudaSetDevice(device) != cudaSuccess) return SD_ERR_NO_DEVICE;
    sd_ctx* c = new sd_ctx();
    c->device = device;
    if (params) c->P = *params; else sd_default_params(&c->P);
    // border: the reference hard-codes the 2-pixel frame in every loop (PM.cc:198-199, 492, 555, 659, 342); the field
    // exists so that the struct mirrors the literal, but no other value has a reference behaviour to match
    if (c->P.border != 2 || c->P.lambdaN < 1 || !(c->P.theta > 0) || !(c->P.lambdaTheta >= 0) || !(c->P.lambdaL >= 0)) {
        delete c;
        return SD_ERR_INVALID_ARG;
    }
    c->chi_fuse_f = geom::round_up_to_float(c->P.chi2_fuse);
    c->chi_inter_f = geom::round_up_to_float(c->P.chi2_inter);
    int prio_least = 0, prio_greatest = 0;
    cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest);
    if (cudaStreamCreateWithFlags(&c->s_up, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithPriority(&c->s_main, cudaStreamNonBlocking, prio_greatest) != cudaSuccess ||
        cudaStreamCreateWithPriority(&c->s_srch, cudaStreamNonBlocking, prio_least) != cudaSuccess ||
        cudaStreamCreateWithPriority(&c->s_chk, cudaStreamNonBlocking, prio_greatest) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_srch[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_srch[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_hypfree[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_hypfree[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->s_dn, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_tmp, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreate(&c->ev_span_a) != cudaSuccess || cudaEventCreate(&c->ev_span_b) != cudaSuccess) {
        delete c;
        return SD_ERR_CUDA;
    }
    {
        cudaDeviceProp prop;
        int per_sm = 0;
        if (cudaGetDeviceProperties(&prop, device) != cudaSuccess ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_search, K1_WARPS * 32, 0) != cudaSuccess || per_sm < 1) {
            delete c;
            return SD_ERR_CUDA;
        }
        if (cudaFuncSetAttribute(k23_intra<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K23_SMEM) != cudaSuccess ||
            cudaFuncSetAttribute(k23_intra<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K23_SMEM) != cudaSuccess ||
            cudaFuncSetAttribute(k23_intra<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K23_SMEM) != cudaSuccess) {
            delete c;
            return SD_ERR_CUDA;
        }
        c->sm_count = prop.multiProcessorCount;
        // SDPM_K1_BLOCKS_PER_SM < occupancy leaves room for the post-search kernels of the previous key frame
        int use = per_sm;
        if (const char* e = getenv("SDPM_K1_BLOCKS_PER_SM")) use = std::max(1, std::min(per_sm, atoi(e)));
        c->k1_grid = c->sm_count * use;
        for (int i = 0; i < sd_ctx::RING; ++i)
            if (cudaEventCreateWithFlags(&c->ev_ring[i], cudaEventDisableTiming) != cudaSuccess) {
                delete c;
                return SD_ERR_CUDA;
            }
    }
    *out = c;
    return SD_OK;
}

int sd_destroy(sd_ctx* c) {
    if (!c) return SD_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    for (auto& kv : c->kfs) free_slot(kv.second);
    for (auto* s : c->free_slots) free_slot(s);
    for (auto& sp : c->spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
    for (auto e : c->ev_pool) cudaEventDestroy(e);
    for (int b = 0; b < 2; ++b) {
        cudaFree(c->tmp_depth[b]);  // tmp_sigma[b] is the second half of the same allocation
    }
    cudaFree(c->hyp[0]);
    cudaFree(c->hyp[1]);
    for (int b = 0; b < 2; ++b) {
        if (c->ev_srch[b]) cudaEventDestroy(c->ev_srch[b]);
        if (c->ev_hypfree[b]) cudaEventDestroy(c->ev_hypfree[b]);
    }
    if (c->ev_tmp) cudaEventDestroy(c->ev_tmp);
    if (c->ev_span_a) cudaEventDestroy(c->ev_span_a);
    if (c->ev_span_b) cudaEventDestroy(c->ev_span_b);
    if (c->ev_fence) cudaEventDestroy(c->ev_fence);
    cudaFree(c->conf_state);
    cudaFree(c->conf_hist);
    cudaFree(c->conf_seg);
    cudaFree(c->conf_out);
    if (c->h_conf) cudaFreeHost(c->h_conf);
    for (auto& ps : c->pack) {
        cudaFree(ps.dev);
        if (ps.h_count) cudaFreeHost(ps.h_count);
        if (ps.ev_cnt) cudaEventDestroy(ps.ev_cnt);
    }
    sd_comm_destroy(c);
    cudaStreamDestroy(c->s_up);
    cudaStreamDestroy(c->s_main);
    cudaStreamDestroy(c->s_srch);
    cudaStreamDestroy(c->s_chk);
    for (int i = 0; i < sd_ctx::RING; ++i)
        if (c->ev_ring[i]) cudaEventDestroy(c->ev_ring[i]);
    cudaStreamDestroy(c->s_dn);
    delete c;
    return SD_OK;
}

int sd_upload_keyframe(sd_ctx* c, int kf_id, const uint8_t* gray, int w, int h, int stride, const float K[4], const float Tcw[12]) {
    return upload_common(c, kf_id, gray, false, w, h, stride, K, Tcw);
}

int sd_upload_keyframe_device(sd_ctx* c, int kf_id, const uint8_t* gray_dev, int w, int h, int stride, const float K[4],
                              const float Tcw[12]) {
    return upload_common(c, kf_id, gray_dev, true, w, h, stride, K, Tcw);
}

int sd_set_pose(sd_ctx* c, int kf_id, const float Tcw[12]) {
    if (!c || !Tcw) return fail(c, SD_ERR_INVALID_ARG, "sd_set_pose: bad argument");
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    set_pose_fields(s, Tcw);
    s->pose_changed = true;
    return SD_OK;
}

int sd_release_keyframe(sd_ctx* c, int kf_id) {
    if (!c) return SD_ERR_INVALID_ARG;
    auto it = c->kfs.find(kf_id);
    if (it == c->kfs.end()) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    // No synchronisation here: queued work may still read the planes, so the slot is only parked; whoever takes it
    // from the free list (sd_upload_keyframe) first orders its writes after everything queued on the other streams.
    c->free_slots.push_back(it->second);
    c->kfs.erase(it);
    return SD_OK;
}

int sd_semidense(sd_ctx* c, int kf_id, const int* nbr_ids, const float* nbr_rot_deg, int n, float min_invdepth, float max_invdepth) {
    if (!c || !nbr_ids || n < 1 || n > SD_MAX_NEIGHBOURS) return fail(c, SD_ERR_INVALID_ARG, "sd_semidense: bad argument (n=%d)", n);
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    CU(c, cudaSetDevice(c->device));
    CU(c, poll_pack(c, false));
    static thread_local SearchArgs A;
    fill_search_common(c, s, A);
    A.n = n;
    A.mind = min_invdepth;
    A.maxd = max_invdepth;
    A.mind_d = (double)min_invdepth;
    A.maxd_d = (double)max_invdepth;
    if (c->hyp_rows < n) {  // grow-only scratch: 2 x n x (W*H) hypotheses
        CU(c, cudaStreamSynchronize(c->s_srch));
        CU(c, cudaStreamSynchronize(c->s_main));
        for (int b = 0; b < 2; ++b) {
            cudaFree(c->hyp[b]);
            c->hyp[b] = nullptr;
            c->hypfree_valid[b] = false;
        }
        c->hyp_rows = 0;
        c->hyp_stride = (int)align_up((size_t)c->W * c->H, 64);
        for (int b = 0; b < 2; ++b) CU(c, cudaMalloc(&c->hyp[b], (size_t)n * c->hyp_stride * sizeof(float2)));
        c->hyp_rows = n;
    }
    const int hb = c->hyp_turn;
    c->hyp_turn ^= 1;
    A.hyp = c->hyp[hb];
    A.hyp_stride = c->hyp_stride;
    // ---- search stream: k1_search of this key frame into hyp[hb] ----
    cudaStream_t s_srch = c->s_srch;
    c->dirty = true;
    const Slot* latest_up = s;  // the search reads the planes of kf and of its neighbours: wait for the last of those uploads
    for (int j = 0; j < n; ++j) {
        Slot* nb = find(c, nbr_ids[j]);
        if (!nb) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown neighbour key frame %d", nbr_ids[j]);
        fill_nbr(s, nb, nbr_rot_deg ? nbr_rot_deg[j] : 0.f, A.nb[j]);
        if (nb->up_seq > latest_up->up_seq) latest_up = nb;
    }
    CU(c, wait_upload(s_srch, c->srch_seen_up, latest_up));
    // buffers [hb] are free once the k23_intra of their previous key frame has read them and they are zero again
    if (c->hypfree_valid[hb]) CU(c, cudaStreamWaitEvent(s_srch, c->ev_hypfree[hb], 0));
    // meta->k1_task is zero here: zeroed with the meta block at upload, and again by the last warp of every k1_search launch
    {
        const int grid = c->k1_grid;
        ScopedTimer t(c, s_srch, T_SEARCH);
        k1_search<<<grid, K1_WARPS * 32, 0, s_srch>>>(A);
    }
    CU(c, cudaEventRecord(c->ev_srch[hb], s_srch));
    // ---- main stream: fusion into the scratch planes [hb] (zero = no depth), then k23_intra reads them with a
    // halo and writes the key frame's own planes ----
    CU(c, cudaStreamWaitEvent(c->s_main, c->ev_srch[hb], 0));
    if (s->dn_pending) {  // a download of the old maps may still be in flight
        CU(c, cudaStreamWaitEvent(c->s_main, s->ev_dn, 0));
        s->dn_pending = false;
    }
    CU(c, wait_map_readers(c, s));  // k23_intra below overwrites this key frame's maps and list (k1_fuse zeroes meta->valid_count)
    const size_t fplane = (size_t)c->pitch_f * c->H * sizeof(float);
    if (c->tmp_dirty[hb])  // an earlier call failed between its fusion and the memset that restores the planes
        CU(c, cudaMemsetAsync(c->tmp_depth[hb], 0, 2 * fplane, c->s_main));
    c->tmp_dirty[hb] = true;
    {
        ScopedTimer t(c, c->s_main, T_FUSE);
        FuseArgs F;
        F.n = n; F.lambdaN = c->P.lambdaN; F.pitch_f = c->pitch_f; F.hyp_stride = c->hyp_stride; F.chi_fuse = c->chi_fuse_f;
        F.hyp = c->hyp[hb]; F.list = s->list; F.meta = s->meta; F.depth = c->tmp_depth[hb]; F.sigma = c->tmp_sigma[hb];
        const int blocks = ((c->W - 2 * c->P.border) * (c->H - 2 * c->P.border) + 127) / 128;  // the list's capacity
        if (n <= 6) k1_fuse<6><<<blocks, 128, 0, c->s_main>>>(F);
        else if (n <= 10) k1_fuse<10><<<blocks, 128, 0, c->s_main>>>(F);
        else if (n <= 16) k1_fuse<16><<<blocks, 128, 0, c->s_main>>>(F);
        else {
            const int cap = (c->W - 2 * c->P.border) * (c->H - 2 * c->P.border);
            k1_fuse_rows<<<(cap + K1FR_THREADS - 1) / K1FR_THREADS, K1FR_THREADS, k1fr_smem_bytes(n), c->s_main>>>(F);
        }
    }
    {
        ScopedTimer t(c, c->s_main, T_INTRA);
        IntraArgs S;
        S.W = c->W; S.H = c->H; S.pitch_in = c->pitch_f; S.pitch_out = c->pitch_f; S.border = c->P.border; S.vec_ok = 1;
        S.lambdaG = c->P.lambdaG; S.chi_fuse = c->chi_fuse_f;
        S.depth_in = c->tmp_depth[hb]; S.sigma_in = c->tmp_sigma[hb]; S.grad = s->grad;
        S.depth_out = s->depth; S.sigma_out = s->sigma; S.depth_copy = s->depth_chk;
        S.vlist = s->vlist; S.vcount = &s->meta->valid_count;
        dim3 grd((c->W + K23_TW - 1) / K23_TW, (c->H + K23_TH - 1) / K23_TH);
        k23_intra<true, true><<<grd, K23_THREADS, K23_SMEM, c->s_main>>>(S);
    }
    CU(c, cudaGetLastError());
    CU(c, cudaEventRecord(s->ev_sd, c->s_main));
    record_maps(c, s);
    // the scratch planes (one allocation per buffer: depth | sigma) go back to all-zero for the next key frame that uses
    // buffers [hb] -- after the maps-ready record, so not on this key frame's dependency chain
    CU(c, cudaMemsetAsync(c->tmp_depth[hb], 0, 2 * fplane, c->s_main));
    c->tmp_dirty[hb] = false;
    CU(c, cudaEventRecord(c->ev_hypfree[hb], c->s_main));
    c->hypfree_valid[hb] = true;
    s->has_semidense = true;
    s->vlist_ok = true;
    s->vlist_from_import = false;
    s->has_checked = false;
    s->has_points = false;
    return SD_OK;
}

static int interkf_common(sd_ctx* c, int kf_id, const int* nbr_ids, int n, bool with_points) {
    if (!c || !nbr_ids || n < 1 || n > SD_MAX_NEIGHBOURS) return fail(c, SD_ERR_INVALID_ARG, "sd_interkf_check: bad argument (n=%d)", n);
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    if (!s->has_semidense) return fail(c, SD_ERR_NOT_READY, "key frame %d has no semi-dense maps", kf_id);
    CU(c, cudaSetDevice(c->device));
    CU(c, poll_pack(c, false));
    static thread_local InterArgs A;
    A.W = c->W; A.H = c->H; A.pitch_f = c->pitch_f; A.n = n; A.lambdaN = c->P.lambdaN;
    A.fx = s->K[0]; A.fy = s->K[1]; A.cx = s->K[2]; A.cy = s->K[3];
    A.chi_inter = c->chi_inter_f;
    A.depth_in = s->depth;
    A.depth_out = s->depth_chk;
    A.vlist = s->vlist;
    A.vcount = &s->meta->valid_count;
    const bool inplace = c->P.interkf_inplace != 0;
    c->dirty = true;
    const Slot* latest_sd = s;  // K4 reads the maps of kf and of its neighbours: wait for the last of them to be ready
    for (int j = 0; j < n; ++j) {
        Slot* nb = find(c, nbr_ids[j]);
        if (!nb) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown neighbour key frame %d", nbr_ids[j]);
        if (!nb->has_semidense) return fail(c, SD_ERR_NOT_READY, "neighbour key frame %d has no semi-dense maps", nbr_ids[j]);
        InterNbr& N = A.nb[j];
        geom::relative_pose(s->Rcw, s->tcw, nb->Rcw, nb->tcw, N.R, N.t);
        for (int k = 0; k < 3; ++k) { N.Rd2[k] = (double)N.R[6 + k]; N.td[k] = (double)N.t[k]; }
        N.fx = nb->K[0]; N.fy = nb->K[1]; N.cx = nb->K[2]; N.cy = nb->K[3];
        // in-place mode (PM.cc:764, 793): a neighbour that has been checked already is read as checked; its K4 ran
        // earlier on this same stream, so the read is ordered after it
        N.depth = (inplace && nb->has_checked && nb != s) ? nb->depth_chk : nb->depth;
        N.sigma = nb->sigma;
        if (nb->sd_seq > latest_sd->sd_seq) latest_sd = nb;
    }
    CU(c, wait_maps(c->s_chk, c->chk_seen_sd, latest_sd));
    if (s->dn_pending) {
        CU(c, cudaStreamWaitEvent(c->s_chk, s->ev_dn, 0));
        s->dn_pending = false;
    }
    if (!s->vlist_ok) {  // maps installed by sd_import_maps: rebuild the list and the plane k4 patches
        const size_t pb = (size_t)c->pitch_f * sizeof(float);
        CU(c, cudaMemsetAsync(&s->meta->valid_count, 0, sizeof(int), c->s_chk));
        CU(c, cudaMemcpy2DAsync(s->depth_chk, pb, s->depth, pb, (size_t)c->W * sizeof(float), c->H, cudaMemcpyDeviceToDevice, c->s_chk));
        dim3 grd((c->W + 31) / 32, (c->H + 7) / 8);
        k_build_vlist<<<grd, 256, 0, c->s_chk>>>(s->depth, s->sigma, c->pitch_f, c->W, c->H, s->vlist, &s->meta->valid_count);
        c->launches++;
        s->vlist_ok = true;
        s->vlist_from_import = true;
    } else if (s->has_checked) {  // a repeated check starts again from the intra-KF map
        const size_t pb = (size_t)c->pitch_f * sizeof(float);
        CU(c, cudaMemcpy2DAsync(s->depth_chk, pb, s->depth, pb, (size_t)c->W * sizeof(float), c->H, cudaMemcpyDeviceToDevice, c->s_chk));
    }
    if (with_points) {
        if (!s->points) CU(c, cudaMalloc(&s->points, (size_t)c->W * c->H * 3 * sizeof(float)));
        A.P.W = c->W; A.P.H = c->H; A.P.pitch_f = c->pitch_f;
        A.P.fx = s->K[0]; A.P.fy = s->K[1]; A.P.cx = s->K[2]; A.P.cy = s->K[3];
        std::memcpy(A.P.Twc, s->Twc, sizeof(A.P.Twc));
        A.P.depth = s->depth_chk;
        A.P.points = s->points;
        CU(c, cudaMemsetAsync(s->points, 0, (size_t)c->W * c->H * 3 * sizeof(float), c->s_chk));
    }
    {
        ScopedTimer t(c, c->s_chk, T_INTER);
        const int worst = (c->W - 4) * (c->H - 4);  // the list holds pixels inside the 2-px frame only
        if (with_points) k4_inter_check<true><<<(worst + 127) / 128, 128, 0, c->s_chk>>>(A);
        else k4_inter_check<false><<<(worst + 127) / 128, 128, 0, c->s_chk>>>(A);
    }
    CU(c, cudaGetLastError());
    {   // this launch read the maps of kf and of its neighbours: writers of those maps wait for it
        CU(c, record_chk(c, s));
        s->rd_idx = s->chk_idx;
        for (int j = 0; j < n; ++j) find(c, nbr_ids[j])->rd_idx = s->chk_idx;
    }
    s->has_checked = true;
    s->has_points = with_points;
    if (with_points) s->pose_changed = false;
    return SD_OK;
}

int sd_interkf_check(sd_ctx* c, int kf_id, const int* nbr_ids, int n) { return interkf_common(c, kf_id, nbr_ids, n, false); }

int sd_interkf_check_points(sd_ctx* c, int kf_id, const int* nbr_ids, int n) { return interkf_common(c, kf_id, nbr_ids, n, true); }

int sd_update_pointset(sd_ctx* c, int kf_id) {
    if (!c) return SD_ERR_INVALID_ARG;
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    if (!s->has_semidense) return fail(c, SD_ERR_NOT_READY, "key frame %d has no semi-dense maps", kf_id);
    CU(c, cudaSetDevice(c->device));
    if (!s->points) CU(c, cudaMalloc(&s->points, (size_t)c->W * c->H * 3 * sizeof(float)));
    c->dirty = true;
    CU(c, wait_maps(c->s_chk, c->chk_seen_sd, s));
    if (s->dn_pending) {
        CU(c, cudaStreamWaitEvent(c->s_chk, s->ev_dn, 0));
        s->dn_pending = false;
    }
    PointArgs A;
    A.W = c->W; A.H = c->H; A.pitch_f = c->pitch_f;
    A.fx = s->K[0]; A.fy = s->K[1]; A.cx = s->K[2]; A.cy = s->K[3];
    std::memcpy(A.Twc, s->Twc, sizeof(A.Twc));
    A.depth = s->has_checked ? s->depth_chk : s->depth;
    A.points = s->points;
    {
        ScopedTimer t(c, c->s_chk, T_POINTS);
        dim3 grd((c->W + 255) / 256, c->H);
        k5_pointset<<<grd, 256, 0, c->s_chk>>>(A);
    }
    CU(c, cudaGetLastError());
    CU(c, record_chk(c, s));
    s->has_points = true;
    s->pose_changed = false;
    return SD_OK;
}

int sd_confident_points_device(sd_ctx* c, int kf_id, double max_sigma, int max_points, const float** xyz_dev, const float** sigma_dev,
                               const uint32_t** pixel_dev, int* count, int* confident) {
    if (!c || !count) return SD_ERR_INVALID_ARG;
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    if (!s->has_semidense) return fail(c, SD_ERR_NOT_READY, "key frame %d has no semi-dense maps", kf_id);
    CU(c, cudaSetDevice(c->device));
    const int npix = c->W * c->H;
    const int cap = (max_points > 0 && max_points < npix) ? max_points : npix;
    const int nseg_row = (c->W + 255) / 256, nseg = nseg_row * c->H;
    if (!c->conf_state) {
        CU(c, cudaMalloc(&c->conf_state, sizeof(K6State)));
        CU(c, cudaMalloc(&c->conf_hist, 2 * 65536 * sizeof(uint32_t)));
        CU(c, cudaMalloc(&c->conf_seg, (size_t)4 * nseg * sizeof(int)));
        CU(c, cudaHostAlloc(&c->h_conf, 2 * sizeof(int), cudaHostAllocDefault));
    }
    if (c->conf_cap < cap) {
        CU(c, cudaStreamSynchronize(c->s_chk));
        cudaFree(c->conf_out);
        c->conf_out = nullptr;
        c->conf_cap = 0;
        c->conf_last_xyz = nullptr;
        CU(c, cudaMalloc(&c->conf_out, (size_t)cap * 5 * sizeof(float)));
        c->conf_cap = cap;
    }
    cudaStream_t st = c->s_chk;
    c->dirty = true;
    CU(c, wait_maps(st, c->chk_seen_sd, s));
    ConfArgs A;
    A.P.W = c->W; A.P.H = c->H; A.P.pitch_f = c->pitch_f;
    A.P.fx = s->K[0]; A.P.fy = s->K[1]; A.P.cx = s->K[2]; A.P.cy = s->K[3];
    std::memcpy(A.P.Twc, s->Twc, sizeof(A.P.Twc));
    A.P.depth = s->has_checked ? s->depth_chk : s->depth;
    A.P.points = nullptr;
    A.sigma = s->sigma;
    A.max_sigma = max_sigma;
    A.max_points = max_points;
    A.nseg_row = nseg_row; A.nseg = nseg;
    A.st = c->conf_state;
    A.hist = c->conf_hist;
    A.seg_lt = c->conf_seg; A.seg_eq = c->conf_seg + nseg; A.seg_off = c->conf_seg + 2 * nseg; A.seg_eqoff = c->conf_seg + 3 * nseg;
    A.out_xyz = (float*)c->conf_out;
    A.out_sigma = A.out_xyz + (size_t)3 * c->conf_cap;
    A.out_pix = (uint32_t*)(A.out_sigma + c->conf_cap);
    A.cap = cap;
    CU(c, cudaMemsetAsync(c->conf_state, 0, sizeof(K6State), st));
    const dim3 blk(256), grd((c->W + 31) / 32, (c->H + 7) / 8);
    CU(c, cudaMemsetAsync(c->conf_hist, 0, 2 * 65536 * sizeof(uint32_t), st));
    k6_hist16<0><<<grd, blk, 0, st>>>(A);
    k6_pick16<0><<<1, 1024, 0, st>>>(A);
    c->launches += 2;
    if (max_points > 0 && max_points < npix) {
        k6_hist16<1><<<grd, blk, 0, st>>>(A);
        k6_pick16<1><<<1, 1024, 0, st>>>(A);
        c->launches += 2;
    }
    k6_count<<<(nseg + 7) / 8, 256, 0, st>>>(A);
    k6_scan<<<1, 1024, 0, st>>>(A);
    k6_write<<<(nseg + 7) / 8, 256, 0, st>>>(A);
    c->launches += 3;
    CU(c, cudaGetLastError());
    CU(c, cudaMemcpyAsync(&c->h_conf[0], &c->conf_state->total, sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(c, cudaMemcpyAsync(&c->h_conf[1], &c->conf_state->confident, sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(c, record_chk(c, s));
    CU(c, cudaStreamSynchronize(st));
    *count = c->h_conf[0];
    c->conf_last_xyz = A.out_xyz; c->conf_last_sigma = A.out_sigma; c->conf_last_pix = A.out_pix; c->conf_last_count = c->h_conf[0];
    if (confident) *confident = c->h_conf[1];
    if (xyz_dev) *xyz_dev = A.out_xyz;
    if (sigma_dev) *sigma_dev = A.out_sigma;
    if (pixel_dev) *pixel_dev = A.out_pix;
    return SD_OK;
}

int sd_confident_points(sd_ctx* c, int kf_id, double max_sigma, int max_points, int capacity, float* xyz, float* sigma, uint32_t* pixel,
                        int* count) {
    if (!c || !count || capacity < 0) return SD_ERR_INVALID_ARG;
    const float *dx = nullptr, *ds = nullptr;
    const uint32_t* dp = nullptr;
    int rc = sd_confident_points_device(c, kf_id, max_sigma, max_points, &dx, &ds, &dp, count, nullptr);
    if (rc != SD_OK) return rc;
    int n = 0;
    return sd_confident_points_fetch(c, capacity, xyz, sigma, pixel, &n);
}

int sd_confident_points_fetch(sd_ctx* c, int capacity, float* xyz, float* sigma, uint32_t* pixel, int* count) {
    if (!c || !count || capacity < 0) return SD_ERR_INVALID_ARG;
    if (!c->conf_last_xyz) return fail(c, SD_ERR_NOT_READY, "sd_confident_points_fetch: no device-resident result");
    CU(c, cudaSetDevice(c->device));
    *count = c->conf_last_count;
    const size_t n = (size_t)std::min(c->conf_last_count, capacity);
    if (n) {
        if (xyz) CU(c, cudaMemcpyAsync(xyz, c->conf_last_xyz, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, c->s_chk));
        if (sigma) CU(c, cudaMemcpyAsync(sigma, c->conf_last_sigma, n * sizeof(float), cudaMemcpyDeviceToHost, c->s_chk));
        if (pixel) CU(c, cudaMemcpyAsync(pixel, c->conf_last_pix, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->s_chk));
        CU(c, cudaStreamSynchronize(c->s_chk));
    }
    return SD_OK;
}

int sd_download_async(sd_ctx* c, int kf_id, float* depth, float* sigma, float* points) {
    if (!c) return SD_ERR_INVALID_ARG;
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    if (!s->has_semidense) return fail(c, SD_ERR_NOT_READY, "key frame %d has no semi-dense maps", kf_id);
    if (points && !s->has_points) return fail(c, SD_ERR_NOT_READY, "key frame %d has no point set", kf_id);
    CU(c, cudaSetDevice(c->device));
    c->dirty = true;
    CU(c, wait_maps(c->s_dn, c->dn_seen_sd, s));
    CU(c, wait_ring(c, c->s_dn, c->dn_seen_chk, s->chk_idx));
    const size_t wb = (size_t)c->W * sizeof(float), pb = (size_t)c->pitch_f * sizeof(float);
    if (depth) CU(c, cudaMemcpy2DAsync(depth, wb, s->has_checked ? s->depth_chk : s->depth, pb, wb, c->H, cudaMemcpyDeviceToHost, c->s_dn));
    if (sigma) CU(c, cudaMemcpy2DAsync(sigma, wb, s->sigma, pb, wb, c->H, cudaMemcpyDeviceToHost, c->s_dn));
    if (points) CU(c, cudaMemcpyAsync(points, s->points, (size_t)c->W * c->H * 3 * sizeof(float), cudaMemcpyDeviceToHost, c->s_dn));
    CU(c, cudaEventRecord(s->ev_dn, c->s_dn));
    s->dn_pending = true;
    return SD_OK;
}

int sd_download(sd_ctx* c, int kf_id, float* depth, float* sigma, float* points) {
    int rc = sd_download_async(c, kf_id, depth, sigma, points);
    if (rc != SD_OK) return rc;
    CU(c, cudaStreamSynchronize(c->s_dn));
    return SD_OK;
}


int sd_download_compact_async(sd_ctx* c, int kf_id, int with_points, int capacity, void* records, int* count) {
    if (!c || capacity < 0 || (capacity > 0 && !records)) return SD_ERR_INVALID_ARG;
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    if (!s->has_semidense) return fail(c, SD_ERR_NOT_READY, "key frame %d has no semi-dense maps", kf_id);
    if (!s->vlist_ok || s->vlist_from_import)
        return fail(c, SD_ERR_NOT_READY, "key frame %d: maps were imported, no compact form (use sd_download)", kf_id);
    if (with_points && !s->has_points) return fail(c, SD_ERR_NOT_READY, "key frame %d has no point set", kf_id);
    CU(c, cudaSetDevice(c->device));
    CU(c, poll_pack(c, false));
    if (c->pack_tail - c->pack_head == sd_ctx::PACK_RING) {  // ring full: the oldest pending download must go out first
        sd_ctx::PackSlot& old = c->pack[c->pack_head % sd_ctx::PACK_RING];
        CU(c, cudaEventSynchronize(old.ev_cnt));
        CU(c, pack_issue(c, c->pack_head));
        c->pack_head++;
    }
    const int cap = (c->W - 4) * (c->H - 4);  // the list holds pixels inside the 2-px frame only
    sd_ctx::PackSlot& ps = c->pack[c->pack_tail % sd_ctx::PACK_RING];
    if (!ps.dev) {
        CU(c, cudaMalloc(&ps.dev, 256 + (size_t)cap * sizeof(SdCompactPoint)));
        CU(c, cudaHostAlloc(&ps.h_count, sizeof(int), cudaHostAllocDefault));
        CU(c, cudaEventCreateWithFlags(&ps.ev_cnt, cudaEventDisableTiming));
    }
    // stream order on s_dn: the copy of this slot's previous use precedes the kernel below
    c->dirty = true;
    CU(c, wait_maps(c->s_dn, c->dn_seen_sd, s));
    CU(c, wait_ring(c, c->s_dn, c->dn_seen_chk, s->chk_idx));
    PackArgs A;
    A.W = c->W; A.pitch_f = c->pitch_f;
    A.vlist = s->vlist; A.vcount = &s->meta->valid_count;
    A.depth = s->has_checked ? s->depth_chk : s->depth;
    A.sigma = s->sigma;
    A.points = with_points ? s->points : nullptr;
    A.with_points = with_points ? 1 : 0;
    A.out = ps.dev + 256;
    A.out_count = reinterpret_cast<int*>(ps.dev);
    A.cap = cap;
    k7_pack<<<(cap + 255) / 256, 256, 0, c->s_dn>>>(A);
    c->launches++;
    CU(c, cudaGetLastError());
    CU(c, cudaMemcpyAsync(ps.h_count, ps.dev, sizeof(int), cudaMemcpyDeviceToHost, c->s_dn));
    CU(c, cudaEventRecord(ps.ev_cnt, c->s_dn));
    CU(c, cudaEventRecord(s->ev_dn, c->s_dn));  // the planes have been read once the kernel is done
    s->dn_pending = true;
    ps.host_records = records;
    ps.capacity = capacity;
    ps.stride = with_points ? (int)sizeof(SdCompactPoint) : (int)sizeof(SdCompactDepth);
    ps.count_out = count;
    c->pack_tail++;
    return SD_OK;
}

int sd_download_compact(sd_ctx* c, int kf_id, int with_points, int capacity, void* records, int* count) {
    int rc = sd_download_compact_async(c, kf_id, with_points, capacity, records, count);
    if (rc != SD_OK) return rc;
    CU(c, poll_pack(c, true));
    CU(c, cudaStreamSynchronize(c->s_dn));
    return SD_OK;
}

int sd_download_plane(sd_ctx* c, int kf_id, int which, float* out) {
    if (!c || !out) return SD_ERR_INVALID_ARG;
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    CU(c, cudaSetDevice(c->device));
    const float* src = nullptr;
    switch (which) {
        case 0: src = s->grad; break;
        case 1: src = s->theta; break;
        case 2: src = s->has_semidense ? s->depth : nullptr; break;
        case 3: src = s->has_semidense ? s->sigma : nullptr; break;
        case 4: src = s->has_checked ? s->depth_chk : nullptr; break;
        case 5: src = s->has_points ? s->points : nullptr; break;
        default: return fail(c, SD_ERR_INVALID_ARG, "sd_download_plane: which=%d", which);
    }
    if (!src) return fail(c, SD_ERR_NOT_READY, "plane %d of key frame %d not computed", which, kf_id);
    CU(c, cudaStreamSynchronize(c->s_up));
    CU(c, cudaStreamSynchronize(c->s_main));
    CU(c, cudaStreamSynchronize(c->s_chk));
    const size_t wb = (size_t)c->W * sizeof(float), pb = (size_t)c->pitch_f * sizeof(float);
    if (which == 5) CU(c, cudaMemcpy(out, src, (size_t)c->W * c->H * 3 * sizeof(float), cudaMemcpyDeviceToHost));
    else CU(c, cudaMemcpy2D(out, wb, src, pb, wb, c->H, cudaMemcpyDeviceToHost));
    return SD_OK;
}

int sd_get_istddev(sd_ctx* c, int kf_id, float* out) {
    if (!c || !out) return SD_ERR_INVALID_ARG;
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->s_up));
    KfMeta m;
    CU(c, cudaMemcpy(&m, s->meta, sizeof(m), cudaMemcpyDeviceToHost));
    *out = m.istddev;
    return SD_OK;
}

int sd_get_active_count(sd_ctx* c, int kf_id, int* out) {
    if (!c || !out) return SD_ERR_INVALID_ARG;
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->s_up));
    KfMeta m;
    CU(c, cudaMemcpy(&m, s->meta, sizeof(m), cudaMemcpyDeviceToHost));
    *out = m.active_count;
    return SD_OK;
}

int sd_device_ptrs(sd_ctx* c, int kf_id, SdDevicePlanes* out) {
    if (!c || !out) return SD_ERR_INVALID_ARG;
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    out->width = c->W; out->height = c->H; out->pitch_f32 = c->pitch_f; out->pitch_u8 = c->pitch_u8;
    out->gray = s->gray; out->grad = s->grad; out->theta = s->theta;
    out->depth = s->has_semidense ? s->depth : nullptr;
    out->sigma = s->has_semidense ? s->sigma : nullptr;
    out->depth_checked = s->has_checked ? s->depth_chk : nullptr;
    out->points = s->has_points ? s->points : nullptr;
    return SD_OK;
}

int sd_import_maps(sd_ctx* c, int kf_id, const float* depth_dev, const float* sigma_dev) {
    if (!c || !depth_dev || !sigma_dev) return SD_ERR_INVALID_ARG;
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    CU(c, cudaSetDevice(c->device));
    const size_t wb = (size_t)c->W * sizeof(float), pb = (size_t)c->pitch_f * sizeof(float);
    // a re-upload of this key frame (or a recycled slot) forgets the slot's reader bookkeeping and makes only s_up
    // wait for the work still queued on the other streams: order the overwrite of the maps after that upload
    c->dirty = true;
    CU(c, wait_upload(c->s_main, c->main_seen_up, s));
    CU(c, wait_map_readers(c, s));
    if (s->dn_pending) {
        CU(c, cudaStreamWaitEvent(c->s_main, s->ev_dn, 0));
        s->dn_pending = false;
    }
    CU(c, cudaMemcpy2DAsync(s->depth, pb, depth_dev, wb, wb, c->H, cudaMemcpyDeviceToDevice, c->s_main));
    CU(c, cudaMemcpy2DAsync(s->sigma, pb, sigma_dev, wb, wb, c->H, cudaMemcpyDeviceToDevice, c->s_main));
    CU(c, cudaEventRecord(s->ev_sd, c->s_main));
    record_maps(c, s);
    s->has_semidense = true;
    s->has_checked = false;
    s->has_points = false;  // the point set on the device belongs to the maps just replaced
    s->vlist_ok = false;
    return SD_OK;
}

int sd_import_maps_host(sd_ctx* c, int kf_id, const float* depth, const float* sigma) {
    if (!c || !depth || !sigma) return SD_ERR_INVALID_ARG;
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    CU(c, cudaSetDevice(c->device));
    const size_t wb = (size_t)c->W * sizeof(float), pb = (size_t)c->pitch_f * sizeof(float);
    c->dirty = true;
    CU(c, wait_upload(c->s_main, c->main_seen_up, s));
    CU(c, wait_map_readers(c, s));
    if (s->dn_pending) {
        CU(c, cudaStreamWaitEvent(c->s_main, s->ev_dn, 0));
        s->dn_pending = false;
    }
    CU(c, cudaMemcpy2DAsync(s->depth, pb, depth, wb, wb, c->H, cudaMemcpyHostToDevice, c->s_main));
    CU(c, cudaMemcpy2DAsync(s->sigma, pb, sigma, wb, wb, c->H, cudaMemcpyHostToDevice, c->s_main));
    CU(c, cudaEventRecord(s->ev_sd, c->s_main));
    record_maps(c, s);
    CU(c, cudaStreamSynchronize(c->s_main));  // the host buffers are borrowed for the call only
    s->has_semidense = true;
    s->has_checked = false;
    s->has_points = false;
    s->vlist_ok = false;
    return SD_OK;
}

// ---- NCCL halo exchange (dlopen; minimal local declarations of the NCCL C API) ----
namespace {
struct NcclUniqueId { char internal[128]; };
typedef int (*nccl_get_unique_id_t)(NcclUniqueId*);
typedef int (*nccl_comm_init_rank_t)(void**, int, NcclUniqueId, int);
typedef int (*nccl_comm_destroy_t)(void*);
typedef int (*nccl_group_t)(void);
typedef int (*nccl_send_t)(const void*, size_t, int, int, void*, cudaStream_t);
typedef int (*nccl_recv_t)(void*, size_t, int, int, void*, cudaStream_t);
typedef const char* (*nccl_err_t)(int);
const int kNcclFloat32 = 7;  // ncclFloat32 in ncclDataType_t

struct NcclApi {
    void* lib = nullptr;
    nccl_get_unique_id_t get_unique_id = nullptr;
    nccl_comm_init_rank_t comm_init_rank = nullptr;
    nccl_comm_destroy_t comm_destroy = nullptr;
    nccl_group_t group_start = nullptr, group_end = nullptr;
    nccl_send_t send = nullptr;
    nccl_recv_t recv = nullptr;
    nccl_err_t err = nullptr;
};

NcclApi* nccl_api() {
    static NcclApi api;
    static std::once_flag once;  // contexts may be created from several threads: one dlopen, finished before anyone reads `api`
    std::call_once(once, [] {
        // a process that already loaded NCCL (torch ships its own libnccl.so.2) gets that copy back
        const char* names[] = {getenv("SDPM_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) {
            if (!nm || !*nm) continue;
            api.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
        }
        if (api.lib) {
            api.get_unique_id = (nccl_get_unique_id_t)dlsym(api.lib, "ncclGetUniqueId");
            api.comm_init_rank = (nccl_comm_init_rank_t)dlsym(api.lib, "ncclCommInitRank");
            api.comm_destroy = (nccl_comm_destroy_t)dlsym(api.lib, "ncclCommDestroy");
            api.group_start = (nccl_group_t)dlsym(api.lib, "ncclGroupStart");
            api.group_end = (nccl_group_t)dlsym(api.lib, "ncclGroupEnd");
            api.send = (nccl_send_t)dlsym(api.lib, "ncclSend");
            api.recv = (nccl_recv_t)dlsym(api.lib, "ncclRecv");
            api.err = (nccl_err_t)dlsym(api.lib, "ncclGetErrorString");
            if (!api.get_unique_id || !api.comm_init_rank || !api.comm_destroy || !api.group_start || !api.group_end || !api.send ||
                !api.recv)
                api.lib = nullptr;
        }
    });
    return api.lib ? &api : nullptr;
}

#define NC(c, api, call)                                                                                  \
    do {                                                                                                  \
        int r__ = (call);                                                                                 \
        if (r__ != 0) return fail(c, SD_ERR_CUDA, "%s: NCCL error %d (%s)", #call, r__, (api)->err ? (api)->err(r__) : "?"); \
    } while (0)
}  // namespace

int sd_nccl_unique_id(void* id128) {
    if (!id128) return SD_ERR_INVALID_ARG;
    NcclApi* api = nccl_api();
    if (!api) return SD_ERR_NOT_READY;
    NcclUniqueId id;
    if (api->get_unique_id(&id) != 0) return SD_ERR_CUDA;
    std::memcpy(id128, &id, sizeof(id));
    return SD_OK;
}

int sd_comm_init(sd_ctx* c, int world, int rank, const void* id128) {
    if (!c || !id128 || world < 1 || rank < 0 || rank >= world) return fail(c, SD_ERR_INVALID_ARG, "sd_comm_init: bad argument");
    NcclApi* api = nccl_api();
    if (!api) return fail(c, SD_ERR_NOT_READY, "sd_comm_init: libnccl.so.2 not found (set SDPM_NCCL_LIB)");
    if (c->nccl_comm) return fail(c, SD_ERR_INVALID_ARG, "sd_comm_init: the context already has a communicator");
    CU(c, cudaSetDevice(c->device));
    NcclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    void* comm = nullptr;
    NC(c, api, api->comm_init_rank(&comm, world, id, rank));
    c->nccl_comm = comm;
    c->nccl_world = world;
    c->nccl_rank = rank;
    return SD_OK;
}

int sd_comm_destroy(sd_ctx* c) {
    if (!c) return SD_ERR_INVALID_ARG;
    if (c->nccl_comm) {
        NcclApi* api = nccl_api();
        cudaSetDevice(c->device);
        cudaStreamSynchronize(c->s_main);
        if (api) api->comm_destroy(c->nccl_comm);
        c->nccl_comm = nullptr;
    }
    return SD_OK;
}

int sd_halo_exchange(sd_ctx* c, const SdHaloPeer* peers, int n_peers) {
    if (!c || n_peers < 0 || (n_peers > 0 && !peers)) return fail(c, SD_ERR_INVALID_ARG, "sd_halo_exchange: bad argument");
    if (n_peers == 0) return SD_OK;
    if (!c->nccl_comm) return fail(c, SD_ERR_NOT_READY, "sd_halo_exchange: no communicator (sd_comm_init)");
    NcclApi* api = nccl_api();
    CU(c, cudaSetDevice(c->device));
    c->dirty = true;
    const size_t plane = (size_t)c->pitch_f * c->H;  // whole pitched planes: every rank of a run has the same geometry
    // order the receives after everything that may still read (or is still producing) the planes they overwrite
    for (int p = 0; p < n_peers; ++p) {
        const SdHaloPeer& P = peers[p];
        if (P.peer < 0 || P.peer >= c->nccl_world || P.peer == c->nccl_rank || P.n_send < 0 || P.n_recv < 0 ||
            (P.n_send && !P.send_ids) || (P.n_recv && !P.recv_ids))
            return fail(c, SD_ERR_INVALID_ARG, "sd_halo_exchange: bad peer entry %d", p);
        for (int i = 0; i < P.n_send; ++i) {
            Slot* s = find(c, P.send_ids[i]);
            if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", P.send_ids[i]);
            if (!s->has_semidense) return fail(c, SD_ERR_NOT_READY, "key frame %d has no semi-dense maps to send", P.send_ids[i]);
            // the maps to send were produced on s_main itself (k23_intra / an earlier import): already in stream order
        }
        for (int i = 0; i < P.n_recv; ++i) {
            Slot* s = find(c, P.recv_ids[i]);
            if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", P.recv_ids[i]);
            CU(c, wait_upload(c->s_main, c->main_seen_up, s));
            CU(c, wait_map_readers(c, s));
            if (s->dn_pending) {
                CU(c, cudaStreamWaitEvent(c->s_main, s->ev_dn, 0));
                s->dn_pending = false;
            }
        }
    }
    NC(c, api, api->group_start());
    // an error inside the bracket must not return with the group still open on this thread (the peers would block in
    // their grouped receives): stop queueing, close the group, then report the first error
    int nerr = 0;
    const char* nwhat = "";
    for (int p = 0; p < n_peers && !nerr; ++p) {
        const SdHaloPeer& P = peers[p];
        for (int i = 0; i < P.n_send && !nerr; ++i) {
            Slot* s = find(c, P.send_ids[i]);
            nerr = api->send(s->depth, plane, kNcclFloat32, P.peer, c->nccl_comm, c->s_main);
            if (!nerr) nerr = api->send(s->sigma, plane, kNcclFloat32, P.peer, c->nccl_comm, c->s_main);
            if (nerr) nwhat = "ncclSend";
        }
        for (int i = 0; i < P.n_recv && !nerr; ++i) {
            Slot* s = find(c, P.recv_ids[i]);
            nerr = api->recv(s->depth, plane, kNcclFloat32, P.peer, c->nccl_comm, c->s_main);
            if (!nerr) nerr = api->recv(s->sigma, plane, kNcclFloat32, P.peer, c->nccl_comm, c->s_main);
            if (nerr) nwhat = "ncclRecv";
        }
    }
    const int gerr = api->group_end();
    if (nerr) return fail(c, SD_ERR_CUDA, "sd_halo_exchange: %s: NCCL error %d (%s)", nwhat, nerr, api->err ? api->err(nerr) : "?");
    if (gerr) return fail(c, SD_ERR_CUDA, "sd_halo_exchange: ncclGroupEnd: NCCL error %d (%s)", gerr, api->err ? api->err(gerr) : "?");
    for (int p = 0; p < n_peers; ++p) {
        const SdHaloPeer& P = peers[p];
        for (int i = 0; i < P.n_recv; ++i) {
            Slot* s = find(c, P.recv_ids[i]);
            CU(c, cudaEventRecord(s->ev_sd, c->s_main));
            record_maps(c, s);
            s->has_semidense = true;
            s->has_checked = false;
            s->has_points = false;
            s->vlist_ok = false;
        }
        // the sends read our boundary maps on s_main: a later sd_semidense of those key frames (next pass) overwrites
        // them on s_main as well, i.e. in stream order after the sends
    }
    return SD_OK;
}

int sd_semidense_loop(sd_ctx* c, const SdLoopItem* items, int n_items, int flags, int* inter_done) {
    if (!c || n_items < 0 || (n_items > 0 && !items)) return fail(c, SD_ERR_INVALID_ARG, "sd_semidense_loop: bad argument");
    // position of every item's key frame in the list, to find when an item's neighbours are all queued
    std::unordered_map<int, int> pos;
    pos.reserve((size_t)n_items * 2);
    for (int i = 0; i < n_items; ++i) {
        if (!items[i].nbr_ids || items[i].n < 1 || items[i].n > SD_MAX_NEIGHBOURS)
            return fail(c, SD_ERR_INVALID_ARG, "sd_semidense_loop: item %d: bad neighbour list", i);
        pos[items[i].kf_id] = i;
    }
    std::vector<int> ready_at(n_items, -1);  // index of the last item whose sd_semidense must be queued first; -2: skip
    for (int i = 0; i < n_items; ++i) {
        if (inter_done) inter_done[i] = 0;
        int last = i;
        for (int j = 0; (flags & SD_LOOP_INTERKF) && j < items[i].n && last != -2; ++j) {
            auto it = pos.find(items[i].nbr_ids[j]);
            if (it != pos.end()) last = std::max(last, it->second);
            else {
                Slot* nb = find(c, items[i].nbr_ids[j]);
                if (!nb || !nb->has_semidense) last = -2;  // PM.cc:289-297: a neighbour without depth -> not checked in this pass
            }
        }
        ready_at[i] = last;
    }
    auto finish = [&](int i) -> int {
        const SdLoopItem& it = items[i];
        if ((flags & SD_LOOP_INTERKF) && ready_at[i] != -2) {
            int rc = sd_interkf_check_points(c, it.kf_id, it.nbr_ids, it.n);
            if (rc != SD_OK) return rc;
            if (inter_done) inter_done[i] = 1;
        }
        switch (it.download) {
            case SD_DL_NONE: return SD_OK;
            case SD_DL_DENSE: return sd_download_async(c, it.kf_id, it.out_depth, it.out_sigma, nullptr);
            case SD_DL_DENSE_POINTS: return sd_download_async(c, it.kf_id, it.out_depth, it.out_sigma, it.out_points);
            case SD_DL_COMPACT: return sd_download_compact_async(c, it.kf_id, 0, it.capacity, it.out_records, it.out_count);
            case SD_DL_COMPACT_POINTS: return sd_download_compact_async(c, it.kf_id, 1, it.capacity, it.out_records, it.out_count);
            default: return fail(c, SD_ERR_INVALID_ARG, "sd_semidense_loop: item %d: download mode %d", i, it.download);
        }
    };
    // items in list order; the finishing stages of item j are queued right after the sd_semidense of the last item it
    // waits for (so downloads and checks overlap the remaining searches); a pending pointer walks the list in order,
    // like PM.cc:262's loop
    // In-place mode (SdParams.interkf_inplace) makes the checks order-dependent: there they are issued strictly in list
    // order once every search has been queued, which is the order of PM.cc:262's loop over vpKFs.
    const bool in_order = (flags & SD_LOOP_INTERKF) && c->P.interkf_inplace != 0;
    std::vector<std::vector<int>> due(n_items);
    std::vector<int> skipped, ordered;
    for (int i = 0; i < n_items; ++i) {
        if (ready_at[i] < 0) skipped.push_back(i);
        else if (in_order) ordered.push_back(i);
        else due[ready_at[i]].push_back(i);
    }
    for (int k = 0; k < n_items; ++k) {
        const SdLoopItem& it = items[k];
        int rc = SD_OK;
        if (!it.finish_only) rc = sd_semidense(c, it.kf_id, it.nbr_ids, it.nbr_rot_deg, it.n, it.min_invdepth, it.max_invdepth);
        else if (!find(c, it.kf_id) || !find(c, it.kf_id)->has_semidense)
            rc = fail(c, SD_ERR_NOT_READY, "sd_semidense_loop: item %d (key frame %d) is finish_only but has no maps", k, it.kf_id);
        if (rc != SD_OK) return rc;
        for (int i : due[k]) {
            rc = finish(i);
            if (rc != SD_OK) return rc;
        }
    }
    for (int i : ordered) {
        int rc = finish(i);
        if (rc != SD_OK) return rc;
    }
    for (int i : skipped) {  // no inter-key-frame stage in this pass; a requested download still delivers the intra maps
        if (items[i].download == SD_DL_NONE || items[i].download == SD_DL_DENSE_POINTS) continue;
        if (items[i].download == SD_DL_COMPACT_POINTS) {  // no point set exists: say so instead of leaving a stale count behind
            if (items[i].out_count) *items[i].out_count = 0;
            continue;
        }
        int rc = finish(i);
        if (rc != SD_OK) return rc;
    }
    return SD_OK;
}

int sd_export_maps(sd_ctx* c, int kf_id, float* depth_dev, float* sigma_dev) {
    if (!c || !depth_dev || !sigma_dev) return SD_ERR_INVALID_ARG;
    Slot* s = find(c, kf_id);
    if (!s) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d", kf_id);
    if (!s->has_semidense) return fail(c, SD_ERR_NOT_READY, "key frame %d has no semi-dense maps", kf_id);
    CU(c, cudaSetDevice(c->device));
    const size_t wb = (size_t)c->W * sizeof(float), pb = (size_t)c->pitch_f * sizeof(float);
    c->dirty = true;
    CU(c, cudaMemcpy2DAsync(depth_dev, wb, s->depth, pb, wb, c->H, cudaMemcpyDeviceToDevice, c->s_main));
    CU(c, cudaMemcpy2DAsync(sigma_dev, wb, s->sigma, pb, wb, c->H, cudaMemcpyDeviceToDevice, c->s_main));
    return SD_OK;
}

int sd_alloc_host(sd_ctx* c, uint64_t bytes, void** out) {
    if (!c || !out || bytes == 0) return fail(c, SD_ERR_INVALID_ARG, "sd_alloc_host: bad argument");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaHostAlloc(out, (size_t)bytes, cudaHostAllocDefault));
    return SD_OK;
}

int sd_free_host(sd_ctx* c, void* ptr) {
    if (!c) return SD_ERR_INVALID_ARG;
    if (!ptr) return SD_OK;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaFreeHost(ptr));
    return SD_OK;
}

int sd_download_fence(sd_ctx* c, uint64_t* ticket) {
    if (!c || !ticket) return SD_ERR_INVALID_ARG;
    *ticket = (uint64_t)c->pack_tail;
    return SD_OK;
}

int sd_download_wait(sd_ctx* c, uint64_t ticket) {
    if (!c || (long long)ticket > c->pack_tail) return fail(c, SD_ERR_INVALID_ARG, "sd_download_wait: ticket from the future");
    CU(c, cudaSetDevice(c->device));
    // compact downloads are two-step (count, then exactly that many records): issue the copies of everything up to the
    // ticket, in order, as their counts arrive
    while (c->pack_head < (long long)ticket) {
        sd_ctx::PackSlot& ps = c->pack[c->pack_head % sd_ctx::PACK_RING];
        CU(c, cudaEventSynchronize(ps.ev_cnt));
        CU(c, pack_issue(c, c->pack_head));
        c->pack_head++;
    }
    // every download runs on the one download stream, in order: an event behind the last copy covers the dense ones too
    if (!c->ev_fence) CU(c, cudaEventCreateWithFlags(&c->ev_fence, cudaEventDisableTiming));
    CU(c, cudaEventRecord(c->ev_fence, c->s_dn));
    CU(c, cudaEventSynchronize(c->ev_fence));
    return SD_OK;
}

int sd_sync(sd_ctx* c) {
    if (!c) return SD_ERR_INVALID_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, poll_pack(c, true));
    CU(c, cudaStreamSynchronize(c->s_up));
    CU(c, cudaStreamSynchronize(c->s_srch));
    CU(c, cudaStreamSynchronize(c->s_main));
    CU(c, cudaStreamSynchronize(c->s_chk));
    CU(c, cudaStreamSynchronize(c->s_dn));
    for (auto& kv : c->kfs) kv.second->dn_pending = false;
    if (c->timing) collect_spans(c);
    return SD_OK;
}

const char* sd_last_error(sd_ctx* c) { return c ? c->err.c_str() : "null context"; }

int sd_compute_stream(sd_ctx* c, void** stream) {
    if (!c || !stream) return SD_ERR_INVALID_ARG;
    *stream = (void*)c->s_main;
    return SD_OK;
}

int sd_launch_count(sd_ctx* c, uint64_t* out) {
    if (!c || !out) return SD_ERR_INVALID_ARG;
    *out = c->launches;
    return SD_OK;
}

int sd_span_begin(sd_ctx* c) {
    int rc = sd_sync(c);
    if (rc != SD_OK) return rc;
    CU(c, cudaEventRecord(c->ev_span_a, c->s_main));
    // later work on the other two streams must not start before the span does
    CU(c, cudaStreamWaitEvent(c->s_up, c->ev_span_a, 0));
    CU(c, cudaStreamWaitEvent(c->s_srch, c->ev_span_a, 0));
    CU(c, cudaStreamWaitEvent(c->s_chk, c->ev_span_a, 0));
    CU(c, cudaStreamWaitEvent(c->s_dn, c->ev_span_a, 0));
    return SD_OK;
}

int sd_span_end(sd_ctx* c, double* ms) {
    if (!c || !ms) return SD_ERR_INVALID_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, poll_pack(c, true));  // the copies of pending compact downloads belong to the span
    CU(c, cudaEventRecord(c->ev_tmp, c->s_up));
    CU(c, cudaStreamWaitEvent(c->s_main, c->ev_tmp, 0));
    CU(c, cudaEventRecord(c->ev_tmp, c->s_srch));
    CU(c, cudaStreamWaitEvent(c->s_main, c->ev_tmp, 0));
    CU(c, cudaEventRecord(c->ev_tmp, c->s_chk));
    CU(c, cudaStreamWaitEvent(c->s_main, c->ev_tmp, 0));
    CU(c, cudaEventRecord(c->ev_tmp, c->s_dn));
    CU(c, cudaStreamWaitEvent(c->s_main, c->ev_tmp, 0));
    CU(c, cudaEventRecord(c->ev_span_b, c->s_main));
    int rc = sd_sync(c);
    if (rc != SD_OK) return rc;
    float f = 0;
    CU(c, cudaEventElapsedTime(&f, c->ev_span_a, c->ev_span_b));
    *ms = f;
    return SD_OK;
}

int sd_timing_enable(sd_ctx* c, int enable) {
    if (!c) return SD_ERR_INVALID_ARG;
    c->timing = enable != 0;
    c->timing_serial = enable == 2;
    return SD_OK;
}

int sd_timing_get(sd_ctx* c, const char* name, double* ms, uint64_t* launches) {
    if (!c || !name) return SD_ERR_INVALID_ARG;
    int rc = sd_sync(c);
    if (rc != SD_OK) return rc;
    collect_spans(c);
    for (int i = 0; i < T_COUNT; ++i)
        if (!std::strcmp(name, kTimerNames[i])) {
            if (ms) *ms = c->t_ms[i];
            if (launches) *launches = c->t_n[i];
            return SD_OK;
        }
    return fail(c, SD_ERR_INVALID_ARG, "unknown timer %s", name);
}

int sd_timing_reset(sd_ctx* c) {
    if (!c) return SD_ERR_INVALID_ARG;
    int rc = sd_sync(c);
    if (rc != SD_OK) return rc;
    collect_spans(c);
    for (int i = 0; i < T_COUNT; ++i) { c->t_ms[i] = 0; c->t_n[i] = 0; }
    return SD_OK;
}

// ---- host-side helpers ----

int sd_stereo_search_constraints(const float* d, int n, float* min_invdepth, float* max_invdepth) {
    if (!d || n < 1 || !min_invdepth || !max_invdepth) return SD_ERR_INVALID_ARG;
    // PM.cc:373-382.  std::accumulate / std::inner_product are seeded with 0.0: double accumulators.
    double acc = 0.0;
    for (int i = 0; i < n; ++i) acc = acc + d[i];
    const float sum = (float)acc;
    const float mean = sum / (float)(size_t)n;
    double vacc = 0.0;
    for (int i = 0; i < n; ++i) {
        const float df = d[i] - mean;
        vacc = vacc + (df * df);
    }
    const float variance = (float)(vacc / (double)(size_t)n);
    const float stdev = std::sqrt(variance);
    *max_invdepth = 1 / (mean + 2 * stdev);
    *min_invdepth = 1 / (mean - 2 * stdev);
    return SD_OK;
}

float sd_rot_in_plane_median(const long long* mp1, const float* ang1, int n1, const long long* mp2, const float* ang2, int n2) {
    // PM.cc:467-484: every (idx1, idx2) pair observing the same map point contributes angle2-angle1
    // unless either angle is negative; then the lower median, PM.cc:172-177.  The pair set is the
    // same as the reference's O(n1*n2) pointer scan; it is built through a hash of kf2's points.
    std::unordered_multimap<long long, int> by_point;
    by_point.reserve((size_t)n2 * 2);
    for (int i = 0; i < n2; ++i)
        if (mp2[i] >= 0) by_point.emplace(mp2[i], i);
    std::vector<float> rot;
    for (int i = 0; i < n1; ++i) {
        if (mp1[i] < 0) continue;
        auto range = by_point.equal_range(mp1[i]);
        for (auto it = range.first; it != range.second; ++it) {
            const float a1 = ang1[i], a2 = ang2[it->second];
            if (a1 < 0 || a2 < 0) continue;
            rot.push_back(a2 - a1);
        }
    }
    if (rot.empty()) return 0.f;
    std::sort(rot.begin(), rot.end());
    return rot[(rot.size() - 1) / 2];
}

int sd_pair_geometry(const float K1[4], const float Tcw1[12], const float K2[4], const float Tcw2[12], float F12[9], float R21[9],
                     float t21[3]) {
    if (!K1 || !Tcw1 || !K2 || !Tcw2 || !F12 || !R21 || !t21) return SD_ERR_INVALID_ARG;
    float R1[9], t1[3], R2[9], t2[3];
    for (int r = 0; r < 3; ++r) {
        for (int col = 0; col < 3; ++col) { R1[r * 3 + col] = Tcw1[r * 4 + col]; R2[r * 3 + col] = Tcw2[r * 4 + col]; }
        t1[r] = Tcw1[r * 4 + 3];
        t2[r] = Tcw2[r * 4 + 3];
    }
    geom::fundamental(K1, R1, t1, K2, R2, t2, F12);
    geom::relative_pose(R1, t1, R2, t2, R21, t21);
    return SD_OK;
}

int sd_search_range(const float K1[4], const float Tcw1[12], const float Tcw2[12], int cols, int px, int py, float mind, float maxd,
                    float* umin_out, float* umax_out) {
    if (!K1 || !Tcw1 || !Tcw2 || !umin_out || !umax_out || cols < 1) return SD_ERR_INVALID_ARG;
    // PM.cc:877-910, same operation order as stage (A) of k1_search
    float R1[9], t1[3], R2[9], t2[3], R21[9], t21[3];
    for (int r = 0; r < 3; ++r) {
        for (int col = 0; col < 3; ++col) { R1[r * 3 + col] = Tcw1[r * 4 + col]; R2[r * 3 + col] = Tcw2[r * 4 + col]; }
        t1[r] = Tcw1[r * 4 + 3];
        t2[r] = Tcw2[r * 4 + 3];
    }
    geom::relative_pose(R1, t1, R2, t2, R21, t21);
    const float fx = K1[0], fy = K1[1], cx = K1[2], cy = K1[3];
    const float xp0 = ((float)px - cx) / fx, xp1 = ((float)py - cy) / fy;
    const float q0 = R21[0] * xp0 + R21[1] * xp1 + R21[2] * 1.0f;
    const float q2 = R21[6] * xp0 + R21[7] * xp1 + R21[8] * 1.0f;
    const float xmin0 = (float)((double)q0 * (double)mind + (double)t21[0]);
    const float xmin2 = (float)((double)q2 * (double)mind + (double)t21[2]);
    const float xmax0 = (float)((double)q0 * (double)maxd + (double)t21[0]);
    const float xmax2 = (float)((double)q2 * (double)maxd + (double)t21[2]);
    float umin = fx * xmin0 / xmin2 + cx;
    float umax = fx * xmax0 / xmax2 + cx;
    if (umin > umax) { float t = umax; umax = umin; umin = t; }
    const float colsf = (float)cols;
    if (umin < 0) umin = 0;
    if (umax < 0) umax = 0;
    if (umin > colsf) umin = colsf;
    if (umax > colsf) umax = colsf;
    *umin_out = umin;
    *umax_out = umax;
    return SD_OK;
}

static int intra_planes(sd_ctx* c, float* depth, float* sigma, const float* grad, int w, int h, bool grow) {
    if (!c || !depth || !sigma || !grad || w < 8 || h < 8 || (long long)(w + 128) * h >= (1ll << 30))
        return fail(c, SD_ERR_INVALID_ARG, "sd_intra_depth_*_planes: bad argument");
    CU(c, cudaSetDevice(c->device));
    const size_t bytes = (size_t)w * h * sizeof(float);
    float* d = nullptr;
    CU(c, cudaMalloc(&d, 5 * bytes));
    float *d_depth = d, *d_sigma = d + (size_t)w * h, *d_grad = d + 2 * (size_t)w * h, *o_depth = d + 3 * (size_t)w * h,
          *o_sigma = d + 4 * (size_t)w * h;
    cudaError_t e = cudaMemcpyAsync(d_depth, depth, bytes, cudaMemcpyHostToDevice, c->s_main);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_sigma, sigma, bytes, cudaMemcpyHostToDevice, c->s_main);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_grad, grad, bytes, cudaMemcpyHostToDevice, c->s_main);
    if (e == cudaSuccess) {
        IntraArgs S;
        S.W = w; S.H = h; S.pitch_in = w; S.pitch_out = w; S.border = c->P.border;
        S.vec_ok = (w % 4 == 0) && (((size_t)w * h) % 4 == 0);  // the five planes are carved out of one allocation
        S.lambdaG = c->P.lambdaG; S.chi_fuse = c->chi_fuse_f; S.grad = d_grad;
        S.depth_in = d_depth; S.sigma_in = d_sigma; S.depth_out = o_depth; S.sigma_out = o_sigma;
        S.depth_copy = nullptr; S.vlist = nullptr; S.vcount = nullptr;
        dim3 grd((w + K23_TW - 1) / K23_TW, (h + K23_TH - 1) / K23_TH);
        if (grow) k23_intra<false, true><<<grd, K23_THREADS, K23_SMEM, c->s_main>>>(S);
        else k23_intra<true, false><<<grd, K23_THREADS, K23_SMEM, c->s_main>>>(S);
        c->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(depth, o_depth, bytes, cudaMemcpyDeviceToHost, c->s_main);
    if (e == cudaSuccess) e = cudaMemcpyAsync(sigma, o_sigma, bytes, cudaMemcpyDeviceToHost, c->s_main);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->s_main);
    cudaFree(d);
    if (e != cudaSuccess) return fail(c, SD_ERR_CUDA, "sd_intra_depth_*_planes: %s", cudaGetErrorString(e));
    return SD_OK;
}

int sd_intra_depth_check_planes(sd_ctx* c, float* depth, float* sigma, const float* grad, int w, int h) {
    return intra_planes(c, depth, sigma, grad, w, h, false);
}

int sd_intra_depth_grow_planes(sd_ctx* c, float* depth, float* sigma, const float* grad, int w, int h) {
    return intra_planes(c, depth, sigma, grad, w, h, true);
}

// ---- the entry point that allocates nothing ----
// Working planes of sd_intra_depth_check_planes_static: three uploaded inputs + two outputs, each up to
// SD_STATIC_MAX_PIXELS floats.  The array belongs to the loaded module, one copy per GPU that loads libsdpm.so, held until
// the process exits.  In host code `g_static_planes` is only a handle for cudaGetSymbolAddress: it is not a device address
// and must never be handed to a kernel or a cudaMemcpy as one.
static __device__ __align__(256) float g_static_planes[5 * (size_t)SD_STATIC_MAX_PIXELS];

namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu;
float* g_static_addr[kStaticMaxDevices] = {};  // device address of g_static_planes on each GPU, looked up once

// The caller has made `device` current.
int static_planes_addr(sd_ctx* c, int device, float** out) {
    if (device < 0 || device >= kStaticMaxDevices) return fail(c, SD_ERR_INVALID_ARG, "sd_intra_depth_check_planes_static: device %d out of range", device);
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_addr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_planes);
        if (e != cudaSuccess || !p)
            return fail(c, SD_ERR_CUDA, "cudaGetSymbolAddress(g_static_planes) on device %d: %s", device, cudaGetErrorString(e));
        g_static_addr[device] = static_cast<float*>(p);
    }
    *out = g_static_addr[device];
    return SD_OK;
}
}  // namespace

int sd_intra_depth_check_planes_static(sd_ctx* c, float* depth, float* sigma, const float* grad, int w, int h) {
    if (!c || !depth || !sigma || !grad || w < 8 || h < 8)
        return fail(c, SD_ERR_INVALID_ARG, "sd_intra_depth_check_planes_static: bad argument");
    if ((long long)w * h > (long long)SD_STATIC_MAX_PIXELS)
        return fail(c, SD_ERR_INVALID_ARG,
                    "sd_intra_depth_check_planes_static: %d x %d = %lld pixels is more than SD_STATIC_MAX_PIXELS = %d, the size of the "
                    "library's built-in working planes; use sd_intra_depth_check_planes for larger planes",
                    w, h, (long long)w * h, (int)SD_STATIC_MAX_PIXELS);
    CU(c, cudaSetDevice(c->device));
    float* d = nullptr;
    const int rc = static_planes_addr(c, c->device, &d);
    if (rc != SD_OK) return rc;
    const size_t npix = (size_t)w * h, bytes = npix * sizeof(float);
    float *d_depth = d, *d_sigma = d + npix, *d_grad = d + 2 * npix, *o_depth = d + 3 * npix, *o_sigma = d + 4 * npix;
    // Nothing here has to start at zero: the uploads overwrite the three input planes and the kernel stores every pixel
    // of both output planes, so whatever the previous call left behind is never read.
    cudaError_t e = cudaMemcpyAsync(d_depth, depth, bytes, cudaMemcpyHostToDevice, c->s_main);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_sigma, sigma, bytes, cudaMemcpyHostToDevice, c->s_main);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_grad, grad, bytes, cudaMemcpyHostToDevice, c->s_main);
    if (e == cudaSuccess) {
        IntraArgs S;
        S.W = w; S.H = h; S.pitch_in = w; S.pitch_out = w; S.border = c->P.border;
        S.vec_ok = (w % 4 == 0) && (npix % 4 == 0);  // the array is 256-byte aligned and the planes start npix floats apart
        S.lambdaG = c->P.lambdaG; S.chi_fuse = c->chi_fuse_f; S.grad = d_grad;
        S.depth_in = d_depth; S.sigma_in = d_sigma; S.depth_out = o_depth; S.sigma_out = o_sigma;
        S.depth_copy = nullptr; S.vlist = nullptr; S.vcount = nullptr;
        dim3 grd((w + K23_TW - 1) / K23_TW, (h + K23_TH - 1) / K23_TH);
        k23_intra<true, false><<<grd, K23_THREADS, K23_SMEM, c->s_main>>>(S);
        c->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(depth, o_depth, bytes, cudaMemcpyDeviceToHost, c->s_main);
    if (e == cudaSuccess) e = cudaMemcpyAsync(sigma, o_sigma, bytes, cudaMemcpyDeviceToHost, c->s_main);
    // wait in every case: the caller's planes must not be the target or source of a copy still in flight on return
    const cudaError_t es = cudaStreamSynchronize(c->s_main);
    if (e == cudaSuccess) e = es;
    if (e != cudaSuccess) return fail(c, SD_ERR_CUDA, "sd_intra_depth_check_planes_static: %s", cudaGetErrorString(e));
    return SD_OK;
}

// ---- probes ----

int sd_probe_epipolar(sd_ctx* c, int kf1, int kf2, const int* xy, int n, float min_invdepth, float max_invdepth, float rot_deg, float* out) {
    if (!c || !xy || !out || n < 1) return fail(c, SD_ERR_INVALID_ARG, "sd_probe_epipolar: bad argument");
    Slot* s = find(c, kf1);
    Slot* nb = find(c, kf2);
    if (!s || !nb) return fail(c, SD_ERR_UNKNOWN_KEYFRAME, "unknown key frame %d or %d", kf1, kf2);
    CU(c, cudaSetDevice(c->device));
    std::vector<uint32_t> entries(n);
    for (int i = 0; i < n; ++i) {
        const int x = xy[2 * i], y = xy[2 * i + 1];
        if (x < 0 || x >= c->W || y < 0 || y >= c->H) return fail(c, SD_ERR_INVALID_ARG, "probe pixel (%d,%d) outside image", x, y);
        entries[i] = ((uint32_t)y << 16) | (uint32_t)x;
    }
    uint32_t* d_list = nullptr;
    float* d_out = nullptr;
    CU(c, cudaMalloc(&d_list, (n + 1) * sizeof(uint32_t)));  // + the task counter
    CU(c, cudaMalloc(&d_out, (size_t)n * 6 * sizeof(float)));
    CU(c, cudaStreamSynchronize(c->s_up));
    CU(c, cudaMemcpyAsync(d_list, entries.data(), n * sizeof(uint32_t), cudaMemcpyHostToDevice, c->s_main));
    static thread_local SearchArgs A;
    fill_search_common(c, s, A);
    A.n = 1;
    A.mind = min_invdepth;
    A.maxd = max_invdepth;
    A.mind_d = (double)min_invdepth;
    A.maxd_d = (double)max_invdepth;
    A.list = d_list;
    A.count_override = n;
    A.task_counter = d_list + n;
    CU(c, cudaMemsetAsync(d_list + n, 0, sizeof(uint32_t), c->s_main));
    A.probe_out = d_out;
    fill_nbr(s, nb, rot_deg, A.nb[0]);
    const int blocks = std::min(c->k1_grid, (n + K1_WARPS * 32 - 1) / (K1_WARPS * 32));
    k1_search<<<blocks, K1_WARPS * 32, 0, c->s_main>>>(A);
    c->launches++;
    CU(c, cudaGetLastError());
    CU(c, cudaMemcpyAsync(out, d_out, (size_t)n * 6 * sizeof(float), cudaMemcpyDeviceToHost, c->s_main));
    CU(c, cudaStreamSynchronize(c->s_main));
    cudaFree(d_list);
    cudaFree(d_out);
    return SD_OK;
}

int sd_probe_fusion(sd_ctx* c, const float* depth, const float* sigma, int n, float* out_depth, float* out_sigma, int* supported) {
    if (!c || !depth || !sigma || n < 0 || n > SD_MAX_NEIGHBOURS || !out_depth || !out_sigma || !supported)
        return fail(c, SD_ERR_INVALID_ARG, "sd_probe_fusion: bad argument");
    CU(c, cudaSetDevice(c->device));
    float* d = nullptr;
    CU(c, cudaMalloc(&d, (2 * SD_MAX_NEIGHBOURS + 4) * sizeof(float)));
    if (n) {
        CU(c, cudaMemcpyAsync(d, depth, n * sizeof(float), cudaMemcpyHostToDevice, c->s_main));
        CU(c, cudaMemcpyAsync(d + SD_MAX_NEIGHBOURS, sigma, n * sizeof(float), cudaMemcpyHostToDevice, c->s_main));
    }
    k1_fusion_probe<<<1, 1, 0, c->s_main>>>(d, d + SD_MAX_NEIGHBOURS, n, c->P.lambdaN, c->chi_fuse_f, d + 2 * SD_MAX_NEIGHBOURS);
    c->launches++;
    CU(c, cudaGetLastError());
    float r[3];
    CU(c, cudaMemcpyAsync(r, d + 2 * SD_MAX_NEIGHBOURS, sizeof(r), cudaMemcpyDeviceToHost, c->s_main));
    CU(c, cudaStreamSynchronize(c->s_main));
    cudaFree(d);
    *out_depth = r[0];
    *out_sigma = r[1];
    *supported = r[2] != 0.f;
    return SD_OK;
}

}  // extern "C"
