// Scene scoring with the per-model top-k reduced on the device, and the pose list sharded over GPUs.
//
// The reference scores a scene's candidate poses in gpu_batch_size batches on ONE GPU (search_env.cpp:2504-2525) and
// then takes the per-model argmin on the host (2542-2583).  Here a scene is one call:
//   H2D of the pose batch -> fused render+cost kernel -> top-k kernels (topk.cu) -> ncclAllGather of k x 16 B per model
//   (only when the pose list is sharded) -> D2H of the gathered records (+ the cost arrays if the caller wants them)
// all on the context's stream with ONE host synchronisation at the end, followed by the lexicographic (cost, pose id)
// merge of the gathered records on the host.  Poses are independent, so ranks score contiguous blocks of the list
// against replicated models / scene / observed cloud and the all-gather is the only exchange.
//
// Two ways to shard, same code underneath:
//   perch_comm_init_rank     one process per GPU (torchrun / MPI): every rank calls perch_score_scene on its block
//   perch_create_multi       one process, n devices: one context + one host thread + one stream per device,
//                            ncclCommInitAll; perch_multi_score cuts the list and runs the blocks in parallel
#include <cuda_runtime.h>
#include <nccl.h>

#include <algorithm>
#include <climits>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "capi_common.hpp"
#include "context.hpp"
#include "kernels.cuh"
#include "perch_b200.h"

namespace {
using perch_detail::fail;
using perch_detail::Scratch;

#define NC(call)                                                                                                       \
    do {                                                                                                               \
        ncclResult_t r__ = (call);                                                                                     \
        if (r__ != ncclSuccess)                                                                                        \
            return fail(PERCH_ERR_CUDA, "%s failed: %s (%s:%d)", #call, ncclGetErrorString(r__), __FILE__, __LINE__); \
    } while (0)

// One rank's share of a scene.  Host pointers; the output pointers are already offset to this rank's block and the three
// count rows lie `count_row_stride` ints apart (the caller's 3 x P_total array).
struct SceneJob {
    const float* poses16 = nullptr;
    int n = 0;
    const int* pmm = nullptr;
    const int* label = nullptr;
    const float* tot = nullptr;
    const perch_params* p = nullptr;
    int pose_id_offset = 0, n_models = 0, k = 0;
    float *rendered = nullptr, *observed = nullptr, *diff = nullptr;
    int32_t* counts = nullptr;
    size_t count_row_stride = 0;
    perch_hit* out_hits = nullptr;  // merged n_models x k (every rank receives the gathered list; null = not wanted)
};

bool hit_before(const perch_hit& a, const perch_hit& b) { return a.cost != b.cost ? a.cost < b.cost : a.pose_id < b.pose_id; }

// gathered: n_ranks x n_models x k records (unordered within a model, INT32_MAX pads) -> out: n_models x k best, ordered
void merge_hits(const perch_hit* gathered, int n_ranks, int n_models, int k, perch_hit* out) {
    std::vector<perch_hit> all;
    all.reserve((size_t)n_ranks * k);
    const perch_hit empty{INT32_MAX, -1, 0.f, 0.f};
    for (int m = 0; m < n_models; ++m) {
        all.clear();
        for (int r = 0; r < n_ranks; ++r) {
            const perch_hit* src = gathered + ((size_t)r * n_models + m) * k;
            for (int j = 0; j < k; ++j)
                if (src[j].cost != INT32_MAX || src[j].pose_id >= 0) all.push_back(src[j]);
        }
        const size_t take = std::min<size_t>(all.size(), (size_t)k);
        std::partial_sort(all.begin(), all.begin() + take, all.end(), hit_before);
        for (size_t j = 0; j < (size_t)k; ++j) out[(size_t)m * k + j] = j < take ? all[j] : empty;
    }
}

int score_scene_impl(perch_ctx* c, const SceneJob& j) {
    if (!c || !j.p) return fail(PERCH_ERR_INVALID, "null argument");
    if (j.n < 0 || (j.n > 0 && (!j.poses16 || !j.pmm))) return fail(PERCH_ERR_INVALID, "bad pose arguments");
    if (j.n_models <= 0 || j.n_models > perch::kTopkMaxModels)
        return fail(PERCH_ERR_CAPACITY, "n_models %d outside [1, %d]", j.n_models, perch::kTopkMaxModels);
    if (j.n > perch::kTopkMaxPoses) return fail(PERCH_ERR_CAPACITY, "more than 2^24 poses in one call (per GPU)");
    if (j.k <= 0 || (size_t)j.k * j.n_models * c->comm_size > ((size_t)1 << 22))
        return fail(PERCH_ERR_CAPACITY, "k = %d: k * n_models * ranks must be in [1, 2^22]", j.k);
    if (j.p->do_icp) return fail(PERCH_ERR_INVALID, "perch_score_scene scores the poses as given (do_icp: use perch_render_cost)");
    perch_detail::DeviceGuard g(c->device);
    perch::PoseArgs a;
    bool color = false;
    a.pose_label = j.label ? reinterpret_cast<const int*>(1) : nullptr;  // 6-DoF or not (the pointer is set below)
    int rc = perch_detail::prepare_fused_cost(c, j.p, a, color);
    if (rc) return rc;
    const size_t P = (size_t)j.n, M = (size_t)j.n_models, K = (size_t)j.k, G = (size_t)c->comm_size;
    const size_t hits_bytes = M * K * sizeof(perch_hit);
    const size_t need = Scratch::align(P * 64) + 3 * Scratch::align(P * 4) + 2 * Scratch::align(P * 12) + Scratch::align(P * 8) +
                        Scratch::align(hits_bytes) + Scratch::align(G * hits_bytes) + 4096;
    if ((rc = perch_detail::ensure_scratch(c, need))) return rc;
    if ((rc = perch_detail::ensure_pinned(c, G * hits_bytes + 64))) return rc;
    Scratch sc(c);
    float* d_poses = sc.take<float>(P * 16);
    int* d_pmm = sc.take<int>(P);
    int* d_label = sc.take<int>(P);
    float* d_tot = sc.take<float>(P);
    float* d_costs = sc.take<float>(P * 3);
    int32_t* d_counts = sc.take<int32_t>(P * 3);
    unsigned long long* d_keys = sc.take<unsigned long long>(P);
    perch_hit* d_hits = sc.take<perch_hit>(M * K);
    perch_hit* d_gathered = sc.take<perch_hit>(G * M * K);
    cudaStream_t st = c->stream;
    if (P > 0) {
        PERCH_CU(cudaMemcpyAsync(d_poses, j.poses16, P * 64, cudaMemcpyHostToDevice, st));
        PERCH_CU(cudaMemcpyAsync(d_pmm, j.pmm, P * 4, cudaMemcpyHostToDevice, st));
        if (j.label) PERCH_CU(cudaMemcpyAsync(d_label, j.label, P * 4, cudaMemcpyHostToDevice, st));
        if (j.tot) PERCH_CU(cudaMemcpyAsync(d_tot, j.tot, P * 4, cudaMemcpyHostToDevice, st));
    }
    a.poses = reinterpret_cast<const perch::Mat4*>(d_poses);
    a.pose_model_map = d_pmm;
    a.pose_label = j.label ? d_label : nullptr;
    a.pose_obs_total = j.tot ? d_tot : nullptr;
    a.n_poses = j.n;
    a.out_costs = d_costs;
    a.out_counts = d_counts;
    PERCH_CU(cudaEventRecord(c->ev0, st));
    PERCH_CU(perch::launch_fused_cost(a, color, st));
    PERCH_CU(cudaEventRecord(c->ev1, st));
    perch::TopkArgs t;
    t.costs = d_costs; t.pose_model_map = d_pmm; t.n_poses = j.n; t.pose_id_offset = j.pose_id_offset;
    t.n_models = j.n_models; t.k = j.k; t.keys = d_keys; t.hits = d_hits;
    PERCH_CU(perch::launch_topk(t, st));
    const perch_hit* d_all = d_hits;
    if (c->comm && G > 1) {
        NC(ncclAllGather(d_hits, d_gathered, hits_bytes, ncclChar, static_cast<ncclComm_t>(c->comm), st));
        d_all = d_gathered;
    }
    char* h = static_cast<char*>(c->h_pinned);
    PERCH_CU(cudaMemcpyAsync(h, d_all, G * hits_bytes, cudaMemcpyDeviceToHost, st));
    PERCH_CU(cudaMemcpyAsync(h + G * hits_bytes, c->d_err_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    if (P > 0) {
        if (j.rendered) PERCH_CU(cudaMemcpyAsync(j.rendered, d_costs, P * 4, cudaMemcpyDeviceToHost, st));
        if (j.p->calculate_observed_cost) {
            if (j.observed) PERCH_CU(cudaMemcpyAsync(j.observed, d_costs + P, P * 4, cudaMemcpyDeviceToHost, st));
            if (j.diff) PERCH_CU(cudaMemcpyAsync(j.diff, d_costs + 2 * P, P * 4, cudaMemcpyDeviceToHost, st));
        }
        if (j.counts)
            for (int r = 0; r < 3; ++r)
                PERCH_CU(cudaMemcpyAsync(j.counts + (size_t)r * j.count_row_stride, d_counts + (size_t)r * P, P * 4,
                                         cudaMemcpyDeviceToHost, st));
    }
    PERCH_CU(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
    c->last_launches = (P > 0 ? 2 : 0) + 1;  // fused kernel + key kernel, select kernel (the all-gather is NCCL's)
    c->timing_pending = false;
    int flag = 0;
    memcpy(&flag, h + G * hits_bytes, sizeof(int));
    if (flag) {
        cudaMemsetAsync(c->d_err_flag, 0, sizeof(int), st);
        return perch_detail::report_err_flag(flag);
    }
    if (j.out_hits) merge_hits(reinterpret_cast<const perch_hit*>(h), (int)G, j.n_models, j.k, j.out_hits);
    return PERCH_OK;
}

}  // namespace

void perch_detail::destroy_comm(perch_ctx* c) {
    if (c->comm && c->comm_owned) ncclCommDestroy(static_cast<ncclComm_t>(c->comm));
    c->comm = nullptr;
    c->comm_rank = 0;
    c->comm_size = 1;
    c->comm_owned = false;
}

// ---- one process, n devices ------------------------------------------------------------------------------------------
struct perch_multi {
    std::vector<perch_ctx*> ctx;
    // one persistent host thread per device; a call hands every worker a closure and waits for all of them
    struct Worker {
        std::thread th;
        std::mutex mu;
        std::condition_variable cv;
        std::function<int()> job;
        bool has_job = false, done = false, quit = false;
        int rc = 0;
        std::string err;
    };
    std::vector<Worker*> workers;

    void worker_loop(Worker* w) {
        for (;;) {
            std::function<int()> job;
            {
                std::unique_lock<std::mutex> lk(w->mu);
                w->cv.wait(lk, [&] { return w->has_job || w->quit; });
                if (w->quit) return;
                job = std::move(w->job);
                w->has_job = false;
            }
            const int rc = job();
            const std::string err = rc ? perch_last_error() : "";
            {
                std::lock_guard<std::mutex> lk(w->mu);
                w->rc = rc;
                w->err = err;
                w->done = true;
            }
            w->cv.notify_all();
        }
    }
    // runs fn(i) on worker i for every device, returns the first failure (its message becomes the caller's last error)
    int run_all(const std::function<int(int)>& fn) {
        for (size_t i = 0; i < workers.size(); ++i) {
            Worker* w = workers[i];
            {
                std::lock_guard<std::mutex> lk(w->mu);
                w->job = [fn, i] { return fn((int)i); };
                w->has_job = true;
                w->done = false;
            }
            w->cv.notify_all();
        }
        int rc = PERCH_OK;
        for (Worker* w : workers) {
            std::unique_lock<std::mutex> lk(w->mu);
            w->cv.wait(lk, [&] { return w->done; });
            if (w->rc && !rc) rc = fail(w->rc, "%s", w->err.c_str());
        }
        return rc;
    }
};

extern "C" {

int perch_comm_unique_id(void* out_id, size_t bytes) {
    if (!out_id || bytes < sizeof(ncclUniqueId))
        return fail(PERCH_ERR_INVALID, "out_id must hold %zu bytes (PERCH_COMM_ID_BYTES)", sizeof(ncclUniqueId));
    static_assert(sizeof(ncclUniqueId) == PERCH_COMM_ID_BYTES, "PERCH_COMM_ID_BYTES must equal sizeof(ncclUniqueId)");
    ncclUniqueId id;
    NC(ncclGetUniqueId(&id));
    memcpy(out_id, &id, sizeof(id));
    return PERCH_OK;
}

int perch_comm_init_rank(perch_ctx* c, int n_ranks, int rank, const void* id_bytes) {
    if (!c || !id_bytes || n_ranks <= 0 || rank < 0 || rank >= n_ranks) return fail(PERCH_ERR_INVALID, "bad comm argument");
    perch_detail::DeviceGuard g(c->device);
    perch_detail::destroy_comm(c);
    if (n_ranks == 1) return PERCH_OK;
    ncclUniqueId id;
    memcpy(&id, id_bytes, sizeof(id));
    ncclComm_t comm = nullptr;
    NC(ncclCommInitRank(&comm, n_ranks, id, rank));
    c->comm = comm;
    c->comm_rank = rank;
    c->comm_size = n_ranks;
    c->comm_owned = true;
    return PERCH_OK;
}

int perch_comm_destroy(perch_ctx* c) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    perch_detail::DeviceGuard g(c->device);
    cudaStreamSynchronize(c->stream);
    perch_detail::destroy_comm(c);
    return PERCH_OK;
}

int perch_score_scene(perch_ctx* c, const float* poses16, int n_poses, const int* pose_model_map, const int* pose_label,
                      const float* pose_obs_total, const perch_params* p, int pose_id_offset, int n_models, int k,
                      perch_outputs* out, perch_hit* out_hits) {
    SceneJob j;
    j.poses16 = poses16; j.n = n_poses; j.pmm = pose_model_map; j.label = pose_label; j.tot = pose_obs_total; j.p = p;
    j.pose_id_offset = pose_id_offset; j.n_models = n_models; j.k = k;
    if (out) {
        j.rendered = out->rendered_cost; j.observed = out->observed_cost; j.diff = out->points_diff_cost;
        j.counts = out->counts; j.count_row_stride = (size_t)std::max(n_poses, 0);
    }
    j.out_hits = out_hits;
    return score_scene_impl(c, j);
}

int perch_create_multi(const int* devices, int n_devices, perch_multi** out) {
    if (!out || !devices || n_devices <= 0 || n_devices > 64) return fail(PERCH_ERR_INVALID, "bad create_multi argument");
    *out = nullptr;
    for (int i = 0; i < n_devices; ++i)
        for (int j = 0; j < i; ++j)
            if (devices[i] == devices[j]) return fail(PERCH_ERR_INVALID, "device %d listed twice", devices[i]);
    perch_multi* m = new perch_multi();
    auto cleanup = [&](int rc) {
        perch_destroy_multi(m);
        return rc;
    };
    for (int i = 0; i < n_devices; ++i) {
        perch_ctx* c = nullptr;
        int rc = perch_create(devices[i], &c);
        if (rc) return cleanup(rc);
        m->ctx.push_back(c);
    }
    if (n_devices > 1) {
        std::vector<ncclComm_t> comms(n_devices);
        ncclResult_t r = ncclCommInitAll(comms.data(), n_devices, devices);
        if (r != ncclSuccess) return cleanup(fail(PERCH_ERR_CUDA, "ncclCommInitAll failed: %s", ncclGetErrorString(r)));
        for (int i = 0; i < n_devices; ++i) {
            m->ctx[i]->comm = comms[i];
            m->ctx[i]->comm_rank = i;
            m->ctx[i]->comm_size = n_devices;
            m->ctx[i]->comm_owned = true;
        }
    }
    for (int i = 0; i < n_devices; ++i) {
        auto* w = new perch_multi::Worker();
        m->workers.push_back(w);
        w->th = std::thread([m, w] { m->worker_loop(w); });
    }
    *out = m;
    return PERCH_OK;
}

void perch_destroy_multi(perch_multi* m) {
    if (!m) return;
    for (auto* w : m->workers) {
        {
            std::lock_guard<std::mutex> lk(w->mu);
            w->quit = true;
        }
        w->cv.notify_all();
        if (w->th.joinable()) w->th.join();
        delete w;
    }
    for (perch_ctx* c : m->ctx) perch_destroy(c);
    delete m;
}

int perch_multi_size(perch_multi* m) { return m ? (int)m->ctx.size() : 0; }

perch_ctx* perch_multi_context(perch_multi* m, int i) {
    if (!m || i < 0 || i >= (int)m->ctx.size()) return nullptr;
    return m->ctx[i];
}

// Replicated state: the same call on every device, in parallel on the worker threads.
int perch_multi_set_camera(perch_multi* m, int width, int height, const float proj16[16], float cx, float cy, float fx, float fy,
                           float depth_factor) {
    if (!m) return fail(PERCH_ERR_INVALID, "null multi context");
    return m->run_all([=](int i) { return perch_set_camera(m->ctx[i], width, height, proj16, cx, cy, fx, fy, depth_factor); });
}
int perch_multi_set_models(perch_multi* m, const perch_triangle* tris, int64_t n_tris, const int* tris_model_count, int n_models) {
    if (!m) return fail(PERCH_ERR_INVALID, "null multi context");
    return m->run_all([=](int i) { return perch_set_models(m->ctx[i], tris, n_tris, tris_model_count, n_models); });
}
int perch_multi_set_scene(perch_multi* m, const int32_t* source_depth, const uint8_t* red, const uint8_t* green,
                          const uint8_t* blue, const uint8_t* label) {
    if (!m) return fail(PERCH_ERR_INVALID, "null multi context");
    return m->run_all([=](int i) { return perch_set_scene(m->ctx[i], source_depth, red, green, blue, label); });
}
int perch_multi_set_observed(perch_multi* m, const float* xyz_aos, const uint8_t* rgb_soa, const int* label, int n_obs) {
    if (!m) return fail(PERCH_ERR_INVALID, "null multi context");
    return m->run_all([=](int i) { return perch_set_observed(m->ctx[i], xyz_aos, rgb_soa, label, n_obs); });
}

int perch_multi_score(perch_multi* m, const float* poses16, int n_poses, const int* pose_model_map, const int* pose_label,
                      const float* pose_obs_total, const perch_params* p, int n_models, int k, perch_outputs* out,
                      perch_hit* out_hits) {
    if (!m || !p) return fail(PERCH_ERR_INVALID, "null argument");
    if (n_poses < 0 || (n_poses > 0 && (!poses16 || !pose_model_map))) return fail(PERCH_ERR_INVALID, "bad pose arguments");
    const int G = (int)m->ctx.size();
    const int base = n_poses / G, extra = n_poses % G;
    // contiguous blocks in rank order: global pose ids stay monotone in (rank, local index), so first-wins ties survive
    return m->run_all([=](int r) {
        const int lo = r * base + std::min(r, extra), n = base + (r < extra ? 1 : 0);
        SceneJob j;
        j.poses16 = poses16 ? poses16 + 16 * (size_t)lo : nullptr;
        j.n = n;
        j.pmm = pose_model_map ? pose_model_map + lo : nullptr;
        j.label = pose_label ? pose_label + lo : nullptr;
        j.tot = pose_obs_total ? pose_obs_total + lo : nullptr;
        j.p = p;
        j.pose_id_offset = lo; j.n_models = n_models; j.k = k;
        if (out) {
            j.rendered = out->rendered_cost ? out->rendered_cost + lo : nullptr;
            j.observed = out->observed_cost ? out->observed_cost + lo : nullptr;
            j.diff = out->points_diff_cost ? out->points_diff_cost + lo : nullptr;
            j.counts = out->counts ? out->counts + lo : nullptr;
            j.count_row_stride = (size_t)n_poses;
        }
        j.out_hits = r == 0 ? out_hits : nullptr;  // every rank holds the gathered list; rank 0 merges it for the caller
        return score_scene_impl(m->ctx[r], j);
    });
}

int perch_multi_last_timing(perch_multi* m, float* max_kernel_ms) {
    if (!m) return fail(PERCH_ERR_INVALID, "null multi context");
    float mx = 0.f;
    for (perch_ctx* c : m->ctx) mx = std::max(mx, c->last_ms);
    if (max_kernel_ms) *max_kernel_ms = mx;
    return PERCH_OK;
}

}  // extern "C"
