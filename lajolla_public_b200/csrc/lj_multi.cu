// Multi-device contexts (lj_init_devices): one render fanned out over the GPUs of one box.
//
// Replaces the reference's unit of parallel work -- parallel_init(num_threads) (main.cpp:31) + the parallel_for
// over 16x16 tiles with one pcg32 stream each (render.cpp:75-98) -- at the scale of whole devices:
//   * the scene is replicated: every device gets its own upload and its own BVH build (lj_scene_create_single);
//   * one persistent host thread per device issues that device's work (each thread has its own default stream:
//     the library is compiled with --default-stream per-thread), so the devices render concurrently;
//   * a render is split by tiles (tile id % n, bit-identical to one device) or by sample slices, each device
//     renders into its own f64 sums, and the sums are added with ONE ncclReduce to the first device
//     (NVLink / NVSwitch); there is no other exchange on this path.
// NCCL is bound at run time (dlopen of libnccl.so.2: the system copy, or the one a host process such as PyTorch
// already mapped), so the library itself has no link-time dependency on it and one-device use never touches it.
#include "lj_internal.h"
#include "lj_alloc.h"

#include <condition_variable>
#include <cstring>
#include <dlfcn.h>
#include <functional>
#include <mutex>
#include <nccl.h>
#include <string>
#include <thread>

namespace {

// ---- NCCL, bound with dlopen (types from the system's nccl.h; no link-time dependency) ----
enum { kNcclSum = ncclSum, kNcclFloat64 = ncclFloat64 };

struct NcclApi {
    void *handle = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    std::string error;
    bool load() {
        if (handle) return true;
        const char *names[] = {getenv("LJ_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {
            if (!n || !*n) continue;
            handle = dlopen(n, RTLD_NOW | RTLD_LOCAL);
            if (handle) break;
            error = dlerror();
        }
        if (!handle) return false;
        bool ok = true;
        auto sym = [&](const char *n) { void *p = dlsym(handle, n); if (!p) { ok = false; error = std::string("missing symbol ") + n; } return p; };
        GetErrorString = (decltype(GetErrorString))sym("ncclGetErrorString");
        GetVersion = (decltype(GetVersion))sym("ncclGetVersion");
        CommInitAll = (decltype(CommInitAll))sym("ncclCommInitAll");
        CommDestroy = (decltype(CommDestroy))sym("ncclCommDestroy");
        GroupStart = (decltype(GroupStart))sym("ncclGroupStart");
        GroupEnd = (decltype(GroupEnd))sym("ncclGroupEnd");
        Reduce = (decltype(Reduce))sym("ncclReduce");
        if (!ok) { dlclose(handle); handle = nullptr; }
        return ok;
    }
};

// ---- one persistent host thread per device ----
class Worker {
public:
    Worker() : thread_([this] { loop(); }) {}
    ~Worker() {
        { std::lock_guard<std::mutex> g(mu_); quit_ = true; }
        cv_.notify_all();
        thread_.join();
    }
    void post(std::function<void()> job) {
        { std::lock_guard<std::mutex> g(mu_); job_ = std::move(job); busy_ = true; }
        cv_.notify_all();
    }
    void wait() { std::unique_lock<std::mutex> l(mu_); cv_.wait(l, [this] { return !busy_; }); }
private:
    void loop() {
        std::unique_lock<std::mutex> l(mu_);
        while (true) {
            cv_.wait(l, [this] { return quit_ || (busy_ && job_); });
            if (quit_) return;
            std::function<void()> job = std::move(job_);
            job_ = nullptr;
            l.unlock();
            job();
            l.lock();
            busy_ = false;
            cv_.notify_all();
        }
    }
    std::mutex mu_;
    std::condition_variable cv_;
    std::function<void()> job_;
    bool busy_ = false, quit_ = false;
    std::thread thread_;
};

}  // namespace

struct LjDeviceGroup {
    std::vector<lj_ctx *> ctx;          // one-device contexts, ctx[0] is the root of the reduction
    std::vector<Worker *> worker;
    std::vector<cudaStream_t> stream;   // per device, for the collective
    std::vector<double *> sum;          // per device f64 sums of the render in flight (grown on demand)
    size_t sum_bytes = 0;
    NcclApi nccl;
    std::vector<ncclComm_t> comm;
    int nccl_version = 0;

    int n() const { return (int)ctx.size(); }
    // run job(i) on every device's thread; returns the first failure (message of that thread kept)
    int on_all(const std::function<int(int)> &job) {
        std::vector<int> rc(n(), LJ_OK);
        std::vector<std::string> msg(n());
        for (int i = 0; i < n(); i++)
            worker[i]->post([&, i] { rc[i] = job(i); if (rc[i] != LJ_OK) msg[i] = lj_error_text(); });
        for (int i = 0; i < n(); i++) worker[i]->wait();
        for (int i = 0; i < n(); i++)
            if (rc[i] != LJ_OK) { lj_set_error("device %d: %s", ctx[i] ? ctx[i]->device : i, msg[i].c_str()); return rc[i]; }
        return LJ_OK;
    }
    int ensure_comms() {
        if (!comm.empty()) return LJ_OK;
        if (!nccl.load()) {
            lj_set_error("multi-device render needs NCCL for the reduction of the radiance sums and libnccl.so.2 could not be loaded (%s)", nccl.error.c_str());
            return LJ_EUNSUPPORTED;
        }
        nccl.GetVersion(&nccl_version);
        std::vector<int> ids;
        for (lj_ctx *c : ctx) ids.push_back(c->device);
        comm.assign(n(), nullptr);
        ncclResult_t r = nccl.CommInitAll(comm.data(), n(), ids.data());
        if (r != ncclSuccess) { comm.clear(); lj_set_error("ncclCommInitAll over %d devices failed: %s", n(), nccl.GetErrorString(r)); return LJ_ECUDA; }
        return LJ_OK;
    }
};

extern "C" int lj_ctx_device_count(lj_ctx *ctx) { return !ctx ? 0 : (ctx->group ? ctx->group->n() : 1); }

extern "C" int lj_init_devices(const int *device_ids, int n_devices, lj_ctx **out) {
    if (!out) { lj_set_error("lj_init_devices: out is null"); return LJ_EINVAL; }
    int visible = 0;
    cudaError_t e = cudaGetDeviceCount(&visible);
    if (e != cudaSuccess || visible == 0) {
        lj_set_error("lj_init_devices: no CUDA device (%s); this library has no CPU path", e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
        return LJ_ENODEVICE;
    }
    std::vector<int> ids;
    if (!device_ids || n_devices <= 0) for (int i = 0; i < visible; i++) ids.push_back(i);
    else ids.assign(device_ids, device_ids + n_devices);
    for (size_t i = 0; i < ids.size(); i++) {
        if (ids[i] < 0 || ids[i] >= visible) { lj_set_error("lj_init_devices: device %d out of range [0,%d)", ids[i], visible); return LJ_EINVAL; }
        for (size_t j = 0; j < i; j++) if (ids[j] == ids[i]) { lj_set_error("lj_init_devices: device %d listed twice", ids[i]); return LJ_EINVAL; }
    }
    if (ids.size() == 1) return lj_init_single(ids[0], out);

    LjDeviceGroup *g = new LjDeviceGroup();
    g->ctx.assign(ids.size(), nullptr);
    g->stream.assign(ids.size(), nullptr);
    g->sum.assign(ids.size(), nullptr);
    for (size_t i = 0; i < ids.size(); i++) g->worker.push_back(new Worker());
    int rc = g->on_all([&](int i) {
        int r = lj_init_single(ids[i], &g->ctx[i]);
        if (r != LJ_OK) return r;
        LJ_CUDA(cudaStreamCreateWithFlags(&g->stream[i], cudaStreamNonBlocking));
        return LJ_OK;
    });
    lj_ctx *handle = new lj_ctx();
    handle->device = ids[0]; handle->sm_count = g->ctx[0] ? g->ctx[0]->sm_count : 0;
    handle->wf_pool = nullptr; handle->wf_pool_bytes = 0; handle->group = g;
    if (rc != LJ_OK) { std::string keep = lj_error_text(); lj_group_shutdown(handle); lj_set_error("%s", keep.c_str()); return rc; }
    *out = handle;
    return LJ_OK;
}

void lj_group_shutdown(lj_ctx *handle) {
    LjDeviceGroup *g = handle->group;
    for (ncclComm_t c : g->comm) if (c) g->nccl.CommDestroy(c);
    g->on_all([&](int i) {
        if (!g->ctx[i]) return LJ_OK;
        cudaSetDevice(g->ctx[i]->device);
        if (g->sum[i]) lj_free(g->sum[i]);
        if (g->stream[i]) cudaStreamDestroy(g->stream[i]);
        lj_shutdown_single(g->ctx[i]);
        return LJ_OK;
    });
    for (Worker *w : g->worker) delete w;
    delete g;
    delete handle;
}

int lj_group_scene_create(lj_ctx *handle, const lj_scene_desc *desc, uint32_t flags, lj_scene **out) {
    LjDeviceGroup *g = handle->group;
    lj_scene *h = new lj_scene();
    h->ctx = handle;
    memset(&h->d, 0, sizeof h->d); memset(&h->bvh, 0, sizeof h->bvh); memset(&h->build_stats, 0, sizeof h->build_stats);
    h->replicas.assign(g->n(), nullptr);
    int rc = g->on_all([&](int i) { return lj_scene_create_single(g->ctx[i], desc, flags, &h->replicas[i]); });
    if (rc != LJ_OK) { std::string keep = lj_error_text(); lj_group_scene_destroy(h); lj_set_error("%s", keep.c_str()); return rc; }
    h->d = h->replicas[0]->d;                       // film size / options for the entry points that only look
    h->build_stats = h->replicas[0]->build_stats;
    for (int i = 1; i < g->n(); i++) {              // the slowest device is what a caller waits for
        const lj_stats &b = h->replicas[i]->build_stats;
        if (b.bvh_build_ms > h->build_stats.bvh_build_ms) h->build_stats.bvh_build_ms = b.bvh_build_ms;
        if (b.upload_ms > h->build_stats.upload_ms) h->build_stats.upload_ms = b.upload_ms;
        if (b.tables_build_ms > h->build_stats.tables_build_ms) h->build_stats.tables_build_ms = b.tables_build_ms;
    }
    *out = h;
    return LJ_OK;
}

void lj_group_scene_destroy(lj_scene *h) {
    LjDeviceGroup *g = h->ctx->group;
    g->on_all([&](int i) { if (h->replicas[i]) lj_scene_destroy_single(h->replicas[i]); return LJ_OK; });
    delete h;
}

// dev_sum lives on the first device.  Every device renders its share into its own buffer (the first one straight
// into dev_sum), then one ncclReduce(sum) with the first device as root.
int lj_group_render_device(lj_scene *h, const lj_render_params *params, double *dev_sum, lj_stats *stats) {
    LjDeviceGroup *g = h->ctx->group;
    const int n = g->n();
    lj_render_params p;
    if (params) p = *params; else { memset(&p, 0, sizeof p); p.max_depth_override = -100; p.rng_mode = LJ_RNG_SAMPLE_STREAMS; }
    const DScene &d0 = h->replicas[0]->d;
    const size_t count = 3 * (size_t)d0.camera.width * d0.camera.height;
    // a caller that partitions by hand (tile_world / sample slice) gets exactly that, on the first device
    const bool manual = p.tile_world > 0 || p.sample_end > p.sample_begin;
    if (manual) {
        int rc = LJ_OK; std::string msg;
        g->worker[0]->post([&] { rc = lj_render_device_single(h->replicas[0], &p, dev_sum, stats); if (rc) msg = lj_error_text(); });
        g->worker[0]->wait();
        if (rc) lj_set_error("%s", msg.c_str());
        return rc;
    }
    const int integ = d0.options.integrator;
    const bool sampling = integ == LJ_INTEGRATOR_PATH || integ == LJ_INTEGRATOR_VOLPATH;
    const int spp = p.spp > 0 ? p.spp : d0.options.samples_per_pixel;
    int partition = p.partition == LJ_PARTITION_SPP ? LJ_PARTITION_SPP : LJ_PARTITION_TILES;
    if (const char *e = getenv("LJ_PARTITION")) if (p.partition == LJ_PARTITION_AUTO) partition = (e[0] == 's' || e[0] == 'S') ? LJ_PARTITION_SPP : LJ_PARTITION_TILES;
    if (partition == LJ_PARTITION_SPP && (!sampling || p.rng_mode != LJ_RNG_SAMPLE_STREAMS || spp < n)) partition = LJ_PARTITION_TILES;
    int rc = g->ensure_comms();
    if (rc != LJ_OK) return rc;

    if (g->sum_bytes < count * sizeof(double)) {
        rc = g->on_all([&](int i) {
            LJ_CUDA(cudaSetDevice(g->ctx[i]->device));
            if (g->sum[i]) lj_free(g->sum[i]);
            g->sum[i] = nullptr;
            if (i > 0) LJ_CUDA(lj_malloc(&g->sum[i], count * sizeof(double)));
            LJ_CUDA(cudaStreamSynchronize(0));
            return LJ_OK;
        });
        if (rc != LJ_OK) { g->sum_bytes = 0; return rc; }
        g->sum_bytes = count * sizeof(double);
    }
    std::vector<lj_stats> st(n);
    rc = g->on_all([&](int i) {
        lj_render_params pi = p;
        if (partition == LJ_PARTITION_TILES) { pi.tile_rank = i; pi.tile_world = n; }
        else {      // contiguous slices whose sizes differ by at most one sample
            const int base = spp / n, extra = spp % n;
            pi.spp = spp;
            pi.sample_begin = i * base + (i < extra ? i : extra);
            pi.sample_end = pi.sample_begin + base + (i < extra ? 1 : 0);
        }
        return lj_render_device_single(h->replicas[i], &pi, i == 0 ? dev_sum : g->sum[i], &st[i]);
    });
    if (rc != LJ_OK) return rc;

    // the one collective of the path: sum over devices, root = first device (in place there)
    cudaEvent_t r0, r1;
    LJ_CUDA(cudaSetDevice(g->ctx[0]->device));
    cudaEventCreate(&r0); cudaEventCreate(&r1);
    cudaEventRecord(r0, g->stream[0]);
    ncclResult_t nr = g->nccl.GroupStart();
    for (int i = 0; i < n && nr == ncclSuccess; i++) {
        double *buf = i == 0 ? dev_sum : g->sum[i];
        nr = g->nccl.Reduce(buf, buf, count, ncclFloat64, ncclSum, 0, g->comm[i], g->stream[i]);
    }
    ncclResult_t ne = g->nccl.GroupEnd();
    if (nr == ncclSuccess) nr = ne;
    cudaSetDevice(g->ctx[0]->device);
    cudaEventRecord(r1, g->stream[0]);
    cudaError_t ce = cudaSuccess;
    for (int i = 0; i < n; i++) { cudaSetDevice(g->ctx[i]->device); cudaError_t e = cudaStreamSynchronize(g->stream[i]); if (ce == cudaSuccess) ce = e; }
    cudaSetDevice(g->ctx[0]->device);
    float reduce_ms = 0;
    if (ce == cudaSuccess) cudaEventElapsedTime(&reduce_ms, r0, r1);
    cudaEventDestroy(r0); cudaEventDestroy(r1);
    if (nr != ncclSuccess) { lj_set_error("ncclReduce of the radiance sums failed: %s", g->nccl.GetErrorString(nr)); return LJ_ECUDA; }
    if (ce != cudaSuccess) { lj_set_error("reduction of the radiance sums failed: %s", cudaGetErrorString(ce)); return LJ_ECUDA; }

    if (stats) {
        lj_stats t = st[0];
        t.device_ms_min = t.device_ms_max = st[0].render_ms;
        for (int i = 1; i < n; i++) {
            const lj_stats &a = st[i];
            if (partition == LJ_PARTITION_SPP) t.samples += a.samples;
            t.rays_primary += a.rays_primary; t.rays_extension += a.rays_extension; t.rays_shadow += a.rays_shadow;
            t.nodes_visited += a.nodes_visited; t.prims_tested += a.prims_tested;
            t.shadow_nodes_visited += a.shadow_nodes_visited; t.shadow_prims_tested += a.shadow_prims_tested;
            t.kernel_launches += a.kernel_launches;
            if (a.iterations > t.iterations) t.iterations = a.iterations;
            if (a.extend_ms > t.extend_ms) t.extend_ms = a.extend_ms;
            if (a.shadow_ms > t.shadow_ms) t.shadow_ms = a.shadow_ms;
            if (a.shade_ms > t.shade_ms) t.shade_ms = a.shade_ms;
            if (a.generate_ms > t.generate_ms) t.generate_ms = a.generate_ms;
            if (a.render_ms < t.device_ms_min) t.device_ms_min = a.render_ms;
            if (a.render_ms > t.device_ms_max) t.device_ms_max = a.render_ms;
        }
        t.render_ms = t.device_ms_max + reduce_ms;
        t.reduce_ms = reduce_ms; t.n_devices = n; t.partition = partition;
        t.upload_ms = h->build_stats.upload_ms; t.bvh_build_ms = h->build_stats.bvh_build_ms;
        *stats = t;
    }
    return LJ_OK;
}
