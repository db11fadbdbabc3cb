// group.cpp -- ONE process driving several GPUs: the multi-device half of the drop-in.
//
// The reference drives one device (`Handler handle(0)`, cuda_sim.cu:916).  Walkers never interact and
// the scene is read-only, so the path shards by walker: the Morton-sorted walker array
// (SimulationBox.cpp:131-147) is cut into contiguous ranges, one per device, the scene is replicated,
// every walker keeps the random stream of its GLOBAL index.  The one exchange is the signal: per sprint
// one in-place SUM all-reduce of sprintSteps/pfreq doubles (NCCL, NVLink/NVSwitch), enqueued on each
// member's side stream so that it overlaps the next sprint's kernels.
//
// NCCL in a single process: the all-reduce calls of all devices go inside one
// ncclGroupStart/ncclGroupEnd, which defers them to the GroupEnd -- hence the three phases of
// ctx_internal.h (kernels | grouped all-reduce | D2H + completion event).
#include <cuda_runtime.h>

#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "checkpoint_internal.h"
#include "ctx_internal.h"
#include "mrisim.h"
#include "nccl_dyn.h"

struct mrisim_group {
    std::vector<mrisim_ctx*> ctx;
    std::vector<int> devices;
    std::vector<ncclComm_t> comms;
    std::vector<uint64_t> lo, count;          // member ranges within the group's walker array
    uint64_t n = 0, first_walker_id = 0, walker_seed = 0;
    std::string err;
};

namespace {

thread_local std::string g_group_create_error;

int gfail(mrisim_group* g, const std::string& m) { g->err = m; return 1; }
int from_ctx(mrisim_group* g, int i, const char* what) {
    g->err = std::string(what) + " (device " + std::to_string(g->devices[i]) + "): " + mrisim_last_error(g->ctx[i]);
    return 1;
}

}  // namespace

extern "C" {

const char* mrisim_group_last_error(const mrisim_group* g) { return g ? g->err.c_str() : g_group_create_error.c_str(); }

int mrisim_group_create(const mrisim_params* p, const int* devices, int n_devices, mrisim_group** out) {
    if (!p || !devices || n_devices < 1 || !out) { g_group_create_error = "group_create: bad argument"; return 1; }
    for (int i = 0; i < n_devices; i++)
        for (int j = 0; j < i; j++)
            if (devices[i] == devices[j]) { g_group_create_error = "group_create: a device is listed twice"; return 1; }
    std::unique_ptr<mrisim_group> g(new mrisim_group);
    g->devices.assign(devices, devices + n_devices);
    auto cleanup = [&]() { for (mrisim_ctx* c : g->ctx) mrisim_destroy(c); g->ctx.clear(); };
    for (int i = 0; i < n_devices; i++) {
        mrisim_ctx* c = nullptr;
        if (mrisim_create(p, devices[i], &c)) { g_group_create_error = mrisim_last_error(nullptr); cleanup(); return 1; }
        g->ctx.push_back(c);
    }
    if (n_devices > 1) {
        if (!mrisim_nccl::ok()) { g_group_create_error = mrisim_nccl::api().error; cleanup(); return 1; }
        g->comms.assign((size_t)n_devices, nullptr);
        const ncclResult_t r = mrisim_nccl::api().CommInitAll(g->comms.data(), n_devices, devices);
        if (r != ncclSuccess) {
            g_group_create_error = std::string("ncclCommInitAll: ") + mrisim_nccl::api().GetErrorString(r);
            cleanup(); return 1;
        }
        for (int i = 0; i < n_devices; i++) mrisim_comm_attach(g->ctx[i], g->comms[i], n_devices, i);
    }
    *out = g.release();
    return 0;
}

void mrisim_group_destroy(mrisim_group* g) {
    if (!g) return;
    for (size_t i = 0; i < g->ctx.size(); i++) {
        // drain what is outstanding so that no all-reduce is in flight when the communicators go
        while (mrisim_pending_sprints(g->ctx[i]) > 0 && mrisim_wait_sprint(g->ctx[i], nullptr) == 0) {}
        mrisim_comm_attach(g->ctx[i], nullptr, 1, 0);
    }
    for (size_t i = 0; i < g->comms.size(); i++)
        if (g->comms[i]) { cudaSetDevice(g->devices[i]); mrisim_nccl::api().CommDestroy(g->comms[i]); }
    for (mrisim_ctx* c : g->ctx) mrisim_destroy(c);
    delete g;
}

int mrisim_group_size(const mrisim_group* g) { return (int)g->ctx.size(); }
mrisim_ctx* mrisim_group_ctx(mrisim_group* g, int i) { return (i >= 0 && i < (int)g->ctx.size()) ? g->ctx[(size_t)i] : nullptr; }

int mrisim_group_set_scene(mrisim_group* g, const mrisim_scene* s) {
    for (size_t i = 0; i < g->ctx.size(); i++)
        if (mrisim_set_scene(g->ctx[i], s)) return from_ctx(g, (int)i, "scene upload");
    return 0;
}

int mrisim_group_set_waters(mrisim_group* g, const mrisim_water* waters, uint64_t n, uint64_t walker_seed, uint64_t first_walker_id) {
    const uint64_t k = g->ctx.size();
    if (!waters || n < k) return gfail(g, "group_set_waters: fewer walkers than devices");
    g->lo.assign(k, 0); g->count.assign(k, 0);
    for (uint64_t i = 0; i < k; i++) {                      // contiguous ranges, sizes differ by at most one
        g->lo[i] = n * i / k;
        g->count[i] = n * (i + 1) / k - g->lo[i];
        if (mrisim_set_waters(g->ctx[i], waters + g->lo[i], g->count[i], walker_seed, first_walker_id + g->lo[i])) return from_ctx(g, (int)i, "water upload");
    }
    g->n = n; g->first_walker_id = first_walker_id; g->walker_seed = walker_seed;
    return 0;
}

int mrisim_group_get_waters(mrisim_group* g, mrisim_water* out, uint64_t n) {
    if (!out || n != g->n) return gfail(g, "group_get_waters: n does not match the walkers held");
    for (size_t i = 0; i < g->ctx.size(); i++)
        if (mrisim_get_waters(g->ctx[i], out + g->lo[i], g->count[i])) return from_ctx(g, (int)i, "get_waters");
    return 0;
}

int mrisim_group_get_rng(mrisim_group* g, mrisim_rng_state* out, uint64_t n) {
    if (!out || n != g->n) return gfail(g, "group_get_rng: n does not match the walkers held");
    for (size_t i = 0; i < g->ctx.size(); i++)
        if (mrisim_get_rng(g->ctx[i], out + g->lo[i], g->count[i])) return from_ctx(g, (int)i, "get_rng");
    return 0;
}

int64_t mrisim_group_get_time(const mrisim_group* g) { return g->ctx.empty() ? 0 : mrisim_get_time(g->ctx[0]); }

int mrisim_group_submit_sprint(mrisim_group* g, int32_t steps) {
    if (!g->n) return gfail(g, "group_submit_sprint: no waters set");
    const size_t k = g->ctx.size();
    for (size_t i = 0; i < k; i++)
        if (mrisim_ctx_phase_compute(g->ctx[i], steps, nullptr)) return from_ctx(g, (int)i, "sprint");
    if (k > 1) {
        const mrisim_nccl::Api& nccl = mrisim_nccl::api();
        ncclResult_t r = nccl.GroupStart();
        if (r != ncclSuccess) return gfail(g, std::string("ncclGroupStart: ") + nccl.GetErrorString(r));
        for (size_t i = 0; i < k; i++)
            if (mrisim_ctx_phase_reduce(g->ctx[i])) { nccl.GroupEnd(); return from_ctx(g, (int)i, "all-reduce"); }
        r = nccl.GroupEnd();
        if (r != ncclSuccess) return gfail(g, std::string("ncclGroupEnd: ") + nccl.GetErrorString(r));
    }
    for (size_t i = 0; i < k; i++)
        if (mrisim_ctx_phase_finish(g->ctx[i])) return from_ctx(g, (int)i, "sprint");
    return 0;
}

int mrisim_group_wait_sprint(mrisim_group* g, double* signal_out) {
    // every member holds the same all-reduced samples; member 0's copy is the one returned, the others
    // are waited for so that their error flags are seen and their ring slots are released
    for (size_t i = 0; i < g->ctx.size(); i++)
        if (mrisim_wait_sprint(g->ctx[i], i == 0 ? signal_out : nullptr)) return from_ctx(g, (int)i, "sprint");
    return 0;
}

int mrisim_group_run_sprint(mrisim_group* g, int32_t steps, double* signal_out) {
    if (mrisim_group_submit_sprint(g, steps)) return 1;
    return mrisim_group_wait_sprint(g, signal_out);
}

uint64_t mrisim_group_kernel_launches(const mrisim_group* g) {
    uint64_t t = 0;
    for (mrisim_ctx* c : g->ctx) t += mrisim_kernel_launches(c);
    return t;
}

int mrisim_group_checkpoint_save(mrisim_group* g, const char* path, uint64_t host_seed, uint64_t csv_bytes) {
    if (!path || !g->n) return gfail(g, "group_checkpoint_save: bad argument or no waters");
    std::vector<mrisim_ckpt::ShardState> shards(g->ctx.size());
    for (size_t i = 0; i < g->ctx.size(); i++) {
        if (mrisim_pending_sprints(g->ctx[i])) return gfail(g, "group_checkpoint_save: submitted sprints are still outstanding");
        if (mrisim_ckpt::gather(g->ctx[i], shards[i])) return gfail(g, mrisim_ckpt::last_error());
    }
    mrisim_checkpoint_info info;
    std::memset(&info, 0, sizeof info);
    info.params_hash = mrisim_params_hash(mrisim_ctx_params(g->ctx[0]));
    info.scene_digest = mrisim_ctx_scene_digest(g->ctx[0]);
    info.host_seed = host_seed; info.walker_seed = g->walker_seed;
    info.time_steps = mrisim_get_time(g->ctx[0]); info.csv_bytes = csv_bytes;
    info.n_walkers = g->n; info.first_walker_id = g->first_walker_id;
    if (mrisim_ckpt::write_file(path, info, shards)) return gfail(g, mrisim_ckpt::last_error());
    return 0;
}

int mrisim_group_checkpoint_load(mrisim_group* g, const char* path) {
    mrisim_checkpoint_info info;
    if (mrisim_ckpt::read_info(path, &info, nullptr)) return gfail(g, mrisim_ckpt::last_error());
    const uint64_t k = g->ctx.size(), n = info.n_walkers;
    if (n < k) return gfail(g, "group_checkpoint_load: fewer walkers in the file than devices");
    g->lo.assign(k, 0); g->count.assign(k, 0);
    for (uint64_t i = 0; i < k; i++) {
        if (mrisim_ckpt::check_compatible(g->ctx[i], info)) return gfail(g, mrisim_ckpt::last_error());
        g->lo[i] = n * i / k;
        g->count[i] = n * (i + 1) / k - g->lo[i];
        mrisim_ckpt::ShardState st;
        if (mrisim_ckpt::read_range(path, g->lo[i], g->count[i], st)) return gfail(g, mrisim_ckpt::last_error());
        if (mrisim_ckpt::scatter(g->ctx[i], info, st)) return gfail(g, mrisim_ckpt::last_error());
    }
    g->n = n; g->first_walker_id = info.first_walker_id; g->walker_seed = info.walker_seed;
    return 0;
}

}  // extern "C"
