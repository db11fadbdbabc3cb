// checkpoint.cpp -- on-disk checkpoint / resume of the walker state.
//
// The reference writes nothing but the signal CSV; its walker state does round-trip through
// d.waters once per sprint (cuda_sim.cu:694,744), so a sprint boundary is the natural checkpoint and
// the CSV row (:1022) the natural unit to resume after.  Everything here sits on the public C ABI
// (get/set_waters, get/set_rng, get/set_order, get/set_time): no device code.
//
// File "MRICKP01" (little-endian):
//   Header   (below, 128 bytes)
//   Shard    n_shards x {u64 first, u64 count, u32 has_order, u32 pad}   first is relative to walker 0 of the file
//   f64      x[n] y[n] z[n] phase[n] spare[n]
//   u64      s0[n] s1[n]
//   i32      nearest[n]
//   u8       in_cell[n], padded to a multiple of 8 bytes
//   per shard with has_order: u32 perm[count] keys[count]   (perm holds indices local to the shard)
#include <cerrno>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "checkpoint_internal.h"
#include "mrisim.h"

namespace {

thread_local std::string g_ckpt_error;

struct Header {
    char     magic[8];
    uint32_t version, n_shards;
    uint64_t params_hash, scene_digest, host_seed, walker_seed;
    int64_t  time_steps;
    uint64_t csv_bytes, n_walkers, first_walker_id;
    uint64_t reserved[6];
};
static_assert(sizeof(Header) == 128, "checkpoint header layout");

struct ShardRec { uint64_t first, count; uint32_t has_order, pad; };
static_assert(sizeof(ShardRec) == 24, "checkpoint shard record layout");

int fail(const std::string& m) { g_ckpt_error = m; return 1; }

uint64_t fnv(uint64_t h, const void* data, size_t bytes) {
    const unsigned char* b = static_cast<const unsigned char*>(data);
    for (size_t i = 0; i < bytes; i++) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}

bool put(std::FILE* f, const void* p, size_t bytes) { return bytes == 0 || std::fwrite(p, 1, bytes, f) == bytes; }
bool get(std::FILE* f, void* p, size_t bytes) { return bytes == 0 || std::fread(p, 1, bytes, f) == bytes; }

struct FileGuard { std::FILE* f = nullptr; ~FileGuard() { if (f) std::fclose(f); } };

// byte offsets of the arrays for a file of n walkers
struct Layout {
    uint64_t n, f64_0, u64_0, nearest_0, in_cell_0, order_0;
    Layout(uint64_t n_, uint32_t n_shards) : n(n_) {
        f64_0 = sizeof(Header) + (uint64_t)n_shards * sizeof(ShardRec);
        u64_0 = f64_0 + 5 * n * 8;
        nearest_0 = u64_0 + 2 * n * 8;
        in_cell_0 = nearest_0 + n * 4;
        order_0 = in_cell_0 + (n + 7) / 8 * 8;
    }
};

}  // namespace

namespace mrisim_ckpt {

const std::string& last_error() { return g_ckpt_error; }
void set_error(const std::string& m) { g_ckpt_error = m; }

int gather(mrisim_ctx* ctx, ShardState& out) {
    const uint64_t n = mrisim_walker_count(ctx);
    if (!n) return fail("checkpoint: the context holds no walkers");
    out.first = mrisim_first_walker_id(ctx);
    out.waters.resize(n); out.rng.resize(n); out.perm.resize(n); out.keys.resize(n);
    if (mrisim_get_waters(ctx, out.waters.data(), n)) return fail(std::string("checkpoint: ") + mrisim_last_error(ctx));
    if (mrisim_get_rng(ctx, out.rng.data(), n)) return fail(std::string("checkpoint: ") + mrisim_last_error(ctx));
    int32_t valid = 0;
    if (mrisim_get_order(ctx, out.perm.data(), out.keys.data(), n, &valid)) return fail(std::string("checkpoint: ") + mrisim_last_error(ctx));
    out.has_order = valid != 0;
    if (!out.has_order) { out.perm.clear(); out.keys.clear(); }
    return 0;
}

int write_file(const char* path, const mrisim_checkpoint_info& info, const std::vector<ShardState>& shards) {
    if (!path || shards.empty()) return fail("checkpoint: bad argument");
    uint64_t n = 0;
    for (const ShardState& s : shards) {
        if (s.first != info.first_walker_id + n) return fail("checkpoint: shards are not contiguous in global walker order");
        n += s.waters.size();
    }
    if (n != info.n_walkers) return fail("checkpoint: shard sizes do not add up");
    const std::string tmp = std::string(path) + ".tmp";
    FileGuard fg;
    fg.f = std::fopen(tmp.c_str(), "wb");
    if (!fg.f) return fail("checkpoint: cannot open " + tmp + ": " + std::strerror(errno));
    Header h;
    std::memset(&h, 0, sizeof h);
    std::memcpy(h.magic, "MRICKP01", 8);
    h.version = 1; h.n_shards = (uint32_t)shards.size();
    h.params_hash = info.params_hash; h.scene_digest = info.scene_digest; h.host_seed = info.host_seed; h.walker_seed = info.walker_seed;
    h.time_steps = info.time_steps; h.csv_bytes = info.csv_bytes; h.n_walkers = n; h.first_walker_id = info.first_walker_id;
    bool ok = put(fg.f, &h, sizeof h);
    uint64_t first = 0;
    for (const ShardState& s : shards) {
        const ShardRec r{first, (uint64_t)s.waters.size(), s.has_order ? 1u : 0u, 0u};
        ok = ok && put(fg.f, &r, sizeof r);
        first += s.waters.size();
    }
    std::vector<double> col;
    auto put_f64 = [&](double mrisim_water::*m) {
        for (const ShardState& s : shards) {
            col.resize(s.waters.size());
            for (size_t i = 0; i < col.size(); i++) col[i] = s.waters[i].*m;
            ok = ok && put(fg.f, col.data(), col.size() * 8);
        }
    };
    put_f64(&mrisim_water::x); put_f64(&mrisim_water::y); put_f64(&mrisim_water::z); put_f64(&mrisim_water::phase);
    for (const ShardState& s : shards) { col.resize(s.rng.size()); for (size_t i = 0; i < col.size(); i++) col[i] = s.rng[i].spare; ok = ok && put(fg.f, col.data(), col.size() * 8); }
    std::vector<uint64_t> u;
    for (int k = 0; k < 2; k++)
        for (const ShardState& s : shards) { u.resize(s.rng.size()); for (size_t i = 0; i < u.size(); i++) u[i] = k ? s.rng[i].s1 : s.rng[i].s0; ok = ok && put(fg.f, u.data(), u.size() * 8); }
    std::vector<int32_t> nr;
    for (const ShardState& s : shards) { nr.resize(s.waters.size()); for (size_t i = 0; i < nr.size(); i++) nr[i] = s.waters[i].nearest; ok = ok && put(fg.f, nr.data(), nr.size() * 4); }
    std::vector<uint8_t> ic;
    for (const ShardState& s : shards) { ic.resize(s.waters.size()); for (size_t i = 0; i < ic.size(); i++) ic[i] = s.waters[i].in_cell; ok = ok && put(fg.f, ic.data(), ic.size()); }
    const uint64_t padb = (n + 7) / 8 * 8 - n;
    const uint8_t zeros[8] = {0};
    ok = ok && put(fg.f, zeros, padb);
    for (const ShardState& s : shards)
        if (s.has_order) { ok = ok && put(fg.f, s.perm.data(), s.perm.size() * 4) && put(fg.f, s.keys.data(), s.keys.size() * 4); }
    ok = ok && std::fflush(fg.f) == 0;
    std::fclose(fg.f); fg.f = nullptr;
    if (!ok) { std::remove(tmp.c_str()); return fail("checkpoint: short write to " + tmp); }
    if (std::rename(tmp.c_str(), path) != 0) return fail(std::string("checkpoint: rename to ") + path + ": " + std::strerror(errno));
    return 0;
}

int read_info(const char* path, mrisim_checkpoint_info* out, std::vector<ShardEntry>* shards) {
    if (!path || !out) return fail("checkpoint: bad argument");
    FileGuard fg;
    fg.f = std::fopen(path, "rb");
    if (!fg.f) return fail(std::string("checkpoint: cannot open ") + path + ": " + std::strerror(errno));
    Header h;
    if (!get(fg.f, &h, sizeof h) || std::memcmp(h.magic, "MRICKP01", 8) != 0 || h.version != 1) return fail(std::string("checkpoint: ") + path + " is not an MRICKP01 file");
    if (h.n_shards == 0 || h.n_shards > 4096 || h.n_walkers == 0 || h.n_walkers >= (1ull << 40) || h.time_steps < 0)
        return fail("checkpoint: implausible header");
    std::vector<ShardRec> recs(h.n_shards);
    if (!get(fg.f, recs.data(), recs.size() * sizeof(ShardRec))) return fail("checkpoint: truncated shard table");
    uint64_t first = 0, order_bytes = 0;
    for (const ShardRec& r : recs) {
        if (r.first != first || r.count == 0 || r.count > h.n_walkers) return fail("checkpoint: inconsistent shard table");
        first += r.count;
        if (r.has_order) order_bytes += r.count * 8;
    }
    if (first != h.n_walkers) return fail("checkpoint: shard table does not cover the walkers");
    const Layout lay(h.n_walkers, h.n_shards);
    std::fseek(fg.f, 0, SEEK_END);
    const long size = std::ftell(fg.f);
    if (size < 0 || (uint64_t)size != lay.order_0 + order_bytes) return fail("checkpoint: file size does not match its header (truncated?)");
    out->params_hash = h.params_hash; out->scene_digest = h.scene_digest; out->host_seed = h.host_seed; out->walker_seed = h.walker_seed;
    out->time_steps = h.time_steps; out->csv_bytes = h.csv_bytes; out->n_walkers = h.n_walkers; out->first_walker_id = h.first_walker_id;
    out->n_shards = h.n_shards; out->version = h.version;
    if (shards) {
        shards->clear();
        uint64_t off = lay.order_0;
        for (const ShardRec& r : recs) {
            shards->push_back(ShardEntry{r.first, r.count, r.has_order != 0, off});
            if (r.has_order) off += r.count * 8;
        }
    }
    return 0;
}

int read_range(const char* path, uint64_t lo, uint64_t count, ShardState& out) {
    mrisim_checkpoint_info info;
    std::vector<ShardEntry> shards;
    if (read_info(path, &info, &shards)) return 1;
    if (count == 0) count = info.n_walkers - (lo < info.n_walkers ? lo : info.n_walkers);
    if (lo + count > info.n_walkers || count == 0) return fail("checkpoint: walker range outside the file");
    FileGuard fg;
    fg.f = std::fopen(path, "rb");
    if (!fg.f) return fail(std::string("checkpoint: cannot open ") + path);
    const Layout lay(info.n_walkers, info.n_shards);
    const uint64_t n = info.n_walkers;
    out.first = info.first_walker_id + lo;
    out.waters.assign(count, mrisim_water{});
    out.rng.assign(count, mrisim_rng_state{});
    std::vector<double> col(count);
    bool ok = true;
    auto col_f64 = [&](int k) { ok = ok && std::fseek(fg.f, (long)(lay.f64_0 + ((uint64_t)k * n + lo) * 8), SEEK_SET) == 0 && get(fg.f, col.data(), count * 8); };
    col_f64(0); for (uint64_t i = 0; i < count; i++) out.waters[i].x = col[i];
    col_f64(1); for (uint64_t i = 0; i < count; i++) out.waters[i].y = col[i];
    col_f64(2); for (uint64_t i = 0; i < count; i++) out.waters[i].z = col[i];
    col_f64(3); for (uint64_t i = 0; i < count; i++) out.waters[i].phase = col[i];
    col_f64(4); for (uint64_t i = 0; i < count; i++) out.rng[i].spare = col[i];
    std::vector<uint64_t> u(count);
    for (int k = 0; k < 2; k++) {
        ok = ok && std::fseek(fg.f, (long)(lay.u64_0 + ((uint64_t)k * n + lo) * 8), SEEK_SET) == 0 && get(fg.f, u.data(), count * 8);
        for (uint64_t i = 0; i < count; i++) (k ? out.rng[i].s1 : out.rng[i].s0) = u[i];
    }
    std::vector<int32_t> nr(count);
    ok = ok && std::fseek(fg.f, (long)(lay.nearest_0 + lo * 4), SEEK_SET) == 0 && get(fg.f, nr.data(), count * 4);
    std::vector<uint8_t> ic(count);
    ok = ok && std::fseek(fg.f, (long)(lay.in_cell_0 + lo), SEEK_SET) == 0 && get(fg.f, ic.data(), count);
    for (uint64_t i = 0; i < count; i++) { out.waters[i].nearest = nr[i]; out.waters[i].in_cell = ic[i]; }
    out.has_order = false; out.perm.clear(); out.keys.clear();
    for (const ShardEntry& s : shards)
        if (s.first == lo && s.count == count && s.has_order) {
            out.perm.resize(count); out.keys.resize(count);
            ok = ok && std::fseek(fg.f, (long)s.order_offset, SEEK_SET) == 0 && get(fg.f, out.perm.data(), count * 4) && get(fg.f, out.keys.data(), count * 4);
            out.has_order = ok;
        }
    if (!ok) return fail("checkpoint: short read");
    return 0;
}

int scatter(mrisim_ctx* ctx, const mrisim_checkpoint_info& info, const ShardState& st) {
    const uint64_t n = st.waters.size();
    if (mrisim_set_waters(ctx, st.waters.data(), n, info.walker_seed, st.first)) return fail(std::string("checkpoint: ") + mrisim_last_error(ctx));
    if (mrisim_set_rng(ctx, st.rng.data(), n)) return fail(std::string("checkpoint: ") + mrisim_last_error(ctx));
    if (mrisim_set_time(ctx, info.time_steps)) return fail(std::string("checkpoint: ") + mrisim_last_error(ctx));
    if (st.has_order && mrisim_set_order(ctx, st.perm.data(), st.keys.data(), n)) return fail(std::string("checkpoint: ") + mrisim_last_error(ctx));
    return 0;
}

int check_compatible(mrisim_ctx* ctx, const mrisim_checkpoint_info& info) {
    if (info.params_hash != mrisim_params_hash(mrisim_ctx_params(ctx)))
        return fail("checkpoint: written with different parameters (parameter hash differs)");
    const uint64_t dig = mrisim_ctx_scene_digest(ctx);
    if (!dig) return fail("checkpoint: set the scene before loading a checkpoint");
    if (info.scene_digest != dig) return fail("checkpoint: written for a different scene (scene digest differs)");
    return 0;
}

}  // namespace mrisim_ckpt

extern "C" {

const char* mrisim_checkpoint_last_error(void) { return g_ckpt_error.c_str(); }

uint64_t mrisim_params_hash(const mrisim_params* p) {
    if (!p) return 0;
    // field by field (the struct has padding); num_threads, totaltime and t do not change what a walker does
    uint64_t h = 1469598103934665603ull;
#define H(f) h = fnv(h, &p->f, sizeof p->f)
    H(g); H(scale); H(border); H(sprintSteps); H(num_cells); H(cell_r); H(mmoment); H(kick_mode);
    H(avoid_intracellular_throw); H(phase_stdev); H(phase_k); H(bound); H(water_start_bound); H(hashDim); H(maxNeighbors);
    H(D_cell); H(D_extra); H(P_expr); H(tau); H(taucp); H(tcp); H(pfreq); H(reflectIO); H(reflectOI); H(in_stdev); H(out_stdev);
#undef H
    return h;
}

int mrisim_checkpoint_save(mrisim_ctx* ctx, const char* path, uint64_t host_seed, uint64_t csv_bytes) {
    if (!ctx || !path) return fail("checkpoint_save: null argument");
    if (mrisim_pending_sprints(ctx)) return fail("checkpoint_save: submitted sprints are still outstanding (mrisim_wait_sprint)");
    std::vector<mrisim_ckpt::ShardState> shards(1);
    if (mrisim_ckpt::gather(ctx, shards[0])) return 1;
    mrisim_checkpoint_info info;
    std::memset(&info, 0, sizeof info);
    info.params_hash = mrisim_params_hash(mrisim_ctx_params(ctx));
    info.scene_digest = mrisim_ctx_scene_digest(ctx);
    info.host_seed = host_seed; info.walker_seed = mrisim_walker_seed(ctx);
    info.time_steps = mrisim_get_time(ctx); info.csv_bytes = csv_bytes;
    info.n_walkers = shards[0].waters.size(); info.first_walker_id = shards[0].first;
    return mrisim_ckpt::write_file(path, info, shards);
}

int mrisim_checkpoint_info_read(const char* path, mrisim_checkpoint_info* out) { return mrisim_ckpt::read_info(path, out, nullptr); }

int mrisim_checkpoint_load(mrisim_ctx* ctx, const char* path, uint64_t lo, uint64_t count) {
    if (!ctx || !path) return fail("checkpoint_load: null argument");
    mrisim_checkpoint_info info;
    if (mrisim_ckpt::read_info(path, &info, nullptr)) return 1;
    if (mrisim_ckpt::check_compatible(ctx, info)) return 1;
    mrisim_ckpt::ShardState st;
    if (mrisim_ckpt::read_range(path, lo, count, st)) return 1;
    return mrisim_ckpt::scatter(ctx, info, st);
}

}  // extern "C"
