// scene.cpp -- host scene builder of the walker path (product code).
//
// Builds, from run-time parameters and one 64-bit seed, exactly what the reference's
//   BacteriaBox simBox(&gen); simBox.populateSimulation();        (src/cuda_sim.cu:920-921)
// hands to its GPU driver: cell centres, dipoles incl. periodic images, the nearest-cell
// lookup table, the hashed octree forest and the Morton-sorted waters.  Every array is
// bit-identical to the reference's for the same seed (tests/test_scene_parity.py compares
// whole MRISCN01 files against dumps of the reference's own host code); the structure is
// ours: flat arrays instead of vector-of-vector-of-pointer, index lists instead of copied
// MNP records during descent, a table-driven periodic-image enumerator, dynamic work
// distribution over all host threads.
//
// Must be compiled like the oracle side compiles the reference (-O2 -ffp-contract=off, same
// glibc): the leaf criterion and the float B samples go through libm pow().
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "mrisim.h"
#include "mrisim_math.h"
#include "scene_internal.h"

static_assert(sizeof(mrisim_params) == 192 && sizeof(mrisim_water) == 40 && sizeof(mrisim_mnp) == 40 &&
              sizeof(mrisim_node) == 48 && sizeof(mrisim_rng_state) == 24, "C ABI layouts (mirrored in _cabi.py)");

namespace {

thread_local std::string g_scene_error;

inline double normsq(double x, double y, double z) { return x * x + y * y + z * z; }

// XORShift<uint64_t> of the reference (src/xorshift.h): same generator the walker stream
// contract uses, here as the HOST stream for cell and water placement.
struct HostRng {
    mrisim_rng r;
    explicit HostRng(uint64_t seed) { mrisim_rng_seed(&r, seed); }
    double rand_pos_double() { return mrisim_rng_u01(&r); }
};

// ---- Morton arithmetic (host flavour: src/octree.cpp:37-72) -------------------------
inline uint64_t spread3(uint32_t v) {            // bit b -> bit 3b, 21 bits
    uint64_t x = v & 0x1FFFFF;
    x = (x | x << 32) & 0x1F00000000FFFFULL;
    x = (x | x << 16) & 0x1F0000FF0000FFULL;
    x = (x | x << 8)  & 0x100F00F00F00F00FULL;
    x = (x | x << 4)  & 0x10C30C30C30C30C3ULL;
    x = (x | x << 2)  & 0x1249249249249249ULL;
    return x;
}

inline uint64_t host_morton(int depth, double x, double y, double z, double bound) {
    double size = std::pow(2, depth);
    uint32_t ix = std::floor(x / bound * size);
    uint32_t iy = std::floor(y / bound * size);
    uint32_t iz = std::floor(z / bound * size);
    // the reference ORs three byte-wide LUT slices per axis; for depth <= 8 only the low
    // byte contributes, for deeper codes the 24-bit groups are concatenated -- which is
    // plain bit interleaving with x in the high position (octree.h:83-94)
    const uint32_t mask = depth > 16 ? 0xFFFFFFu : (depth > 8 ? 0xFFFFu : 0xFFu);
    return (spread3(ix & mask) << 2) | (spread3(iy & mask) << 1) | spread3(iz & mask);
}

// ---- octree build -------------------------------------------------------------------
struct SubtreeOut {
    std::vector<mrisim_node> nodes;     // DFS order, child offsets relative to the parent
    std::vector<mrisim_mnp>  residents; // leaf resident lists, back to back in node order
};

class OctreeBuilder {
public:
    OctreeBuilder(const mrisim_params& p, const std::vector<mrisim_mnp>& mnps,
                  double max_prod, double max_g, double min_g)
        : p_(p), mnps_(mnps), max_prod_(max_prod), min_g_(min_g) {
        // octree.cpp:400-401
        min_depth = std::ceil(std::log(p.bound / max_g) / std::log(2));
        max_depth = std::ceil(std::log(p.bound / min_g) / std::log(2));
    }
    int min_depth, max_depth;

    // octree.cpp:344-371 (one hashed subtree; `offset` is its index in the root table)
    void build_subtree_at(int offset, SubtreeOut& out) const {
        double x, y, z, g; uint64_t mc;
        root_of(offset, &x, &y, &z, &g, &mc);
        std::vector<int> all(mnps_.size());
        for (size_t i = 0; i < all.size(); i++) all[i] = (int)i;
        std::vector<int> root_res = filter(all, x, y, z, g);
        out.nodes.clear(); out.residents.clear();
        descend(out, mc, root_res, x, y, z, g);
    }

    // seams for the GPU builder (scene_internal.h): the same expressions, one value at a time
    double space_limit(double g) const { return std::pow(g / 2 * std::sqrt(3) + p_.scale * p_.cell_r, 2); }
    double criterion(double x, double y, double z, double g) const {
        return g / 2 * 1e-6 * grad(x + g / 2, y + g / 2, z + g / 2);
    }
    double leaf_B(const int32_t* res, size_t n_res, double xn, double yn, double zn) const {
        double B = field(xn, yn, zn);
        for (size_t k = 0; k < n_res; k++) {
            const mrisim_mnp& m = mnps_[(size_t)res[k]];
            double dx = xn - m.x, dy = yn - m.y, dz = zn - m.z;
            B -= dipole(dx, dy, dz, m.M);
        }
        return B;
    }
    void root_of(int offset, double* xo, double* yo, double* zo, double* go, uint64_t* mco) const {
        double x = 0, y = 0, z = 0, g = p_.bound;
        int temp = offset;
        for (int j = min_depth - 1; j >= 0; j--) {
            int pw = (int)std::pow(8, j);
            int oct_id = temp / pw;
            double dx = (double)(oct_id >> 2) * g / 2;
            double dy = (double)((oct_id >> 1) & 1) * g / 2;
            double dz = (double)(oct_id & 1) * g / 2;
            x += dx; y += dy; z += dz;
            temp %= pw;
            g /= 2;
        }
        uint64_t mc = host_morton(min_depth, x, y, z, p_.bound);
        mc |= (uint64_t)(int64_t)(1 << (3 * min_depth));       // int shift, octree.cpp:367
        *xo = x; *yo = y; *zo = z; *go = g; *mco = mc;
    }

private:
    const mrisim_params& p_;
    const std::vector<mrisim_mnp>& mnps_;
    double max_prod_, min_g_;

    // octree.cpp:168-174 MNP_in_space + :183-201 transfer_MNPs (index lists, same order)
    std::vector<int> filter(const std::vector<int>& src, double x, double y, double z, double g) const {
        std::vector<int> dst;
        const double lim = space_limit(g);
        for (int i : src) {
            const mrisim_mnp& m = mnps_[i];
            double dx = x + g / 2 - m.x;
            double dy = y + g / 2 - m.y;
            double dz = z + g / 2 - m.z;
            if (normsq(dx, dy, dz) < lim) dst.push_back(i);
        }
        return dst;
    }

    // octree.cpp:206-210 (HOST dipole: unmasked, pow(r^2, 2.5))
    static double dipole(double dx, double dy, double dz, double M) {
        double divisor = std::pow(normsq(dx, dy, dz), 2.5);
        return M * 1e11 * (2 * dz * dz - dx * dx - dy * dy) / divisor;
    }
    // octree.cpp:215-222
    double field(double x, double y, double z) const {
        double ret = 0;
        for (const mrisim_mnp& m : mnps_) ret += dipole(x - m.x, y - m.y, z - m.z, m.M);
        return ret;
    }
    // octree.cpp:228-252
    double grad(double x, double y, double z) const {
        double rx = 0, ry = 0, rz = 0;
        const double lim = std::pow(p_.scale * p_.cell_r, 2);
        for (const mrisim_mnp& m : mnps_) {
            double dx = x - m.x, dy = y - m.y, dz = z - m.z, M = m.M;
            if (normsq(dx, dy, dz) > lim) {
                double divisor = std::pow(normsq(dx, dy, dz), 3.5);
                rx += 3e17 * M * dx * (1 * (dx * dx + dy * dy) - 4 * dz * dz) / divisor;
                ry += 3e17 * M * dy * (1 * (dx * dx + dy * dy) - 4 * dz * dz) / divisor;
                rz += 3e17 * M * dz * (3 * (dx * dx + dy * dy) - 2 * dz * dz) / divisor;
            }
        }
        return std::sqrt(normsq(rx, ry, rz));
    }

    // octree.cpp:262-333, written as pre-order emission into one flat vector: a node's
    // index is fixed when it is pushed, children follow contiguously, and
    // child[i] = index(child) - index(parent) is the reference's `curr->size()` offset.
    void descend(SubtreeOut& out, uint64_t mc, const std::vector<int>& res,
                 double x, double y, double z, double g) const {
        const size_t self = out.nodes.size();
        // res_start is the running resident cursor for every node (leaf or not), n_res = 0
        // for internal nodes: the canonical MRISCN01 form
        out.nodes.push_back(mrisim_node{mc, {0, 0, 0, 0, 0, 0, 0, 0}, (uint32_t)out.residents.size(), 0});

        if (g < min_g_ || (g / 2 * 1e-6 * grad(x + g / 2, y + g / 2, z + g / 2)) < max_prod_) {
            mrisim_node leaf = out.nodes[self];
            leaf.mc |= (1ULL << 63);
            leaf.n_res = (uint32_t)res.size();
            for (int i = 0; i < 8; i++) {
                // sic: g/4 in x but g/2 in y and z (octree.cpp:276-278)
                double xn = x + g / 4 + (double)((i >> 2) & 1) * g / 2;
                double yn = y + g / 2 + (double)((i >> 1) & 1) * g / 2;
                double zn = z + g / 2 + (double)((i >> 0) & 1) * g / 2;
                double B = field(xn, yn, zn);
                for (int r : res) {
                    const mrisim_mnp& m = mnps_[r];
                    double dx = xn - m.x, dy = yn - m.y, dz = zn - m.z;
                    B -= dipole(dx, dy, dz, m.M);
                }
                float Bf = (float)B;
                std::memcpy(&leaf.child[i], &Bf, sizeof Bf);
            }
            for (int r : res) out.residents.push_back(mnps_[r]);
            out.nodes[self] = leaf;
            return;
        }
        for (int i = 0; i < 8; i++) {
            double xn = x + (double)((i >> 2) & 1) * g / 2;
            double yn = y + (double)((i >> 1) & 1) * g / 2;
            double zn = z + (double)((i >> 0) & 1) * g / 2;
            std::vector<int> sub = filter(res, xn, yn, zn, g / 2);
            out.nodes[self].child[i] = (uint32_t)(out.nodes.size() - self);
            descend(out, (mc << 3) | (uint64_t)i, sub, xn, yn, zn, g / 2);
        }
    }
};

}  // namespace

// =====================================================================================
// mrisim_scene
// =====================================================================================
struct mrisim_scene {
    mrisim_params p;
    uint64_t seed = 0;
    std::vector<double> cells;            // Triple[num_cells]
    std::vector<mrisim_mnp> mnps;
    std::vector<int32_t> lookup;          // [hashDim^3][maxNeighbors]
    int32_t min_depth = 0, max_depth = 0;
    std::vector<uint32_t> subtree_sizes;
    std::vector<mrisim_node> nodes;
    std::vector<mrisim_mnp> residents;
    std::vector<mrisim_water> waters;
    std::vector<uint64_t> prefix;         // subtree roots in `nodes`
};

namespace mrisim_host {

// octree.cpp:99-128 host traversal; returns the flat node index of the leaf
uint64_t scene_find_leaf(const mrisim_scene& s, double wx, double wy, double wz) {
    uint64_t i = host_morton(s.min_depth, wx, wy, wz, s.p.bound);
    uint64_t n = s.prefix[i];
    int d = s.min_depth;
    while (!(s.nodes[n].mc >> 63)) {
        unsigned c = host_morton(d + 1, wx, wy, wz, s.p.bound) & 7;
        n += s.nodes[n].child[c];
        d++;
    }
    return n;
}

static double wall_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static int hw_threads(int n_threads) {
    if (n_threads > 0) return n_threads;
    unsigned h = std::thread::hardware_concurrency();
    return h ? (int)h : 1;
}

template <typename F>
static void parallel_for(int64_t n, int n_threads, F&& body) {
    n_threads = (int)std::min<int64_t>(hw_threads(n_threads), std::max<int64_t>(n, 1));
    if (n_threads <= 1) { for (int64_t i = 0; i < n; i++) body(i); return; }
    std::atomic<int64_t> next{0};
    std::vector<std::thread> th;
    for (int t = 0; t < n_threads; t++)
        th.emplace_back([&] { for (int64_t i; (i = next.fetch_add(1)) < n;) body(i); });
    for (auto& t : th) t.join();
}

// BacteriaBox.cpp:49-72: 3 draws per attempt (x, y, z), rejection on overlap with any
// earlier cell, no early exit, no periodic-image check.
static void throw_cells(mrisim_scene& s, HostRng& gen) {
    const mrisim_params& p = s.p;
    s.cells.reserve((size_t)p.num_cells * 3);
    for (int i = 0; i < p.num_cells; i++) {
        double x, y, z;
        for (;;) {
            x = p.cell_r + gen.rand_pos_double() * (p.bound - 2 * p.cell_r);
            y = p.cell_r + gen.rand_pos_double() * (p.bound - 2 * p.cell_r);
            z = p.cell_r + gen.rand_pos_double() * (p.bound - 2 * p.cell_r);
            bool clash = false;
            for (int j = 0; j < i; j++) {
                double dx = s.cells[3 * j] - x, dy = s.cells[3 * j + 1] - y, dz = s.cells[3 * j + 2] - z;
                if (normsq(dx, dy, dz) < 4 * p.cell_r * p.cell_r) clash = true;
            }
            if (!clash) break;
        }
        s.cells.push_back(x); s.cells.push_back(y); s.cells.push_back(z);
    }
}

// BacteriaBox.cpp:114-121: one zero-radius dipole of moment mmoment per cell
static void place_dipoles(mrisim_scene& s) {
    for (int i = 0; i < s.p.num_cells; i++)
        s.mnps.push_back(mrisim_mnp{s.cells[3 * i], s.cells[3 * i + 1], s.cells[3 * i + 2], 0, s.p.mmoment});
}

// SimulationBox.cpp:161-173 (only the sign of the result is ever used)
static bool inside_any_cell(const mrisim_scene& s, double x, double y, double z) {
    const double r2 = std::pow(s.p.cell_r, 2);
    bool in = false;
    for (int i = 0; i < s.p.num_cells; i++) {
        double dx = x - s.cells[3 * i], dy = y - s.cells[3 * i + 1], dz = z - s.cells[3 * i + 2];
        if (normsq(dx, dy, dz) < r2) in = true;
    }
    return in;
}

// BacteriaBox.cpp:80-105.  Every attempt consumes exactly three draws whatever its fate,
// so attempt j owns draws 3j..3j+2 of the stream: attempts are generated in batches,
// tested in parallel, and accepted in stream order -- identical to the sequential loop.
static void throw_waters(mrisim_scene& s, HostRng& gen, int64_t num_water, int n_threads) {
    const mrisim_params& p = s.p;
    const double offset = (p.bound - p.water_start_bound) / 2.0;
    s.waters.clear();
    s.waters.reserve((size_t)num_water);
    std::vector<double> cand;
    std::vector<uint8_t> bad;
    while ((int64_t)s.waters.size() < num_water) {
        const int64_t want = num_water - (int64_t)s.waters.size();
        const int64_t batch = p.avoid_intracellular_throw ? want + want / 8 + 64 : want;
        cand.resize((size_t)batch * 3);
        // draws must stay in stream order; if the batch overshoots we must not consume
        // draws the sequential loop would not have consumed, so remember the state
        std::vector<mrisim_rng> state_before((size_t)batch);
        for (int64_t j = 0; j < batch; j++) {
            state_before[j] = gen.r;
            cand[3 * j + 0] = offset + gen.rand_pos_double() * p.water_start_bound;
            cand[3 * j + 1] = offset + gen.rand_pos_double() * p.water_start_bound;
            cand[3 * j + 2] = offset + gen.rand_pos_double() * p.water_start_bound;
        }
        bad.assign((size_t)batch, 0);
        if (p.avoid_intracellular_throw) {
            const int64_t chunk = 256, n_chunks = (batch + chunk - 1) / chunk;
            parallel_for(n_chunks, n_threads, [&](int64_t c) {
                for (int64_t j = c * chunk; j < std::min(batch, (c + 1) * chunk); j++)
                    bad[j] = inside_any_cell(s, cand[3 * j], cand[3 * j + 1], cand[3 * j + 2]);
            });
        }
        int64_t j = 0;
        for (; j < batch && (int64_t)s.waters.size() < num_water; j++) {
            if (bad[j]) continue;
            mrisim_water w;
            std::memset(&w, 0, sizeof w);
            w.x = cand[3 * j]; w.y = cand[3 * j + 1]; w.z = cand[3 * j + 2];
            w.phase = 0; w.nearest = -1;
            s.waters.push_back(w);
        }
        if (j < batch) gen.r = state_before[j];     // rewind to just after the last attempt used
    }
}

// SimulationBox.cpp:262-290
static int build_lookup(mrisim_scene& s, int n_threads) {
    const mrisim_params& p = s.p;
    const double cubeLength = p.bound / p.hashDim;
    const double diagonal = std::sqrt(3) * cubeLength;
    const int64_t h3 = (int64_t)p.hashDim * p.hashDim * p.hashDim;
    s.lookup.assign((size_t)h3 * p.maxNeighbors, -1);
    std::atomic<int> overflow{0};
    const int64_t chunk = 64, n_chunks = (h3 + chunk - 1) / chunk;
    parallel_for(n_chunks, n_threads, [&](int64_t c) {
        for (int64_t i = c * chunk; i < std::min(h3, (c + 1) * chunk); i++) {
            double x = (i % p.hashDim) * cubeLength;
            double y = ((i / p.hashDim) % (p.hashDim)) * cubeLength;
            double z = i / (p.hashDim * p.hashDim) * cubeLength;
            int cnt = 0;
            for (int j = 0; j < p.num_cells; j++) {
                double dx = x - s.cells[3 * j], dy = y - s.cells[3 * j + 1], dz = z - s.cells[3 * j + 2];
                if (std::sqrt(normsq(dx, dy, dz)) < p.cell_r + diagonal) {
                    if (cnt < p.maxNeighbors) s.lookup[(size_t)i * p.maxNeighbors + cnt] = j;
                    cnt++;
                }
            }
            // the reference asserts ncells.size() < maxNeighbors (SimulationBox.cpp:281)
            if (cnt >= p.maxNeighbors) overflow = 1;
        }
    });
    return overflow;
}

// SimulationBox.cpp:348-527: images of the dipoles that lie within `border` of a face.
// Per axis an image goes to the one side the `if / else if` pair selects; the emplace
// order of the nested branches is the fixed sequence X, XY, Y, XYZ, XZ, YZ, Z restricted
// to the subsets of the active axes.
static void add_periodic_images(mrisim_scene& s) {
    const mrisim_params& p = s.p;
    static const int order[7] = {0b100, 0b110, 0b010, 0b111, 0b101, 0b011, 0b001};   // bit2 = x, bit1 = y, bit0 = z
    const size_t n0 = s.mnps.size();
    for (size_t i = 0; i < n0; i++) {
        const mrisim_mnp np = s.mnps[i];
        auto side = [&](double c) -> int { return (c + p.border > p.bound) ? -1 : ((c - p.border < 0) ? +1 : 0); };
        const int sx = side(np.x), sy = side(np.y), sz = side(np.z);
        const int active = (sx ? 4 : 0) | (sy ? 2 : 0) | (sz ? 1 : 0);
        for (int m : order) {
            if ((m & active) != m) continue;
            mrisim_mnp img = np;
            if (m & 4) img.x = sx < 0 ? np.x - p.bound : np.x + p.bound;
            if (m & 2) img.y = sy < 0 ? np.y - p.bound : np.y + p.bound;
            if (m & 1) img.z = sz < 0 ? np.z - p.bound : np.z + p.bound;
            s.mnps.push_back(img);
        }
    }
}

// SimulationBox.cpp:86-88 + octree.cpp:388-450
static void build_octree(mrisim_scene& s, int n_threads) {
    OctreeBuilder ob(s.p, s.mnps, kMaxProduct, kMaxG, kMinG);
    s.min_depth = ob.min_depth; s.max_depth = ob.max_depth;
    const int n_sub = 1 << (3 * s.min_depth);
    std::vector<SubtreeOut> parts((size_t)n_sub);
    parallel_for(n_sub, n_threads, [&](int64_t i) { ob.build_subtree_at((int)i, parts[(size_t)i]); });
    s.subtree_sizes.resize((size_t)n_sub);
    s.prefix.assign((size_t)n_sub + 1, 0);
    size_t n_nodes = 0, n_res = 0;
    for (int i = 0; i < n_sub; i++) { n_nodes += parts[i].nodes.size(); n_res += parts[i].residents.size(); }
    s.nodes.reserve(n_nodes); s.residents.reserve(n_res);
    for (int i = 0; i < n_sub; i++) {
        const uint32_t res_base = (uint32_t)s.residents.size();
        s.subtree_sizes[i] = (uint32_t)parts[i].nodes.size();
        s.prefix[i + 1] = s.prefix[i] + parts[i].nodes.size();
        for (mrisim_node n : parts[i].nodes) {
            n.res_start += res_base;
            s.nodes.push_back(n);
        }
        s.residents.insert(s.residents.end(), parts[i].residents.begin(), parts[i].residents.end());
        parts[i] = SubtreeOut();
    }
}

// ---- seams for the GPU octree builder (scene_internal.h) -----------------------------
void octree_depths(const mrisim_params& p, int32_t* min_depth, int32_t* max_depth) {
    static const std::vector<mrisim_mnp> none;
    OctreeBuilder ob(p, none, kMaxProduct, kMaxG, kMinG);
    *min_depth = ob.min_depth; *max_depth = ob.max_depth;
}
void octree_root(const mrisim_params& p, int min_depth, int offset,
                 double* x, double* y, double* z, double* g, uint64_t* mc) {
    static const std::vector<mrisim_mnp> none;
    OctreeBuilder ob(p, none, kMaxProduct, kMaxG, kMinG);
    ob.min_depth = min_depth;
    ob.root_of(offset, x, y, z, g, mc);
}
double octree_space_limit(const mrisim_params& p, double g) {
    static const std::vector<mrisim_mnp> none;
    return OctreeBuilder(p, none, kMaxProduct, kMaxG, kMinG).space_limit(g);
}
double octree_grad_limit(const mrisim_params& p) { return std::pow(p.scale * p.cell_r, 2); }
double octree_criterion(const mrisim_params& p, const std::vector<mrisim_mnp>& mnps,
                        double x, double y, double z, double g) {
    return OctreeBuilder(p, mnps, kMaxProduct, kMaxG, kMinG).criterion(x, y, z, g);
}
double octree_leaf_B(const mrisim_params& p, const std::vector<mrisim_mnp>& mnps, const int32_t* res, size_t n_res,
                     double xn, double yn, double zn) {
    return OctreeBuilder(p, mnps, kMaxProduct, kMaxG, kMinG).leaf_B(res, n_res, xn, yn, zn);
}

// the GPU builder's output, moved into the scene (same canonical form as build_octree above)
static void adopt_octree(mrisim_scene& s, OctreeOut& o) {
    s.min_depth = o.min_depth; s.max_depth = o.max_depth;
    s.subtree_sizes = std::move(o.subtree_sizes);
    s.nodes = std::move(o.nodes);
    s.residents = std::move(o.residents);
    s.prefix.assign(s.subtree_sizes.size() + 1, 0);
    for (size_t i = 0; i < s.subtree_sizes.size(); i++) s.prefix[i + 1] = s.prefix[i] + s.subtree_sizes[i];
}

// SimulationBox.cpp:131-147: std::sort (unstable) on the Morton code of the starting leaf,
// leaf bit included.  The permutation produced by libstdc++'s introsort depends only on
// the comparison results, so sorting (key, index) pairs reproduces the reference's order.
static int sort_waters(mrisim_scene& s, int n_threads, int device = -1) {
    struct Key { uint64_t mc; uint32_t idx; };
    const int64_t n = (int64_t)s.waters.size();
    std::vector<Key> keys((size_t)n);
    if (device < 0) {
        const int64_t chunk = 1024, n_chunks = (n + chunk - 1) / chunk;
        parallel_for(n_chunks, n_threads, [&](int64_t c) {
            for (int64_t i = c * chunk; i < std::min(n, (c + 1) * chunk); i++) {
                const mrisim_water& w = s.waters[(size_t)i];
                keys[(size_t)i] = Key{s.nodes[scene_find_leaf(s, w.x, w.y, w.z)].mc, (uint32_t)i};
            }
        });
    } else {                       // leaf codes from the GPU descent (scene_gpu.cu), same values
        std::vector<uint64_t> mc;
        std::string err;
        if (gpu_leaf_keys(s.p, s.min_depth, s.prefix, s.nodes, s.waters, device, mc, err)) {
            g_scene_error = "gpu_leaf_keys: " + err;
            return 1;
        }
        for (int64_t i = 0; i < n; i++) keys[(size_t)i] = Key{mc[(size_t)i], (uint32_t)i};
    }
    std::sort(keys.begin(), keys.end(), [](const Key& a, const Key& b) { return a.mc < b.mc; });
    std::vector<mrisim_water> sorted((size_t)n);
    for (int64_t i = 0; i < n; i++) sorted[(size_t)i] = s.waters[keys[(size_t)i].idx];
    s.waters.swap(sorted);
    return 0;
}

static int validate(const mrisim_params& p) {
    if (p.num_cells < 0 || p.num_water < 0 || p.hashDim <= 0 || p.maxNeighbors <= 0 ||
        !(p.bound > 0) || !(p.cell_r > 0) || !(p.water_start_bound > 0) || p.water_start_bound > p.bound) {
        g_scene_error = "mrisim_scene_build: invalid parameters";
        return 1;
    }
    return 0;
}

}  // namespace mrisim_host

// =====================================================================================
// C ABI
// =====================================================================================
using namespace mrisim_host;

extern "C" {

void mrisim_params_default(mrisim_params* p) {
    std::memset(p, 0, sizeof *p);
    p->g = 42.5781e6;
    p->scale = 2;
    p->border = 6;
    p->num_threads = 16;
    p->sprintSteps = 20000;
    p->num_water = 4032;
    p->num_cells = 270;
    p->cell_r = .55;
    p->mmoment = 3.5e-17;
    p->kick_mode = MRISIM_KICK_RANDOM;
    p->avoid_intracellular_throw = 1;
    p->phase_stdev = 2 * 3.14 * 1.5 * 1e-3 * 42 * 12 * 10e-3;
    p->phase_k = 5;
    p->bound = 30;
    p->water_start_bound = 10;
    p->hashDim = 20;
    p->maxNeighbors = 13;
    p->D_cell = .5547;
    p->D_extra = 1.6642;
    p->P_expr = 0.2;
    p->tau = 1e-6;
    p->totaltime = 40;
    p->taucp = 5.5;
    p->reflectIO = 0;
    p->reflectOI = 0;
    mrisim_params_derive(p);
}

void mrisim_params_derive(mrisim_params* p) {
    p->t = (int)(p->totaltime / p->tau);               // parameters.h:93
    p->tcp = (int)(p->taucp / p->tau);                 // parameters.h:95
    const double pInt = 1e-3;                          // cuda_sim.cu:44
    p->pfreq = (int)(pInt / p->tau);                   // cuda_sim.cu:45
    p->in_stdev = std::sqrt(M_PI * p->D_cell * p->tau);   // parameters.h:120
    p->out_stdev = std::sqrt(M_PI * p->D_extra * p->tau); // parameters.h:121
}

void mrisim_params_low_permeability(mrisim_params* p) {
    p->reflectIO = 1 - std::sqrt(p->tau / (6 * p->D_cell)) * 4 * p->P_expr;          // parameters.h:116
    p->reflectOI = 1 - ((1 - p->reflectIO) * std::sqrt(p->D_cell / p->D_extra));     // parameters.h:117
}

const char* mrisim_scene_last_error(void) { return g_scene_error.c_str(); }

// device < 0: octree on the host threads; device >= 0: octree_build.cu on that GPU
static int scene_build(const mrisim_params* p, uint64_t host_seed, int n_threads, int device,
                       mrisim_scene** out, mrisim_octree_build_stats* stats) {
    if (!p || !out) { g_scene_error = "mrisim_scene_build: null argument"; return 1; }
    if (validate(*p)) return 1;
    std::unique_ptr<mrisim_scene> s(new mrisim_scene);
    s->p = *p; s->seed = host_seed;
    HostRng gen(host_seed);
    // SimulationBox.cpp:77-97, in the reference's order (it fixes the RNG consumption order)
    throw_cells(*s, gen);
    place_dipoles(*s);
    std::string err;
    const double t0 = wall_ms();
    if (device < 0) {
        throw_waters(*s, gen, p->num_water, n_threads);
    } else if (gpu_throw_waters(s->p, s->cells, &gen.r, p->num_water, device, s->waters, err)) {
        g_scene_error = "mrisim_scene_build_gpu: " + err;
        return 3;
    }
    const double t1 = wall_ms();
    int overflow = 0;
    if (device < 0) {
        overflow = build_lookup(*s, n_threads);
    } else if (gpu_build_lookup(s->p, s->cells, device, s->lookup, &overflow, err)) {
        g_scene_error = "mrisim_scene_build_gpu: " + err;
        return 3;
    }
    if (overflow) {
        g_scene_error = "mrisim_scene_build: a lookup-table row holds >= maxNeighbors cells "
                        "(the reference asserts here, SimulationBox.cpp:281); raise maxNeighbors or hashDim";
        return 2;
    }
    const double t2 = wall_ms();
    add_periodic_images(*s);
    if (device < 0) {
        build_octree(*s, n_threads);
    } else {
        OctreeOut o;
        if (build_octree_gpu(s->p, s->mnps, device, n_threads, o, stats, err)) {
            g_scene_error = "mrisim_scene_build_gpu: " + err;
            return 3;
        }
        adopt_octree(*s, o);
    }
    const double t3 = wall_ms();
    if (sort_waters(*s, n_threads, device)) return 3;
    if (stats && device >= 0) {
        stats->scene_throw_ms = t1 - t0; stats->scene_lookup_ms = t2 - t1; stats->scene_sort_ms = wall_ms() - t3;
    }
    *out = s.release();
    return 0;
}

int mrisim_scene_build(const mrisim_params* p, uint64_t host_seed, int n_threads, mrisim_scene** out) {
    return scene_build(p, host_seed, n_threads, -1, out, nullptr);
}

int mrisim_scene_build_gpu(const mrisim_params* p, uint64_t host_seed, int n_threads, int device,
                           mrisim_scene** out, mrisim_octree_build_stats* stats) {
    if (device < 0) { g_scene_error = "mrisim_scene_build_gpu: device must be >= 0 (no CPU fallback on this entry point)"; return 1; }
    return scene_build(p, host_seed, n_threads, device, out, stats);
}

static int rethrow(mrisim_scene* s, const mrisim_params* p, int64_t num_water, uint64_t seed, int sort_by_morton, int device) {
    if (!s || num_water < 0) { g_scene_error = "mrisim_scene_rethrow_waters: bad argument"; return 1; }
    if (p) { s->p.avoid_intracellular_throw = p->avoid_intracellular_throw; s->p.water_start_bound = p->water_start_bound; }
    HostRng gen(seed);
    if (device < 0) {
        throw_waters(*s, gen, num_water, 0);
    } else {
        std::string err;
        if (gpu_throw_waters(s->p, s->cells, &gen.r, num_water, device, s->waters, err)) {
            g_scene_error = "mrisim_scene_rethrow_waters_gpu: " + err;
            return 3;
        }
    }
    if (sort_by_morton && sort_waters(*s, 0, device)) return 3;
    s->p.num_water = (int32_t)std::min<int64_t>(num_water, INT32_MAX);
    return 0;
}

int mrisim_scene_rethrow_waters(mrisim_scene* s, const mrisim_params* p, int64_t num_water, uint64_t seed, int sort_by_morton) {
    return rethrow(s, p, num_water, seed, sort_by_morton, -1);
}

int mrisim_scene_rethrow_waters_gpu(mrisim_scene* s, const mrisim_params* p, int64_t num_water, uint64_t seed,
                                    int sort_by_morton, int device) {
    if (device < 0) { g_scene_error = "mrisim_scene_rethrow_waters_gpu: device must be >= 0 (no CPU fallback on this entry point)"; return 1; }
    return rethrow(s, p, num_water, seed, sort_by_morton, device);
}

void mrisim_scene_free(mrisim_scene* s) { delete s; }

const double* mrisim_scene_cells(const mrisim_scene* s, int32_t* n) { if (n) *n = s->p.num_cells; return s->cells.data(); }
const mrisim_mnp* mrisim_scene_mnps(const mrisim_scene* s, int32_t* n) { if (n) *n = (int32_t)s->mnps.size(); return s->mnps.data(); }
const int32_t* mrisim_scene_lookup(const mrisim_scene* s, int32_t* h, int32_t* m) {
    if (h) *h = s->p.hashDim;
    if (m) *m = s->p.maxNeighbors;
    return s->lookup.data();
}
const mrisim_water* mrisim_scene_waters(const mrisim_scene* s, int32_t* n) { if (n) *n = (int32_t)s->waters.size(); return s->waters.data(); }
int mrisim_scene_octree(const mrisim_scene* s, int32_t* min_depth, int32_t* max_depth,
                        const uint32_t** subtree_sizes, int32_t* n_subtrees,
                        const mrisim_node** nodes, uint64_t* n_nodes,
                        const mrisim_mnp** residents, uint64_t* n_res) {
    if (!s) return 1;
    if (min_depth) *min_depth = s->min_depth;
    if (max_depth) *max_depth = s->max_depth;
    if (subtree_sizes) *subtree_sizes = s->subtree_sizes.data();
    if (n_subtrees) *n_subtrees = (int32_t)s->subtree_sizes.size();
    if (nodes) *nodes = s->nodes.data();
    if (n_nodes) *n_nodes = s->nodes.size();
    if (residents) *residents = s->residents.data();
    if (n_res) *n_res = s->residents.size();
    return 0;
}

// ---- MRISCN01 file ------------------------------------------------------------------
namespace {
struct FileHeader {
    char magic[8];
    uint64_t seed;
    int32_t num_cells, num_mnps, hashDim, maxNeighbors, min_depth, max_depth, num_water, n_subtrees;
    uint64_t n_nodes, n_res;
    double bound, cell_r, mmoment, water_start_bound;
    int64_t avoid;
};
static_assert(sizeof(FileHeader) == 8 + 8 + 32 + 16 + 32 + 8, "MRISCN01 header layout");
}

int mrisim_scene_save(const mrisim_scene* s, const char* path) {
    FILE* f = std::fopen(path, "wb");
    if (!f) { g_scene_error = std::string("mrisim_scene_save: cannot open ") + path; return 1; }
    FileHeader h;
    std::memcpy(h.magic, "MRISCN01", 8);
    h.seed = s->seed;
    h.num_cells = s->p.num_cells; h.num_mnps = (int32_t)s->mnps.size();
    h.hashDim = s->p.hashDim; h.maxNeighbors = s->p.maxNeighbors;
    h.min_depth = s->min_depth; h.max_depth = s->max_depth;
    h.num_water = (int32_t)s->waters.size(); h.n_subtrees = (int32_t)s->subtree_sizes.size();
    h.n_nodes = s->nodes.size(); h.n_res = s->residents.size();
    h.bound = s->p.bound; h.cell_r = s->p.cell_r; h.mmoment = s->p.mmoment; h.water_start_bound = s->p.water_start_bound;
    h.avoid = s->p.avoid_intracellular_throw ? 1 : 0;
    bool ok = std::fwrite(&h, sizeof h, 1, f) == 1;
    auto put = [&](const void* d, size_t bytes) { if (bytes) ok = ok && std::fwrite(d, 1, bytes, f) == bytes; };
    put(s->cells.data(), s->cells.size() * sizeof(double));
    put(s->mnps.data(), s->mnps.size() * sizeof(mrisim_mnp));
    put(s->lookup.data(), s->lookup.size() * sizeof(int32_t));
    put(s->subtree_sizes.data(), s->subtree_sizes.size() * sizeof(uint32_t));
    put(s->nodes.data(), s->nodes.size() * sizeof(mrisim_node));
    put(s->residents.data(), s->residents.size() * sizeof(mrisim_mnp));
    for (const mrisim_water& w : s->waters) { double v[4] = {w.x, w.y, w.z, w.phase}; put(v, sizeof v); }
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) { g_scene_error = "mrisim_scene_save: short write"; return 1; }
    return 0;
}

int mrisim_scene_load(const char* path, mrisim_scene** out) {
    if (!path || !out) { g_scene_error = "mrisim_scene_load: null argument"; return 1; }
    FILE* f = std::fopen(path, "rb");
    if (!f) { g_scene_error = std::string("mrisim_scene_load: cannot open ") + path; return 1; }
    struct Closer { FILE* f; ~Closer() { std::fclose(f); } } closer{f};
    auto bad = [&](const char* why) { g_scene_error = std::string("mrisim_scene_load: malformed scene file ") + path + " (" + why + ")"; return 1; };
    FileHeader h;
    if (std::fread(&h, sizeof h, 1, f) != 1 || std::memcmp(h.magic, "MRISCN01", 8) != 0) return bad("not an MRISCN01 header");
    // Nothing in the header is trusted: every count is checked for sign and for the bounds the rest of the
    // library assumes, and the payload size they imply must equal the file's before anything is allocated.
    if (h.num_cells < 0 || h.num_mnps < 0 || h.num_water < 0 || h.maxNeighbors <= 0 || h.maxNeighbors > 255 ||
        h.hashDim <= 0 || h.hashDim > 1290 || h.min_depth < 0 || h.min_depth > 8 || h.max_depth < h.min_depth || h.max_depth > 20 ||
        h.n_subtrees != (1 << (3 * h.min_depth)) || h.n_nodes >= (1ull << 31) || h.n_res >= (1ull << 32) ||
        !(h.bound > 0.0) || !(h.cell_r > 0.0))
        return bad("header field out of range");
    const uint64_t h3 = (uint64_t)h.hashDim * h.hashDim * h.hashDim;
    const uint64_t payload = (uint64_t)h.num_cells * 3 * sizeof(double) + (uint64_t)h.num_mnps * sizeof(mrisim_mnp) +
                             h3 * (uint64_t)h.maxNeighbors * sizeof(int32_t) + (uint64_t)h.n_subtrees * sizeof(uint32_t) +
                             h.n_nodes * sizeof(mrisim_node) + h.n_res * sizeof(mrisim_mnp) + (uint64_t)h.num_water * 4 * sizeof(double);
    if (std::fseek(f, 0, SEEK_END) != 0) return bad("cannot seek");
    const long size = std::ftell(f);
    if (size < 0 || (uint64_t)size != sizeof h + payload) return bad("file size does not match the header");
    if (std::fseek(f, (long)sizeof h, SEEK_SET) != 0) return bad("cannot seek");
    try {
        std::unique_ptr<mrisim_scene> s(new mrisim_scene);
        bool ok = true;
        auto get = [&](void* d, size_t bytes) { if (bytes) ok = ok && std::fread(d, 1, bytes, f) == bytes; };
        mrisim_params_default(&s->p);
        s->seed = h.seed;
        s->p.num_cells = h.num_cells; s->p.hashDim = h.hashDim; s->p.maxNeighbors = h.maxNeighbors;
        s->p.num_water = h.num_water; s->p.bound = h.bound; s->p.cell_r = h.cell_r; s->p.mmoment = h.mmoment;
        s->p.water_start_bound = h.water_start_bound; s->p.avoid_intracellular_throw = (int32_t)h.avoid;
        s->min_depth = h.min_depth; s->max_depth = h.max_depth;
        s->cells.resize((size_t)h.num_cells * 3); get(s->cells.data(), s->cells.size() * sizeof(double));
        s->mnps.resize((size_t)h.num_mnps); get(s->mnps.data(), s->mnps.size() * sizeof(mrisim_mnp));
        s->lookup.resize((size_t)(h3 * (uint64_t)h.maxNeighbors)); get(s->lookup.data(), s->lookup.size() * sizeof(int32_t));
        s->subtree_sizes.resize((size_t)h.n_subtrees); get(s->subtree_sizes.data(), s->subtree_sizes.size() * sizeof(uint32_t));
        s->nodes.resize(h.n_nodes); get(s->nodes.data(), s->nodes.size() * sizeof(mrisim_node));
        s->residents.resize(h.n_res); get(s->residents.data(), s->residents.size() * sizeof(mrisim_mnp));
        s->waters.resize((size_t)h.num_water);
        for (mrisim_water& w : s->waters) {
            double v[4]; get(v, sizeof v);
            std::memset(&w, 0, sizeof w); w.x = v[0]; w.y = v[1]; w.z = v[2]; w.phase = v[3]; w.nearest = -1;
        }
        if (!ok) return bad("short read");
        s->prefix.assign(s->subtree_sizes.size() + 1, 0);
        for (size_t i = 0; i < s->subtree_sizes.size(); i++) s->prefix[i + 1] = s->prefix[i] + s->subtree_sizes[i];
        if (s->prefix.back() != s->nodes.size()) return bad("subtree sizes do not cover the node array");
        // node links: resident ranges inside the residents array, child offsets inside their subtree
        for (size_t t = 0; t < s->subtree_sizes.size(); t++)
            for (uint64_t j = s->prefix[t]; j < s->prefix[t + 1]; j++) {
                const mrisim_node& nd = s->nodes[j];
                if (nd.mc >> 63) { if ((uint64_t)nd.res_start + nd.n_res > h.n_res) return bad("resident range out of bounds"); }
                else for (int c = 0; c < 8; c++) if (nd.child[c] == 0 || j + nd.child[c] >= s->prefix[t + 1]) return bad("child offset leaves its subtree");
            }
        for (int32_t v : s->lookup) if (v < -1 || v >= h.num_cells) return bad("lookup table names a cell outside the cells set");
        *out = s.release();
    } catch (const std::exception& e) {                      // bad_alloc and friends must not cross the C boundary
        g_scene_error = std::string("mrisim_scene_load: ") + e.what();
        return 1;
    }
    return 0;
}

}  // extern "C"
