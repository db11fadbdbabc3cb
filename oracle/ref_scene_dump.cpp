// TEST INFRASTRUCTURE (oracle side) -- not product code.
//
// Driver around the UNMODIFIED reference host scene code
// (/root/reference/src/{BacteriaBox,SimulationBox,octree}.cpp, xorshift.h), compiled in
// place by oracle/Makefile into oracle/_ref/.  It replaces the non-reproducible seed of
// simulateWaters() (cuda_sim.cu:911-912, time(NULL)+rd()) with argv[1], runs
// BacteriaBox::populateSimulation() (SimulationBox.cpp:77-97) and dumps every array the
// GPU driver consumes (cuda_sim.cu:926-964) in the "MRISCN01" interchange layout that the
// product scene builder (mri_nanoparticle_simulation_b200/csrc/scene.cpp) also writes, so
// parity of the scene is a byte comparison of two files.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <vector>
#include <iostream>
#include "parameters.h"
#include "BacteriaBox.h"
#include "octree.h"

ParameterStruct p(0);   // the reference defines this global in cuda_sim.cu:41

template <typename T> static void put(FILE* f, const T& v) { fwrite(&v, sizeof(T), 1, f); }

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s <seed> <out.scn>\n", argv[0]); return 2; }
    uint64_t seed = strtoull(argv[1], nullptr, 0);
    XORShift<uint64_t> gen(seed);
    BacteriaBox box(&gen);
    box.populateSimulation();

    Octree* oct = box.getOctree();
    const int n_sub = 1 << (3 * oct->min_depth);
    uint64_t n_nodes = 0, n_res = 0;
    for (int i = 0; i < n_sub; i++) {
        n_nodes += oct->space[i].size();
        for (auto& n : oct->space[i]) if (n.resident) n_res += n.resident->size();
    }

    FILE* f = fopen(argv[2], "wb");
    if (!f) { perror("fopen"); return 1; }
    fwrite("MRISCN01", 1, 8, f);
    put<uint64_t>(f, seed);
    put<int32_t>(f, p.num_cells);
    put<int32_t>(f, box.getMNPCount());
    put<int32_t>(f, p.hashDim);
    put<int32_t>(f, p.maxNeighbors);
    put<int32_t>(f, oct->min_depth);
    put<int32_t>(f, oct->max_depth);
    put<int32_t>(f, p.num_water);
    put<int32_t>(f, n_sub);
    put<uint64_t>(f, n_nodes);
    put<uint64_t>(f, n_res);
    put<double>(f, p.bound);
    put<double>(f, p.cell_r);
    put<double>(f, p.mmoment);
    put<double>(f, p.water_start_bound);
#ifdef AVOID_INTRACELLULAR_THROW
    put<int64_t>(f, 1);
#else
    put<int64_t>(f, 0);
#endif

    Triple* cells = box.getCells();
    for (int i = 0; i < p.num_cells; i++) { put(f, cells[i].x); put(f, cells[i].y); put(f, cells[i].z); }
    MNP_info* m = box.getMNPs();
    for (int i = 0; i < box.getMNPCount(); i++) { put(f, m[i].x); put(f, m[i].y); put(f, m[i].z); put(f, m[i].r); put(f, m[i].M); }
    int** lut = box.getLookupTable();
    for (int i = 0; i < p.hashDim * p.hashDim * p.hashDim; i++) fwrite(lut[i], sizeof(int), p.maxNeighbors, f);
    for (int i = 0; i < n_sub; i++) put<uint32_t>(f, (uint32_t)oct->space[i].size());
    uint32_t res_cursor = 0;
    for (int i = 0; i < n_sub; i++)
        for (auto& n : oct->space[i]) {
            put<uint64_t>(f, n.mc);
            for (int k = 0; k < 8; k++) put<uint32_t>(f, n.child[k].idx);
            uint32_t nr = n.resident ? (uint32_t)n.resident->size() : 0;
            put<uint32_t>(f, res_cursor);
            put<uint32_t>(f, nr);
            res_cursor += nr;
        }
    for (int i = 0; i < n_sub; i++)
        for (auto& n : oct->space[i])
            if (n.resident)
                for (auto& r : *n.resident) { put(f, r.x); put(f, r.y); put(f, r.z); put(f, r.r); put(f, r.M); }
    water_info* w = box.getWaters();
    for (int i = 0; i < p.num_water; i++) { put(f, w[i].x); put(f, w[i].y); put(f, w[i].z); put(f, w[i].phase); }
    fclose(f);

    // Probes of the reference's own host traversal: argv[3] = file of (x,y,z) double
    // triples, argv[4] = output of (uint64 leaf mc, double Octree::get_field) pairs.
    if (argc > 4) {
        FILE* in = fopen(argv[3], "rb");
        FILE* out = fopen(argv[4], "wb");
        if (!in || !out) { perror("probe files"); return 1; }
        double q3[3];
        while (fread(q3, sizeof(double), 3, in) == 3) {
            water_info q; q.x = q3[0]; q.y = q3[1]; q.z = q3[2];
            oct_node* leaf = oct->get_voxel(&q);
            put<uint64_t>(out, leaf->mc);
            put<double>(out, oct->get_field(&q, leaf));
        }
        fclose(in); fclose(out);
    } else if (argc > 3) {
        // A.6-style probe: leaf holding (15,15,15) and the reference host field there.
        water_info q; q.x = q.y = q.z = 15.0;
        oct_node* leaf = oct->get_voxel(&q);
        printf("PROBE mc=%016llx field=%.17g\n", (unsigned long long)leaf->mc, oct->get_field(&q, leaf));
    }
    return 0;
}
