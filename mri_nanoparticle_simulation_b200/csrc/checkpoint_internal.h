// checkpoint_internal.h -- seam between checkpoint.cpp and group.cpp.  Not part of the C ABI.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "mrisim.h"

namespace mrisim_ckpt {

// host copy of the state of a contiguous walker range
struct ShardState {
    uint64_t first = 0;                       // GLOBAL id of its first walker
    std::vector<mrisim_water> waters;
    std::vector<mrisim_rng_state> rng;
    bool has_order = false;
    std::vector<uint32_t> perm, keys;         // walker order of the fused path, indices local to the range
};
struct ShardEntry { uint64_t first, count; bool has_order; uint64_t order_offset; };

int gather(mrisim_ctx* ctx, ShardState& out);                                     // device -> host
int scatter(mrisim_ctx* ctx, const mrisim_checkpoint_info& info, const ShardState& st);   // host -> device (set_waters, set_rng, set_time, set_order)
int write_file(const char* path, const mrisim_checkpoint_info& info, const std::vector<ShardState>& shards);
int read_info(const char* path, mrisim_checkpoint_info* out, std::vector<ShardEntry>* shards);
int read_range(const char* path, uint64_t lo, uint64_t count, ShardState& out);   // count == 0: to the end
int check_compatible(mrisim_ctx* ctx, const mrisim_checkpoint_info& info);
const std::string& last_error();
void set_error(const std::string& m);

}  // namespace mrisim_ckpt
