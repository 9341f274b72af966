// Parameter block for the tcgen05 attention kernel.
#pragma once
#include <cuda.h>
#include <stdint.h>

namespace b200 {

// One attention problem: queries are rows [q_row0, q_row0+q_len) of the Q matrix, keys/values are rows
// [kv_row0, kv_row0+kv_len) of the K / V matrices.  When AttnParams::len_override is set the lengths are
// read on the device from len_override[*_len_slot] (adaptive width pruning, no host synchronisation).
struct AttnSeq {
    int q_row0, q_len, kv_row0, kv_len;
    int q_len_slot, kv_len_slot;
};

struct AttnParams {
    const AttnSeq* seqs;      // device array, one entry per blockIdx.z
    const int* len_override;  // optional device array of current sequence lengths
    int q_col0, k_col0, v_col0;  // first column of head 0 inside the Q / K / V matrices
    float scale;                 // logits are scale * (q . k)
    void* out;                   // bf16 [tokens, ld_out]
    long long ld_out;
    int out_col0;
    // filled by the launcher: work items = (pair of query tiles, head, sequence)
    int q_tiles, heads, num_items;
};

}  // namespace b200
