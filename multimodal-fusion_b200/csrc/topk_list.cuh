// Ascending (value, arrival-order) top-k list held in registers.
#pragma once

namespace hg {

// Insert (s, idx) after every entry <= s.  Precondition: s < v[KL-1].
// Fully predicated: slot q takes its predecessor if that one is larger than s,
// takes s if it is the first slot larger than s, and is kept otherwise.
template <int KL>
__device__ __forceinline__ void list_insert(float (&v)[KL], int (&ix)[KL], float s, int idx) {
#pragma unroll
    for (int q = KL - 1; q > 0; --q) {
        const bool from_prev = v[q - 1] > s;
        const bool here = v[q] > s;
        ix[q] = from_prev ? ix[q - 1] : (here ? idx : ix[q]);
        v[q] = from_prev ? v[q - 1] : (here ? s : v[q]);
    }
    if (v[0] > s) { v[0] = s; ix[0] = idx; }
}

}  // namespace hg
