// Test harness (CPU): the assembly kernel's literal restatement of libstdc++'s std::sort
// (tf-pose-estimation_b200/csrc/std_sort_literal.h) against the REAL std::sort with the reference's comparator
// (comp_candidate, tf_pose/pafprocess/pafprocess.cpp:239-241: a.score > b.score), on the kernel's key layout
// (limb << 56 | ~score_bits << 24 | idx1 << 12 | idx2).  Built by tests/test_host.py with g++.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

#define SS_TEST_HOOK 1
#include "../../tf-pose-estimation_b200/csrc/std_sort_literal.h"

static inline float score_of(uint64_t key) {
    const uint32_t bits = 0xffffffffu - (uint32_t)((key >> 24) & 0xffffffffu);
    float f;
    std::memcpy(&f, &bits, 4);
    return f;
}

extern "C" {
// how often the depth-limit heap-sort fallback of the restatement has run so far
int harness_heap_sort_calls(void) { return tfpd::ss_heap_sort_calls; }
// sorts `keys` with the literal restatement
void harness_literal_sort(uint64_t *keys, int n) { tfpd::std_sort_literal((tfpd::CandKey *)keys, n); }
// sorts `keys` with std::sort and the reference's comparator on the float scores
void harness_std_sort(uint64_t *keys, int n) {
    std::vector<uint64_t> v(keys, keys + n);
    std::sort(v.begin(), v.end(), [](uint64_t a, uint64_t b) { return score_of(a) > score_of(b); });
    std::copy(v.begin(), v.end(), keys);
}
}
