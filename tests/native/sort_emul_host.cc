// Test helper: the device-side std::sort emulation (mam3slam_b200/csrc/introsort_emul.h) compiled for the host, so the
// CPU suite can compare its element movement with the real libstdc++ std::sort without a GPU.
#include "../../mam3slam_b200/csrc/introsort_emul.h"
#include <vector>
extern "C" void emul_sort_nodes(const int* size, const int* ulx, int n, int* perm, int heap_only) {
    std::vector<ise::Elem> e(n > 0 ? n : 1);
    for (int i = 0; i < n; ++i) { e[i].key = ((uint64_t)(uint32_t)size[i] << 32) | (uint32_t)ulx[i]; e[i].id = i; e[i].pad = 0; }
    // heap_only: 0 = std::sort emulation, 1 = heap sort alone, 2 / 3 = the round-based schedule the GPU warp runs (two work orders)
    if (heap_only == 1) ise::heap_sort(e.data(), 0, n);
    else if (heap_only >= 2) ise::sort_rounds(e.data(), n, heap_only == 3);
    else ise::sort(e.data(), n);
    for (int i = 0; i < n; ++i) perm[i] = e[i].id;
}
