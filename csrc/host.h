// hetseq_b200 -- host-side helpers shared by the launchers in csrc/*.cu.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace hs {

void count_launch(int n);
// Bumps the launch counter and converts a failed launch into a non-zero return + hs_last_error().
int post_launch(const char* kernel);
int num_sms();
// Row slabs used by the two-stage column sums (scratch = slabs * N floats).
int colsum_slabs(int M);

// 2-D tiled tensor map (128-byte swizzle) over a row-major 16-bit matrix:
// inner dimension `inner` elements (contiguous), `outer` rows of pitch `ld` elements.
int make_tmap_2d_16b(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t ld,
                     uint32_t box_inner, uint32_t box_outer);

}  // namespace hs
