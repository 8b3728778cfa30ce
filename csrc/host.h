// hetseq_b200 -- host-side helpers shared by the launchers in csrc/*.cu.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <utility>

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

// 3-D variant: [outer][mid][inner], box = box_inner x box_mid x 1; rows >= mid are zero-filled (out of bounds).
int make_tmap_3d_16b(CUtensorMap* out, const void* base, uint64_t inner, uint64_t mid, uint64_t outer,
                     uint64_t ld_mid, uint64_t ld_outer, uint32_t box_inner, uint32_t box_mid);

// HETSEQ_PDL=0 switches programmatic dependent launch off (every launch then serialises like <<<>>>).
bool pdl_enabled();

// <<<grid, block, smem, stream>>> with the programmatic-stream-serialization attribute: the kernel may start
// (up to its pdl_wait()) while the previous kernel of the stream drains. Only for kernels whose device code
// calls pdl_wait() before its first global memory access. Compile-time __cluster_dims__ are honoured.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

}  // namespace hs
